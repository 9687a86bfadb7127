// Pixel-GEMM engine: executes the layer programs of psob/lowering.py on sm_100a.
//
//   one persistent CTA per SM (or a cluster of two: tcgen05 cta_group::2) over work items (particle tile of 128 / 256
//   rows, chain of jobs); blockDim = 32 * (4 * epilogue groups + 1 + MMA issuers)
//   warps 0..7   epilogue, 1-2 groups x four TMEM lane quadrants: tcgen05.ld -> bias + activation -> bf16 -> swizzled
//                staging tile -> TMA store per 64-column chunk; or the fused-head reduction (kDot); or direct stores
//                (image / score outputs)
//   next warp    TMA producer: A tile [128 particles x 64 ch] (+ the weight tile or, in a pair, this CTA's half of it,
//                when weights are streamed) per step, SWIZZLE_128B, n_stages-deep mbarrier ring
//   last warp(s) MMA issuer (pair: leader CTA only): 4 x tcgen05.mma (K = 16) per op into a TMEM accumulator
//                (2 buffers of nt columns), tcgen05.commit frees the stage
//
// Replaces the cuDNN calls behind nn.ConvTranspose2d / nn.Conv2d in Generator.forward /
// Discriminator.forward (src/utils/util_dcgan.py:103-149).  A SIMT interpreter of the same tables
// is kept for bring-up and tests (PSOB_IMPL_SIMT); the product path is PSOB_IMPL_TCGEN05.
#include "psob_common.cuh"
#include "umma.cuh"

#include <cuda.h>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include <new>
#include <type_traits>

namespace psob {

// A/B tuning knobs (PSOB_* environment variables) are read only when PSOB_TUNING=1: a stray PSOB_* variable in a user's
// shell must not change how the product path executes.  tools/ and the mode-equivalence tests set PSOB_TUNING=1.
static const char* knob(const char* name) {
  const char* t = getenv("PSOB_TUNING");
  if (!t || t[0] != '1' || t[1] != 0) return nullptr;
  return getenv(name);
}

constexpr int kMaxStages = 8;
constexpr int kTileM = 128;
constexpr int kKChunk = 64;
constexpr int kATileBytes = kTileM * kKChunk * 2;   // 16 KB
constexpr int kSmemLimit = 232448;                  // 227 KB per CTA on sm_100
constexpr int kBarBytes = 256;
constexpr int kMaxBiasRowsSmem = 8;
constexpr int kMaxStepsSmem = 2048;                // step descriptors staged in SMEM (16 B each)
constexpr int kMaxJobsSmem = 512;
constexpr int kMaxSoloResident = 96 * 1024;       // resident weights of one CTA before the pair mode takes over
constexpr int kMaxPairResident = 128 * 1024;      // per CTA of a pair
constexpr int kBandPitch = 30;                    // accumulator columns per padded image row of an image band (psob/lowering.py)
// step word 1 (meta): bits 0-2 number of MMA ops (1..4), bit 3 weights resident, bit 4 some op targets the NEXT job
constexpr int kMetaOpsMask = 7, kMetaResident = 8, kMetaHasNext = 16, kMetaEye = 32;
constexpr int kEyeBytes = 64 * 128;   // built-in 64 x 64 bf16 identity tile (SWIZZLE_128B rows): the B operand of an identity op
// op word: bits 0-15 weight row / 16, bits 16-23 accumulator column / 2, bit 24 NEXT job's buffer, bit 25 accumulate
// job word 7: bit 0 = the job has a chain predecessor, bits 1.. = (base column of the job's step list) / 8 (psob/lowering.py pack_tables)
// x / d for x < 2^31 as (x * magic) >> shift: L = ceil(log2 d), magic = ceil(2^(31 + L) / d) < 2^32, shift = 31 + L.
// Exact because magic * d - 2^(31 + L) < d <= 2^L and x < 2^31.
__host__ __device__ __forceinline__ uint32_t fast_div(uint32_t x, uint32_t magic, uint32_t shift) {
  return (uint32_t)(((unsigned long long)x * magic) >> shift);
}
static inline void fast_div_magic(uint32_t d, uint32_t* magic, uint32_t* shift) {
  uint32_t L = 0;
  while ((1ull << L) < d) ++L;
  const unsigned long long num = 1ull << (31 + L);
  *magic = (uint32_t)((num + d - 1) / d);
  *shift = 31 + L;
}
// j1.w = chain flag | (base column / 8) << 1; the DEVICE copy of the table carries the job's valid 64-column chunks in bits 28..31
// (filled in by psob_layer_create from the column map and the pixel mask, so the epilogue does not look them up per job)
__host__ __device__ __forceinline__ int job_abase(const int4& j1) { return (int)((((unsigned)j1.w & 0x0FFFFFFFu) >> 1) << 3); }
__host__ __device__ __forceinline__ int op_b_row(int op) { return (op & 0xFFFF) << 4; }
__host__ __device__ __forceinline__ int op_d_col(int op) { return ((op >> 16) & 0xFF) << 1; }
__host__ __device__ __forceinline__ int op_next(int op) { return (op >> 24) & 1; }
__host__ __device__ __forceinline__ int op_acc(int op) { return (op >> 25) & 1; }
__host__ __device__ __forceinline__ int op_eye(int op) { return (op >> 26) & 1; }
// op width: the op covers (1 + w) * mma_n accumulator columns and weight rows (bits 28..29; resident weights, solo CTAs) --
// neighbouring output pixels fed by one input pixel through neighbouring taps are one wider tcgen05.mma
__host__ __device__ __forceinline__ int op_width(int op) { return 1 + (((unsigned)op >> 28) & 3); }
__host__ __device__ __forceinline__ int op_region(int op) { return (op >> 27) & 1; }   // set by psob_layer_create: parity of the op's column region (2 issuers)   // D[:, d_col : d_col + 64] (+)= A (N = 64 against the identity tile)                  // job descriptors staged in SMEM (32 B each)
constexpr int kMaxEpiGroups = 3;                  // epilogue groups of 4 warps (one per TMEM lane quadrant); 3: lean kernels only
constexpr int kMaxThreads = 32 * (4 * 2 + 3);     // + TMA warp + 1 or 2 MMA warps; blockDim = 32 * (4 * epi_groups + 1 + issuers)
// lean kernels: up to 3 groups + TMA warp + ONE MMA warp = 14 warps.  ptxas keeps the lean epilogue in 128 registers without a
// spill under this bound (it takes 168 under the 352-thread bound because it may, not because it must).
constexpr int kMaxThreadsLean = 32 * (4 * kMaxEpiGroups + 2);
// The SMSP arbiter favours the highest warp id: the two single-thread roles that must never wait for
// an issue slot (TMA producer, MMA issuer) get the top ids; epilogue warps 0..7 map to TMEM quadrant w % 4.
// (warp ids 4 * epi_groups and 4 * epi_groups + 1, see kTmaWarp / kMmaWarp in the kernel)

struct EngineParams {
  const int4* steps;         // 2 x int4 per step: {a_col, meta, op0, op1} {op2, op3, 0, 0}
  const int4* jobs;          // 2 x int4 per job
  const int2* chains;        // {first job, job count}: jobs of a chain run back to back on one CTA (pair)
  const int* col_map;
  const int* col_map32;
  const float* bias;
  __nv_bfloat16* y;
  float* y32;
  const __nv_bfloat16* res;  // optional residual, same layout as y (TMA-store epilogue): out = act(acc + bias + res)
  long long res_pitch;
  long long P;
  long long n_items;
  int n_jobs, n_chains;
  int mma_n, nt, cc, act, out_pitch, out32_pitch, vec8;
  int b_resident, n_wrows, n_bias_rows, n_steps;
  int n_stages, stage_bytes, res_off, bar_off, stg_off;
  int vec4, strided;
  int no_bias;               // every bias entry is 0 (e.g. folded into the K padding): the chunked epilogue skips the loads and adds
  uint32_t nc_magic, nc_shift, rot_magic, rot_shift;   // x / n_chains, x / rot_div for x < 2^31 as (x * magic) >> shift (fast_div_magic)
  int rot_div;               // chain rotation: item (tile t, slot r) runs chain (r + t / rot_div) % n_chains (0: chain r) -- see pick_rotation
  int chain_ident;           // every chain is one job (chain c == job c): the chain table is not read
  int step_stride;           // int4 words per step in the shared-memory copy of the step table (1 when no step has > 2 ops)
  int warp_store;            // every epilogue warp stores its own 32 rows of a chunk (32-row TMA box): no barriers inside a group
  int issuers;               // MMA-issuing warps (1 or 2): one thread sustains only ~1 tcgen05.mma per 85-94 cycles (tools/umma_bench.cu)
  int issue_split;           // 2 issuers: ownership of an op by (accumulator buffer + column region) parity (1) or by buffer only (0)
  int half_store;            // warp stores with a single staging tile: the warp's 4 KB slice is two 32-column halves (SWIZZLE_64B);
                             // 2: QUARTER stores -- a 2 KB slice per warp, two 16-column quarters (SWIZZLE_32B), 8 KB of staging per group
  int stg_group_bytes;       // staging bytes per epilogue group (stg_bufs * 16 KB, quarter stores: 8 KB)
                             // published by separate TMA stores, so one half drains while the other is written
  int band;                  // image-band epilogue (lower_convt_image_band): 4 chunks of two padded image rows each; 2: tiled band
  int band_w;                // tiled band: width of the image in pixels (row pitch of the fp32 image / the target)
  int epi_groups;            // TMA-store epilogue: 2 groups of 4 warps alternate chunks, or 1 group (frees a staging tile for the ring)
  int in_tiled, out_tiled, n_in_chunks, n_out_chunks;   // tile-major activations: [particle tile][64-col chunk][128][64]
  int bias_off, steps_off, jobs_off, chains_off, stg_bufs;   // SMEM table offsets from the table base (-1: stays in global)
  unsigned long long* prof;   // optional [gridDim][32] role wait-cycle counters (debug)
  uint32_t idesc, tmem_cols;
  uint32_t idesc_eye;        // instruction descriptor of identity ops (M = 128, N = 64)
  int eye;                   // the layer has identity ops: the kernel builds the identity tile at res_off
  // dot epilogue (kDot): the layer's bf16 output never leaves the SM; every output row is reduced against dot_w
  // (bf16 [out_pitch], the weights of a following full-map conv with ONE output channel) as it comes out of TMEM
  const uint16_t* dot_w;
  float* dot_part;           // fp32 [n_jobs * epi_groups][P]: one partial sum per (job, epilogue group) and particle
  // reconstruction epilogue (image layers): every thread owns one particle row of the image, so the per-particle
  // sum((target - G(z))^2) of Discovery.fitness_pso_inverter (src/pso/util_discovery.py:97) is accumulated from the fp32
  // pixel it already holds in a register; partial sums per (job, epilogue group), fixed order -> deterministic
  const void* recon_target;  // [P, recon_pitch] fp32 or bf16, laid out like the fp32 image output
  float* recon_part;         // fp32 [n_jobs * (band ? epi_groups : 1)][P]
  int recon_bf16, recon_pitch;
};

__device__ __forceinline__ float recon_load(const void* base, int is_bf16, long long idx) {
  return is_bf16 ? __bfloat162float(__ldg(reinterpret_cast<const __nv_bfloat16*>(base) + idx))
                 : __ldg(reinterpret_cast<const float*>(base) + idx);
}

__device__ __forceinline__ float apply_act(float v, int act) {
  switch (act) {
    case PSOB_ACT_RELU: return fmaxf(v, 0.f);
    case PSOB_ACT_LEAKY: return v > 0.f ? v : 0.2f * v;
    case PSOB_ACT_TANH: return tanhf(v);
    case PSOB_ACT_SIGMOID: return 1.f / (1.f + expf(-v));
    default: return v;
  }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

// timed mbarrier wait: adds the cycles spent waiting to `acc` when role profiling is on
__device__ __forceinline__ void mbar_wait_t(uint32_t bar, uint32_t parity, bool prof, unsigned long long& acc) {
  if (!prof) { ptx::mbar_wait(bar, parity); return; }
  long long t0 = clock64();
  ptx::mbar_wait(bar, parity);
  acc += (unsigned long long)(clock64() - t0);
}

// tanh(v) = 1 - 2 / (exp(2v) + 1) on the SFU (ex2.approx + rcp.approx, 5 instructions, no branch).  libdevice's tanhf is
// ~40 instructions with a divergent small-argument branch, and the image layer's epilogue evaluates 224 of them per
// thread and job next to the single MMA-issuing thread on the same scheduler.  Absolute error <= ~3e-7 over the whole
// range (saturates to +-1 through exp -> inf / 0), far below the bf16 image's 4e-3 and the 2e-2 image gate.
__device__ __forceinline__ float tanh_sfu(float v) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(v * 2.8853900817779268f));     // exp(2v) = 2^(2v log2 e)
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.f));
  return fmaf(-2.f, r, 1.f);
}

template <int kAct>
__device__ __forceinline__ float act_fn(float v) {
  if (kAct == PSOB_ACT_RELU) return fmaxf(v, 0.f);
  if (kAct == PSOB_ACT_LEAKY) return fmaxf(v, 0.2f * v);      // max(v, 0.2v) == leaky_relu(v, 0.2)
  if (kAct == PSOB_ACT_TANH) return tanh_sfu(v);
  if (kAct == PSOB_ACT_SIGMOID) return 1.f / (1.f + expf(-v));
  return v;
}

// Two accumulator values -> bias -> activation -> one packed bf16x2 word, on Blackwell's packed fp32 pipe:
// add.rn.f32x2 (FADD2) for the bias, mul.rn.f32x2 (FMUL2) for the leaky slope, cvt.rn.relu.bf16x2.f32 folds ReLU into
// the conversion.  Each lane is an ordinary round-to-nearest fp32 op, so the results are bit-identical to the
// scalar form; the epilogue is issue-bound (2 warps per scheduler), and this halves its arithmetic instructions.
template <int kAct>
__device__ __forceinline__ uint32_t bias_act_pack(uint32_t v0, uint32_t v1, float b0, float b1) {
  if (kAct == PSOB_ACT_TANH || kAct == PSOB_ACT_SIGMOID)
    return pack_bf16x2(act_fn<kAct>(__uint_as_float(v0) + b0), act_fn<kAct>(__uint_as_float(v1) + b1));
  unsigned long long s;
  float lo, hi;
  uint32_t d;
  asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%1, %2};\n\tmov.b64 b, {%3, %4};\n\tadd.rn.f32x2 %0, a, b;\n\t}"
      : "=l"(s) : "r"(v0), "r"(v1), "f"(b0), "f"(b1));
  if (kAct == PSOB_ACT_LEAKY) {
    unsigned long long t;
    float tl, th;
    asm("{\n\t.reg .b64 c;\n\tmov.b64 c, {%1, %1};\n\tmul.rn.f32x2 %0, %2, c;\n\t}" : "=l"(t) : "f"(0.2f), "l"(s));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(s));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(tl), "=f"(th) : "l"(t));
    lo = fmaxf(lo, tl);
    hi = fmaxf(hi, th);
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
  }
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(s));
  if (kAct == PSOB_ACT_RELU) asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// the same without a bias (all-zero bias rows): ReLU is a bare cvt.rn.relu.bf16x2.f32
template <int kAct>
__device__ __forceinline__ uint32_t act_pack(uint32_t v0, uint32_t v1) {
  uint32_t d;
  if (kAct == PSOB_ACT_RELU) { asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(__uint_as_float(v1)), "f"(__uint_as_float(v0))); return d; }
  if (kAct == PSOB_ACT_NONE) { asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(__uint_as_float(v1)), "f"(__uint_as_float(v0))); return d; }
  if (kAct == PSOB_ACT_LEAKY) {           // max(v, 0.2 v): one packed multiply, two max, one convert
    unsigned long long t;
    float tl, th;
    asm("{\n\t.reg .b64 a, c;\n\tmov.b64 a, {%1, %2};\n\tmov.b64 c, {%3, %3};\n\tmul.rn.f32x2 %0, a, c;\n\t}" : "=l"(t) : "r"(v0), "r"(v1), "f"(0.2f));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(tl), "=f"(th) : "l"(t));
    const float lo = fmaxf(__uint_as_float(v0), tl), hi = fmaxf(__uint_as_float(v1), th);
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
  }
  return bias_act_pack<kAct>(v0, v1, 0.f, 0.f);
}

// kTmaStore: bf16 NHWC outputs whose 64-column chunks are contiguous 128-byte rows leave through a
// swizzled shared-memory tile and one TMA store per chunk (two epilogue groups alternate chunks);
// otherwise (image / score outputs, nt = 16..64) group 0 stores directly from registers.
//
// kPair: the CTA is one half of a cluster of two (tcgen05 cta_group::2).  The pair works on 256 particles
// (rank r owns rows [r*128, r*128+128) of the pair tile and the matching 128 TMEM lanes); every weight
// tile is split by rows between the two CTAs and stays RESIDENT in shared memory, so the only per-step
// L2 traffic is each CTA's own 16 KB A tile (the L2 slices deliver ~42 B/clk/SM with all SMs pulling:
// streaming a 32 KB B tile per 512-cycle step had every heavy layer L2-bound at ~45 % of the MMA rate).
// The leader (rank 0) issues all MMAs; its commits arrive on the barriers of both CTAs.
//
// kDot: the chunked epilogue without the store.  Each thread owns one particle row of the accumulator, so the
// per-particle reduction of a following Cout = 1 full-map conv (Discriminator's last layer, src/utils/util_dcgan.py:117)
// is a thread-local dot product of the bf16-rounded activations with that conv's weights (kept in the unused
// staging tiles): no staging, no TMA store, no HBM round trip of the activation, and the next layer disappears.
// Partial sums are written per (job, epilogue group) in a fixed order, so the result is deterministic.
// kLean: the common chunked bf16 epilogue as a compile-time property -- no image band, per-warp stores, tile-major output, no
// role profiling.  ncu (r02_ncu_step_v22): the epilogue warps of the short-chain layers executed 288 instructions per 64-column
// chunk, 32 conversions and 8 stores among them; the rest were uniform branches over these run-time modes.
template <int kAct, bool kTmaStore, bool kPair, bool kRes = false, bool kDot = false, bool kLean = false>
__global__ void __launch_bounds__(kLean ? kMaxThreadsLean : kMaxThreads, 1)
pixel_gemm_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmY, const EngineParams p) {
  extern __shared__ uint8_t smem_raw[];
  // Hot scalar parameters in registers: the kernel parameters live in constant memory, and every asm volatile with a
  // "memory" clobber (mbarrier, TMA, tcgen05 wrappers) makes the compiler re-load whatever it needs next from there --
  // ncu showed LDCU -> short-scoreboard chains of ~400 cycles in the head of every epilogue chunk.
  const int q_n_stages = p.n_stages;
  const int q_stage_bytes = p.stage_bytes;
  const int q_res_off = p.res_off;
  const int q_nt = p.nt;
  const int q_mma_n = p.mma_n;
  const int q_epi_groups = p.epi_groups;
  const int q_stg_bufs = p.stg_bufs;
  const int q_band = kLean ? 0 : p.band;
  const int q_in_tiled = p.in_tiled;
  const int q_out_tiled = kLean ? 1 : p.out_tiled;
  const bool q_warp_store = kLean || (kTmaStore && p.warp_store != 0);
  const bool q_prof = !kLean && p.prof != nullptr;
  const bool q_half = q_warp_store && p.half_store != 0;
  const bool q_quarter = q_warp_store && p.half_store == 2;
  const int q_n_in_chunks = p.n_in_chunks;
  const int q_n_out_chunks = p.n_out_chunks;
  const uint32_t q_idesc = p.idesc;
  const int q_strided = p.strided;
  const int q_n_chains = p.n_chains;
  const int q_b_resident = p.b_resident;
  const long long q_P = p.P;
  const int q_out32_pitch = p.out32_pitch;
  const int q_out_pitch = p.out_pitch;
  const int q_vec4 = p.vec4;
  const int q_cc = p.cc;
  const long long q_n_items = p.n_items;
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - raw);
  const uint32_t bars = base + p.bar_off;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int b) { return bars + 8u * (2 * kMaxStages + b); };
  auto tempty_bar = [&](int b) { return bars + 8u * (2 * kMaxStages + 2 + b); };
  const uint32_t res_bar = bars + 8u * (2 * kMaxStages + 4);
  const uint32_t tmem_slot = bars + 8u * (2 * kMaxStages + 5);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int kThreads = (int)blockDim.x;
  const int kTmaWarp = 4 * q_epi_groups, kMmaWarp = kTmaWarp + 1;
  const uint32_t crank = kPair ? ptx::cluster_ctarank() : 0u;
  const bool leader = crank == 0u;
  // Contiguous item ranges: a CTA (or CTA pair) walks the jobs of one particle tile in order, so the input
  // pixels shared by neighbouring jobs (2.25x - 4x re-reads) are re-read from L2 moments later, not from DRAM.
  const long long n_workers = kPair ? (gridDim.x >> 1) : gridDim.x;
  const long long worker = kPair ? (blockIdx.x >> 1) : blockIdx.x;
  // Item order.  strided: worker w takes items w, w + W, w + 2W, ... so that at any moment all workers sit on
  // neighbouring jobs of the same one or two particle tiles (re-reads of shared input pixels hit L2);
  // contiguous: every worker walks its own particle tiles (no sharing between workers).
  const long long per_cta = (q_n_items + n_workers - 1) / n_workers;
  const long long item_begin = q_strided ? worker : (per_cta * worker < q_n_items ? per_cta * worker : q_n_items);
  const long long item_end = q_strided ? q_n_items : (item_begin + per_cta < q_n_items ? item_begin + per_cta : q_n_items);
  const long long item_step = q_strided ? n_workers : 1;
  const uint32_t n_chains_u = (uint32_t)q_n_chains;
  // item -> (particle tile, chain).  With the strided item order worker w sees the slots (w + k W) % n_chains only; when
  // gcd(W, n_chains) > 1 some workers would get nothing but the short chains of every tile (a border column group with
  // half the pixels, the last image band) and idle at the end.  Rotating the slot -> chain map by the tile index
  // (host-picked divisor) spreads every chain over every worker and keeps the chains of a tile on neighbouring workers.
  const uint32_t q_rot_div = (uint32_t)p.rot_div;
  const uint32_t q_nc_magic = p.nc_magic, q_nc_shift = p.nc_shift, q_rot_magic = p.rot_magic, q_rot_shift = p.rot_shift;
  auto item_chain = [&](long long item, int& ptile) -> uint32_t {
    // 32-bit: n_items < 2^31 (checked on the host); divisions by multiply + shift -- the emulated 32-bit division is ~25
    // dependent instructions, three of them per item and role sat in the epilogue's per-job prelude (ncu: 15 % of its samples)
    const uint32_t t = fast_div((uint32_t)item, q_nc_magic, q_nc_shift);
    uint32_t c = (uint32_t)item - t * n_chains_u;
    if (q_rot_div) {
      const uint32_t r = fast_div(t, q_rot_magic, q_rot_shift);
      c += r - fast_div(r, q_nc_magic, q_nc_shift) * n_chains_u;
      if (c >= n_chains_u) c -= n_chains_u;
    }
    ptile = (int)t;
    return c;
  };
  // particle-tile index of this CTA for TMA coordinates: pair tile t covers tiles 2t (leader) and 2t+1
  auto tile_idx = [&](int ptile) { return kPair ? 2 * ptile + (int)crank : ptile; };
  auto tile_row = [&](int ptile) { return tile_idx(ptile) * kTileM; };
  // TMA coordinates of the 64-column chunk starting at column `col` of particle tile `ptile`.  Tile-major
  // activations keep every [128 particles x 64 channels] operand tile in ONE contiguous 16 KB block (whole
  // DRAM pages per access); in the row-major layout the same tile is 128 rows of 128 B, one DRAM page each.
  auto in_c0 = [&](int col) { return q_in_tiled ? 0 : col; };
  auto in_c1 = [&](int ptile, int col) { return q_in_tiled ? (tile_idx(ptile) * q_n_in_chunks + (col >> 6)) * kTileM : tile_row(ptile); };
  auto out_c0 = [&](int col) { return q_out_tiled ? 0 : col; };
  auto out_c1 = [&](int ptile, int col) { return q_out_tiled ? (tile_idx(ptile) * q_n_out_chunks + (col >> 6)) * kTileM : tile_row(ptile); };

  if (threadIdx.x == 0) {
    for (int s = 0; s < q_n_stages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), (uint32_t)p.issuers);      // every issuing warp commits every step
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(tfull_bar(b), (uint32_t)p.issuers);
      ptx::mbar_init(tempty_bar(b), (kTmaStore ? 4 * q_epi_groups : 4) * (kPair ? 2 : 1));   // pair: both CTAs' epilogue warps
    }
    ptx::mbar_init(res_bar, 1);
    ptx::fence_barrier_init();
  }
  if (warp == kMmaWarp) {
    if (kPair) { ptx::tmem_alloc_pair(tmem_slot, p.tmem_cols); ptx::tmem_relinquish_pair(); }
    else { ptx::tmem_alloc(tmem_slot, p.tmem_cols); ptx::tmem_relinquish(); }
  }
  if (warp == kTmaWarp && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    if (kTmaStore) ptx::prefetch_tmap(&tmY);
  }
  // Epilogue tables live in shared memory: with ~220 KB of the SM's L1/SMEM carved out for the
  // ring, every __ldg would be an L2 round trip to the SAME few lines from all 148 SMs (measured:
  // ~4k cycles per 16-column epilogue iteration before this staging).
  uint8_t* tab = gbase + p.bar_off + kBarBytes;
  int* s_cm = reinterpret_cast<int*>(tab);
  int* s_cm32 = s_cm + 256;
  uint8_t* s_pix = reinterpret_cast<uint8_t*>(s_cm32 + 256);
  float* s_bias = reinterpret_cast<float*>(tab + (p.bias_off >= 0 ? p.bias_off : 0));
  const bool bias_in_smem = p.bias_off >= 0;
  const uint32_t s_bias_addr = base + p.bar_off + kBarBytes + (uint32_t)(p.bias_off >= 0 ? p.bias_off : 0);
  // Step / job descriptors too: a __ldg per step is an L2 round trip (~700-1000 cycles under load) that a
  // one-step prefetch cannot hide — it was the per-step floor of every layer (profiles/r01_roles_v3c.log).
  const int4* steps = p.steps_off >= 0 ? reinterpret_cast<const int4*>(tab + p.steps_off) : p.steps;
  const int4* jobs = p.jobs_off >= 0 ? reinterpret_cast<const int4*>(tab + p.jobs_off) : p.jobs;
  const int2* chains = p.chains_off >= 0 ? reinterpret_cast<const int2*>(tab + p.chains_off) : p.chains;
  if (p.chains_off >= 0)
    for (int i = threadIdx.x; i < q_n_chains; i += kThreads) reinterpret_cast<int2*>(tab + p.chains_off)[i] = p.chains[i];
  if (p.steps_off >= 0)
    for (int i = threadIdx.x; i < p.step_stride * p.n_steps; i += kThreads)     // <= 2 ops per step: first word only (16 B)
      reinterpret_cast<int4*>(tab + p.steps_off)[i] = p.step_stride == 2 ? p.steps[i] : p.steps[2 * i];
  const int ss = p.steps_off >= 0 ? p.step_stride : 2;
  // chain c of an item.  Layers without chained jobs (chain c == job c) never touch the chain table; when the job table is too
  // large for shared memory (thousands of jobs: G256's late layers) the NEXT item's job descriptor is pulled into L1 while the
  // current item runs -- from L2 a descriptor costs ~800 cycles per item and role, exposed because the ring holds about one job.
  const bool q_chain_ident = p.chain_ident != 0;
  const bool q_jobs_global = p.jobs_off < 0;
  auto chain_of = [&](uint32_t c) { return q_chain_ident ? make_int2((int)c, 1) : chains[c]; };
  auto prefetch_next_job = [&](long long item_next) {
    if (!q_jobs_global || !q_chain_ident || item_next >= q_n_items) return;
    int pt;
    const uint32_t c = item_chain(item_next, pt);
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p.jobs + 2 * (size_t)c));
  };
  if (p.jobs_off >= 0)
    for (int i = threadIdx.x; i < 2 * p.n_jobs; i += kThreads) reinterpret_cast<int4*>(tab + p.jobs_off)[i] = p.jobs[i];
  for (int i = threadIdx.x; i < q_nt; i += kThreads) {
    s_cm[i] = p.col_map[i];
    s_cm32[i] = p.col_map32[i];
    int px = i / q_cc;
    s_pix[i] = (uint8_t)(px < 64 ? px : 64);
  }
  if (bias_in_smem)
    for (int i = threadIdx.x; i < p.n_bias_rows * q_nt; i += kThreads) s_bias[i] = p.bias[i];
  if (kDot) {
    // weights of the fused head, indexed by this layer's output column (16-byte copies; out_pitch % 8 == 0)
    const uint4* src = reinterpret_cast<const uint4*>(p.dot_w);
    uint4* dst = reinterpret_cast<uint4*>(gbase + p.stg_off);
    for (int i = threadIdx.x; i < (q_out_pitch >> 3); i += kThreads) dst[i] = __ldg(src + i);
  }
  if (p.eye && threadIdx.x < 64) {
    // identity tile, K-major SWIZZLE_128B: row r is 128 B, its 16-byte slot j lives at j ^ (r % 8); element (r, r) = 1
    const uint32_t r = threadIdx.x;
    uint8_t* erow = gbase + p.res_off + r * 128u;
#pragma unroll
    for (int j = 0; j < 8; ++j) reinterpret_cast<uint4*>(erow)[j] = make_uint4(0u, 0u, 0u, 0u);
    *reinterpret_cast<uint16_t*>(erow + (((r >> 3) ^ (r & 7u)) << 4) + ((2u * r) & 15u)) = 0x3F80u;     // bf16 1.0
    ptx::fence_proxy_async();              // generic-proxy writes -> visible to the tensor core's operand reads
  }
  ptx::tc_fence_before();
  if (kPair) ptx::cluster_sync_all(); else __syncthreads();     // pair: the peer's barriers are initialised too
  ptx::tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(gbase + p.bar_off + 8 * (2 * kMaxStages + 5));

  if (kPair && q_b_resident) {
    // each CTA loads ITS half of every weight tile (rows [crank*mma_n/2, +mma_n/2) of the tile); the leader's
    // MMAs read both halves, so nobody starts before both CTAs have seen their loads complete
    if (warp == kTmaWarp) {
      if (ptx::elect_one()) {
        const int half = q_mma_n >> 1;
        ptx::mbar_arrive_expect_tx(res_bar, (uint32_t)p.n_wrows * 64u);
        for (int r = 0; r < p.n_wrows; r += q_mma_n)
          ptx::tma_load_2d(base + q_res_off + (uint32_t)r * 64u, &tmB, 0, r + (int)crank * half, res_bar);
      }
      __syncwarp();
    }
    ptx::mbar_wait(res_bar, 0);
    ptx::cluster_sync_all();
  }

  if (warp == kTmaWarp) {
    // ------------------------------------------------------------------ TMA producer
    {
      if (!kPair && q_b_resident && ptx::elect_one()) {
        ptx::mbar_arrive_expect_tx(res_bar, (uint32_t)p.n_wrows * 128u);
        for (int r = 0; r < p.n_wrows; r += q_mma_n)
          ptx::tma_load_2d(base + q_res_off + (uint32_t)r * 128u, &tmB, 0, r, res_bar);
      }
      __syncwarp();
      const uint32_t full0_leader = kPair ? ptx::map_to_cta(full_bar(0), 0u) : 0u;
      int stage = 0;
      uint32_t phase = 0;
      const bool prof = q_prof;
      unsigned long long w_empty = 0;
      const long long t_start = clock64();
      for (long long item = item_begin; item < item_end; item += item_step) {
        int ptile;
        const int2 ch = chain_of(item_chain(item, ptile));
        prefetch_next_job(item + item_step);
        for (int job = ch.x; job < ch.x + ch.y; ++job) {
          const int4 j0 = jobs[2 * job];
          const int abase = job_abase(jobs[2 * job + 1]);       // step columns are relative to the job's base column
          int4 st_next = steps[ss * j0.x];
          for (int s = j0.x; s < j0.x + j0.y; ++s) {
            int4 st = st_next;
            st.x += abase;
            if (s + 1 < j0.x + j0.y) st_next = steps[ss * (s + 1)];     // descriptor fetch overlaps the wait
            mbar_wait_t(empty_bar(stage), phase ^ 1u, prof, w_empty);
            const bool res = (st.y & (kMetaResident | kMetaEye)) != 0;      // no weight tile to stream for this step
            if (ptx::elect_one()) {
              const uint32_t a_dst = base + (uint32_t)stage * q_stage_bytes;
              if (kPair) {
                // both CTAs' tiles are counted on the leader's barrier (the leader issues the pair MMA).  Streamed weights:
                // each CTA fetches ITS half of the weight tile (rows [crank * mma_n / 2, + mma_n / 2)), so a step pulls
                // 16 KB + half a tile through this SM's L2 port instead of 16 KB + the whole tile
                const uint32_t half_b = res ? 0u : (uint32_t)q_mma_n * 64u;
                if (leader) ptx::mbar_arrive_expect_tx(full_bar(stage), 2u * ((uint32_t)kATileBytes + half_b));
                ptx::tma_load_2d_pair(a_dst, &tmA, in_c0(st.x), in_c1(ptile, st.x), full0_leader + 8u * (uint32_t)stage);
                if (!res) ptx::tma_load_2d_pair(a_dst + kATileBytes, &tmB, 0, op_b_row(st.z) + (int)crank * (q_mma_n >> 1), full0_leader + 8u * (uint32_t)stage);
              } else {
                ptx::mbar_arrive_expect_tx(full_bar(stage), (uint32_t)kATileBytes + (res ? 0u : (uint32_t)q_mma_n * 128u));
                ptx::tma_load_2d(a_dst, &tmA, in_c0(st.x), in_c1(ptile, st.x), full_bar(stage));
                if (!res) ptx::tma_load_2d(a_dst + kATileBytes, &tmB, 0, op_b_row(st.z), full_bar(stage));
              }
            }
            __syncwarp();
            if (++stage == q_n_stages) { stage = 0; phase ^= 1u; }
          }
        }
      }
      if (prof && lane == 0) { p.prof[blockIdx.x * 32 + 0] = w_empty; p.prof[blockIdx.x * 32 + 1] = (unsigned long long)(clock64() - t_start); }
    }
  } else if (warp >= kMmaWarp && leader) {
    // ------------------------------------------------------------------ MMA issuer (pair: leader CTA only)
    if (!kPair && q_b_resident) ptx::mbar_wait(res_bar, 0);
    int stage = 0;
    uint32_t phase = 0;
    uint32_t it = 0;
    const bool prof = q_prof && warp == kMmaWarp;
    unsigned long long w_full = 0, w_tempty = 0;
    const long long t_start = clock64();
    bool next_acquired = false;          // the next job's accumulator buffer was already acquired by a NEXT op
    // Two issuing warps: warp w owns the ops whose (accumulator buffer + column region) parity is w, so every
    // accumulator column has ONE issuer for its whole life (deterministic accumulation order); both warps walk all
    // steps and both commit on every stage / accumulator barrier.
    const uint32_t wsel = (uint32_t)(warp - kMmaWarp);
    const bool two = p.issuers == 2;
    for (long long item = item_begin; item < item_end; item += item_step) {
      int ptile_unused;
      const int2 ch = chain_of(item_chain(item, ptile_unused));
      prefetch_next_job(item + item_step);
      for (int job = ch.x; job < ch.x + ch.y; ++job, ++it) {
        const int4 j0 = jobs[2 * job];
        const uint32_t buf = it & 1u, use = it >> 1;
        int4 st_next = steps[ss * j0.x];
        if (!next_acquired) {
          if (kPair) {                                                              // arrivals come from both CTAs
            const long long t0 = prof ? clock64() : 0;
            ptx::mbar_wait_cluster(tempty_bar(buf), (use & 1u) ^ 1u);
            if (prof) w_tempty += (unsigned long long)(clock64() - t0);
          } else {
            mbar_wait_t(tempty_bar(buf), (use & 1u) ^ 1u, prof, w_tempty);
          }
        }
        next_acquired = false;
        ptx::tc_fence_after();
        for (int s = j0.x; s < j0.x + j0.y; ++s) {
          const int4 st = st_next;
          const int n_ops = st.y & kMetaOpsMask;
          int4 st2 = make_int4(0, 0, 0, 0);
          if (n_ops > 2) st2 = steps[ss * s + 1];
          if (s + 1 < j0.x + j0.y) st_next = steps[ss * (s + 1)];
          if ((st.y & kMetaHasNext) && !next_acquired) {
            // this A tile also feeds the NEXT job of the chain: its accumulator buffer (the other one) must be drained
            const uint32_t nuse = (it + 1u) >> 1;
            const long long t0 = prof ? clock64() : 0;
            if (kPair) ptx::mbar_wait_cluster(tempty_bar(buf ^ 1u), (nuse & 1u) ^ 1u);
            else ptx::mbar_wait(tempty_bar(buf ^ 1u), (nuse & 1u) ^ 1u);
            if (prof) w_tempty += (unsigned long long)(clock64() - t0);
            next_acquired = true;
          }
          mbar_wait_t(full_bar(stage), phase, prof, w_full);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            const uint32_t a_addr = base + (uint32_t)stage * q_stage_bytes;
            const uint64_t adesc = ptx::smem_desc_sw128(a_addr);
            const bool res = (st.y & kMetaResident) != 0;
            auto b_addr_of = [&](int op) {
              return !res ? a_addr + kATileBytes
                          : (kPair ? base + q_res_off + (uint32_t)op_b_row(op) * 64u : base + q_res_off + (uint32_t)op_b_row(op) * 128u);
            };
            auto d_of = [&](int op) { return tmem_base + (buf ^ (uint32_t)op_next(op)) * (uint32_t)q_nt + (uint32_t)op_d_col(op); };
            auto mine = [&](int op) { return !two || (((buf ^ (uint32_t)op_next(op)) + (uint32_t)op_region(op)) & 1u) == wsel; };
            // instruction descriptor of an op: N = width * mma_n (bits 17..22 hold N / 8)
            auto idesc_of = [&](int op) { return (q_idesc & ~(0x3Fu << 17)) | ((uint32_t)((op_width(op) * q_mma_n) >> 3) << 17); };
            auto commit_only = [&]() { if (kPair) ptx::umma_commit_pair(empty_bar(stage)); else ptx::umma_commit(empty_bar(stage)); };
            if (n_ops == 1 && !mine(st.z)) {
              commit_only();
            } else if (n_ops == 1) {
              // 4 x (K = 16) with +32 B descriptor advances inside the 128 B swizzle row, then commit -> stage free
              const bool eye = !kPair && (st.y & kMetaEye) != 0;
              const uint64_t bdesc = ptx::smem_desc_sw128(eye ? base + q_res_off : b_addr_of(st.z));
              if (kPair) ptx::umma_bf16_k64_commit_pair(d_of(st.z), adesc, bdesc, q_idesc, (uint32_t)op_acc(st.z), empty_bar(stage));
              else ptx::umma_bf16_k64_commit(d_of(st.z), adesc, bdesc, eye ? p.idesc_eye : idesc_of(st.z), (uint32_t)op_acc(st.z), empty_bar(stage));
            } else {
              // one A tile, several weight tiles: own / next job's accumulators (resident weights only)
              const int all_ops[4] = {st.z, n_ops > 1 ? st.w : st.z, n_ops > 2 ? st2.x : st.z, n_ops > 3 ? st2.y : st.z};
              int ops[4] = {all_ops[0], all_ops[1], all_ops[2], all_ops[3]};
              int m = n_ops;
              if (two) {
                // this warp's ops, in table order, compacted to the front with register selects (no indexed local array)
                const bool k0 = mine(all_ops[0]), k1 = n_ops > 1 && mine(all_ops[1]), k2 = n_ops > 2 && mine(all_ops[2]),
                           k3 = n_ops > 3 && mine(all_ops[3]);
                m = 0;
                int c0 = st.z, c1 = st.z, c2 = st.z, c3 = st.z;
                auto push = [&](bool k, int op) {
                  if (!k) return;
                  if (m == 0) c0 = op; else if (m == 1) c1 = op; else if (m == 2) c2 = op; else c3 = op;
                  ++m;
                };
                push(k0, all_ops[0]); push(k1, all_ops[1]); push(k2, all_ops[2]); push(k3, all_ops[3]);
                ops[0] = c0; ops[1] = c1; ops[2] = c2; ops[3] = c3;
              }
              if (m == 0) {
                commit_only();
              } else {
                uint32_t dd[4], aa[4];
                uint64_t bb[4];
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                  dd[o] = d_of(ops[o]); aa[o] = (uint32_t)op_acc(ops[o]); bb[o] = ptx::smem_desc_sw128(b_addr_of(ops[o]));
                }
                if (kPair) {
                  ptx::umma_bf16_k64x4_commit_pair(dd, adesc, bb, q_idesc, aa, m, empty_bar(stage));
                } else {
                  const uint32_t ii[4] = {idesc_of(ops[0]), idesc_of(ops[1]), idesc_of(ops[2]), idesc_of(ops[3])};
                  ptx::umma_bf16_k64x4w_commit(dd, adesc, bb, ii, aa, m, empty_bar(stage));
                }
              }
            }
          }
          __syncwarp();
          if (++stage == q_n_stages) { stage = 0; phase ^= 1u; }
        }
        if (ptx::elect_one()) { if (kPair) ptx::umma_commit_pair(tfull_bar(buf)); else ptx::umma_commit(tfull_bar(buf)); }
        __syncwarp();
      }
    }
    if (prof && lane == 0) {       // first issuing warp only
      p.prof[blockIdx.x * 32 + 2] = w_full; p.prof[blockIdx.x * 32 + 3] = w_tempty;
      p.prof[blockIdx.x * 32 + 4] = (unsigned long long)(clock64() - t_start);
    }
  } else if (warp < kTmaWarp) {
    // ------------------------------------------------------------------ epilogue
    // 8 warps = 2 groups x 4 TMEM lane quadrants (quadrant = warp % 4, the hardware's access rule)
    const uint32_t tempty0_leader = kPair ? ptx::map_to_cta(tempty_bar(0), 0u) : 0u;
    auto release_tmem = [&](uint32_t buf) {
      // the accumulator has been read into registers (tcgen05.wait::ld returned): a pure signal, no memory to publish
      if (kPair) ptx::mbar_arrive_cluster_relaxed(tempty0_leader + 8u * buf); else ptx::mbar_arrive_relaxed(tempty_bar(buf));
    };
    const int q = warp & 3;
    const int grp = warp >> 2;
    const int row = q * 32 + lane;
    const bool issuer_warp = (warp & 3) == 0;          // its elected lane issues / tracks the group's TMA stores
    const uint32_t sbuf0 = base + p.stg_off + (uint32_t)grp * (uint32_t)p.stg_group_bytes;
    uint32_t sflip = 0;                       // staging tile in use (double-buffered when stg_bufs == 2)
    const uint32_t sw = (uint32_t)(row & 7);
    uint32_t it = 0;
    // chunks of a job this group handles: grp, grp + groups, ...  With THREE groups and 4 chunks per job that would leave
    // group 0 with two chunks of every job (no faster than two groups): the assignment rotates with the job counter instead,
    // chunk ch of this CTA's it-th job goes to group (ch + rot) % 3, rot = (it * n_chunks) % 3 -- every group gets 4 chunks
    // of every 3 jobs, and a group that is done with its chunk of job j moves on to job j + 1 in the other accumulator.
    uint32_t grp_chunks = 0, grp_chunks1 = 0, grp_chunks2 = 0;
    const bool q_rot3 = kLean && q_epi_groups == 3;
    if (q_rot3) {
      for (int ch = 0; ch < (q_nt >> 6); ++ch) {
        if ((ch + 0) % 3 == grp) grp_chunks |= 1u << ch;
        if ((ch + 1) % 3 == grp) grp_chunks1 |= 1u << ch;
        if ((ch + 2) % 3 == grp) grp_chunks2 |= 1u << ch;
      }
    } else {
      for (int ch = grp; ch < (q_nt >> 6); ch += q_epi_groups) grp_chunks |= 1u << ch;
    }
    const uint32_t rot_inc = (uint32_t)(q_nt >> 6) % 3u;
    uint32_t rot3 = 0;
    const bool prof = q_prof;
    constexpr bool has_res = kRes;      // residual epilogue: its own instantiation (32 more live registers)
    unsigned long long w_tfull = 0, e_ld = 0, e_free = 0, e_math = 0, e_pub = 0;
    long long tq = 0;
    const long long t_start = clock64();
    if (kTmaStore ? grp < q_epi_groups : grp == 0) {
      for (long long item = item_begin; item < item_end; item += item_step) {
        int ptile;
        const int2 chn = chain_of(item_chain(item, ptile));
        prefetch_next_job(item + item_step);
        for (int job = chn.x; job < chn.x + chn.y; ++job, ++it) {
        const int4 j0 = jobs[2 * job];
        const int4 j1 = jobs[2 * job + 1];
        const unsigned long long mask = (unsigned long long)(uint32_t)j1.x | ((unsigned long long)(uint32_t)j1.y << 32);
        const float* bias = bias_in_smem ? s_bias + j1.z * q_nt : p.bias + (size_t)j1.z * q_nt;
        const uint32_t buf = it & 1u, use = it >> 1;
        // Reconstruction epilogue of a direct-store image layer with one 16-column accumulator per job (the 2x2-block image
        // layers): this row's 16 target pixels are fetched BEFORE waiting for the accumulator, so that their DRAM latency
        // hides behind the job's MMAs instead of serialising 16 dependent load -> subtract -> fma chains behind the TMEM load.
        float tv[16];
        const bool rpre = !kTmaStore && p.recon_part != nullptr && q_nt <= 16;
        if (!kTmaStore && rpre) {
          const long long particle = (long long)tile_row(ptile) + row;
          const long long tb = particle * (long long)p.recon_pitch + j0.w;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int pix = s_pix[i], cm = i < q_nt ? s_cm32[i] : -1;
            const bool ok = particle < q_P && cm >= 0 && !(pix < 64 && !((mask >> pix) & 1ull));
            tv[i] = ok ? recon_load(p.recon_target, p.recon_bf16, tb + cm) : 0.f;
          }
        }
        mbar_wait_t(tfull_bar(buf), use & 1u, prof, w_tfull);
        ptx::tc_fence_after();
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + buf * (uint32_t)q_nt;
        float dacc0 = 0.f, dacc1 = 0.f, dacc2 = 0.f, dacc3 = 0.f;      // kDot: this row's partial dot product over the job
        float racc = 0.f;                                              // reconstruction: this row's sum of squared differences over the job
        if (kTmaStore) {
          const int n_chunks = q_nt >> 6;
          // this group's valid chunks of the job: bits 28..31 of the job word (psob_layer_create), all chunks for an image band
          const uint32_t my_chunks = !q_rot3 ? grp_chunks : (rot3 == 0 ? grp_chunks : (rot3 == 1 ? grp_chunks1 : grp_chunks2));
          if (q_rot3) { rot3 += rot_inc; if (rot3 >= 3u) rot3 -= 3u; }
          const uint32_t vchunks = (q_band ? 0xFu : (uint32_t)j1.w >> 28) & my_chunks;
          const int last = 31 - __clz((int)vchunks);      // last chunk this group will read: release TMEM right after it (-1: none)
          if (last < 0) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) release_tmem(buf);
          }
          for (int ch = q_rot3 ? 0 : grp; ch < n_chunks; ch += q_rot3 ? 1 : q_epi_groups) {
            // image band: chunk ch = padded image rows 2ch, 2ch+1 of the band = accumulator columns [60 ch, 60 ch + 60)
            const int n0 = q_band ? ch * 2 * kBandPitch : ch << 6;
            if (!((vchunks >> ch) & 1u)) continue;                                    // uniform over the group
            const int cm = q_band ? ch << 6 : s_cm[n0];
            uint32_t v[64];
            uint4 rres[8];
            if (has_res) {
              // residual chunk of this row, issued ahead of the TMEM load so that its latency hides behind it
              const long long particle = (long long)tile_row(ptile) + row;
              const int col = j0.z + cm;
              const __nv_bfloat16* rp = q_out_tiled ? p.res + ((size_t)out_c1(ptile, col) + row) * kKChunk
                                                    : p.res + particle * p.res_pitch + col;
#pragma unroll
              for (int j = 0; j < 8; ++j) rres[j] = particle < q_P ? __ldg(reinterpret_cast<const uint4*>(rp) + j) : make_uint4(0u, 0u, 0u, 0u);
            }
            __syncwarp();
            if (prof) tq = clock64();
            ptx::tmem_ld_x16(tbase + (uint32_t)n0, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
            ptx::tmem_ld_x16(tbase + (uint32_t)n0 + 16u, *reinterpret_cast<uint32_t(*)[16]>(&v[16]));
            ptx::tmem_ld_x16(tbase + (uint32_t)n0 + 32u, *reinterpret_cast<uint32_t(*)[16]>(&v[32]));
            ptx::tmem_ld_x16(tbase + (uint32_t)n0 + 48u, *reinterpret_cast<uint32_t(*)[16]>(&v[48]));
            ptx::tmem_ld_wait();
            if (prof) { const long long t = clock64(); e_ld += (unsigned long long)(t - tq); tq = t; }
            if (ch == last) {
              ptx::tc_fence_before();
              __syncwarp();
              if (lane == 0) release_tmem(buf);
            }
            if (kDot) {
              // all lanes read the same 16 bytes of weights (shared-memory broadcast); bf16 -> fp32 is a shift / mask
              const uint32_t waddr = base + (uint32_t)p.stg_off + (uint32_t)(j0.z + cm) * 2u;
              auto dot_slots = [&](auto kSmemBiasTag) {
              constexpr bool kSmemBias = decltype(kSmemBiasTag)::value;      // compile-time inside the loop: one basic block of 8 slots
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 b0, b1;
                if (kSmemBias) {
                  const uint32_t baddr = s_bias_addr + (uint32_t)(j1.z * q_nt + n0 + 8 * j) * 4u;
                  b0 = ptx::ld_shared_f4(baddr); b1 = ptx::ld_shared_f4(baddr + 16u);
                } else {
                  b0 = __ldg(reinterpret_cast<const float4*>(bias + n0 + 8 * j)); b1 = __ldg(reinterpret_cast<const float4*>(bias + n0 + 8 * j + 4));
                }
                const float4 wf = ptx::ld_shared_f4(waddr + 16u * (uint32_t)j);
                const uint32_t w[4] = {__float_as_uint(wf.x), __float_as_uint(wf.y), __float_as_uint(wf.z), __float_as_uint(wf.w)};
                const uint32_t o[4] = {bias_act_pack<kAct>(v[8 * j + 0], v[8 * j + 1], b0.x, b0.y), bias_act_pack<kAct>(v[8 * j + 2], v[8 * j + 3], b0.z, b0.w),
                                       bias_act_pack<kAct>(v[8 * j + 4], v[8 * j + 5], b1.x, b1.y), bias_act_pack<kAct>(v[8 * j + 6], v[8 * j + 7], b1.z, b1.w)};
                dacc0 = fmaf(__uint_as_float(o[0] << 16), __uint_as_float(w[0] << 16), dacc0);
                dacc1 = fmaf(__uint_as_float(o[0] & 0xFFFF0000u), __uint_as_float(w[0] & 0xFFFF0000u), dacc1);
                dacc2 = fmaf(__uint_as_float(o[1] << 16), __uint_as_float(w[1] << 16), dacc2);
                dacc3 = fmaf(__uint_as_float(o[1] & 0xFFFF0000u), __uint_as_float(w[1] & 0xFFFF0000u), dacc3);
                dacc0 = fmaf(__uint_as_float(o[2] << 16), __uint_as_float(w[2] << 16), dacc0);
                dacc1 = fmaf(__uint_as_float(o[2] & 0xFFFF0000u), __uint_as_float(w[2] & 0xFFFF0000u), dacc1);
                dacc2 = fmaf(__uint_as_float(o[3] << 16), __uint_as_float(w[3] << 16), dacc2);
                dacc3 = fmaf(__uint_as_float(o[3] & 0xFFFF0000u), __uint_as_float(w[3] & 0xFFFF0000u), dacc3);
              }
              };
              if (bias_in_smem) dot_slots(std::true_type{}); else dot_slots(std::false_type{});
              if (prof) { const long long t = clock64(); e_math += (unsigned long long)(t - tq); tq = t; }
              continue;
            }
            if (q_band == 2) {
              // tiled image band (wide images): nothing is staged; every thread finishes its particle's two image rows of the
              // tile straight from the accumulator -- fp32 store and / or (target - tanh(acc))^2
              const float bz = bias_in_smem ? s_bias[0] : __ldg(p.bias);
              const long long particle = (long long)tile_row(ptile) + row;
              const bool pv = particle < q_P;
              const int n_cols = (int)((mask >> 8) & 0xFFull);
              const bool recon = p.recon_part != nullptr;
#pragma unroll
              for (int rr = 0; rr < 2; ++rr) {
                const bool rowok = pv && ((mask >> (2 * ch + rr)) & 1ull) != 0;
                const long long pix0 = (long long)j0.w + (long long)(2 * ch + rr) * p.band_w;      // image pixel of padded column 1
                float* y32row = (p.y32 && rowok) ? p.y32 + particle * q_out32_pitch + pix0 : nullptr;
                const long long tidx = particle * (long long)p.recon_pitch + pix0;
#pragma unroll
                for (int px = 1; px <= 28; ++px) {
                  const bool ok = rowok && px <= n_cols;
                  const float f = act_fn<kAct>(__uint_as_float(v[rr * kBandPitch + px]) + bz);
                  if (ok && y32row) y32row[px - 1] = f;
                  if (ok && recon) { const float dlt = recon_load(p.recon_target, p.recon_bf16, tidx + px - 1) - f; racc = fmaf(dlt, dlt, racc); }
                }
              }
              if (prof) { const long long t = clock64(); e_math += (unsigned long long)(t - tq); tq = t; }
              continue;
            }
            const uint32_t sbuf = sbuf0 + sflip * kATileBytes;
            const uint32_t srow = sbuf + (uint32_t)row * 128u;
            // half mode: this warp's two 2 KB halves; quarter mode: two 1 KB quarters used alternately by the chunk's four quarters
            const uint32_t hbase = sbuf0 + (uint32_t)q * (q_quarter ? 2048u : 4096u);
            auto half_open = [&](int) {                                  // at most the OTHER half's store may still be reading
              if (lane == 0) ptx::bulk_wait_read1();
              __syncwarp();
            };
            auto half_publish = [&](int h) {
              ptx::fence_proxy_async();
              __syncwarp();
              if (lane == 0) {
                if (q_quarter) ptx::tma_store_2d(&tmY, hbase + (uint32_t)(h & 1) * 1024u, out_c0(j0.z + cm) + 16 * h, out_c1(ptile, j0.z + cm) + q * 32);
                else ptx::tma_store_2d(&tmY, hbase + (uint32_t)h * 2048u, out_c0(j0.z + cm) + 32 * h, out_c1(ptile, j0.z + cm) + q * 32);
                ptx::bulk_commit();
              }
              __syncwarp();
            };
            if (q_half) {
              // nothing to wait for here: every half waits for its own previous store just before it is rewritten
            } else if (q_warp_store) {
              // the warp's own slice (32 rows = 4 KB) of the staging tile, tracked by ITS lane 0's bulk groups
              if (lane == 0) { if (q_stg_bufs == 3) ptx::bulk_wait_read2(); else if (q_stg_bufs == 2) ptx::bulk_wait_read1(); else ptx::bulk_wait_read0(); }
              __syncwarp();
            } else {
              if (issuer_warp) {
                if (ptx::elect_one()) { if (q_stg_bufs == 3) ptx::bulk_wait_read2(); else if (q_stg_bufs == 2) ptx::bulk_wait_read1(); else ptx::bulk_wait_read0(); }
                __syncwarp();
              }   // the store that last used this tile has drained it
              ptx::named_bar_sync(1 + grp, 128);
            }
            if (prof) { const long long t = clock64(); e_free += (unsigned long long)(t - tq); tq = t; }
            // kHalfTag (std::true_type / false_type): the half-store mode as a COMPILE-TIME property of the slot loop.  As a
            // runtime flag it put a branch into every one of the 8 unrolled slots, the loop body fell apart into 8+ basic
            // blocks and ptxas could not interleave the slots' bias -> activation -> pack chains: the epilogue warps issued
            // one instruction per ~5 cycles (ncu: 37 % fixed-latency 'wait' stalls, 20 % issuing).
            auto emit_slot = [&](auto kHalfTag, auto kNoBiasTag, int j, const float4& b0, const float4& b1) {
              constexpr int kHalfMode = (int)decltype(kHalfTag)::value;      // 0 full tile, 1 halves, 2 quarters
              constexpr bool kHalf = kHalfMode == 1;
              constexpr bool kNoBias = decltype(kNoBiasTag)::value && !has_res;
              if (has_res) {
                const uint32_t rw[4] = {rres[j].x, rres[j].y, rres[j].z, rres[j].w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {     // bf16 -> fp32 is a 16-bit shift
                  v[8 * j + 2 * i] = __float_as_uint(__uint_as_float(v[8 * j + 2 * i]) + __uint_as_float(rw[i] << 16));
                  v[8 * j + 2 * i + 1] = __float_as_uint(__uint_as_float(v[8 * j + 2 * i + 1]) + __uint_as_float(rw[i] & 0xFFFF0000u));
                }
              }
              uint4 o;
              if (kNoBias) {
                o.x = act_pack<kAct>(v[8 * j + 0], v[8 * j + 1]); o.y = act_pack<kAct>(v[8 * j + 2], v[8 * j + 3]);
                o.z = act_pack<kAct>(v[8 * j + 4], v[8 * j + 5]); o.w = act_pack<kAct>(v[8 * j + 6], v[8 * j + 7]);
              } else {
              o.x = bias_act_pack<kAct>(v[8 * j + 0], v[8 * j + 1], b0.x, b0.y);
              o.y = bias_act_pack<kAct>(v[8 * j + 2], v[8 * j + 3], b0.z, b0.w);
              o.z = bias_act_pack<kAct>(v[8 * j + 4], v[8 * j + 5], b1.x, b1.y);
              o.w = bias_act_pack<kAct>(v[8 * j + 6], v[8 * j + 7], b1.z, b1.w);
              }
              if (kHalfMode == 2) {
                // SWIZZLE_32B quarter tile [32 rows x 32 B]: 16-byte slot jj of row r lives at jj ^ ((r / 4) % 2)
                if ((j & 1) == 0) half_open(j >> 1);
                ptx::st_shared_v4(hbase + (uint32_t)((j >> 1) & 1) * 1024u + (uint32_t)lane * 32u + ((((uint32_t)j & 1u) ^ (((uint32_t)lane >> 2) & 1u)) << 4), o);
                if ((j & 1) == 1) half_publish(j >> 1);
              } else if (kHalf) {
                // SWIZZLE_64B half tile [32 rows x 64 B]: 16-byte slot jj of row r lives at jj ^ ((r / 2) % 4)
                if ((j & 3) == 0) half_open(j >> 2);
                ptx::st_shared_v4(hbase + (uint32_t)(j >> 2) * 2048u + (uint32_t)lane * 64u + ((((uint32_t)j & 3u) ^ (((uint32_t)lane >> 1) & 3u)) << 4), o);
                if ((j & 3) == 3) half_publish(j >> 2);
              } else {
                ptx::st_shared_v4(srow + (((uint32_t)j ^ sw) << 4), o);   // SWIZZLE_128B: 16-byte slot j of row r lives at j ^ (r % 8)
              }
            };
            if (q_band) {
              // out element e = 32 * rr + px of the 128-byte row <- column rr * 30 + px; the padding columns and the
              // padding rows of the image (mask) are written as zeros
              const float bz = bias_in_smem ? s_bias[0] : __ldg(p.bias);
              const long long particle = (long long)tile_row(ptile) + row;
              float* y32row = (p.y32 && particle < q_P) ? p.y32 + particle * q_out32_pitch + j0.w + ch * 2 * 28 : nullptr;
              const bool recon = p.recon_part != nullptr && particle < q_P;
              const long long tbase_idx = particle * (long long)p.recon_pitch + j0.w + ch * 2 * 28;
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const int rr = j >> 2;
                const bool rowok = ((mask >> (2 * ch + rr)) & 1ull) != 0;
                float f[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  const int px = (8 * j + i) & 31;
                  f[i] = (px >= 1 && px <= 28 && rowok) ? act_fn<kAct>(__uint_as_float(v[(rr * kBandPitch + px) & 63]) + bz) : 0.f;
                  if (px >= 1 && px <= 28 && rowok && y32row) y32row[rr * 28 + px - 1] = f[i];
                  if (px >= 1 && px <= 28 && rowok && recon) {
                    const float dlt = recon_load(p.recon_target, p.recon_bf16, tbase_idx + rr * 28 + px - 1) - f[i];
                    racc = fmaf(dlt, dlt, racc);
                  }
                }
                uint4 o;
                o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]); o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
                ptx::st_shared_v4(srow + (((uint32_t)j ^ sw) << 4), o);
              }
            } else {
              auto run_slots = [&](auto kHalfTag) {
                if (p.no_bias && !has_res) {
                  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                  for (int j = 0; j < 8; ++j) emit_slot(kHalfTag, std::true_type{}, j, z4, z4);
                } else if (bias_in_smem) {
                  // explicit ld.shared, software-pipelined one slot ahead: through a generic pointer every slot was a
                  // dependent LD -> math -> STS chain of ~110 cycles
                  const uint32_t baddr = s_bias_addr + (uint32_t)(j1.z * q_nt + n0) * 4u;
                  float4 b0 = ptx::ld_shared_f4(baddr), b1 = ptx::ld_shared_f4(baddr + 16u);
#pragma unroll
                  for (int j = 0; j < 8; ++j) {
                    float4 c0 = b0, c1 = b1;
                    if (j < 7) { b0 = ptx::ld_shared_f4(baddr + 32u * (j + 1)); b1 = ptx::ld_shared_f4(baddr + 32u * (j + 1) + 16u); }
                    emit_slot(kHalfTag, std::false_type{}, j, c0, c1);
                  }
                } else {
#pragma unroll
                  for (int j = 0; j < 8; ++j)
                    emit_slot(kHalfTag, std::false_type{}, j, __ldg(reinterpret_cast<const float4*>(bias + n0 + 8 * j)), __ldg(reinterpret_cast<const float4*>(bias + n0 + 8 * j + 4)));
                }
              };
              if (q_quarter) run_slots(std::integral_constant<int, 2>{});
              else if (q_half) run_slots(std::integral_constant<int, 1>{});
              else run_slots(std::integral_constant<int, 0>{});
            }
            if (prof) { const long long t = clock64(); e_math += (unsigned long long)(t - tq); tq = t; }
            if (q_half) {
              if (prof) { const long long t = clock64(); e_pub += (unsigned long long)(t - tq); tq = t; }
              continue;                                  // both halves were published from inside the slot loop
            }
            ptx::fence_proxy_async();                    // generic-proxy writes -> visible to the TMA (async proxy)
            if (q_warp_store) {
              __syncwarp();
              if (lane == 0) {
                ptx::tma_store_2d(&tmY, sbuf + (uint32_t)q * 4096u, out_c0(j0.z + cm), out_c1(ptile, j0.z + cm) + q * 32);
                ptx::bulk_commit();
              }
              __syncwarp();
            } else {
              ptx::named_bar_sync(1 + grp, 128);
              if (issuer_warp) {
                if (ptx::elect_one()) {
                  ptx::tma_store_2d(&tmY, sbuf, out_c0(j0.z + cm), out_c1(ptile, j0.z + cm));   // row-major: rows >= P are clipped by the tensor map
                  ptx::bulk_commit();
                }
                __syncwarp();
              }
            }
            if (q_stg_bufs > 1) { if (++sflip == (uint32_t)q_stg_bufs) sflip = 0; }
            if (prof) { const long long t = clock64(); e_pub += (unsigned long long)(t - tq); tq = t; }
          }
          if (kDot) {
            // one coalesced 128-byte store per warp; slots without a valid chunk hold 0
            const long long particle = (long long)tile_row(ptile) + row;
            if (particle < q_P) p.dot_part[(size_t)(job * q_epi_groups + grp) * (size_t)q_P + particle] = (dacc0 + dacc1) + (dacc2 + dacc3);
          }
          if (!kDot && q_band && p.recon_part != nullptr) {
            const long long particle = (long long)tile_row(ptile) + row;
            if (particle < q_P) p.recon_part[(size_t)(job * q_epi_groups + grp) * (size_t)q_P + particle] = racc;
          }
        } else {
          const long long particle = (long long)tile_row(ptile) + row;
          const bool valid = particle < q_P;
          __nv_bfloat16* yrow = p.y ? p.y + particle * q_out_pitch + j0.z : nullptr;
          float* y32row = p.y32 ? p.y32 + particle * q_out32_pitch + j0.w : nullptr;
          const bool recon = p.recon_part != nullptr;
          const long long tbase_idx = particle * (long long)p.recon_pitch + j0.w;
          for (int c0 = 0; c0 < q_nt; c0 += 16) {
            uint32_t v[16];
            __syncwarp();                       // tcgen05.ld is .sync.aligned: reconverge after the masked stores
            ptx::tmem_ld_x16(tbase + (uint32_t)c0, v);
            ptx::tmem_ld_wait();
            if (!valid) continue;               // rows past P hold zero-filled TMA data: nothing to store
            if (q_vec4 && yrow) {
#pragma unroll
              for (int g = 0; g < 4; ++g) {     // cc == 1: 4 consecutive pixels -> one 8-byte store
                const int n = c0 + 4 * g;
                const int cm = s_cm[n];
                const uint32_t m4 = n < 64 ? (uint32_t)((mask >> n) & 0xFull) : 0u;
                if (cm < 0 || m4 == 0u) continue;
                float f[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) f[i] = act_fn<kAct>(__uint_as_float(v[4 * g + i]) + bias[n + i]);
                if (m4 == 0xFu) {
                  uint2 o;
                  o.x = pack_bf16x2(f[0], f[1]);
                  o.y = pack_bf16x2(f[2], f[3]);
                  *reinterpret_cast<uint2*>(yrow + cm) = o;
                } else {
#pragma unroll
                  for (int i = 0; i < 4; ++i) if ((m4 >> i) & 1u) yrow[cm + i] = __float2bfloat16_rn(f[i]);
                }
                if (y32row) {
#pragma unroll
                  for (int i = 0; i < 4; ++i) { const int c32 = s_cm32[n + i]; if (c32 >= 0 && ((m4 >> i) & 1u)) y32row[c32] = f[i]; }
                }
                if (recon) {
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const int c32 = s_cm32[n + i];
                    if (c32 >= 0 && ((m4 >> i) & 1u)) { const float dlt = recon_load(p.recon_target, p.recon_bf16, tbase_idx + c32) - f[i]; racc = fmaf(dlt, dlt, racc); }
                  }
                }
              }
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int n = c0 + i;
                const int pix = s_pix[n];
                if (pix < 64 && !((mask >> pix) & 1ull)) continue;
                const float f = act_fn<kAct>(__uint_as_float(v[i]) + bias[n]);
                if (yrow) { const int cm = s_cm[n]; if (cm >= 0) yrow[cm] = __float2bfloat16_rn(f); }
                if (y32row) { const int cm = s_cm32[n]; if (cm >= 0) y32row[cm] = f; }
                if (recon) {
                  const int cm = s_cm32[n];
                  if (cm >= 0) { const float dlt = (rpre ? tv[i] : recon_load(p.recon_target, p.recon_bf16, tbase_idx + cm)) - f; racc = fmaf(dlt, dlt, racc); }
                }
              }
            }
          }
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) release_tmem(buf);
          if (recon && valid) p.recon_part[(size_t)job * (size_t)q_P + particle] = racc;
        }
        }   // jobs of the chain
      }
      if (kTmaStore && q_warp_store) { if (lane == 0) ptx::bulk_wait_all0(); }
      else if (kTmaStore && issuer_warp && ptx::elect_one()) ptx::bulk_wait_all0();      // staging tiles must outlive their TMA stores
      if (prof && warp < 4 && lane == 0) {        // per lane-quadrant warp of group 0: {tmem load, staging/barrier wait, math, publish}
        unsigned long long* q4 = p.prof + blockIdx.x * 32 + 16 + 4 * warp;
        q4[0] = e_ld; q4[1] = e_free; q4[2] = e_math; q4[3] = e_pub;
      }
      if (prof && warp == 0 && lane == 0) {
        p.prof[blockIdx.x * 32 + 5] = w_tfull; p.prof[blockIdx.x * 32 + 6] = (unsigned long long)(clock64() - t_start);
        p.prof[blockIdx.x * 32 + 7] = (unsigned long long)it;
        p.prof[blockIdx.x * 32 + 8] = e_ld; p.prof[blockIdx.x * 32 + 9] = e_free;
        p.prof[blockIdx.x * 32 + 10] = e_math; p.prof[blockIdx.x * 32 + 11] = e_pub;
      }
    }
  }

  ptx::tc_fence_before();
  if (kPair) ptx::cluster_sync_all(); else __syncthreads();     // pair: the peer's TMEM / barriers are in use until both are done
  if (warp == kMmaWarp) {
    ptx::tc_fence_after();
    if (kPair) ptx::tmem_dealloc_pair(tmem_base, p.tmem_cols); else ptx::tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

typedef void (*UmmaKernelFn)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const EngineParams);
static UmmaKernelFn umma_kernel_for(int act, bool tma_store, bool pair, bool res, bool dot = false, bool lean = false) {
  if (lean && tma_store && !res) {      // the hot hidden layers: chunked epilogue, per-warp stores, tile-major output (kLean)
#define PSOB_L(A, D) if (act == A && dot == D) return pair ? pixel_gemm_umma_kernel<A, true, true, false, D, true> : pixel_gemm_umma_kernel<A, true, false, false, D, true>;
    PSOB_L(PSOB_ACT_RELU, false) PSOB_L(PSOB_ACT_LEAKY, false) PSOB_L(PSOB_ACT_NONE, false)
    PSOB_L(PSOB_ACT_RELU, true) PSOB_L(PSOB_ACT_LEAKY, true)
#undef PSOB_L
  }
  if (dot) {            // fused Cout = 1 head: chunked epilogue layers with LeakyReLU / ReLU (discriminator / encoder-style stacks)
    if (res || !tma_store) return nullptr;
    if (act == PSOB_ACT_LEAKY) return pair ? pixel_gemm_umma_kernel<PSOB_ACT_LEAKY, true, true, false, true> : pixel_gemm_umma_kernel<PSOB_ACT_LEAKY, true, false, false, true>;
    if (act == PSOB_ACT_RELU) return pair ? pixel_gemm_umma_kernel<PSOB_ACT_RELU, true, true, false, true> : pixel_gemm_umma_kernel<PSOB_ACT_RELU, true, false, false, true>;
    return nullptr;
  }
  if (res) {            // residual add: solo CTAs, TMA-store epilogue, the bottleneck's ReLU (or none)
    if (pair || !tma_store) return nullptr;
    if (act == PSOB_ACT_RELU) return pixel_gemm_umma_kernel<PSOB_ACT_RELU, true, false, true>;
    if (act == PSOB_ACT_NONE) return pixel_gemm_umma_kernel<PSOB_ACT_NONE, true, false, true>;
    return nullptr;
  }
  if (pair) {           // pair mode: hidden conv layers (TMA-store epilogue, ReLU / LeakyReLU / none) and the tanh image band
    switch (act) {
      case PSOB_ACT_NONE: return pixel_gemm_umma_kernel<PSOB_ACT_NONE, true, true>;
      case PSOB_ACT_RELU: return pixel_gemm_umma_kernel<PSOB_ACT_RELU, true, true>;
      case PSOB_ACT_LEAKY: return pixel_gemm_umma_kernel<PSOB_ACT_LEAKY, true, true>;
      case PSOB_ACT_TANH: return pixel_gemm_umma_kernel<PSOB_ACT_TANH, true, true>;
    }
    return nullptr;
  }
  switch (act) {
#define PSOB_K(A) case A: return tma_store ? pixel_gemm_umma_kernel<A, true, false> : pixel_gemm_umma_kernel<A, false, false>;
    PSOB_K(PSOB_ACT_NONE) PSOB_K(PSOB_ACT_RELU) PSOB_K(PSOB_ACT_LEAKY) PSOB_K(PSOB_ACT_TANH) PSOB_K(PSOB_ACT_SIGMOID)
#undef PSOB_K
  }
  return nullptr;
}

// ------------------------------------------------------------------------------------------
// SIMT interpreter of the same tables (bring-up / tests).  grid = (n_jobs, n_ptiles).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
pixel_gemm_simt_kernel(const __nv_bfloat16* __restrict__ x, long long x_pitch, const __nv_bfloat16* __restrict__ w, const EngineParams p) {
  const int job = blockIdx.x;
  const long long ptile = blockIdx.y;
  const int4 j0 = p.jobs[2 * job];
  const int4 j1 = p.jobs[2 * job + 1];
  const unsigned long long mask = (unsigned long long)(uint32_t)j1.x | ((unsigned long long)(uint32_t)j1.y << 32);
  const float* bias = p.bias + (size_t)j1.z * p.nt;
  for (int idx = threadIdx.x; idx < kTileM * p.nt; idx += 256) {
    const int row = idx / p.nt, n = idx - row * p.nt;
    const long long particle = ptile * kTileM + row;
    if (particle >= p.P) continue;
    const int pix = n / p.cc;
    if (pix < 64 && !((mask >> pix) & 1ull)) continue;
    if (p.band_w) { const int px = n - pix * p.cc; if (px < 1 || px > (int)((mask >> 8) & 0xFFull)) continue; }      // tiled band: right border
    float acc = 0.f;
    // pass 0: the previous job of the chain (its NEXT ops accumulate into this job), pass 1: this job's own ops
    for (int pass = ((j1.w & 1) ? 0 : 1); pass < 2; ++pass) {
      const int4 js = pass == 0 ? p.jobs[2 * (job - 1)] : j0;
      const int abase = job_abase(pass == 0 ? p.jobs[2 * (job - 1) + 1] : j1);
      for (int s = js.x; s < js.x + js.y; ++s) {
        const int4 st = p.steps[2 * s];
        const int4 st2 = p.steps[2 * s + 1];
        const int ops[4] = {st.z, st.w, st2.x, st2.y};
        const int a_col = st.x + abase;
        const __nv_bfloat16* a = p.in_tiled ? x + (((size_t)ptile * p.n_in_chunks + (a_col >> 6)) * kTileM + row) * kKChunk
                                            : x + particle * x_pitch + a_col;
        for (int o = 0; o < (st.y & kMetaOpsMask); ++o) {
          const int op = ops[o];
          if (op_next(op) != (pass == 0)) continue;
          const int dc = op_d_col(op);
          if (op_eye(op)) {
            if (n >= dc && n < dc + 64) acc = (op_acc(op) ? acc : 0.f) + __bfloat162float(a[n - dc]);
            continue;
          }
          if (n < dc || n >= dc + p.mma_n * op_width(op)) continue;
          if (!op_acc(op)) acc = 0.f;
          const __nv_bfloat16* b = w + (size_t)(op_b_row(op) + n - dc) * kKChunk;
          for (int k = 0; k < kKChunk; ++k) acc += __bfloat162float(a[k]) * __bfloat162float(b[k]);
        }
      }
    }
    float pre = acc + bias[n];
    if (p.res && p.col_map[n] >= 0) {
      const int col = j0.z + p.col_map[n];
      pre += __bfloat162float(p.res[p.out_tiled ? (((size_t)ptile * p.n_out_chunks + (col >> 6)) * kTileM + row) * kKChunk + (col & 63)
                                                : (size_t)particle * p.res_pitch + col]);
    }
    const float f = apply_act(pre, p.act);
    if (p.y) {
      const int cm = p.col_map[n];
      if (cm >= 0) {
        const int col = j0.z + cm;
        const size_t o = p.out_tiled ? (((size_t)ptile * p.n_out_chunks + (col >> 6)) * kTileM + row) * kKChunk + (col & 63)
                                     : (size_t)particle * p.out_pitch + col;
        p.y[o] = __float2bfloat16_rn(f);
      }
    }
    if (p.y32) { const int cm = p.col_map32[n]; if (cm >= 0) p.y32[particle * p.out32_pitch + j0.w + cm] = f; }
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// bf16 matrix [rows, cols] with row pitch `pitch` elements; box = [box_rows x 64], SWIZZLE_128B
static int make_tmap(CUtensorMap* m, const void* ptr, uint64_t cols, uint64_t rows, uint64_t pitch, uint32_t box_rows,
                     uint32_t box_cols = (uint32_t)kKChunk, CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail(PSOB_E_DEVICE, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail((int)r, "cuTensorMapEncodeTiled failed with CUresult %d (cols=%llu rows=%llu pitch=%llu box_rows=%u)",
                                     (int)r, (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)pitch, box_rows);
  return 0;
}

}  // namespace psob

using namespace psob;

struct psob_layer {
  psob_layer_desc d;          // scalar fields only (host pointers are not kept)
  int4* steps = nullptr;
  int4* jobs = nullptr;
  int2* chains = nullptr;
  int* col_map = nullptr;
  int* col_map32 = nullptr;
  float* bias = nullptr;
  __nv_bfloat16* wpack = nullptr;
  CUtensorMap tmB;
  bool have_tmB = false;
  bool warp_store = false;         // per-warp 32-row TMA stores in the chunked epilogue (PSOB_WARP_STORE)
  bool lean = true;                // the lean kernel instantiation where its conditions hold (PSOB_LEAN=0: A/B)
  bool quarter = false;            // quarter stores: two groups on the staging bytes of one (8 KB each), see psob_layer_create
  int issuers = 1;                 // MMA-issuing warps
  bool issue_split = true;
  bool eye = false;                // some step is an identity op (streamed layers): 8 KB identity tile at res_off
  const void* attr_fn = nullptr;   // kernel instantiation whose dynamic-SMEM limit has been raised
  bool tma_store = false, vec4 = false, pair = false, strided = false;
  int epi_groups = 2, step_stride = 2;
  bool band = false;
  bool band_wide = false;          // tiled image band (no bf16 output; fp32 / reconstruction epilogue from registers)
  bool chain_ident = false;        // chains[c] == {c, 1} for every c
  bool no_bias = false;            // all bias entries are 0
  bool dot_head = false;           // dot-only layer: staging area = the head's weights, 2 epilogue groups
  int stg_bytes = 0;               // bytes of the staging area in the shared-memory plan
  unsigned long long* prof = nullptr;
  int bias_off = -1, steps_off = -1, jobs_off = -1, chains_off = -1, tab_bytes = 0, stg_bufs = 0;
  int n_stages = 0, stage_bytes = 0, res_off = 0, bar_off = 0, stg_off = 0, smem_bytes = 0;
  uint32_t tmem_cols = 0;
  int sm_count = 0;
  int device = -1;
  std::vector<int> chain_steps, chain_ops, chain_chunks;   // per chain: A-tile loads, MMA ops, stored 64-column chunks
  std::mutex rot_mu;
  long long rot_tiles = -1;        // tile count the cached rotation divisor was picked for
  int rot_div = 0;
};

// Rotation divisor of the slot -> chain map (EngineParams::rot_div) for `tiles` (pair) tiles on `workers` CTAs (pairs):
// the candidate with the best worst-case balance over three cost proxies (loads, MMA ops, epilogue chunks) of the
// strided item order; 0 (no rotation) unless a candidate is better by more than 1 %.
static int pick_rotation(psob_layer* l, long long tiles, int workers) {
  std::lock_guard<std::mutex> g(l->rot_mu);
  if (l->rot_tiles == tiles) return l->rot_div;
  const int C = (int)l->chain_steps.size();
  int best = 0;
  if (C > 1 && workers > 0 && tiles * C > workers && !knob("PSOB_NO_ROTATE")) {
    const std::vector<int>* cost[3] = {&l->chain_steps, &l->chain_ops, &l->chain_chunks};
    auto efficiency = [&](int div) {
      double worst = 1.0;
      for (int k = 0; k < 3; ++k) {
        std::vector<double> load((size_t)workers, 0.0);
        double total = 0.0;
        for (long long i = 0; i < tiles * C; ++i) {
          const long long t = i / C;
          int c = (int)(i - t * C);
          if (div) c = (int)((c + (t / div) % C) % C);
          load[(size_t)(i % workers)] += (*cost[k])[(size_t)c];
          total += (*cost[k])[(size_t)c];
        }
        double mx = 0.0;
        for (double v : load) mx = v > mx ? v : mx;
        if (mx > 0.0 && total / workers / mx < worst) worst = total / workers / mx;
      }
      return worst;
    };
    double best_eff = efficiency(0) + 0.01;
    const int per_round = workers / C > 0 ? workers / C : 1;      // tiles a round of the strided order advances by
    for (int div : {1, per_round, per_round + 1, 2 * per_round + 1, 3, 7})
      if (div > 0) { const double e = efficiency(div); if (e > best_eff) { best_eff = e; best = div; } }
  }
  l->rot_tiles = tiles;
  l->rot_div = best;
  return best;
}

struct ReconArgs { const void* target = nullptr; int is_bf16 = 0; int n_pix = 0; float* part = nullptr; };
static int layer_run_impl(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32, const void* res,
                          int64_t res_pitch, const void* dot_w, float* dot_part, int32_t impl, psob_stream_t s,
                          const ReconArgs& rcn = ReconArgs());

extern "C" {

int psob_layer_destroy(psob_layer* l) {
  if (!l) return 0;
  cudaFree(l->steps); cudaFree(l->jobs); cudaFree(l->chains); cudaFree(l->col_map); cudaFree(l->col_map32); cudaFree(l->bias); cudaFree(l->wpack);
  delete l;
  return 0;
}

int psob_layer_create(const psob_layer_desc* d, psob_layer** out) {
  if (!d || !out) return fail(PSOB_E_ARG, "psob_layer_create: null argument");
  *out = nullptr;
  if (d->mma_n % 16 || d->mma_n < 16 || d->mma_n > 256 || d->nt % 16 || d->nt < d->mma_n || d->nt > 256)
    return fail(PSOB_E_UNSUPPORTED, "psob_layer_create: need 16 <= mma_n <= nt <= 256, multiples of 16 (mma_n=%d nt=%d)", d->mma_n, d->nt);
  if (d->n_steps <= 0 || d->n_jobs <= 0 || d->n_wrows <= 0 || d->n_bias_rows <= 0 || d->cc <= 0)
    return fail(PSOB_E_ARG, "psob_layer_create: empty program");
  if (!d->wpack || !d->steps || !d->jobs || !d->col_map || !d->col_map32 || !d->bias)
    return fail(PSOB_E_ARG, "psob_layer_create: null table");
  if (d->in_pitch % 8 || (d->out_pitch == 0 && d->out32_pitch == 0))
    return fail(PSOB_E_ARG, "psob_layer_create: in_pitch %% 8 != 0 or no output");
  if (d->in_tiled && d->in_pitch % kKChunk) return fail(PSOB_E_ARG, "psob_layer_create: tile-major input needs in_pitch %% 64 == 0");
  if (d->out_tiled && (d->out_pitch % kKChunk || d->out32_pitch)) return fail(PSOB_E_ARG, "psob_layer_create: tile-major output needs a bf16-only output with out_pitch %% 64 == 0");
  if (d->vec8 && (d->out32_pitch != 0 || d->out_pitch % 8 || d->cc % 8))
    return fail(PSOB_E_ARG, "psob_layer_create: vec8 needs bf16-only output with 16-byte aligned rows");
  // validate the tables so a bad program can never fault on the device
  if (d->n_chains <= 0 || !d->chains) return fail(PSOB_E_ARG, "psob_layer_create: no chains");
  bool has_wide = false;          // ops wider than mma_n: solo CTAs only (a pair splits every weight tile by rows between its CTAs)
  for (int s = 0; s < d->n_steps; ++s) {
    const int32_t* st = d->steps + 8 * s;
    if (st[0] < 0 || st[0] % 8 || (d->in_tiled && st[0] % kKChunk))       // the range is checked per job below (a_col = job base + st[0])
      return fail(PSOB_E_ARG, "step %d: relative a_col %d negative or misaligned (chunk alignment for a tile-major input)", s, st[0]);
    const int n_ops = st[1] & kMetaOpsMask;
    if (n_ops < 1 || n_ops > 4) return fail(PSOB_E_ARG, "step %d: %d ops", s, n_ops);
    if (((st[1] & kMetaResident) != 0) != (d->b_resident != 0)) return fail(PSOB_E_ARG, "step %d: resident flag disagrees with the layer", s);
    if (n_ops > 1 && !d->b_resident) return fail(PSOB_E_ARG, "step %d: several ops per A tile need resident weights", s);
    bool has_next = false;
    const int32_t ops[4] = {st[2], st[3], st[4], st[5]};
    const bool eye = (st[1] & kMetaEye) != 0;
    if (eye != (op_eye(ops[0]) != 0)) return fail(PSOB_E_ARG, "step %d: identity flag disagrees with its op", s);
    if (eye) {
      if (n_ops != 1 || d->b_resident || op_next(ops[0]) || op_d_col(ops[0]) + 64 > d->nt)
        return fail(PSOB_E_ARG, "step %d: an identity op is the only op of a streamed-weight step and stays inside the job's columns", s);
      continue;
    }
    for (int o = 0; o < n_ops; ++o) {
      if (op_eye(ops[o])) return fail(PSOB_E_ARG, "step %d op %d: identity ops cannot share a step", s, o);
      const int br = op_b_row(ops[o]), dc = op_d_col(ops[o]), wn = op_width(ops[o]) * d->mma_n;
      if (op_width(ops[o]) > 1 && (!d->b_resident || wn > 256 || d->solo))
        return fail(PSOB_E_ARG, "step %d op %d: a wide op (%d columns) needs resident weights, N <= 256 and no residual run", s, o, wn);
      if (op_width(ops[o]) > 1) has_wide = true;
      if (br % 8 || br + wn > d->n_wrows) return fail(PSOB_E_ARG, "step %d op %d: b_row %d out of range", s, o, br);
      if (dc + wn > d->nt) return fail(PSOB_E_ARG, "step %d op %d: d_col %d out of range", s, o, dc);
      has_next = has_next || op_next(ops[o]);
    }
    if (has_next != ((st[1] & kMetaHasNext) != 0)) return fail(PSOB_E_ARG, "step %d: NEXT flag disagrees with its ops", s);
  }
  {
    // chains partition the jobs in order; only the last job of a chain may not use NEXT ops
    int expect = 0;
    for (int c = 0; c < d->n_chains; ++c) {
      const int jb = d->chains[2 * c], jc = d->chains[2 * c + 1];
      if (jb != expect || jc <= 0) return fail(PSOB_E_ARG, "chain %d: jobs must be consecutive and non-empty", c);
      expect = jb + jc;
      for (int j = jb; j < jb + jc; ++j) {
        if (((d->jobs[8 * j + 7] & 1) != 0) != (j > jb)) return fail(PSOB_E_ARG, "job %d: chain-predecessor flag is wrong", j);
        if (j == jb + jc - 1) {
          const int sb = d->jobs[8 * j], sc = d->jobs[8 * j + 1];
          for (int s = sb; s >= 0 && s < sb + sc && s < d->n_steps; ++s)
            if (d->steps[8 * s + 1] & kMetaHasNext) return fail(PSOB_E_ARG, "job %d ends a chain but has NEXT ops", j);
        }
      }
    }
    if (expect != d->n_jobs) return fail(PSOB_E_ARG, "chains cover %d of %d jobs", expect, d->n_jobs);
  }
  long long max_cm = -1, max_cm32 = -1;
  for (int n = 0; n < d->nt; ++n) {
    if (d->col_map[n] > max_cm) max_cm = d->col_map[n];
    if (d->col_map32[n] > max_cm32) max_cm32 = d->col_map32[n];
    if (d->vec8 && d->col_map[n] >= 0 && n % 8 == 0 && d->col_map[n] % 8) return fail(PSOB_E_ARG, "col_map[%d] not 8-aligned for vec8", n);
  }
  for (int j = 0; j < d->n_jobs; ++j) {
    const int32_t* jb = d->jobs + 8 * j;
    if (jb[0] < 0 || jb[1] <= 0 || jb[0] + jb[1] > d->n_steps) return fail(PSOB_E_ARG, "job %d: step range out of bounds", j);
    {
      const long long abase = (long long)((uint32_t)jb[7] >> 1) * 8;
      if (d->in_tiled && abase % kKChunk) return fail(PSOB_E_ARG, "job %d: base column %lld not chunk-aligned for a tile-major input", j, abase);
      for (int st = jb[0]; st < jb[0] + jb[1]; ++st)
        if (abase + d->steps[8 * st] + kKChunk > d->in_pitch)
          return fail(PSOB_E_ARG, "job %d step %d: a_col %lld out of range", j, st, abase + d->steps[8 * st]);
    }
    if (jb[6] < 0 || jb[6] >= d->n_bias_rows) return fail(PSOB_E_ARG, "job %d: bias row out of range", j);
    if (d->vec8 && jb[2] % 8) return fail(PSOB_E_ARG, "job %d: out_base not 8-aligned for vec8", j);
    // stores are masked per pixel, so only enabled pixels must stay inside the row
    const unsigned long long mask = (unsigned long long)(uint32_t)jb[4] | ((unsigned long long)(uint32_t)jb[5] << 32);
    for (int n = 0; n < d->nt; ++n) {
      int pix = n / d->cc;
      if (pix < 64 && !((mask >> pix) & 1ull)) continue;
      if (d->band_width) {        // tiled image band: mask bits 8-15 = valid image columns of the tile (right border)
        const int px = n - pix * d->cc;
        if (px < 1 || px > (int)((mask >> 8) & 0xFFull)) continue;
      }
      if (d->out_pitch && d->col_map[n] >= 0) {
        long long o = (long long)jb[2] + d->col_map[n];
        if (o < 0 || o >= d->out_pitch) return fail(PSOB_E_ARG, "job %d col %d: bf16 store offset %lld outside [0,%d)", j, n, o, d->out_pitch);
      }
      if (d->out32_pitch && d->col_map32[n] >= 0) {
        long long o = (long long)jb[3] + d->col_map32[n];
        if (o < 0 || o >= d->out32_pitch) return fail(PSOB_E_ARG, "job %d col %d: fp32 store offset %lld outside [0,%d)", j, n, o, d->out32_pitch);
      }
    }
  }
  (void)max_cm; (void)max_cm32;

  psob_layer* l = new (std::nothrow) psob_layer();
  if (!l) return fail(PSOB_E_ARG, "psob_layer_create: out of host memory");
  l->d = *d;
  l->d.wpack = nullptr; l->d.steps = nullptr; l->d.jobs = nullptr; l->d.chains = nullptr; l->d.col_map = nullptr; l->d.col_map32 = nullptr; l->d.bias = nullptr;
#define PSOB_CUDA_L(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { psob_layer_destroy(l); \
    return fail((int)e__, "%s failed: %s", #call, cudaGetErrorString(e__)); } } while (0)
  PSOB_CUDA_L(cudaGetDevice(&l->device));
  PSOB_CUDA_L(cudaDeviceGetAttribute(&l->sm_count, cudaDevAttrMultiProcessorCount, l->device));
  PSOB_CUDA_L(cudaMalloc(&l->steps, sizeof(int4) * 2 * d->n_steps));
  PSOB_CUDA_L(cudaMalloc(&l->chains, sizeof(int2) * d->n_chains));
  PSOB_CUDA_L(cudaMemcpy(l->chains, d->chains, sizeof(int2) * d->n_chains, cudaMemcpyHostToDevice));
  PSOB_CUDA_L(cudaMalloc(&l->jobs, sizeof(int4) * 2 * d->n_jobs));
  PSOB_CUDA_L(cudaMalloc(&l->col_map, sizeof(int) * d->nt));
  PSOB_CUDA_L(cudaMalloc(&l->col_map32, sizeof(int) * d->nt));
  PSOB_CUDA_L(cudaMalloc(&l->bias, sizeof(float) * d->nt * d->n_bias_rows));
  PSOB_CUDA_L(cudaMalloc(&l->wpack, (size_t)d->n_wrows * kKChunk * 2));
  {
    // Two MMA-issuing warps (PSOB_ISSUERS, default 2) pay off when a job's ops cover several accumulator regions (N < 256
    // ops, NEXT ops of chained jobs).  Ops are split by column region when every op sits on a multiple of mma_n (identity
    // ops fall inside a region); layers with overlapping column windows (image band) are split by buffer only.
    const char* e = knob("PSOB_ISSUERS");
    l->issuers = e ? (atoi(e) == 2 ? 2 : 1) : 1;
    if (has_wide) l->issuers = 1;          // a wide op spans both issuers' column regions
    l->issue_split = true;
    std::vector<int32_t> st(d->steps, d->steps + (size_t)8 * d->n_steps);
    for (int s = 0; s < d->n_steps; ++s)
      for (int o = 0; o < (st[8 * s + 1] & kMetaOpsMask); ++o)
        if (!op_eye(st[8 * s + 2 + o]) && op_d_col(st[8 * s + 2 + o]) % d->mma_n) l->issue_split = false;
    if (l->issue_split)
      for (int s = 0; s < d->n_steps; ++s)
        for (int o = 0; o < (st[8 * s + 1] & kMetaOpsMask); ++o)
          st[8 * s + 2 + o] |= ((op_d_col(st[8 * s + 2 + o]) / d->mma_n) & 1) << 27;
    PSOB_CUDA_L(cudaMemcpy(l->steps, st.data(), sizeof(int4) * 2 * d->n_steps, cudaMemcpyHostToDevice));
  }
  {
    // valid 64-column chunks of every job (column map entry >= 0 and the chunk's pixel set in the job's mask) -> bits 28..31 of
    // job word 7 in the device copy; the chunked epilogue reads them instead of looking both up per job and chunk
    std::vector<int32_t> jb(d->jobs, d->jobs + (size_t)8 * d->n_jobs);
    for (int j = 0; j < d->n_jobs; ++j) {
      const unsigned long long mask = (unsigned long long)(uint32_t)jb[8 * j + 4] | ((unsigned long long)(uint32_t)jb[8 * j + 5] << 32);
      uint32_t bits = 0;
      for (int ch = 0; ch < 4 && ch * 64 < d->nt; ++ch) {
        const int pix = (ch * 64) / d->cc;
        if (d->col_map[ch * 64] >= 0 && pix < 64 && ((mask >> pix) & 1ull)) bits |= 1u << ch;
      }
      if ((uint32_t)jb[8 * j + 7] >> 28) { psob_layer_destroy(l); return fail(PSOB_E_ARG, "psob_layer_create: job %d: base column beyond 2^30", j); }
      jb[8 * j + 7] = (int32_t)((uint32_t)jb[8 * j + 7] | (bits << 28));
    }
    PSOB_CUDA_L(cudaMemcpy(l->jobs, jb.data(), sizeof(int4) * 2 * d->n_jobs, cudaMemcpyHostToDevice));
  }
  PSOB_CUDA_L(cudaMemcpy(l->col_map, d->col_map, sizeof(int) * d->nt, cudaMemcpyHostToDevice));
  PSOB_CUDA_L(cudaMemcpy(l->col_map32, d->col_map32, sizeof(int) * d->nt, cudaMemcpyHostToDevice));
  PSOB_CUDA_L(cudaMemcpy(l->bias, d->bias, sizeof(float) * d->nt * d->n_bias_rows, cudaMemcpyHostToDevice));
  PSOB_CUDA_L(cudaMemcpy(l->wpack, d->wpack, (size_t)d->n_wrows * kKChunk * 2, cudaMemcpyHostToDevice));
#undef PSOB_CUDA_L

  // epilogue flavour: TMA store when every 64-column chunk is one contiguous, 16-byte aligned 128-byte row
  l->band = d->band_pitch != 0;
  l->band_wide = l->band && d->band_width != 0;
  if (l->band_wide && (d->out_pitch != 0 || d->out32_pitch <= 0 || d->band_width % 4)) {
    psob_layer_destroy(l);
    return fail(PSOB_E_ARG, "psob_layer_create: a tiled image band has no bf16 output, an fp32 image and a width that is a multiple of 4");
  }
  if (l->band && (d->band_pitch != kBandPitch || d->nt != 256 || d->cc != kBandPitch || d->out_pitch % 8 || d->in_tiled < 0 || d->out_tiled)) {
    psob_layer_destroy(l);
    return fail(PSOB_E_ARG, "psob_layer_create: an image-band layer has nt = 256, cc = band_pitch = %d and a row-major output", kBandPitch);
  }
  l->tma_store = l->band || (d->vec8 && d->nt % 64 == 0 && d->cc % 64 == 0 && d->out32_pitch == 0);
  for (int c = 0; c < d->nt && l->tma_store && !l->band; c += 64)
    for (int i = 1; i < 64; ++i)
      if ((d->col_map[c] >= 0) != (d->col_map[c + i] >= 0) || (d->col_map[c] >= 0 && d->col_map[c + i] != d->col_map[c] + i)) l->tma_store = false;
  if (d->out_tiled) {
    bool ok = l->tma_store;
    for (int j = 0; j < d->n_jobs && ok; ++j)
      for (int c = 0; c < d->nt && ok; c += 64)
        if (d->col_map[c] >= 0 && (d->jobs[8 * j + 2] + d->col_map[c]) % kKChunk) ok = false;
    if (!ok) { psob_layer_destroy(l); return fail(PSOB_E_ARG, "psob_layer_create: tile-major output needs chunk-aligned TMA-store jobs"); }
  }
  // direct epilogue: 4 single-pixel columns -> one 8-byte bf16 store
  l->vec4 = !d->vec8 && d->cc == 1 && d->out_pitch > 0 && d->nt <= 64;
  for (int c = 0; c < d->nt && l->vec4; c += 4) {
    for (int i = 1; i < 4; ++i)
      if ((d->col_map[c] >= 0) != (d->col_map[c + i] >= 0) || (d->col_map[c] >= 0 && d->col_map[c + i] != d->col_map[c] + i)) l->vec4 = false;
    for (int j = 0; j < d->n_jobs && l->vec4 && d->col_map[c] >= 0; ++j)
      if ((d->jobs[8 * j + 2] + d->col_map[c]) % 4) l->vec4 = false;
  }

  // shared-memory plan.  Resident weights beyond what one CTA can hold next to a useful ring are split over a
  // CTA pair (cta_group::2): each CTA keeps half of every weight tile.
  const int b_tile = d->mma_n * 128;
  l->eye = false;
  for (int s = 0; s < d->n_steps; ++s) if (d->steps[8 * s + 1] & kMetaEye) l->eye = true;
  const int full_res_bytes = d->b_resident ? d->n_wrows * 128 : (l->eye ? kEyeBytes : 0);
  l->pair = d->b_resident && full_res_bytes > kMaxSoloResident;
  // The image band is bound by the single thread that issues its N = 96 instructions (~840 cycles per step against 4 x ~100
  // of issue floor and an HBM floor below that): as a CTA pair one instruction covers 256 particles, which halves the issue
  // cost per particle.  PSOB_BAND_PAIR=0 keeps the solo form.
  if (l->band && d->b_resident && d->mma_n % 32 == 0 && d->n_wrows % d->mma_n == 0 && l->sm_count >= 2) {
    const char* e = knob("PSOB_BAND_PAIR");
    if (e ? atoi(e) != 0 : true) l->pair = true;
  }
  if (l->pair && (has_wide || !l->tma_store || full_res_bytes / 2 > kMaxPairResident || d->mma_n % 32 || d->n_wrows % d->mma_n ||
                  (d->act != PSOB_ACT_NONE && d->act != PSOB_ACT_RELU && d->act != PSOB_ACT_LEAKY && !(l->band && d->act == PSOB_ACT_TANH)) || l->sm_count < 2)) {
    psob_layer_destroy(l);
    return fail(PSOB_E_UNSUPPORTED, "psob_layer_create: %d B of resident weights need the CTA-pair mode, which this layer cannot use", full_res_bytes);
  }
  // Streamed weights as a CTA pair: every step pulls A (16 KB) + the weight tile through the SM's L2 port, which delivers
  // ~60 B/clk; an N = 256 tile (32 KB per 512 tensor cycles) overruns it.  In a pair each CTA fetches half of the tile and one
  // instruction covers 256 particles.  Not with identity ops (their built-in tile is a solo construct); the residual
  // epilogue (psob_layer_run_residual) stays solo as well, which is checked at run time.  PSOB_STREAM_PAIR=0/1 overrides.
  if (!l->pair && !d->b_resident && !l->eye && l->tma_store && !l->band && d->mma_n % 32 == 0 && l->sm_count >= 2 &&
      (d->act == PSOB_ACT_NONE || d->act == PSOB_ACT_RELU || d->act == PSOB_ACT_LEAKY) && !d->solo) {
    // Only where the MMA chain of a job outweighs its epilogue: in a pair an accumulator buffer is released by BOTH CTAs'
    // epilogue warps, which costs epilogue-bound layers (few K-steps per 256 output columns) more than the halved weight
    // traffic gives back (G256's 64 -> 64 layer: 8.98 -> 9.7 ms as a pair; its 512 -> 256 layer: 7.5 -> 5.7 ms).
    long long n_ops = 0, n_chunks = 0;      // summed per JOB: jobs share step records (pack_tables), the table itself is not the work
    for (int j = 0; j < d->n_jobs; ++j)
      for (int s = d->jobs[8 * j]; s < d->jobs[8 * j] + d->jobs[8 * j + 1]; ++s) n_ops += d->steps[8 * s + 1] & kMetaOpsMask;
    for (int j = 0; j < d->n_jobs; ++j) {
      const unsigned long long mask = (unsigned long long)(uint32_t)d->jobs[8 * j + 4] | ((unsigned long long)(uint32_t)d->jobs[8 * j + 5] << 32);
      for (int n = 0; n < d->nt; n += 64) {
        const int pix = n / d->cc;
        if (d->col_map[n] >= 0 && (pix >= 64 || ((mask >> pix) & 1ull))) ++n_chunks;
      }
    }
    const long long mma_cycles = n_ops * 4 * (d->mma_n / 2 > 90 ? d->mma_n / 2 : 90);     // per instruction: N / 2 cycles, issue floor ~90
    const long long epi_cycles = n_chunks * 500;                                          // ~1000 cycles per chunk and group, 2 groups
    const char* e = knob("PSOB_STREAM_PAIR");
    l->pair = e ? atoi(e) != 0 : (d->mma_n >= 128 && mma_cycles * 100 >= epi_cycles * (knob("PSOB_PAIR_RATIO") ? atoi(knob("PSOB_PAIR_RATIO")) : 100));   // 150 before the lean epilogue; at 100 the assessor's stem pairs up: 0.70 -> 0.61 ms (profiles/r02_pair_ratio_ab.log)
  }
  l->stage_bytes = kATileBytes + (d->b_resident ? 0 : (l->pair ? b_tile / 2 : b_tile));
  const int res_bytes = l->pair ? full_res_bytes / 2 : full_res_bytes;
  // table area: col_map + col_map32 + pix (fixed 2304 B), then optional bias rows, steps, jobs
  int tab = 256 * 4 * 2 + 256;
  l->bias_off = l->steps_off = l->jobs_off = -1;
  if (d->n_bias_rows <= kMaxBiasRowsSmem) { l->bias_off = tab; tab += d->n_bias_rows * d->nt * 4; }
  tab = (tab + 15) & ~15;
  { const char* e2 = knob("PSOB_MAX_JOBS_SMEM"); const int mj = e2 ? atoi(e2) : kMaxJobsSmem;
    if (d->n_chains <= mj && d->n_chains <= kMaxJobsSmem) { l->chains_off = tab; tab += (d->n_chains * 8 + 15) & ~15; } }
  l->step_stride = 1;
  for (int s = 0; s < d->n_steps; ++s) if ((d->steps[8 * s + 1] & kMetaOpsMask) > 2) l->step_stride = 2;
  {
    const char* e1 = knob("PSOB_MAX_STEPS_SMEM");
    const char* e2 = knob("PSOB_MAX_JOBS_SMEM");
    const int max_steps = e1 ? atoi(e1) : kMaxStepsSmem, max_jobs = e2 ? atoi(e2) : kMaxJobsSmem;
    if (d->n_steps <= max_steps && d->n_steps <= kMaxStepsSmem && d->n_steps * 16 * l->step_stride <= 32 * 1024) { l->steps_off = tab; tab += d->n_steps * 16 * l->step_stride; }
    if (d->n_jobs <= max_jobs && d->n_jobs <= kMaxJobsSmem) { l->jobs_off = tab; tab += d->n_jobs * 32; }
  }
  l->tab_bytes = (tab + 127) & ~127;
  // staging tiles for the TMA-store epilogue: 2 per group if that still leaves a 3-stage ring
  l->stg_bufs = 0;
  // a CTA-pair layer with long MMA chains per item (>= ~6000 tensor cycles) keeps one epilogue group busy at most
  // half of the time: give the second group's staging tile to the A ring (3 -> 4 stages next to 128 KB of weights)
  // layers with short chains (K = 128 per 256 output columns) are bound by the epilogue's latency chain
  // (TMEM load -> bias/activation -> swizzled st.shared -> TMA store): a third group hides more of it
  long long total_steps = 0;                 // A-tile loads of one particle tile (per job, not per shared step record)
  for (int j = 0; j < d->n_jobs; ++j) total_steps += d->jobs[8 * j + 1];
  // threshold 3000 tensor cycles per job (it was 6000): with the lean epilogue one group keeps up with shorter chains, and the ring's
  // extra stage is worth more -- G256's 128 -> 64 layer (7.76 loads per job) 8.81 -> 7.72 ms, g.l2 (6.1 loads per job) 0.686 -> 0.678,
  // assessor unchanged (profiles/r02_group1_threshold_ab.log); PSOB_GROUP1_CYCLES overrides
  long long group1_cycles = 3000;
  { const char* e = knob("PSOB_GROUP1_CYCLES"); if (e && atoll(e) > 0) group1_cycles = atoll(e); }
  l->epi_groups = (l->pair && !l->band && total_steps * 2 * d->mma_n >= group1_cycles * d->n_jobs) ? 1 : 2;
  { const char* e = knob("PSOB_LEAN"); l->lean = e ? atoi(e) != 0 : true; }
  { const char* e = knob("PSOB_WARP_STORE"); l->warp_store = e ? atoi(e) != 0 : true; }   // measured -4 % on the forward pass vs group stores     // tuning knob (default: strided)
  // three groups: lean kernels only (14 warps), i.e. no band, tile-major output, never a residual run (solo hint) or a fused head
  const bool can3 = l->lean && l->warp_store && l->tma_store && !l->band && d->out_tiled && !d->solo && !d->dot_head && l->issuers == 1;
  // ... chosen for layers with at most 4 A-tile loads per job (the MMA chain of a job is shorter than two chunks of epilogue)
  // when the third group's staging tiles do not cost the ring a stage it needs (g.l2, 128 KB resident: 0.69 -> 0.80 ms with 3).
  // Measured (profiles/r02_epi_groups3_ab.log): d.c1 0.428 -> 0.382 ms; g.l1 0.170 -> 0.171 (bound by the mixed TMA load + store
  // traffic of its streamed weights, profiles/r01_tma_bench.log).  PSOB_EPI3=0 keeps two groups.
  {
    const int room = kSmemLimit - 1024 - kBarBytes - l->tab_bytes - res_bytes;
    auto stages_for = [&](int groups) {       // ring depth next to `groups` groups' staging tiles (same rule as below)
      const int bufs = (room - 2 * groups * kATileBytes) / l->stage_bytes >= (l->pair ? 4 : 3) ? 2 : 1;
      const int n = (room - groups * bufs * kATileBytes) / l->stage_bytes;
      return n > kMaxStages ? kMaxStages : n;
    };
    const int s2 = stages_for(2), s3 = stages_for(3);
    const char* e = knob("PSOB_EPI3");
    // resident weights only: with streamed weights the mixed TMA load + store traffic is the bound (profiles/r01_tma_bench.log)
    // and a third group changes nothing or costs 1-3 % (g.l1, G256 l2, the assessor's 1x1 convs: profiles/r02_epi3_heuristic_ab.log)
    if (can3 && l->epi_groups == 2 && d->b_resident && total_steps <= 4ll * d->n_jobs && s3 >= (s2 < 4 ? s2 : 4) && !(e && atoi(e) == 0)) l->epi_groups = 3;
  }
  { const char* e = knob("PSOB_EPI_GROUPS"); if (e && atoi(e) >= 1 && atoi(e) <= (can3 ? 3 : 2)) l->epi_groups = atoi(e); }
  if (!l->tma_store) l->epi_groups = 2;             // the direct-store epilogue uses group 0 only (thread layout of 2 groups)
  l->dot_head = d->dot_head != 0;
  if (l->dot_head && (!l->tma_store || l->band || d->out32_pitch)) {
    psob_layer_destroy(l);
    return fail(PSOB_E_UNSUPPORTED, "psob_layer_create: dot_head needs a layer with the chunked bf16 epilogue");
  }
  // Quarter stores (A/B knob, OFF by default): where the rule above gave a layer ONE epilogue group to buy a ring stage, two groups
  // fit into the same 16 KB when every warp stages 16-column quarters (2 x 1 KB per warp, SWIZZLE_32B, four 32-row TMA stores per
  // chunk) instead of 32-column halves.  Bit-identical (tests), but measured no faster: g.l2 0.670 vs 0.672 ms, G256's 128 -> 64
  // layer 7.9 vs 7.8 (profiles/r02_quarter_store_ab.log) -- with the lean epilogue one group is not on the critical path there.
  // PSOB_QUARTER=1 switches it on for those layers, =2 forces it on every layer that can take it.
  {
    const char* e = knob("PSOB_QUARTER");
    l->quarter = l->lean && l->warp_store && l->tma_store && !l->band && !l->dot_head && d->out_tiled && l->pair && l->epi_groups == 1 &&
                 !knob("PSOB_EPI_GROUPS") && e && atoi(e) >= 1;
    if (e && atoi(e) == 2 && l->lean && l->warp_store && l->tma_store && !l->band && !l->dot_head && d->out_tiled && l->epi_groups <= 2) l->quarter = true;   // A/B: force
    if (l->quarter) l->epi_groups = 2;
  }
  if (l->tma_store && !l->dot_head && !l->band_wide) {      // a tiled band stages nothing: its epilogue works from registers
    const int room = kSmemLimit - 1024 - kBarBytes - l->tab_bytes - res_bytes;
    l->stg_bufs = (room - 2 * l->epi_groups * kATileBytes) / l->stage_bytes >= (l->pair ? 4 : 3) ? 2 : 1;
    if (l->quarter) l->stg_bufs = 1;
    // a third staging tile per group where the ring stays >= 4 deep (PSOB_STG3=1; A/B knob): the epilogue warps of
    // store-bound layers wait ~10 % of their time for the store of two chunks ago to release its tile
    const char* e = knob("PSOB_STG3");
    if (l->stg_bufs == 2 && !l->band && e && atoi(e) != 0 && (room - 3 * l->epi_groups * kATileBytes) / l->stage_bytes >= 4) l->stg_bufs = 3;
  }
  int stg_bytes = l->quarter ? l->epi_groups * 8192 : l->epi_groups * l->stg_bufs * kATileBytes;
  if (l->dot_head) {
    // nothing is staged: the area holds the head's bf16 weights, and the second epilogue group is free -- it halves the
    // time an accumulator buffer stays occupied, which chained jobs wait for (they acquire the next buffer early)
    const char* e = knob("PSOB_EPI_GROUPS");
    l->epi_groups = (e && atoi(e) == 1) ? 1 : 2;
    l->stg_bufs = 1;
    stg_bytes = (d->out_pitch * 2 + 1023) & ~1023;
  }
  l->stg_bytes = stg_bytes;
  int avail = kSmemLimit - 1024 - kBarBytes - l->tab_bytes - res_bytes - stg_bytes;
  l->n_stages = avail / l->stage_bytes;
  if (l->n_stages > kMaxStages) l->n_stages = kMaxStages;
  // The tiled image band streams a 2 MB-per-particle input from DRAM with 1.5x re-reads between neighbouring jobs: a deep ring
  // puts too many loads in flight and DRAM efficiency drops -- measured on G256's 128^2 -> 256^2 layer (P = 8192, same box):
  // 8 stages 6.91 ms, 6: 5.08, 4: 4.91, 3: 5.27 (profiles/r02_g256_ring_depth_ab.log).  The 28 x 28 band (25 KB per particle,
  // at its HBM floor) is the opposite: 8 stages 0.311 ms, 4 stages 0.355 (profiles/r02_config2_ring_depth_ab.log).
  if (l->band_wide && l->n_stages > 4) l->n_stages = 4;
  if (knob("PSOB_VERBOSE"))
    fprintf(stderr, "psob_layer_create: nt %d jobs %d steps %d (%lld loads per tile) pair %d resident %d in_tiled %d out_tiled %d tma_store %d band %d dot %d | "
            "epi groups %d staging bufs %d quarter %d ring %d x %d B\n", d->nt, d->n_jobs, d->n_steps, total_steps, (int)l->pair, d->b_resident, d->in_tiled, d->out_tiled,
            (int)l->tma_store, (int)l->band, d->dot_head, l->epi_groups, l->stg_bufs, (int)l->quarter, l->n_stages, l->stage_bytes);
  { const char* e = knob("PSOB_MAX_STAGES"); if (e && atoi(e) >= 2 && l->n_stages > atoi(e)) l->n_stages = atoi(e); }
  if (l->n_stages < 2) {
    psob_layer_destroy(l);
    return fail(PSOB_E_UNSUPPORTED, "psob_layer_create: resident weights (%d B) leave no room for a 2-stage ring", res_bytes);
  }
  l->res_off = l->n_stages * l->stage_bytes;
  l->stg_off = l->res_off + res_bytes;
  l->bar_off = l->stg_off + stg_bytes;
  l->smem_bytes = l->bar_off + kBarBytes + l->tab_bytes + 1024;
  { const char* e = knob("PSOB_STRIDED"); l->strided = e ? atoi(e) != 0 : true; }
  l->chain_ident = d->n_chains == d->n_jobs && !(knob("PSOB_CHAIN_IDENT") && atoi(knob("PSOB_CHAIN_IDENT")) == 0);
  for (int c = 0; c < d->n_chains && l->chain_ident; ++c) if (d->chains[2 * c] != c || d->chains[2 * c + 1] != 1) l->chain_ident = false;
  l->no_bias = true;
  for (int i = 0; i < d->nt * d->n_bias_rows; ++i) if (d->bias[i] != 0.f) { l->no_bias = false; break; }
  for (int c = 0; c < d->n_chains; ++c) {
    int n_st = 0, n_op = 0, n_ck = 0;
    for (int j = d->chains[2 * c]; j < d->chains[2 * c] + d->chains[2 * c + 1]; ++j) {
      const int32_t* jb = d->jobs + 8 * j;
      n_st += jb[1];
      for (int st = jb[0]; st < jb[0] + jb[1]; ++st) n_op += d->steps[8 * st + 1] & kMetaOpsMask;
      const unsigned long long mask = (unsigned long long)(uint32_t)jb[4] | ((unsigned long long)(uint32_t)jb[5] << 32);
      for (int n = 0; n < d->nt; n += 64) {
        const int pix = n / d->cc;
        if (l->band || (d->col_map[n] >= 0 && (pix >= 64 || ((mask >> pix) & 1ull)))) ++n_ck;
      }
    }
    l->chain_steps.push_back(n_st); l->chain_ops.push_back(n_op); l->chain_chunks.push_back(n_ck);
  }
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * d->nt)) cols <<= 1;
  l->tmem_cols = cols;
  *out = l;
  return 0;
}

int psob_layer_profile(psob_layer* l, void* counters) {
  if (!l) return fail(PSOB_E_ARG, "psob_layer_profile: null layer");
  l->prof = reinterpret_cast<unsigned long long*>(counters);
  return 0;
}

int psob_layer_run(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32, int32_t impl, psob_stream_t s) {
  return layer_run_impl(l, x, x_pitch, P, y, y32, nullptr, 0, nullptr, nullptr, impl, s);
}

int psob_layer_run_residual(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32, const void* res,
                            int64_t res_pitch, int32_t impl, psob_stream_t s) {
  return layer_run_impl(l, x, x_pitch, P, y, y32, res, res_pitch, nullptr, nullptr, impl, s);
}

int psob_layer_dot_slots(const psob_layer* l) {
  if (!l) return fail(PSOB_E_ARG, "psob_layer_dot_slots: null layer");
  if (!l->tma_store || l->band) return fail(PSOB_E_UNSUPPORTED, "psob_layer_dot_slots: the layer has no chunked bf16 epilogue");
  return l->d.n_jobs * l->epi_groups;
}

int psob_layer_recon_slots(const psob_layer* l) {
  if (!l) return fail(PSOB_E_ARG, "psob_layer_recon_slots: null layer");
  if (!l->d.out32_pitch || (l->tma_store && !l->band)) return fail(PSOB_E_UNSUPPORTED, "psob_layer_recon_slots: not an image layer");
  return l->d.n_jobs * (l->band ? l->epi_groups : 1);
}

int psob_layer_run_recon(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32, const void* target,
                         int32_t target_is_bf16, int32_t n_pix, float* part, psob_stream_t s) {
  if (!target || !part) return fail(PSOB_E_ARG, "psob_layer_run_recon: null target or partial-sum buffer");
  ReconArgs rc;
  rc.target = target; rc.is_bf16 = target_is_bf16 ? 1 : 0; rc.n_pix = n_pix; rc.part = part;
  return layer_run_impl(l, x, x_pitch, P, y, y32, nullptr, 0, nullptr, nullptr, PSOB_IMPL_TCGEN05, s, rc);
}

int psob_layer_run_dot(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, const void* dot_w, float* part, psob_stream_t s) {
  if (!dot_w || !part) return fail(PSOB_E_ARG, "psob_layer_run_dot: null weights or partial-sum buffer");
  return layer_run_impl(l, x, x_pitch, P, nullptr, nullptr, nullptr, 0, dot_w, part, PSOB_IMPL_TCGEN05, s);
}

}  // extern "C"

static int layer_run_impl(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32, const void* res,
                          int64_t res_pitch, const void* dot_w, float* dot_part, int32_t impl, psob_stream_t s,
                          const ReconArgs& rcn) {
  if (!l || !x || P <= 0) return fail(PSOB_E_ARG, "psob_layer_run: bad arguments");
  const psob_layer_desc& d = l->d;
  const bool dot = dot_w != nullptr;
  if (l->dot_head && !dot) return fail(PSOB_E_ARG, "psob_layer_run: the layer was created with dot_head and only runs through psob_layer_run_dot");
  if (dot) {
    if (!l->tma_store || l->band || d.out32_pitch)       // the output layout (row- / tile-major) is irrelevant: nothing is stored
      return fail(PSOB_E_UNSUPPORTED, "psob_layer_run_dot: needs a layer with the chunked bf16 epilogue");
    if (d.out_pitch * 2 > l->stg_bytes)
      return fail(PSOB_E_UNSUPPORTED, "psob_layer_run_dot: %d B of head weights do not fit the layer's %d B of staging tiles", d.out_pitch * 2,
                  l->stg_bytes);
    if (!aligned16(dot_w) || (reinterpret_cast<uintptr_t>(dot_part) & 3)) return fail(PSOB_E_ARG, "psob_layer_run_dot: misaligned weights / partial sums");
  }
  if (res) {
    if (l->pair) return fail(PSOB_E_UNSUPPORTED, "psob_layer_run_residual: the layer runs as a CTA pair (create it with PSOB_STREAM_PAIR=0 for the residual epilogue)");
    if (!l->tma_store || l->band || !y) return fail(PSOB_E_UNSUPPORTED, "psob_layer_run_residual: the residual add needs the chunked bf16 (TMA-store) epilogue");
    if (!aligned16(res) || (d.out_tiled ? res_pitch != d.out_pitch : (res_pitch < d.out_pitch || res_pitch % 8)))
      return fail(PSOB_E_ARG, "psob_layer_run_residual: res must have the output's layout (pitch %lld vs out_pitch %d) and 16-byte alignment", (long long)res_pitch, d.out_pitch);
  }
  if (d.in_tiled && x_pitch != d.in_pitch) return fail(PSOB_E_ARG, "psob_layer_run: a tile-major input has x_pitch == in_pitch (%d), got %lld", d.in_pitch, (long long)x_pitch);
  if (x_pitch < d.in_pitch || x_pitch % 8 || !aligned16(x)) return fail(PSOB_E_ARG, "psob_layer_run: x_pitch %lld < in_pitch %d, or misaligned", (long long)x_pitch, d.in_pitch);
  const bool recon = rcn.part != nullptr;
  if (recon) {
    if (impl != PSOB_IMPL_TCGEN05 || dot || res) return fail(PSOB_E_UNSUPPORTED, "psob_layer_run_recon: tcgen05 path only, no residual / fused head");
    if (!d.out32_pitch || rcn.n_pix != d.out32_pitch || !rcn.target) return fail(PSOB_E_ARG, "psob_layer_run_recon: the layer must be an image layer with %d fp32 pixels per particle (target has %d)", d.out32_pitch, rcn.n_pix);
    if (l->tma_store && !l->band) return fail(PSOB_E_UNSUPPORTED, "psob_layer_run_recon: the layer has neither the image-band nor the direct fp32 epilogue");
    if (l->band && !l->band_wide && !y) return fail(PSOB_E_ARG, "psob_layer_run_recon: an image-band layer needs its bf16 output buffer");
  }
  if (!dot && !recon && ((d.out_pitch && !y && !(d.out32_pitch && y32)) || (!d.out_pitch && !y32))) return fail(PSOB_E_ARG, "psob_layer_run: missing output buffer");
  if (y && !aligned16(y)) return fail(PSOB_E_ARG, "psob_layer_run: y must be 16-byte aligned");
  EngineParams p;
  p.steps = l->steps; p.jobs = l->jobs; p.chains = l->chains; p.n_chains = d.n_chains; p.chains_off = l->chains_off; p.col_map = l->col_map; p.col_map32 = l->col_map32; p.bias = l->bias;
  p.y = d.out_pitch ? reinterpret_cast<__nv_bfloat16*>(y) : nullptr;
  p.y32 = d.out32_pitch ? y32 : nullptr;
  p.res = reinterpret_cast<const __nv_bfloat16*>(res); p.res_pitch = res_pitch;
  p.dot_w = reinterpret_cast<const uint16_t*>(dot_w); p.dot_part = dot_part;
  p.recon_target = rcn.target; p.recon_part = rcn.part; p.recon_bf16 = rcn.is_bf16; p.recon_pitch = rcn.n_pix;
  p.P = P;
  const long long n_ptiles = (P + kTileM - 1) / kTileM;
  p.n_items = n_ptiles * d.n_chains;
  p.n_jobs = d.n_jobs;
  p.mma_n = d.mma_n; p.nt = d.nt; p.cc = d.cc; p.act = d.act; p.out_pitch = d.out_pitch; p.out32_pitch = d.out32_pitch;
  p.vec8 = (d.vec8 && p.y) ? 1 : 0;
  p.b_resident = d.b_resident; p.n_wrows = d.n_wrows; p.n_bias_rows = d.n_bias_rows;
  p.n_stages = l->n_stages; p.stage_bytes = l->stage_bytes; p.res_off = l->res_off; p.bar_off = l->bar_off;
  p.in_tiled = d.in_tiled; p.out_tiled = d.out_tiled; p.n_in_chunks = d.in_pitch / kKChunk; p.n_out_chunks = d.out_pitch / kKChunk;
  p.stg_off = l->stg_off; p.vec4 = l->vec4 ? 1 : 0; p.prof = l->prof; p.strided = l->strided ? 1 : 0; p.epi_groups = l->epi_groups; p.step_stride = l->step_stride; p.band = l->band_wide ? 2 : ((l->band && p.y != nullptr) ? 1 : 0); p.band_w = l->band_wide ? d.band_width : 0;
  p.issuers = l->issuers; p.issue_split = l->issue_split ? 1 : 0;
  p.warp_store = l->warp_store ? 1 : 0;
  { const char* e = knob("PSOB_HALF_STORE"); p.half_store = (l->warp_store && l->stg_bufs == 1 && !l->band && l->tma_store && (e ? atoi(e) != 0 : true)) ? 1 : 0; }   // measured: d.c2 -2.7 %, g.l2 -1 %
  if (l->quarter) p.half_store = 2;
  p.stg_group_bytes = l->quarter ? 8192 : l->stg_bufs * kATileBytes;
  if (p.n_items >= (1ll << 31)) return fail(PSOB_E_UNSUPPORTED, "psob_layer_run: %lld work items exceed 2^31", p.n_items);
  p.bias_off = l->bias_off; p.steps_off = l->steps_off; p.jobs_off = l->jobs_off; p.stg_bufs = l->stg_bufs; p.n_steps = d.n_steps;
  p.idesc = make_idesc_bf16(kTileM, d.mma_n);
  p.idesc_eye = make_idesc_bf16(kTileM, 64);
  p.eye = l->eye ? 1 : 0;
  p.no_bias = l->no_bias ? 1 : 0;
  p.rot_div = 0;
  p.chain_ident = l->chain_ident ? 1 : 0;
  p.tmem_cols = l->tmem_cols;

  if (impl == PSOB_IMPL_SIMT) {
    if (n_ptiles > 65535) return fail(PSOB_E_UNSUPPORTED, "psob_layer_run: SIMT bring-up path is limited to 65535 particle tiles");
    dim3 grid((unsigned)d.n_jobs, (unsigned)n_ptiles);
    pixel_gemm_simt_kernel<<<grid, 256, 0, (cudaStream_t)s>>>(reinterpret_cast<const __nv_bfloat16*>(x), x_pitch, l->wpack, p);
    PSOB_LAUNCH_CHECK("pixel_gemm_simt_kernel");
    return 0;
  }
  if (impl != PSOB_IMPL_TCGEN05) return fail(PSOB_E_ARG, "psob_layer_run: unknown impl %d", impl);

  const bool tma_store = l->tma_store && (p.y != nullptr || dot || l->band_wide);      // a tiled band has the chunked epilogue without a store
  if (d.out_tiled && !tma_store) return fail(PSOB_E_ARG, "psob_layer_run: a tile-major output needs its bf16 output buffer");
  if (l->pair && !tma_store) return fail(PSOB_E_ARG, "psob_layer_run: a CTA-pair layer needs its bf16 output buffer");
  // lean instantiation: every mode the generic kernel tests per chunk is the common one (A/B: PSOB_LEAN=0)
  const bool lean = l->lean && tma_store && !res && !l->band && !l->band_wide && l->warp_store && (d.out_tiled || dot) && !l->prof;   // a fused head stores nothing
  if (l->epi_groups == 3 && !lean)
    return fail(PSOB_E_UNSUPPORTED, "psob_layer_run: a layer with three epilogue groups runs the lean kernel only (no residual, no role counters)");
  UmmaKernelFn fn = umma_kernel_for(d.act, tma_store, l->pair, res != nullptr, dot, lean);
  if (!fn) return fail(PSOB_E_ARG, "psob_layer_run: no kernel for activation %d%s", d.act,
                       res ? " with a residual (solo CTAs, ReLU / none only)" : (dot ? " with a fused head (ReLU / LeakyReLU only)" : ""));
  if (l->pair) {
    p.n_items = ((n_ptiles + 1) / 2) * d.n_chains;                // items are (pair of particle tiles, chain of jobs)
    p.idesc = make_idesc_bf16(2 * kTileM, d.mma_n);
  }
  p.rot_div = l->strided ? pick_rotation(l, p.n_items / d.n_chains, l->pair ? l->sm_count / 2 : l->sm_count) : 0;
  fast_div_magic((uint32_t)d.n_chains, &p.nc_magic, &p.nc_shift);
  fast_div_magic(p.rot_div > 0 ? (uint32_t)p.rot_div : 1u, &p.rot_magic, &p.rot_shift);
  if (!l->have_tmB) {
    int rc = make_tmap(&l->tmB, l->wpack, kKChunk, (uint64_t)d.n_wrows, kKChunk, (uint32_t)(l->pair ? d.mma_n / 2 : d.mma_n));
    if (rc) return rc;
    l->have_tmB = true;
  }
  if (l->attr_fn != (const void*)fn) {
    PSOB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    l->attr_fn = (const void*)fn;
  }
  CUtensorMap tmA, tmY;
  // tile-major buffers hold n_ptiles full tiles: [n_ptiles][pitch / 64][128][64] == a [rows, 64] matrix with 128 B rows
  int rc = d.in_tiled ? make_tmap(&tmA, x, kKChunk, (uint64_t)n_ptiles * (d.in_pitch / kKChunk) * kTileM, kKChunk, kTileM)
                      : make_tmap(&tmA, x, (uint64_t)x_pitch, (uint64_t)P, (uint64_t)x_pitch, kTileM);
  if (rc) return rc;
  if (tma_store && !dot && !l->band_wide) {
    const uint32_t box_rows = l->warp_store ? 32u : (uint32_t)kTileM;
    const uint32_t box_cols = p.half_store == 2 ? 16u : (p.half_store ? 32u : (uint32_t)kKChunk);
    const CUtensorMapSwizzle sw = p.half_store == 2 ? CU_TENSOR_MAP_SWIZZLE_32B : (p.half_store ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
    rc = d.out_tiled ? make_tmap(&tmY, y, kKChunk, (uint64_t)n_ptiles * (d.out_pitch / kKChunk) * kTileM, kKChunk, box_rows, box_cols, sw)
                     : make_tmap(&tmY, y, (uint64_t)d.out_pitch, (uint64_t)P, (uint64_t)d.out_pitch, box_rows, box_cols, sw);
    if (rc) return rc;
  } else {
    tmY = tmA;
  }
  if (l->pair) {
    const long long pairs = p.n_items < l->sm_count / 2 ? p.n_items : l->sm_count / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * pairs));
    cfg.blockDim = dim3(32 * (4 * l->epi_groups + 1 + l->issuers));
    cfg.dynamicSmemBytes = (size_t)l->smem_bytes;
    cfg.stream = (cudaStream_t)s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, tmA, l->tmB, tmY, p);
    if (e != cudaSuccess) return fail((int)e, "cluster launch of pixel_gemm_umma_kernel failed: %s", cudaGetErrorString(e));
    return 0;
  }
  long long grid = p.n_items < l->sm_count ? p.n_items : l->sm_count;
  fn<<<(unsigned)grid, 32 * (4 * l->epi_groups + 1 + l->issuers), l->smem_bytes, (cudaStream_t)s>>>(tmA, l->tmB, tmY, p);
  PSOB_LAUNCH_CHECK("pixel_gemm_umma_kernel");
  return 0;
}

// out[p] = act(sum over slots of part[slot][p] + bias): the tail of the fused head (fixed summation order)
__global__ void __launch_bounds__(256) dot_finish_kernel(const float* __restrict__ part, int n_slots, long long P, float bias, int act,
                                                         float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  float acc = 0.f;
  for (int sl = 0; sl < n_slots; ++sl) acc += part[(size_t)sl * (size_t)P + i];
  out[i] = apply_act(acc + bias, act);
}

extern "C" int psob_dot_finish(const float* part, int32_t n_slots, int64_t P, float bias, int32_t act, float* out, psob_stream_t s) {
  if (!part || !out || n_slots <= 0 || P <= 0) return fail(PSOB_E_ARG, "psob_dot_finish: bad arguments");
  dot_finish_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)s>>>(part, n_slots, P, bias, act, out);
  PSOB_LAUNCH_CHECK("dot_finish_kernel");
  return 0;
}

// ------------------------------------------------------------------------------------------
// Fused single-channel head WITHOUT any allocation by the caller or the library (psob_layer_run_dot_static): the partial
// sums of psob_layer_run_dot -- pure scratch between the layer's epilogue and psob_dot_finish -- live in a file-scope
// device array instead of a caller-allocated buffer.  Every GPU that loads libpsob.so holds PSOB_STATIC_SCRATCH_BYTES of it
// until the process exits.  No counterpart in the reference; added for callers that
// cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// In HOST code `g_dot_part_static` is only a stand-in for the symbol, not a device address: the kernels get the pointer
// that cudaGetSymbolAddress returns, looked up once per device with that device current.
// ------------------------------------------------------------------------------------------
constexpr size_t kStaticPartFloats = PSOB_STATIC_SCRATCH_BYTES / sizeof(float);
static __device__ __align__(256) float g_dot_part_static[kStaticPartFloats];

static int static_part_address(float** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static float* addr[kMaxDevices] = {};
  int dev = 0;
  PSOB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(PSOB_E_UNSUPPORTED, "psob_layer_run_dot_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> g(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_dot_part_static);      // the CURRENT device's instance of the array
    if (e != cudaSuccess || !p)
      return fail((int)(e != cudaSuccess ? e : cudaErrorInvalidSymbol), "cudaGetSymbolAddress(g_dot_part_static) failed on device %d: %s", dev,
                  cudaGetErrorString(e));
    addr[dev] = reinterpret_cast<float*>(p);
  }
  *out = addr[dev];
  return 0;
}

extern "C" int psob_layer_run_dot_static(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, const void* dot_w, float bias,
                                         int32_t act, float* out, psob_stream_t s) {
  if (!l || !x || !dot_w || !out || P <= 0) return fail(PSOB_E_ARG, "psob_layer_run_dot_static: bad arguments");
  const int slots = psob_layer_dot_slots(l);
  if (slots <= 0) return slots;                                         // not a layer with the chunked bf16 epilogue (message set)
  // the limit is checked before anything touches the device
  if ((unsigned long long)slots * (unsigned long long)P > (unsigned long long)kStaticPartFloats)
    return fail(PSOB_E_UNSUPPORTED,
                "psob_layer_run_dot_static: %d partial-sum slots x %lld particles = %llu bytes exceed the static scratch of %llu bytes "
                "(at most %lld particles per call for this layer); split the batch or use psob_layer_run_dot with your own buffer",
                slots, (long long)P, (unsigned long long)slots * (unsigned long long)P * 4ull, (unsigned long long)PSOB_STATIC_SCRATCH_BYTES,
                (long long)(kStaticPartFloats / (size_t)slots));
  float* part = nullptr;
  if (int rc = static_part_address(&part)) return rc;
  // every slot of rows < P is rewritten by the layer's epilogue on every run (slots without a valid chunk store 0), so the
  // array's previous contents -- the last call's partial sums, possibly for another P -- never reach the result
  if (int rc = layer_run_impl(l, x, x_pitch, P, nullptr, nullptr, nullptr, 0, dot_w, part, PSOB_IMPL_TCGEN05, s)) return rc;
  dot_finish_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)s>>>(part, slots, P, bias, act, out);
  PSOB_LAUNCH_CHECK("dot_finish_kernel");
  return 0;
}
