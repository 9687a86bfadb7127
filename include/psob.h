/*
 * psob.h — C ABI of libpsob.so: the B200 (sm_100a) swarm-evaluation hot path of
 * ltronchin/Gan-discovery-PSO.
 *
 * The reference is pure Python/PyTorch and has no FFI; the interface each entry point
 * replaces is therefore a *method body* of the reference, cited as file:line relative to
 * the reference root.  The Python host (`gan-discovery-pso_b200/psob/_ffi.py`) binds these
 * symbols with ctypes; INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - return 0 on success; negative = argument error (PSOB_E_*); positive = cudaError_t /
 *     CUresult of the failing call.  psob_last_error() returns a thread-local message.
 *   - no C++ exception crosses the boundary; the caller owns every buffer; nothing is
 *     allocated per call (layer handles own their device tables, created once).
 *   - all work is enqueued on the caller's stream, no internal synchronisation: every
 *     call is CUDA-graph capturable.  Handles are immutable after creation.
 *   - pointers are DEVICE pointers unless the parameter is documented as host.
 *   - one process per GPU; `index_offset` carries the rank's first global particle index.
 */
#ifndef PSOB_H
#define PSOB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PSOB_VERSION 200

#define PSOB_E_ARG        (-1)   /* bad argument (null pointer, misaligned, size) */
#define PSOB_E_UNSUPPORTED (-2)  /* shape outside what the kernels implement */
#define PSOB_E_DEVICE     (-3)   /* not an sm_100 device / driver entry point missing */

typedef void* psob_stream_t;     /* cudaStream_t */

int         psob_version(void);
const char* psob_last_error(void);
/* compute capability and SM count of the current device */
int         psob_device_info(int* cc_major, int* cc_minor, int* sm_count);

/* ------------------------------------------------------------------------------------------
 * PSO state update — replaces Particle.update_velocity + Particle.update_position driven by
 * Swarm.move_particles (src/pso/util_pso.py:39-50, 153-157), fused with the deferred
 * personal-best position copy of Swarm.update_personal_best (src/pso/util_pso.py:122-124):
 *
 *   if improved[p]: pbest_pos[p,:] = pos[p,:]
 *   vel = (w*vel + c_soci*r2[p]*(pbest_pos - pos)) + c_cogn*r1[p]*(gbest_pos - pos)
 *   vel = clamp(vel, -v_max, v_max)            (extension; off when v_max = +inf)
 *   pos = clamp(pos + vel, lo, hi)             (extension; off when lo/hi = -/+inf)
 *
 * NB the reference's naming: c_cogn scales the pull to the GLOBAL best, c_soci the pull to
 * the PERSONAL best (src/pso/util_pso.py:46,48).  r1/r2 are one scalar per particle; when
 * both pointers are NULL they are drawn in-kernel: Philox4x32-10, key = seed,
 * counter = (index_offset + p, iteration, 0, 0), r1 = (x0>>8)*2^-24, r2 = (x1>>8)*2^-24.
 * Optionally emits the new position as a zero-padded bf16 row (generator input).
 * State is fp32 [P, D] row-major, D % 4 == 0, pointers 16-byte aligned.
 * ---------------------------------------------------------------------------------------- */
/* Device-resident iteration state, so that a captured CUDA graph replays correctly:
 * psob_gbest_apply advances it, psob_update reads it. */
typedef struct {
  uint32_t iteration;   /* 1-based index of the iteration whose move is about to run */
  float    w;           /* current inertia weight */
} psob_iter_state;

typedef struct {
  float*         pos;
  float*         vel;
  float*         pbest_pos;
  const float*   gbest_pos;     /* [D] */
  const uint8_t* improved;      /* [P] or NULL */
  const float*   r1;            /* [P] or NULL (Philox) */
  const float*   r2;            /* [P] or NULL (Philox) */
  void*          z_bf16;        /* [P, z_pitch] bf16 or NULL; columns D..z_pitch-1 untouched */
  const psob_iter_state* dyn;   /* device, or NULL; when set, overrides `w` and `iteration` below */
  int64_t        P;
  int32_t        D;
  int32_t        z_pitch;
  float          w, c_cogn, c_soci;
  float          v_max, lo, hi;
  uint64_t       seed;
  uint64_t       index_offset;
  uint32_t       iteration;
  uint32_t       reserved;
} psob_update_args;

int psob_update(const psob_update_args* a, psob_stream_t s);

/* x[p, col0 .. col0 + n) = value for every row of a bf16 matrix [P, pitch]: the constant-1 columns in the K padding of the
 * latent rows that carry the first generator layer's bias (Generator.forward's first ConvTranspose2d + BatchNorm shift,
 * src/utils/util_dcgan.py:133-135, folded into the MMA) */
int psob_fill_columns_bf16(void* x, int64_t P, int64_t pitch, int32_t col0, int32_t n, float value, psob_stream_t s);
/* fp32 NCHW latent grid [P, C, HW] -> bf16 NHWC rows [P, pitch >= HW * cp] with channel pitch cp (padding channels 0), relu != 0
 * applies ReLU first: the input of VectorQuantizedVAE.decoder, whose first module is an in-place ReLU
 * (src/inverter/utils_vq_vae/util_model.py:151-164,194-203) */
int psob_pack_grid_nchw(const float* x, int64_t P, int32_t C, int32_t HW, int32_t cp, int32_t relu, void* out_bf16, int64_t pitch,
                        psob_stream_t s);
/* z[p, 0:D] = bf16(codebook[idx[p], :]) (columns D..z_pitch-1 untouched): the generator input of the VQ-snapped swarm -- particles
 * are decoded at their nearest code (VQEmbedding.forward + embedding lookup, src/inverter/utils_vq_vae/util_model.py:133-136,
 * after psob_vq_nearest); codebook fp32 [K, D] dense, idx int64 [P] */
int psob_gather_pack_latent(const float* codebook, const int64_t* idx, int64_t P, int32_t D, void* z_bf16, int32_t z_pitch,
                            psob_stream_t s);
/* fp32 [P, D] -> zero-padded bf16 [P, z_pitch] (initial generator input) */
int psob_pack_latent(const float* pos, int64_t P, int32_t D, void* z_bf16, int32_t z_pitch, psob_stream_t s);

/* ------------------------------------------------------------------------------------------
 * Personal best + local min-loc — replaces the compare in Swarm.update_personal_best
 * (src/pso/util_pso.py:122-124) and the scan of Swarm.update_global_best (:139-143):
 *   improved[p] = fit[p] < pbest_val[p]  (strict; NaN never improves);  pbest_val[p] = min
 *   *key = min over p of (order_key(pbest_val[p]) << 32 | index_offset + p)
 * order_key is the order-preserving float->uint map, so the lowest global index wins ties
 * exactly like the reference's in-order scan.  *key is reset by this call.
 * ---------------------------------------------------------------------------------------- */
int psob_pbest_minloc(const float* fit, float* pbest_val, uint8_t* improved, int64_t P,
                      uint64_t index_offset, uint64_t* key, psob_stream_t s);

/* Candidate record of this rank: rec = { key_lo, key_hi, value_bits, 0, pos[D] } (4+D words),
 * pos = the winner's personal-best position (its current pos when improved this iteration). */
int psob_gbest_candidate(const uint64_t* key, const float* pos, const float* pbest_pos,
                         const uint8_t* improved, const float* pbest_val, int64_t P,
                         uint64_t index_offset, int32_t D, float* rec, psob_stream_t s);

/* Global best — replaces the tail of Swarm.update_global_best (src/pso/util_pso.py:144-151)
 * and the inertia schedule of Swarm.optimize (:72-74, 178-179): picks the minimum key among
 * n_rec records (all-gathered over ranks) and replaces the incumbent only on a strict
 * improvement.  hist (device [hist_rows, 2], may be NULL): row dyn->iteration (row 0 without
 * dyn) gets { gbest value after the call, 1.0 if replaced else 0.0 }.  Then, when dyn is
 * given: dyn->iteration += 1 and, if the new value is > 1, dyn->w *= inertia_decay. */
int psob_gbest_apply(const float* recs, int32_t n_rec, int32_t D, float* gbest_val,
                     float* gbest_pos, float* hist, int32_t hist_rows, psob_iter_state* dyn,
                     float inertia_decay, psob_stream_t s);

/* Candidate record like psob_gbest_candidate, and *key is re-armed (reset to all ones) for the next iteration's
 * psob_fitness_tail -- the multi-rank iteration needs no memset node. */
int psob_gbest_record(uint64_t* key, const float* pos, const float* pbest_pos, const uint8_t* improved,
                      const float* pbest_val, int64_t P, uint64_t index_offset, int32_t D, float* rec, psob_stream_t s);

/* Single-rank global best: psob_gbest_candidate + psob_gbest_apply (Swarm.update_global_best, src/pso/util_pso.py:135-151,
 * and the inertia schedule :72-74,178-179) in one launch; *key is consumed and re-armed. */
int psob_gbest_local(uint64_t* key, const float* pos, const float* pbest_pos, const uint8_t* improved, const float* pbest_val,
                     int64_t P, uint64_t index_offset, int32_t D, float* gbest_val, float* gbest_pos, float* hist,
                     int32_t hist_rows, psob_iter_state* dyn, float inertia_decay, psob_stream_t s);

/* ------------------------------------------------------------------------------------------
 * Fused fitness tail — replaces, in ONE launch, the arithmetic of Discovery.fitness after the assessor network
 * (softmax pick, clamp / affine: src/pso/util_discovery.py:70-80), the inverter loss's weighted sum (:97-101) and the
 * strict-< personal-best compare of Swarm.update_personal_best (src/pso/util_pso.py:122-124) with the local min-loc of
 * Swarm.update_global_best (:139-143).  Exactly one score source:
 *   part   fp32 [n_slots][P]   partial sums of a fused single-channel head (psob_layer_run_dot): score = act(sum + bias)
 *   logits fp32 [P, n_class]   score = softmax(dim 1)[class_idx]
 *   score  fp32 [P * stride]   a ready probability
 *   fit_in fp32 [P]            a ready fitness value (generic objectives; no clamp / affine)
 * (none of them with recon_part: pure reconstruction loss, w_ass ignored).  Then
 *   fit = mode(min(score + thr, 1)) + eps                                                       (as psob_score_fitness)
 *   fit = w_ass * fit + w_rec * (sum over slots of recon_part[slot][p]) / n_pix + eps_rec       (when recon_part != NULL)
 *   improved[p] = fit < pbest_val[p]; pbest_val[p] = min; atomicMin(*key, order_key(pbest_val) << 32 | index_offset + p)
 * fit (optional) receives the fitness.  *key must be all ones on entry (psob_gbest_local / psob_gbest_record re-arm it).
 * Bit-identical to the chain psob_dot_finish -> psob_score_fitness -> psob_pbest_minloc.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
  const float* part;   int32_t n_slots; float bias; int32_t act;
  const float* score;  int32_t score_stride;
  const float* logits; int32_t n_class, class_idx;
  const float* fit_in;
  int32_t mode; float thr, eps;
  const float* recon_part; int32_t n_recon_slots, n_pix; float w_ass, w_rec, eps_rec;
  float* fit; float* pbest_val; uint8_t* improved;
  int64_t P; uint64_t index_offset; uint64_t* key;
} psob_fitness_tail_args;

int psob_fitness_tail(const psob_fitness_tail_args* a, psob_stream_t s);

/* ------------------------------------------------------------------------------------------
 * Fitness transforms — replace Discovery.fitness / fitness_pso_inverter arithmetic
 * (src/pso/util_discovery.py:70-80, 97-101) and util_data.rescale_torch (src/utils/util_data.py:129-135)
 * ---------------------------------------------------------------------------------------- */
#define PSOB_MODE_IN_TRAINING  0  /* fit = min(p+thr,1) + eps      */
#define PSOB_MODE_OUT_TRAINING 1  /* fit = 1 - min(p+thr,1) + eps  */

/* score[p*stride] is the assessor probability */
int psob_score_fitness(const float* score, int32_t stride, int64_t P, int32_t mode, float thr,
                       float eps, float* fit, psob_stream_t s);
/* logits [P, n_class] -> softmax(dim 1)[class_idx] */
int psob_softmax_select(const float* logits, int32_t n_class, int32_t class_idx, int64_t P,
                        float* prob, psob_stream_t s);
/* fit[p] = w_ass*fit_ass[p] + w_rec*mean((target[p,:]-img[p,:])^2) + eps ; fit_ass may be NULL (0);
 * target is fp32 or bf16 [P, n] */
int psob_recon_fitness(const float* img, const void* target, int32_t target_is_bf16, int64_t P,
                       int32_t n, const float* fit_ass, float w_ass, float w_rec, float eps,
                       float* fit, psob_stream_t s);
/* per-image (img-min)/(max-min), fp32 [P, n] */
int psob_rescale_minmax(const float* img, int64_t P, int32_t n, float* out, psob_stream_t s);
/* fp32 [P, 784] image -> padded bf16 [P, 31*32] (real (y,x) at (y+1,x+1), zeros elsewhere) */
int psob_pack_image(const float* img, int64_t P, void* out_bf16, psob_stream_t s);

/* ------------------------------------------------------------------------------------------
 * "Next" rows of the hot-path scope
 * ---------------------------------------------------------------------------------------- */
/* Assessor input (src/pso/util_discovery.py:45,61-66): fp32 [P, 784] image -> bf16 [P, 36 * 40] rows, zero-padded by 3
 * with a 40-element row pitch (the stem's banded-GEMM input); rescale != 0 applies the per-image min-max rescale
 * (src/utils/util_data.py:129-135) first; out32 (optional, fp32 [P, 784]) receives the rescaled image. */
int psob_pack_image_stem(const float* img, int64_t P, int32_t rescale, void* out_bf16, float* out32, psob_stream_t s);
/* MaxPool2d(k, stride, pad) on a bf16 NHWC activation [P, h * w * cp] (src/pso/util_cnn.py:88), cp % 8 == 0.
 * y_tiled != 0: y is tile-major [ceil(P/128)][y_pitch/64][128][64] (the engine's internal activation layout). */
int psob_maxpool_nhwc(const void* x, int64_t x_pitch, int64_t P, int32_t h, int32_t w, int32_t cp, int32_t k, int32_t stride,
                      int32_t pad, void* y, int64_t y_pitch, int32_t y_tiled, psob_stream_t s);
/* Swarm.mse (src/pso/util_pso.py:76-86): *sum_out (device double) = sum over unordered pairs i < j of |pos_i - pos_j|_2;
 * the caller divides by P (P - 1) / 2.  pos: fp32 [P, pitch], D <= pitch. */
int psob_pairwise_distance_sum(const float* pos, int64_t P, int32_t D, int32_t pitch, double* sum_out, psob_stream_t s);
/* vq (src/inverter/utils_vq_vae/util_function.py:4-23): idx_out[p] = argmin_k |x_p - c_k|^2 (lowest k on ties),
 * computed as (|c_k|^2 + |x_p|^2) - 2 x_p.c_k like the reference; dist_out (optional) receives that minimum.
 * x: fp32 [P, x_pitch]; codebook: fp32 [K, D] dense; idx_out: int64 [P]. */
int psob_vq_nearest(const float* x, int64_t P, int32_t D, int32_t x_pitch, const float* codebook, int32_t K,
                    int64_t* idx_out, float* dist_out, psob_stream_t s);

/* ------------------------------------------------------------------------------------------
 * Pixel-GEMM layers — replace Generator.forward / Discriminator.forward / Encoder.forward
 * (src/utils/util_dcgan.py:103-149, src/inverter/utils_ae/util_inverter.py:164-184): every
 * ConvTranspose2d / Conv2d (+ folded BatchNorm + activation) is a table of tcgen05 MMA
 * chains over 128-particle x 64-channel TMA tiles (see psob/lowering.py for the tables).
 * ---------------------------------------------------------------------------------------- */
#define PSOB_ACT_NONE 0
#define PSOB_ACT_RELU 1
#define PSOB_ACT_LEAKY 2   /* slope 0.2 */
#define PSOB_ACT_TANH 3
#define PSOB_ACT_SIGMOID 4

#define PSOB_IMPL_TCGEN05 0   /* product path */
#define PSOB_IMPL_SIMT    1   /* bring-up interpreter of the same tables (tests only) */

typedef struct {
  int32_t mma_n, nt, cc, act;
  int32_t in_pitch, out_pitch, out32_pitch;
  int32_t b_resident, vec8;
  int32_t n_steps, n_jobs, n_wrows, n_bias_rows, n_chains;
  int32_t in_tiled, out_tiled; /* tile-major bf16 activations: [ceil(P/128)][pitch/64][128][64] (each 128-particle x
                                * 64-column operand tile is one contiguous 16 KB block); 0 = row-major [P, pitch] */
  int32_t band_pitch;          /* != 0: image-band layer (psob/lowering.py lower_convt_image_band): the job's accumulator is
                                * 8 padded image rows x band_pitch columns; stored as four 64-element chunks (2 rows each) */
  int32_t dot_head;            /* != 0: the layer will only run through psob_layer_run_dot (fused single-channel head): no staging
                                * tiles are planned (the head weights take their place), two epilogue groups drain the accumulator */
  int32_t solo;                /* != 0: never run as a CTA pair (layers that are run through psob_layer_run_residual) */
  int32_t band_width;          /* != 0 (with band_pitch): TILED image band for wide images (lower_convt_image_band_tiled): a job is a
                                * tile of 8 padded image rows x 28 image columns of an image `band_width` pixels wide; no bf16 output,
                                * the epilogue stores fp32 pixels and / or accumulates the reconstruction loss from the accumulator;
                                * job mask bits 8-15 = valid image columns of the tile */
  const uint16_t* wpack;     /* HOST bf16 bits [n_wrows, 64] */
  const int32_t*  steps;     /* HOST [n_steps, 8]: a_col, meta, op0, op1, op2, op3, 0, 0.  One A tile (TMA load) feeds 1..4 MMA ops.
                              * meta: bits 0-2 op count, bit 3 weights resident, bit 4 some op targets the NEXT job of the chain,
                              *       bit 5 identity step (streamed-weight layers: no weight tile is loaded).
                              * op:   bits 0-15 weight row / 16, bits 16-23 accumulator column / 2, bit 24 NEXT job, bit 25 accumulate,
                              *       bit 26 identity: D[:, col : col + 64] (+)= A, an N = 64 MMA against a built-in identity tile
                              *       (the bottleneck's `x += identity`, src/pso/util_cnn.py:185, through the operand ring) */
  const int32_t*  jobs;      /* HOST [n_jobs, 8]: step_begin, step_count, out_base, out32_base, mask_lo, mask_hi, bias_row,
                              * has_chain_predecessor */
  const int32_t*  chains;    /* HOST [n_chains, 2]: first job, job count.  The jobs of a chain run back to back on one CTA, so an
                              * A tile loaded by job j may also accumulate into job j+1 (input rows shared by neighbouring outputs) */
  const int32_t*  col_map;   /* HOST [nt] */
  const int32_t*  col_map32; /* HOST [nt] */
  const float*    bias;      /* HOST [n_bias_rows, nt] */
} psob_layer_desc;

typedef struct psob_layer psob_layer;

int psob_layer_create(const psob_layer_desc* d, psob_layer** out);
int psob_layer_destroy(psob_layer* l);
/* x: bf16 [P, x_pitch] (x_pitch >= in_pitch, x_pitch % 8 == 0, 16-byte aligned);
 * y: bf16 [P, out_pitch] or NULL; y32: fp32 [P, out32_pitch] or NULL */
int psob_layer_run(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32,
                   int32_t impl, psob_stream_t s);
/* Same, with a residual operand: out = act(acc + bias + res) -- the bottleneck's `x += identity; relu`
 * (src/pso/util_cnn.py:185-187).  res: bf16 in the OUTPUT's layout (row-major [P, res_pitch >= out_pitch], or the
 * same tile-major buffer shape when the layer stores tile-major); only layers with the chunked bf16 epilogue. */
int psob_layer_run_residual(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32,
                            const void* res, int64_t res_pitch, int32_t impl, psob_stream_t s);

/* Fused single-channel head -- replaces the LAST TWO modules of Discriminator.forward (src/utils/util_dcgan.py:115-118:
 * the hidden Conv2d + BatchNorm + LeakyReLU block and the final Conv2d(features, 1, k = whole map) + Sigmoid) by one
 * launch plus a tiny reduction.  The layer's bf16 activation is never stored: every epilogue thread owns one particle
 * row of the accumulator and reduces it, bf16-rounded exactly as the stored activation would be, against dot_w
 * (device bf16 [out_pitch], the head conv's weight indexed by this layer's output column).  part: device fp32
 * [psob_layer_dot_slots(l)][P], one partial sum per (job, epilogue group) and particle, every slot rewritten by every run;
 * psob_dot_finish adds the slots in index order, then bias and activation: out[p] = act(sum + bias).  Deterministic.
 * Only layers with the chunked bf16 epilogue, ReLU / LeakyReLU, out_pitch * 2 bytes <= the layer's staging tiles. */
int psob_layer_dot_slots(const psob_layer* l);
int psob_layer_run_dot(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, const void* dot_w, float* part,
                       psob_stream_t s);
int psob_dot_finish(const float* part, int32_t n_slots, int64_t P, float bias, int32_t act, float* out, psob_stream_t s);

/* The fused head with NO allocation anywhere: psob_layer_run_dot + psob_dot_finish in one call, the partial sums in a static device
 * array inside libpsob.so instead of a caller-allocated `part` buffer.  This entry point has no counterpart in the reference; it was
 * added for callers that cannot allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  out[p] = act(sum of the layer's partial sums + bias), bit-identical to the two
 * existing calls.  Limits and rules:
 *   - psob_layer_dot_slots(l) * P * 4 bytes must not exceed PSOB_STATIC_SCRATCH_BYTES (8 MiB: 37 449 particles per call for the
 *     DCGAN discriminator's 56 slots); larger calls are rejected with PSOB_E_UNSUPPORTED before anything is launched;
 *   - every GPU that loads the library holds the array until the process exits;
 *   - all calls share the array and it keeps the last call's contents (every slot a call reads is rewritten by that call first, so
 *     nothing needs clearing); two calls that OVERLAP on one GPU (different streams, or threads) race on it -- serialise them;
 *   - the call itself performs no cudaMalloc / cudaFree and calls no allocating library routine; the layer handle and the caller's
 *     x / dot_w / out buffers are made beforehand. */
#define PSOB_STATIC_SCRATCH_BYTES (8u << 20)
int psob_layer_run_dot_static(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, const void* dot_w, float bias, int32_t act,
                              float* out, psob_stream_t s);

/* Reconstruction loss in the image layer's epilogue — replaces `torch.mean((img - img_rec) ** 2)` of
 * Discovery.fitness_pso_inverter (src/pso/util_discovery.py:97) without materialising the fp32 image: every epilogue thread
 * owns one particle row of the generator's last layer and accumulates (target - tanh(acc))^2 from the fp32 pixel it holds in a
 * register.  target: device fp32 or bf16 [P, n_pix] laid out like the fp32 image (n_pix == out32_pitch); part: device fp32
 * [psob_layer_recon_slots(l)][P], every slot rewritten by every run, summed in slot order by psob_fitness_tail.  y / y32 as in
 * psob_layer_run (y32 may be NULL: the image is then never written; an image-band layer still needs y).  Only image layers
 * (image band or direct fp32 epilogue). */
int psob_layer_recon_slots(const psob_layer* l);
int psob_layer_run_recon(psob_layer* l, const void* x, int64_t x_pitch, int64_t P, void* y, float* y32, const void* target,
                         int32_t target_is_bf16, int32_t n_pix, float* part, psob_stream_t s);

/* Debug: when `counters` (device, u64 [n_SMs][32]) is non-NULL, later runs of the layer record per CTA
 * { producer wait-empty, producer total, mma wait-full, mma wait-tmem-empty, mma total,
 *   epilogue wait-tmem-full, epilogue total, items, epilogue tmem-load, staging-free wait,
 *   bias/activation/st.shared, publish (fence + barrier + TMA store issue), 4 spare, then the last four for each
 *   of the warps 0..3 (TMEM lane quadrants of epilogue group 0) } in SM cycles.
 *   NULL switches it off. */
int psob_layer_profile(psob_layer* l, void* counters);

#ifdef __cplusplus
}
#endif
#endif /* PSOB_H */
