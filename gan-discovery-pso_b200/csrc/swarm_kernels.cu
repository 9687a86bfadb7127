// PSO state kernels of libpsob.so: velocity/position update with in-kernel Philox, personal
// best + min-loc, global-best record/apply, fitness transforms.  All HBM-bound elementwise or
// reduction work on fp32 SoA state [P, D]; see include/psob.h for the reference lines each
// entry point replaces.
#include "psob_common.cuh"

#include <cfloat>
#include <cmath>
#include <cstring>

namespace psob {
thread_local char g_err[512] = "";

// ------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11).  counter = (particle, iteration, 0, 0), key = seed.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t& o0, uint32_t& o1) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  o0 = c0; o1 = c1;
}

struct UpdateParams {
  float4* pos; float4* vel; float4* pbest;
  const float4* gbest;
  const uint8_t* improved;
  const float* r1; const float* r2;
  __nv_bfloat16* z;
  const psob_iter_state* dyn;
  uint32_t total4, d4;
  int32_t z_pitch;
  float w, c_cogn, c_soci, v_max, lo, hi;
  uint32_t seed_lo, seed_hi, iteration;
  uint64_t index_offset;
  int clamp_v, clamp_x;
};

__device__ __forceinline__ float move1(float x, float v, float pb, float g, float w, float c1, float c2,
                                       const UpdateParams& a, float& nv_out) {
  // (w*vel + c_soci*r2*(pbest-pos)) + c_cogn*r1*(gbest-pos) — same association as
  // src/pso/util_pso.py:44-49; explicit _rn ops keep ptxas from contracting into FMAs so the
  // result is bit-identical to the fp32 restatement in oracle/port.py.
  float inertia = __fmul_rn(w, v);
  float cog = __fmul_rn(c1, __fsub_rn(g, x));
  float soc = __fmul_rn(c2, __fsub_rn(pb, x));
  float nv = __fadd_rn(__fadd_rn(inertia, soc), cog);
  if (a.clamp_v) nv = fminf(fmaxf(nv, -a.v_max), a.v_max);
  float nx = __fadd_rn(x, nv);
  if (a.clamp_x) nx = fminf(fmaxf(nx, a.lo), a.hi);
  nv_out = nv;
  return nx;
}

template <bool kPhilox>
__global__ void __launch_bounds__(256) pso_update_kernel(const UpdateParams a) {
  uint32_t i = blockIdx.x * 256u + threadIdx.x;
  if (i >= a.total4) return;
  uint32_t p = i / a.d4;
  uint32_t j = i - p * a.d4;
  const float w = a.dyn ? a.dyn->w : a.w;
  float r1, r2;
  if (kPhilox) {
    uint64_t gidx = a.index_offset + p;
    uint32_t x0, x1;
    philox4x32_10((uint32_t)gidx, a.dyn ? a.dyn->iteration : a.iteration, 0u, 0u, a.seed_lo, a.seed_hi, x0, x1);
    r1 = (float)(x0 >> 8) * 5.9604644775390625e-08f;   // 2^-24
    r2 = (float)(x1 >> 8) * 5.9604644775390625e-08f;
  } else {
    r1 = __ldg(a.r1 + p);
    r2 = __ldg(a.r2 + p);
  }
  float4 x = __ldcs(a.pos + i);
  float4 v = __ldcs(a.vel + i);
  float4 g = __ldg(a.gbest + j);
  float4 pb;
  if (a.improved != nullptr && a.improved[p]) {
    pb = x;
    __stcs(a.pbest + i, x);
  } else {
    pb = __ldcs(a.pbest + i);
  }
  float c1 = __fmul_rn(a.c_cogn, r1);
  float c2 = __fmul_rn(a.c_soci, r2);
  float4 nx, nv;
  nx.x = move1(x.x, v.x, pb.x, g.x, w, c1, c2, a, nv.x);
  nx.y = move1(x.y, v.y, pb.y, g.y, w, c1, c2, a, nv.y);
  nx.z = move1(x.z, v.z, pb.z, g.z, w, c1, c2, a, nv.z);
  nx.w = move1(x.w, v.w, pb.w, g.w, w, c1, c2, a, nv.w);
  __stcs(a.vel + i, nv);
  __stcs(a.pos + i, nx);
  if (a.z != nullptr) {
    __nv_bfloat162 lo2 = __floats2bfloat162_rn(nx.x, nx.y);
    __nv_bfloat162 hi2 = __floats2bfloat162_rn(nx.z, nx.w);
    uint2 packed;
    packed.x = *reinterpret_cast<uint32_t*>(&lo2);
    packed.y = *reinterpret_cast<uint32_t*>(&hi2);
    *reinterpret_cast<uint2*>(a.z + (size_t)p * a.z_pitch + 4 * j) = packed;
  }
}

__global__ void pack_latent_kernel(const float4* pos, uint32_t total4, uint32_t d4, __nv_bfloat16* z, int z_pitch) {
  uint32_t i = blockIdx.x * 256u + threadIdx.x;
  if (i >= total4) return;
  uint32_t p = i / d4, j = i - p * d4;
  float4 x = __ldcs(pos + i);
  __nv_bfloat162 lo2 = __floats2bfloat162_rn(x.x, x.y), hi2 = __floats2bfloat162_rn(x.z, x.w);
  uint2 packed;
  packed.x = *reinterpret_cast<uint32_t*>(&lo2);
  packed.y = *reinterpret_cast<uint32_t*>(&hi2);
  *reinterpret_cast<uint2*>(z + (size_t)p * z_pitch + 4 * j) = packed;
}

__global__ void pack_latent_scalar_kernel(const float* pos, long long total, int D, __nv_bfloat16* z, int z_pitch) {
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  long long p = i / D;
  int j = (int)(i - p * D);
  z[p * z_pitch + j] = __float2bfloat16_rn(pos[i]);
}

// ------------------------------------------------------------------------------------------
// personal best + min-loc
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long k) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long other = __shfl_xor_sync(0xffffffffu, k, o);
    k = other < k ? other : k;
  }
  return k;
}

__global__ void __launch_bounds__(256) pbest_minloc_kernel(const float* __restrict__ fit, float* __restrict__ pbest_val,
                                                           uint8_t* __restrict__ improved, long long P,
                                                           unsigned long long index_offset,
                                                           unsigned long long* __restrict__ key) {
  __shared__ unsigned long long s_k[8];
  unsigned long long k = ~0ull;
  for (long long p = (long long)blockIdx.x * 256 + threadIdx.x; p < P; p += (long long)gridDim.x * 256) {
    float f = fit[p];
    float pv = pbest_val[p];
    bool imp = f < pv;                       // strict; NaN never improves (src/pso/util_pso.py:122)
    if (imp) { pbest_val[p] = f; pv = f; }
    improved[p] = imp ? 1 : 0;
    unsigned long long kk = ((unsigned long long)order_key(pv) << 32) | (unsigned long long)(uint32_t)(index_offset + p);
    k = kk < k ? kk : k;
  }
  k = warp_min_u64(k);
  if ((threadIdx.x & 31) == 0) s_k[threadIdx.x >> 5] = k;
  __syncthreads();
  if (threadIdx.x < 32) {
    k = threadIdx.x < 8 ? s_k[threadIdx.x] : ~0ull;
    k = warp_min_u64(k);
    if (threadIdx.x == 0) atomicMin(key, k);
  }
}

__global__ void gbest_candidate_kernel(const unsigned long long* key, const float* pos, const float* pbest_pos,
                                       const uint8_t* improved, const float* pbest_val, long long P,
                                       unsigned long long index_offset, int D, float* rec, unsigned long long* rearm) {
  unsigned long long k = *key;
  __syncthreads();                                  // every thread has read the key before it is re-armed
  if (rearm != nullptr && threadIdx.x == 0) *rearm = ~0ull;
  long long local = (long long)(uint32_t)(k & 0xffffffffull) - (long long)(uint32_t)index_offset;
  bool ok = local >= 0 && local < P;
  if (threadIdx.x == 0) {
    reinterpret_cast<uint32_t*>(rec)[0] = (uint32_t)(k & 0xffffffffull);
    reinterpret_cast<uint32_t*>(rec)[1] = (uint32_t)(k >> 32);
    rec[2] = ok ? pbest_val[local] : INFINITY;
    rec[3] = 0.f;
  }
  if (ok) {
    const float* src = (improved != nullptr && improved[local]) ? pos + local * D : pbest_pos + local * D;
    for (int d = threadIdx.x; d < D; d += blockDim.x) rec[4 + d] = src[d];
  } else {
    for (int d = threadIdx.x; d < D; d += blockDim.x) rec[4 + d] = 0.f;
  }
}

__global__ void gbest_apply_kernel(const float* recs, int n_rec, int D, float* gbest_val, float* gbest_pos, float* hist,
                                   int hist_rows, psob_iter_state* dyn, float inertia_decay) {
  __shared__ int s_best;
  __shared__ int s_replace;
  if (threadIdx.x == 0) {
    unsigned long long best = ~0ull;
    int bi = 0;
    for (int r = 0; r < n_rec; ++r) {
      const uint32_t* w = reinterpret_cast<const uint32_t*>(recs + (size_t)r * (4 + D));
      unsigned long long k = ((unsigned long long)w[1] << 32) | w[0];
      if (k < best) { best = k; bi = r; }
    }
    float cand = recs[(size_t)bi * (4 + D) + 2];
    float inc = *gbest_val;
    bool rep = cand < inc;                    // equal never replaces (src/pso/util_pso.py:144)
    if (rep) *gbest_val = cand;
    if (hist != nullptr) {
      // rows past the recorder's capacity are dropped (never clamped onto the last row): SwarmEngine.history() returns
      // min(iteration, hist_rows) rows and SwarmEngine.run refuses to go past max_iterations with a recorder attached
      const int hrow = dyn ? (int)dyn->iteration : 0;
      if (hrow < hist_rows) {
        hist[2 * hrow] = rep ? cand : inc;
        hist[2 * hrow + 1] = rep ? 1.f : 0.f;
      }
    }
    if (dyn != nullptr) {
      // `if i > 1 and schedule_inertia: w = 0.99*w` runs after the global best and before the move
      uint32_t i_now = dyn->iteration + 1;
      dyn->iteration = i_now;
      if (i_now > 1 && inertia_decay != 1.0f) dyn->w = inertia_decay * dyn->w;
    }
    s_best = bi;
    s_replace = rep ? 1 : 0;
  }
  __syncthreads();
  if (s_replace) {
    const float* src = recs + (size_t)s_best * (4 + D) + 4;
    for (int d = threadIdx.x; d < D; d += blockDim.x) gbest_pos[d] = src[d];
  }
}

// Single-rank global best: psob_gbest_candidate + psob_gbest_apply in one launch, and the min-loc key is re-armed
// (reset to ~0) for the next iteration's psob_fitness_tail, so no memset node sits in the iteration.
__global__ void gbest_local_kernel(unsigned long long* key, const float* pos, const float* pbest_pos, const uint8_t* improved,
                                   const float* pbest_val, long long P, unsigned long long index_offset, int D, float* gbest_val,
                                   float* gbest_pos, float* hist, int hist_rows, psob_iter_state* dyn, float inertia_decay) {
  __shared__ long long s_local;
  __shared__ int s_replace;
  if (threadIdx.x == 0) {
    const unsigned long long k = *key;
    *key = ~0ull;
    const long long local = (long long)(uint32_t)(k & 0xffffffffull) - (long long)(uint32_t)index_offset;
    const bool ok = local >= 0 && local < P;
    const float cand = ok ? pbest_val[local] : INFINITY;
    const float inc = *gbest_val;
    const bool rep = cand < inc;                    // equal never replaces (src/pso/util_pso.py:144)
    if (rep) *gbest_val = cand;
    if (hist != nullptr) {
      const int hrow = dyn ? (int)dyn->iteration : 0;
      if (hrow < hist_rows) {
        hist[2 * hrow] = rep ? cand : inc;
        hist[2 * hrow + 1] = rep ? 1.f : 0.f;
      }
    }
    if (dyn != nullptr) {
      const uint32_t i_now = dyn->iteration + 1;
      dyn->iteration = i_now;
      if (i_now > 1 && inertia_decay != 1.0f) dyn->w = inertia_decay * dyn->w;
    }
    s_local = local;
    s_replace = (rep && ok) ? 1 : 0;
  }
  __syncthreads();
  if (s_replace) {
    const long long local = s_local;
    const float* src = (improved != nullptr && improved[local]) ? pos + local * D : pbest_pos + local * D;
    for (int d = threadIdx.x; d < D; d += blockDim.x) gbest_pos[d] = src[d];
  }
}

// ------------------------------------------------------------------------------------------
// Fused fitness tail: everything between the last network layer and the global best in ONE launch --
//   fused-head finish (partial sums + bias + sigmoid)  |  softmax pick  |  ready score
//   -> clamp / affine (src/pso/util_discovery.py:75-80)
//   -> + w_rec * mean((target - G(z))^2) + eps from the image epilogue's partial sums (:97-101)
//   -> strict-< personal-best compare (src/pso/util_pso.py:122-124) -> block min-loc -> atomicMin(key)
// Every arithmetic step is the one the separate kernels (dot_finish, score_fitness, softmax_select, recon_fitness,
// pbest_minloc) perform, in the same order, so the results are bit-identical to the unfused chain.
// ------------------------------------------------------------------------------------------
struct TailParams {
  const float* part; int n_slots; float bias; int act;
  const float* score; int score_stride;
  const float* logits; int n_class, class_idx;
  const float* fit_in;
  int mode; float thr, eps;
  const float* recon_part; int n_recon_slots; float n_pix; float w_ass, w_rec, eps_rec;
  float* fit; float* pbest_val; uint8_t* improved; long long P; unsigned long long index_offset; unsigned long long* key;
};

__global__ void __launch_bounds__(256) fitness_tail_kernel(const TailParams a) {
  __shared__ unsigned long long s_k[8];
  unsigned long long k = ~0ull;
  for (long long p = (long long)blockIdx.x * 256 + threadIdx.x; p < a.P; p += (long long)gridDim.x * 256) {
    float f;
    bool have_score = true;
    float sc = 0.f;
    if (a.part != nullptr) {
      float acc = 0.f;
      for (int sl = 0; sl < a.n_slots; ++sl) acc += a.part[(size_t)sl * (size_t)a.P + p];
      const float v = acc + a.bias;
      sc = a.act == PSOB_ACT_SIGMOID ? 1.f / (1.f + expf(-v)) : v;
    } else if (a.logits != nullptr) {
      const float* l = a.logits + p * a.n_class;
      float m = -INFINITY;
      for (int c = 0; c < a.n_class; ++c) m = fmaxf(m, l[c]);
      float sum = 0.f;
      for (int c = 0; c < a.n_class; ++c) sum += expf(l[c] - m);
      sc = expf(l[a.class_idx] - m) / sum;
    } else if (a.score != nullptr) {
      sc = a.score[p * a.score_stride];
    } else {
      have_score = false;
    }
    if (have_score) {
      const float s = fminf(__fadd_rn(sc, a.thr), 1.0f);
      f = a.mode == PSOB_MODE_IN_TRAINING ? __fadd_rn(s, a.eps) : __fadd_rn(__fsub_rn(1.0f, s), a.eps);
    } else {
      f = a.fit_in != nullptr ? a.fit_in[p] : 0.f;
    }
    if (a.recon_part != nullptr) {
      float acc = 0.f;
      for (int sl = 0; sl < a.n_recon_slots; ++sl) acc += a.recon_part[(size_t)sl * (size_t)a.P + p];
      const float mse = acc / a.n_pix;
      const float ass = (have_score || a.fit_in != nullptr) ? a.w_ass * f : 0.f;
      f = ass + a.w_rec * mse + a.eps_rec;
    }
    if (a.fit != nullptr) a.fit[p] = f;
    float pv = a.pbest_val[p];
    const bool imp = f < pv;                       // strict; NaN never improves (src/pso/util_pso.py:122)
    if (imp) { a.pbest_val[p] = f; pv = f; }
    a.improved[p] = imp ? 1 : 0;
    const unsigned long long kk = ((unsigned long long)order_key(pv) << 32) | (unsigned long long)(uint32_t)(a.index_offset + p);
    k = kk < k ? kk : k;
  }
  k = warp_min_u64(k);
  if ((threadIdx.x & 31) == 0) s_k[threadIdx.x >> 5] = k;
  __syncthreads();
  if (threadIdx.x < 32) {
    k = threadIdx.x < 8 ? s_k[threadIdx.x] : ~0ull;
    k = warp_min_u64(k);
    if (threadIdx.x == 0) atomicMin(a.key, k);
  }
}

// ------------------------------------------------------------------------------------------
// fitness transforms
// ------------------------------------------------------------------------------------------
__global__ void score_fitness_kernel(const float* score, int stride, long long P, int mode, float thr, float eps, float* fit) {
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  float s = fminf(__fadd_rn(score[p * stride], thr), 1.0f);     // torch.min(p + thr, 1)  (util_discovery.py:76,78)
  fit[p] = mode == PSOB_MODE_IN_TRAINING ? __fadd_rn(s, eps) : __fadd_rn(__fsub_rn(1.0f, s), eps);
}

__global__ void softmax_select_kernel(const float* logits, int n_class, int class_idx, long long P, float* prob) {
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const float* l = logits + p * n_class;
  float m = -INFINITY;
  for (int c = 0; c < n_class; ++c) m = fmaxf(m, l[c]);
  float sum = 0.f;
  for (int c = 0; c < n_class; ++c) sum += expf(l[c] - m);
  prob[p] = expf(l[class_idx] - m) / sum;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// one group of `tpp` threads (32 or 256) per particle
template <bool kBf16Target>
__global__ void __launch_bounds__(256) recon_fitness_kernel(const float* __restrict__ img, const void* __restrict__ target,
                                                            long long P, int n, int tpp, const float* __restrict__ fit_ass,
                                                            float w_ass, float w_rec, float eps, float* __restrict__ fit) {
  __shared__ float s_part[8];
  int groups = 256 / tpp;
  long long p = (long long)blockIdx.x * groups + threadIdx.x / tpp;
  int t = threadIdx.x % tpp;
  float acc = 0.f;
  if (p < P) {
    const float* x = img + p * n;
    if (kBf16Target) {
      const __nv_bfloat16* tg = reinterpret_cast<const __nv_bfloat16*>(target) + p * n;
      for (int i = t; i < n; i += tpp) { float d = __bfloat162float(tg[i]) - x[i]; acc += d * d; }
    } else {
      const float* tg = reinterpret_cast<const float*>(target) + p * n;
      for (int i = t; i < n; i += tpp) { float d = tg[i] - x[i]; acc += d * d; }
    }
  }
  acc = warp_sum(acc);
  if (tpp == 256) {
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
      acc = threadIdx.x < 8 ? s_part[threadIdx.x] : 0.f;
      acc = warp_sum(acc);
    }
  }
  if (t == 0 && p < P) {
    float mse = acc / (float)n;
    float ass = fit_ass != nullptr ? w_ass * fit_ass[p] : 0.f;
    fit[p] = ass + w_rec * mse + eps;        // util_discovery.py:97-101 (the caller's fit_ass already holds one eps)
  }
}

__global__ void __launch_bounds__(256) rescale_minmax_kernel(const float* __restrict__ img, long long P, int n, float* __restrict__ out) {
  // one warp per image (src/utils/util_data.py:129-135 with batch 1)
  long long p = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (p >= P) return;
  int lane = threadIdx.x & 31;
  const float* x = img + p * n;
  float lo = INFINITY, hi = -INFINITY;
  for (int i = lane; i < n; i += 32) { float v = x[i]; lo = fminf(lo, v); hi = fmaxf(hi, v); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  float range = __fsub_rn(hi, lo);
  float* y = out + p * n;
  for (int i = lane; i < n; i += 32) y[i] = __fdiv_rn(__fsub_rn(x[i], lo), range);
}

__global__ void pack_image_kernel(const float* __restrict__ img, long long P, __nv_bfloat16* __restrict__ out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long total = P * 992;
  if (i >= total) return;
  long long p = i / 992;
  int r = (int)(i - p * 992);
  int y = r / 32 - 1, x = r % 32 - 1;
  float v = (y >= 0 && y < 28 && x >= 0 && x < 28) ? img[p * 784 + y * 28 + x] : 0.f;
  out[i] = __float2bfloat16_rn(v);
}

}  // namespace psob

using namespace psob;

// ------------------------------------------------------------------------------------------
// Swarm diversity: mean Euclidean distance over all unordered particle pairs -- `Swarm.mse`
// (src/pso/util_pso.py:76-86; reporting only, O(P^2 D)).  One block per 128 x 128 tile of the upper
// triangle; both 128-particle slabs sit in shared memory (the second one transposed, so that two neighbouring
// particles of one dimension are one 8-byte load), every thread owns an 8 x 8 patch of pairs and advances two
// pairs per instruction on the packed fp32 pipe (sub.f32x2 + fma.f32x2: one instruction per pair and dimension).
// Differences are squared directly (no Gram-matrix cancellation), distances are summed per block in double and
// added with one atomicAdd(double).
// ------------------------------------------------------------------------------------------
constexpr int kPairTile = 128;

// packed fp32 pair: {lo, hi} in one 64-bit register (Blackwell FADD2 / FFMA2: two IEEE fp32 lanes per instruction)
__device__ __forceinline__ unsigned long long f2_pack(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ unsigned long long f2_sqdiff_acc(unsigned long long a, unsigned long long b, unsigned long long acc) {
  unsigned long long d, r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(r) : "l"(d), "l"(acc));
  return r;
}

__global__ void __launch_bounds__(256)
pairwise_distance_sum_kernel(const float* __restrict__ pos, long long P, int D, int pitch, int n_tiles, double* __restrict__ out) {
  extern __shared__ __align__(16) float sm[];
  // linear block index -> (ti <= tj) of the upper triangle
  long long b = blockIdx.x;
  int ti = 0;
  while (b >= (long long)(n_tiles - ti)) { b -= (n_tiles - ti); ++ti; }
  const int tj = ti + (int)b;
  const int lda = D + 1;                  // A: [128 rows][D + 1]  (row reads are warp broadcasts)
  const int ldb = kPairTile + 2;          // B: transposed [D][128 + 2]: two neighbouring particles of one dimension = one 8-byte load
  float* A = sm;
  float* Bt = sm + ((kPairTile * lda + 1) & ~1);
  for (int idx = threadIdx.x; idx < kPairTile * D; idx += 256) {
    const int r = idx / D, c = idx - r * D;
    const long long ia = (long long)ti * kPairTile + r, ib = (long long)tj * kPairTile + r;
    A[r * lda + c] = ia < P ? pos[ia * pitch + c] : 0.f;
    Bt[c * ldb + r] = ib < P ? pos[ib * pitch + c] : 0.f;
  }
  __syncthreads();
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  // thread tile: rows ty + 16 i (i < 8) x column pairs (2 tx + 32 j, 2 tx + 32 j + 1) (j < 4)
  unsigned long long acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;
  for (int c = 0; c < D; ++c) {
    unsigned long long bb[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) bb[j] = *reinterpret_cast<const unsigned long long*>(Bt + c * ldb + 2 * tx + 32 * j);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float a = A[(ty + 16 * i) * lda + c];
      const unsigned long long aa = f2_pack(a, a);
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = f2_sqdiff_acc(aa, bb[j], acc[i][j]);
    }
  }
  double sum = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float lo, hi;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i][j]));
      const long long gi = (long long)ti * kPairTile + ty + 16 * i, gj = (long long)tj * kPairTile + 2 * tx + 32 * j;
      if (gi < gj && gj < P) sum += (double)sqrtf(lo);
      if (gi < gj + 1 && gj + 1 < P) sum += (double)sqrtf(hi);
    }
  __shared__ double red[8];
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    atomicAdd(out, t);
  }
}

// ------------------------------------------------------------------------------------------
// Nearest codebook entry -- `vq` (src/inverter/utils_vq_vae/util_function.py:4-23):
// argmin_k (|c_k|^2 + |x|^2 - 2 x.c_k), lowest k on ties (torch.min).  A block owns 64 input rows (shared memory,
// pitch D + 1) and walks the codebook in slabs of 64 codes; thread (ty, tx) of 16 x 16 keeps a 4 x 4 register tile of
// dot products (rows ty + 16 i, codes tx + 16 j: conflict-free shared loads, 8 loads per 16 FMAs), folds each slab into
// its running (distance, index) minimum, and the 16 threads of a row reduce lexicographically at the end.
// ------------------------------------------------------------------------------------------
constexpr int kVqTile = 64;

__global__ void __launch_bounds__(256)
vq_nearest_kernel(const float* __restrict__ x, long long P, int D, int x_pitch, const float* __restrict__ cb, int K,
                  long long* __restrict__ idx_out, float* __restrict__ dist_out) {
  extern __shared__ float sm[];
  const int ld = D + 1;
  float* X = sm;                          // [64][ld]
  float* C = X + kVqTile * ld;            // [64][ld]
  float* X2 = C + kVqTile * ld;           // [64]
  float* C2 = X2 + kVqTile;               // [64]
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const long long row0 = (long long)blockIdx.x * kVqTile;
  for (int idx = threadIdx.x; idx < kVqTile * D; idx += 256) {
    const int r = idx / D, c = idx - r * D;
    X[r * ld + c] = row0 + r < P ? x[(row0 + r) * x_pitch + c] : 0.f;
  }
  __syncthreads();
  if (threadIdx.x < kVqTile) {
    float s2 = 0.f;
    for (int c = 0; c < D; ++c) s2 = fmaf(X[threadIdx.x * ld + c], X[threadIdx.x * ld + c], s2);
    X2[threadIdx.x] = s2;
  }
  float best[4];
  int best_k[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { best[i] = INFINITY; best_k[i] = 0x7fffffff; }
  for (int k0 = 0; k0 < K; k0 += kVqTile) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < kVqTile * D; idx += 256) {
      const int r = idx / D, c = idx - r * D;
      C[r * ld + c] = k0 + r < K ? cb[(size_t)(k0 + r) * D + c] : 0.f;
    }
    __syncthreads();
    if (threadIdx.x < kVqTile) {
      float s2 = 0.f;
      for (int c = 0; c < D; ++c) s2 = fmaf(C[threadIdx.x * ld + c], C[threadIdx.x * ld + c], s2);
      C2[threadIdx.x] = s2;
    }
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int c = 0; c < D; ++c) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = X[(ty + 16 * i) * ld + c]; b[i] = C[(tx + 16 * i) * ld + c]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();                       // C2 of this slab is complete
#pragma unroll
    for (int j = 0; j < 4; ++j) {          // ascending code index per thread: strict < keeps the lowest index
      const int k = k0 + tx + 16 * j;
      if (k >= K) continue;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float d = (C2[tx + 16 * j] + X2[ty + 16 * i]) - 2.f * acc[i][j];
        if (d < best[i]) { best[i] = d; best_k[i] = k; }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float bd = best[i];
    int bk = best_k[i];
    for (int o = 8; o > 0; o >>= 1) {      // the 16 threads of a row are 16 consecutive lanes
      const float od = __shfl_xor_sync(0xffffffffu, bd, o);
      const int ok = __shfl_xor_sync(0xffffffffu, bk, o);
      if (od < bd || (od == bd && ok < bk)) { bd = od; bk = ok; }
    }
    const long long row = row0 + ty + 16 * i;
    if (tx == 0 && row < P) {
      idx_out[row] = bk;
      if (dist_out) dist_out[row] = bd;
    }
  }
}

// Wide rows (2 x 64 x (D + 1) floats no longer fit in shared memory): one warp per input row, lane l scans codes
// l, l + 32, ... of a shared-memory codebook slab.
__global__ void __launch_bounds__(256)
vq_nearest_rowwarp_kernel(const float* __restrict__ x, long long P, int D, int x_pitch, const float* __restrict__ cb, int K, int slab,
                  long long* __restrict__ idx_out, float* __restrict__ dist_out) {
  extern __shared__ float sm[];
  const int ld = D + 1;
  float* C = sm;                       // [slab][ld]
  float* C2 = sm + (size_t)slab * ld;  // [slab]
  float* X = C2 + slab;                // [8 warps][D]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + warp;
  float x2 = 0.f;
  if (row < P) {
    for (int c = lane; c < D; c += 32) { const float v = x[row * x_pitch + c]; X[warp * D + c] = v; x2 = fmaf(v, v, x2); }
  }
  for (int o = 16; o > 0; o >>= 1) x2 += __shfl_xor_sync(0xffffffffu, x2, o);
  float best = INFINITY;
  int best_k = 0x7fffffff;
  for (int k0 = 0; k0 < K; k0 += slab) {
    const int n = min(slab, K - k0);
    __syncthreads();
    for (int idx = threadIdx.x; idx < n * D; idx += 256) { const int r = idx / D, c = idx - r * D; C[r * ld + c] = cb[(size_t)(k0 + r) * D + c]; }
    __syncthreads();
    for (int r = threadIdx.x; r < n; r += 256) {
      float s2 = 0.f;
      for (int c = 0; c < D; ++c) s2 = fmaf(C[r * ld + c], C[r * ld + c], s2);
      C2[r] = s2;
    }
    __syncthreads();
    if (row < P) {
      for (int r = lane; r < n; r += 32) {
        float dot = 0.f;
        for (int c = 0; c < D; ++c) dot = fmaf(X[warp * D + c], C[r * ld + c], dot);
        const float d = (C2[r] + x2) - 2.f * dot;
        if (d < best) { best = d; best_k = k0 + r; }      // ascending k per lane: strict < keeps the lowest index
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int ok = __shfl_xor_sync(0xffffffffu, best_k, o);
    if (ob < best || (ob == best && ok < best_k)) { best = ob; best_k = ok; }
  }
  if (row < P && lane == 0) {
    idx_out[row] = best_k;
    if (dist_out) dist_out[row] = best;
  }
}

// ------------------------------------------------------------------------------------------
// Assessor input: fp32 28x28 image -> the stem's bf16 rows (zero padding 3, row pitch 40, 36 rows; see
// lower_conv_image_k7s2p3), optionally through the per-image min-max rescale of Discovery.particles_to_img
// (src/pso/util_discovery.py:45 -> src/utils/util_data.py:129-135) in the same pass.  One warp per image.
// ------------------------------------------------------------------------------------------
constexpr int kStemPad = 3, kStemPitch = 40, kStemRows = 36;

__global__ void __launch_bounds__(256)
pack_image_stem_kernel(const float* __restrict__ img, long long P, int rescale, __nv_bfloat16* __restrict__ out, float* __restrict__ out32) {
  // the image goes through shared memory once: float4 loads, min/max from registers, 16-byte bf16 stores
  __shared__ __align__(16) float simg[8][784];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long p = (long long)blockIdx.x * 8 + w;
  if (p >= P) return;
  const float4* x4 = reinterpret_cast<const float4*>(img + p * 784);
  float mn = INFINITY, mx = -INFINITY;
  for (int i = lane; i < 196; i += 32) {
    const float4 v = __ldg(x4 + i);
    reinterpret_cast<float4*>(simg[w])[i] = v;
    mn = fminf(fminf(mn, fminf(v.x, v.y)), fminf(v.z, v.w));
    mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
  }
  float lo = 0.f, range = 1.f;
  if (rescale) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    lo = mn;
    range = __fsub_rn(mx, mn);
  }
  __syncwarp();
  if (rescale) {
    for (int i = lane; i < 784; i += 32) simg[w][i] = __fdiv_rn(__fsub_rn(simg[w][i], lo), range);
    __syncwarp();
    if (out32)
      for (int i = lane; i < 196; i += 32) reinterpret_cast<float4*>(out32 + p * 784)[i] = reinterpret_cast<const float4*>(simg[w])[i];
  } else if (out32) {
    for (int i = lane; i < 196; i += 32) reinterpret_cast<float4*>(out32 + p * 784)[i] = reinterpret_cast<const float4*>(simg[w])[i];
  }
  uint4* y = reinterpret_cast<uint4*>(out + p * (kStemRows * kStemPitch));       // 2880 B per image: 180 x 16 B
  for (int v = lane; v < kStemRows * kStemPitch / 8; v += 32) {
    const int r = v / (kStemPitch / 8) - kStemPad, c0 = (v % (kStemPitch / 8)) * 8 - kStemPad;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = c0 + e;
      f[e] = (r >= 0 && r < 28 && c >= 0 && c < 28) ? simg[w][r * 28 + c] : 0.f;
    }
    uint4 o;
    __nv_bfloat162 t;
    t = __floats2bfloat162_rn(f[0], f[1]); o.x = *reinterpret_cast<uint32_t*>(&t);
    t = __floats2bfloat162_rn(f[2], f[3]); o.y = *reinterpret_cast<uint32_t*>(&t);
    t = __floats2bfloat162_rn(f[4], f[5]); o.z = *reinterpret_cast<uint32_t*>(&t);
    t = __floats2bfloat162_rn(f[6], f[7]); o.w = *reinterpret_cast<uint32_t*>(&t);
    y[v] = o;
  }
}

// The assessor's own pool (3x3, stride 2, pad 1): one thread per (particle, output row, 8 channels) walks the
// input columns once -- column maxima over the <= 3 rows (all loads independent), each feeding the one or two
// windows it belongs to -- 6 loads per output instead of 9, with 3 W loads in flight per thread.
// element offset of (particle p, column col) in a row-major [P, pitch] or tile-major [P/128][pitch/64][128][64] activation
__device__ __forceinline__ long long out_offset(long long p, long long col, long long pitch, int tiled) {
  return tiled ? (((p >> 7) * (pitch >> 6) + (col >> 6)) * 128 + (p & 127)) * 64 + (col & 63) : p * pitch + col;
}

template <int W>
__global__ void __launch_bounds__(256)
maxpool3x3s2_rows_kernel(const __nv_bfloat16* __restrict__ x, long long P, int h, int cp, int ho, long long x_pitch, long long y_pitch,
                         int y_tiled, __nv_bfloat16* __restrict__ y) {
  constexpr int WO = (W + 2 - 3) / 2 + 1;
  const int c8 = cp >> 3;
  const long long per = (long long)ho * c8;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P * per) return;
  const long long p = i / per;
  const int r = (int)(i - p * per);
  const int c = r % c8, oy = r / c8;
  const __nv_bfloat162 ninf = __float2bfloat162_rn(-INFINITY);
  __nv_bfloat162 cm[W][4];
#pragma unroll
  for (int ix = 0; ix < W; ++ix)
#pragma unroll
    for (int e = 0; e < 4; ++e) cm[ix][e] = ninf;
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const int iy = 2 * oy - 1 + ky;
    if (iy < 0 || iy >= h) continue;
    const uint4* row = reinterpret_cast<const uint4*>(x + p * x_pitch + (long long)iy * W * cp) + c;
#pragma unroll
    for (int ix = 0; ix < W; ++ix) {
      const uint4 q = __ldg(row + ix * c8);
      const __nv_bfloat162* v = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
      for (int e = 0; e < 4; ++e) cm[ix][e] = __hmax2(cm[ix][e], v[e]);
    }
  }
#pragma unroll
  for (int ox = 0; ox < WO; ++ox) {
    __nv_bfloat162 m[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      m[e] = cm[2 * ox][e];
      if (2 * ox - 1 >= 0) m[e] = __hmax2(m[e], cm[2 * ox - 1][e]);
      if (2 * ox + 1 < W) m[e] = __hmax2(m[e], cm[2 * ox + 1][e]);
    }
    const long long col = ((long long)oy * WO + ox) * cp + 8 * c;
    *reinterpret_cast<uint4*>(y + out_offset(p, col, y_pitch, y_tiled)) = *reinterpret_cast<const uint4*>(m);
  }
}

// MaxPool2d(k, stride, pad) on a bf16 NHWC activation [P, h * w * cp] (the assessor's 3x3 stride-2 pool,
// src/pso/util_cnn.py:88).  One thread per 8 channels of one output pixel; padding never wins (-inf).
__global__ void __launch_bounds__(256)
maxpool_nhwc_kernel(const __nv_bfloat16* __restrict__ x, long long P, int h, int w, int cp, int k, int stride, int pad, int ho, int wo,
                    long long x_pitch, long long y_pitch, int y_tiled, __nv_bfloat16* __restrict__ y) {
  const int c8 = cp >> 3;
  const long long per = (long long)ho * wo * c8;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P * per) return;
  const long long p = i / per;
  int r = (int)(i - p * per);
  const int c = r % c8; r /= c8;
  const int ox = r % wo, oy = r / wo;
  float m[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) m[e] = -INFINITY;
  for (int ky = 0; ky < k; ++ky) {
    const int iy = oy * stride - pad + ky;
    if (iy < 0 || iy >= h) continue;
    for (int kx = 0; kx < k; ++kx) {
      const int ix = ox * stride - pad + kx;
      if (ix < 0 || ix >= w) continue;
      const uint4 q = __ldg(reinterpret_cast<const uint4*>(x + p * x_pitch + ((long long)iy * w + ix) * cp) + c);
      const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        m[2 * e] = fmaxf(m[2 * e], __uint_as_float(qw[e] << 16));
        m[2 * e + 1] = fmaxf(m[2 * e + 1], __uint_as_float(qw[e] & 0xFFFF0000u));
      }
    }
  }
  uint4 o;      // the maxima are bf16 values already: truncation is exact
  o.x = (__float_as_uint(m[0]) >> 16) | (__float_as_uint(m[1]) & 0xFFFF0000u);
  o.y = (__float_as_uint(m[2]) >> 16) | (__float_as_uint(m[3]) & 0xFFFF0000u);
  o.z = (__float_as_uint(m[4]) >> 16) | (__float_as_uint(m[5]) & 0xFFFF0000u);
  o.w = (__float_as_uint(m[6]) >> 16) | (__float_as_uint(m[7]) & 0xFFFF0000u);
  *reinterpret_cast<uint4*>(y + out_offset(p, ((long long)oy * wo + ox) * cp + 8 * c, y_pitch, y_tiled)) = o;
}

// fp32 NCHW latent grid [P, C, HW] -> bf16 NHWC rows [P, pitch] with channel pitch cp (padding channels 0), optional ReLU:
// the input of VectorQuantizedVAE.decoder, whose first module is an in-place ReLU (src/inverter/utils_vq_vae/util_model.py:151-164)
__global__ void __launch_bounds__(256) pack_grid_nchw_kernel(const float* __restrict__ x, long long P, int C, int HW, int cp, int relu,
                                                             __nv_bfloat16* __restrict__ out, long long pitch) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = (long long)HW * cp;
  if (i >= P * per) return;
  const long long p = i / per;
  const int r = (int)(i - p * per);
  const int pix = r / cp, c = r - pix * cp;
  float v = c < C ? x[(p * C + c) * HW + pix] : 0.f;
  if (relu) v = fmaxf(v, 0.f);
  out[p * pitch + r] = __float2bfloat16_rn(v);
}

// z[p, :] = bf16(codebook[idx[p], :]) zero-padded to z_pitch: the VQ-snapped generator input (vq + embedding lookup,
// src/inverter/utils_vq_vae/util_model.py:133-136, util_function.py:4-23)
__global__ void __launch_bounds__(256) gather_pack_latent_kernel(const float* __restrict__ codebook, const long long* __restrict__ idx, long long P,
                                                                 int D, __nv_bfloat16* __restrict__ z, int z_pitch) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P * D) return;
  const long long p = i / D;
  const int j = (int)(i - p * D);
  z[p * z_pitch + j] = __float2bfloat16_rn(codebook[idx[p] * D + j]);
}

__global__ void __launch_bounds__(256) fill_columns_bf16_kernel(__nv_bfloat16* __restrict__ x, long long P, long long pitch, int col0, int n,
                                                                __nv_bfloat16 value) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P * n) return;
  x[(i / n) * pitch + col0 + (i % n)] = value;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is raised once per (kernel, device) to the largest size seen so far, not on
// every call (the entry points promise to be plain stream work)
#include <mutex>
#include <map>
#include <utility>
static int ensure_dyn_smem(const void* fn, size_t bytes) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, size_t> seen;
  int dev = 0;
  PSOB_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> g(mu);
  size_t& cur = seen[std::make_pair(fn, dev)];
  if (bytes > cur) {
    PSOB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    cur = bytes;
  }
  return 0;
}

extern "C" {

int psob_version(void) { return PSOB_VERSION; }
const char* psob_last_error(void) { return g_err; }

int psob_device_info(int* cc_major, int* cc_minor, int* sm_count) {
  int dev = 0;
  PSOB_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  PSOB_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (sm_count) *sm_count = prop.multiProcessorCount;
  return 0;
}

int psob_update(const psob_update_args* a, psob_stream_t s) {
  if (!a || !a->pos || !a->vel || !a->pbest_pos || !a->gbest_pos) return fail(PSOB_E_ARG, "psob_update: null state pointer");
  if (a->P <= 0 || a->D <= 0 || a->D % 4) return fail(PSOB_E_ARG, "psob_update: need P > 0 and D %% 4 == 0 (D=%d)", a->D);
  if (!aligned16(a->pos) || !aligned16(a->vel) || !aligned16(a->pbest_pos) || !aligned16(a->gbest_pos))
    return fail(PSOB_E_ARG, "psob_update: state pointers must be 16-byte aligned");
  if ((a->r1 == nullptr) != (a->r2 == nullptr)) return fail(PSOB_E_ARG, "psob_update: give both r1 and r2 or neither");
  unsigned long long total4 = (unsigned long long)a->P * (a->D / 4);
  if (total4 >= (1ull << 32)) return fail(PSOB_E_UNSUPPORTED, "psob_update: P*D/4 must be < 2^32");
  if (a->z_bf16 && (a->z_pitch < a->D || a->z_pitch % 4 || (reinterpret_cast<uintptr_t>(a->z_bf16) & 7)))
    return fail(PSOB_E_ARG, "psob_update: z_pitch must be >= D, a multiple of 4, z 8-byte aligned");
  UpdateParams u;
  u.pos = reinterpret_cast<float4*>(a->pos); u.vel = reinterpret_cast<float4*>(a->vel);
  u.pbest = reinterpret_cast<float4*>(a->pbest_pos); u.gbest = reinterpret_cast<const float4*>(a->gbest_pos);
  u.improved = a->improved; u.r1 = a->r1; u.r2 = a->r2; u.z = reinterpret_cast<__nv_bfloat16*>(a->z_bf16);
  u.dyn = a->dyn;
  u.total4 = (uint32_t)total4; u.d4 = (uint32_t)(a->D / 4); u.z_pitch = a->z_pitch;
  u.w = a->w; u.c_cogn = a->c_cogn; u.c_soci = a->c_soci; u.v_max = a->v_max; u.lo = a->lo; u.hi = a->hi;
  u.seed_lo = (uint32_t)(a->seed & 0xffffffffull); u.seed_hi = (uint32_t)(a->seed >> 32); u.iteration = a->iteration;
  u.index_offset = a->index_offset;
  u.clamp_v = std::isfinite(a->v_max) ? 1 : 0;
  u.clamp_x = (std::isfinite(a->lo) || std::isfinite(a->hi)) ? 1 : 0;
  unsigned blocks = (unsigned)((total4 + 255) / 256);
  if (a->r1) pso_update_kernel<false><<<blocks, 256, 0, (cudaStream_t)s>>>(u);
  else       pso_update_kernel<true><<<blocks, 256, 0, (cudaStream_t)s>>>(u);
  PSOB_LAUNCH_CHECK("pso_update_kernel");
  return 0;
}

int psob_fill_columns_bf16(void* x, int64_t P, int64_t pitch, int32_t col0, int32_t n, float value, psob_stream_t s) {
  if (!x || P <= 0 || n <= 0 || col0 < 0 || col0 + n > pitch) return fail(PSOB_E_ARG, "psob_fill_columns_bf16: bad arguments");
  const long long total = (long long)P * n;
  fill_columns_bf16_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)s>>>(
      reinterpret_cast<__nv_bfloat16*>(x), P, pitch, col0, n, __float2bfloat16_rn(value));
  PSOB_LAUNCH_CHECK("fill_columns_bf16_kernel");
  return 0;
}

int psob_pack_grid_nchw(const float* x, int64_t P, int32_t C, int32_t HW, int32_t cp, int32_t relu, void* out_bf16, int64_t pitch, psob_stream_t s) {
  if (!x || !out_bf16 || P <= 0 || C <= 0 || HW <= 0 || cp < C || pitch < (int64_t)HW * cp) return fail(PSOB_E_ARG, "psob_pack_grid_nchw: bad arguments");
  const long long total = (long long)P * HW * cp;
  pack_grid_nchw_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)s>>>(x, P, C, HW, cp, relu, reinterpret_cast<__nv_bfloat16*>(out_bf16), pitch);
  PSOB_LAUNCH_CHECK("pack_grid_nchw_kernel");
  return 0;
}

int psob_gather_pack_latent(const float* codebook, const int64_t* idx, int64_t P, int32_t D, void* z, int32_t z_pitch, psob_stream_t s) {
  if (!codebook || !idx || !z || P <= 0 || D <= 0 || z_pitch < D) return fail(PSOB_E_ARG, "psob_gather_pack_latent: bad arguments");
  const long long total = (long long)P * D;
  gather_pack_latent_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)s>>>(codebook, reinterpret_cast<const long long*>(idx), P, D,
                                                                                     reinterpret_cast<__nv_bfloat16*>(z), z_pitch);
  PSOB_LAUNCH_CHECK("gather_pack_latent_kernel");
  return 0;
}

int psob_pack_latent(const float* pos, int64_t P, int32_t D, void* z, int32_t z_pitch, psob_stream_t s) {
  if (!pos || !z || P <= 0 || D <= 0 || z_pitch < D) return fail(PSOB_E_ARG, "psob_pack_latent: bad arguments");
  if (D % 4 || z_pitch % 4 || !aligned16(pos) || (reinterpret_cast<uintptr_t>(z) & 7)) {
    long long total = (long long)P * D;
    pack_latent_scalar_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)s>>>(pos, total, D, reinterpret_cast<__nv_bfloat16*>(z), z_pitch);
    PSOB_LAUNCH_CHECK("pack_latent_scalar_kernel");
    return 0;
  }
  unsigned long long total4 = (unsigned long long)P * (D / 4);
  if (total4 >= (1ull << 32)) return fail(PSOB_E_UNSUPPORTED, "psob_pack_latent: too large");
  pack_latent_kernel<<<(unsigned)((total4 + 255) / 256), 256, 0, (cudaStream_t)s>>>(
      reinterpret_cast<const float4*>(pos), (uint32_t)total4, (uint32_t)(D / 4), reinterpret_cast<__nv_bfloat16*>(z), z_pitch);
  PSOB_LAUNCH_CHECK("pack_latent_kernel");
  return 0;
}

int psob_pbest_minloc(const float* fit, float* pbest_val, uint8_t* improved, int64_t P, uint64_t index_offset,
                      uint64_t* key, psob_stream_t s) {
  if (!fit || !pbest_val || !improved || !key || P <= 0) return fail(PSOB_E_ARG, "psob_pbest_minloc: bad arguments");
  if (index_offset + (uint64_t)P > (1ull << 32)) return fail(PSOB_E_UNSUPPORTED, "psob_pbest_minloc: global index must fit 32 bits");
  PSOB_CUDA(cudaMemsetAsync(key, 0xFF, sizeof(uint64_t), (cudaStream_t)s));
  long long blocks = (P + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  pbest_minloc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)s>>>(fit, pbest_val, improved, P, index_offset,
                                                                    reinterpret_cast<unsigned long long*>(key));
  PSOB_LAUNCH_CHECK("pbest_minloc_kernel");
  return 0;
}

int psob_gbest_candidate(const uint64_t* key, const float* pos, const float* pbest_pos, const uint8_t* improved,
                         const float* pbest_val, int64_t P, uint64_t index_offset, int32_t D, float* rec, psob_stream_t s) {
  if (!key || !pos || !pbest_pos || !pbest_val || !rec || P <= 0 || D <= 0) return fail(PSOB_E_ARG, "psob_gbest_candidate: bad arguments");
  gbest_candidate_kernel<<<1, 128, 0, (cudaStream_t)s>>>(reinterpret_cast<const unsigned long long*>(key), pos, pbest_pos,
                                                        improved, pbest_val, P, index_offset, D, rec, nullptr);
  PSOB_LAUNCH_CHECK("gbest_candidate_kernel");
  return 0;
}

int psob_gbest_record(uint64_t* key, const float* pos, const float* pbest_pos, const uint8_t* improved,
                      const float* pbest_val, int64_t P, uint64_t index_offset, int32_t D, float* rec, psob_stream_t s) {
  if (!key || !pos || !pbest_pos || !pbest_val || !rec || P <= 0 || D <= 0) return fail(PSOB_E_ARG, "psob_gbest_record: bad arguments");
  gbest_candidate_kernel<<<1, 128, 0, (cudaStream_t)s>>>(reinterpret_cast<const unsigned long long*>(key), pos, pbest_pos,
                                                        improved, pbest_val, P, index_offset, D, rec,
                                                        reinterpret_cast<unsigned long long*>(key));
  PSOB_LAUNCH_CHECK("gbest_candidate_kernel");
  return 0;
}

int psob_gbest_local(uint64_t* key, const float* pos, const float* pbest_pos, const uint8_t* improved, const float* pbest_val,
                     int64_t P, uint64_t index_offset, int32_t D, float* gbest_val, float* gbest_pos, float* hist,
                     int32_t hist_rows, psob_iter_state* dyn, float inertia_decay, psob_stream_t s) {
  if (!key || !pos || !pbest_pos || !pbest_val || !gbest_val || !gbest_pos || P <= 0 || D <= 0 || (hist && hist_rows <= 0))
    return fail(PSOB_E_ARG, "psob_gbest_local: bad arguments");
  gbest_local_kernel<<<1, 128, 0, (cudaStream_t)s>>>(reinterpret_cast<unsigned long long*>(key), pos, pbest_pos, improved, pbest_val,
                                                    P, index_offset, D, gbest_val, gbest_pos, hist, hist_rows, dyn, inertia_decay);
  PSOB_LAUNCH_CHECK("gbest_local_kernel");
  return 0;
}

int psob_fitness_tail(const psob_fitness_tail_args* a, psob_stream_t s) {
  if (!a || !a->pbest_val || !a->improved || !a->key || a->P <= 0) return fail(PSOB_E_ARG, "psob_fitness_tail: null state pointer or P <= 0");
  const int n_src = (a->part != nullptr) + (a->logits != nullptr) + (a->score != nullptr) + (a->fit_in != nullptr);
  if (n_src > 1 || (n_src == 0 && a->recon_part == nullptr))
    return fail(PSOB_E_ARG, "psob_fitness_tail: give exactly one of part / logits / score / fit_in (or only recon_part)");
  if (a->part && a->n_slots <= 0) return fail(PSOB_E_ARG, "psob_fitness_tail: n_slots <= 0");
  if (a->part && a->act != PSOB_ACT_SIGMOID && a->act != PSOB_ACT_NONE) return fail(PSOB_E_UNSUPPORTED, "psob_fitness_tail: head activation %d", a->act);
  if (a->logits && (a->n_class <= 0 || a->class_idx < 0 || a->class_idx >= a->n_class)) return fail(PSOB_E_ARG, "psob_fitness_tail: class index out of range");
  if (a->score && a->score_stride <= 0) return fail(PSOB_E_ARG, "psob_fitness_tail: score_stride <= 0");
  if ((a->part || a->logits || a->score) && a->mode != PSOB_MODE_IN_TRAINING && a->mode != PSOB_MODE_OUT_TRAINING)
    return fail(PSOB_E_ARG, "psob_fitness_tail: unknown mode %d", a->mode);
  if (a->recon_part && (a->n_recon_slots <= 0 || a->n_pix <= 0)) return fail(PSOB_E_ARG, "psob_fitness_tail: recon partial sums need n_recon_slots > 0 and n_pix > 0");
  if (a->index_offset + (uint64_t)a->P > (1ull << 32)) return fail(PSOB_E_UNSUPPORTED, "psob_fitness_tail: global index must fit 32 bits");
  TailParams t;
  t.part = a->part; t.n_slots = a->n_slots; t.bias = a->bias; t.act = a->act;
  t.score = a->score; t.score_stride = a->score_stride;
  t.logits = a->logits; t.n_class = a->n_class; t.class_idx = a->class_idx;
  t.fit_in = a->fit_in;
  t.mode = a->mode; t.thr = a->thr; t.eps = a->eps;
  t.recon_part = a->recon_part; t.n_recon_slots = a->n_recon_slots; t.n_pix = (float)a->n_pix;
  t.w_ass = a->w_ass; t.w_rec = a->w_rec; t.eps_rec = a->eps_rec;
  t.fit = a->fit; t.pbest_val = a->pbest_val; t.improved = a->improved; t.P = a->P; t.index_offset = a->index_offset;
  t.key = reinterpret_cast<unsigned long long*>(a->key);
  long long blocks = (a->P + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  fitness_tail_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)s>>>(t);
  PSOB_LAUNCH_CHECK("fitness_tail_kernel");
  return 0;
}

int psob_gbest_apply(const float* recs, int32_t n_rec, int32_t D, float* gbest_val, float* gbest_pos, float* hist,
                     int32_t hist_rows, psob_iter_state* dyn, float inertia_decay, psob_stream_t s) {
  if (!recs || !gbest_val || !gbest_pos || n_rec <= 0 || D <= 0 || (hist && hist_rows <= 0)) return fail(PSOB_E_ARG, "psob_gbest_apply: bad arguments");
  gbest_apply_kernel<<<1, 128, 0, (cudaStream_t)s>>>(recs, n_rec, D, gbest_val, gbest_pos, hist, hist_rows, dyn, inertia_decay);
  PSOB_LAUNCH_CHECK("gbest_apply_kernel");
  return 0;
}

int psob_score_fitness(const float* score, int32_t stride, int64_t P, int32_t mode, float thr, float eps, float* fit, psob_stream_t s) {
  if (!score || !fit || P <= 0 || stride <= 0) return fail(PSOB_E_ARG, "psob_score_fitness: bad arguments");
  if (mode != PSOB_MODE_IN_TRAINING && mode != PSOB_MODE_OUT_TRAINING) return fail(PSOB_E_ARG, "psob_score_fitness: unknown mode %d", mode);
  score_fitness_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)s>>>(score, stride, P, mode, thr, eps, fit);
  PSOB_LAUNCH_CHECK("score_fitness_kernel");
  return 0;
}

int psob_softmax_select(const float* logits, int32_t n_class, int32_t class_idx, int64_t P, float* prob, psob_stream_t s) {
  if (!logits || !prob || P <= 0 || n_class <= 0 || class_idx < 0 || class_idx >= n_class) return fail(PSOB_E_ARG, "psob_softmax_select: bad arguments");
  softmax_select_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)s>>>(logits, n_class, class_idx, P, prob);
  PSOB_LAUNCH_CHECK("softmax_select_kernel");
  return 0;
}

int psob_recon_fitness(const float* img, const void* target, int32_t target_is_bf16, int64_t P, int32_t n, const float* fit_ass,
                       float w_ass, float w_rec, float eps, float* fit, psob_stream_t s) {
  if (!img || !target || !fit || P <= 0 || n <= 0) return fail(PSOB_E_ARG, "psob_recon_fitness: bad arguments");
  int tpp = n >= 4096 ? 256 : 32;
  long long groups = 256 / tpp;
  unsigned blocks = (unsigned)((P + groups - 1) / groups);
  if (target_is_bf16) recon_fitness_kernel<true><<<blocks, 256, 0, (cudaStream_t)s>>>(img, target, P, n, tpp, fit_ass, w_ass, w_rec, eps, fit);
  else                recon_fitness_kernel<false><<<blocks, 256, 0, (cudaStream_t)s>>>(img, target, P, n, tpp, fit_ass, w_ass, w_rec, eps, fit);
  PSOB_LAUNCH_CHECK("recon_fitness_kernel");
  return 0;
}

int psob_rescale_minmax(const float* img, int64_t P, int32_t n, float* out, psob_stream_t s) {
  if (!img || !out || P <= 0 || n <= 0) return fail(PSOB_E_ARG, "psob_rescale_minmax: bad arguments");
  rescale_minmax_kernel<<<(unsigned)((P + 7) / 8), 256, 0, (cudaStream_t)s>>>(img, P, n, out);
  PSOB_LAUNCH_CHECK("rescale_minmax_kernel");
  return 0;
}

int psob_pack_image(const float* img, int64_t P, void* out_bf16, psob_stream_t s) {
  if (!img || !out_bf16 || P <= 0) return fail(PSOB_E_ARG, "psob_pack_image: bad arguments");
  long long total = P * 992;
  pack_image_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)s>>>(img, P, reinterpret_cast<__nv_bfloat16*>(out_bf16));
  PSOB_LAUNCH_CHECK("pack_image_kernel");
  return 0;
}

int psob_pairwise_distance_sum(const float* pos, int64_t P, int32_t D, int32_t pitch, double* sum_out, psob_stream_t s) {
  if (!pos || !sum_out || P <= 0 || D <= 0 || pitch < D) return fail(PSOB_E_ARG, "psob_pairwise_distance_sum: bad arguments");
  const long long n_tiles = (P + kPairTile - 1) / kPairTile;
  const long long n_blocks = n_tiles * (n_tiles + 1) / 2;
  if (n_blocks >= (1ll << 31)) return fail(PSOB_E_UNSUPPORTED, "psob_pairwise_distance_sum: P = %lld is too large (sub-sample the swarm)", (long long)P);
  const size_t smem = ((size_t)((kPairTile * (D + 1) + 1) & ~1) + (size_t)D * (kPairTile + 2)) * sizeof(float);
  if (smem > 200 * 1024) return fail(PSOB_E_UNSUPPORTED, "psob_pairwise_distance_sum: D = %d is too large", D);
  if (int rc = ensure_dyn_smem((const void*)pairwise_distance_sum_kernel, smem)) return rc;
  PSOB_CUDA(cudaMemsetAsync(sum_out, 0, sizeof(double), (cudaStream_t)s));
  pairwise_distance_sum_kernel<<<(unsigned)n_blocks, 256, smem, (cudaStream_t)s>>>(pos, P, D, pitch, (int)n_tiles, sum_out);
  PSOB_LAUNCH_CHECK("pairwise_distance_sum_kernel");
  return 0;
}

int psob_vq_nearest(const float* x, int64_t P, int32_t D, int32_t x_pitch, const float* codebook, int32_t K,
                    int64_t* idx_out, float* dist_out, psob_stream_t s) {
  if (!x || !codebook || !idx_out || P <= 0 || D <= 0 || K <= 0 || x_pitch < D) return fail(PSOB_E_ARG, "psob_vq_nearest: bad arguments");
  const size_t smem = ((size_t)2 * kVqTile * (D + 1) + 2 * kVqTile) * sizeof(float);
  if (smem > 160 * 1024) {
    int slab = K;
    auto bytes = [&](int sl) { return ((size_t)sl * (D + 1) + sl + (size_t)8 * D) * sizeof(float); };
    while (slab > 32 && bytes(slab) > 160 * 1024) slab = (slab + 1) / 2;
    if (bytes(slab) > 200 * 1024) return fail(PSOB_E_UNSUPPORTED, "psob_vq_nearest: D = %d is too large", D);
    if (int rc = ensure_dyn_smem((const void*)vq_nearest_rowwarp_kernel, bytes(slab))) return rc;
    vq_nearest_rowwarp_kernel<<<(unsigned)((P + 7) / 8), 256, bytes(slab), (cudaStream_t)s>>>(
        x, P, D, x_pitch, codebook, K, slab, reinterpret_cast<long long*>(idx_out), dist_out);
    PSOB_LAUNCH_CHECK("vq_nearest_rowwarp_kernel");
    return 0;
  }
  if (int rc = ensure_dyn_smem((const void*)vq_nearest_kernel, smem)) return rc;
  vq_nearest_kernel<<<(unsigned)((P + kVqTile - 1) / kVqTile), 256, smem, (cudaStream_t)s>>>(
      x, P, D, x_pitch, codebook, K, reinterpret_cast<long long*>(idx_out), dist_out);
  PSOB_LAUNCH_CHECK("vq_nearest_kernel");
  return 0;
}

int psob_pack_image_stem(const float* img, int64_t P, int32_t rescale, void* out_bf16, float* out32, psob_stream_t s) {
  if (!img || !out_bf16 || P <= 0 || ((uintptr_t)img & 15) || ((uintptr_t)out_bf16 & 15) || ((uintptr_t)out32 & 15))
    return fail(PSOB_E_ARG, "psob_pack_image_stem: bad arguments (pointers must be 16-byte aligned)");
  pack_image_stem_kernel<<<(unsigned)((P + 7) / 8), 256, 0, (cudaStream_t)s>>>(img, P, rescale, reinterpret_cast<__nv_bfloat16*>(out_bf16), out32);
  PSOB_LAUNCH_CHECK("pack_image_stem_kernel");
  return 0;
}

int psob_maxpool_nhwc(const void* x, int64_t x_pitch, int64_t P, int32_t h, int32_t w, int32_t cp, int32_t k, int32_t stride, int32_t pad,
                      void* y, int64_t y_pitch, int32_t y_tiled, psob_stream_t s) {
  if (!x || !y || P <= 0 || h <= 0 || w <= 0 || cp <= 0 || cp % 8 || k <= 0 || stride <= 0 || pad < 0 || 2 * pad > k)
    return fail(PSOB_E_ARG, "psob_maxpool_nhwc: bad arguments");
  const int ho = (h + 2 * pad - k) / stride + 1, wo = (w + 2 * pad - k) / stride + 1;
  if (ho <= 0 || wo <= 0 || x_pitch < (int64_t)h * w * cp || y_pitch < (int64_t)ho * wo * cp || x_pitch % 8 || y_pitch % 8 ||
      ((uintptr_t)x & 15) || ((uintptr_t)y & 15) || (y_tiled && (y_pitch % 64 || cp % 64)))
    return fail(PSOB_E_ARG, "psob_maxpool_nhwc: pitches must cover the maps and keep 16-byte alignment");
  if (k == 3 && stride == 2 && pad == 1 && w == 14) {
    const long long rows = (long long)P * ho * (cp / 8);
    maxpool3x3s2_rows_kernel<14><<<(unsigned)((rows + 255) / 256), 256, 0, (cudaStream_t)s>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), P, h, cp, ho, x_pitch, y_pitch, y_tiled, reinterpret_cast<__nv_bfloat16*>(y));
    PSOB_LAUNCH_CHECK("maxpool3x3s2_rows_kernel");
    return 0;
  }
  const long long total = (long long)P * ho * wo * (cp / 8);
  maxpool_nhwc_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)s>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), P, h, w, cp, k, stride, pad, ho, wo, x_pitch, y_pitch, y_tiled, reinterpret_cast<__nv_bfloat16*>(y));
  PSOB_LAUNCH_CHECK("maxpool_nhwc_kernel");
  return 0;
}

}  // extern "C"
