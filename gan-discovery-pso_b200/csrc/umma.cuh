// Inline-PTX primitives for sm_100a: mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma /
// commit / ld / fences) and the UMMA shared-memory + instruction descriptors used by the
// pixel-GEMM engine.  Encodings follow the PTX ISA 8.6 tcgen05 chapter; field layouts were
// cross-checked against the CUTLASS headers shipped in the image (cute/arch/mma_sm100_desc.hpp)
// — nothing from CUTLASS is compiled into this library.
#pragma once
#include <cstdint>
#include <cuda.h>

namespace psob {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// One elected lane of a CONVERGED warp.  Unlike `if (lane == 0)`, this tells ptxas the warp is converged,
// so operands of the single-thread instructions (UTCHMMA / UTMALDG / UTCBAR) are moved to uniform
// registers with plain R2UR instead of a per-instruction ELECT/R2UR/BRA.U.ANY waterfall loop.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_relaxed(uint32_t bar) {      // pure signal, see mbar_arrive_cluster_relaxed
  asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded spin: a protocol bug traps (cudaErrorLaunchFailure) instead of hanging the GPU.  try_wait
// suspends the thread for a HW-defined interval, so 2^28 failed polls is many seconds.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins == (1u << 28)) __trap();
  }
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load: coordinates (c0 = innermost element index, c1 = row)
__device__ __forceinline__ void tma_load_2d(uint32_t dst_smem, const CUtensorMap* m, int32_t c0, int32_t c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

// L2 prefetch of a 2-D tile (no shared memory, no barrier): the tile is in L2 by the time its real TMA load is issued
__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* m, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1)
               : "memory");
}

// 2-D tiled store shared -> global (bulk async group); rows/columns outside the tensor are clipped
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, uint32_t src_smem, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(src_smem), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read2() { asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// ---------------------------------------------------------------- CTA pair (cluster of 2)
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank`
__device__ __forceinline__ uint32_t map_to_cta(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// Signal only (no data is published through memory): e.g. "my tcgen05.ld of this accumulator have completed"
// (tcgen05.wait::ld has already returned).  The release form costs a cluster-scope fence per arrive.
__device__ __forceinline__ void mbar_arrive_cluster_relaxed(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait_cluster(bar, parity)) {
    if (++spins == (1u << 28)) __trap();
  }
}
// TMA load issued by either CTA of a pair: data lands in the issuing CTA's shared memory, the bytes are
// counted on `bar_cluster_addr`, which may live in the peer (the pair leader's full barrier).
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst_smem, const CUtensorMap* m, int32_t c0, int32_t c1, uint32_t bar_cluster_addr) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}

// No "memory" clobber on purpose: the bias loads of the next 16-byte slot must be free to move above this
// store (with the clobber every slot serialised LDS -> math -> STS, ~110 cycles each).  The staging tile is only
// read by the TMA store, which is ordered by the fence.proxy.async + bar.sync (both clobber memory) after the loop.
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w));
}

// volatile (ordered against the other volatile asm of the epilogue) but no "memory" clobber: schedulable
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// ---------------------------------------------------------------- tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 x bf16 -> fp32, issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// One K-chunk (64 bf16 = 4 MMAs of K=16) plus the commit that frees the ring stage, as ONE asm block:
// the operands are not provably warp-uniform for nvcc (they come from the step table), so every
// separate tcgen05 statement costs an ELECT/R2UR "waterfall" (~100+ cycles of scalar latency, measured
// ~140 cycles per MMA regardless of N).  Inside one block the +32-byte descriptor advances stay in the
// uniform datapath and the waterfall is paid once per step.
__device__ __forceinline__ void umma_bf16_k64_commit(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                     uint32_t accumulate, uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred p, t;\n\t.reg .b64 a1, b1, a2, b2, a3, b3;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.u32 t, 0, 0;\n\t"
      "add.u64 a1, %1, 2;\n\tadd.u64 b1, %2, 2;\n\t"
      "add.u64 a2, %1, 4;\n\tadd.u64 b2, %2, 4;\n\t"
      "add.u64 a3, %1, 6;\n\tadd.u64 b3, %2, 6;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %3, t;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a2, b2, %3, t;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a3, b3, %3, t;\n\t"
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%5];\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(bar)
      : "memory");
}
// The same K-chunk without the commit (steps that feed several weight tiles commit once after the last one).
#define PSOB_UMMA_K64_BODY(GROUP)                                                                        \
  asm volatile(                                                                                          \
      "{\n\t.reg .pred p, t;\n\t.reg .b64 a1, b1, a2, b2, a3, b3;\n\t"                                   \
      "setp.ne.b32 p, %4, 0;\n\t"                                                                        \
      "setp.eq.u32 t, 0, 0;\n\t"                                                                         \
      "add.u64 a1, %1, 2;\n\tadd.u64 b1, %2, 2;\n\t"                                                    \
      "add.u64 a2, %1, 4;\n\tadd.u64 b2, %2, 4;\n\t"                                                    \
      "add.u64 a3, %1, 6;\n\tadd.u64 b3, %2, 6;\n\t"                                                    \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], %1, %2, %3, p;\n\t"                               \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a1, b1, %3, t;\n\t"                               \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a2, b2, %3, t;\n\t"                               \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a3, b3, %3, t;\n\t}"                              \
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)                               \
      : "memory")
__device__ __forceinline__ void umma_bf16_k64(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  PSOB_UMMA_K64_BODY("1");
}
__device__ __forceinline__ void umma_bf16_k64_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  PSOB_UMMA_K64_BODY("2");
}
#undef PSOB_UMMA_K64_BODY
// One A tile against up to four weight tiles (ops 1..3 enabled by en1..en3) and the stage-freeing commit, as ONE
// asm block: each separate tcgen05 statement costs a uniform-register "waterfall", which made multi-op steps
// issue-bound when every op was its own block.
#define PSOB_UMMA_K64X4_BODY(GROUP, COMMIT)                                                              \
  asm volatile(                                                                                          \
      "{\n\t.reg .pred p0, p1, p2, p3, e1, e2, e3, t;\n\t.reg .b64 a1, a2, a3, x1, x2, x3;\n\t"           \
      "setp.ne.b32 p0, %10, 0;\n\tsetp.ne.b32 p1, %11, 0;\n\tsetp.ne.b32 p2, %12, 0;\n\tsetp.ne.b32 p3, %13, 0;\n\t" \
      "setp.ne.b32 e1, %14, 0;\n\tsetp.ne.b32 e2, %15, 0;\n\tsetp.ne.b32 e3, %16, 0;\n\t"                \
      "setp.eq.u32 t, 0, 0;\n\t"                                                                         \
      "add.u64 a1, %4, 2;\n\tadd.u64 a2, %4, 4;\n\tadd.u64 a3, %4, 6;\n\t"                              \
      "add.u64 x1, %5, 2;\n\tadd.u64 x2, %5, 4;\n\tadd.u64 x3, %5, 6;\n\t"                              \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], %4, %5, %9, p0;\n\t"                              \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a1, x1, %9, t;\n\t"                               \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a2, x2, %9, t;\n\t"                               \
      "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], a3, x3, %9, t;\n\t"                               \
      "add.u64 x1, %6, 2;\n\tadd.u64 x2, %6, 4;\n\tadd.u64 x3, %6, 6;\n\t"                              \
      "@e1 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%1], %4, %6, %9, p1;\n\t"                          \
      "@e1 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%1], a1, x1, %9, t;\n\t"                           \
      "@e1 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%1], a2, x2, %9, t;\n\t"                           \
      "@e1 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%1], a3, x3, %9, t;\n\t"                           \
      "add.u64 x1, %7, 2;\n\tadd.u64 x2, %7, 4;\n\tadd.u64 x3, %7, 6;\n\t"                              \
      "@e2 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%2], %4, %7, %9, p2;\n\t"                          \
      "@e2 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%2], a1, x1, %9, t;\n\t"                           \
      "@e2 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%2], a2, x2, %9, t;\n\t"                           \
      "@e2 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%2], a3, x3, %9, t;\n\t"                           \
      "add.u64 x1, %8, 2;\n\tadd.u64 x2, %8, 4;\n\tadd.u64 x3, %8, 6;\n\t"                              \
      "@e3 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%3], %4, %8, %9, p3;\n\t"                          \
      "@e3 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%3], a1, x1, %9, t;\n\t"                           \
      "@e3 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%3], a2, x2, %9, t;\n\t"                           \
      "@e3 tcgen05.mma.cta_group::" GROUP ".kind::f16 [%3], a3, x3, %9, t;\n\t"                           \
      COMMIT "\n\t}"                                                                                     \
      ::"r"(d[0]), "r"(d[1]), "r"(d[2]), "r"(d[3]), "l"(a_desc), "l"(b[0]), "l"(b[1]), "l"(b[2]), "l"(b[3]),   \
        "r"(idesc), "r"(acc[0]), "r"(acc[1]), "r"(acc[2]), "r"(acc[3]), "r"((uint32_t)(n_ops > 1)),          \
        "r"((uint32_t)(n_ops > 2)), "r"((uint32_t)(n_ops > 3)), "r"(bar), "h"((uint16_t)3)                  \
      : "memory")
__device__ __forceinline__ void umma_bf16_k64x4_commit(const uint32_t (&d)[4], uint64_t a_desc, const uint64_t (&b)[4], uint32_t idesc,
                                                       const uint32_t (&acc)[4], int n_ops, uint32_t bar) {
  PSOB_UMMA_K64X4_BODY("1", "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%17];");
}
// The same with one instruction descriptor PER OP (ops of different widths N share an A tile; solo CTAs).
__device__ __forceinline__ void umma_bf16_k64x4w_commit(const uint32_t (&d)[4], uint64_t a_desc, const uint64_t (&b)[4], const uint32_t (&idesc)[4],
                                                        const uint32_t (&acc)[4], int n_ops, uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred p0, p1, p2, p3, e1, e2, e3, t;\n\t.reg .b64 a1, a2, a3, x1, x2, x3;\n\t"
      "setp.ne.b32 p0, %10, 0;\n\tsetp.ne.b32 p1, %11, 0;\n\tsetp.ne.b32 p2, %12, 0;\n\tsetp.ne.b32 p3, %13, 0;\n\t"
      "setp.ne.b32 e1, %14, 0;\n\tsetp.ne.b32 e2, %15, 0;\n\tsetp.ne.b32 e3, %16, 0;\n\t"
      "setp.eq.u32 t, 0, 0;\n\t"
      "add.u64 a1, %4, 2;\n\tadd.u64 a2, %4, 4;\n\tadd.u64 a3, %4, 6;\n\t"
      "add.u64 x1, %5, 2;\n\tadd.u64 x2, %5, 4;\n\tadd.u64 x3, %5, 6;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %4, %5, %9, p0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, x1, %9, t;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a2, x2, %9, t;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a3, x3, %9, t;\n\t"
      "add.u64 x1, %6, 2;\n\tadd.u64 x2, %6, 4;\n\tadd.u64 x3, %6, 6;\n\t"
      "@e1 tcgen05.mma.cta_group::1.kind::f16 [%1], %4, %6, %18, p1;\n\t"
      "@e1 tcgen05.mma.cta_group::1.kind::f16 [%1], a1, x1, %18, t;\n\t"
      "@e1 tcgen05.mma.cta_group::1.kind::f16 [%1], a2, x2, %18, t;\n\t"
      "@e1 tcgen05.mma.cta_group::1.kind::f16 [%1], a3, x3, %18, t;\n\t"
      "add.u64 x1, %7, 2;\n\tadd.u64 x2, %7, 4;\n\tadd.u64 x3, %7, 6;\n\t"
      "@e2 tcgen05.mma.cta_group::1.kind::f16 [%2], %4, %7, %19, p2;\n\t"
      "@e2 tcgen05.mma.cta_group::1.kind::f16 [%2], a1, x1, %19, t;\n\t"
      "@e2 tcgen05.mma.cta_group::1.kind::f16 [%2], a2, x2, %19, t;\n\t"
      "@e2 tcgen05.mma.cta_group::1.kind::f16 [%2], a3, x3, %19, t;\n\t"
      "add.u64 x1, %8, 2;\n\tadd.u64 x2, %8, 4;\n\tadd.u64 x3, %8, 6;\n\t"
      "@e3 tcgen05.mma.cta_group::1.kind::f16 [%3], %4, %8, %20, p3;\n\t"
      "@e3 tcgen05.mma.cta_group::1.kind::f16 [%3], a1, x1, %20, t;\n\t"
      "@e3 tcgen05.mma.cta_group::1.kind::f16 [%3], a2, x2, %20, t;\n\t"
      "@e3 tcgen05.mma.cta_group::1.kind::f16 [%3], a3, x3, %20, t;\n\t"
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%17];\n\t}"
      ::"r"(d[0]), "r"(d[1]), "r"(d[2]), "r"(d[3]), "l"(a_desc), "l"(b[0]), "l"(b[1]), "l"(b[2]), "l"(b[3]),
        "r"(idesc[0]), "r"(acc[0]), "r"(acc[1]), "r"(acc[2]), "r"(acc[3]), "r"((uint32_t)(n_ops > 1)),
        "r"((uint32_t)(n_ops > 2)), "r"((uint32_t)(n_ops > 3)), "r"(bar), "r"(idesc[1]), "r"(idesc[2]), "r"(idesc[3])
      : "memory");
}
__device__ __forceinline__ void umma_bf16_k64x4_commit_pair(const uint32_t (&d)[4], uint64_t a_desc, const uint64_t (&b)[4], uint32_t idesc,
                                                            const uint32_t (&acc)[4], int n_ops, uint32_t bar) {
  PSOB_UMMA_K64X4_BODY("2", "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%17], %18;");
}
#undef PSOB_UMMA_K64X4_BODY
// CTA-pair form (issued by the pair leader only): M = 256 over both CTAs' A tiles, each CTA supplies half
// of the B rows from the same shared-memory offset; the commit arrives on `bar` in BOTH CTAs.
__device__ __forceinline__ void umma_bf16_k64_commit_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                          uint32_t accumulate, uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred p, t;\n\t.reg .b64 a1, b1, a2, b2, a3, b3;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.u32 t, 0, 0;\n\t"
      "add.u64 a1, %1, 2;\n\tadd.u64 b1, %2, 2;\n\t"
      "add.u64 a2, %1, 4;\n\tadd.u64 b2, %2, 4;\n\t"
      "add.u64 a3, %1, 6;\n\tadd.u64 b3, %2, 6;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a1, b1, %3, t;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a2, b2, %3, t;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a3, b3, %3, t;\n\t"
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%5], %6;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(bar), "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}
// arrive on an mbarrier when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns -> 16 registers per thread (lane i reads TMEM lane base+i)
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------- descriptors
// K-major operand tile, 64 bf16 (128 B) per row, TMA SWIZZLE_128B: 8-row groups of 1024 B.
//   bits [0,14)  start address >> 4        bits [16,30) leading byte offset >> 4 (unused for SW128 K-major)
//   bits [32,46) stride byte offset >> 4   bits [46,48) descriptor version (1 on sm_100)
//   bits [61,64) layout: 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

}  // namespace ptx

// kind::f16 instruction descriptor: D fp32 (bit 4), A/B bf16 (bits 7, 10), both K-major (bits 15/16 = 0),
// N>>3 at [17,23), M>>4 at [24,29).
inline uint32_t make_idesc_bf16(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

}  // namespace psob
