// Micro-benchmark: sustained tcgen05.mma rate per instruction shape on B200, all SMs busy, operands resident in shared
// memory (SWIZZLE_128B K-major tiles, contents irrelevant), no loads, no epilogue.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o build/umma_bench tools/umma_bench.cu
// One elected thread per CTA (pair: of the leader CTA) issues `reps` K=64 ops (4 x K=16 instructions + commit to a
// barrier, exactly the engine's issue block) round-robin over `n_b` weight tiles and 2 accumulator buffers, then waits
// for the last commit.  Reports cycles per K=16 instruction and the implied MAC / clock / SM.
#include "../gan-discovery-pso_b200/csrc/umma.cuh"
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

using namespace psob;

template <bool kPair>
__global__ void __launch_bounds__(128, 1) umma_rate_kernel(int n, int reps, int n_b, int per_commit, uint32_t idesc, int issuers, unsigned long long* cycles, int coloff = 0) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t a_addr = base;                                  // one A tile [128 x 64] bf16 = 16 KB
  const uint32_t b_addr = base + 16384u;                         // n_b B tiles of [n (pair: n/2) x 64]
  const uint32_t b_bytes = (uint32_t)(kPair ? n / 2 : n) * 128u;
  const uint32_t bar = b_addr + (uint32_t)n_b * b_bytes;         // 8-byte aligned (tiles are multiples of 1 KB)
  const uint32_t slot = bar + 16u;
  const int warp = threadIdx.x >> 5;
  const uint32_t crank = kPair ? ptx::cluster_ctarank() : 0u;
  const uint32_t n_commits = (uint32_t)((reps + per_commit - 1) / per_commit) * (uint32_t)issuers;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, n_commits); ptx::fence_barrier_init(); }     // ONE phase: completes after the last commit
  for (uint32_t i = threadIdx.x; i < (16384u + (uint32_t)n_b * b_bytes) / 4u; i += 128) reinterpret_cast<uint32_t*>(smem_raw + (base - raw))[i] = 0x3c003c00u;
  ptx::fence_proxy_async();
  if (warp == 1) {
    if (kPair) { ptx::tmem_alloc_pair(slot, 512); ptx::tmem_relinquish_pair(); } else { ptx::tmem_alloc(slot, 512); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  if (kPair) ptx::cluster_sync_all(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (slot - raw));
  if ((warp == 0 || (warp == 2 && issuers == 2)) && crank == 0) {      // two issuers: warp 0 -> TMEM buffer 0, warp 2 -> buffer 1
    const bool lane0 = (threadIdx.x & 31) == 0;
    const uint64_t adesc = ptx::smem_desc_sw128(a_addr);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (ptx::elect_one()) {
        const uint64_t bdesc = ptx::smem_desc_sw128(b_addr + (uint32_t)(r % n_b) * b_bytes);
        const uint32_t d = tmem_base + (issuers == 2 ? (uint32_t)(warp >> 1) : (uint32_t)(r & 1)) * 256u + (uint32_t)(((r >> 1) % 5) * coloff);   // coloff: accumulator windows at arbitrary column offsets (image band)
        const bool commit = (r % per_commit) == per_commit - 1 || r == reps - 1;
        if (commit) {
          if (kPair) ptx::umma_bf16_k64_commit_pair(d, adesc, bdesc, idesc, 1u, bar);
          else ptx::umma_bf16_k64_commit(d, adesc, bdesc, idesc, 1u, bar);
        } else {
          if (kPair) ptx::umma_bf16_k64_pair(d, adesc, bdesc, idesc, 1u);
          else ptx::umma_bf16_k64(d, adesc, bdesc, idesc, 1u);
        }
      }
      __syncwarp();
    }
    const long long t_issue = clock64();
    ptx::mbar_wait(bar, 0u);
    const long long t1 = clock64();
    if (warp == 0 && lane0) { cycles[2 * blockIdx.x] = (unsigned long long)(t1 - t0); cycles[2 * blockIdx.x + 1] = (unsigned long long)(t_issue - t0); }
  }
  ptx::tc_fence_before();
  if (kPair) ptx::cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    if (kPair) ptx::tmem_dealloc_pair(tmem_base, 512); else ptx::tmem_dealloc(tmem_base, 512);
  }
}

int main(int argc, char** argv) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  unsigned long long* d_cyc;
  cudaMalloc(&d_cyc, 2 * sms * sizeof(unsigned long long));
  const int reps = 4000;
  printf("%d SMs, %d K=64 ops per CTA (4 x tcgen05.mma K=16 each)\n", sms, reps);
  printf("%-6s %4s %4s %10s | %12s %12s %14s | %12s\n", "mode", "N", "n_b", "ops/commit", "cyc/instr", "issue-only", "MAC/clk/SM", "% of 4096");
  for (int issuers = 1; issuers <= 2; ++issuers)
  for (int pair = 0; pair < 2; ++pair)
    for (int n : {64, 96, 128, 256})
      for (int per_commit : {1, 4}) {
        if (issuers == 2 && per_commit == 1) continue;
        const int n_b = 4;
        if (pair && n % 32) continue;
        const size_t smem = 1024 + 16384 + (size_t)n_b * (pair ? n / 2 : n) * 128 + 64;
        int grid = pair ? (sms / 2) * 2 : sms;
        cudaMemset(d_cyc, 0, 2 * sms * sizeof(unsigned long long));
        cudaError_t e;
        if (pair) {
          cudaFuncSetAttribute(umma_rate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
          cudaLaunchConfig_t cfg = {};
          cfg.gridDim = dim3(grid); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
          cfg.attrs = at; cfg.numAttrs = 1;
          e = cudaLaunchKernelEx(&cfg, umma_rate_kernel<true>, n, reps, n_b, per_commit, make_idesc_bf16(256, n), issuers, d_cyc, 0);
        } else {
          cudaFuncSetAttribute(umma_rate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
          umma_rate_kernel<false><<<grid, 128, smem>>>(n, reps, n_b, per_commit, make_idesc_bf16(128, n), issuers, d_cyc);
          e = cudaGetLastError();
        }
        cudaError_t e2 = cudaDeviceSynchronize();
        if (e != cudaSuccess || e2 != cudaSuccess) { printf("launch failed: %s / %s\n", cudaGetErrorString(e), cudaGetErrorString(e2)); return 1; }
        unsigned long long h[2 * 160];
        cudaMemcpy(h, d_cyc, 2 * sms * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        double tot = 0, iss = 0; int cnt = 0;
        for (int b = 0; b < grid; ++b) if (h[2 * b]) { tot += (double)h[2 * b]; iss += (double)h[2 * b + 1]; ++cnt; }
        tot /= cnt; iss /= cnt;
        const double per_instr = tot / (reps * 4.0 * issuers);
        const double macs = 128.0 * n * 16.0;           // per SM per instruction (pair: each SM does 128 rows x n)
        printf("%-6s %4d %4d %10d | %12.1f %12.1f %14.0f | %11.1f%%\n", pair ? (issuers == 2 ? "pair x2" : "pair") : (issuers == 2 ? "solo x2" : "solo"), n, n_b,
               per_commit, per_instr, iss / (reps * 4.0 * issuers), macs / per_instr, 100.0 * macs / per_instr / 4096.0);
      }
  // accumulator windows that start at arbitrary (even) columns, as the image-band layers use them: N = 96, solo
  printf("solo N = 96, 4 ops per commit, accumulator window start = buffer + ((op / 2) %% 5) * coloff\n");
  for (int coloff : {0, 2, 8, 16, 30, 32}) {
    const int n = 96, n_b = 4;
    const size_t smem = 1024 + 16384 + (size_t)n_b * n * 128 + 64;
    cudaMemset(d_cyc, 0, 2 * sms * sizeof(unsigned long long));
    cudaFuncSetAttribute(umma_rate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    umma_rate_kernel<false><<<sms, 128, smem>>>(n, reps, n_b, 4, make_idesc_bf16(128, n), 1, d_cyc, coloff);
    cudaError_t e = cudaGetLastError(), e2 = cudaDeviceSynchronize();
    if (e != cudaSuccess || e2 != cudaSuccess) { printf("launch failed: %s / %s\n", cudaGetErrorString(e), cudaGetErrorString(e2)); return 1; }
    unsigned long long h[2 * 160];
    cudaMemcpy(h, d_cyc, 2 * sms * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    double tot = 0; int cnt = 0;
    for (int b = 0; b < sms; ++b) if (h[2 * b]) { tot += (double)h[2 * b]; ++cnt; }
    printf("coloff %3d | %8.1f cycles per instruction\n", coloff, tot / cnt / (reps * 4.0));
  }
  return 0;
}
