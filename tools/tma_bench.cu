// Micro-benchmark: per-SM TMA load / store throughput for the engine's operand tiles on B200.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tma_bench tools/tma_bench.cu -lcuda
// One CTA per SM.  Thread 0 issues TMA loads of a [box_rows x 64] bf16 box (SWIZZLE_128B) into an S-stage ring;
// thread 32 consumes (waits the full barrier, frees the stage).  Reports bytes / clock / SM.
#include "../gan-discovery-pso_b200/csrc/umma.cuh"
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace psob;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn enc;

static CUtensorMap make_map(void* ptr, uint64_t cols, uint64_t rows, uint64_t pitch_elems, uint32_t box_rows) {
  CUtensorMap m;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch_elems * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

// mode 0: every load reads the same box (pure TMA / L2-hit throughput)
// mode 1: CTA b streams through its own region: box k at row (b * n_loads + k) * box_rows   (DRAM stream)
// mode 2: like the engine, row-major activation: box k at column (k % n_chunks) * 64 of the CTA's 128 rows (scattered rows)
__global__ void __launch_bounds__(64, 1)
load_kernel(const __grid_constant__ CUtensorMap tm, int n_loads, int stages, int box_rows, int mode, int n_chunks,
            unsigned long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t stage_bytes = (uint32_t)box_rows * 128u;
  const uint32_t bars = base + (uint32_t)stages * stage_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { ptx::mbar_init(bars + 8u * s, 1); ptx::mbar_init(bars + 8u * (32 + s), 1); }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int k = 0; k < n_loads; ++k) {
      ptx::mbar_wait(bars + 8u * (32 + stage), phase ^ 1u);
      ptx::mbar_arrive_expect_tx(bars + 8u * stage, stage_bytes);
      int c0 = 0, c1 = 0;
      if (mode == 1) c1 = (int)(((long long)(blockIdx.x * n_loads + k) * box_rows) % ((1 << 25) - 256));
      if (mode == 2) { c0 = (k % n_chunks) * 64; c1 = ((blockIdx.x + (k / n_chunks) * gridDim.x) * box_rows) % (300000 - 256); }
      ptx::tma_load_2d(base + (uint32_t)stage * stage_bytes, &tm, c0, c1, bars + 8u * stage);
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (threadIdx.x == 32) {
    int stage = 0; uint32_t phase = 0;
    for (int k = 0; k < n_loads; ++k) {
      ptx::mbar_wait(bars + 8u * stage, phase);
      ptx::mbar_arrive(bars + 8u * (32 + stage));
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
    cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
}

// TMA stores of a [box_rows x 64] tile from shared memory, `bufs` staging tiles in rotation
__global__ void __launch_bounds__(32, 1)
store_kernel(const __grid_constant__ CUtensorMap tm, int n_stores, int bufs, int box_rows, int mode, int n_chunks,
             unsigned long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t tile_bytes = (uint32_t)box_rows * 128u;
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    for (int k = 0; k < n_stores; ++k) {
      if (bufs == 1) ptx::bulk_wait_read0(); else ptx::bulk_wait_read1();
      int c0 = 0, c1 = 0;
      if (mode == 1) c1 = (int)(((long long)(blockIdx.x * n_stores + k) * box_rows) % ((1 << 25) - 256));
      if (mode == 2) { c0 = (k % n_chunks) * 64; c1 = ((blockIdx.x + (k / n_chunks) * gridDim.x) * box_rows) % (300000 - 256); }
      ptx::tma_store_2d(&tm, base + (uint32_t)(k % bufs) * tile_bytes, c0, c1);
      ptx::bulk_commit();
    }
    ptx::bulk_wait_all0();
    cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
}

// loads (threads 0 / 32, as load_kernel) and DRAM-stream stores (thread 64) running concurrently in one CTA
__global__ void __launch_bounds__(96, 1)
mixed_kernel(const __grid_constant__ CUtensorMap tl, const __grid_constant__ CUtensorMap ts, int n_loads, int n_stores, int stages,
             int load_mode, unsigned long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t stage_bytes = 128u * 128u;
  const uint32_t bars = base + (uint32_t)(stages + 2) * stage_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { ptx::mbar_init(bars + 8u * s, 1); ptx::mbar_init(bars + 8u * (32 + s), 1); }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int k = 0; k < n_loads; ++k) {
      ptx::mbar_wait(bars + 8u * (32 + stage), phase ^ 1u);
      ptx::mbar_arrive_expect_tx(bars + 8u * stage, stage_bytes);
      int c1 = load_mode == 1 ? (int)(((long long)(blockIdx.x * n_loads + k) * 128) % ((1 << 24) - 256)) : 0;
      ptx::tma_load_2d(base + (uint32_t)stage * stage_bytes, &tl, 0, c1, bars + 8u * stage);
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (threadIdx.x == 32) {
    int stage = 0; uint32_t phase = 0;
    for (int k = 0; k < n_loads; ++k) {
      ptx::mbar_wait(bars + 8u * stage, phase);
      ptx::mbar_arrive(bars + 8u * (32 + stage));
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
    cycles[2 * blockIdx.x] = (unsigned long long)(clock64() - t0);
  } else if (threadIdx.x == 64) {
    for (int k = 0; k < n_stores; ++k) {
      ptx::bulk_wait_read1();
      int c1 = (1 << 24) + (int)(((long long)(blockIdx.x * n_stores + k) * 128) % ((1 << 24) - 256));
      ptx::tma_store_2d(&ts, base + (uint32_t)(stages + (k & 1)) * stage_bytes, 0, c1);
      ptx::bulk_commit();
    }
    ptx::bulk_wait_all0();
    cycles[2 * blockIdx.x + 1] = (unsigned long long)(clock64() - t0);
  }
}

// 1-D bulk copies (no tensor map): contiguous `bytes` per op.  mode 0: loads through an S-stage ring, mode 1: stores, 2 buffers
__global__ void __launch_bounds__(64, 1)
bulk1d_kernel(const uint8_t* g, int n_ops, int stages, int bytes, int mode, int same, unsigned long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + (uint32_t)stages * (uint32_t)bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { ptx::mbar_init(bars + 8u * s, 1); ptx::mbar_init(bars + 8u * (32 + s), 1); }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  const unsigned long long span = (3ull << 30);
  if (mode == 0) {
    if (threadIdx.x == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int k = 0; k < n_ops; ++k) {
        ptx::mbar_wait(bars + 8u * (32 + stage), phase ^ 1u);
        ptx::mbar_arrive_expect_tx(bars + 8u * stage, (uint32_t)bytes);
        const uint8_t* src = g + (same ? 0ull : ((unsigned long long)(blockIdx.x * n_ops + k) * (unsigned long long)bytes) % span);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(base + (uint32_t)stage * (uint32_t)bytes), "l"(src), "r"((uint32_t)bytes), "r"(bars + 8u * stage) : "memory");
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    } else if (threadIdx.x == 32) {
      int stage = 0; uint32_t phase = 0;
      for (int k = 0; k < n_ops; ++k) {
        ptx::mbar_wait(bars + 8u * stage, phase);
        ptx::mbar_arrive(bars + 8u * (32 + stage));
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
      cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
    }
  } else if (threadIdx.x == 0) {
    for (int k = 0; k < n_ops; ++k) {
      ptx::bulk_wait_read1();
      uint8_t* dst = const_cast<uint8_t*>(g) + ((unsigned long long)(blockIdx.x * n_ops + k) * (unsigned long long)bytes) % span;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                   ::"l"(dst), "r"(base + (uint32_t)(k & 1) * (uint32_t)bytes), "r"((uint32_t)bytes) : "memory");
      ptx::bulk_commit();
    }
    ptx::bulk_wait_all0();
    cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
}

int main() {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q);
  enc = (EncodeTiledFn)sym;
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const size_t bytes = 4ull << 30;
  void* buf; cudaMalloc(&buf, bytes); cudaMemset(buf, 0, bytes);
  unsigned long long* cyc; cudaMalloc(&cyc, sizeof(unsigned long long) * sms);
  std::vector<unsigned long long> h(sms);
  cudaFuncSetAttribute(load_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  cudaFuncSetAttribute(store_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  const int n = 2000;
  auto report = [&](const char* what, int box_rows) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: CUDA error %s\n", what, cudaGetErrorString(e)); exit(1); }
    cudaMemcpy(h.data(), cyc, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost);
    double mean = 0; for (auto v : h) mean += (double)v; mean /= sms;
    printf("%-78s %7.1f cycles/op  %6.1f B/clk/SM\n", what, mean / n, (double)n * box_rows * 128 / mean);
  };
  char name[160];
  for (int box_rows : {128, 256}) {
    for (int stages : {3, 6}) {
      if (stages * box_rows * 128 > 200 * 1024) continue;
      const int smem = stages * box_rows * 128 + 2048;
      // mode 0: same box
      CUtensorMap t0 = make_map(buf, 64, 1 << 25, 64, box_rows);
      load_kernel<<<sms, 64, smem>>>(t0, n, stages, box_rows, 0, 1, cyc);
      snprintf(name, sizeof name, "load  box %3dx128B stages %d  same box (L2 hit), contiguous rows", box_rows, stages); report(name, box_rows);
      // mode 1: DRAM stream, contiguous rows (tile-major)
      load_kernel<<<sms, 64, smem>>>(t0, n, stages, box_rows, 1, 1, cyc);
      snprintf(name, sizeof name, "load  box %3dx128B stages %d  DRAM stream, contiguous rows (tile-major)", box_rows, stages); report(name, box_rows);
      // mode 2: row-major activations, pitch 12544 B (49 px x 128 ch), 98 chunks per row
      CUtensorMap t2 = make_map(buf, 6272, 300000, 6272, box_rows);
      load_kernel<<<sms, 64, smem>>>(t2, n, stages, box_rows, 2, 98, cyc);
      snprintf(name, sizeof name, "load  box %3dx128B stages %d  DRAM stream, rows 12544 B apart (row-major)", box_rows, stages); report(name, box_rows);
    }
  }
  for (int box_rows : {32, 128}) {
    for (int bufs : {1, 2}) {
      const int smem = bufs * box_rows * 128 + 2048;
      CUtensorMap t0 = make_map(buf, 64, 1 << 25, 64, box_rows);
      store_kernel<<<sms, 32, smem>>>(t0, n, bufs, box_rows, 0, 1, cyc);
      snprintf(name, sizeof name, "store box %3dx128B bufs %d    same box, contiguous rows", box_rows, bufs); report(name, box_rows);
      store_kernel<<<sms, 32, smem>>>(t0, n, bufs, box_rows, 1, 1, cyc);
      snprintf(name, sizeof name, "store box %3dx128B bufs %d    DRAM stream, contiguous rows (tile-major)", box_rows, bufs); report(name, box_rows);
      CUtensorMap t2 = make_map(buf, 6272, 300000, 6272, box_rows);
      store_kernel<<<sms, 32, smem>>>(t2, n, bufs, box_rows, 2, 98, cyc);
      snprintf(name, sizeof name, "store box %3dx128B bufs %d    DRAM stream, rows 12544 B apart (row-major)", box_rows, bufs); report(name, box_rows);
    }
  }
  cudaFuncSetAttribute(bulk1d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  for (int g : {19, 148}) {
    for (int kb : {16, 32, 64}) {
      const int bytes1 = kb * 1024, stages = kb == 64 ? 3 : 4;
      auto run = [&](int mode, int same, const char* what) {
        bulk1d_kernel<<<g, 64, stages * bytes1 + 2048>>>((const uint8_t*)buf, n, stages, bytes1, mode, same, cyc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("bulk1d: %s\n", cudaGetErrorString(e)); exit(1); }
        cudaMemcpy(h.data(), cyc, sizeof(unsigned long long) * g, cudaMemcpyDeviceToHost);
        double mean = 0; for (int i = 0; i < g; ++i) mean += (double)h[i]; mean /= g;
        printf("bulk1d %-28s %2d KB/op %3d CTAs: %7.1f cycles/op %6.1f B/clk/SM\n", what, kb, g, mean / n, (double)n * bytes1 / mean);
      };
      run(0, 1, "load, same block (L2 hit)");
      run(0, 0, "load, DRAM stream");
      run(1, 0, "store, DRAM stream");
    }
  }
  for (int g : {19, 148}) {
    CUtensorMap t0 = make_map(buf, 64, 1 << 25, 64, 256);
    store_kernel<<<g, 32, 2 * 32768 + 2048>>>(t0, n, 2, 256, 1, 1, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), cyc, sizeof(unsigned long long) * g, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < g; ++i) mean += (double)h[i]; mean /= g;
    printf("tensor store box 256x128B bufs 2, DRAM stream, %3d CTAs: %7.1f cycles/op %6.1f B/clk/SM\n", g, mean / n, (double)n * 32768 / mean);
  }
  // fewer SMs storing: is ~26 B/clk/SM a per-SM limit or the shared HBM limit?
  for (int g : {19, 37, 74, 148}) {
    CUtensorMap t0 = make_map(buf, 64, 1 << 25, 64, 128);
    store_kernel<<<g, 32, 2 * 16384 + 2048>>>(t0, n, 2, 128, 1, 1, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), cyc, sizeof(unsigned long long) * g, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < g; ++i) mean += (double)h[i]; mean /= g;
    printf("store box 128x128B bufs 2, DRAM stream, %3d CTAs: %7.1f cycles/op %6.1f B/clk/SM\n", g, mean / n, (double)n * 16384 / mean);
    load_kernel<<<g, 64, 6 * 16384 + 2048>>>(t0, n, 6, 128, 1, 1, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), cyc, sizeof(unsigned long long) * g, cudaMemcpyDeviceToHost);
    mean = 0; for (int i = 0; i < g; ++i) mean += (double)h[i]; mean /= g;
    printf("load  box 128x128B stages 6, DRAM stream, %3d CTAs: %7.1f cycles/op %6.1f B/clk/SM\n", g, mean / n, (double)n * 16384 / mean);
  }
  // loads and stores together
  cudaFuncSetAttribute(mixed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  unsigned long long* cyc2; cudaMalloc(&cyc2, sizeof(unsigned long long) * 2 * sms);
  std::vector<unsigned long long> h2(2 * sms);
  for (int lm : {0, 1}) {
    for (int ratio : {1, 2}) {
      CUtensorMap t0 = make_map(buf, 64, 1 << 25, 64, 128);
      mixed_kernel<<<sms, 96, 8 * 16384 + 2048>>>(t0, t0, n * ratio, n, 6, lm, cyc2);
      cudaDeviceSynchronize();
      cudaMemcpy(h2.data(), cyc2, sizeof(unsigned long long) * 2 * sms, cudaMemcpyDeviceToHost);
      double ml = 0, ms = 0; for (int i = 0; i < sms; ++i) { ml += (double)h2[2 * i]; ms += (double)h2[2 * i + 1]; } ml /= sms; ms /= sms;
      printf("mixed: %d loads (%s) per store: loads done after %8.0f cycles (%5.1f B/clk/SM), stores after %8.0f (%5.1f B/clk/SM)\n", ratio,
             lm ? "DRAM stream" : "L2 hit", ml, (double)n * ratio * 16384 / ml, ms, (double)n * 16384 / ms);
    }
  }
  return 0;
}
