// Microbenchmark: cost of one tcgen05.mma (kind::f16, SS) as a function of N, CTA group and A-row shift,
// with every SM busy.  Operands sit in shared memory (garbage values are fine), no TMA in the loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o mma_bench tools/mma_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../expecto_b200/csrc/ptx.cuh"
using namespace bx;

template <int CG>
__global__ void __launch_bounds__(128, 1) bench_kernel(int n, int m, int iters, int shift_rows, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t sA = base, sB = base + 32 * 1024, bar = base + 32 * 1024 + 64 * 1024, slot = bar + 16;
  // zero operands so accumulators stay finite
  for (uint32_t i = threadIdx.x; i < (96 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - raw))[i] = 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 1 ? 0 : cluster_ctarank();
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc<CG>(slot, 512);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  if (CG == 1) __syncthreads(); else cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem_raw + (slot - raw));
  if (warp == 0 && lane == 0 && rank == 0) {
    const uint32_t idesc = make_idesc_f16(0, m, n);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_f16<CG>(tmem, make_sw128_desc(sA + (shift_rows ? t * 128 : 0) + k * 32), make_sw128_desc(sB + k * 32), idesc, 1);
      }
    }
    umma_commit<CG>(bar);
    mbar_wait(bar, 0, 99);
    long long t1 = clock64();
    if (blockIdx.x == 0) cycles[0] = t1 - t0;
  }
  __syncwarp();
  tc_fence_before();
  if (CG == 1) __syncthreads(); else cluster_sync_all();
  if (warp == 1) { tc_fence_after(); tmem_dealloc<CG>(tmem, 512); }
}

template <int CG>
void run(int n, int m, int shift) {
  const int iters = 200;
  long long* d; cudaMalloc(&d, 8);
  auto k = bench_kernel<CG>;
  const int smem = 100 * 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(148); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaLaunchKernelEx(&cfg, k, n, m, iters, shift, d);  // warm-up
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  cudaLaunchKernelEx(&cfg, k, n, m, iters, shift, d);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
  long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
  const double instr = iters * 32.0;
  const double flops = 2.0 * m * n * 16 * instr * (148 / CG);
  printf("cg=%d M=%3d N=%3d shift=%d : %7.1f cycles/mma  (ideal M*N/(256*cg)=%5.1f)  %.0f TFLOP/s  %.3f ms  %s\n", CG, m, n, shift,
         c / instr, (double)(m < 128 * CG ? 128 * CG : m) * n / (256.0 * CG) / 1.0 / (CG == 2 ? 1 : 1), flops / ms / 1e9, ms, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  const int ns[] = {64, 96, 128, 160, 192, 208, 224, 240, 256};
  for (int n : ns) run<1>(n, 128, 0);
  for (int n : ns) run<1>(n, 128, 1);
  for (int n : ns) run<2>(n, 256, 0);
  for (int n : ns) run<2>(n, 256, 1);
  for (int n : {128, 256}) run<2>(n, 128, 0);
  for (int n : {64, 128, 256}) run<1>(n, 64, 0);
  return 0;
}
