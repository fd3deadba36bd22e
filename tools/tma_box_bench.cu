// Micro-benchmark behind score_tma_kernel's x staging (kernel 5): how fast one SM's TMA unit streams boxes
// [F features x R rows x 9 shifts] of fp32 out of a [9, n, pitch] tensor when the box is 36 KB and S boxes are in
// flight, for row widths of 32 / 64 / 128 bytes.  Every CTA walks its own rows through all 2002 features, as the
// scoring kernel does; nothing is computed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_box_bench tools/tma_box_bench.cu -lcuda
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(32, 1) walk(const __grid_constant__ CUtensorMap tm, int F, int R, int stages, int tiles,
                                              int box_bytes) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  const uint32_t bars = base + stages * box_bytes;
  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8 * i));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncwarp();
  if (threadIdx.x != 0) return;
  const int chunks = (2002 + F - 1) / F;
  int issued = 0, waited = 0;
  const int rows_per_tile = R;
  for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
    for (int ch = 0; ch < chunks; ++ch) {
      if (issued - waited == stages) {   // oldest box must have landed before its stage is reused
        const int s = waited % stages, ph = (waited / stages) & 1;
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                       : "=r"(ok) : "r"(bars + 8 * s), "r"(ph) : "memory");
        ++waited;
      }
      const int s = issued % stages;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bars + 8 * s), "r"(box_bytes) : "memory");
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                   ::"r"(base + s * box_bytes), "l"(reinterpret_cast<uint64_t>(&tm)), "r"(bars + 8 * s), "r"(ch * F),
                   "r"(t * rows_per_tile), "r"(0) : "memory");
      ++issued;
    }
  }
  while (waited < issued) {
    const int s = waited % stages, ph = (waited / stages) & 1;
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(ok) : "r"(bars + 8 * s), "r"(ph) : "memory");
    ++waited;
  }
}

int main(int argc, char** argv) {
  const long long n = 148LL * 128 * 4, pitch = argc > 1 ? atoll(argv[1]) : 2016;
  float* x;
  CK(cudaMalloc(&x, 9 * n * pitch * sizeof(float)));
  CK(cudaMemset(x, 0, 9 * n * pitch * sizeof(float)));
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  auto encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  CK(cudaFuncSetAttribute(walk, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int geo[][2] = {{8, 128}, {16, 64}, {32, 32}, {8, 64}, {16, 128}, {32, 64}};
  for (auto& g : geo) {
    const int F = g[0], R = g[1];
    const int box_bytes = F * R * 9 * 4;
    for (int stages = 1; stages <= 4 && stages * box_bytes <= 160 * 1024; ++stages) {
      CUtensorMap tm;
      cuuint64_t dims[3] = {2002, (cuuint64_t)n, 9};
      cuuint64_t strides[2] = {(cuuint64_t)pitch * 4, (cuuint64_t)n * pitch * 4};
      cuuint32_t box[3] = {(cuuint32_t)F, (cuuint32_t)R, 9}, es[3] = {1, 1, 1};
      CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, x, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
      const int tiles = (int)(n / R);
      const size_t smem = (size_t)stages * box_bytes + 2048;
      float best = 1e9f;
      for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        walk<<<148, 32, smem>>>(tm, F, R, stages, tiles, box_bytes);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        best = ms < best ? ms : best;
      }
      const double bytes = 9.0 * n * 2002 * 4;
      printf("box %2d feat x %3d rows x 9 (%5.1f KB, %4d-byte rows)  stages %d  in flight %5.1f KB  %.3f ms  %.2f TB/s\n", F, R,
             box_bytes / 1024.0, F * 4, stages, stages * box_bytes / 1024.0, best, bytes / best / 1e9);
    }
  }
  return 0;
}
