// mma_probe.cu - cycles per tcgen05.mma (kind::f16, SS) for several (M, N), K=16, operands resident in smem.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include "../image_caption_b200/csrc/ptx_sm100.cuh"
using namespace capdec;
__device__ __host__ constexpr uint32_t idesc(int m, int n) { return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24); }
__global__ void __launch_bounds__(64, 1) probe(int m, int n, int iters, int nacc, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t* done = (uint64_t*)sm;
  uint32_t* slot = (uint32_t*)(sm + 16);
  uint8_t* a = sm + 1024;            // up to 128 rows x 128 B
  uint8_t* b = sm + 1024 + 16384;    // up to 256 rows x 128 B
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 64) ((uint32_t*)a)[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(done, 1); ptx::fence_barrier_init(); }
  if ((threadIdx.x >> 5) == 1) ptx::tmem_alloc(slot, 512);
  ptx::fence_proxy_async(); ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tm = *slot;
  if (threadIdx.x == 32) {
    const uint32_t id = idesc(m, n);
    const uint32_t aa = ptx::smem_u32(a), bb = ptx::smem_u32(b);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
      ptx::umma_f16(tm + (uint32_t)((i % nacc) * n), ptx::umma_desc_sw128(aa + (i & 3) * 32), ptx::umma_desc_sw128(bb + (i & 3) * 32), id, 1u);
    ptx::umma_commit(done);
    ptx::mbar_wait(done, 0);
    out[0] = clock64() - t0;
  }
  ptx::tc_fence_before(); __syncthreads();
  if ((threadIdx.x >> 5) == 1) ptx::tmem_dealloc(tm, 512);
}
int main() {
  long long* d; cudaMalloc(&d, 8);
  size_t smem = 2048 + 16384 + 32768;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int ms[] = {64, 128}, ns[] = {16, 32, 64, 128, 256};
  for (int m : ms) for (int n : ns) for (int nacc : {1, 2}) {
    if (nacc * n > 512) continue;
    long long h = 0;
    for (int rep = 0; rep < 2; ++rep) { probe<<<1, 64, smem>>>(m, n, 2048, nacc, d); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost); }
    cudaError_t e = cudaGetLastError();
    printf("M=%3d N=%3d nacc=%d: %.1f cycles per mma  (B bytes/clk %.1f, A bytes/clk %.1f) %s\n", m, n, nacc, h / 2048.0, n * 32.0 / (h / 2048.0), m * 32.0 / (h / 2048.0),
           e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
