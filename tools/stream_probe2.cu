// stream_probe2.cu - bulk-copy ring feeding tcgen05.mma (M=64, N=16/32) the way rollout_cluster does.
// usage: stream_probe2 <ctas> <unit_bytes: 8192|16384|32768> <stages> <n> <hint 0|1> <nacc> <nofence>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include "../image_caption_b200/csrc/ptx_sm100.cuh"
using namespace capdec;
__device__ __forceinline__ void bulk_hint(void* dst, const void* src, uint32_t n, uint64_t* b, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(ptx::smem_u32(dst)), "l"(src), "r"(n), "r"(ptx::smem_u32(b)), "l"(pol) : "memory");
}
__device__ __host__ constexpr uint32_t idesc64(int n) { return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(64 >> 4) << 24); }
template <int N>
__global__ void __launch_bounds__(64, 1) probe(const uint8_t* src, size_t per_cta, int unit, int stages, int hint, long long* cycles, int nacc, int nofence) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)sm;
  uint64_t* empty = full + 32;
  uint64_t* done = full + 64;
  uint32_t* slot = (uint32_t*)(full + 65);
  uint8_t* bop = sm + 1024;           // B operand: N rows x 128 B (one k-block), zeros
  uint8_t* ring = sm + 1024 + 8192;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 8192 / 4; i += 64) ((uint32_t*)bop)[i] = 0;
  if (threadIdx.x == 0) { for (int s = 0; s < stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); } ptx::mbar_init(done, 1); ptx::fence_barrier_init(); }
  if (warp == 1) ptx::tmem_alloc(slot, 512);
  ptx::fence_proxy_async();
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tm = *slot;
  const int units = (int)(per_cta / unit);
  const int tiles = unit / 8192;  // 64x64 bf16 tiles per unit
  long long t0 = clock64();
  if (threadIdx.x == 0) {
    const uint64_t pol = ptx::policy_evict_last();
    for (int u = 0; u < units; ++u) {
      int s = u % stages;
      ptx::mbar_wait(&empty[s], ((u / stages) & 1) ^ 1);
      ptx::mbar_expect_tx(&full[s], unit);
      if (hint) bulk_hint(ring + (size_t)s * unit, src + (size_t)u * unit, unit, &full[s], pol);
      else ptx::bulk_load_1d(ring + (size_t)s * unit, src + (size_t)u * unit, unit, &full[s]);
    }
  } else if (threadIdx.x == 32) {
    for (int u = 0; u < units; ++u) {
      int s = u % stages;
      ptx::mbar_wait(&full[s], (u / stages) & 1);
      if (!nofence) ptx::tc_fence_after();
      for (int t = 0; t < tiles; ++t) {
        const uint32_t a = ptx::smem_u32(ring + (size_t)s * unit + t * 8192), b = ptx::smem_u32(bop);
#pragma unroll
        for (int k = 0; k < 4; ++k) ptx::umma_f16(tm + (uint32_t)(((t * 4 + k) % nacc) * N), ptx::umma_desc_sw128(a + k * 32), ptx::umma_desc_sw128(b + k * 32), idesc64(N), 1u);
      }
      ptx::umma_commit(&empty[s]);
    }
    ptx::umma_commit(done);
    ptx::mbar_wait(done, 0);
    cycles[blockIdx.x] = clock64() - t0;
  }
  ptx::tc_fence_before(); __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tm, 512);
}
int main(int argc, char** argv) {
  int ctas = atoi(argv[1]), unit = atoi(argv[2]), stages = atoi(argv[3]), n = atoi(argv[4]), hint = atoi(argv[5]), nacc = atoi(argv[6]), nofence = atoi(argv[7]);
  size_t per_cta = 4 << 20;
  uint8_t* src; cudaMalloc(&src, per_cta); cudaMemset(src, 0, per_cta);
  long long* cyc; cudaMalloc(&cyc, ctas * sizeof(long long));
  size_t smem = 2048 + 8192 + (size_t)unit * stages;
  auto k = n == 16 ? probe<16> : probe<32>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k<<<ctas, 64, smem>>>(src, per_cta, unit, stages, hint, cyc, nacc, nofence);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep > 0 && ms < best) best = ms;
  }
  cudaError_t e = cudaGetLastError();
  long long h[1024]; cudaMemcpy(h, cyc, ctas * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < ctas; ++i) if (h[i] > mx) mx = h[i];
  printf("mma nacc=%d nofence=%d ctas=%3d unit=%6d stages=%2d N=%d hint=%d: %.3f ms per-CTA %.1f GB/s  cycles/unit %.0f (%.1f B/clk) %s\n", nacc, nofence, ctas, unit, stages, n, hint,
         best, per_cta / (best * 1e6), (double)mx / (per_cta / unit), (double)per_cta / mx, e == cudaSuccess ? "" : cudaGetErrorString(e));
  return 0;
}
