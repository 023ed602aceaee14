// stream_probe.cu - how fast can one CTA pull an L2/HBM-resident stream through a cp.async.bulk ring?
// usage: stream_probe <ctas> <unit_bytes> <stages> <shared 0|1> <mb_per_cta>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mb_expect(uint64_t* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mb_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mb_wait(uint64_t* b, uint32_t par) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(s32(b)), "r"(par) : "memory");
}
__device__ __forceinline__ void bulk(void* dst, const void* src, uint32_t n, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)), "l"(src), "r"(n), "r"(s32(b)) : "memory");
}
__global__ void __launch_bounds__(64, 1) probe(const uint8_t* src, size_t per_cta, int shared_src, int unit, int stages, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t sm[];
  uint64_t* full = (uint64_t*)sm;
  uint64_t* empty = full + 32;
  uint8_t* ring = sm + 1024;
  if (threadIdx.x == 0) { for (int s = 0; s < stages; ++s) { mb_init(&full[s], 1); mb_init(&empty[s], 1); } asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  const uint8_t* base = src + (shared_src ? 0 : (size_t)blockIdx.x * per_cta);
  const int units = (int)(per_cta / unit);
  long long t0 = clock64();
  if (threadIdx.x == 0) {  // producer
    for (int u = 0; u < units; ++u) {
      int s = u % stages;
      mb_wait(&empty[s], ((u / stages) & 1) ^ 1);
      mb_expect(&full[s], unit);
      bulk(ring + (size_t)s * unit, base + (size_t)u * unit, unit, &full[s]);
    }
  } else if (threadIdx.x == 32) {  // consumer
    for (int u = 0; u < units; ++u) {
      int s = u % stages;
      mb_wait(&full[s], (u / stages) & 1);
      mb_arrive(&empty[s]);
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}
int main(int argc, char** argv) {
  int ctas = atoi(argv[1]), unit = atoi(argv[2]), stages = atoi(argv[3]), shared_src = atoi(argv[4]);
  size_t per_cta = (size_t)atoi(argv[5]) << 20;
  size_t total = shared_src ? per_cta : per_cta * ctas;
  uint8_t* src; cudaMalloc(&src, total); cudaMemset(src, 1, total);
  long long* cyc; cudaMalloc(&cyc, ctas * sizeof(long long));
  size_t smem = 1024 + (size_t)unit * stages;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    probe<<<ctas, 64, smem>>>(src, per_cta, shared_src, unit, stages, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep > 0 && ms < best) best = ms;
  }
  cudaError_t e = cudaGetLastError();
  long long h[1024]; cudaMemcpy(h, cyc, ctas * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < ctas; ++i) if (h[i] > mx) mx = h[i];
  printf("ctas=%3d unit=%6d stages=%2d shared=%d per_cta=%zuMB: %.3f ms  per-CTA %.1f GB/s  aggregate %.1f GB/s  max cycles %lld (%.1f B/clk/CTA) %s\n",
         ctas, unit, stages, shared_src, per_cta >> 20, best, per_cta / (best * 1e6), per_cta * (double)ctas / (best * 1e6), mx,
         (double)per_cta / mx, e == cudaSuccess ? "" : cudaGetErrorString(e));
  return 0;
}
