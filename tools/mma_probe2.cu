// mma_probe2.cu - does the ~148-cycle cost of one tcgen05.mma (kind::f16, SS, K=16) belong to the issuing thread,
// to the CTA, or to the SM?  Variants: several warps of one CTA issuing into their own accumulators, and two CTAs
// resident on the same SM.  Reports cycles per MMA seen by one issuer and the aggregate MMAs per 1000 cycles per SM.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#include "../image_caption_b200/csrc/ptx_sm100.cuh"
using namespace capdec;
__device__ __host__ constexpr uint32_t idesc(int m, int n) { return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24); }

// block = 32 * (1 + nissue) threads: warp 0 allocates TMEM, warps 1..nissue each issue `iters` MMAs
__global__ void __launch_bounds__(160, 2) probe(int m, int n, int iters, int nissue, long long* out, int alloc_warp, int ncols, int nacc, int flags) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t* done = (uint64_t*)sm;  // [4]
  uint32_t* slot = (uint32_t*)(sm + 64);
  uint8_t* a = sm + 1024;          // 128 rows x 128 B
  uint8_t* b = sm + 1024 + 16384;  // 128 rows x 128 B
  for (int i = threadIdx.x; i < (16384 + 16384 + 8 * 20480) / 4; i += blockDim.x)
    ((uint32_t*)a)[i] = (flags & 32) ? (0x3c003c00u | ((uint32_t)(i * 2654435761u) >> 24) | (((uint32_t)(i * 40503u) & 0x7fu) << 16)) : 0u;  // bf16 pairs near 0.0078
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) ptx::mbar_init(&done[i], 1); ptx::fence_barrier_init(); }
  if (warp == alloc_warp) ptx::tmem_alloc(slot, ncols);
  ptx::fence_proxy_async(); ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t tm = *slot;
  if ((flags & 8) && warp >= 1 && warp <= nissue) {  // whole warp converged, MMA under elect.sync
    const int w = warp - 1;
    const uint32_t id = idesc(m, n);
    const uint32_t aa = ptx::smem_u32(a), bb = ptx::smem_u32(b);
    long long t0 = clock64();
    for (int i = 0; i < iters; i += 4) {
      if (flags & 2) ptx::tc_fence_after();
      if (ptx::elect_one()) {
#pragma unroll
        const uint32_t off = (flags & 16) ? (uint32_t)(((i >> 2) & 7) * 20480) : 0u;
        for (int k = 0; k < 4; ++k)
          ptx::umma_f16(tm + (uint32_t)(w * n), ptx::umma_desc_sw128(aa + off + k * 32), ptx::umma_desc_sw128(bb + off + k * 32), id, (uint32_t)((i | k) != 0));
        if (flags & 1) ptx::umma_commit(&done[3]);
      }
      __syncwarp();
      if (flags & 4) ptx::mbar_wait(&done[3], (uint32_t)((i >> 2) & 1));
    }
    if (ptx::elect_one()) ptx::umma_commit(&done[w]);
    __syncwarp();
    ptx::mbar_wait(&done[w], 0);
    if ((threadIdx.x & 31) == 0) out[blockIdx.x * 4 + w] = clock64() - t0;
  } else if (warp >= 1 && warp <= nissue && (threadIdx.x & 31) == 0) {
    const int w = warp - 1;
    const uint32_t id = idesc(m, n);
    const uint32_t aa = ptx::smem_u32(a), bb = ptx::smem_u32(b);
    long long t0 = clock64();
    // flags: 1 = tcgen05.commit to a scratch mbarrier after every 4th MMA, 2 = tcgen05.fence::after_thread_sync before
    // every 4th, 4 = wait for that commit (parity) before the next group (serialises issue on completion)
    for (int i = 0; i < iters; ++i) {
      if ((flags & 2) && (i & 3) == 0) ptx::tc_fence_after();
      ptx::umma_f16(tm + (uint32_t)(w * n) + (nacc > 0 ? (uint32_t)((i % nacc) * n) : 0u), ptx::umma_desc_sw128(aa + (i & 3) * 32), ptx::umma_desc_sw128(bb + (i & 3) * 32), id, (uint32_t)(i != 0));
      if ((flags & 1) && (i & 3) == 3) ptx::umma_commit(&done[3]);
      if ((flags & 4) && (i & 3) == 3) ptx::mbar_wait(&done[3], (uint32_t)((i >> 2) & 1));
    }
    ptx::umma_commit(&done[w]);
    ptx::mbar_wait(&done[w], 0);
    out[blockIdx.x * 4 + w] = clock64() - t0;
  }
  ptx::tc_fence_before(); __syncthreads();
  if (warp == alloc_warp) ptx::tmem_dealloc(tm, ncols);
}

int main() {
  long long* d; cudaMalloc(&d, 8 * 4 * 296);
  size_t smem = 2048 + 32768 + 8 * 20480;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, probe, 160, smem);
  printf("resident CTAs per SM: %d\n", occ);
  const int iters = 2048;
  static long long h[4 * 296];
  struct V { int n, nissue, grid, alloc_warp, ncols, nacc, flags; };
  const V vs[] = {{32, 1, 1, 1, 32, 0, 27}, {32, 1, 1, 1, 32, 0, 59}, {32, 2, 1, 0, 64, 0, 27}, {32, 2, 1, 0, 64, 0, 59}, {32, 4, 1, 0, 128, 0, 59},
                  {128, 1, 1, 1, 128, 0, 59}, {128, 2, 1, 0, 256, 0, 59}};
  for (const V& v : vs) {
    for (int rep = 0; rep < 2; ++rep) { cudaMemset(d, 0, sizeof h); probe<<<v.grid, 32 * (1 + v.nissue), smem>>>(128, v.n, iters, v.nissue, d, v.alloc_warp, v.ncols, v.nacc, v.flags); cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost); }
    cudaError_t e = cudaGetLastError();
    long long mx = 0; double sum = 0; int cnt = 0;
    for (int i = 0; i < v.grid * 4; ++i) if (h[i]) { mx = h[i] > mx ? h[i] : mx; sum += (double)h[i]; ++cnt; }
    printf("N=%3d issuers/CTA=%d grid=%3d alloc_warp=%d tmem_cols=%3d modulo_acc=%d flags=%d: %.1f cycles per mma per issuer (mean), slowest %.1f %s\n", v.n, v.nissue,
           v.grid, v.alloc_warp, v.ncols, v.nacc, v.flags, sum / cnt / iters, (double)mx / iters, e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
