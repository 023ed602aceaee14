// gemm_simt.cuh - CUDA-core GEMM, C[M,N] = A[M,K] * W[N,K]^T, fp32 accumulate.
// The strict-parity (fp32) mode runs every projection through this kernel; the bf16 mode
// uses it only for the 109-way vocabulary projection (kept in fp32) and as a debug route.
#pragma once
#include "common.cuh"

namespace capdec {

constexpr int kSimtBM = 64, kSimtBN = 64, kSimtBK = 16, kSimtThreads = 256;

template <typename T> __device__ __forceinline__ float4 load4_f32(const T* p);
template <> __device__ __forceinline__ float4 load4_f32<float>(const float* p) {
  return __ldg(reinterpret_cast<const float4*>(p));
}
template <> __device__ __forceinline__ float4 load4_f32<bf16>(const bf16* p) {
  const uint2 r = __ldg(reinterpret_cast<const uint2*>(p));
  return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u),
                     __uint_as_float(r.y << 16), __uint_as_float(r.y & 0xffff0000u));
}

// A: [M, K] row-major (lda), W: [N, K] row-major (ldw).  K % 16 == 0, N % 4 == 0.
// If st != null the launch no-ops once the rollout is done.
template <typename TA, typename TW, typename Epi>
__global__ void __launch_bounds__(kSimtThreads)
gemm_simt_kernel(const TA* __restrict__ A, int lda, const TW* __restrict__ W, int ldw,
                 int M, int N, int K, Epi epi, const DecState* st) {
  if (st != nullptr && st->done) return;
  __shared__ __align__(16) float As[2][kSimtBK][kSimtBM + 4];
  __shared__ __align__(16) float Ws[2][kSimtBK][kSimtBN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * kSimtBM, n0 = blockIdx.x * kSimtBN;
  const int lr = tid >> 2, lk = (tid & 3) * 4;  // tile row / k offset this thread loads
  const int ty = tid >> 4, tx = tid & 15;
  const bool a_ok = (m0 + lr) < M, w_ok = (n0 + lr) < N;
  const TA* ap = A + (size_t)(m0 + lr) * lda + lk;
  const TW* wp = W + (size_t)(n0 + lr) * ldw + lk;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 ra = a_ok ? load4_f32<TA>(ap) : z;
  float4 rw = w_ok ? load4_f32<TW>(wp) : z;
  const int nk = K / kSimtBK;
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    As[buf][lk + 0][lr] = ra.x; As[buf][lk + 1][lr] = ra.y; As[buf][lk + 2][lr] = ra.z; As[buf][lk + 3][lr] = ra.w;
    Ws[buf][lk + 0][lr] = rw.x; Ws[buf][lk + 1][lr] = rw.y; Ws[buf][lk + 2][lr] = rw.z; Ws[buf][lk + 3][lr] = rw.w;
    __syncthreads();
    if (kt + 1 < nk) {  // prefetch the next k-slab while computing this one
      ra = a_ok ? load4_f32<TA>(ap + (kt + 1) * kSimtBK) : z;
      rw = w_ok ? load4_f32<TW>(wp + (kt + 1) * kSimtBK) : z;
    }
#pragma unroll
    for (int k = 0; k < kSimtBK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      const float4 w = *reinterpret_cast<const float4*>(&Ws[buf][k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    // double-buffered smem: the next iteration writes the other buffer, so one barrier per slab
  }
  const int col0 = n0 + tx * 4;
  if (col0 + 4 <= N) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int row = m0 + ty * 4 + i;
      if (row < M) epi.template store<4>(row, col0, acc[i]);
    }
  }
}

}  // namespace capdec
