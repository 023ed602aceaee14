// epilogue.cuh - GEMM epilogue functors shared by the tcgen05 and the CUDA-core GEMM.
// Contract: store<NV>(row, col0, v) receives NV (multiple of 4) raw accumulators of one
// output row starting at column col0 (multiple of 4, never straddling a 64-column head
// boundary inside a group of 4).  The caller guarantees row < M and col0 + NV <= N.
#pragma once
#include "common.cuh"

namespace capdec {

// out[row, col] = act(acc + bias[col])
template <typename OutT, bool kRelu>
struct EpiLinear {
  OutT* out;
  const float* bias;  // may be null
  int ldo;
  template <int NV>
  __device__ __forceinline__ void store(int row, int col0, const float* v) const {
    OutT* dst = out + (size_t)row * ldo + col0;
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
      float4 b = bias ? __ldg(reinterpret_cast<const float4*>(bias + col0 + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
      float r0 = v[i] + b.x, r1 = v[i + 1] + b.y, r2 = v[i + 2] + b.z, r3 = v[i + 3] + b.w;
      if (kRelu) { r0 = fmaxf(r0, 0.f); r1 = fmaxf(r1, 0.f); r2 = fmaxf(r2, 0.f); r3 = fmaxf(r3, 0.f); }
      store4(dst + i, r0, r1, r2, r3);
    }
  }
};

// Self-attention in_proj: columns [0,512) -> q buffer, [512,1024) -> K cache,
// [1024,1536) -> V cache, both appended at position st->step.
// Cache layout per layer: [M][H][t_pad][64] so that one (sequence, head) history is contiguous.
template <typename Elem>
struct EpiQKV {
  Elem* q;             // [M, 512]
  Elem* kcache;        // this layer
  Elem* vcache;
  const float* bias;   // [1536]
  const DecState* st;
  int t_pad;
  template <int NV>
  __device__ __forceinline__ void store(int row, int col0, const float* v) const {
    const int n = st->step;
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
      const int col = col0 + i;
      const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col));
      const float r0 = v[i] + b.x, r1 = v[i + 1] + b.y, r2 = v[i + 2] + b.z, r3 = v[i + 3] + b.w;
      Elem* dst;
      if (col < kD) {
        dst = q + (size_t)row * kD + col;
      } else {
        const int c = (col - kD) & (kD - 1);
        Elem* base = (col < 2 * kD) ? kcache : vcache;
        dst = base + (((size_t)row * kH + (c >> 6)) * t_pad + n) * kHD + (c & 63);
      }
      store4(dst, r0, r1, r2, r3);
    }
  }
};

// Prefill: row = image*S + s, column = layer*1024 + kv*512 + head*64 + d.
// Destination layout [L][2][B_img][H][S][64]: one (layer, kv, image, head) tile is contiguous
// so the cross-attention kernel can fetch it with a single bulk (TMA) copy.
template <typename Elem>
struct EpiCrossKV {
  Elem* out;
  const float* bias;  // [L*1024]
  int n_images;       // images of the whole prefill (the cache's image stride)
  int mem_len;
  int img0 = 0;       // first image of this launch (chunked prefill: host copies pipelined with the projection)
  template <int NV>
  __device__ __forceinline__ void store(int row, int col0, const float* v) const {
    const int il = row / mem_len, s = row - il * mem_len, img = img0 + il;
#pragma unroll
    for (int i = 0; i < NV; i += 4) {
      const int col = col0 + i;
      const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col));
      const int lk = col >> 9;  // layer*2 + kv
      const int c = col & (kD - 1);
      Elem* dst = out + ((((size_t)lk * n_images + img) * kH + (c >> 6)) * mem_len + s) * kHD + (c & 63);
      store4(dst, v[i] + b.x, v[i + 1] + b.y, v[i + 2] + b.z, v[i + 3] + b.w);
    }
  }
};

}  // namespace capdec
