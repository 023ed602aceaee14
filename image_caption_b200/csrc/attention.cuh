// attention.cuh - single-query attention kernels of the decode step.
//   self_attn_kernel     one warp per (sequence, head) over the cached prefix
//   cross_attn_kernel    one warp per (image, head, query-tile): the K/V rows of an
//                                   image are read once for all GT rollouts that share it
// Both are HBM-bound streams: 16-byte vector loads, 8 lanes per 64-element row so every
// warp-wide load instruction covers 4 whole rows (512 contiguous bytes in bf16).
// Softmax scale 1/sqrt(64) = 0.125 is folded into q (exact: power of two).
#pragma once
#include "common.cuh"

namespace capdec {

constexpr int kAttnWarps = 4;

// q, out: [M, 512]; kc/vc: this layer's cache [M][H][t_pad][64]; positions 0..st->step valid.
template <typename Elem, int NIT>
__global__ void __launch_bounds__(kAttnWarps * 32)
self_attn_kernel(const Elem* __restrict__ q, const Elem* __restrict__ kc, const Elem* __restrict__ vc,
                 Elem* __restrict__ out, int M, int t_pad, const DecState* st) {
  asm volatile("griddepcontrol.wait;" ::: "memory");  // q and the appended K/V rows (PDL launches)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (st->done) return;
  const int gw = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
  if (gw >= M * kH) return;
  const int lane = threadIdx.x & 31, g = lane >> 3, c = lane & 7;
  const int n = st->step;  // last valid position
  const size_t row_base = (size_t)gw * t_pad * kHD;  // gw == m*H + h
  const Elem* kb = kc + row_base + c * 8;
  const Elem* vb = vc + row_base + c * 8;

  float qv[8];
  load8(q + (size_t)gw * kHD + c * 8, qv);  // (m*512 + h*64) == gw*64
#pragma unroll
  for (int i = 0; i < 8; ++i) qv[i] *= 0.125f;

  float s[NIT];
#pragma unroll
  for (int it = 0; it < NIT; ++it) {
    const int j = it * 4 + g;
    float d = 0.f;
    if (it * 4 <= n) {  // warp-uniform
      if (j <= n) {
        float kv[8];
        load8(kb + (size_t)j * kHD, kv);
#pragma unroll
        for (int i = 0; i < 8; ++i) d = fmaf(qv[i], kv[i], d);
      }
      d += __shfl_xor_sync(0xffffffffu, d, 1);
      d += __shfl_xor_sync(0xffffffffu, d, 2);
      d += __shfl_xor_sync(0xffffffffu, d, 4);
    }
    s[it] = (j <= n) ? d : -INFINITY;
  }
  float mx = s[0];
#pragma unroll
  for (int it = 1; it < NIT; ++it) mx = fmaxf(mx, s[it]);
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));

  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float sum = 0.f;
#pragma unroll
  for (int it = 0; it < NIT; ++it) {
    const int j = it * 4 + g;
    if (it * 4 <= n && j <= n) {
      const float p = expf(s[it] - mx);
      sum += p;
      float vv[8];
      load8(vb + (size_t)j * kHD, vv);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = fmaf(p, vv[i], acc[i]);
    }
  }
  sum += __shfl_xor_sync(0xffffffffu, sum, 8);
  sum += __shfl_xor_sync(0xffffffffu, sum, 16);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
  }
  if (g == 0) {
    const float inv = 1.f / sum;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= inv;
    store8(out + (size_t)gw * kHD + c * 8, acc);
  }
}

// q, out: [n_images*group, 512] image-major; ck/cv: this layer's [n_images][H][S][64].
// Work item = (image, head, query tile of GT sequences); group % GT == 0.
template <typename Elem, int GT>
__global__ void __launch_bounds__(kAttnWarps * 32)
cross_attn_kernel(const Elem* __restrict__ q, const Elem* __restrict__ ck, const Elem* __restrict__ cv,
                  Elem* __restrict__ out, int n_images, int group, int S, const DecState* st) {
  if (st->done) return;
  const int tiles = group / GT;
  const int gw = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
  if (gw >= n_images * kH * tiles) return;
  const int tile = gw % tiles, ih = gw / tiles;  // ih == image*H + head
  const int img = ih / kH, h = ih % kH;
  const int lane = threadIdx.x & 31, g = lane >> 3, c = lane & 7;
  const int m0 = img * group + tile * GT;
  const Elem* kb = ck + (size_t)ih * S * kHD + c * 8;
  const Elem* vb = cv + (size_t)ih * S * kHD + c * 8;

  float qv[GT][8], acc[GT][8], mrun[GT], lrun[GT];
#pragma unroll
  for (int t = 0; t < GT; ++t) {
    load8(q + (size_t)(m0 + t) * kD + h * kHD + c * 8, qv[t]);
    mrun[t] = -INFINITY;
    lrun[t] = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { qv[t][i] *= 0.125f; acc[t][i] = 0.f; }
  }

  for (int j0 = 0; j0 < S; j0 += 8) {
    float kv[2][8], vv[2][8];
    bool ok[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {  // issue all loads of this slab before any math
      const int j = j0 + u * 4 + g;
      ok[u] = j < S;
      if (ok[u]) {
        load8(kb + (size_t)j * kHD, kv[u]);
        load8(vb + (size_t)j * kHD, vv[u]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) { kv[u][i] = 0.f; vv[u][i] = 0.f; }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
#pragma unroll
      for (int t = 0; t < GT; ++t) {
        float d = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) d = fmaf(qv[t][i], kv[u][i], d);
        d += __shfl_xor_sync(0xffffffffu, d, 1);
        d += __shfl_xor_sync(0xffffffffu, d, 2);
        d += __shfl_xor_sync(0xffffffffu, d, 4);
        if (ok[u]) {  // online softmax, private to this lane group
          const float mn = fmaxf(mrun[t], d);
          const float corr = expf(mrun[t] - mn);  // exp(-inf) = 0 on the first row
          const float p = expf(d - mn);
          lrun[t] = lrun[t] * corr + p;
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[t][i] = fmaf(p, vv[u][i], acc[t][i] * corr);
          mrun[t] = mn;
        }
      }
    }
  }
  // merge the four lane groups
#pragma unroll
  for (int t = 0; t < GT; ++t) {
    float mx = fmaxf(mrun[t], __shfl_xor_sync(0xffffffffu, mrun[t], 8));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
    const float sc = (mrun[t] == -INFINITY) ? 0.f : expf(mrun[t] - mx);
    float l = lrun[t] * sc;
    l += __shfl_xor_sync(0xffffffffu, l, 8);
    l += __shfl_xor_sync(0xffffffffu, l, 16);
    const float inv = 1.f / l;
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float a = acc[t][i] * sc;
      a += __shfl_xor_sync(0xffffffffu, a, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16);
      o[i] = a * inv;
    }
    if (g == 0) store8(out + (size_t)(m0 + t) * kD + h * kHD + c * 8, o);
  }
}

}  // namespace capdec
