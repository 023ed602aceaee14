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
// anc (beam search, beam.cuh; else null): [2][M][t_pad], buffer (step & 1): the cache row that holds position j of
// the hypothesis in row m -- re-ranked hypotheses read their ancestors' rows instead of copying them.
template <typename Elem, int NIT, bool kAnc>
__global__ void __launch_bounds__(kAttnWarps * 32)
self_attn_kernel(const Elem* __restrict__ q, const Elem* __restrict__ kc, const Elem* __restrict__ vc,
                 Elem* __restrict__ out, int M, int t_pad, const DecState* st, const int* __restrict__ anc) {
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
  const int* arow = kAnc ? anc + ((size_t)(n & 1) * M + gw / kH) * t_pad : nullptr;
  // offset of position j's row relative to kb / vb: 0 without ancestry
  auto row_shift = [&](int j) -> long long {
    if constexpr (kAnc) return ((long long)arow[j] - gw / kH) * (long long)kH * t_pad * kHD;
    else return 0ll;
  };

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
        load8(kb + row_shift(j) + (size_t)j * kHD, kv);
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
      load8(vb + row_shift(j) + (size_t)j * kHD, vv);
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

// bf16 engine, v2 step path.  Same work split as self_attn_kernel (one warp per (sequence, head), 8 lanes per
// 128-byte row, 4 rows per warp-wide load) but flash-style over chunks of 32 positions: the K AND V rows of a
// chunk are requested together (16 independent 16-byte loads per lane in flight, streaming, no L1 allocation)
// before any arithmetic, so a warp waits for HBM once per chunk instead of twice per rollout position range.
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
template <bool kAnc, int kIts = 8>  // kIts: positions per chunk / 4 (registers: 8 per it); kAnc: beam search, rows are read through the ancestry table (a separate instantiation: the
                      // indirection costs ~50 registers, which the plain rollout's occupancy should not pay)
__global__ void __launch_bounds__(kAttnWarps * 32)
self_attn_stream_kernel(const bf16* __restrict__ q, const bf16* __restrict__ kc, const bf16* __restrict__ vc,
                        bf16* __restrict__ out, int M, int t_pad, const DecState* st, const int* __restrict__ anc) {
  // DecState was last written by the head kernel of the previous step, at least two kernels upstream (the QKV
  // projection sits in between), so it is complete before this grid may start under programmatic launch: reading it
  // before the wait takes an L2 round trip off the critical path (the K/V addresses depend on `n`).
  const int n = st->step;  // last valid position
  const bool done = st->done != 0;
  asm volatile("griddepcontrol.wait;" ::: "memory");  // q and the appended K/V rows (PDL launches)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (done) return;
  const int gw = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
  if (gw >= M * kH) return;
  const int lane = threadIdx.x & 31, g = lane >> 3, c = lane & 7;
  const size_t row_base = (size_t)gw * t_pad * kHD + c * 8;  // gw == m*H + h
  const bf16* kb = kc + row_base;
  const bf16* vb = vc + row_base;
  // beam search (beam.cuh): position j of the hypothesis in row m lives in cache row anc[step & 1][m][j]
  const int* arow = kAnc ? anc + ((size_t)(n & 1) * M + gw / kH) * t_pad : nullptr;
  auto row_shift = [&](int pos) -> long long {
    if constexpr (kAnc) return ((long long)arow[pos] - gw / kH) * (long long)kH * t_pad * kHD;
    else return 0ll;
  };
  float qv[8];
  {
    const uint4 r = *reinterpret_cast<const uint4*>(q + (size_t)gw * kHD + c * 8);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { qv[2 * i] = __uint_as_float(w[i] << 16) * 0.125f; qv[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u) * 0.125f; }
  }
  float m_run = -INFINITY, l_run = 0.f;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int p0 = 0; p0 <= n; p0 += 4 * kIts) {
    uint4 kr[kIts], vr[kIts];
#pragma unroll
    for (int it = 0; it < kIts; ++it) {
      const int pos = p0 + it * 4 + g;
      if (pos <= n) {
        const long long sh = row_shift(pos);
        kr[it] = ld_stream16(kb + sh + (size_t)pos * kHD);
        vr[it] = ld_stream16(vb + sh + (size_t)pos * kHD);
      } else {
        kr[it] = make_uint4(0u, 0u, 0u, 0u);
        vr[it] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    float s[kIts];
    float mc = -INFINITY;
#pragma unroll
    for (int it = 0; it < kIts; ++it) {
      const uint32_t w[4] = {kr[it].x, kr[it].y, kr[it].z, kr[it].w};
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        d = fmaf(qv[2 * i], __uint_as_float(w[i] << 16), d);
        d = fmaf(qv[2 * i + 1], __uint_as_float(w[i] & 0xffff0000u), d);
      }
      d += __shfl_xor_sync(0xffffffffu, d, 1);
      d += __shfl_xor_sync(0xffffffffu, d, 2);
      d += __shfl_xor_sync(0xffffffffu, d, 4);
      s[it] = (p0 + it * 4 + g <= n) ? d : -INFINITY;
      mc = fmaxf(mc, s[it]);
    }
    mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 8));
    mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 16));
    const float m_new = fmaxf(m_run, mc);          // finite: position p0 <= n is in this chunk
    const float corr = __expf(m_run - m_new);      // exp(-inf) = 0 on the first chunk
    l_run *= corr;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= corr;
#pragma unroll
    for (int it = 0; it < kIts; ++it) {
      const float pr = __expf(s[it] - m_new);      // 0 for masked positions
      l_run += pr;
      const uint32_t w[4] = {vr[it].x, vr[it].y, vr[it].z, vr[it].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] = fmaf(pr, __uint_as_float(w[i] << 16), acc[2 * i]);
        acc[2 * i + 1] = fmaf(pr, __uint_as_float(w[i] & 0xffff0000u), acc[2 * i + 1]);
      }
    }
    m_run = m_new;
  }
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 8);
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 16);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
  }
  if (g == 0) {
    const float inv = 1.f / l_run;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= inv;
    store8(out + (size_t)gw * kHD + c * 8, acc);
  }
}

// bf16 engine, v2 step path, one sequence per image (greedy decoding: cfg2, cfg5).  One warp per (image, head), the
// same lane layout and online softmax as self_attn_stream_kernel, over the projected memory
// ckv = [L][2][n_images][H][S][64].  The memory was projected at prefill, so the first chunk's K and V rows are
// requested (and the second chunk's lines pulled into L2) BEFORE the programmatic-launch wait: only the query
// depends on the previous kernel of the step.  With one query per (image, head) the tensor-core kernel
// (cross_attn_tma.cuh) spends its time on CTA scheduling -- 126 registers x 128 threads allow 4 CTAs per SM,
// so 2048 CTAs run in 3.5 waves -- while here every (image, head) is resident at once.
__global__ void __launch_bounds__(kAttnWarps * 32)
cross_attn_stream_kernel(const bf16* __restrict__ q, const bf16* __restrict__ ckv, bf16* __restrict__ out, int layer,
                         int n_images, int S, const DecState* st) {
  const int gw = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);  // image * H + head
  const bool live = gw < n_images * kH;
  const int lane = threadIdx.x & 31, g = lane >> 3, c = lane & 7;
  const size_t plane = (size_t)n_images * kH * S * kHD;  // one of K / V of one layer
  const bf16* kb = ckv + (size_t)(2 * layer) * plane + (size_t)(live ? gw : 0) * S * kHD + c * 8;
  const bf16* vb = kb + plane;
  uint4 kr[8], vr[8];
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int pos = it * 4 + g;
    if (live && pos < S) {
      kr[it] = ld_stream16(kb + (size_t)pos * kHD);
      vr[it] = ld_stream16(vb + (size_t)pos * kHD);
    } else {
      kr[it] = make_uint4(0u, 0u, 0u, 0u);
      vr[it] = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  if (live && 32 + lane < S) {  // one 128-byte row per lane
    asm volatile("prefetch.global.L2 [%0];" ::"l"(kb - c * 8 + (size_t)(32 + lane) * kHD));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(vb - c * 8 + (size_t)(32 + lane) * kHD));
  }
  const bool done = st->done != 0;  // written by the previous step's head kernel, long complete (see self_attn_stream_kernel)
  asm volatile("griddepcontrol.wait;" ::: "memory");  // the query projection of this step
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (done || !live) return;
  float qv[8];
  {
    const uint4 r = *reinterpret_cast<const uint4*>(q + (size_t)gw * kHD + c * 8);  // group == 1: row == image
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { qv[2 * i] = __uint_as_float(w[i] << 16) * 0.125f; qv[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u) * 0.125f; }
  }
  float m_run = -INFINITY, l_run = 0.f;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int p0 = 0; p0 < S; p0 += 32) {
    if (p0 > 0) {
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int pos = p0 + it * 4 + g;
        if (pos < S) {
          kr[it] = ld_stream16(kb + (size_t)pos * kHD);
          vr[it] = ld_stream16(vb + (size_t)pos * kHD);
        } else {
          kr[it] = make_uint4(0u, 0u, 0u, 0u);
          vr[it] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      if (p0 + 32 + lane < S) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(kb - c * 8 + (size_t)(p0 + 32 + lane) * kHD));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(vb - c * 8 + (size_t)(p0 + 32 + lane) * kHD));
      }
    }
    float s[8];
    float mc = -INFINITY;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const uint32_t w[4] = {kr[it].x, kr[it].y, kr[it].z, kr[it].w};
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        d = fmaf(qv[2 * i], __uint_as_float(w[i] << 16), d);
        d = fmaf(qv[2 * i + 1], __uint_as_float(w[i] & 0xffff0000u), d);
      }
      d += __shfl_xor_sync(0xffffffffu, d, 1);
      d += __shfl_xor_sync(0xffffffffu, d, 2);
      d += __shfl_xor_sync(0xffffffffu, d, 4);
      s[it] = (p0 + it * 4 + g < S) ? d : -INFINITY;
      mc = fmaxf(mc, s[it]);
    }
    mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 8));
    mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 16));
    const float m_new = fmaxf(m_run, mc);      // finite: position p0 < S is in this chunk
    const float corr = __expf(m_run - m_new);  // exp(-inf) = 0 on the first chunk
    l_run *= corr;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= corr;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const float pr = __expf(s[it] - m_new);  // 0 for masked positions
      l_run += pr;
      const uint32_t w[4] = {vr[it].x, vr[it].y, vr[it].z, vr[it].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] = fmaf(pr, __uint_as_float(w[i] << 16), acc[2 * i]);
        acc[2 * i + 1] = fmaf(pr, __uint_as_float(w[i] & 0xffff0000u), acc[2 * i + 1]);
      }
    }
    m_run = m_new;
  }
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 8);
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 16);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
  }
  if (g == 0) {
    const float inv = 1.f / l_run;
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
