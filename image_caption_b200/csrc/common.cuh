// common.cuh - shared device types/helpers of the caption decode engine (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace capdec {

constexpr int kD = 512;       // d_model (fixed: every shipped reference config)
constexpr int kH = 8;         // heads
constexpr int kHD = 64;       // head dim
constexpr int kVPad = 128;    // vocabulary padded to one GEMM tile
constexpr int kWarp = 32;

typedef __nv_bfloat16 bf16;

// Device-resident rollout state.  One CUDA graph (= one decode step) is replayed max_len-1
// times; kernels read the current position from here instead of from launch arguments.
struct DecState {
  int step;            // n: index of the token being consumed == number of cached self positions
  int done;            // set when the rollout's stop condition was met; later launches no-op
  int greedy_done;     // latched: all greedy rows emitted <end> on one step (vit:322)
  int sample_done;     // all sampled rows finished (scst:248)
  int steps_executed;  // number of steps that produced tokens
  unsigned ticket;     // last-block election in the head kernel
  int cnt_end_now;     // greedy rows that emitted <end> this step
  int cnt_finished;    // sampled rows finished so far (recomputed every step)
};

// Per-call parameters, device resident so that the captured graph never changes.
struct CallParams {
  long long* ids;          // [M, max_len]
  float* logp;             // [M, max_len-1] or null
  float* logits_dump;      // [M, max_len-1, V] or null
  unsigned long long seed;
  unsigned long long seq_id_base;
  int max_len;
  int group;
  int n_sample;
  int start_token, end_token, pad_token;
  int n_greedy_rows, n_sample_rows;
  int no_early_stop;
  int vocab;               // rows of the embedding table
  int forced;              // teacher forcing: ids holds the whole sequence, the head scores ids[:, n+1] instead of picking
};

// Beam-search state (beam.cuh), device resident; pointers are engine-owned buffers sized for the handle's capacity.
struct BeamState {
  float* score;      // [M] running score of the hypothesis in each slot
  int* live;         // [M] 1 = alive
  int* k;            // [n_images] hypotheses to select at the next step (vit:381 `beam_size`)
  int* anc;          // [2][M][t_pad] cache row holding position q of the hypothesis in slot m (double-buffered by step parity)
  int* par;          // [M][max_len] slot at position q-1 that the hypothesis in slot m at position q extended
  float* comp_score; // [M] completed hypotheses of an image, in completion order
  int* comp_slot;    // [M]
  int* comp_pos;     // [M] position of their <end>
  int* comp_n;       // [n_images]
  int* images_done;  // [1] images whose beam is empty
};

// ---------------------------------------------------------------- element helpers
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

// load 8 consecutive elements (32 B fp32 / 16 B bf16) as fp32
__device__ __forceinline__ void load8(const float* __restrict__ p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void load8(const bf16* __restrict__ p, float (&f)[8]) {
  const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
// same, from shared memory / generic pointers (no __ldg)
__device__ __forceinline__ void load8s(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void load8s(const bf16* p, float (&f)[8]) {
  const uint4 r = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void store8(float* p, const float (&f)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void store8(bf16* p, const float (&f)[8]) {
  uint4 r;
  r.x = pack_bf16x2(f[0], f[1]); r.y = pack_bf16x2(f[2], f[3]);
  r.z = pack_bf16x2(f[4], f[5]); r.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = r;
}
// store 4 consecutive fp32 values as T
__device__ __forceinline__ void store4(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void store4(bf16* p, float a, float b, float c, float d) {
  uint2 r; r.x = pack_bf16x2(a, b); r.y = pack_bf16x2(c, d);
  *reinterpret_cast<uint2*>(p) = r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------- Philox4x32-10
// Scheme documented in oracle/philox.py and DESIGN.md: key=(seed lo,hi), ctr=(seq lo,hi,step,0).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float philox_uniform(unsigned long long seed, unsigned long long seq, int step) {
  const uint4 r = philox4x32_10(make_uint4((uint32_t)seq, (uint32_t)(seq >> 32), (uint32_t)step, 0u),
                                make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  return (float)(r.x >> 8) * (1.0f / 16777216.0f);
}


// x[row] <- LayerNorm(x[row] + bias + sum_z part[z][row]) * gamma + beta, and its bf16 mirror; one warp per row of 512.
// Shared by sum_ln_kernel and by the split-K projection that closes itself (gemm_swap.cuh, SwapPartialLN) so that
// both produce the same bits.  Partials and x are read through L2 (ld.global.cg): they were written by other CTAs.
template <int NZ>
__device__ __forceinline__ void sum_ln_row(float* __restrict__ x, const float* __restrict__ part, int M, int row, int lane,
                                           const float4 (&bi)[4], const float4 (&ga)[4], const float4 (&be)[4],
                                           __nv_bfloat16* __restrict__ xb) {
  float4 xin[4], pz[NZ][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 128 + lane * 4;
    xin[i] = __ldcg(reinterpret_cast<const float4*>(x + (size_t)row * kD + c));
#pragma unroll
    for (int z = 0; z < NZ; ++z) pz[z][i] = __ldcg(reinterpret_cast<const float4*>(part + ((size_t)z * M + row) * kD + c));
  }
  float v[16];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float4 a = xin[i];
    a.x += bi[i].x; a.y += bi[i].y; a.z += bi[i].z; a.w += bi[i].w;
#pragma unroll
    for (int z = 0; z < NZ; ++z) { a.x += pz[z][i].x; a.y += pz[z][i].y; a.z += pz[z][i].z; a.w += pz[z][i].w; }
    v[4 * i] = a.x; v[4 * i + 1] = a.y; v[4 * i + 2] = a.z; v[4 * i + 3] = a.w;
    s += (a.x + a.y) + (a.z + a.w);
  }
  const float mean = warp_sum(s) * (1.f / kD);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { v[i] -= mean; q = fmaf(v[i], v[i], q); }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / kD) + 1e-5f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 128 + lane * 4;
    const float r0 = v[4 * i] * rstd * ga[i].x + be[i].x, r1 = v[4 * i + 1] * rstd * ga[i].y + be[i].y;
    const float r2 = v[4 * i + 2] * rstd * ga[i].z + be[i].z, r3 = v[4 * i + 3] * rstd * ga[i].w + be[i].w;
    *reinterpret_cast<float4*>(x + (size_t)row * kD + c) = make_float4(r0, r1, r2, r3);
    store4(xb + (size_t)row * kD + c, r0, r1, r2, r3);
  }
}

}  // namespace capdec
