// attention_dense.cuh - attention over whole sequences for the full-sequence (teacher-forced) pass (SURVEY 8f-4;
// reference: nn.TransformerDecoderLayer's self- and cross-attention blocks as called at vit:172-181 / scst:223 with
// tgt = the whole caption and tgt_mask = causal).
//
// One CTA per (unit, head).  A unit is
//   self-attention : one sequence.   T queries, keys/values = that sequence's T cached positions, causal.
//   cross-attention: one image.      group * T queries (all rollouts of the image), keys/values = its S memory rows.
// so the K/V block of a unit is read from HBM once for all of its queries (the step kernels re-read it at every step).
// The K/V block ([keys][64] bf16, contiguous in both caches) is staged in shared memory with a 16-byte XOR swizzle; each
// warp takes 16 query rows at a time: S = Q K^T and O = P V on mma.sync.m16n8k16 (bf16 in, fp32 accumulate), the softmax
// in registers between them (scores never leave the register file).  Head dim 64, keys <= 256.
//
// This is the legacy tensor path on purpose: per (unit, head) the two products are 49..294 x 64 x 49..196 - a fraction of
// one tcgen05 128 x 256 tile - and the pass is dominated by the projections (gemm_ws.cuh, tcgen05).
#pragma once
#include "common.cuh"

namespace capdec {

constexpr int kAdWarps = 4;

struct DenseAttnArgs {
  const bf16* q;        // [rows, ldq] queries, head h at columns [64 h, 64 h + 64)
  bf16* out;            // [rows, ldo]
  const bf16* k;        // K block of (unit u, head h) at k + ((unit0 + u) * kH + h) * unit_stride, [keys][64]
  const bf16* v;
  long long unit_stride;
  int ldq, ldo;
  int unit0;            // first unit of this launch inside the cache (cache row / image index)
  int nq;               // query rows per unit; unit u owns rows [u * nq, (u + 1) * nq)
  int nk;               // keys per unit
  int causal_t;         // 0 = every key; else query row i of the unit sees keys j <= i % causal_t
};

namespace ad {
__device__ __forceinline__ void ldsm4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm4_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&p);
}
}  // namespace ad

// NKT = key tiles of 8 held in registers (nk <= 8 NKT, NKT even).  grid = (units, heads), block = kAdWarps warps.
// dynamic shared memory: 2 * NKT * 8 * 128 bytes.
template <int NKT>
__global__ void __launch_bounds__(kAdWarps * 32)
dense_attn_kernel(const DenseAttnArgs A) {
  extern __shared__ __align__(128) uint8_t ad_smem[];
  const int unit = blockIdx.x, head = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, qq = lane & 3;
  const int nk = A.nk, nk16 = (nk + 15) & ~15;
  uint8_t* Ks = ad_smem;
  uint8_t* Vs = ad_smem + (size_t)NKT * 8 * 128;

  {  // K and V blocks -> shared memory, 16-byte chunk c of row r at r * 128 + ((c ^ (r & 7)) << 4); rows >= nk are zero
    const size_t blk = ((size_t)(A.unit0 + unit) * kH + head) * (size_t)A.unit_stride;
    const uint4* kg = reinterpret_cast<const uint4*>(A.k + blk);
    const uint4* vg = reinterpret_cast<const uint4*>(A.v + blk);
    for (int i = threadIdx.x; i < nk16 * 8; i += kAdWarps * 32) {
      const int r = i >> 3, c = i & 7;
      uint4 kv = make_uint4(0u, 0u, 0u, 0u), vv = kv;
      if (r < nk) { kv = __ldg(kg + i); vv = __ldg(vg + i); }
      const int at = r * 128 + ((c ^ (r & 7)) << 4);
      *reinterpret_cast<uint4*>(Ks + at) = kv;
      *reinterpret_cast<uint4*>(Vs + at) = vv;
    }
  }
  __syncthreads();
  const uint32_t ks_u = (uint32_t)__cvta_generic_to_shared(Ks), vs_u = (uint32_t)__cvta_generic_to_shared(Vs);
  const int nkt = nk16 >> 3;  // key tiles in use (even)
  const float c2 = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) * log2(e)

  const size_t row_base = (size_t)unit * A.nq;
  for (int qt = warp; qt * 16 < A.nq; qt += kAdWarps) {
    const int r0 = qt * 16 + g, r1 = r0 + 8;  // this thread's two query rows inside the unit
    // ---- Q fragments (A operand, 4 k-steps of 16 dims)
    uint32_t qa[4][4];
    {
      const bf16* q0 = A.q + (row_base + r0) * A.ldq + head * kHD + qq * 2;
      const bf16* q1 = A.q + (row_base + r1) * A.ldq + head * kHD + qq * 2;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        qa[ks][0] = r0 < A.nq ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16) : 0u;
        qa[ks][1] = r1 < A.nq ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16) : 0u;
        qa[ks][2] = r0 < A.nq ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16 + 8) : 0u;
        qa[ks][3] = r1 < A.nq ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16 + 8) : 0u;
      }
    }
    // ---- S = Q K^T
    float s[NKT][4];
#pragma unroll
    for (int nt = 0; nt < NKT; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      if (nt < nkt) {
        const int krow = nt * 8 + (lane & 7);
#pragma unroll
        for (int kp = 0; kp < 2; ++kp) {  // dims [32 kp, 32 kp + 32): two k-steps per ldmatrix.x4
          uint32_t b[4];
          const int c = kp * 4 + (lane >> 3);
          ad::ldsm4(ks_u + krow * 128 + ((c ^ (krow & 7)) << 4), b);
          ad::mma16816(s[nt], qa[kp * 2], b[0], b[1]);
          ad::mma16816(s[nt], qa[kp * 2 + 1], b[2], b[3]);
        }
      }
    }
    // ---- mask + softmax (rows r0 and r1; a row's 8 NKT scores are spread over the 4 threads of its quad)
    const int lim0 = A.causal_t ? min(nk - 1, r0 % A.causal_t) : nk - 1;  // last visible key
    const int lim1 = A.causal_t ? min(nk - 1, r1 % A.causal_t) : nk - 1;
    float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < NKT; ++nt) {
      const int key = nt * 8 + qq * 2;
      if (key > lim0) s[nt][0] = -INFINITY;
      if (key + 1 > lim0) s[nt][1] = -INFINITY;
      if (key > lim1) s[nt][2] = -INFINITY;
      if (key + 1 > lim1) s[nt][3] = -INFINITY;
      m0 = fmaxf(m0, fmaxf(s[nt][0], s[nt][1]));
      m1 = fmaxf(m1, fmaxf(s[nt][2], s[nt][3]));
    }
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < NKT; ++nt) {
      s[nt][0] = exp2f((s[nt][0] - m0) * c2); s[nt][1] = exp2f((s[nt][1] - m0) * c2);  // key 0 is always visible: m finite
      s[nt][2] = exp2f((s[nt][2] - m1) * c2); s[nt][3] = exp2f((s[nt][3] - m1) * c2);
      l0 += s[nt][0] + s[nt][1];
      l1 += s[nt][2] + s[nt][3];
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    // ---- O = P V (P as bf16 A fragments straight from the score registers)
    float o[8][4];
#pragma unroll
    for (int dt = 0; dt < 8; ++dt) o[dt][0] = o[dt][1] = o[dt][2] = o[dt][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < NKT / 2; ++kk) {
      if (kk * 2 < nkt) {
        uint32_t pa[4];
        pa[0] = ad::pack_bf16(s[2 * kk][0], s[2 * kk][1]);
        pa[1] = ad::pack_bf16(s[2 * kk][2], s[2 * kk][3]);
        pa[2] = ad::pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
        pa[3] = ad::pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
        const int vrow = kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7);
#pragma unroll
        for (int dp = 0; dp < 4; ++dp) {  // dims [16 dp, 16 dp + 16): two 8-dim tiles per ldmatrix.x4.trans
          uint32_t b[4];
          const int c = dp * 2 + (lane >> 4);
          ad::ldsm4_t(vs_u + vrow * 128 + ((c ^ (vrow & 7)) << 4), b);
          ad::mma16816(o[dp * 2], pa, b[0], b[1]);
          ad::mma16816(o[dp * 2 + 1], pa, b[2], b[3]);
        }
      }
    }
    // ---- normalise and store
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    bf16* o0 = A.out + (row_base + r0) * A.ldo + head * kHD + qq * 2;
    bf16* o1 = A.out + (row_base + r1) * A.ldo + head * kHD + qq * 2;
#pragma unroll
    for (int dt = 0; dt < 8; ++dt) {
      if (r0 < A.nq) *reinterpret_cast<uint32_t*>(o0 + dt * 8) = ad::pack_bf16(o[dt][0] * i0, o[dt][1] * i0);
      if (r1 < A.nq) *reinterpret_cast<uint32_t*>(o1 + dt * 8) = ad::pack_bf16(o[dt][2] * i1, o[dt][3] * i1);
    }
  }
}

constexpr size_t dense_attn_smem(int nkt) { return (size_t)2 * nkt * 8 * 128; }

}  // namespace capdec
