// cross_attn_tma.cuh - cross-attention of the decode step, bf16 engine.
//
// One CTA per (image, head).  The image's K and V tiles for this layer and head ([S,64] bf16,
// contiguous in the cross cache) are staged into shared memory by two TMA tensor loads
// (128-byte swizzle, completion on an mbarrier), then shared by ALL `group` rollouts of the
// image: the K/V stream is read from HBM once per image and step, not once per sequence.
// Math: flash-style, one 16-row query tile (rows >= group are zero) on mma.sync m16n8k16
// (bf16 x bf16 -> fp32); the 4 warps split the keys in 16-key blocks, keep their scores in
// registers (QK^T C-fragment == PV A-fragment), and merge (max, sum, partial O) through shared
// memory.  Probabilities enter the PV product as P_hi + P_lo (two bf16 terms) so that P keeps
// ~16 mantissa bits, matching the fp32 softmax of the reference to well below the bf16 K/V rounding.
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace capdec {

constexpr int kCaThreads = 128;
constexpr int kCaQStride = 72;   // bf16 elements per q row in smem (144 B: conflict-free fragment loads)
constexpr int kCaOStride = 68;   // fp32 elements per partial-output row

__host__ __device__ constexpr int ca_key_rows(int S) { return (S + 15) / 16 * 16; }
// rows of the 16-row query tile that are real (the partial-output scratch is sized by it: greedy decoding
// with one rollout per image needs 1 KB instead of 17 KB, so ~10 instead of 6 CTAs share an SM)
__host__ __device__ constexpr int ca_q_rows(int group) { return group < 16 ? group : 16; }
__host__ __device__ constexpr int ca_smem_bytes(int S, int group) {
  return 1024 + 2 * ca_key_rows(S) * 128 + 16 * kCaQStride * 2 + 4 * ca_q_rows(group) * kCaOStride * 4 + 4 * 16 * 2 * 4 + 64;
}
// Compact form: the warps' partial outputs alias the K tile (dead once every warp has its scores; one query tile per
// item, the K tile large enough) and, when the dynamic shared memory is known to start 1 KB-aligned (probed once per
// process, capdec.cu), no alignment slack: 56.1 KB instead of 63.7 KB at S = 196, so four CTAs share an SM, not three.
__host__ __device__ constexpr bool ca_compact_ok(int S, int group) {
  return group <= 16 && ca_key_rows(S) * 128 >= 4 * ca_q_rows(group) * kCaOStride * 4;
}
__host__ __device__ constexpr int ca_smem_bytes_compact(int S, bool aligned) {
  return (aligned ? 0 : 1024) + 2 * ca_key_rows(S) * 128 + 16 * kCaQStride * 2 + 4 * 16 * 2 * 4 + 64;
}
__global__ void smem_base_probe_kernel(unsigned* out) {
  extern __shared__ uint8_t probe_raw[];
  if (threadIdx.x == 0) *out = (unsigned)(reinterpret_cast<uintptr_t>(probe_raw) & 1023u);
}

namespace ptx {
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
}  // namespace ptx

// tm_kv: 2-D view of the whole cross cache, [L*2*n_images*H*S rows, 64] bf16, box = 64 x ca_key_rows(S),
// SWIZZLE_128B.  q, out: [n_images*group, 512] bf16 image-major.  kKT = max 16-key blocks per warp.
template <int kKT>
__global__ void __launch_bounds__(kCaThreads)
cross_attn_tma_kernel(const __grid_constant__ CUtensorMap tm_kv, const bf16* __restrict__ q, bf16* __restrict__ out,
                      int layer, int n_images, int group, int S, const DecState* st, int compact) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int SB = ca_key_rows(S);
  uint8_t* k_s = smem;
  uint8_t* v_s = smem + (size_t)SB * 128;
  bf16* q_s = reinterpret_cast<bf16*>(v_s + (size_t)SB * 128);
  const int GP = ca_q_rows(group);
  // [4][GP][kCaOStride]; compact: over the K tile, written only after the block-wide barrier that follows the scores
  float* o_part = compact ? reinterpret_cast<float*>(k_s) : reinterpret_cast<float*>(q_s + 16 * kCaQStride);
  float* ml_part = compact ? reinterpret_cast<float*>(q_s + 16 * kCaQStride) : o_part + 4 * GP * kCaOStride;  // [4][16][2]
  uint64_t* bar = reinterpret_cast<uint64_t*>(ml_part + 4 * 16 * 2);  // [2]: K landed, V landed

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ih = blockIdx.x;  // image * H + head
  const int img = ih / kH, h = ih - img * kH;

  if (tid == 0) {
    ptx::prefetch_tensormap(&tm_kv);
    ptx::mbar_init(&bar[0], 1);
    ptx::mbar_init(&bar[1], 1);
    ptx::fence_barrier_init();
    // the projected memory was written at prefill, long before the previous kernel of the step: request the
    // K and V tiles before the programmatic-launch wait; only the queries have to wait for the step's chain.
    // K/V of one step are streamed once: do not let them displace the weights in L2
    const uint64_t pol = ptx::policy_evict_first();
    const int row_k = (((2 * layer + 0) * n_images + img) * kH + h) * S;
    const int row_v = (((2 * layer + 1) * n_images + img) * kH + h) * S;
    ptx::mbar_expect_tx(&bar[0], (uint32_t)SB * 128);
    ptx::tma_load_2d_hint(k_s, &tm_kv, &bar[0], 0, row_k, pol);
    ptx::mbar_expect_tx(&bar[1], (uint32_t)SB * 128);
    ptx::tma_load_2d_hint(v_s, &tm_kv, &bar[1], 0, row_v, pol);
  }
  const bool done = st->done != 0;  // written by the previous step's head kernel, long complete
  __syncthreads();
  ptx::pdl_wait();
  ptx::pdl_trigger();
  if (done) {  // rollout over: nothing to compute, but no copy may outlive the CTA
    if (tid == 0) { ptx::mbar_wait(&bar[0], 0); ptx::mbar_wait(&bar[1], 0); }
    return;
  }

  const int g = lane >> 2, t = lane & 3;
  const int KT = SB / 16;  // 16-key blocks; warp w owns blocks w, w+4, ...
  const uint32_t k_base = ptx::smem_u32(k_s), v_base = ptx::smem_u32(v_s);

  for (int g0 = 0; g0 < group; g0 += 16) {
    const int G = min(16, group - g0);
    const int m0 = img * group + g0;
    if (g0 > 0) __syncthreads();  // previous query tile fully merged before q_s / o_part are reused
    {  // stage the query tile: row = tid/8, 16-byte chunk = tid%8; rows >= G are zero
      const int r = tid >> 3, c = tid & 7;
      uint4 val = make_uint4(0u, 0u, 0u, 0u);
      if (r < G) val = *reinterpret_cast<const uint4*>(q + (size_t)(m0 + r) * kD + h * kHD + c * 8);
      *reinterpret_cast<uint4*>(q_s + r * kCaQStride + c * 8) = val;
    }
    __syncthreads();
    // A fragments of Q for the 4 k-steps over the head dim
    uint32_t qa[4][4];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const bf16* p0 = q_s + g * kCaQStride + kk * 16 + 2 * t;
      qa[kk][0] = *reinterpret_cast<const uint32_t*>(p0);
      qa[kk][1] = *reinterpret_cast<const uint32_t*>(p0 + 8 * kCaQStride);
      qa[kk][2] = *reinterpret_cast<const uint32_t*>(p0 + 8);
      qa[kk][3] = *reinterpret_cast<const uint32_t*>(p0 + 8 * kCaQStride + 8);
    }
    if (g0 == 0) ptx::mbar_wait(&bar[0], 0);

    // ---- scores: sc[i][nt][4] for key block (warp + 4i), n-tile nt (8 keys)
    float sc[kKT][2][4];
    float mx0 = -INFINITY, mx1 = -INFINITY;  // rows g and g+8
#pragma unroll
    for (int i = 0; i < kKT; ++i) {
      const int kb = warp + 4 * i;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int x = 0; x < 4; ++x) sc[i][nt][x] = 0.f;
      if (kb < KT) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          // matrices: (keys +0..7, chunk 2kk), (keys +0..7, chunk 2kk+1), (keys +8..15, chunk 2kk), (keys +8..15, chunk 2kk+1)
          const int key = kb * 16 + (lane & 7) + ((lane >> 4) << 3);
          const int chunk = 2 * kk + ((lane >> 3) & 1);
          uint32_t b[4];
          ptx::ldsm_x4(k_base + key * 128 + ((chunk ^ (key & 7)) << 4), b);
          ptx::mma_bf16_16816(sc[i][0], qa[kk], b[0], b[1]);
          ptx::mma_bf16_16816(sc[i][1], qa[kk], b[2], b[3]);
        }
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int x = 0; x < 4; ++x) {
            const int key = kb * 16 + nt * 8 + 2 * t + (x & 1);
            const float s = key < S ? sc[i][nt][x] * 0.125f : -INFINITY;
            sc[i][nt][x] = s;
            if (x < 2) mx0 = fmaxf(mx0, s); else mx1 = fmaxf(mx1, s);
          }
      } else {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int x = 0; x < 4; ++x) sc[i][nt][x] = -INFINITY;
      }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float sub0 = mx0 == -INFINITY ? 0.f : mx0, sub1 = mx1 == -INFINITY ? 0.f : mx1;

    // ---- probabilities (hi/lo bf16 split) and partial output over this warp's keys
    float l0 = 0.f, l1 = 0.f;
    float o[8][4];
#pragma unroll
    for (int d = 0; d < 8; ++d)
#pragma unroll
      for (int x = 0; x < 4; ++x) o[d][x] = 0.f;
    if (g0 == 0) ptx::mbar_wait(&bar[1], 0);
#pragma unroll
    for (int i = 0; i < kKT; ++i) {
      const int kb = warp + 4 * i;
      if (kb < KT) {
        uint32_t ph[4], pl[4];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const float p0 = __expf(sc[i][nt][0] - sub0), p1 = __expf(sc[i][nt][1] - sub0);
          const float p2 = __expf(sc[i][nt][2] - sub1), p3 = __expf(sc[i][nt][3] - sub1);
          l0 += p0 + p1;
          l1 += p2 + p3;
          const __nv_bfloat162 h01 = __floats2bfloat162_rn(p0, p1), h23 = __floats2bfloat162_rn(p2, p3);
          const __nv_bfloat162 l01 = __floats2bfloat162_rn(p0 - __low2float(h01), p1 - __high2float(h01));
          const __nv_bfloat162 l23 = __floats2bfloat162_rn(p2 - __low2float(h23), p3 - __high2float(h23));
          // A fragment: a0=(g, k 2t..) a1=(g+8, k 2t..) from n-tile 0; a2, a3 from n-tile 1
          ph[nt * 2 + 0] = *reinterpret_cast<const uint32_t*>(&h01);
          ph[nt * 2 + 1] = *reinterpret_cast<const uint32_t*>(&h23);
          pl[nt * 2 + 0] = *reinterpret_cast<const uint32_t*>(&l01);
          pl[nt * 2 + 1] = *reinterpret_cast<const uint32_t*>(&l23);
        }
#pragma unroll
        for (int dp = 0; dp < 4; ++dp) {  // pairs of 8-wide d tiles
          // matrices: (keys +0..7, chunk 2dp), (keys +8..15, chunk 2dp), (keys +0..7, chunk 2dp+1), (keys +8..15, chunk 2dp+1)
          const int key = kb * 16 + (lane & 7) + (((lane >> 3) & 1) << 3);
          const int chunk = 2 * dp + (lane >> 4);
          uint32_t b[4];
          ptx::ldsm_x4_t(v_base + key * 128 + ((chunk ^ (key & 7)) << 4), b);
          ptx::mma_bf16_16816(o[2 * dp], ph, b[0], b[1]);
          ptx::mma_bf16_16816(o[2 * dp], pl, b[0], b[1]);
          ptx::mma_bf16_16816(o[2 * dp + 1], ph, b[2], b[3]);
          ptx::mma_bf16_16816(o[2 * dp + 1], pl, b[2], b[3]);
        }
      }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

    // ---- merge the 4 warps
    if (compact) __syncthreads();  // every warp has left the K tile
    float* op = o_part + (size_t)warp * GP * kCaOStride;
#pragma unroll
    for (int d = 0; d < 8; ++d) {
      if (g < GP) *reinterpret_cast<float2*>(op + g * kCaOStride + d * 8 + 2 * t) = make_float2(o[d][0], o[d][1]);
      if (g + 8 < GP) *reinterpret_cast<float2*>(op + (g + 8) * kCaOStride + d * 8 + 2 * t) = make_float2(o[d][2], o[d][3]);
    }
    if (t == 0) {
      ml_part[(warp * 16 + g) * 2 + 0] = mx0; ml_part[(warp * 16 + g) * 2 + 1] = l0;
      ml_part[(warp * 16 + g + 8) * 2 + 0] = mx1; ml_part[(warp * 16 + g + 8) * 2 + 1] = l1;
    }
    __syncthreads();
    {
      const int r = tid >> 3, c = tid & 7;  // row, 8-wide dim chunk
      if (r < G) {
        float m = -INFINITY;
#pragma unroll
        for (int w = 0; w < 4; ++w) m = fmaxf(m, ml_part[(w * 16 + r) * 2]);
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float l = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const float mw = ml_part[(w * 16 + r) * 2];
          const float scl = mw == -INFINITY ? 0.f : __expf(mw - m);
          l += scl * ml_part[(w * 16 + r) * 2 + 1];
          const float4 a = *reinterpret_cast<const float4*>(o_part + ((size_t)w * GP + r) * kCaOStride + c * 8);
          const float4 b = *reinterpret_cast<const float4*>(o_part + ((size_t)w * GP + r) * kCaOStride + c * 8 + 4);
          acc[0] += scl * a.x; acc[1] += scl * a.y; acc[2] += scl * a.z; acc[3] += scl * a.w;
          acc[4] += scl * b.x; acc[5] += scl * b.y; acc[6] += scl * b.z; acc[7] += scl * b.w;
        }
        const float inv = 1.f / l;
#pragma unroll
        for (int x = 0; x < 8; ++x) acc[x] *= inv;
        store8(out + (size_t)(m0 + r) * kD + h * kHD + c * 8, acc);
      }
    }
  }
}

}  // namespace capdec
