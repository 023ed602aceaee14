// rollout_cluster.cuh - the whole decode loop as ONE kernel: a thread-block cluster of 8 CTAs owns
// a block of RB sequences for every layer of every step of the rollout chunk.
//
// Why: with a few hundred live sequences every projection of the step is a latency-bound kernel
// (launch + prologue + one TMA round trip + drain ~ 9 us, 50 of them per step).  Sequences never
// interact, so all dependencies of a decode step are local to the rows' owner.  Here the owner is
// a cluster: CTA c (= attention head c) holds output features [64c, 64c+64) of every projection,
// streams ITS eighth of the layer's weights through a TMA ring (5.5 MB per step per CTA, laid out
// in HBM in streaming order, L2-resident across clusters), multiplies on tcgen05 (M = 64 weight
// rows x N = RB sequences, accumulators in TMEM) and exchanges activations with its 7 peers through
// distributed shared memory:
//   QKV (head c) -> self-attention over the global KV cache -> all-gather of the head outputs
//   -> out-proj + residual + LayerNorm (row statistics all-gathered) -> cross-q -> cross-attention
//   (K/V tiles by TMA, mma.sync flash kernel) -> all-gather -> out-proj + LN -> FFN1 (256 hidden
//   features per CTA, stay in this CTA's smem) -> FFN2 partial over that K slice -> reduce-scatter
//   -> LN -> ... -> vocabulary slice -> all-gather logits -> argmax / Philox sample -> embedding.
// No grid-wide barrier, no kernel boundary, no host round trip inside a chunk of steps.
//
// Warp roles (320 threads): warp 0 weight producer (cp.async.bulk, 8 KB units), warp 1 MMA issuer,
// warps 2-9 compute (TMEM epilogues, attention, LayerNorm, head).  Only the compute warps take
// part in the cluster barrier (an mbarrier with one remote arrive per CTA).
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "cross_attn_tma.cuh"

namespace capdec {

constexpr int kRcCluster = 8;
constexpr int kRcThreads = 320;
constexpr int kRcCompute = 256;          // warps 2..9
constexpr int kRcUnitBytes = 8192;       // 64 weight rows x 64 k, bf16, 128B-swizzled image
constexpr int kRcUnitsPerLayer = 112;    // qkv 24 + so 8 + cq 8 + co 8 + ffn1 32 + ffn2 32  (d_ff = 2048)
constexpr int kRcUnitsHead = 8;          // vocabulary tile (64 padded rows) x 8 k-blocks, after the last layer
constexpr int kRcVocPerCta = 16;         // vocabulary entries per CTA in the head (8 x 16 = kVPad)

// Per-layer fp32 vectors the epilogues read (device pointers).
struct RcLayer {
  const float *b_qkv, *b_so, *b_cq, *b_co, *b_1, *b_2;
  const float *ln_w[3], *ln_b[3];
  bf16 *kcache, *vcache;  // self K/V cache of this layer, [M][H][t_pad][64]
};

struct RcParams {
  RcLayer layer[CAPDEC_MAX_LAYERS];
  const uint8_t* wstream;   // [8 ctas][L*112 + 8][8192]: per-CTA weight units in consumption order
  const float *emb, *pe;    // [V,512], [pe_rows,512]
  const float *fc_b;        // [kVPad] (entries >= V are zero)
  const CallParams* cp;
  unsigned char* fin;       // [M] sampled row already finished
  int* counters;            // [max_len][4]: per step {rows reported, greedy rows that emitted <end>, sampled rows finished}
  const DecState* st;       // st->done: the rollout already met its stop rule in an earlier chunk
  long long* trace;         // optional: clock64 stamps of CTA 0 / block 0, second step of the chunk, layer 1 (tuning)
  int n_layer, M, group, S, t_pad, vocab, n_images;
  int step_begin, step_end;
};

template <int RB> struct RcSmem {
  static constexpr int kBTile = RB * 128;                 // one k-block of a B operand (RB rows x 64 k bf16)
  static constexpr int kXb = 0;                           // 8 k-blocks
  static constexpr int kA = kXb + 8 * kBTile;             // 8 k-blocks
  static constexpr int kH = kA + 8 * kBTile;              // 4 k-blocks
  static constexpr int kKv = kH + 4 * kBTile;             // 4 slots x (K 8 KB + V 8 KB)   (S <= 64)
  static constexpr int kU = kKv + 4 * 16384;              // union: red_s [8][RB][64] f32 | cross-attention scratch (2 groups)
  static constexpr int kUBytes = 8 * RB * 64 * 4;
  static constexpr int kXs = kU + (kUBytes > 2 * 20480 ? kUBytes : 2 * 20480);  // x_s  [RB][64] f32 (own residual slice)
  static constexpr int kVs = kXs + RB * 64 * 4;           // vs_s [RB][64] f32 staging (TMEM tile / LN input)
  static constexpr int kQf = kVs + RB * 64 * 4;           // qf_s [RB][64] f32 (self-attention query, pre-scaled)
  static constexpr int kQb = kQf + RB * 64 * 4;           // qb_s [RB][72] bf16 (cross-attention query)
  static constexpr int kStage = kQb + RB * kCaQStride * 2;  // stage_s [RB][64] bf16 (slice to publish)
  static constexpr int kStat = kStage + RB * 64 * 2;      // stat_s [8][RB][2] f32
  static constexpr int kLg = kStat + 8 * RB * 2 * 4;      // lg_s [RB][kVPad] f32
  static constexpr int kTok = kLg + RB * kVPad * 4;       // tok_s [RB] int
  static constexpr int kBars = kTok + RB * 4 + 64;        // mbarriers (8-byte aligned)
  static constexpr int kRing = (kBars + 512 + 1023) / 1024 * 1024;
  static constexpr int bytes(int stages) { return kRing + stages * kRcUnitBytes + 1024; }
};

namespace ptx {
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void bulk_load_1d_hint(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(gmem_src)), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait_cluster(bar, parity)) return;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  uint32_t spins = 0;
  while (!mbar_try_wait_cluster(bar, parity)) {
    if ((++spins & 0x3ff) == 0) {
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 4000000000ull) {
        printf("capdec: cluster barrier timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
        __trap();
      }
    }
  }
}
__device__ __forceinline__ void st_cluster_u4(uint32_t addr, uint4 v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_cluster_f2(uint32_t addr, float a, float b) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// kind::f16 instruction descriptor, D=f32, A=B=bf16 K-major, M=64, N=n
__device__ __host__ constexpr uint32_t umma_idesc_bf16_m64(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(64 >> 4) << 24);
}
}  // namespace ptx

// byte offset of element (row j, k index kk in [0,64)) inside one 128B-swizzled K-major tile
__device__ __forceinline__ int rc_boff(int j, int kk) { return j * 128 + ((((kk) >> 3) ^ (j & 7)) << 4) + ((kk) & 7) * 2; }


// ---------------------------------------------------------------- cross-attention of one image tile
// Same math as cross_attn_tma_kernel (kKT = 1: S <= 64, one 16-key block per warp), run by a group of
// 4 warps on a K/V tile already staged in shared memory.  q rows: qb[g][kCaQStride] bf16, rows >= G
// are treated as zero.  Output rows g < G go to out[g][64] bf16 (the slice that is then published).
__device__ __forceinline__ void rc_cross_tile(uint32_t k_base, uint32_t v_base, int S, int KT, const bf16* qb, int G,
                                              bf16* out, float* o_part, float* ml_part, int gw, int lane, int gtid,
                                              int bar_id) {
  const int g = lane >> 2, t = lane & 3;
  uint32_t qa[4][4];
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    const bf16* p0 = qb + g * kCaQStride + kk * 16 + 2 * t;
    qa[kk][0] = g < G ? *reinterpret_cast<const uint32_t*>(p0) : 0u;
    qa[kk][1] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(p0 + 8 * kCaQStride) : 0u;
    qa[kk][2] = g < G ? *reinterpret_cast<const uint32_t*>(p0 + 8) : 0u;
    qa[kk][3] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(p0 + 8 * kCaQStride + 8) : 0u;
  }
  float sc[2][4];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int x = 0; x < 4; ++x) sc[nt][x] = 0.f;
  float mx0 = -INFINITY, mx1 = -INFINITY;
  const int kb = gw;  // this warp's 16-key block
  if (kb < KT) {
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const int key = kb * 16 + (lane & 7) + ((lane >> 4) << 3);
      const int chunk = 2 * kk + ((lane >> 3) & 1);
      uint32_t b[4];
      ptx::ldsm_x4(k_base + key * 128 + ((chunk ^ (key & 7)) << 4), b);
      ptx::mma_bf16_16816(sc[0], qa[kk], b[0], b[1]);
      ptx::mma_bf16_16816(sc[1], qa[kk], b[2], b[3]);
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const int key = kb * 16 + nt * 8 + 2 * t + (x & 1);
        const float s = key < S ? sc[nt][x] : -INFINITY;  // q already carries the 1/8 scale
        sc[nt][x] = s;
        if (x < 2) mx0 = fmaxf(mx0, s); else mx1 = fmaxf(mx1, s);
      }
  } else {
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int x = 0; x < 4; ++x) sc[nt][x] = -INFINITY;
  }
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
  const float sub0 = mx0 == -INFINITY ? 0.f : mx0, sub1 = mx1 == -INFINITY ? 0.f : mx1;
  float l0 = 0.f, l1 = 0.f;
  float o[8][4];
#pragma unroll
  for (int d = 0; d < 8; ++d)
#pragma unroll
    for (int x = 0; x < 4; ++x) o[d][x] = 0.f;
  if (kb < KT) {
    uint32_t ph[4], pl[4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const float p0 = __expf(sc[nt][0] - sub0), p1 = __expf(sc[nt][1] - sub0);
      const float p2 = __expf(sc[nt][2] - sub1), p3 = __expf(sc[nt][3] - sub1);
      l0 += p0 + p1;
      l1 += p2 + p3;
      const __nv_bfloat162 h01 = __floats2bfloat162_rn(p0, p1), h23 = __floats2bfloat162_rn(p2, p3);
      const __nv_bfloat162 l01 = __floats2bfloat162_rn(p0 - __low2float(h01), p1 - __high2float(h01));
      const __nv_bfloat162 l23 = __floats2bfloat162_rn(p2 - __low2float(h23), p3 - __high2float(h23));
      ph[nt * 2 + 0] = *reinterpret_cast<const uint32_t*>(&h01);
      ph[nt * 2 + 1] = *reinterpret_cast<const uint32_t*>(&h23);
      pl[nt * 2 + 0] = *reinterpret_cast<const uint32_t*>(&l01);
      pl[nt * 2 + 1] = *reinterpret_cast<const uint32_t*>(&l23);
    }
#pragma unroll
    for (int dp = 0; dp < 4; ++dp) {
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
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  float* op = o_part + (size_t)gw * 16 * kCaOStride;
#pragma unroll
  for (int d = 0; d < 8; ++d) {
    *reinterpret_cast<float2*>(op + g * kCaOStride + d * 8 + 2 * t) = make_float2(o[d][0], o[d][1]);
    *reinterpret_cast<float2*>(op + (g + 8) * kCaOStride + d * 8 + 2 * t) = make_float2(o[d][2], o[d][3]);
  }
  if (t == 0) {
    ml_part[(gw * 16 + g) * 2 + 0] = mx0; ml_part[(gw * 16 + g) * 2 + 1] = l0;
    ml_part[(gw * 16 + g + 8) * 2 + 0] = mx1; ml_part[(gw * 16 + g + 8) * 2 + 1] = l1;
  }
  ptx::bar_sync(bar_id, 128);
  {
    const int r = gtid >> 3, c8 = gtid & 7;
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
        const float4 a = *reinterpret_cast<const float4*>(o_part + ((size_t)w * 16 + r) * kCaOStride + c8 * 8);
        const float4 b = *reinterpret_cast<const float4*>(o_part + ((size_t)w * 16 + r) * kCaOStride + c8 * 8 + 4);
        acc[0] += scl * a.x; acc[1] += scl * a.y; acc[2] += scl * a.z; acc[3] += scl * a.w;
        acc[4] += scl * b.x; acc[5] += scl * b.y; acc[6] += scl * b.z; acc[7] += scl * b.w;
      }
      const float inv = 1.f / l;
#pragma unroll
      for (int x = 0; x < 8; ++x) acc[x] *= inv;
      store8(out + r * 64 + c8 * 8, acc);
    }
  }
  ptx::bar_sync(bar_id, 128);  // o_part / ml_part and the K/V slot may be reused
}


namespace ptx {
__device__ __forceinline__ uint32_t map_addr_to_cta(uint32_t cta_addr, uint32_t rank) {
  uint32_t out;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(cta_addr), "r"(rank));
  return out;
}
}  // namespace ptx

// grid = 8 * ceil(M / RB) CTAs, cluster (8,1,1); requires S <= 64, t_pad == 64, RB % group == 0, d_ff == 2048.
template <int RB>
__global__ void __cluster_dims__(kRcCluster, 1, 1) __launch_bounds__(kRcThreads, 1)
rollout_cluster_kernel(const __grid_constant__ CUtensorMap tm_kv, const __grid_constant__ RcParams P, int stages) {
  using SM = RcSmem<RB>;
  constexpr int EPT = RB / 4;        // LayerNorm: elements per compute thread (RB*64 / 256)
  constexpr int TPR = 64 / EPT;      // threads per row
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  if (P.st->done) return;  // uniform over the grid: nothing has been initialised yet

  float* x_s = reinterpret_cast<float*>(smem + SM::kXs);
  float* vs_s = reinterpret_cast<float*>(smem + SM::kVs);
  float* qf_s = reinterpret_cast<float*>(smem + SM::kQf);
  bf16* qb_s = reinterpret_cast<bf16*>(smem + SM::kQb);
  bf16* stage_s = reinterpret_cast<bf16*>(smem + SM::kStage);
  float* stat_s = reinterpret_cast<float*>(smem + SM::kStat);
  float* lg_s = reinterpret_cast<float*>(smem + SM::kLg);
  int* tok_s = reinterpret_cast<int*>(smem + SM::kTok);
  float* red_s = reinterpret_cast<float*>(smem + SM::kU);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::kBars);
  uint64_t* full = bars;             // [stages] (<= 16)
  uint64_t* empty = bars + 16;       // [stages]
  uint64_t* op_ready = bars + 32;    // compute -> MMA: B operand of the next GEMM is in place, TMEM drained
  uint64_t* acc_full = bars + 33;    // MMA -> compute
  uint64_t* kv_full = bars + 34;     // [4] cross K/V slots
  uint64_t* cb = bars + 38;          // cluster barrier: one arrive per CTA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 39);
  uint8_t* ring = smem + SM::kRing;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t c = ptx::cluster_ctarank();
  const int blk = blockIdx.x / kRcCluster;
  const int r0 = blk * RB;
  const int L = P.n_layer;
  const int n_steps = P.step_end - P.step_begin;
  constexpr int kTmemCols = 8 * RB < 32 ? 32 : 8 * RB;

  if (tid == 0) {
    ptx::prefetch_tensormap(&tm_kv);
    for (int s = 0; s < stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::mbar_init(op_ready, 1);
    ptx::mbar_init(acc_full, 1);
    for (int s = 0; s < 4; ++s) ptx::mbar_init(&kv_full[s], 1);
    ptx::mbar_init(cb, kRcCluster);
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // every CTA's barriers are initialised before any peer may arrive on them / store into its smem
  ptx::cluster_sync();

  if (warp == 0) {
    // ===================== weight producer: one linear stream of 8 KB units per step =====================
    if (lane == 0) {
      const uint64_t keep = ptx::policy_evict_last();
      const int units_per_step = L * kRcUnitsPerLayer + kRcUnitsHead;
      const uint8_t* src0 = P.wstream + (size_t)c * units_per_step * kRcUnitBytes;
      uint32_t U = 0;
      for (int n = 0; n < n_steps; ++n)
        for (int u = 0; u < units_per_step; ++u, ++U) {
          const int s = U % stages;
          ptx::mbar_wait(&empty[s], ((U / stages) & 1) ^ 1);
          ptx::mbar_expect_tx(&full[s], kRcUnitBytes);
          ptx::bulk_load_1d_hint(ring + (size_t)s * kRcUnitBytes, src0 + (size_t)u * kRcUnitBytes, kRcUnitBytes, &full[s], keep);
        }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16_m64(RB);
      uint32_t U = 0, par = 0;
      auto gemm = [&](int tiles, int kblocks, int b_off) {
        ptx::mbar_wait(op_ready, par);
        par ^= 1;
        ptx::tc_fence_after();
        const uint32_t b_base = ptx::smem_u32(smem + b_off);
        for (int t = 0; t < tiles; ++t)
          for (int kb = 0; kb < kblocks; ++kb, ++U) {
            const int s = U % stages;
            ptx::mbar_wait(&full[s], (U / stages) & 1);
            ptx::tc_fence_after();
            const uint32_t a_addr = ptx::smem_u32(ring + (size_t)s * kRcUnitBytes);
            const uint32_t b_addr = b_base + kb * SM::kBTile;
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f16(tmem_base + (uint32_t)(t * RB), ptx::umma_desc_sw128(a_addr + k * 32),
                            ptx::umma_desc_sw128(b_addr + k * 32), idesc, (kb | k) != 0 ? 1u : 0u);
            ptx::umma_commit(&empty[s]);
          }
        ptx::umma_commit(acc_full);
      };
      for (int n = 0; n < n_steps; ++n) {
        for (int l = 0; l < L; ++l) {
          gemm(3, 8, SM::kXb);  // q, k, v of head c
          gemm(1, 8, SM::kA);   // self out-proj
          gemm(1, 8, SM::kXb);  // cross q
          gemm(1, 8, SM::kA);   // cross out-proj
          gemm(4, 8, SM::kXb);  // FFN1: hidden [256c, 256c+256)
          gemm(8, 4, SM::kH);   // FFN2 partial over that hidden slice
        }
        gemm(1, 8, SM::kXb);    // vocabulary tile (c & 1)
      }
    }
  } else {
    // ===================== compute warps =====================
    const int ct = tid - 64, cw = ct >> 5;
    const CallParams cp = *P.cp;
    const int G = P.group, S = P.S, SB = ca_key_rows(S), KT = SB / 16;
    const int M = P.M;
    uint32_t cpar = 0, apar = 0;
    // TMEM epilogue identity: warps cw < 4 hold one 64-row tile, lanes 0..15 of each lane quarter
    const int qrt = warp & 3;
    const bool epi_warp = cw < 4, epi_lane = epi_warp && lane < 16;
    const int em = qrt * 16 + lane;  // tile row (feature within the CTA's 64-slice) of this thread
    const uint32_t t_lane = tmem_base + ((uint32_t)(qrt * 32) << 16);
    // LayerNorm / elementwise identity: row lj, features lm0 .. lm0+EPT-1 of the slice
    const int lj = ct / TPR, lm0 = (ct % TPR) * EPT;
    // cross-attention groups: compute warps 0-3 and 4-7
    const int gq = cw >> 2, gw = cw & 3, gtid = ct & 127;
    const int img0 = r0 / G;
    const int NI = min(RB / G, P.n_images - img0);         // images of this block
    const int NIg = (NI - gq + 1) / 2;                      // tiles per layer of this group (i = gq, gq+2, ..)
    uint8_t* kv_slots = smem + SM::kKv + gq * 2 * 16384;    // this group's two slots
    float* o_part = reinterpret_cast<float*>(smem + SM::kU + gq * 20480);
    float* ml_part = o_part + 4 * 16 * kCaOStride;
    const uint64_t kv_pol = ptx::policy_evict_first();
    long long kv_issued = 0, kv_done = 0;                   // per-group tile counters (same in all its threads)
    const long long kv_total = (long long)n_steps * L * NIg;
    auto kv_issue = [&](long long e) {  // elected thread: tile e of this group's sequence -> slot e & 1
      const int within = (int)(e % ((long long)L * NIg));
      const int l = within / NIg, i = gq + 2 * (within % NIg);
      const int img = img0 + i;
      uint8_t* slot = kv_slots + (e & 1) * 16384;
      uint64_t* bar = &kv_full[gq * 2 + (e & 1)];
      ptx::mbar_expect_tx(bar, (uint32_t)(2 * SB * 128));
      ptx::tma_load_2d_hint(slot, &tm_kv, bar, 0, (((2 * l + 0) * P.n_images + img) * kH + (int)c) * S, kv_pol);
      ptx::tma_load_2d_hint(slot + 8192, &tm_kv, bar, 0, (((2 * l + 1) * P.n_images + img) * kH + (int)c) * S, kv_pol);
    };
    if (gtid == 0) { for (; kv_issued < 2 && kv_issued < kv_total; ++kv_issued) kv_issue(kv_issued); }

    auto lsync = [&]() {  // compute-wide barrier; generic smem writes become visible to the tensor core
      ptx::tc_fence_before();
      ptx::fence_proxy_async_all();
      ptx::bar_sync(1, kRcCompute);
    };
    auto csync = [&]() {  // cluster-wide barrier of the compute warps (DSMEM stores before it are visible after it)
      lsync();
      if (ct == 0) {
#pragma unroll
        for (uint32_t d = 0; d < kRcCluster; ++d) ptx::mbar_arrive_remote(ptx::map_to_cta(cb, d));
      }
      ptx::mbar_wait_cluster(cb, cpar);
      cpar ^= 1;
    };
    auto signal_mma = [&]() { if (ct == 0) ptx::mbar_arrive(op_ready); };  // call right after lsync()/csync()
    auto wait_acc = [&]() { ptx::mbar_wait(acc_full, apar); apar ^= 1; ptx::tc_fence_after(); };
    int tr_n = 0;
    bool tr_on = false;
    auto stamp = [&]() { if (tr_on && tr_n < 64) P.trace[tr_n++] = clock64(); };
    // stage_s [RB][64] bf16 -> k-block c of the B-operand buffer at smem offset buf_off, in all 8 CTAs
    auto publish = [&](int buf_off) {
      for (int idx = ct; idx < RB * 8; idx += kRcCompute) {
        const int j = idx >> 3, ch = idx & 7;
        const uint4 val = *reinterpret_cast<const uint4*>(stage_s + j * 64 + ch * 8);
        const uint32_t local = ptx::smem_u32(smem + buf_off + (int)c * SM::kBTile + j * 128 + ((ch ^ (j & 7)) << 4));
#pragma unroll
        for (uint32_t d = 0; d < kRcCluster; ++d) ptx::st_cluster_u4(ptx::map_addr_to_cta(local, d), val);
      }
    };
    // x = emb[tok] * sqrt(512) + pe[pos]: full bf16 row into xb_s (local), own 64-feature slice into x_s
    auto embed = [&](int pos) {
      const float sc = 22.627416997969522f;
      for (int idx = ct; idx < RB * 128; idx += kRcCompute) {
        const int j = idx >> 7, f = (idx & 127) * 4;
        const int tok = tok_s[j];
        const float4 e = __ldg(reinterpret_cast<const float4*>(P.emb + (size_t)tok * kD + f));
        const float4 pp = __ldg(reinterpret_cast<const float4*>(P.pe + (size_t)pos * kD + f));
        const float r0f = __fadd_rn(__fmul_rn(e.x, sc), pp.x), r1f = __fadd_rn(__fmul_rn(e.y, sc), pp.y);
        const float r2f = __fadd_rn(__fmul_rn(e.z, sc), pp.z), r3f = __fadd_rn(__fmul_rn(e.w, sc), pp.w);
        uint2 pk;
        pk.x = pack_bf16x2(r0f, r1f);
        pk.y = pack_bf16x2(r2f, r3f);
        *reinterpret_cast<uint2*>(smem + SM::kXb + (f >> 6) * SM::kBTile + rc_boff(j, f & 63)) = pk;
        if ((f >> 6) == (int)c) *reinterpret_cast<float4*>(x_s + j * 64 + (f & 63)) = make_float4(r0f, r1f, r2f, r3f);
      }
    };
    // vs_s holds (projection + bias + residual) of this CTA's slice: LayerNorm over the 512 features of each
    // row (statistics all-gathered), result to x_s (fp32) and, as bf16, to k-block c of xb_s everywhere.
    auto layer_norm = [&](float (&v)[EPT], const float* gamma, const float* beta) {
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int e = 0; e < EPT; ++e) { s1 += v[e]; s2 = fmaf(v[e], v[e], s2); }
#pragma unroll
      for (int o = 1; o < TPR; o <<= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      }
      if ((ct % TPR) == 0) {
        const uint32_t local = ptx::smem_u32(stat_s + ((int)c * RB + lj) * 2);
#pragma unroll
        for (uint32_t d = 0; d < kRcCluster; ++d) ptx::st_cluster_f2(ptx::map_addr_to_cta(local, d), s1, s2);
      }
      csync();
      float t1 = 0.f, t2 = 0.f;
#pragma unroll
      for (int src = 0; src < kRcCluster; ++src) { t1 += stat_s[(src * RB + lj) * 2]; t2 += stat_s[(src * RB + lj) * 2 + 1]; }
      const float mean = t1 * (1.f / kD);
      const float rstd = rsqrtf(fmaxf(t2 * (1.f / kD) - mean * mean, 0.f) + 1e-5f);
#pragma unroll
      for (int e = 0; e < EPT; ++e) {
        const int f = (int)c * 64 + lm0 + e;
        const float y = (v[e] - mean) * rstd * __ldg(gamma + f) + __ldg(beta + f);
        x_s[lj * 64 + lm0 + e] = y;
        stage_s[lj * 64 + lm0 + e] = __float2bfloat16_rn(y);
      }
      lsync();
      publish(SM::kXb);
      csync();
    };

    // ---- chunk start: tokens consumed by the first step
    if (ct < RB) tok_s[ct] = (r0 + ct < M) ? (int)cp.ids[(size_t)(r0 + ct) * cp.max_len + P.step_begin] : 0;
    ptx::bar_sync(1, kRcCompute);
    embed(P.step_begin);
    lsync();
    signal_mma();

    for (int n = P.step_begin; n < P.step_end; ++n) {
      for (int l = 0; l < L; ++l) {
        const RcLayer& lp = P.layer[l];
        tr_on = P.trace != nullptr && ct == 0 && blockIdx.x == 0 && n == P.step_begin + 1 && l == (L > 1 ? 1 : 0);
        stamp();  // 0: layer start
        // ---------- P1: q, k, v of head c ----------
        wait_acc();
        stamp();  // 1: qkv accumulators ready
        if (epi_warp) {
#pragma unroll
          for (int t = 0; t < 3; ++t) {
            const float b = __ldg(lp.b_qkv + t * kD + (int)c * 64 + (em & 63));
#pragma unroll
            for (int cc = 0; cc < RB / 16; ++cc) {
              float v[16];
              ptx::tmem_ld16(t_lane + (uint32_t)(t * RB + cc * 16), v);
              if (epi_lane) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const int row = cc * 16 + j;
                  const float val = v[j] + b;
                  if (t == 0) {
                    qf_s[row * 64 + em] = val * 0.125f;
                  } else if (r0 + row < M) {
                    bf16* cache = t == 1 ? lp.kcache : lp.vcache;
                    cache[(((size_t)(r0 + row) * kH + c) * P.t_pad + n) * kHD + em] = __float2bfloat16_rn(val);
                  }
                }
              }
            }
          }
        }
        __threadfence_block();
        lsync();
        stamp();  // 2: qkv epilogue done
        // ---------- P2: self-attention of (row, head c) over cache positions 0..n ----------
        for (int j = cw; j < RB; j += 8) {
          const int g = lane >> 3, cc = lane & 7;
          float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          if (r0 + j < M) {
            const size_t base = (((size_t)(r0 + j) * kH + c) * P.t_pad) * kHD + cc * 8;
            const bf16* kb = lp.kcache + base;
            const bf16* vb = lp.vcache + base;
            float qv[8];
            load8s(qf_s + j * 64 + cc * 8, qv);
            float sc[16];
#pragma unroll
            for (int it = 0; it < 16; ++it) {
              const int pos = it * 4 + g;
              float d = 0.f;
              if (it * 4 <= n) {
                if (pos <= n) {
                  float kv[8];
                  const uint4 raw = *reinterpret_cast<const uint4*>(kb + (size_t)pos * kHD);
                  const uint32_t w4[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
                  for (int i = 0; i < 4; ++i) { kv[2 * i] = __uint_as_float(w4[i] << 16); kv[2 * i + 1] = __uint_as_float(w4[i] & 0xffff0000u); }
#pragma unroll
                  for (int i = 0; i < 8; ++i) d = fmaf(qv[i], kv[i], d);
                }
                d += __shfl_xor_sync(0xffffffffu, d, 1);
                d += __shfl_xor_sync(0xffffffffu, d, 2);
                d += __shfl_xor_sync(0xffffffffu, d, 4);
              }
              sc[it] = (pos <= n) ? d : -INFINITY;
            }
            float mx = sc[0];
#pragma unroll
            for (int it = 1; it < 16; ++it) mx = fmaxf(mx, sc[it]);
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
            float sum = 0.f;
#pragma unroll
            for (int it = 0; it < 16; ++it) {
              const int pos = it * 4 + g;
              if (it * 4 <= n && pos <= n) {
                const float p = __expf(sc[it] - mx);
                sum += p;
                const uint4 raw = *reinterpret_cast<const uint4*>(vb + (size_t)pos * kHD);
                const uint32_t w4[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  acc[2 * i] = fmaf(p, __uint_as_float(w4[i] << 16), acc[2 * i]);
                  acc[2 * i + 1] = fmaf(p, __uint_as_float(w4[i] & 0xffff0000u), acc[2 * i + 1]);
                }
              }
            }
            sum += __shfl_xor_sync(0xffffffffu, sum, 8);
            sum += __shfl_xor_sync(0xffffffffu, sum, 16);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
              acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
            }
            const float inv = 1.f / sum;
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] *= inv;
          }
          if (g == 0) store8(stage_s + j * 64 + cc * 8, acc);
        }
        lsync();
        stamp();  // 3 / 9: attention done
        publish(SM::kA);
        csync();
        stamp();  // 4 / 10: heads all-gathered
        signal_mma();
        // ---------- P3 / P6 helper: out-proj tile + bias + residual -> vs_s ----------
        auto outproj_to_vs = [&](const float* bias) {
          wait_acc();
          if (epi_warp) {
            const float b = __ldg(bias + (int)c * 64 + (em & 63));
#pragma unroll
            for (int cc = 0; cc < RB / 16; ++cc) {
              float v[16];
              ptx::tmem_ld16(t_lane + (uint32_t)(cc * 16), v);
              if (epi_lane) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const int row = cc * 16 + j;
                  vs_s[row * 64 + em] = v[j] + b + x_s[row * 64 + em];
                }
              }
            }
          }
          lsync();
        };
        {  // ---------- P3: self out-proj + LayerNorm 1 ----------
          outproj_to_vs(lp.b_so);
          stamp();  // 5: self out-proj in vs_s
          float v[EPT];
#pragma unroll
          for (int e = 0; e < EPT; ++e) v[e] = vs_s[lj * 64 + lm0 + e];
          layer_norm(v, lp.ln_w[0], lp.ln_b[0]);
          signal_mma();
          stamp();  // 6: LN1 done (2 cluster barriers)
        }
        // ---------- P4: cross-attention query of head c ----------
        wait_acc();
        if (epi_warp) {
          const float b = __ldg(lp.b_cq + (int)c * 64 + (em & 63));
#pragma unroll
          for (int cc = 0; cc < RB / 16; ++cc) {
            float v[16];
            ptx::tmem_ld16(t_lane + (uint32_t)(cc * 16), v);
            if (epi_lane) {
#pragma unroll
              for (int j = 0; j < 16; ++j) qb_s[(cc * 16 + j) * kCaQStride + em] = __float2bfloat16_rn((v[j] + b) * 0.125f);
            }
          }
        }
        stamp();  // 7: cross q ready
        for (int idx = ct; idx < RB * 8; idx += kRcCompute)  // rows without an image publish zeros
          if ((idx >> 3) >= NI * G) *reinterpret_cast<uint4*>(stage_s + idx * 8) = make_uint4(0u, 0u, 0u, 0u);
        lsync();
        stamp();  // 8: cross q epilogue done
        // ---------- P5: cross-attention, one (image, head c) K/V tile per 4-warp group at a time ----------
        for (int ii = 0; ii < NIg; ++ii) {
          const int i = gq + 2 * ii;
          const long long e = kv_done;
          ptx::mbar_wait(&kv_full[gq * 2 + (e & 1)], (uint32_t)((e >> 1) & 1));
          const uint32_t k_base = ptx::smem_u32(kv_slots + (e & 1) * 16384);
          rc_cross_tile(k_base, k_base + 8192, S, KT, qb_s + (size_t)i * G * kCaQStride, G, stage_s + (size_t)i * G * 64,
                        o_part, ml_part, gw, lane, gtid, 2 + gq);
          ++kv_done;
          if (gtid == 0 && kv_issued < kv_total) { kv_issue(kv_issued); }
          if (kv_issued < kv_total) ++kv_issued;
        }
        lsync();
        stamp();  // 3 / 9: attention done
        publish(SM::kA);
        csync();
        stamp();  // 4 / 10: heads all-gathered
        signal_mma();
        {  // ---------- P6: cross out-proj + LayerNorm 2 ----------
          outproj_to_vs(lp.b_co);
          stamp();  // 11: cross out-proj in vs_s
          float v[EPT];
#pragma unroll
          for (int e = 0; e < EPT; ++e) v[e] = vs_s[lj * 64 + lm0 + e];
          layer_norm(v, lp.ln_w[1], lp.ln_b[1]);
          signal_mma();
          stamp();  // 12: LN2 done
        }
        // ---------- P7: FFN1, hidden features [256c, 256c+256) stay in this CTA as FFN2's K slice ----------
        wait_acc();
        stamp();  // 13: FFN1 accumulators ready
        if (epi_warp) {
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const float b = __ldg(lp.b_1 + (int)c * 256 + t * 64 + (em & 63));
#pragma unroll
            for (int cc = 0; cc < RB / 16; ++cc) {
              float v[16];
              ptx::tmem_ld16(t_lane + (uint32_t)(t * RB + cc * 16), v);
              if (epi_lane) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const int row = cc * 16 + j;
                  *reinterpret_cast<bf16*>(smem + SM::kH + t * SM::kBTile + rc_boff(row, em)) =
                      __float2bfloat16_rn(fmaxf(v[j] + b, 0.f));
                }
              }
            }
          }
        }
        lsync();
        signal_mma();
        stamp();  // 14: FFN1 epilogue done
        // ---------- P8: FFN2 partial sums, reduce-scatter: tile t belongs to CTA t ----------
        wait_acc();
        stamp();  // 15: FFN2 accumulators ready
        if (epi_warp) {
#pragma unroll 1
          for (int t = 0; t < 8; ++t) {
#pragma unroll
            for (int cc = 0; cc < RB / 16; ++cc) {
              float v[16];
              ptx::tmem_ld16(t_lane + (uint32_t)(t * RB + cc * 16), v);
              if (epi_lane) {
                const uint32_t local = ptx::smem_u32(red_s + ((int)c * RB + cc * 16) * 64 + em);
                const uint32_t remote = ptx::map_addr_to_cta(local, (uint32_t)t);
#pragma unroll
                for (int j = 0; j < 16; ++j) ptx::st_cluster_f32(remote + j * 64 * 4, v[j]);
              }
            }
          }
        }
        csync();
        stamp();  // 16: reduce-scatter done
        {  // ---------- LayerNorm 3 on (sum of the 8 partials + bias + residual) ----------
          float v[EPT];
#pragma unroll
          for (int e = 0; e < EPT; ++e) v[e] = __ldg(lp.b_2 + (int)c * 64 + lm0 + e) + x_s[lj * 64 + lm0 + e];
#pragma unroll
          for (int src = 0; src < kRcCluster; ++src)
#pragma unroll
            for (int e = 0; e < EPT; ++e) v[e] += red_s[(src * RB + lj) * 64 + lm0 + e];
          layer_norm(v, lp.ln_w[2], lp.ln_b[2]);
          signal_mma();
          stamp();  // 17: LN3 done
        }
      }
      // ---------- head: vocabulary tile (c & 1) on the tensor core; CTAs 0 and 1 all-gather the logits ----------
      wait_acc();
      if (epi_warp) {
        const int vv = (int)(c & 1) * 64 + (em & 63);
        const float b = __ldg(P.fc_b + vv);
#pragma unroll
        for (int cc = 0; cc < RB / 16; ++cc) {
          float v[16];
          ptx::tmem_ld16(t_lane + (uint32_t)(cc * 16), v);
          if (epi_lane && c < 2) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const uint32_t local = ptx::smem_u32(lg_s + (cc * 16 + j) * kVPad + vv);
#pragma unroll
              for (uint32_t d = 0; d < kRcCluster; ++d) ptx::st_cluster_f32(ptx::map_addr_to_cta(local, d), v[j] + b);
            }
          }
        }
      }
      csync();
      // every CTA picks the block's tokens (identical results); CTA 0 records them
      for (int j = cw; j < RB; j += 8) {
        const int m = r0 + j;
        int tok = 0;
        if (m < M) {
          float lgt[4];
          const float4 r4 = *reinterpret_cast<const float4*>(lg_s + j * kVPad + lane * 4);
          lgt[0] = r4.x; lgt[1] = r4.y; lgt[2] = r4.z; lgt[3] = r4.w;
#pragma unroll
          for (int i = 0; i < 4; ++i) if (lane * 4 + i >= P.vocab) lgt[i] = -INFINITY;
          if (c == 0 && cp.logits_dump != nullptr) {
            float* d = cp.logits_dump + ((size_t)m * (cp.max_len - 1) + n) * P.vocab;
#pragma unroll
            for (int i = 0; i < 4; ++i) if (lane * 4 + i < P.vocab) d[lane * 4 + i] = lgt[i];
          }
          const bool sample = (m % cp.group) < cp.n_sample;
          const float u = sample ? philox_uniform(cp.seed, cp.seq_id_base + (unsigned long long)m, n) : 0.f;
          float lpv;
          tok = warp_pick(lgt, lane, sample, u, P.vocab, &lpv);
          if (c == 0 && lane == 0) {
            cp.ids[(size_t)m * cp.max_len + n + 1] = tok;
            const bool ended = tok == cp.end_token;
            atomicAdd(&P.counters[n * 4 + 0], 1);
            if (sample) {
              const bool was = P.fin[m] != 0;
              if (cp.logp != nullptr) cp.logp[(size_t)m * (cp.max_len - 1) + n] = was ? 0.f : lpv;  // scst:239
              if (ended) P.fin[m] = 1;
              if (was || ended) atomicAdd(&P.counters[n * 4 + 2], 1);
            } else if (ended) {
              atomicAdd(&P.counters[n * 4 + 1], 1);
            }
          }
        }
        if (lane == 0) tok_s[j] = tok;
      }
      ptx::bar_sync(1, kRcCompute);
      if (n + 1 < P.step_end) {
        embed(n + 1);
        lsync();
        signal_mma();
      }
    }
  }
  // teardown: no peer may still address this CTA's shared memory or barriers after this point
  __syncwarp();
  ptx::tc_fence_before();
  ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, kTmemCols);
}


// ---------------------------------------------------------------- weight stream packing
struct RcPackLayer { const float *w_qkv, *w_so, *w_cq, *w_co, *w_1, *w_2; };
struct RcPackSrc {
  RcPackLayer layer[CAPDEC_MAX_LAYERS];
  const float* fc_w;  // [V, 512]
  int n_layer, vocab;
};

// One thread per 16-byte chunk (8 consecutive k of one weight row) of one 8 KB unit.
// dst: [8 ctas][L*112 + 8 units][8192 B]; unit layout = 128B-swizzled [64 rows][64 k] bf16 tile image.
__global__ void pack_wstream_kernel(const __grid_constant__ RcPackSrc src, uint8_t* __restrict__ dst) {
  const int units_per_cta = src.n_layer * kRcUnitsPerLayer + kRcUnitsHead;
  const long long total = (long long)kRcCluster * units_per_cta * 512;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int chunk = (int)(gid & 511);
  const long long uu = gid >> 9;
  const int u_all = (int)(uu % units_per_cta), c = (int)(uu / units_per_cta);
  const int r = chunk >> 3, ch = chunk & 7;  // tile row, 8-wide k chunk
  const float* w = nullptr;
  int ld = kD, row = 0, k0 = 0, rows_valid = 1 << 30;
  if (u_all < src.n_layer * kRcUnitsPerLayer) {
    const RcPackLayer& lp = src.layer[u_all / kRcUnitsPerLayer];
    int u = u_all % kRcUnitsPerLayer;
    if (u < 24) { w = lp.w_qkv; row = (u >> 3) * kD + c * 64 + r; k0 = (u & 7) * 64; }
    else if (u < 32) { u -= 24; w = lp.w_so; row = c * 64 + r; k0 = u * 64; }
    else if (u < 40) { u -= 32; w = lp.w_cq; row = c * 64 + r; k0 = u * 64; }
    else if (u < 48) { u -= 40; w = lp.w_co; row = c * 64 + r; k0 = u * 64; }
    else if (u < 80) { u -= 48; w = lp.w_1; row = c * 256 + (u >> 3) * 64 + r; k0 = (u & 7) * 64; }
    else { u -= 80; w = lp.w_2; ld = 4 * kD; row = (u >> 2) * 64 + r; k0 = c * 256 + (u & 3) * 64; }
  } else {
    const int u = u_all - src.n_layer * kRcUnitsPerLayer;
    w = src.fc_w; row = (c & 1) * 64 + r; k0 = u * 64; rows_valid = src.vocab;
  }
  float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (row < rows_valid) load8(w + (size_t)row * ld + k0 + ch * 8, f);
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]); o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(dst + (size_t)uu * kRcUnitBytes + r * 128 + ((ch ^ (r & 7)) << 4)) = o;
}

// After each chunk of steps: apply the reference's stop rules to the per-step counters
// (vit:322 all greedy rows emit <end> on one step; scst:248 all sampled rows finished) and blank
// what the cluster kernel generated speculatively past the stop.  One block.
__global__ void rollout_check_kernel(DecState* st, const CallParams* cp, const int* __restrict__ counters, int M,
                                     int step_begin, int step_end) {
  __shared__ int s_stop;  // number of executed steps if the stop fell inside this chunk, else -1
  const CallParams p = *cp;
  if (threadIdx.x == 0) {
    s_stop = -1;
    if (!st->done) {
      for (int n = step_begin; n < step_end; ++n) {
        if (p.n_greedy_rows > 0 && counters[n * 4 + 1] == p.n_greedy_rows) st->greedy_done = 1;
        if (p.n_sample_rows > 0 && counters[n * 4 + 2] == p.n_sample_rows) st->sample_done = 1;
        st->steps_executed = n + 1;
        st->step = n + 1;
        if (st->greedy_done && st->sample_done && !p.no_early_stop) { st->done = 1; s_stop = n + 1; break; }
      }
    }
  }
  __syncthreads();
  const int stop = s_stop;
  if (stop < 0 || stop >= step_end) return;
  const int extra = step_end - stop;  // speculative steps stop .. step_end-1
  for (int i = threadIdx.x; i < M * extra; i += blockDim.x) {
    const int m = i / extra, n = stop + i % extra;
    p.ids[(size_t)m * p.max_len + n + 1] = p.pad_token;
    if (p.logp != nullptr) p.logp[(size_t)m * (p.max_len - 1) + n] = 0.f;
  }
}

}  // namespace capdec
