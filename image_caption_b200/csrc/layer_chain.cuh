// layer_chain.cuh - the projections of a decoder layer as cluster-resident CHAINS (bf16 engine, v2 step).
//
// A decode step is a strict dependency chain per sequence row: out-proj -> +residual -> LayerNorm -> next projection
// -> ... Run as one kernel per projection, every link pays a grid-wide hand-over (last CTA exit -> dependent's
// griddepcontrol.wait ~1.3 us, activation tile back from L2 ~0.8 us) that is longer than its arithmetic.  Sequences
// never interact, so here a thread-block cluster of CS CTAs owns a tile of BNB sequence rows for a whole chain and the
// hand-over between links is a DSMEM all-gather + one cluster barrier:
//
//   kChainQKV      x -> self-attention in_proj -> q, K/V cache append                       (layer 0 of a step)
//   kChainAttnOut  a -> out_proj + residual + LayerNorm1 -> cross-attention q projection     (K5 + K6)
//   kChainFFN      a -> out_proj + residual + LayerNorm2 -> linear1 + ReLU -> linear2 (K split over the cluster,
//                  partials reduced through L2) + residual + LayerNorm3 [-> next layer's in_proj]   (K9 + K10 [+ K3])
//
// CTA c of the cluster owns output features [FS*c, FS*c + FS) of every 512-wide projection (FS = 512 / CS), hidden
// features [2048/CS * c, ...) of the FFN (which never leave its shared memory) and rows [1536/CS * c, ...) of the
// in_proj.  Its share of the weights is pre-packed (capdec_set_weights) as ONE linear stream of 16 KB units, each the
// 128B-swizzled shared-memory image of a UMMA A tile, in consumption order: the producer warp is a plain ring of
// cp.async.bulk copies that starts before the programmatic-launch wait (weights do not depend on the step).
// Swap-AB form as in gemm_swap.cuh: weight rows are the UMMA M dimension (128, or 64 when FS = 64), the BNB sequences
// are UMMA N; accumulators in TMEM, one column block per tile.
//
// Warp roles (320 threads): warp 0 weight / operand producer, warp 1 MMA issuer (elect.sync), warps 2-9 epilogue
// (TMEM -> registers -> bias / residual / LayerNorm / ReLU -> shared memory or global).  Only the epilogue warps take
// part in the in-chain cluster barrier (an mbarrier with one remote arrive per CTA).
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace capdec {

constexpr int kChThreads = 320;
constexpr int kChEpi = 256;      // warps 2..9
constexpr int kChUnit = 16384;   // packing granule: one k-block of a 128-row weight tile, or two k-blocks of a 64-row tile
// Ring slot = two granules, fetched by ONE bulk copy: the TMA unit keeps ~2 bulk copies in flight per SM at ~650 cycles
// each whatever their size, so 16 KB copies deliver ~50 B/clk and 32 KB copies ~115 B/clk; and the MMA
// issuer pays its barrier poll / election / commit once per 8 (16) instructions instead of once per 4 (8).
constexpr int kChSlot = 2 * kChUnit;
constexpr int kChFF = 2048;      // dim_feedforward the chain kernels are built for (every shipped reference config)

enum ChainKind { kChainQKV = 0, kChainAttnOut = 1, kChainFFN = 2 };

template <int CS> struct ChainShape {
  static constexpr int FS = kD / CS;                 // feature slice of the 512-wide projections (64 or 128)
  static constexpr bool kM64 = FS == 64;
  static constexpr int KBS = FS / 64;                // k-blocks of the gathered operand this CTA publishes
  static constexpr int UP = kM64 ? 4 : 8;            // ring units of one 512-wide projection
  static constexpr int HS = kChFF / CS;              // hidden features per CTA
  static constexpr int NT1 = HS / 128;               // linear1 tiles
  static constexpr int KB2 = HS / 64;                // k-blocks of this CTA's linear2 slice
  static constexpr int RQ = 3 * kD / CS;             // in_proj rows per CTA
  static constexpr int NTQ = RQ / 128;               // full in_proj tiles
  static constexpr bool kQ64 = (RQ % 128) != 0;      // plus one 64-row tile
  static constexpr int UQ = NTQ * 8 + (kQ64 ? 4 : 0);
  static constexpr int UC = 2 * UP;                  // out_proj + cross q
  static constexpr int UE = UP + NT1 * 8 + 4 * KB2;  // out_proj + linear1 + linear2
  static constexpr int UL = UQ + UC + UE;            // units per layer and CTA, stream order: in_proj | attn-out chain | FFN chain
};

// What one launch needs besides the operand tensor map.  Device pointers; unused ones may be null.
struct ChainParams {
  const uint8_t* wstream;     // first unit of this launch's stream for cluster rank 0
  long long rank_stride;      // bytes between the streams of consecutive cluster ranks
  float* x;                   // [M,512] fp32 residual stream: read as the residual, written with the chain's last LayerNorm
  bf16* q;                    // [M,512] bf16: cross-attention queries (AttnOut) / self-attention queries (in_proj)
  bf16* kcache;               // self K/V cache of the layer whose in_proj this launch computes: [M][H][t_pad][64]
  bf16* vcache;
  float* part;                // [row tiles][CS][BNB][512] fp32: linear2 partial products of the cluster's k-slices
  const float* b_proj;        // out_proj bias
  const float* ln_g;          // LayerNorm after the out_proj (norm1 / norm2)
  const float* ln_b;
  const float* b_cq;          // cross-attention q bias (AttnOut)
  const float* b_1;           // linear1 bias
  const float* b_2;           // linear2 bias
  const float* ln3_g;         // norm3
  const float* ln3_b;
  const float* b_qkv;         // in_proj bias (QKV, FFN with has_qkv)
  const DecState* st;
  int M, t_pad, has_qkv;
};

template <int BNB, int CS> struct ChSmem {
  using SH = ChainShape<CS>;
  static constexpr int kBTile = BNB * 128;                   // one k-block of a B operand: BNB rows x 64 k, bf16, 128B swizzle
  static constexpr int kBop = 0;                             // gathered / loaded operand, 8 k-blocks
  static constexpr int kH = kBop + 8 * kBTile;               // hidden slice (KB2 k-blocks) | LayerNorm input | output staging
  static constexpr int kHBytes = SH::KB2 * kBTile;           // >= BNB*FS*4 and >= BNB*RQ*2 (checked below)
  static constexpr int kXs = kH + kHBytes;                   // [BNB][FS] f32: this CTA's slice of the residual stream
  static constexpr int kStage = kXs + BNB * SH::FS * 4;      // [BNB][FS] bf16: slice to publish
  static constexpr int kStat = kStage + BNB * SH::FS * 2;    // [CS][BNB][2] f32 row statistics
  static constexpr int kBars = kStat + CS * BNB * 8;
  static constexpr int kRing = (kBars + 512 + 1023) / 1024 * 1024;  // 256 B of barriers + 96 B of trace stamps
  static_assert(kHBytes >= BNB * SH::FS * 4 && kHBytes >= BNB * SH::RQ * 2, "staging aliases the hidden slice");
  static constexpr int bytes(int stages) { return kRing + stages * kChSlot + 1024; }
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
      if (t1 - t0 > 2000000000ull) {  // 2 s: a protocol bug must surface as a launch failure, never as a hung GPU
        printf("capdec: cluster barrier timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
        __trap();
      }
    }
  }
}
// shared::cta -> shared::cluster bulk copy (the copy engine moves the bytes; completion on the DESTINATION CTA's mbarrier)
__device__ __forceinline__ void bulk_copy_to_cta(uint32_t dst_cluster_addr, const void* smem_src, uint32_t bytes, uint32_t bar_cluster_addr) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster_addr), "r"(smem_u32(smem_src)), "r"(bytes), "r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void st_cluster_u4(uint32_t addr, uint4 v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_cluster_f2(uint32_t addr, float a, float b) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ uint32_t map_addr_to_cta(uint32_t cta_addr, uint32_t rank) {
  uint32_t out;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(cta_addr), "r"(rank));
  return out;
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

// tuning aid (env CAPDEC_TRACE_CHAIN): CTA 0 of every launch records 64 stamps: [0] globaltimer at entry, [1] kind,
// [2..39] clock64 of epilogue thread 0 at numbered points, [40..51] clock64 of the MMA issuer, [62] globaltimer at exit
__device__ long long* g_chain_trace = nullptr;
__device__ unsigned int g_chain_trace_n = 0;

// byte offset of element (row j, k index kk in [0,64)) inside one 128B-swizzled K-major tile
__device__ __forceinline__ int ch_boff(int j, int kk) { return j * 128 + ((((kk) >> 3) ^ (j & 7)) << 4) + ((kk) & 7) * 2; }

// grid = CS * ceil(M / BNB) CTAs, cluster (CS,1,1) by launch attribute.  tm_b: [64, M, 8] view of the chain's input
// operand ([M,512] bf16: xb for kChainQKV, the attention output for the others), box = 64 x BNB x 8.
template <int BNB, int CS, int KIND>
__global__ void __launch_bounds__(kChThreads, 1)
chain_kernel(const __grid_constant__ CUtensorMap tm_b, const __grid_constant__ ChainParams P, int stages) {
  using SM = ChSmem<BNB, CS>;
  using SH = ChainShape<CS>;
  constexpr int FS = SH::FS;
  constexpr int EPT = BNB * FS / kChEpi;   // LayerNorm: elements per epilogue thread
  constexpr int TPR = FS / EPT;            // threads per row
  constexpr int HC = BNB / 2;              // TMEM columns (= sequence rows) per epilogue warp: two warps share a lane quarter
  static_assert(BNB == 32 || BNB == 64, "BNB");
  static_assert(EPT % 4 == 0 && TPR <= 32 && (32 % TPR) == 0, "LayerNorm thread map");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);

  float* x_s = reinterpret_cast<float*>(smem + SM::kXs);
  float* v_s = reinterpret_cast<float*>(smem + SM::kH);      // aliases the hidden slice (dead whenever a LayerNorm runs)
  bf16* out_s = reinterpret_cast<bf16*>(smem + SM::kH);      // aliases it too (projection outputs leave after the last MMA)
  bf16* stage_s = reinterpret_cast<bf16*>(smem + SM::kStage);
  float* stat_s = reinterpret_cast<float*>(smem + SM::kStat);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::kBars);
  uint64_t* full = bars;             // [stages] (<= 10)
  uint64_t* empty = bars + 10;       // [stages]
  uint64_t* b_full = bars + 20;      // the chain's input operand landed (TMA)
  uint64_t* op_ready = bars + 21;    // epilogue -> MMA: next B operand in place, TMEM drained
  uint64_t* acc_full = bars + 22;    // [4] MMA -> epilogue, one per accumulator tile
  uint64_t* cb = bars + 26;          // cluster barrier: one arrive per CTA
  uint64_t* g_full = bars + 28;      // the all-gathered operand (CS slices, bulk copies from the peers) landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 27);
  uint8_t* ring = smem + SM::kRing;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t c = ptx::cluster_ctarank();
  const int blk = blockIdx.x / CS;
  const int r0 = blk * BNB;
  constexpr int kTmemCols = 4 * BNB;
  const bool tracing = g_chain_trace != nullptr && blockIdx.x == 0;
  long long* mtr = reinterpret_cast<long long*>(smem + SM::kBars + 256);  // MMA issuer stamps (behind the barriers)
  long long tr[64];
  if (tracing && tid == 64) {
#pragma unroll
    for (int i = 0; i < 64; ++i) tr[i] = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr[0]));
    tr[1] = KIND; tr[2] = clock64();
  }
  if (tracing && tid == 32) { for (int i = 0; i < 12; ++i) mtr[i] = 0; }
  int tn = 3, mn = 0;
  auto stamp = [&]() { if (tracing && tid == 64 && tn < 40) tr[tn++] = clock64(); };
  auto mstamp = [&]() { if (tracing && mn < 12) mtr[mn++] = clock64(); };
  const int total_units = KIND == kChainQKV ? SH::UQ : (KIND == kChainAttnOut ? SH::UC : SH::UE + (P.has_qkv ? SH::UQ : 0));

  if (tid == 0) {
    ptx::prefetch_tensormap(&tm_b);
    for (int s = 0; s < stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::mbar_init(b_full, 1);
    ptx::mbar_init(op_ready, 1);
    for (int t = 0; t < 4; ++t) ptx::mbar_init(&acc_full[t], 1);
    ptx::mbar_init(cb, CS);
    ptx::mbar_init(g_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // every CTA's barriers are initialised before any peer may arrive on them / store into its shared memory
  ptx::cluster_sync();

  if (warp == 0) {
    // ===================== producer: the weight stream, and the chain's input operand =====================
    const uint8_t* src = P.wstream + (size_t)c * P.rank_stride;
    const uint64_t keep = ptx::policy_evict_last();  // re-read every step by every row tile: keep in L2
    const int total_slots = total_units / 2;
    const int n_pre = total_slots < stages ? total_slots : stages;
    if (ptx::elect_one()) {  // weights do not depend on the previous kernel of the step: fill the ring before the wait
      for (int u = 0; u < n_pre; ++u) {
        ptx::mbar_expect_tx(&full[u], kChSlot);
        ptx::bulk_load_1d_hint(ring + (size_t)u * kChSlot, src + (size_t)u * kChSlot, kChSlot, &full[u], keep);
      }
    }
    __syncwarp();
    ptx::pdl_wait();
    const bool done = P.st->done != 0;
    if (!done && ptx::elect_one()) {
      ptx::mbar_expect_tx(b_full, 8 * SM::kBTile);
      ptx::tma_load_3d(smem + SM::kBop, &tm_b, b_full, 0, r0, 0);  // rows >= M arrive as zeros
    }
    __syncwarp();
    ptx::pdl_trigger();
    if (done) {
      for (int u = 0; u < n_pre; ++u) ptx::mbar_wait(&full[u], 0);  // no copy may outlive the CTA
    } else {
      int s = n_pre % stages;
      uint32_t ph = n_pre >= stages ? 1u : 0u;
      for (int u = n_pre; u < total_slots; ++u, ++s) {
        if (s == stages) { s = 0; ph ^= 1u; }
        ptx::mbar_wait(&empty[s], ph ^ 1u);
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&full[s], kChSlot);
          ptx::bulk_load_1d_hint(ring + (size_t)s * kChSlot, src + (size_t)u * kChSlot, kChSlot, &full[s], keep);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    ptx::pdl_wait();
    const bool done = P.st->done != 0;
    ptx::pdl_trigger();
    if (!done) {
      constexpr uint32_t idesc128 = ptx::umma_idesc_bf16(BNB);
      constexpr uint32_t idesc64 = ptx::umma_idesc_bf16_m64(BNB);
      const uint32_t ring_lo = ptx::umma_desc_lo(ptx::smem_u32(ring));
      const uint32_t bop_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + SM::kBop));
      const uint32_t h_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + SM::kH));
      int s = 0;         // ring slot of the next unit
      uint32_t ph = 0u;  // its phase
      uint32_t opar = 0u;
      // one accumulator tile: `kblocks` k-blocks of the B operand at descriptor low word b_lo.  A ring slot holds two
      // k-blocks of a 128-row tile (16 KB each) or four of a 64-row tile (8 KB each): one barrier poll, one election and
      // one commit per slot
      auto tile = [&](int t, bool m64, int kblocks, uint32_t b_lo) {
        const uint32_t d = tmem_base + (uint32_t)(t * BNB);
        const int slots = m64 ? kblocks / 4 : kblocks / 2;
        for (int u = 0; u < slots; ++u, ++s) {
          if (s == stages) { s = 0; ph ^= 1u; }
          ptx::mbar_wait(&full[s], ph);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            const uint32_t a_lo = ring_lo + (uint32_t)s * (kChSlot >> 4);
            if (m64) {
#pragma unroll
              for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  ptx::umma_f16_lohi(d, a_lo + (uint32_t)j * (8192 >> 4) + k * 2, b_lo + (uint32_t)(4 * u + j) * (SM::kBTile >> 4) + k * 2,
                                     idesc64, (u | j | k) != 0 ? 1u : 0u);
            } else {
#pragma unroll
              for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  ptx::umma_f16_lohi(d, a_lo + (uint32_t)j * (kChUnit >> 4) + k * 2, b_lo + (uint32_t)(2 * u + j) * (SM::kBTile >> 4) + k * 2,
                                     idesc128, (u | j | k) != 0 ? 1u : 0u);
            }
            ptx::umma_commit(&empty[s]);                        // ring slot reusable once these MMAs retire
            if (u == slots - 1) ptx::umma_commit(&acc_full[t]);  // (same thread as the MMAs it tracks)
          }
          __syncwarp();
        }
      };
      uint32_t gpar = 0u;
      auto wait_operand = [&]() { ptx::mbar_wait(op_ready, opar); opar ^= 1u; ptx::tc_fence_after(); };
      auto wait_gathered = [&]() {  // TMEM drained (op_ready) and every peer's slice of the operand landed (bulk copies)
        ptx::mbar_wait(op_ready, opar); opar ^= 1u;
        ptx::mbar_wait(g_full, gpar); gpar ^= 1u;
        ptx::tc_fence_after();
      };
      auto in_proj = [&]() {
#pragma unroll
        for (int t = 0; t < SH::NTQ; ++t) tile(t, false, 8, bop_lo);
        if (SH::kQ64) tile(SH::NTQ, true, 8, bop_lo);
      };
      mstamp();  // 0: dependency wait returned
      ptx::mbar_wait(b_full, 0);
      ptx::tc_fence_after();
      mstamp();  // 1: input operand landed
      if (KIND == kChainQKV) {
        in_proj();
        mstamp();
      } else {
        tile(0, SH::kM64, 8, bop_lo);  // attention out_proj
        mstamp();  // 2: out_proj issued
        wait_gathered();
        mstamp();  // 3: gathered operand ready
        if (KIND == kChainAttnOut) {
          tile(0, SH::kM64, 8, bop_lo);  // cross-attention q
          mstamp();  // 4
        } else {
#pragma unroll
          for (int t = 0; t < SH::NT1; ++t) tile(t, false, 8, bop_lo);   // linear1: this CTA's hidden features
          mstamp();  // 4: linear1 issued
          wait_operand();
          mstamp();  // 5: hidden slice ready
#pragma unroll
          for (int t = 0; t < 4; ++t) tile(t, false, SH::KB2, h_lo);     // linear2 over this CTA's k-slice, all 512 outputs
          mstamp();  // 6: linear2 issued
          if (P.has_qkv) {
            wait_gathered();
            mstamp();  // 7
            in_proj();
            mstamp();  // 8
          }
        }
      }
    }
  } else {
    // ===================== epilogue warps =====================
    const int ct = tid - 64, cw = ct >> 5;
    const int qrt = warp & 3;             // TMEM lane quarter this warp may read
    const int col0 = (cw >> 2) * HC;      // its half of the tile's columns (= sequence rows)
    const uint32_t t_lane = tmem_base + ((uint32_t)(qrt * 32) << 16);
    const int lj = ct / TPR, lm0 = (ct % TPR) * EPT;  // LayerNorm identity: row lj, features lm0 .. lm0+EPT-1 of the slice
    const int M = P.M;
    stamp();  // 3: prologue done (cluster sync)
    ptx::pdl_wait();
    const bool done = P.st->done != 0;
    const int n = P.st->step;
    ptx::pdl_trigger();
    stamp();  // 4: dependency wait returned
    if (!done) {
      uint32_t cpar = 0u;
      uint32_t apar[4] = {0u, 0u, 0u, 0u};
      auto lsync = [&]() {  // epilogue-wide barrier; generic smem writes become visible to the tensor core
        ptx::tc_fence_before();
        ptx::fence_proxy_async_all();
        ptx::bar_sync(1, kChEpi);
      };
      // Cluster-wide barrier of the epilogue warps: DSMEM / global stores before it are visible after it.  `proxy`: the
      // stores are operands of a later tcgen05.mma (generic -> async proxy).  The CS remote arrives are issued by CS
      // lanes of one warp at once and only that warp polls the barrier: with one thread issuing the arrives one after the
      // other (each a cluster-scope release) and all 256 threads polling with cluster-scope acquires the barrier cost
      // ~4.3k cycles (measured, CAPDEC_TRACE_CHAIN), most of a LayerNorm link.
      auto csync = [&](bool proxy) {
        ptx::tc_fence_before();
        if (proxy) ptx::fence_proxy_async_all();
        ptx::bar_sync(1, kChEpi);
        if (cw == 0) {
          if (lane < CS) ptx::mbar_arrive_remote(ptx::map_to_cta(cb, (uint32_t)lane));
          ptx::mbar_wait_cluster(cb, cpar);
        }
        cpar ^= 1u;
        ptx::bar_sync(1, kChEpi);
      };
      auto signal_mma = [&]() { if (ct == 0) ptx::mbar_arrive(op_ready); };  // right after lsync() / csync()
      auto wait_acc = [&](int t) { ptx::mbar_wait(&acc_full[t], apar[t]); apar[t] ^= 1u; ptx::tc_fence_after(); };
      // accumulator tile t, this warp's columns: f(first row, 16 values of this lane's feature)
      auto for_acc = [&](int t, auto&& f) {
#pragma unroll
        for (int cc = 0; cc < HC / 16; ++cc) {
          float v[16];
          ptx::tmem_ld16(t_lane + (uint32_t)(t * BNB + col0 + cc * 16), v);
          f(col0 + cc * 16, v);
        }
      };
      // feature (within a tile) held by this thread: 128-row tiles use all 32 lanes of the quarter, 64-row tiles lanes 0-15
      const int f128 = qrt * 32 + lane;
      const int f64 = qrt * 16 + (lane & 15);
      const bool on64 = lane < 16;

      // stage_s (KBS k-block images, 128B-swizzled like the operand buffer) -> k-blocks [c*KBS, (c+1)*KBS) of the operand
      // buffer of every CTA of the cluster: one bulk copy per peer, issued by CS lanes at once, completing on the PEER's
      // g_full barrier -- its MMA issuer waits there, no cluster barrier follows.  (Per-thread st.shared::cluster stores
      // + a cluster barrier took ~3.5-5k cycles per LayerNorm link: measured, CAPDEC_TRACE_CHAIN.)
      auto publish = [&]() {
        if (cw == 0 && lane < CS) {
          const uint32_t dst = ptx::map_to_cta(smem + SM::kBop + (int)c * SH::KBS * SM::kBTile, (uint32_t)lane);
          ptx::bulk_copy_to_cta(dst, stage_s, SH::KBS * SM::kBTile, ptx::map_to_cta(g_full, (uint32_t)lane));
        }
      };
      // v[] = (projection + bias + residual) of this thread's EPT features of row lj: LayerNorm over the row's 512
      // features (statistics all-gathered), result to x_s, to the global residual stream (kStoreX) and, as bf16, to the
      // operand buffer of every CTA (kPublish).
      auto layer_norm = [&](float (&v)[EPT], const float* gamma, const float* beta, bool store_x, bool pub) {
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int e = 0; e < EPT; ++e) { s1 += v[e]; s2 = fmaf(v[e], v[e], s2); }
#pragma unroll
        for (int o = 1; o < TPR; o <<= 1) {
          s1 += __shfl_xor_sync(0xffffffffu, s1, o);
          s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        }
        if (pub && ct == 0) ptx::mbar_expect_tx(g_full, 8 * SM::kBTile);  // the CS slices the peers (and this CTA) will send
        if ((ct % TPR) == 0) {
          const uint32_t local = ptx::smem_u32(stat_s + ((int)c * BNB + lj) * 2);
#pragma unroll
          for (uint32_t d = 0; d < (uint32_t)CS; ++d) ptx::st_cluster_f2(ptx::map_addr_to_cta(local, d), s1, s2);
        }
        stamp();  // LN a: statistics sent
        csync(false);
        stamp();  // LN b: statistics gathered
        float t1 = 0.f, t2 = 0.f;
#pragma unroll
        for (int src = 0; src < CS; ++src) { t1 += stat_s[(src * BNB + lj) * 2]; t2 += stat_s[(src * BNB + lj) * 2 + 1]; }
        const float mean = t1 * (1.f / kD);
        const float rstd = rsqrtf(fmaxf(t2 * (1.f / kD) - mean * mean, 0.f) + 1e-5f);
        const int fg = (int)c * FS + lm0;  // first global feature of this thread
#pragma unroll
        for (int e = 0; e < EPT; e += 4) {
          const float4 g4 = __ldg(reinterpret_cast<const float4*>(gamma + fg + e));
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(beta + fg + e));
          float4 y;
          y.x = (v[e] - mean) * rstd * g4.x + b4.x;
          y.y = (v[e + 1] - mean) * rstd * g4.y + b4.y;
          y.z = (v[e + 2] - mean) * rstd * g4.z + b4.z;
          y.w = (v[e + 3] - mean) * rstd * g4.w + b4.w;
          *reinterpret_cast<float4*>(x_s + lj * FS + lm0 + e) = y;
          if (pub) {
            const int fl = lm0 + e;  // feature within the slice: k-block fl / 64 of the published tile(s)
            store4(reinterpret_cast<bf16*>(reinterpret_cast<uint8_t*>(stage_s) + (fl >> 6) * SM::kBTile + ch_boff(lj, fl & 63)), y.x, y.y, y.z, y.w);
          }
          if (store_x && r0 + lj < M) *reinterpret_cast<float4*>(P.x + (size_t)(r0 + lj) * kD + fg + e) = y;
        }
        stamp();  // LN c: normalised, stored
        if (pub) {
          ptx::fence_proxy_async();  // this thread's stage_s stores, before the copy engine reads them
          ptx::tc_fence_before();    // (its TMEM loads of the finished accumulators: the MMA issuer is released below)
          ptx::bar_sync(1, kChEpi);
          stamp();  // LN d
          publish();
          stamp();  // LN e: slice on its way to the peers
        }
      };
      // in_proj accumulators -> q / K cache / V cache.  Staged [BNB][RQ] bf16, then 16-byte stores: 64 consecutive
      // features share (q | k | v, head), i.e. one 128-byte run of the destination row.
      auto in_proj_epilogue = [&]() {
        constexpr int RQ = SH::RQ;
#pragma unroll
        for (int t = 0; t < SH::NTQ; ++t) {
          const float b = __ldg(P.b_qkv + (int)c * RQ + t * 128 + f128);
          wait_acc(t);
          for_acc(t, [&](int rb, const float (&v)[16]) {
#pragma unroll
            for (int j = 0; j < 16; ++j) out_s[(rb + j) * RQ + t * 128 + f128] = __float2bfloat16_rn(v[j] + b);
          });
        }
        if (SH::kQ64) {
          const float b = __ldg(P.b_qkv + (int)c * RQ + SH::NTQ * 128 + f64);
          wait_acc(SH::NTQ);
          for_acc(SH::NTQ, [&](int rb, const float (&v)[16]) {
            if (on64) {
#pragma unroll
              for (int j = 0; j < 16; ++j) out_s[(rb + j) * RQ + SH::NTQ * 128 + f64] = __float2bfloat16_rn(v[j] + b);
            }
          });
        }
        ptx::bar_sync(1, kChEpi);
        constexpr int CPR = RQ / 8;
        for (int idx = ct; idx < BNB * CPR; idx += kChEpi) {
          const int j = idx / CPR, ch = idx % CPR;
          const int row = r0 + j;
          if (row >= M) continue;
          const int f = (int)c * RQ + ch * 8;  // global in_proj feature of the chunk's first element
          const int which = f >> 9, col = f & (kD - 1);
          const uint4 val = *reinterpret_cast<const uint4*>(out_s + j * RQ + ch * 8);
          bf16* dst;
          if (which == 0) dst = P.q + (size_t)row * kD + col;
          else dst = (which == 1 ? P.kcache : P.vcache) + (((size_t)row * kH + (col >> 6)) * P.t_pad + n) * kHD + (col & 63);
          *reinterpret_cast<uint4*>(dst) = val;
        }
      };

      if (KIND == kChainQKV) {
        in_proj_epilogue();
        stamp();  // 5
      } else {
        // this CTA's slice of the residual rows (rows >= M read as zero)
        for (int idx = ct; idx < BNB * FS / 4; idx += kChEpi) {
          const int j = idx / (FS / 4), c4 = idx % (FS / 4);
          float4 v4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (r0 + j < M) v4 = *reinterpret_cast<const float4*>(P.x + (size_t)(r0 + j) * kD + (int)c * FS + c4 * 4);
          *reinterpret_cast<float4*>(x_s + j * FS + c4 * 4) = v4;
        }
        const int fp = SH::kM64 ? f64 : f128;
        const bool onp = SH::kM64 ? on64 : true;
        const float bp = __ldg(P.b_proj + (int)c * FS + fp);
        ptx::bar_sync(1, kChEpi);
        // ---------- attention out_proj + bias + residual -> v_s, LayerNorm ----------
        wait_acc(0);
        stamp();  // 5: out_proj accumulator ready
        for_acc(0, [&](int rb, const float (&v)[16]) {
          if (onp) {
#pragma unroll
            for (int j = 0; j < 16; ++j) v_s[(rb + j) * FS + fp] = v[j] + bp + x_s[(rb + j) * FS + fp];
          }
        });
        lsync();
        stamp();  // 6: out_proj epilogue done
        {
          float v[EPT];
#pragma unroll
          for (int e = 0; e < EPT; e += 4) {
            const float4 t4 = *reinterpret_cast<const float4*>(v_s + lj * FS + lm0 + e);
            v[e] = t4.x; v[e + 1] = t4.y; v[e + 2] = t4.z; v[e + 3] = t4.w;
          }
          layer_norm(v, P.ln_g, P.ln_b, KIND == kChainAttnOut, true);
        }
        signal_mma();
        stamp();  // 7: LayerNorm + all-gather done
        if (KIND == kChainAttnOut) {
          // ---------- cross-attention query (the attention kernel applies the 1/8 scale) ----------
          const float bq = __ldg(P.b_cq + (int)c * FS + fp);
          wait_acc(0);
          stamp();  // 8: cross-q accumulator ready
          for_acc(0, [&](int rb, const float (&v)[16]) {
            if (onp) {
#pragma unroll
              for (int j = 0; j < 16; ++j) out_s[(rb + j) * FS + fp] = __float2bfloat16_rn(v[j] + bq);
            }
          });
          ptx::bar_sync(1, kChEpi);
          constexpr int CPR = FS / 8;
          for (int idx = ct; idx < BNB * CPR; idx += kChEpi) {
            const int j = idx / CPR, ch = idx % CPR;
            if (r0 + j < M)
              *reinterpret_cast<uint4*>(P.q + (size_t)(r0 + j) * kD + (int)c * FS + ch * 8) = *reinterpret_cast<const uint4*>(out_s + j * FS + ch * 8);
          }
          stamp();  // 9
        } else {
          // ---------- linear1 + ReLU: hidden features [HS*c, HS*c + HS) stay here as linear2's k-slice ----------
#pragma unroll
          for (int t = 0; t < SH::NT1; ++t) {
            const int fh = t * 128 + f128;
            const float b = __ldg(P.b_1 + (int)c * SH::HS + fh);
            uint8_t* hb = smem + SM::kH + (fh >> 6) * SM::kBTile;
            wait_acc(t);
            if (t == 0) stamp();  // 8: first linear1 tile ready
            for_acc(t, [&](int rb, const float (&v)[16]) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                *reinterpret_cast<bf16*>(hb + ch_boff(rb + j, fh & 63)) = __float2bfloat16_rn(fmaxf(v[j] + b, 0.f));
            });
          }
          lsync();
          signal_mma();
          stamp();  // 9: linear1 epilogue done
          // ---------- linear2 partial products of this k-slice -> L2, reduced by the owner of each feature slice ----------
          float* pw = P.part + ((size_t)blk * CS + c) * BNB * kD;
#pragma unroll 1
          for (int t = 0; t < 4; ++t) {
            wait_acc(t);
            if (t == 0) stamp();  // 10: first linear2 tile ready
            for_acc(t, [&](int rb, const float (&v)[16]) {
#pragma unroll
              for (int j = 0; j < 16; ++j) pw[(size_t)(rb + j) * kD + t * 128 + f128] = v[j];
            });
          }
          stamp();  // 11: partials stored
          __threadfence();
          csync(false);
          stamp();  // 12: partials visible
          {
            float v[EPT];
            const int fg = (int)c * FS + lm0;
#pragma unroll
            for (int e = 0; e < EPT; e += 4) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(P.b_2 + fg + e));
              const float4 x4 = *reinterpret_cast<const float4*>(x_s + lj * FS + lm0 + e);
              v[e] = b4.x + x4.x; v[e + 1] = b4.y + x4.y; v[e + 2] = b4.z + x4.z; v[e + 3] = b4.w + x4.w;
            }
            const float* pr = P.part + ((size_t)blk * CS * BNB + lj) * kD + fg;
#pragma unroll
            for (int src = 0; src < CS; ++src)  // fixed order: bit-reproducible
#pragma unroll
              for (int e = 0; e < EPT; e += 4) {
                const float4 p4 = __ldcg(reinterpret_cast<const float4*>(pr + (size_t)src * BNB * kD + e));
                v[e] += p4.x; v[e + 1] += p4.y; v[e + 2] += p4.z; v[e + 3] += p4.w;
              }
            layer_norm(v, P.ln3_g, P.ln3_b, true, P.has_qkv != 0);
          }
          stamp();  // 13: LayerNorm3 done
          if (P.has_qkv) {
            signal_mma();
            in_proj_epilogue();
            stamp();  // 14
          }
        }
      }
    }
  }
  // teardown: no peer may still address this CTA's shared memory or barriers after this point
  __syncwarp();
  ptx::tc_fence_before();
  ptx::cluster_sync();
  if (tracing && tid == 64) {
    for (int i = 0; i < 12; ++i) tr[40 + i] = mtr[i];
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr[62]));
    tr[63] = clock64();
    const unsigned slot = atomicAdd(&g_chain_trace_n, 1u) & 255u;
    for (int i = 0; i < 64; ++i) g_chain_trace[slot * 64 + i] = tr[i];
  }
  if (warp == 1) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------- weight stream packing
struct ChainPackLayer { const float *w_qkv, *w_so, *w_cq, *w_co, *w_1, *w_2; };
struct ChainPackSrc {
  ChainPackLayer layer[CAPDEC_MAX_LAYERS];
  int n_layer;
};

// One thread per 16-byte chunk (8 consecutive k of one weight row) of one 16 KB unit.
// dst: [CS ranks][n_layer][UL units][16384 B]; a unit is the 128B-swizzled image of a [128 rows][64 k] bf16 tile, or of
// two consecutive k-blocks of a [64 rows][64 k] tile (8 KB each).
template <int CS>
__global__ void pack_chain_kernel(const __grid_constant__ ChainPackSrc src, uint8_t* __restrict__ dst) {
  using SH = ChainShape<CS>;
  const long long total = (long long)CS * src.n_layer * SH::UL * 1024;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int chunk = (int)(gid & 1023);
  const long long uu = gid >> 10;
  const int u = (int)(uu % SH::UL);
  const int l = (int)((uu / SH::UL) % src.n_layer);
  const int c = (int)(uu / ((long long)SH::UL * src.n_layer));
  const ChainPackLayer& lp = src.layer[l];
  const float* w;
  int ld = kD, row0, kb0;  // first weight row of the tile, first k-block of the unit
  bool m64;
  auto proj = [&](const float* wp, int v) { w = wp; row0 = c * SH::FS; m64 = SH::kM64; kb0 = SH::kM64 ? 2 * v : v; };
  if (u < SH::UQ) {
    w = lp.w_qkv;
    if (u < SH::NTQ * 8) { row0 = c * SH::RQ + (u >> 3) * 128; kb0 = u & 7; m64 = false; }
    else { row0 = c * SH::RQ + SH::NTQ * 128; kb0 = 2 * (u - SH::NTQ * 8); m64 = true; }
  } else if (u < SH::UQ + SH::UC) {
    const int v = u - SH::UQ;
    if (v < SH::UP) proj(lp.w_so, v); else proj(lp.w_cq, v - SH::UP);
  } else {
    const int v = u - SH::UQ - SH::UC;
    if (v < SH::UP) proj(lp.w_co, v);
    else if (v < SH::UP + SH::NT1 * 8) { const int t = v - SH::UP; w = lp.w_1; row0 = c * SH::HS + (t >> 3) * 128; kb0 = t & 7; m64 = false; }
    else { const int t = v - SH::UP - SH::NT1 * 8; w = lp.w_2; ld = kChFF; row0 = (t / SH::KB2) * 128; kb0 = c * SH::KB2 + t % SH::KB2; m64 = false; }
  }
  int r, ch, kb, base;
  if (m64) { const int sub = chunk >> 9; r = (chunk & 511) >> 3; ch = chunk & 7; kb = kb0 + sub; base = sub * 8192; }
  else { r = chunk >> 3; ch = chunk & 7; kb = kb0; base = 0; }
  float f[8];
  load8(w + (size_t)(row0 + r) * ld + kb * 64 + ch * 8, f);
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]); o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(dst + (size_t)uu * kChUnit + base + r * 128 + ((ch ^ (r & 7)) << 4)) = o;
}

}  // namespace capdec
