// gemm_ws.cuh - large-batch decode-step projection (K = 512) as a PERSISTENT, WEIGHT-STATIONARY tcgen05 GEMM.
//
// Same swap-AB product as gemm_swap.cuh,  C^T[n_out, batch] = W[n_out, 512] * X[batch, 512]^T,  for batches of thousands
// of sequences (cfg3: 3072 rows), where gemm_swap_kernel runs at ~20 % of the tensor pipe because
//   * its main loop is bound by SHARED-MEMORY bandwidth: per k-block the TMA unit writes 32 KB (W + X tiles) and the
//     four SS-mode MMAs read the same 32 KB back = 512 cycles of the SM's 128 B/clk against 256 cycles of tensor pipe
//     (two CTAs per SM make that ~1000 cycles per k-block, measured: profiles/r01_v2_gemm_phase_trace_cfg3.txt);
//   * every 128 x 128 output tile re-reads its 128 KB weight slice from L2 (FFN1: 384 CTAs x 256 KB = 98 MB per launch);
//   * one tile per CTA: prologue, TMEM allocation and epilogue of a tile never overlap the next tile's loads, and
//     384 tiles on 296 CTA slots leave a 30 % second wave.
// Here a CTA owns ONE 128-feature weight tile for the whole launch and keeps it in TENSOR MEMORY as the MMA's A operand
// (tcgen05.mma with A from TMEM: 128 lanes x 256 columns hold the 128 x 512 bf16 slice, written once with tcgen05.st
// before the programmatic-launch wait: weights do not depend on the step).  It then walks its share of the row tiles:
// only X tiles travel through shared memory (TMA ring), the MMA reads only its B operand from there (half the
// shared-memory traffic per flop), accumulators are double-buffered in TMEM so the epilogue of tile i overlaps the MMAs
// of tile i+1, and grid = (feature tiles) x (148 / feature tiles) row groups fills the GPU in one wave.
//
// The weight slice reaches tensor memory through the (still idle) operand ring: eight TMA boxes, then every epilogue
// thread copies its row shared memory -> registers -> tcgen05.st (a per-thread global read of a 1 KB row took ~6 us).
// Outputs leave through shared memory and bulk tensor stores: the epilogue thread owns a FEATURE (TMEM lane), so direct
// global stores are 64-byte pieces per warp instruction and cost ~4 us per launch (measured, CAPDEC_WS_DBG); staged as a
// row-major [BNB][128] bf16 tile they leave as ONE cp.async.bulk.tensor store (for the K/V cache append: a 4-D box
// {64, position, 2 heads, BNB rows}, which has the same shared-memory image).
//
// TMEM map (512 columns): [0, 2*BNB) two fp32 accumulators, [256, 512) the weight slice (two bf16 per column).
// Warps: 0 = TMA producer, 1 = MMA issuer (elect.sync), 2-9 = epilogue (two per TMEM lane quarter).
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "gemm_swap.cuh"

namespace capdec {

constexpr int kWsThreads = 320;
constexpr int kWsEpi = 256;
constexpr int kWsK = 512;            // the only K this kernel is built for (d_model)
constexpr int kWsKB = kWsK / 64;     // k-blocks per row tile
constexpr int kWsACol = 256;         // first TMEM column of the weight slice

constexpr int kWsRing = 10;          // operand ring stages (the weight prologue needs 8)
__host__ __device__ constexpr int ws_stage_bytes(int bnb) { return bnb * 64 * 2; }
__host__ __device__ constexpr int ws_out_bytes(int bnb) { return bnb * kSwapBF * 2; }  // one staged output tile (bf16)
__host__ __device__ constexpr int ws_smem_bytes(int bnb) { return kWsRing * ws_stage_bytes(bnb) + 2 * ws_out_bytes(bnb) + 1024 + 512; }

namespace ptx {
// D[tmem] (+)= A[tmem] * B[smem]: the A operand is read from tensor memory (lane = row, two 16-bit elements per column)
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "mov.b64 db, {%2, %5};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(kUmmaDescHiSw128)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc512(uint32_t* smem_slot) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(smem_slot)) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st32_u(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
        "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
        "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
}  // namespace ptx

// ---------------------------------------------------------------- epilogues: value(f, acc) per element, store() per tile
// out[row, f] = act(acc + bias[f]);  ts: [M, n_out] bf16, box = 128 features x BNB rows
template <bool kRelu>
struct WsLinear {
  alignas(64) CUtensorMap ts;
  const float* bias;
  __device__ __forceinline__ SwapEpiPre prepare(int f) const { return SwapEpiPre{bias ? __ldg(bias + f) : 0.f, 0}; }
  __device__ __forceinline__ float value(float acc, const SwapEpiPre& pre) const {
    const float r = acc + pre.bias;
    return kRelu ? fmaxf(r, 0.f) : r;
  }
  __device__ __forceinline__ int row0(int tile) const { return tile * 128; }
  __device__ __forceinline__ int stage_index(int j, int fl, int) const { return j * kSwapBF + fl; }  // row-major [128 rows][128 features]
  __device__ __forceinline__ void store(const void* stage, int f0, int tile, const SwapEpiPre&) const { ptx::tma_store_2d(&ts, stage, f0, tile * 128); }
};
// self-attention in_proj: features [0,512) -> q, [512,1024) -> K cache position st->step, [1024,1536) -> V cache.
// ts_k / ts_v: this layer's cache [rows][H][t_pad][64] as a 4-D tensor, box = {64, 1 position, 2 heads, BNB rows}.
struct WsQKV {
  alignas(64) CUtensorMap ts_q;
  alignas(64) CUtensorMap ts_k;
  alignas(64) CUtensorMap ts_v;
  const float* bias;
  const DecState* st;
  __device__ __forceinline__ SwapEpiPre prepare(int f) const { return SwapEpiPre{__ldg(bias + f), st->step}; }
  __device__ __forceinline__ float value(float acc, const SwapEpiPre& pre) const { return acc + pre.bias; }
  __device__ __forceinline__ int row0(int tile) const { return tile * 128; }
  __device__ __forceinline__ int stage_index(int j, int fl, int) const { return j * kSwapBF + fl; }
  __device__ __forceinline__ void store(const void* stage, int f0, int tile, const SwapEpiPre& pre) const {
    const int b0 = tile * 128;
    const int which = f0 >> 9, c = f0 & (kD - 1);
    if (which == 0) ptx::tma_store_2d(&ts_q, stage, c, b0);
    else ptx::tma_store_4d(which == 1 ? &ts_k : &ts_v, stage, 0, pre.step, c >> 6, b0);
  }
};

// Prefill: rows = (image, memory position), features = layer*1024 + kv*512 + head*64 + d; destination = the
// cross cache [L*2][n_images][H][S][64].  A tile is 128 consecutive rows, staged as [head][128 rows][64]; it overlaps a few
// images, and each overlap (image i, positions [pos, pos + n)) is, per head, n * 128 contiguous bytes both in the staged
// plane and in the cache: one 1-D bulk store each.  (First attempt, kept as a measurement: 5-D bulk tensor stores.  A box
// at a negative position faults on B200 - illegal instruction - so partial overlaps were decomposed into exact boxes of
// 64, 32 ... 1 positions; ~16 small tensor stores per tile cost 137 of 470 us at 1024 x 49.)
struct WsCrossKV {
  bf16* cache;
  const float* bias;
  int S, n_images, img0, rows;  // img0 = first image of this chunk; rows = n_chunk * S rows of X
  __device__ __forceinline__ SwapEpiPre prepare(int f) const { return SwapEpiPre{__ldg(bias + f), 0}; }
  __device__ __forceinline__ float value(float acc, const SwapEpiPre& pre) const { return acc + pre.bias; }
  __device__ __forceinline__ int row0(int tile) const { return tile * 128; }
  __device__ __forceinline__ int stage_index(int j, int fl, int) const { return ((fl >> 6) * 128 + j) * 64 + (fl & 63); }
  __device__ __forceinline__ void store(const void* stage, int f0, int tile, const SwapEpiPre&) const {
    const int lk = f0 >> 9, head0 = (f0 & (kD - 1)) >> 6;
    const int r0 = tile * 128, r1 = min(r0 + 128, rows);
    const uint32_t src = ptx::smem_u32(stage);
    for (int i = r0 / S; i * S < r1; ++i) {
      const int a = max(i * S, r0), b = min((i + 1) * S, r1);  // rows [a, b) of the tile belong to image i
      const int n = b - a, pos = a - i * S, jrow = a - r0;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        bf16* dst = cache + ((((size_t)lk * n_images + img0 + i) * kH + head0 + hh) * S + pos) * kHD;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     ::"l"(dst), "r"(src + (uint32_t)((hh * 128 + jrow) * 128)), "r"(n * 128) : "memory");
      }
    }
  }
};

// Full-sequence pass, self-attention in_proj over rows = (sequence, position), sequence-major: features
// [0,512) -> q [rows,512] (2-D tensor store); [512,1024) / [1024,1536) -> this layer's self K / V cache
// [cache rows][H][t_pad][64] at (seq0 + sequence, head, position): like the prefill, each overlap of the tile with a
// sequence is n * 128 contiguous bytes per head on both sides, one 1-D bulk store each.
struct WsQKVFull {
  alignas(64) CUtensorMap ts_q;
  bf16* kc;
  bf16* vc;
  const float* bias;
  int T, t_pad, seq0, rows;  // T positions per sequence; rows = sequences * T
  __device__ __forceinline__ SwapEpiPre prepare(int f) const { return SwapEpiPre{__ldg(bias + f), 0}; }
  __device__ __forceinline__ float value(float acc, const SwapEpiPre& pre) const { return acc + pre.bias; }
  __device__ __forceinline__ int row0(int tile) const { return tile * 128; }
  __device__ __forceinline__ int stage_index(int j, int fl, int f0) const {
    return f0 < kD ? j * kSwapBF + fl : ((fl >> 6) * 128 + j) * 64 + (fl & 63);
  }
  __device__ __forceinline__ void store(const void* stage, int f0, int tile, const SwapEpiPre&) const {
    const int which = f0 >> 9, c = f0 & (kD - 1);
    if (which == 0) { ptx::tma_store_2d(&ts_q, stage, c, tile * 128); return; }
    bf16* cache = which == 1 ? kc : vc;
    const int head0 = c >> 6;
    const int r0 = tile * 128, r1 = min(r0 + 128, rows);
    const uint32_t src = ptx::smem_u32(stage);
    for (int i = r0 / T; i * T < r1; ++i) {
      const int a = max(i * T, r0), b = min((i + 1) * T, r1);  // rows [a, b) of the tile belong to sequence i
      const int n = b - a, pos = a - i * T, jrow = a - r0;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        bf16* dst = cache + (((size_t)(seq0 + i) * kH + head0 + hh) * t_pad + pos) * kHD;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     ::"l"(dst), "r"(src + (uint32_t)((hh * 128 + jrow) * 128)), "r"(n * 128) : "memory");
      }
    }
  }
};

// grid = (n_out / 128, row groups); CTA (x, y) owns features [128 x, 128 x + 128) and row tiles y, y + gridDim.y, ...
// tm_w: the packed bf16 weight matrix [n_out, 512], box = 64 columns x 128 rows; tm_x: [M, 512] activations, box = 64 x BNB.
template <int BNB, typename Epi>
__global__ void __launch_bounds__(kWsThreads, 1)
gemm_ws_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x, int n_tiles,
               const __grid_constant__ Epi epi, const DecState* st, int dbg) {
  static_assert(BNB == 128, "the weight prologue borrows the operand ring: stage bytes must equal one 128-row weight box");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int kXBytes = ws_stage_bytes(BNB);
  constexpr int stages = kWsRing;
  uint8_t* out_s = smem + (size_t)stages * kXBytes;  // [2][BNB][128] bf16
  uint64_t* bars = reinterpret_cast<uint64_t*>(out_s + 2 * ws_out_bytes(BNB));
  uint64_t* full = bars;            // [stages]
  uint64_t* empty = bars + 14;      // [stages]
  uint64_t* acc_full = bars + 28;   // [2] MMA -> epilogue
  uint64_t* acc_empty = bars + 30;  // [2] epilogue -> MMA (one arrive per epilogue warp)
  uint64_t* w_full = bars + 32;     // the weight slice landed in the ring
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 33);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int f0 = blockIdx.x * kSwapBF;
  const int my_tiles = ((int)blockIdx.y < n_tiles) ? (n_tiles - 1 - (int)blockIdx.y) / (int)gridDim.y + 1 : 0;

  if (warp == 0) {
    if (ptx::elect_one()) {
      ptx::prefetch_tensormap(&tm_w);
      ptx::prefetch_tensormap(&tm_x);
      for (int s = 0; s < stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
      for (int b = 0; b < 2; ++b) { ptx::mbar_init(&acc_full[b], 1); ptx::mbar_init(&acc_empty[b], kWsEpi / 32); }
      ptx::mbar_init(w_full, 1);
      ptx::fence_barrier_init();
      if (!(dbg & 2)) {  // the weight slice, k-block kb into ring stage kb (weights do not depend on the step)
        const uint64_t keep = ptx::policy_evict_last();
        ptx::mbar_expect_tx(w_full, kWsKB * kXBytes);
        for (int kb = 0; kb < kWsKB; ++kb) ptx::tma_load_2d_hint(smem + (size_t)kb * kXBytes, &tm_w, w_full, kb * 64, f0, keep);
      }
    }
    __syncwarp();
  }
  if (warp == 1) ptx::tmem_alloc512(tmem_slot);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= 2 && !(dbg & 2)) {
    // ---- weight slice: shared memory (128B-swizzled k-block tiles) -> registers -> tensor memory.  Thread = weight row
    // (TMEM lane); warps 2-5 take k-blocks 0-3, warps 6-9 k-blocks 4-7: 128 bytes = 32 columns per tcgen05.st.
    const int quarter = warp & 3;
    const int khalf = (warp - 2) >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t a_lane = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(kWsACol + khalf * 128);
    ptx::mbar_wait(w_full, 0);
#pragma unroll 1
    for (int p = 0; p < 4; ++p) {
      const uint8_t* tile = smem + (size_t)(khalf * 4 + p) * kXBytes + r * 128;
      uint32_t v[32];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 q4 = *reinterpret_cast<const uint4*>(tile + ((i ^ (r & 7)) << 4));
        v[4 * i] = q4.x; v[4 * i + 1] = q4.y; v[4 * i + 2] = q4.z; v[4 * i + 3] = q4.w;
      }
      ptx::tmem_st32_u(a_lane + (uint32_t)(p * 32), v);
    }
    ptx::tmem_wait_st();
    ptx::tc_fence_before();
  }
  __syncthreads();  // the slice is in tensor memory and the ring is free again before the first X tile / MMA
  ptx::tc_fence_after();

  ptx::pdl_wait();     // activations (and DecState) of the previous kernel of the step
  ptx::pdl_trigger();
  const bool active = !(st != nullptr && st->done) && my_tiles > 0;

  if (warp == 0) {
    if (active) {
      int s = 0;
      uint32_t ph = 0u;
      for (int i = 0; i < my_tiles; ++i) {
        const int b0 = epi.row0((int)blockIdx.y + i * (int)gridDim.y);
        for (int kb = 0; kb < kWsKB; ++kb, ++s) {
          if (s == stages) { s = 0; ph ^= 1u; }
          ptx::mbar_wait(&empty[s], ph ^ 1u);
          if (ptx::elect_one()) {
            if ((dbg & 4) && i > 0) ptx::mbar_arrive(&full[s]);  // tuning aid: no activation traffic after the first tile (wrong results)
            else {
              ptx::mbar_expect_tx(&full[s], kXBytes);
              ptx::tma_load_2d(smem + (size_t)s * kXBytes, &tm_x, &full[s], kb * 64, b0);  // rows beyond the matrix arrive as zeros
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    if (active) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(BNB);
      const uint32_t x_lo0 = ptx::umma_desc_lo(ptx::smem_u32(smem));
      const uint32_t a0 = tmem_base + (uint32_t)kWsACol;
      int s = 0;
      uint32_t ph = 0u;
      for (int i = 0; i < my_tiles; ++i) {
        const int buf = i & 1;
        ptx::mbar_wait(&acc_empty[buf], (((uint32_t)i >> 1) & 1u) ^ 1u);  // (first use of a buffer passes at once)
        ptx::tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(buf * BNB);
        for (int kb = 0; kb < kWsKB; ++kb, ++s) {
          if (s == stages) { s = 0; ph ^= 1u; }
          ptx::mbar_wait(&full[s], ph);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            const uint32_t x_lo = x_lo0 + (uint32_t)s * (kXBytes >> 4);
#pragma unroll
            for (int k = 0; k < 4; ++k)  // 16 bf16 of K per instruction: 8 TMEM columns of A, 32 bytes of the B swizzle atom
              ptx::umma_f16_ts(d, a0 + (uint32_t)(kb * 32 + k * 8), x_lo + k * 2, idesc, (kb | k) != 0 ? 1u : 0u);
            ptx::umma_commit(&empty[s]);
            if (kb == kWsKB - 1) ptx::umma_commit(&acc_full[buf]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    const int quarter = warp & 3;
    const int chalf = (warp - 2) >> 2;   // which half of the tile's columns (= sequence rows) this warp drains
    const int fl = quarter * 32 + lane;  // feature within the tile
    const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const bool issuer = threadIdx.x == 64;
    if (active) {
      const SwapEpiPre pre = epi.prepare(f0 + fl);
      for (int i = 0; i < my_tiles; ++i) {
        const int buf = i & 1;
        const int tile = (int)blockIdx.y + i * (int)gridDim.y;
        bf16* stage = reinterpret_cast<bf16*>(out_s + (size_t)buf * ws_out_bytes(BNB));
        if (issuer && i >= 2) ptx::bulk_wait_read1();  // the store of tile i-2 has read this staging buffer
        ptx::bar_sync(1, kWsEpi);
        ptx::mbar_wait(&acc_full[buf], ((uint32_t)i >> 1) & 1u);
        ptx::tc_fence_after();
#pragma unroll 1
        for (int c0 = chalf * (BNB / 2); c0 < (chalf + 1) * (BNB / 2); c0 += 32) {
          float v[32];
          ptx::tmem_ld32(t_lane + (uint32_t)(buf * BNB + c0), v);
          if (!(dbg & 17)) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int at = epi.stage_index(c0 + j, fl, f0);  // where (row j of the tile, feature fl) sits in the staged box
              if (at >= 0) stage[at] = __float2bfloat16_rn(epi.value(v[j], pre));
            }
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&acc_empty[buf]);  // the accumulator may be overwritten
        ptx::fence_proxy_async();
        ptx::bar_sync(1, kWsEpi);
        if (issuer && !(dbg & 9)) {  // rows >= M are clipped by the store maps
          epi.store(stage, f0, tile, pre);
          ptx::bulk_commit();
        }
      }
      if (issuer) ptx::bulk_wait_all();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
}

}  // namespace capdec
