// gemm_swap.cuh - decode-step projection on tcgen05 in "swap-AB" form:
//     C^T[n_out, batch] = W[n_out, K] * X[batch, K]^T
// The weight rows fill the 128 TMEM lanes (UMMA M = 128), the live sequences are the UMMA N
// dimension (BNB = 32/64/128 per CTA), so a step with few sequences still launches
// (n_out/128) x (M/BNB) CTAs, every CTA streams only its 128 x K weight slice once, and the
// epilogue thread that owns output feature f writes 32 consecutive features per warp for each
// sequence row (coalesced 128 B fp32 / 64 B bf16 stores).
//
// 6 warps: warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2-5 epilogue.
// kLN variant: launched as clusters of 4 CTAs along n_out (= 512 features of the same BNB
// rows); the epilogue adds bias + residual, exchanges per-row (sum, sum of squares) through
// distributed shared memory and writes LayerNorm(x + proj) in place (fp32 x and its bf16 mirror):
// SURVEY K5/K9/K10 "GEMM with residual+LayerNorm epilogue".
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace capdec {

constexpr int kSwapBF = 128;       // output features per CTA (UMMA M)
constexpr int kSwapBK = 64;        // K per pipeline stage (one 128 B swizzle atom row)
// warp 0 = TMA producer, warp 1 = MMA issuer, then the epilogue warps: 4 for a 32-column tile (one per TMEM lane
// quarter), 8 for wider tiles (two per quarter, each taking every other 32-column chunk) -- at BNB = 128 the
// epilogue is the longest phase of the kernel (measured: 17-19k of ~27k cycles with 4 warps).
// The last warp is a second MMA issuer: one warp sustains ~75-130 cycles per tcgen05.mma of this size, the tensor
// pipe ~40 (tools/mma_probe2.cu), so even and odd k-blocks are issued by two warps into two accumulators that the
// epilogue adds.
// (Wide tiles without LayerNorm keep one issuer: they run two CTAs per SM, whose register budget the second
// accumulator's epilogue would exceed, and their MMA issue already overlaps the other CTA's epilogue.)
__host__ __device__ constexpr int swap_epi_threads(int bnb) { return bnb == 32 ? 128 : 256; }
#ifndef CAPDEC_SWAP_ISSUERS32
#define CAPDEC_SWAP_ISSUERS32 2
#endif
__host__ __device__ constexpr int swap_issuers(int bnb, bool ln) { return bnb == 32 ? CAPDEC_SWAP_ISSUERS32 : (ln ? 2 : 1); }
__host__ __device__ constexpr int swap_threads(int bnb, bool ln) { return 64 + swap_epi_threads(bnb) + 32 * (swap_issuers(bnb, ln) - 1); }

constexpr int kSwapWholeKB = 8;  // k-blocks of a "whole K" launch: K = 512 per CTA
__host__ __device__ constexpr int swap_stage_bytes(int bnb) { return (kSwapBF + bnb) * kSwapBK * 2; }
// pipeline stages + 1 KB alignment slack + barriers/tmem slot (256 B) + LN scratch (16 * bnb floats)
// + split-K partial tiles received from the cluster peers ((ksplit-1) x 128 x bnb floats)
// LayerNorm epilogue: the normalised tile (fp32 + bf16 copies, 128 features x bnb rows) is staged in the operand ring
__host__ __device__ constexpr int swap_ln_stage_out_bytes(int bnb) { return bnb * kSwapBF * (4 + 2); }
__host__ __device__ constexpr int swap_smem_bytes(int bnb, int stages, int ksplit = 1) {
  return stages * swap_stage_bytes(bnb) + 1024 + 256 + 16 * bnb * 4 + (ksplit - 1) * kSwapBF * bnb * 4;
}

// Sum over the 32 lanes of v[j], for every j: afterwards lane l holds the total of column l.
// 31 shuffles instead of 32 x 5 (halving exchange).  Destroys v.
__device__ __forceinline__ float lane_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const bool up = (lane & o) != 0;
#pragma unroll
    for (int j = 0; j < o; ++j) {
      const float send = up ? v[j] : v[j + o];
      const float keep = up ? v[j + o] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  return v[0];
}

// ---------------------------------------------------------------- epilogues (feature-major)
// chunk(): called by the thread owning output feature f with 32 accumulators for sequence
// rows r0 .. r0+31 (rows >= M must be skipped).
// prepare(): called by the same thread while the MMAs are still running (after the programmatic-launch wait): the
// bias and the step index are global loads; fetched inside chunk() they would add an L2 round trip to the end of
// every kernel of the chain.
struct SwapEpiPre {
  float bias;
  int step;
};

// out[row, f] = act(acc + bias[f])
template <typename OutT, bool kRelu>
struct SwapLinear {
  OutT* out;
  const float* bias;
  int ldo;
  long long zstride = 0;  // split-K launches (gridDim.z > 1): k-slice z writes its partial product to out + z * zstride
  __device__ __forceinline__ SwapEpiPre prepare(int f) const { return SwapEpiPre{bias ? __ldg(bias + f) : 0.f, 0}; }
  __device__ __forceinline__ void chunk(int f, int r0, int M, const float (&v)[32], const SwapEpiPre& pre) const {
    const float b = pre.bias;
    OutT* o = out + (size_t)blockIdx.z * zstride;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (r0 + j < M) {
        float r = v[j] + b;
        if (kRelu) r = fmaxf(r, 0.f);
        o[(size_t)(r0 + j) * ldo + f] = from_f32<OutT>(r);
      }
    }
  }
};

// Split-K projection that closes itself (FFN2 at small batch): every k-slice CTA writes its fp32 partial tile
// part[z][row][f]; the last of the (feature tiles x k-slices) CTAs of a row tile to finish -- elected by a ticket
// counter in global memory -- adds the slices, the bias and the residual and writes LayerNorm(x + proj) for the
// tile's rows, one warp per row (sum_ln_row, the arithmetic of sum_ln_kernel).  This removes the sum_ln launch
// from the step's dependent chain (measured no-work floor of a chained kernel: ~4.2 us).
struct SwapPartialLN {
  float* part;          // [gridDim.z][M][512]
  long long zstride;    // M * 512
  float* x;             // [M, 512] residual stream, updated in place
  bf16* xb;             // bf16 mirror
  const float* bias;    // [512] projection bias
  const float* gamma;
  const float* beta;
  unsigned* tickets;    // [row tiles], zero between launches
  static constexpr bool kHasFinish = true;
  __device__ __forceinline__ SwapEpiPre prepare(int) const { return SwapEpiPre{0.f, 0}; }
  __device__ __forceinline__ void chunk(int f, int r0, int M, const float (&v)[32], const SwapEpiPre&) const {
    float* o = part + (size_t)blockIdx.z * zstride;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (r0 + j < M) o[(size_t)(r0 + j) * kD + f] = v[j];
  }
  // every thread of the CTA, after the epilogue warps have stored their partial tile
  template <int NZ>
  __device__ __forceinline__ void finish(int b0, int bnb, int M, int* flag) const {
    __threadfence();  // this thread's partial-tile stores
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned total = gridDim.x * gridDim.z;  // CTAs contributing to this row tile
      const unsigned t = atomicAdd(&tickets[blockIdx.y], 1u);
      const bool last = t == total - 1;
      if (last) tickets[blockIdx.y] = 0u;  // nobody else touches it before the next launch
      __threadfence();
      *flag = last ? 1 : 0;
    }
    __syncthreads();
    if (*flag == 0) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    float4 bi[4], ga[4], be[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = i * 128 + lane * 4;
      bi[i] = __ldg(reinterpret_cast<const float4*>(bias + c));
      ga[i] = __ldg(reinterpret_cast<const float4*>(gamma + c));
      be[i] = __ldg(reinterpret_cast<const float4*>(beta + c));
    }
    const int rend = b0 + bnb < M ? b0 + bnb : M;
    for (int row = b0 + warp; row < rend; row += nwarps) sum_ln_row<NZ>(x, part, M, row, lane, bi, ga, be, xb);
  }
};

template <typename Epi, typename = void>
struct swap_epi_has_finish { static constexpr bool value = false; };
template <typename Epi>
struct swap_epi_has_finish<Epi, decltype((void)Epi::kHasFinish)> { static constexpr bool value = Epi::kHasFinish; };

// self-attention in_proj: f in [0,512) -> q, [512,1024) -> K cache row st->step, [1024,1536) -> V cache
template <typename Elem>
struct SwapQKV {
  Elem* q;
  Elem* kcache;  // this layer, [M][H][t_pad][64]
  Elem* vcache;
  const float* bias;
  const DecState* st;
  int t_pad;
  __device__ __forceinline__ SwapEpiPre prepare(int f) const { return SwapEpiPre{__ldg(bias + f), st->step}; }
  __device__ __forceinline__ void chunk(int f, int r0, int M, const float (&v)[32], const SwapEpiPre& pre) const {
    const float b = pre.bias;
    const int n = pre.step;
    const int which = f >> 9, c = f & (kD - 1);
    Elem* base;
    size_t row_stride, off;
    if (which == 0) { base = q; row_stride = kD; off = c; }
    else {
      base = which == 1 ? kcache : vcache;
      row_stride = (size_t)kH * t_pad * kHD;
      off = ((size_t)(c >> 6) * t_pad + n) * kHD + (c & 63);
    }
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (r0 + j < M) base[(size_t)(r0 + j) * row_stride + off] = from_f32<Elem>(v[j] + b);
  }
};

// x <- LayerNorm(x + acc + bias) * gamma + beta, in place; xb = bf16 mirror (next GEMM's operand)
struct SwapLNParams {
  float* x;            // [M, 512] fp32 residual stream
  bf16* xb;            // [M, 512]
  const float* bias;   // [512]
  const float* gamma;
  const float* beta;
  // store views of x (fp32) and xb (bf16): [M, 512], box = 128 features x BNB rows, no swizzle.  The normalised
  // tile leaves through shared memory and two bulk tensor stores: per-thread global stores ran at ~10 cycles
  // per warp instruction and were the longest part of the LayerNorm epilogue (measured, CAPDEC_TRACE_GEMM).
  alignas(64) CUtensorMap ts_x;
  alignas(64) CUtensorMap ts_xb;
  int tma_store;  // 1: output tile staged in shared memory and written by two bulk tensor stores; 0: per-thread stores
};


struct SwapNoLN {};

// tuning aid (env CAPDEC_TRACE_GEMM): CTA (0,0) of every launch records 12 stamps
// {globaltimer at entry, clock at entry, after prologue, after pdl wait, accumulator ready, epilogue done, globaltimer at exit, kLN}
__device__ long long* g_swap_trace = nullptr;
__device__ unsigned int g_swap_trace_n = 0;

// grid = (n_out / 128, ceil(M / BNB)); kLN: n_out == 512, grid.x = cluster dims x = 4 * KSPLIT.
// KSPLIT > 1 (kLN only): cluster rank r = feature tile (r & 3) + 4 * k-slice (r >> 2).  tcgen05.mma costs
// ~148 cycles per instruction whatever its shape (measured, tools/mma_probe.cu), so a CTA's critical path
// is K/16 instructions; the K range is therefore divided over KSPLIT CTAs whose partial accumulators are
// added through distributed shared memory by the k-slice-0 CTA of each feature tile.
template <int BNB, bool kLN, typename Epi, int KSPLIT = 1>
__global__ void __launch_bounds__(swap_threads(BNB, kLN), (kLN && BNB > 32) ? 1 : 2)
gemm_swap_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
                 const __grid_constant__ CUtensorMap tm_w3, const __grid_constant__ CUtensorMap tm_x3, int whole_k,
                 int M, int K, int stages, Epi epi, const __grid_constant__ SwapLNParams ln, const DecState* st) {
  static_assert(BNB == 32 || BNB == 64 || BNB == 128, "BNB");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int kWBytes = kSwapBF * kSwapBK * 2;
  constexpr int kXBytes = BNB * kSwapBK * 2;
  uint8_t* smem_w = smem;
  uint8_t* smem_x = smem + (size_t)stages * kWBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_x + (size_t)stages * kXBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + stages;
  uint64_t* tmem_full = bars + 2 * stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * stages + 1);
  float* red = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);  // [2][4][BNB]; later stat[2][BNB]
  float* clu = red + 8 * BNB;                                                      // [4 ranks][2][BNB]
  float* part = red + 16 * BNB;                                                    // [KSPLIT-1][128][BNB] partial tiles

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kIssuers = swap_issuers(BNB, kLN);
  constexpr int kIssuer2 = 2 + swap_epi_threads(BNB) / 32;  // first of the kIssuers - 1 extra issuer warps (the last warps)
  const int f0 = (kLN ? (int)(blockIdx.x & 3) : (int)blockIdx.x) * kSwapBF, b0 = blockIdx.y * BNB;
  const int ks = kLN ? (int)(blockIdx.x >> 2) : (int)blockIdx.z;  // k-slice of this CTA
  const int num_kb = K / kSwapBK / (kLN ? KSPLIT : (int)gridDim.z);
  const int kb0 = ks * num_kb;
  const int n_acc = num_kb < kIssuers ? num_kb : kIssuers;  // issuers (= accumulators) that have k-blocks at all
  const int n_pre = num_kb < stages ? num_kb : stages;  // k-blocks whose weight tiles are requested in the prologue
  long long tr[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  const bool trace_cta = g_swap_trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  const bool tracing = trace_cta && threadIdx.x == 64;
  if (tracing) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr[0])); tr[1] = clock64(); }

  // TMA and tcgen05 instructions take warp-uniform operands.  Issued under `if (lane == 0)` the compiler cannot prove
  // uniformity and wraps each one in an elect/broadcast/branch loop (measured: ~150 cycles per MMA, ~150-220 per bulk
  // copy); issued by the converged warp under elect.sync they cost ~40-75 cycles (tools/mma_probe2.cu).
  if (warp == 0) {
    if (ptx::elect_one()) {
      ptx::prefetch_tensormap(&tm_w);
      ptx::prefetch_tensormap(&tm_x);
      for (int s = 0; s < stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
      ptx::mbar_init(tmem_full, n_acc);  // one arrival per MMA issuer
      ptx::fence_barrier_init();
      // the weights do not depend on the previous kernel of the step: request the first ring of weight tiles now,
      // so that only the activation tiles are still to come once the programmatic-launch wait returns
      const uint64_t keep = ptx::policy_evict_last();  // weights are re-read every step: keep them in L2
      if (whole_k) {
        // whole K in three copies: full[0] / full[1] = the two halves of the weight slice, full[2] = the activations.
        // (The TMA unit keeps ~2 copies in flight, ~650 cycles each whatever their size: one box per k-block and
        // operand delivered a pair of k-blocks every ~650 cycles and was the longest part of the kernel.)
        ptx::prefetch_tensormap(&tm_w3);
        ptx::prefetch_tensormap(&tm_x3);
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          ptx::mbar_expect_tx(&full[hf], (kSwapWholeKB / 2) * kWBytes);
          ptx::tma_load_3d_hint(smem_w + (size_t)hf * (kSwapWholeKB / 2) * kWBytes, &tm_w3, &full[hf], 0, f0,
                                kb0 + hf * (kSwapWholeKB / 2), keep);
        }
      } else {
        for (int kb = 0; kb < n_pre; ++kb) {
          ptx::mbar_expect_tx(&full[kb], kWBytes + kXBytes);
          ptx::tma_load_2d_hint(smem_w + (size_t)kb * kWBytes, &tm_w, &full[kb], (kb0 + kb) * kSwapBK, f0, keep);
        }
      }
    }
    __syncwarp();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, kIssuers * BNB);  // one accumulator per issuer (even / odd k-blocks)
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (kLN) ptx::cluster_arrive();  // #1: every CTA of the cluster is resident before any DSMEM store
  if (tracing) tr[2] = clock64();

  ptx::pdl_wait();     // outputs of the previous kernel of the step (activations, DecState) are visible
  if (tracing) tr[3] = clock64();
  if (warp == 0) {       // the activation tiles are on the critical path: request them before anything else
    if (ptx::elect_one()) {  // (also when the rollout is over: the prologue's barriers must complete)
      if (whole_k) {
        ptx::mbar_expect_tx(&full[2], kSwapWholeKB * kXBytes);
        ptx::tma_load_3d(smem_x, &tm_x3, &full[2], 0, b0, kb0);
      } else {
        for (int kb = 0; kb < n_pre; ++kb)
          ptx::tma_load_2d(smem_x + (size_t)kb * kXBytes, &tm_x, &full[kb], (kb0 + kb) * kSwapBK, b0);
      }
      if (!kLN && trace_cta) reinterpret_cast<long long*>(red)[0] = clock64();  // activation tiles requested
    }
    __syncwarp();
  }
  ptx::pdl_trigger();  // let the next kernel's prologue overlap this kernel
  const bool active = !(st != nullptr && st->done);

  if (warp == 0) {
    if (whole_k) {
      if (!active) { ptx::mbar_wait(&full[0], 0); ptx::mbar_wait(&full[1], 0); ptx::mbar_wait(&full[2], 0); }  // no copy may outlive the CTA
    } else if (active) {
      const uint64_t keep = ptx::policy_evict_last();
      // ring slot and phase advance by counters: `stages` is a run-time value, and a division per k-block in this
      // loop (and in the issue loop below) is latency on the way to the first MMA of the block
      int s = n_pre % stages;      // n_pre == stages (wraps to 0) or num_kb < stages (loop does not run)
      uint32_t ph = n_pre >= stages ? 1u : 0u;
      for (int kb = n_pre; kb < num_kb; ++kb, ++s) {
        if (s == stages) { s = 0; ph ^= 1u; }
        ptx::mbar_wait(&empty[s], ph ^ 1);
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(&full[s], kWBytes + kXBytes);
          ptx::tma_load_2d_hint(smem_w + (size_t)s * kWBytes, &tm_w, &full[s], (kb0 + kb) * kSwapBK, f0, keep);
          ptx::tma_load_2d(smem_x + (size_t)s * kXBytes, &tm_x, &full[s], (kb0 + kb) * kSwapBK, b0);
        }
        __syncwarp();
      }
    } else {
      for (int kb = 0; kb < n_pre; ++kb) ptx::mbar_wait(&full[kb], 0);  // no copy may outlive the CTA
    }
    __syncwarp();
    if (kLN) { ptx::cluster_wait(); ptx::cluster_arrive(); ptx::cluster_wait(); }
    if (kLN && KSPLIT > 1) ptx::cluster_sync();
  } else if (warp == 1 || warp >= kIssuer2) {
    if (active) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(BNB);
      const int first = warp == 1 ? 0 : warp - kIssuer2 + 1;  // this issuer's k-blocks: first, first + kIssuers, ...
      const uint32_t acc = tmem_base + (uint32_t)(first * BNB);
      const uint32_t w_lo0 = ptx::umma_desc_lo(ptx::smem_u32(smem_w)), x_lo0 = ptx::umma_desc_lo(ptx::smem_u32(smem_x));
      if (whole_k) {
        // Whole K resident: one straight run of MMAs with compile-time descriptor offsets.  The per-k-block loop below
        // spends ~550 cycles per iteration of this warp (barrier poll, fence, election, descriptor arithmetic in the
        // uniform datapath) against ~160 cycles of tensor pipe for its four N = 32 MMAs.
        ptx::mbar_wait(&full[0], 0);  // the weight halves were requested in the prologue, before the dependency wait
        ptx::mbar_wait(&full[1], 0);
        ptx::mbar_wait(&full[2], 0);  // the activations
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          if (!kLN && trace_cta && first == 0) reinterpret_cast<long long*>(red)[1] = clock64();  // first tile landed
          const uint32_t w_lo = w_lo0 + (uint32_t)first * (kWBytes >> 4);  // descriptor low words: address >> 4
          const uint32_t x_lo = x_lo0 + (uint32_t)first * (kXBytes >> 4);
#pragma unroll
          for (int i = 0; i < kSwapWholeKB / kIssuers; ++i)  // k-blocks first, first + kIssuers, ...
#pragma unroll
            for (int k = 0; k < kSwapBK / 16; ++k)  // 16 bf16 = 32 bytes along K inside the swizzle atom
              ptx::umma_f16_lohi(acc, w_lo + (uint32_t)(i * kIssuers) * (kWBytes >> 4) + k * 2,
                                 x_lo + (uint32_t)(i * kIssuers) * (kXBytes >> 4) + k * 2, idesc, (i != 0 || k != 0) ? 1u : 0u);
          ptx::umma_commit(tmem_full);  // same thread as the MMAs it tracks
          if (!kLN && trace_cta && first == kIssuers - 1) reinterpret_cast<long long*>(red)[2] = clock64();  // all issued
        }
        __syncwarp();
      } else {
        int s = first;  // ring slot of this issuer's next k-block
        uint32_t ph = 0u;
        for (int kb = first; kb < num_kb; kb += kIssuers, s += kIssuers) {
          while (s >= stages) { s -= stages; ph ^= 1u; }
          ptx::mbar_wait(&full[s], ph);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            if (!kLN && trace_cta && kb == 0) reinterpret_cast<long long*>(red)[1] = clock64();  // first tile landed
            const uint32_t w_lo = w_lo0 + (uint32_t)s * (kWBytes >> 4);  // descriptor low words: address >> 4
            const uint32_t x_lo = x_lo0 + (uint32_t)s * (kXBytes >> 4);
#pragma unroll
            for (int k = 0; k < kSwapBK / 16; ++k)  // 16 bf16 = 32 bytes along K inside the swizzle atom
              ptx::umma_f16_lohi(acc, w_lo + k * 2, x_lo + k * 2, idesc, (kb != first || k != 0) ? 1u : 0u);
            ptx::umma_commit(&empty[s]);
            if (kb + kIssuers >= num_kb) ptx::umma_commit(tmem_full);  // same thread as the MMAs it tracks
            if (!kLN && trace_cta && kb == num_kb - 1) reinterpret_cast<long long*>(red)[2] = clock64();  // all issued
          }
          __syncwarp();
        }
      }
    }
    __syncwarp();
    if (kLN) { ptx::cluster_wait(); ptx::cluster_arrive(); ptx::cluster_wait(); }
    if (kLN && KSPLIT > 1) ptx::cluster_sync();
  } else {
    constexpr int kEpi = swap_epi_threads(BNB);    // epilogue threads
    constexpr int kCStep = kEpi / 4;               // columns covered by one round of the epilogue warps (32 or 64)
    const int quarter = warp & 3;                  // TMEM lane quarter this warp may read
    const int cbeg = ((warp - 2) >> 2) * 32;       // first column chunk of this warp
    const int f = f0 + quarter * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    constexpr int kChunks = BNB / kCStep;  // 32-column chunks per epilogue warp (1 or 2)
    float xres[kLN ? kChunks : 1][kLN ? 32 : 1];  // kLN: residual rows of this thread's feature, fetched while the MMAs run
    SwapEpiPre epi_pre{0.f, 0};
    float ln_g = 0.f, ln_be = 0.f, ln_bias = 0.f;  // fetched while the MMAs run
    if constexpr (kLN) { ln_g = __ldg(ln.gamma + f); ln_be = __ldg(ln.beta + f); ln_bias = __ldg(ln.bias + f); }
    if constexpr (!kLN) epi_pre = epi.prepare(f);
    if constexpr (kLN) {
      if (active && ks == 0) {
#pragma unroll
        for (int ch = 0; ch < kChunks; ++ch)
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int row = b0 + cbeg + ch * kCStep + j;
            xres[ch][j] = row < M ? ln.x[(size_t)row * kD + f] : 0.f;
          }
      }
    }
    if (active) {
      ptx::mbar_wait(tmem_full, 0);
      ptx::tc_fence_after();
    }
    auto load_acc = [n_acc](uint32_t addr, float (&v)[32]) {  // sum of the issuers' accumulators
      ptx::tmem_ld32(addr, v);
#pragma unroll
      for (int i = 1; i < kIssuers; ++i) {
        if (i >= n_acc) break;
        float w[32];
        ptx::tmem_ld32(addr + (uint32_t)(i * BNB), w);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] += w[j];
      }
    };
    if (tracing) tr[4] = clock64();
    if (!kLN && tracing) {  // mainloop stamps left in shared memory by the producer and MMA warps
      tr[8] = reinterpret_cast<long long*>(red)[0];
      tr[9] = reinterpret_cast<long long*>(red)[1];
      tr[10] = reinterpret_cast<long long*>(red)[2];
    }
    if constexpr (!kLN) {
      if (active) {
#pragma unroll 1
        for (int c0 = cbeg; c0 < BNB; c0 += kCStep) {
          float v[32];
          load_acc(taddr + (uint32_t)c0, v);
          epi.chunk(f, b0 + c0, M, v, epi_pre);
        }
      }
    } else {
      const int e = threadIdx.x - 64;  // 0..kEpi-1
      const int fl = quarter * 32 + lane;  // feature within the tile
      const float bias_f = ln_bias;
      const bool lead = ks == 0;           // this CTA owns the epilogue of its feature tile
      if constexpr (KSPLIT > 1) {
        ptx::cluster_wait();  // #1
        if (active && !lead) {  // hand the partial accumulators of this k-slice to the k-slice-0 CTA
          const uint32_t dst0 = ptx::map_to_cta(part + ((size_t)(ks - 1) * kSwapBF + fl) * BNB, (uint32_t)(blockIdx.x & 3));
#pragma unroll 1
          for (int c0 = cbeg; c0 < BNB; c0 += kCStep) {
            float v[32];
            load_acc(taddr + (uint32_t)c0, v);
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              ptx::st_cluster_v4(dst0 + (uint32_t)(c0 + j) * 4, make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
          }
        }
        ptx::cluster_arrive();  // #2: partial tiles delivered
        ptx::cluster_wait();
      }
      if (active && lead) {
#pragma unroll
        for (int ch = 0; ch < kChunks; ++ch) {
          const int c0 = cbeg + ch * kCStep;
          float v[32], a[32], q[32];
          load_acc(taddr + (uint32_t)c0, v);
          if constexpr (KSPLIT > 1) {
#pragma unroll
            for (int p = 0; p < KSPLIT - 1; ++p) {
              const float4* src = reinterpret_cast<const float4*>(part + ((size_t)p * kSwapBF + fl) * BNB + c0);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 t4 = src[j];
                v[4 * j] += t4.x; v[4 * j + 1] += t4.y; v[4 * j + 2] += t4.z; v[4 * j + 3] += t4.w;
              }
            }
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            v[j] += bias_f + xres[ch][j];
            a[j] = v[j];
            q[j] = v[j] * v[j];
          }
          ptx::tmem_st32(taddr + (uint32_t)c0, v);  // park x + proj + bias for pass 2
          const float s1 = lane_transpose_sum(a, lane);
          const float s2 = lane_transpose_sum(q, lane);
          red[(0 * 4 + quarter) * BNB + c0 + lane] = s1;
          red[(1 * 4 + quarter) * BNB + c0 + lane] = s2;
        }
      }
      if (tracing) tr[8] = clock64();
      ptx::bar_sync(1, kEpi);
      if constexpr (KSPLIT == 1) ptx::cluster_wait();  // #1
      if (active && lead) {
        const uint32_t my_rank = blockIdx.x & 3;  // feature tile == cluster rank of the k-slice-0 CTAs
        for (int idx = e; idx < 2 * BNB; idx += kEpi) {
          const int which = idx / BNB, c = idx - which * BNB;
          const float tot = (red[(which * 4 + 0) * BNB + c] + red[(which * 4 + 1) * BNB + c]) +
                            (red[(which * 4 + 2) * BNB + c] + red[(which * 4 + 3) * BNB + c]);
          float* slot = clu + (my_rank * 2 + which) * BNB + c;
#pragma unroll
          for (uint32_t r = 0; r < 4; ++r) ptx::st_cluster_f32(ptx::map_to_cta(slot, r), tot);
        }
      }
      ptx::cluster_arrive();  // #2 / #3 (release: the DSMEM stores above)
      ptx::cluster_wait();
      if (active && lead) {
        for (int c = e; c < BNB; c += kEpi) {
          const float s1 = (clu[(0 * 2 + 0) * BNB + c] + clu[(1 * 2 + 0) * BNB + c]) +
                           (clu[(2 * 2 + 0) * BNB + c] + clu[(3 * 2 + 0) * BNB + c]);
          const float s2 = (clu[(0 * 2 + 1) * BNB + c] + clu[(1 * 2 + 1) * BNB + c]) +
                           (clu[(2 * 2 + 1) * BNB + c] + clu[(3 * 2 + 1) * BNB + c]);
          const float mean = s1 * (1.f / kD);
          const float var = fmaxf(s2 * (1.f / kD) - mean * mean, 0.f);
          const float rstd = rsqrtf(var + 1e-5f);
          reinterpret_cast<float2*>(red)[c] = make_float2(rstd, -mean * rstd);  // normalised = v * rstd - mean * rstd
        }
      }
      ptx::bar_sync(1, kEpi);
      if (tracing) tr[9] = clock64();
      const bool staged = ln.tma_store != 0;
      if (active && lead) {
        // staged: every MMA has completed (tmem_full), so the operand ring is free to hold the output tile
        float* x_stage = reinterpret_cast<float*>(smem);                             // [BNB][128] fp32
        bf16* xb_stage = reinterpret_cast<bf16*>(smem + (size_t)BNB * kSwapBF * 4);  // [BNB][128] bf16
        const float g = ln_g, be = ln_be;
#pragma unroll 1
        for (int c0 = cbeg; c0 < BNB; c0 += kCStep) {
          // the row statistics first, as 16 vector loads: read one by one between the stores below they would be
          // serialised behind them (the compiler must assume the stores alias the statistics)
          float4 rs[16];
          const float4* sp = reinterpret_cast<const float4*>(red) + (c0 >> 1);
#pragma unroll
          for (int i = 0; i < 16; ++i) rs[i] = sp[i];
          float v[32];
          ptx::tmem_ld32(taddr + (uint32_t)c0, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float sc = (j & 1) ? rs[j >> 1].z : rs[j >> 1].x, sh = (j & 1) ? rs[j >> 1].w : rs[j >> 1].y;
            v[j] = fmaf(fmaf(v[j], sc, sh), g, be);
          }
          if (staged) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              x_stage[(c0 + j) * kSwapBF + fl] = v[j];
              xb_stage[(c0 + j) * kSwapBF + fl] = __float2bfloat16_rn(v[j]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int row = b0 + c0 + j;
              if (row < M) {
                ln.x[(size_t)row * kD + f] = v[j];
                ln.xb[(size_t)row * kD + f] = __float2bfloat16_rn(v[j]);
              }
            }
          }
        }
        if (staged) ptx::fence_proxy_async();
      }
      if (tracing) tr[10] = clock64();
      if (staged) {
        ptx::bar_sync(1, kEpi);
        if (tracing) tr[11] = clock64();
        if (active && lead && e == 0) {  // rows >= M are clipped by the maps
          ptx::tma_store_2d(&ln.ts_x, smem, f0, b0);
          ptx::tma_store_2d(&ln.ts_xb, smem + (size_t)BNB * kSwapBF * 4, f0, b0);
          ptx::bulk_commit();
          ptx::bulk_wait_read0();
        }
      } else if (tracing) {
        tr[11] = tr[10];
      }
    }
  }
  if (tracing) {
    tr[5] = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tr[6]));
    tr[7] = kLN ? 1 : 0;
    const unsigned slot = atomicAdd(&g_swap_trace_n, 1u) & 255u;
    for (int i = 0; i < 12; ++i) g_swap_trace[slot * 12 + i] = tr[i];
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, kIssuers * BNB);
  if constexpr (swap_epi_has_finish<Epi>::value) {
    if (active) epi.template finish<4>(b0, BNB, M, reinterpret_cast<int*>(red));  // (launched with 4 k-slices)
  }
}

}  // namespace capdec
