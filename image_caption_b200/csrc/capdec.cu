// capdec.cu - C ABI + host runtime of the B200-native caption decode engine.
// See include/capdec.h for the boundary and DESIGN.md for the data layout.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <cstdlib>
#include <utility>

#include "engine.h"
#include "epilogue.cuh"
#include "gemm_simt.cuh"
#include "gemm_umma.cuh"
#include "gemm_swap.cuh"
#include "gemm_ws.cuh"
#include "beam.cuh"
#include "cross_attn_tma.cuh"
#include "attention.cuh"
#include "pointwise.cuh"
#include "layer_chain.cuh"
#include "cider.cuh"
#include "attention_dense.cuh"
#include "full_pass.cuh"

using namespace capdec;

// ------------------------------------------------------------------ errors
static std::string g_create_error;

static int fail(capdec_handle h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h) h->err = buf; else g_create_error = buf;
  return code;
}

#define CUDA_TRY(h, expr)                                                                     \
  do {                                                                                        \
    cudaError_t e_ = (expr);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return fail(h, CAPDEC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                  __FILE__, __LINE__);                                                        \
  } while (0)

// ------------------------------------------------------------------ per-launch timing
// When profiling is on (direct launches only), every kernel launch is bracketed by a pair of
// events on the launching stream; capdec_profile_report() reduces them per tag.
struct ProfScope {
  capdec_handle h;
  cudaStream_t s;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  ProfScope(capdec_handle h_, cudaStream_t s_) : h(h_), s(s_) {
    if (!h->profiling) return;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, s);
  }
  ~ProfScope() {
    if (!e0) return;
    cudaEventRecord(e1, s);
    h->prof.push_back({h->tag, e0, e1, h->tag_bytes});
  }
};
static inline void set_tag(capdec_handle h, const char* tag, double bytes) { h->tag = tag; h->tag_bytes = bytes; }

// Algorithmic bytes of one projection: weights + input rows + output rows (SURVEY 8d terms).
static inline double gemm_bytes(int M, int N, int K, size_t ew, size_t ea, size_t eo) {
  return (double)N * K * ew + (double)M * K * ea + (double)M * N * eo + (double)N * 4;
}

// ------------------------------------------------------------------ TMA tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// [rows, cols] row-major matrix, bf16 (or fp32 for the TF32 route); box = box_rows x one 128-byte swizzle row
// (64 bf16 / 32 fp32 columns = one UMMA atom row).
static int make_map(capdec_handle h, CUtensorMap* m, const void* ptr, int rows, int cols, int box_rows, bool f32 = false) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  const cuuint64_t es = f32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * es};
  cuuint32_t box[2] = {(cuuint32_t)(128 / es), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) rows=%d cols=%d", (int)r, rows, cols);
  return 0;
}

// Store view of a [rows, cols] activation matrix: box = 128 columns x box_rows, dense (unswizzled) in shared memory.
static int make_store_map(capdec_handle h, CUtensorMap* m, const void* ptr, int rows, int cols, int box_rows, bool f32) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  const cuuint64_t es = f32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * es};
  cuuint32_t box[2] = {(cuuint32_t)kSwapBF, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled (store view) failed (%d) rows=%d cols=%d", (int)r, rows, cols);
  return 0;
}

// bf16 [rows, cols] row-major matrix seen as [64, rows, cols/64]: one copy brings box_kb k-blocks of box_rows rows,
// laid out in shared memory as box_kb consecutive 128B-swizzled tiles (exactly the per-stage tiles of make_map).
// The TMA unit keeps about two bulk copies in flight per SM and each takes an L2 round trip (~650 cycles), so its
// throughput is ~2 x copy size / 650 cycles: few large copies instead of one small box per k-block.
static int make_map3(capdec_handle h, CUtensorMap* m, const void* ptr, int rows, int cols, int box_rows, int box_kb) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[3] = {(cuuint64_t)kUmmaBK, (cuuint64_t)rows, (cuuint64_t)(cols / kUmmaBK)};
  cuuint64_t strides[2] = {(cuuint64_t)cols * 2, (cuuint64_t)kUmmaBK * 2};
  cuuint32_t box[3] = {(cuuint32_t)kUmmaBK, (cuuint32_t)box_rows, (cuuint32_t)box_kb};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled (3-D view) failed (%d) rows=%d cols=%d", (int)r, rows, cols);
  return 0;
}

static int make_weight_maps(capdec_handle h, WeightMaps* wm, const void* ptr, int n, int k, bool f32 = false) {
  wm->raw = ptr; wm->rows = n; wm->k = k;
  const int bns[3] = {32, 64, 128};
  for (int i = 0; i < 3; ++i) {
    int rc = make_map(h, &wm->bn[i], ptr, n, k, bns[i], f32);
    if (rc) return rc;
  }
  if (!f32 && n % kSwapBF == 0 && (k / kUmmaBK) % kSwapWholeKB == 0) {
    return make_map3(h, &wm->w3, ptr, n, k, kSwapBF, kSwapWholeKB / 2);
  }
  return 0;
}

// ------------------------------------------------------------------ GEMM dispatch
static const int kMaxDynSmem = 227 * 1024;

template <int BN, typename Epi, typename In = bf16>
static int launch_umma_bn(capdec_handle h, cudaStream_t s, const CUtensorMap& ta, const CUtensorMap& tw, int M, int N,
                          int K, const Epi& epi, const DecState* st) {
  auto kern = gemm_umma_kernel<BN, Epi, In>;  // dynamic-smem limit raised once in configure_umma_kernels()
  const int num_kb = K / (128 / (int)sizeof(In));
  int stages = std::min(num_kb, 8);
  dim3 grid(N / BN, (M + kUmmaBM - 1) / kUmmaBM);
  // many tiles per SM (prefill): two CTAs per SM, so that one's epilogue overlaps the other's main loop
  const int budget = (long long)grid.x * grid.y > 2 * 148 ? 110 * 1024 : kMaxDynSmem;
  while (stages > 2 && umma_smem_bytes(BN, stages) > budget) --stages;
  ProfScope ps(h, s);
  kern<<<grid, kUmmaThreads, umma_smem_bytes(BN, stages), s>>>(ta, tw, M, N, K, stages, epi, st);
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  return 0;
}

// Pick the widest N tile that still yields at least one CTA per SM (the step GEMMs have few
// M tiles: M = number of live sequences).
template <typename Epi, typename In = bf16>
static int launch_umma(capdec_handle h, cudaStream_t s, const CUtensorMap& ta, const WeightMaps& wm, int M, int N, int K,
                       const Epi& epi, const DecState* st) {
  if (K % kUmmaBK != 0 || N % 128 != 0) return fail(h, CAPDEC_ERR_INVALID, "tcgen05 GEMM needs K%%64==0, N%%128==0 (N=%d K=%d)", N, K);
  const int tiles_m = (M + kUmmaBM - 1) / kUmmaBM;
  if (tiles_m * (N / 128) >= 148) return launch_umma_bn<128, Epi, In>(h, s, ta, wm.bn[2], M, N, K, epi, st);
  if (tiles_m * (N / 64) >= 148) return launch_umma_bn<64, Epi, In>(h, s, ta, wm.bn[1], M, N, K, epi, st);
  return launch_umma_bn<32, Epi, In>(h, s, ta, wm.bn[0], M, N, K, epi, st);
}

template <typename TA, typename TW, typename Epi>
static int launch_simt(capdec_handle h, cudaStream_t s, const TA* A, int lda, const TW* W, int ldw, int M, int N, int K,
                       const Epi& epi, const DecState* st) {
  if (K % kSimtBK != 0 || N % 4 != 0) return fail(h, CAPDEC_ERR_INVALID, "GEMM needs K%%16==0, N%%4==0 (N=%d K=%d)", N, K);
  dim3 grid((N + kSimtBN - 1) / kSimtBN, (M + kSimtBM - 1) / kSimtBM);
  ProfScope ps(h, s);
  gemm_simt_kernel<TA, TW, Epi><<<grid, kSimtThreads, 0, s>>>(A, lda, W, ldw, M, N, K, epi, st);
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  return 0;
}

template <typename Epi>
static cudaError_t configure_umma_epi() {
  cudaError_t e = cudaFuncSetAttribute(gemm_umma_kernel<32, Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_umma_kernel<64, Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_umma_kernel<128, Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  return e;
}
template <typename Epi>
static cudaError_t configure_umma_epi_tf32() {
  cudaError_t e = cudaFuncSetAttribute(gemm_umma_kernel<32, Epi, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_umma_kernel<64, Epi, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_umma_kernel<128, Epi, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  return e;
}
static cudaError_t configure_umma_kernels_tf32() {
  cudaError_t e = configure_umma_epi_tf32<EpiLinear<float, false>>();
  if (e == cudaSuccess) e = configure_umma_epi_tf32<EpiLinear<float, true>>();
  if (e == cudaSuccess) e = configure_umma_epi_tf32<EpiQKV<float>>();
  if (e == cudaSuccess) e = configure_umma_epi_tf32<EpiCrossKV<float>>();
  return e;
}
static cudaError_t configure_umma_kernels() {
  cudaError_t e = configure_umma_epi<EpiLinear<float, false>>();
  if (e == cudaSuccess) e = configure_umma_epi<EpiLinear<float, true>>();
  if (e == cudaSuccess) e = configure_umma_epi<EpiLinear<bf16, false>>();
  if (e == cudaSuccess) e = configure_umma_epi<EpiLinear<bf16, true>>();
  if (e == cudaSuccess) e = configure_umma_epi<EpiQKV<bf16>>();
  if (e == cudaSuccess) e = configure_umma_epi<EpiCrossKV<bf16>>();
  return e;
}

// One projection of the decode path: routes to tcgen05 (bf16 mode) or CUDA cores.
// `a_map` is the TMA view of A (bf16 mode only), `a_ptr` the raw pointer.
template <typename EpiF32, typename EpiB16>
static int gemm(capdec_handle h, cudaStream_t s, const void* a_ptr, const CUtensorMap* a_map, const void* w_ptr,
                const WeightMaps* w_maps, int M, int N, int K, const EpiF32& e32, const EpiB16& e16, const DecState* st) {
  if (!h->bf16) {
    // fp32 storage: CUDA cores (the strict-parity mode), or tcgen05 kind::tf32 straight from the fp32 buffers
    if (h->umma_f32 && a_map != nullptr && w_maps != nullptr && N % 128 == 0)
      return launch_umma<EpiF32, float>(h, s, *a_map, *w_maps, M, N, K, e32, st);
    return launch_simt<float, float, EpiF32>(h, s, (const float*)a_ptr, K, (const float*)w_ptr, K, M, N, K, e32, st);
  }
  if (h->use_umma) return launch_umma<EpiB16>(h, s, *a_map, *w_maps, M, N, K, e16, st);
  return launch_simt<bf16, bf16, EpiB16>(h, s, (const bf16*)a_ptr, K, (const bf16*)w_ptr, K, M, N, K, e16, st);
}

// ------------------------------------------------------------------ v2 launchers (bf16 + tcgen05)
// All v2 step kernels begin with griddepcontrol.wait, so they may be chained with programmatic
// dependent launch: the next kernel's prologue (barrier init, TMEM alloc, descriptor prefetch)
// overlaps the tail of the current one.
// INVARIANT the attention / sum_ln / head kernels rely on: they read DecState (st->step, st->done) BEFORE their wait.
// DecState is written only by the head kernel, at the end of a step.  That is safe only because every kernel of the
// chain executes griddepcontrol.wait in all of its CTAs before anything else of the step can have completed, and at
// least one full kernel (the layer-0 in_proj) separates the head from the first early reader: a grid can start under
// programmatic launch only after ALL CTAs of its predecessor have triggered, each of which waited for ITS predecessor
// to complete -- so when kernel k+2 starts, kernel k has completed and flushed.  A new kernel placed directly after the
// head must therefore read DecState after its wait (the projection chains in layer_chain.cuh do).  The PDL on/off
// parity test (test_v2_step_matches_v1_step) guards the ordering.
template <typename... KArgs, typename... Args>
static cudaError_t launch_ex(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, int cluster_x,
                             bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (cluster_x > 1) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = cluster_x;
    attrs[na].val.clusterDim.y = 1;
    attrs[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}

static int swap_stages(capdec_handle h, int bnb, int K, int ctas, bool ln = false) {
  const int num_kb = K / kSwapBK;
  int stages;
  if (h->force_stages > 0) {
    stages = std::min(num_kb, std::min(h->force_stages, 12));
  } else {
    // few CTAs: everything in flight at once; more CTAs than SMs: leave room for two per SM (the LayerNorm kernel is
    // one per SM by its registers anyway)
    const int budget = (ctas > 148 && !ln ? 100 : 200) * 1024;
    stages = std::min(num_kb, 12);
    while (stages > 2 && swap_smem_bytes(bnb, stages) > budget) --stages;
  }
  return stages;
}

// With two MMA issuer warps (even / odd k-blocks) the ring depth must be EVEN whenever the ring wraps.  Each issuer waits
// on full[slot] by phase PARITY.  With an even depth a slot is always drained by the same issuer, which therefore sees
// every phase of it.  With an odd depth the slot's fills alternate between the issuers, so an issuer waits for phase
// p + 2 having last seen phase p - and parity cannot tell "p + 1 still in flight" from "p + 2 complete".  The producer
// issues the copies in order, but they may LAND out of order: if k-block 4 lands before k-block 3 (3 stages), issuer 0
// drains 4, reaches 6, finds full[0] in (incomplete) phase 1, takes parity 0 as complete and issues MMAs on a slot still
// being written, then arrives on empty[0] a second time in one phase: an mbarrier over-arrival, which the hardware
// reports as "unspecified launch failure".  Found by the full-sequence pass (the first caller with ~400 such CTAs: 98 row
// tiles always failed at depths 3 and 5, 80 sometimes, and 4 / 6 never); a 12 544-row decode batch hit it too, and
// every 128-row LayerNorm GEMM of round 1-2 (depth 5) carried the race at low odds.
static int even_ring_for_two_issuers(int stages, int num_kb, int issuers, int bnb, int ksplit = 1) {
  if (issuers != 2 || stages >= num_kb || (stages & 1) == 0) return stages;
  // one deeper when it fits (B200, 128-row LayerNorm tiles: depth 6 beats 4 by 4 % over the ~400-CTA launches of the
  // full-sequence pass and is level at cfg3's 96), else one shallower
  if (stages + 1 <= num_kb && swap_smem_bytes(bnb, stages + 1, ksplit) <= kMaxDynSmem) return stages + 1;
  return stages > 2 ? stages - 1 : stages + 1;
}

// All of a CTA's K range in three copies (two weight halves + the activations) when it is exactly 8 k-blocks and the
// ring holds them all (env CAPDEC_NO_WHOLE_K=1 keeps one copy per k-block and operand).
static int swap_whole_k(capdec_handle h, int bnb, int num_kb, int stages) {
  return (!h->no_whole_k && num_kb == kSwapWholeKB && stages >= kSwapWholeKB && bnb <= 64) ? 1 : 0;
}

template <int BNB, bool kLN, typename Epi, int KSPLIT>
static int launch_swap_ks(capdec_handle h, cudaStream_t s, const WeightMaps& tw, const ActMaps& tx, int M, int n_out,
                          int K, const Epi& epi, const SwapLNParams& ln, const DecState* st) {
  dim3 grid(n_out / kSwapBF * KSPLIT, (M + BNB - 1) / BNB);
  int stages = swap_stages(h, BNB, K / KSPLIT, grid.x * grid.y, kLN);
  while (stages > 2 && swap_smem_bytes(BNB, stages, KSPLIT) > kMaxDynSmem) --stages;
  while (kLN && stages * swap_stage_bytes(BNB) < swap_ln_stage_out_bytes(BNB)) ++stages;  // the output tile is staged in the ring
  stages = even_ring_for_two_issuers(stages, K / kSwapBK / KSPLIT, swap_issuers(BNB, kLN), BNB, KSPLIT);
  const int whole = swap_whole_k(h, BNB, K / kSwapBK / KSPLIT, stages);
  ProfScope ps(h, s);
  cudaError_t e = launch_ex(gemm_swap_kernel<BNB, kLN, Epi, KSPLIT>, grid, dim3(swap_threads(BNB, kLN)), swap_smem_bytes(BNB, stages, KSPLIT), s,
                            kLN ? 4 * KSPLIT : 1, h->use_pdl, tw.bn[2], tx.tile, tw.w3, tx.whole,
                            whole, M, K, stages, epi, ln, st);
  if (e != cudaSuccess) return fail(h, CAPDEC_ERR_CUDA, "gemm_swap launch failed: %s (BNB=%d n_out=%d K=%d stages=%d ksplit=%d)",
                                    cudaGetErrorString(e), BNB, n_out, K, stages, KSPLIT);
  h->launches_this_enqueue++;
  return 0;
}

// LayerNorm GEMMs may divide K over the cluster (4 feature tiles x ksplit CTAs).
static int ln_ksplit(capdec_handle h, int K, int row_tiles) {
  // History (B200, tools/step_time.py): while every MMA cost ~150 cycles to issue (per-lane issue, see gemm_swap.cuh)
  // 8-CTA clusters paid for one small wave (cfg2: 476 -> 459 us/step).  With elect.sync issue the K = 512 chain is
  // ~2.4k cycles and the DSMEM hand-over costs more than it saves (cfg2: 368.6 with 2 slices, 361.2 with 1), so the
  // default is no split; CAPDEC_KSPLIT keeps the path testable.
  (void)row_tiles;
  int ks = h->force_ksplit > 0 ? h->force_ksplit : 1;
  if (h->bnb == 128 && ks > 2) ks = 2;  // the received partial tiles must fit beside the pipeline stages
  if (ks != 1 && ks != 2 && ks != 4) ks = 1;
  return (K / kSwapBK) % ks == 0 ? ks : 1;
}

template <int BNB, bool kLN, typename Epi>
static int launch_swap_bnb(capdec_handle h, cudaStream_t s, const WeightMaps& tw, const ActMaps& tx, int M, int n_out,
                           int K, const Epi& epi, const SwapLNParams& ln, const DecState* st, int zsplit = 1) {
  if constexpr (kLN) {
    const int ks = ln_ksplit(h, K, (M + BNB - 1) / BNB);
    if (ks == 4) return launch_swap_ks<BNB, kLN, Epi, 4>(h, s, tw, tx, M, n_out, K, epi, ln, st);
    if (ks == 2) return launch_swap_ks<BNB, kLN, Epi, 2>(h, s, tw, tx, M, n_out, K, epi, ln, st);
  }
  dim3 grid(n_out / kSwapBF, (M + BNB - 1) / BNB, zsplit);
  int stages = swap_stages(h, BNB, K / zsplit, grid.x * grid.y * grid.z, kLN);
  while (kLN && stages * swap_stage_bytes(BNB) < swap_ln_stage_out_bytes(BNB)) ++stages;  // the output tile is staged in the ring
  stages = even_ring_for_two_issuers(stages, K / kSwapBK / zsplit, swap_issuers(BNB, kLN), BNB);
  const int whole = swap_whole_k(h, BNB, K / kSwapBK / zsplit, stages);
  ProfScope ps(h, s);
  cudaError_t e = launch_ex(gemm_swap_kernel<BNB, kLN, Epi>, grid, dim3(swap_threads(BNB, kLN)), swap_smem_bytes(BNB, stages), s,
                            kLN ? 4 : 1, h->use_pdl, tw.bn[2], tx.tile, tw.w3, tx.whole,
                            whole, M, K, stages, epi, ln, st);
  if (e != cudaSuccess) return fail(h, CAPDEC_ERR_CUDA, "gemm_swap launch failed: %s (BNB=%d n_out=%d K=%d stages=%d)",
                                    cudaGetErrorString(e), BNB, n_out, K, stages);
  h->launches_this_enqueue++;
  return 0;
}

// Large batches, K = 512, no LayerNorm: persistent weight-stationary GEMM (gemm_ws.cuh), one wave of CTAs.
template <typename Epi>
static int launch_ws(capdec_handle h, cudaStream_t s, const WeightMaps& tw, const CUtensorMap& tx, int M, int n_out, const Epi& epi,
                     const DecState* st, int n_tiles_override = 0) {
  constexpr int BNB = 128;
  const int f_tiles = n_out / kSwapBF, n_tiles = n_tiles_override > 0 ? n_tiles_override : (M + BNB - 1) / BNB;
  const int groups = std::max(1, std::min(n_tiles, 148 / f_tiles));
  ProfScope ps(h, s);
  cudaError_t e = launch_ex(gemm_ws_kernel<BNB, Epi>, dim3(f_tiles, groups), dim3(kWsThreads), ws_smem_bytes(BNB), s, 1, h->use_pdl,
                            tw.bn[2], tx, n_tiles, epi, st, h->ws_dbg);
  if (e != cudaSuccess) return fail(h, CAPDEC_ERR_CUDA, "gemm_ws launch failed: %s (n_out=%d M=%d)", cudaGetErrorString(e), n_out, M);
  h->launches_this_enqueue++;
  return 0;
}
template <typename Epi>
static cudaError_t configure_ws_epi() {
  return cudaFuncSetAttribute(gemm_ws_kernel<128, Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
}
// In use when the batch gives a full wave of CTAs: (feature tiles) x (row groups) >= 96
static bool use_ws(capdec_handle h, int n_out, int M) {
  const int f_tiles = n_out / kSwapBF, n_tiles = (M + 127) / 128;
  if (h->no_ws || M < h->ws_min_m || f_tiles > 148) return false;
  return f_tiles * std::min(n_tiles, 148 / f_tiles) >= h->ws_min_ctas;
}

// bf16 [rows][H][t_pad][64] self K/V cache of one layer as a 4-D tensor for the in_proj epilogue's bulk store:
// box = {64, 1 position, 2 heads, 128 rows} (shared-memory image: [128 rows][128 features] bf16)
static int make_cache_store_map(capdec_handle h, CUtensorMap* m, const void* ptr, size_t rows, int t_pad) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[4] = {(cuuint64_t)kHD, (cuuint64_t)t_pad, (cuuint64_t)kH, (cuuint64_t)rows};
  cuuint64_t strides[3] = {(cuuint64_t)kHD * 2, (cuuint64_t)t_pad * kHD * 2, (cuuint64_t)kH * t_pad * kHD * 2};
  cuuint32_t box[4] = {(cuuint32_t)kHD, 1, 2, 128};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(h, CAPDEC_ERR_CUDA, "cuTensorMapEncodeTiled (cache store view) failed (%d)", (int)r);
  return 0;
}

template <bool kLN, typename Epi>
static int launch_swap(capdec_handle h, cudaStream_t s, const WeightMaps& tw, const ActMaps& tx, int M, int n_out, int K,
                       const Epi& epi, const SwapLNParams& ln, const DecState* st, int zsplit = 1) {
  if (K % kSwapBK != 0 || n_out % kSwapBF != 0 || (kLN && n_out != kD))
    return fail(h, CAPDEC_ERR_INVALID, "swap GEMM shape (n_out=%d K=%d ln=%d)", n_out, K, (int)kLN);
  switch (h->bnb) {
    case 32: return launch_swap_bnb<32, kLN, Epi>(h, s, tw, tx, M, n_out, K, epi, ln, st, zsplit);
    case 64: return launch_swap_bnb<64, kLN, Epi>(h, s, tw, tx, M, n_out, K, epi, ln, st, zsplit);
    default: return launch_swap_bnb<128, kLN, Epi>(h, s, tw, tx, M, n_out, K, epi, ln, st, zsplit);
  }
}

template <bool kLN, typename Epi>
static cudaError_t configure_swap_epi() {
  cudaError_t e = cudaFuncSetAttribute(gemm_swap_kernel<32, kLN, Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_swap_kernel<64, kLN, Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_swap_kernel<128, kLN, Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  return e;
}
template <int BNB, int KSPLIT>
static cudaError_t configure_swap_ln_ks() {
  auto k = gemm_swap_kernel<BNB, true, SwapNoLN, KSPLIT>;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess && KSPLIT * 4 > 8) e = cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  return e;
}
static cudaError_t configure_fast_kernels() {
  cudaError_t e = configure_swap_epi<false, SwapLinear<bf16, false>>();
  if (e == cudaSuccess) e = configure_swap_epi<false, SwapLinear<bf16, true>>();
  if (e == cudaSuccess) e = configure_swap_epi<false, SwapLinear<float, false>>();
  if (e == cudaSuccess) e = configure_swap_epi<false, SwapLinear<float, true>>();
  if (e == cudaSuccess) e = configure_swap_epi<false, SwapQKV<bf16>>();
  if (e == cudaSuccess) e = configure_swap_epi<false, SwapPartialLN>();
  if (e == cudaSuccess) e = configure_swap_epi<true, SwapNoLN>();
  if (e == cudaSuccess) e = configure_ws_epi<WsLinear<false>>();
  if (e == cudaSuccess) e = configure_ws_epi<WsLinear<true>>();
  if (e == cudaSuccess) e = configure_ws_epi<WsQKV>();
  if (e == cudaSuccess) e = configure_ws_epi<WsCrossKV>();
  if (e == cudaSuccess) e = configure_ws_epi<WsQKVFull>();
  if (e == cudaSuccess) e = cudaFuncSetAttribute(dense_attn_kernel<26>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dense_attn_smem(26));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(dense_attn_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dense_attn_smem(32));
  if (e == cudaSuccess) e = configure_swap_ln_ks<32, 2>();
  if (e == cudaSuccess) e = configure_swap_ln_ks<64, 2>();
  if (e == cudaSuccess) e = configure_swap_ln_ks<128, 2>();
  if (e == cudaSuccess) e = configure_swap_ln_ks<32, 4>();
  if (e == cudaSuccess) e = configure_swap_ln_ks<64, 4>();
  if (e == cudaSuccess) e = configure_swap_ln_ks<128, 4>();
  if (e == cudaSuccess) e = cudaFuncSetAttribute(cross_attn_tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(cross_attn_tma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(cross_attn_tma_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(vocab_head_kernel<1, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, vh_smem_bytes(4, 6));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(vocab_head_kernel<2, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, vh_smem_bytes(8, 6));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(vocab_head_kernel<4, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, vh_smem_bytes(16, 6));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(vocab_head_kernel<6, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, vh_smem_bytes(24, 5));
  return e;
}

// ------------------------------------------------------------------ cluster-resident projection chains (layer_chain.cuh)
template <int BNB, int CS>
static int chain_stages() {
  return std::min(10, (kMaxDynSmem - 1024 - ChSmem<BNB, CS>::kRing) / kChSlot);
}

template <int BNB, int CS, int KIND>
static int launch_chain_t(capdec_handle h, cudaStream_t s, const CUtensorMap& tm_b, const ChainParams& P, int M) {
  const int stages = chain_stages<BNB, CS>();
  const int tiles = (M + BNB - 1) / BNB;
  ProfScope ps(h, s);
  cudaError_t e = launch_ex(chain_kernel<BNB, CS, KIND>, dim3(CS * tiles), dim3(kChThreads), ChSmem<BNB, CS>::bytes(stages), s, CS,
                            h->use_pdl, tm_b, P, stages);
  if (e != cudaSuccess)
    return fail(h, CAPDEC_ERR_CUDA, "chain kernel launch failed: %s (BNB=%d CS=%d kind=%d stages=%d)", cudaGetErrorString(e), BNB, CS,
                KIND, stages);
  h->launches_this_enqueue++;
  return 0;
}

template <int KIND>
static int launch_chain(capdec_handle h, cudaStream_t s, const CUtensorMap& tm_b, const ChainParams& P, int M) {
  if (h->chain_cs == 8) {
    if (h->chain_bnb == 32) return launch_chain_t<32, 8, KIND>(h, s, tm_b, P, M);
    return launch_chain_t<64, 8, KIND>(h, s, tm_b, P, M);
  }
  if (h->chain_bnb == 32) return launch_chain_t<32, 4, KIND>(h, s, tm_b, P, M);
  return launch_chain_t<64, 4, KIND>(h, s, tm_b, P, M);
}

template <int BNB, int CS>
static cudaError_t configure_chain_t(int* max_clusters) {
  cudaError_t e = cudaSuccess;
  auto set = [&](auto kern) {
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  };
  set(chain_kernel<BNB, CS, kChainQKV>);
  set(chain_kernel<BNB, CS, kChainAttnOut>);
  set(chain_kernel<BNB, CS, kChainFFN>);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(CS * 64);
  cfg.blockDim = dim3(kChThreads);
  cfg.dynamicSmemBytes = ChSmem<BNB, CS>::bytes(chain_stages<BNB, CS>());
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  e = cudaOccupancyMaxActiveClusters(&n, chain_kernel<BNB, CS, kChainFFN>, &cfg);
  *max_clusters = n;
  return e;
}

// Units of the packed stream before the first unit of (layer l, kind): [rank][layer][in_proj | attn-out | ffn]
template <int CS>
static size_t chain_unit_offset(int l, int kind) {
  using SH = ChainShape<CS>;
  return (size_t)l * SH::UL + (kind == kChainQKV ? 0 : (kind == kChainAttnOut ? SH::UQ : SH::UQ + SH::UC));
}

// ------------------------------------------------------------------ allocation helpers
template <typename T>
static int dev_alloc(capdec_handle h, T** out, size_t count) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
  if (e != cudaSuccess)
    return fail(h, CAPDEC_ERR_CUDA, "cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e));
  h->allocs.push_back(p);
  *out = static_cast<T*>(p);
  return 0;
}
static int dev_alloc_bytes(capdec_handle h, void** out, size_t bytes) {
  unsigned char* p = nullptr;
  int rc = dev_alloc<unsigned char>(h, &p, bytes);
  *out = p;
  return rc;
}
#define TRY(expr) do { int rc_ = (expr); if (rc_) return rc_; } while (0)

// ------------------------------------------------------------------ lifecycle
extern "C" int capdec_abi_version(void) { return CAPDEC_ABI_VERSION; }

extern "C" const char* capdec_last_error(capdec_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

extern "C" int64_t capdec_launch_count(capdec_handle h) { return h ? h->launches : 0; }

static void free_all(capdec_handle h) {
  for (auto& r : h->prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  h->prof.clear();
  for (auto& kv : h->graphs) {
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    if (kv.second.graph) cudaGraphDestroy(kv.second.graph);
  }
  h->graphs.clear();
  for (void* p : h->allocs) cudaFree(p);
  h->allocs.clear();
  if (h->copy_stream) { cudaStreamDestroy(h->copy_stream); h->copy_stream = nullptr; }
  for (auto& slot : h->copy_ev)
    for (auto& e : slot) if (e) { cudaEventDestroy(e); e = nullptr; }
  if (h->copy_start) { cudaEventDestroy(h->copy_start); h->copy_start = nullptr; }
}

extern "C" int capdec_destroy(capdec_handle h) {
  if (!h) return CAPDEC_OK;
  cudaSetDevice(h->device);
  if (getenv("CAPDEC_TRACE_GEMM")) {  // tuning aid: stamps of the last swap-GEMM launches (gemm_swap.cuh)
    long long* dptr = nullptr;
    unsigned n = 0;
    cudaMemcpyFromSymbol(&dptr, g_swap_trace, sizeof dptr);
    cudaMemcpyFromSymbol(&n, g_swap_trace_n, sizeof n);
    if (dptr) {
      static long long t[256 * 12];
      cudaMemcpy(t, dptr, sizeof t, cudaMemcpyDeviceToHost);
      const unsigned first = n > 48 ? n - 48 : 0;
      long long prev_exit = 0;
      for (unsigned i = first; i < n; ++i) {
        const long long* r = t + (i & 255u) * 12;
        fprintf(stderr, "gemm#%u ln=%lld gap_ns=%lld  cycles: prologue %lld  pdl_wait %lld  load+mma %lld  epilogue %lld (pass1 %lld, exchange %lld, pass2 %lld: stage %lld, barrier %lld, store %lld)  | in-kernel ns %lld\n", i, r[7],
                prev_exit ? r[0] - prev_exit : 0, r[2] - r[1], r[3] - r[2], r[4] - r[3], r[5] - r[4],
                r[7] ? r[8] - r[4] : 0, r[7] ? r[9] - r[8] : 0, r[7] ? r[5] - r[9] : 0, r[7] ? r[10] - r[9] : 0, r[7] ? r[11] - r[10] : 0, r[7] ? r[5] - r[11] : 0, r[6] - r[0]);
        if (!r[7] && r[8])
          fprintf(stderr, "          mainloop, cycles since the pdl wait returned: activation tiles requested %lld, first tile landed %lld, all mma issued %lld, accumulator seen %lld\n",
                  r[8] - r[3], r[9] - r[3], r[10] - r[3], r[4] - r[3]);
        prev_exit = r[6];
      }
    }
  }
  if (getenv("CAPDEC_TRACE_CHAIN")) {  // tuning aid: stamps of the last projection-chain launches (layer_chain.cuh)
    long long* dptr = nullptr;
    unsigned n = 0;
    cudaMemcpyFromSymbol(&dptr, g_chain_trace, sizeof dptr);
    cudaMemcpyFromSymbol(&n, g_chain_trace_n, sizeof n);
    if (dptr) {
      static long long t[256 * 64];
      cudaMemcpy(t, dptr, sizeof t, cudaMemcpyDeviceToHost);
      const unsigned first = n > 30 ? n - 30 : 0;
      long long prev_exit = 0;
      for (unsigned i = first; i < n; ++i) {
        const long long* r = t + (i & 255u) * 64;
        fprintf(stderr, "chain#%u kind=%lld gap_ns=%lld in-kernel ns %lld | epilogue cycles since entry:", i, r[1], prev_exit ? r[0] - prev_exit : 0, r[62] - r[0]);
        for (int j = 3; j < 40 && r[j]; ++j) fprintf(stderr, " [%d]%lld", j, r[j] - r[2]);
        fprintf(stderr, " end %lld | mma:", r[63] - r[2]);
        for (int j = 40; j < 52 && r[j]; ++j) fprintf(stderr, " [%d]%lld", j - 40, r[j] - r[2]);
        fprintf(stderr, "\n");
        prev_exit = r[62];
      }
    }
  }
  free_all(h);
  delete h;
  return CAPDEC_OK;
}

static int create_impl(capdec_handle h) {
  const capdec_config& c = h->cfg;
  const int L = c.n_layer, F = c.d_ff;
  const size_t M = (size_t)c.max_images * c.max_group;
  const size_t es = h->esz;
  LayerParams* lp = h->layer;
  for (int l = 0; l < L; ++l) {
    TRY(dev_alloc_bytes(h, &lp[l].w_qkv, (size_t)3 * kD * kD * es));
    TRY(dev_alloc_bytes(h, &lp[l].w_so, (size_t)kD * kD * es));
    TRY(dev_alloc_bytes(h, &lp[l].w_cq, (size_t)kD * kD * es));
    TRY(dev_alloc_bytes(h, &lp[l].w_co, (size_t)kD * kD * es));
    TRY(dev_alloc_bytes(h, &lp[l].w_1, (size_t)F * kD * es));
    TRY(dev_alloc_bytes(h, &lp[l].w_2, (size_t)kD * F * es));
    TRY(dev_alloc(h, &lp[l].b_qkv, 3 * kD));
    TRY(dev_alloc(h, &lp[l].b_so, kD));
    TRY(dev_alloc(h, &lp[l].b_cq, kD));
    TRY(dev_alloc(h, &lp[l].b_co, kD));
    TRY(dev_alloc(h, &lp[l].b_1, F));
    TRY(dev_alloc(h, &lp[l].b_2, kD));
    for (int j = 0; j < 3; ++j) {
      TRY(dev_alloc(h, &lp[l].ln_w[j], kD));
      TRY(dev_alloc(h, &lp[l].ln_b[j], kD));
    }
  }
  TRY(dev_alloc_bytes(h, &h->w_ckv, (size_t)L * 2 * kD * kD * es));
  TRY(dev_alloc(h, &h->b_ckv, (size_t)L * 2 * kD));
  TRY(dev_alloc(h, &h->emb, (size_t)c.vocab * kD));
  TRY(dev_alloc(h, &h->pe, (size_t)c.pe_rows * kD));
  TRY(dev_alloc(h, &h->fc_w, (size_t)kVPad * kD));
  TRY(dev_alloc(h, &h->fc_b, kVPad));

  h->self_layer_elems = M * kH * h->t_pad * kHD;
  TRY(dev_alloc_bytes(h, &h->self_k, (size_t)L * h->self_layer_elems * es));
  TRY(dev_alloc_bytes(h, &h->self_v, (size_t)L * h->self_layer_elems * es));
  TRY(dev_alloc_bytes(h, &h->cross_kv, (size_t)L * 2 * c.max_images * c.max_mem * kD * es));

  if (h->fast)
    for (int l = 0; l < L; ++l) {
      TRY(make_cache_store_map(h, &lp[l].ts_k, (char*)h->self_k + (size_t)l * h->self_layer_elems * es, M, h->t_pad));
      TRY(make_cache_store_map(h, &lp[l].ts_v, (char*)h->self_v + (size_t)l * h->self_layer_elems * es, M, h->t_pad));
    }
  TRY(dev_alloc(h, &h->x, M * kD));
  TRY(dev_alloc(h, &h->t, M * kD));
  TRY(dev_alloc(h, &h->logits, M * kVPad));
  if (h->bf16) {
    TRY(dev_alloc_bytes(h, &h->xb, M * kD * es));
    TRY(dev_alloc_bytes(h, &h->mem_b, (size_t)c.max_images * c.max_mem * kD * es));
  }
  TRY(dev_alloc_bytes(h, &h->q, M * kD * es));
  TRY(dev_alloc_bytes(h, &h->a, M * kD * es));
  TRY(dev_alloc_bytes(h, &h->hbuf, M * F * es));
  TRY(dev_alloc(h, &h->fin, M));
  TRY(dev_alloc(h, &h->st, 1));
  TRY(dev_alloc(h, &h->cp, 1));
  CUDA_TRY(h, cudaMemset(h->st, 0, sizeof(DecState)));
  if (h->use_umma) CUDA_TRY(h, configure_umma_kernels());
  if (h->umma_f32 || h->use_umma) CUDA_TRY(h, configure_umma_kernels_tf32());
  if (h->fast) {
    {  // does dynamic shared memory start on a 1 KB boundary?  (lets the cross-attention drop its alignment slack)
      unsigned* d = nullptr;
      unsigned off = 1;
      if (cudaMalloc(&d, sizeof(unsigned)) == cudaSuccess) {
        smem_base_probe_kernel<<<1, 32, 4096>>>(d);
        if (cudaMemcpy(&off, d, sizeof off, cudaMemcpyDeviceToHost) != cudaSuccess) off = 1;
        cudaFree(d);
      }
      cudaGetLastError();
      h->dyn_smem_aligned = off == 0;
    }
    if (const char* e = getenv("CAPDEC_NO_CA_COMPACT")) h->no_ca_compact = atoi(e) != 0;
    TRY(dev_alloc(h, &h->fc_wT, (size_t)kD * kVPad));
    h->part_rows = (int)std::min<size_t>(M, 512);
    TRY(dev_alloc(h, &h->part, (size_t)4 * h->part_rows * kD));
    TRY(dev_alloc_bytes(h, (void**)&h->tickets, 64 * sizeof(unsigned)));  // one per 32-row tile of <= 512 rows (and slack)
    CUDA_TRY(h, cudaMemset(h->tickets, 0, 64 * sizeof(unsigned)));
    CUDA_TRY(h, configure_fast_kernels());
  }
  if (h->chain) {
    const size_t units = (size_t)h->cfg.n_layer * (h->chain_cs == 8 ? ChainShape<8>::UL : ChainShape<4>::UL);
    TRY(dev_alloc(h, &h->chain_w, (size_t)h->chain_cs * units * kChUnit));
    const size_t Mpad = (M + 63) / 64 * 64;
    TRY(dev_alloc(h, &h->chain_part, (size_t)h->chain_cs * Mpad * kD));
    if (h->chain_cs == 8) {
      CUDA_TRY(h, (configure_chain_t<32, 8>(&h->chain_clusters[0])));
      CUDA_TRY(h, (configure_chain_t<64, 8>(&h->chain_clusters[1])));
    } else {
      CUDA_TRY(h, (configure_chain_t<32, 4>(&h->chain_clusters[0])));
      CUDA_TRY(h, (configure_chain_t<64, 4>(&h->chain_clusters[1])));
    }
    if (getenv("CAPDEC_CHAIN_INFO"))
      fprintf(stderr, "capdec: projection chains, %d CTAs per cluster, co-resident clusters: %d (32 rows) / %d (64 rows)\n", h->chain_cs,
              h->chain_clusters[0], h->chain_clusters[1]);
  }
  return 0;
}

extern "C" int capdec_create(const capdec_config* cfg, int device, capdec_handle* out) {
  if (!cfg || !out) return fail(nullptr, CAPDEC_ERR_INVALID, "null argument");
  *out = nullptr;
  if (cfg->d_model != kD || cfg->n_head != kH)
    return fail(nullptr, CAPDEC_ERR_INVALID, "only d_model=512, nhead=8 are built (got %d, %d)", cfg->d_model, cfg->n_head);
  if (cfg->n_layer < 1 || cfg->n_layer > CAPDEC_MAX_LAYERS) return fail(nullptr, CAPDEC_ERR_INVALID, "n_layer %d out of range", cfg->n_layer);
  if (cfg->d_ff < 128 || cfg->d_ff % 128 != 0) return fail(nullptr, CAPDEC_ERR_INVALID, "d_ff %d must be a multiple of 128", cfg->d_ff);
  if (cfg->vocab < 2 || cfg->vocab > kVPad) return fail(nullptr, CAPDEC_ERR_INVALID, "vocab %d must be in [2,%d]", cfg->vocab, kVPad);
  if (cfg->max_len < 2 || cfg->max_len > 128) return fail(nullptr, CAPDEC_ERR_INVALID, "max_len %d must be in [2,128]", cfg->max_len);
  if (cfg->pe_rows < 1 || cfg->max_mem < 1 || cfg->max_images < 1 || cfg->max_group < 1)
    return fail(nullptr, CAPDEC_ERR_INVALID, "sizes must be positive");
  if (cfg->precision != CAPDEC_PREC_FP32 && cfg->precision != CAPDEC_PREC_BF16 && cfg->precision != CAPDEC_PREC_TF32)
    return fail(nullptr, CAPDEC_ERR_INVALID, "unknown precision %d", cfg->precision);

  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    return fail(nullptr, CAPDEC_ERR_NO_DEVICE, "CUDA device %d not available (%d visible)", device, ndev);
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10)
    return fail(nullptr, CAPDEC_ERR_NO_DEVICE, "device %d is not sm_100-class (this library is built for sm_100a only)", device);
  if (cudaSetDevice(device) != cudaSuccess) return fail(nullptr, CAPDEC_ERR_CUDA, "cudaSetDevice(%d) failed", device);

  capdec_handle h = new capdec_engine();
  h->cfg = *cfg;
  h->device = device;
  h->bf16 = cfg->precision == CAPDEC_PREC_BF16;
  h->use_umma = h->bf16 && !(cfg->flags & CAPDEC_FLAG_NO_UMMA);
  h->umma_f32 = cfg->precision == CAPDEC_PREC_TF32;  // fp32 storage and step kernels, projections on tcgen05 kind::tf32
  h->use_graph = !(cfg->flags & CAPDEC_FLAG_NO_GRAPH);
  h->esz = h->bf16 ? 2 : 4;
  h->t_pad = cfg->max_len <= 64 ? 64 : 128;
  h->fast = h->use_umma && !(cfg->flags & CAPDEC_FLAG_V1_STEP);
  h->use_pdl = h->fast && !(cfg->flags & CAPDEC_FLAG_NO_PDL);
  h->chain = h->fast && cfg->d_ff == kChFF;
  if (const char* e = getenv("CAPDEC_CHAIN")) { if (atoi(e) == 0) h->chain = false; h->chain_always = atoi(e) == 2; }
  if (const char* e = getenv("CAPDEC_CHAIN_CS")) h->chain_cs = atoi(e) == 4 ? 4 : 8;
  if (const char* e = getenv("CAPDEC_CHAIN_BNB")) h->force_chain_bnb = atoi(e) == 64 ? 64 : (atoi(e) == 32 ? 32 : 0);
  if (getenv("CAPDEC_TRACE_GEMM")) {
    long long* buf = nullptr;
    if (cudaMalloc(&buf, 256 * 12 * sizeof(long long)) == cudaSuccess) {
      cudaMemset(buf, 0, 256 * 12 * sizeof(long long));
      cudaMemcpyToSymbol(g_swap_trace, &buf, sizeof buf);
    }
  }
  if (getenv("CAPDEC_TRACE_CHAIN")) {
    long long* buf = nullptr;
    if (cudaMalloc(&buf, 256 * 64 * sizeof(long long)) == cudaSuccess) {
      cudaMemset(buf, 0, 256 * 64 * sizeof(long long));
      cudaMemcpyToSymbol(g_chain_trace, &buf, sizeof buf);
    }
  }
  if (const char* e = getenv("CAPDEC_LN_TMA_STORE")) h->force_ln_tma_store = atoi(e) != 0 ? 1 : 0;
  if (const char* e = getenv("CAPDEC_NO_CROSS_STREAM")) h->no_cross_stream = atoi(e) != 0;
  if (const char* e = getenv("CAPDEC_NO_WS")) h->no_ws = atoi(e) != 0;
  if (const char* e = getenv("CAPDEC_SELF_ATTN_ITS")) h->self_attn_its = atoi(e);
  if (const char* e = getenv("CAPDEC_NO_TF32_HEAD")) h->no_tf32_head = atoi(e) != 0;
  if (const char* e = getenv("CAPDEC_TF32_HEAD_MIN_M")) h->tf32_head_min_m = std::max(1, atoi(e));
  if (const char* e = getenv("CAPDEC_NO_WS_PREFILL")) h->no_ws_prefill = atoi(e) != 0;
  if (const char* e = getenv("CAPDEC_WS_MIN_CTAS")) h->ws_min_ctas = atoi(e);
  if (const char* e = getenv("CAPDEC_WS_MIN_M")) h->ws_min_m = std::max(128, atoi(e));
  if (const char* e = getenv("CAPDEC_WS_DBG")) h->ws_dbg = atoi(e);  // tuning aid (wrong results): 1 = no output staging/stores, 2 = no weight load, 4 = no activation loads after tile 0, 8 = no bulk stores, 16 = no staging
  if (const char* e = getenv("CAPDEC_NO_WHOLE_K")) h->no_whole_k = atoi(e) != 0;
  if (const char* e = getenv("CAPDEC_FFN2_FUSE")) h->no_ffn2_fuse = atoi(e) == 0;
  if (const char* e = getenv("CAPDEC_HEAD_RW")) { const int v = atoi(e); h->force_head_rw = (v == 1 || v == 2 || v == 4 || v == 6) ? v : 0; }
  if (const char* e = getenv("CAPDEC_BEAM_V1")) h->beam_v1 = atoi(e) != 0;
  if (const char* e = getenv("CAPDEC_GRAPH_STEPS")) h->graph_steps = std::max(1, atoi(e));
  if (const char* e = getenv("CAPDEC_BNB")) h->force_bnb = atoi(e);
  if (const char* e = getenv("CAPDEC_STAGES")) h->force_stages = atoi(e);
  if (const char* e = getenv("CAPDEC_KSPLIT")) h->force_ksplit = atoi(e);
  if (const char* e = getenv("CAPDEC_NO_FFN2_SPLIT")) h->no_ffn2_split = atoi(e) != 0;
  if (const char* e = getenv("CAPDEC_NO_PDL")) { if (atoi(e)) h->use_pdl = false; }
  if (cfg->flags & CAPDEC_FLAG_BATCH_INVARIANT) {  // one route for every batch size (see capdec.h)
    h->chain = false;
    h->no_ws = true;
    h->force_bnb = 64;
    h->no_ffn2_split = true;
    h->no_tf32_head = true;  // (the fused CUDA-core head at every batch size)
  }
  int rc = create_impl(h);
  if (rc) {
    g_create_error = h->err;
    free_all(h);
    delete h;
    return rc;
  }
  *out = h;
  return CAPDEC_OK;
}

// ------------------------------------------------------------------ weights
static int pack_matrix(capdec_handle h, cudaStream_t s, const float* src, void* dst, int rows, int cols, int rows_out) {
  const size_t n = (size_t)rows_out * cols;
  const int threads = 256;
  const unsigned blocks = (unsigned)((n + threads - 1) / threads);
  if (h->bf16) pack_rows_kernel<bf16><<<blocks, threads, 0, s>>>(src, (bf16*)dst, rows, cols, rows_out);
  else pack_rows_kernel<float><<<blocks, threads, 0, s>>>(src, (float*)dst, rows, cols, rows_out);
  CUDA_TRY(h, cudaGetLastError());
  h->launches++;
  return 0;
}
static int copy_f32(capdec_handle h, cudaStream_t s, const float* src, float* dst, size_t n) {
  CUDA_TRY(h, cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
  return 0;
}

extern "C" int capdec_set_weights(capdec_handle h, const capdec_weights* w, void* stream) {
  if (!h || !w) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  const int L = h->cfg.n_layer, F = h->cfg.d_ff, V = h->cfg.vocab;
  if (!w->embedding || !w->pe || !w->fc_out_w || !w->fc_out_b) return fail(h, CAPDEC_ERR_INVALID, "null weight pointer");
  TRY(copy_f32(h, s, w->embedding, h->emb, (size_t)V * kD));
  TRY(copy_f32(h, s, w->pe, h->pe, (size_t)h->cfg.pe_rows * kD));
  // vocabulary projection stays fp32 in both modes; rows V..127 are zero
  pack_rows_kernel<float><<<(kVPad * kD + 255) / 256, 256, 0, s>>>(w->fc_out_w, h->fc_w, V, kD, kVPad);
  pack_rows_kernel<float><<<1, 256, 0, s>>>(w->fc_out_b, h->fc_b, 1, V, 1);
  CUDA_TRY(h, cudaMemsetAsync(h->fc_b + V, 0, (kVPad - V) * sizeof(float), s));
  h->launches += 2;
  if (h->fast) {
    transpose_pad_kernel<<<(kD * kVPad + 255) / 256, 256, 0, s>>>(w->fc_out_w, h->fc_wT, V, kD, kVPad);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
  }
  for (int l = 0; l < L; ++l) {
    const capdec_layer_weights& lw = w->layer[l];
    LayerParams& lp = h->layer[l];
    const void* ptrs[] = {lw.self_in_w, lw.self_in_b, lw.self_out_w, lw.self_out_b, lw.cross_in_w, lw.cross_in_b,
                          lw.cross_out_w, lw.cross_out_b, lw.lin1_w, lw.lin1_b, lw.lin2_w, lw.lin2_b,
                          lw.norm_w[0], lw.norm_w[1], lw.norm_w[2], lw.norm_b[0], lw.norm_b[1], lw.norm_b[2]};
    for (const void* p : ptrs) if (!p) return fail(h, CAPDEC_ERR_INVALID, "null weight pointer in layer %d", l);
    TRY(pack_matrix(h, s, lw.self_in_w, lp.w_qkv, 3 * kD, kD, 3 * kD));
    TRY(pack_matrix(h, s, lw.self_out_w, lp.w_so, kD, kD, kD));
    TRY(pack_matrix(h, s, lw.cross_in_w, lp.w_cq, kD, kD, kD));  // rows [0,512): the q projection
    TRY(pack_matrix(h, s, lw.cross_out_w, lp.w_co, kD, kD, kD));
    TRY(pack_matrix(h, s, lw.lin1_w, lp.w_1, F, kD, F));
    TRY(pack_matrix(h, s, lw.lin2_w, lp.w_2, kD, F, kD));
    // rows [512,1536) of the cross in_proj: K then V projection, concatenated over layers
    TRY(pack_matrix(h, s, lw.cross_in_w + (size_t)kD * kD, (char*)h->w_ckv + (size_t)l * 2 * kD * kD * h->esz, 2 * kD, kD, 2 * kD));
    TRY(copy_f32(h, s, lw.cross_in_b + kD, h->b_ckv + (size_t)l * 2 * kD, 2 * kD));
    TRY(copy_f32(h, s, lw.self_in_b, lp.b_qkv, 3 * kD));
    TRY(copy_f32(h, s, lw.self_out_b, lp.b_so, kD));
    TRY(copy_f32(h, s, lw.cross_in_b, lp.b_cq, kD));
    TRY(copy_f32(h, s, lw.cross_out_b, lp.b_co, kD));
    TRY(copy_f32(h, s, lw.lin1_b, lp.b_1, F));
    TRY(copy_f32(h, s, lw.lin2_b, lp.b_2, kD));
    for (int j = 0; j < 3; ++j) {
      TRY(copy_f32(h, s, lw.norm_w[j], lp.ln_w[j], kD));
      TRY(copy_f32(h, s, lw.norm_b[j], lp.ln_b[j], kD));
    }
    if (h->use_umma || h->umma_f32) {
      const bool f32 = h->umma_f32;
      TRY(make_weight_maps(h, &lp.m_qkv, lp.w_qkv, 3 * kD, kD, f32));
      TRY(make_weight_maps(h, &lp.m_so, lp.w_so, kD, kD, f32));
      TRY(make_weight_maps(h, &lp.m_cq, lp.w_cq, kD, kD, f32));
      TRY(make_weight_maps(h, &lp.m_co, lp.w_co, kD, kD, f32));
      TRY(make_weight_maps(h, &lp.m_1, lp.w_1, F, kD, f32));
      TRY(make_weight_maps(h, &lp.m_2, lp.w_2, kD, F, f32));
    }
  }
  if (h->use_umma || h->umma_f32) TRY(make_weight_maps(h, &h->m_ckv, h->w_ckv, L * 2 * kD, kD, h->umma_f32));
  if (h->fast || h->umma_f32) TRY(make_weight_maps(h, &h->m_fc, h->fc_w, kVPad, kD, true));  // vocabulary projection on kind::tf32
  if (h->chain) {
    ChainPackSrc src{};
    src.n_layer = L;
    for (int l = 0; l < L; ++l) {
      const capdec_layer_weights& lw = w->layer[l];
      src.layer[l] = ChainPackLayer{lw.self_in_w, lw.self_out_w, lw.cross_in_w, lw.cross_out_w, lw.lin1_w, lw.lin2_w};
    }
    const long long total = (long long)h->chain_cs * L * (h->chain_cs == 8 ? ChainShape<8>::UL : ChainShape<4>::UL) * 1024;
    if (h->chain_cs == 8) pack_chain_kernel<8><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(src, h->chain_w);
    else pack_chain_kernel<4><<<(unsigned)((total + 255) / 256), 256, 0, s>>>(src, h->chain_w);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
  }
  h->have_weights = true;
  return CAPDEC_OK;
}

// ------------------------------------------------------------------ prefill
static int finish_prefill(capdec_handle h, int n_images, int mem_len) {
  h->launches += h->launches_this_enqueue;
  if (h->fast && (h->n_images != n_images || h->mem_len != mem_len || !h->prefilled_once)) {
    if (mem_len > 256) return fail(h, CAPDEC_ERR_CAPACITY, "memory length %d > 256 is not supported by the TMA cross-attention", mem_len);
    TRY(make_map(h, &h->tm_ckv, h->cross_kv, h->cfg.n_layer * 2 * n_images * kH * mem_len, kHD, ca_key_rows(mem_len)));
    h->prefilled_once = true;
  }
  h->n_images = n_images;
  h->mem_len = mem_len;
  h->prefilled = true;
  return CAPDEC_OK;
}

// Images [img0, img0 + n_chunk) of a prefill over n_images: operand conversion (bf16 mode) + ONE projection GEMM into the
// cross cache.  capdec_prefill runs it once over everything; capdec_generate_host once per host-copy chunk.
static int prefill_chunk(capdec_handle h, cudaStream_t s, const float* memory_chunk, int img0, int n_chunk, int n_images,
                         int mem_len) {
  const int M = n_chunk * mem_len, N = h->cfg.n_layer * 2 * kD;
  const size_t off = (size_t)img0 * mem_len * kD;  // element offset of the chunk inside the staged operand
  if (h->bf16) {
    const size_t n4 = (size_t)M * kD / 4;
    set_tag(h, "prefill_convert", (double)M * kD * 6);
    ProfScope ps(h, s);
    convert_kernel<bf16><<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(memory_chunk, (bf16*)h->mem_b + off, n4);
    CUDA_TRY(h, cudaGetLastError());
    h->launches_this_enqueue++;
  }
  if (h->fast && !h->no_ws && !h->no_ws_prefill) {
    // persistent weight-stationary GEMM (gemm_ws.cuh): 128-row tiles, per-image bulk stores straight into the cross
    // cache layout (WsCrossKV)
    WsCrossKV epi{(bf16*)h->cross_kv, h->b_ckv, mem_len, n_images, img0, M};
    CUtensorMap tm_x;
    TRY(make_map(h, &tm_x, (bf16*)h->mem_b + off, M, kD, 128));
    set_tag(h, "prefill_gemm_cross_kv", gemm_bytes(M, N, kD, h->esz, h->esz, h->esz));
    return launch_ws(h, s, h->m_ckv, tm_x, M, N, epi, nullptr);
  }
  EpiCrossKV<float> e32{(float*)h->cross_kv, h->b_ckv, n_images, mem_len, img0};
  EpiCrossKV<bf16> e16{(bf16*)h->cross_kv, h->b_ckv, n_images, mem_len, img0};
  CUtensorMap tm_mem;
  if (h->use_umma) TRY(make_map(h, &tm_mem, (bf16*)h->mem_b + off, M, kD, kUmmaBM));
  if (h->umma_f32) TRY(make_map(h, &tm_mem, memory_chunk, M, kD, kUmmaBM, true));  // TF32 route: straight from the fp32 memory
  set_tag(h, "prefill_gemm_cross_kv", gemm_bytes(M, N, kD, h->esz, h->esz, h->esz));
  return gemm(h, s, h->bf16 ? (const void*)((bf16*)h->mem_b + off) : (const void*)memory_chunk, &tm_mem, h->w_ckv, &h->m_ckv, M, N,
              kD, e32, e16, nullptr);
}

static int check_prefill_args(capdec_handle h, const void* memory, int n_images, int mem_len) {
  if (!h || !memory) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  if (!h->have_weights) return fail(h, CAPDEC_ERR_STATE, "prefill before capdec_set_weights");
  if (n_images < 1 || mem_len < 1) return fail(h, CAPDEC_ERR_INVALID, "n_images and mem_len must be positive");
  if (n_images > h->cfg.max_images || mem_len > h->cfg.max_mem)
    return fail(h, CAPDEC_ERR_CAPACITY, "prefill of %d images x %d exceeds capacity %d x %d", n_images, mem_len,
                h->cfg.max_images, h->cfg.max_mem);
  return 0;
}

extern "C" int capdec_prefill(capdec_handle h, const float* memory, int32_t n_images, int32_t mem_len, void* stream) {
  TRY(check_prefill_args(h, memory, n_images, mem_len));
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  h->prefilled = false;
  h->launches_this_enqueue = 0;
  TRY(prefill_chunk(h, s, memory, 0, n_images, n_images, mem_len));
  return finish_prefill(h, n_images, mem_len);
}

// ------------------------------------------------------------------ encoder hand-off
// memory = tokens * P^T + pb is only ever used as the operand of the cross K/V projection, so
//   K|V = memory * Wkv^T + bkv = tokens * (Wkv P)^T + (Wkv pb + bkv):
// the product Wkv P ([L*1024, in_dim]) is formed once per weight update in fp32, and prefill runs ONE GEMM from the
// encoder's tokens without materialising the fp32 memory tensor (vit:98) or running the 768 -> 512 projection.
extern "C" int capdec_set_memory_projection(capdec_handle h, const float* proj_w, const float* proj_b, int32_t in_dim,
                                            const float* const* cross_in_w, const float* const* cross_in_b, void* stream) {
  if (!h || !proj_w || !proj_b || !cross_in_w || !cross_in_b) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  if (in_dim < kUmmaBK || in_dim % kUmmaBK != 0 || in_dim > 4096)
    return fail(h, CAPDEC_ERR_INVALID, "projection input width %d must be a multiple of %d in [%d, 4096]", in_dim, kUmmaBK, kUmmaBK);
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  const int L = h->cfg.n_layer, N = L * 2 * kD;
  if (h->proj_dim != in_dim) {  // (buffers of an earlier width stay allocated until destroy)
    TRY(dev_alloc_bytes(h, &h->w_ckv_f, (size_t)N * in_dim * h->esz));
    TRY(dev_alloc(h, &h->b_ckv_f, (size_t)N));
    TRY(dev_alloc(h, &h->fold_tmp, (size_t)in_dim * kD + (size_t)N * in_dim));
    TRY(dev_alloc_bytes(h, &h->tok_b, (size_t)h->cfg.max_images * h->cfg.max_mem * in_dim * h->esz));
    h->proj_dim = in_dim;
  }
  float* pt = h->fold_tmp;                          // [in_dim][512] = proj_w^T
  float* prod = h->fold_tmp + (size_t)in_dim * kD;  // [L*1024][in_dim]
  transpose_pad_kernel<<<(in_dim * kD + 255) / 256, 256, 0, s>>>(proj_w, pt, kD, in_dim, kD);
  CUDA_TRY(h, cudaGetLastError());
  h->launches++;
  h->launches_this_enqueue = 0;
  for (int l = 0; l < L; ++l) {
    if (!cross_in_w[l] || !cross_in_b[l]) return fail(h, CAPDEC_ERR_INVALID, "null cross-attention weight pointer in layer %d", l);
    const float* wkv = cross_in_w[l] + (size_t)kD * kD;  // rows [512,1536) of in_proj_weight: K then V
    EpiLinear<float, false> ew{prod + (size_t)l * 2 * kD * in_dim, nullptr, in_dim};
    TRY((launch_simt<float, float, EpiLinear<float, false>>(h, s, wkv, kD, pt, kD, 2 * kD, in_dim, kD, ew, nullptr)));
    EpiLinear<float, false> eb{h->b_ckv_f + (size_t)l * 2 * kD, cross_in_b[l] + kD, 2 * kD};
    TRY((launch_simt<float, float, EpiLinear<float, false>>(h, s, proj_b, kD, wkv, kD, 1, 2 * kD, kD, eb, nullptr)));
  }
  h->launches += h->launches_this_enqueue;
  TRY(pack_matrix(h, s, prod, h->w_ckv_f, N, in_dim, N));
  if (h->use_umma || h->umma_f32) {
    const int bns[3] = {32, 64, 128};
    for (int i = 0; i < 3; ++i) TRY(make_map(h, &h->m_ckv_f.bn[i], h->w_ckv_f, N, in_dim, bns[i], h->umma_f32));
  }
  return CAPDEC_OK;
}

extern "C" int capdec_prefill_tokens(capdec_handle h, const float* tokens, int32_t n_images, int32_t tokens_per_image,
                                     int32_t first_token, int32_t mem_len, void* stream) {
  if (!h || !tokens) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  if (!h->have_weights) return fail(h, CAPDEC_ERR_STATE, "capdec_prefill_tokens before capdec_set_weights");
  if (h->proj_dim == 0) return fail(h, CAPDEC_ERR_STATE, "capdec_prefill_tokens before capdec_set_memory_projection");
  if (n_images < 1 || mem_len < 1 || first_token < 0 || first_token + mem_len > tokens_per_image)
    return fail(h, CAPDEC_ERR_INVALID, "token range [%d, %d) outside the %d tokens of an image", first_token, first_token + mem_len,
                tokens_per_image);
  if (n_images > h->cfg.max_images || mem_len > h->cfg.max_mem)
    return fail(h, CAPDEC_ERR_CAPACITY, "prefill of %d images x %d exceeds capacity %d x %d", n_images, mem_len,
                h->cfg.max_images, h->cfg.max_mem);
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  h->prefilled = false;
  h->launches_this_enqueue = 0;
  const int E = h->proj_dim, M = n_images * mem_len, N = h->cfg.n_layer * 2 * kD;
  {
    const size_t n4 = (size_t)M * E / 4;
    set_tag(h, "prefill_gather_tokens", (double)M * E * (4 + h->esz));
    ProfScope ps(h, s);
    if (h->bf16) gather_tokens_kernel<bf16><<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(tokens, (bf16*)h->tok_b, n_images, tokens_per_image, first_token, mem_len, E / 4);
    else gather_tokens_kernel<float><<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(tokens, (float*)h->tok_b, n_images, tokens_per_image, first_token, mem_len, E / 4);
    CUDA_TRY(h, cudaGetLastError());
    h->launches_this_enqueue++;
  }
  EpiCrossKV<float> e32{(float*)h->cross_kv, h->b_ckv_f, n_images, mem_len};
  EpiCrossKV<bf16> e16{(bf16*)h->cross_kv, h->b_ckv_f, n_images, mem_len};
  CUtensorMap tm_tok;
  if (h->use_umma || h->umma_f32) TRY(make_map(h, &tm_tok, h->tok_b, M, E, kUmmaBM, h->umma_f32));
  set_tag(h, "prefill_gemm_cross_kv", gemm_bytes(M, N, E, h->esz, h->esz, h->esz));
  TRY(gemm(h, s, h->tok_b, &tm_tok, h->w_ckv_f, &h->m_ckv_f, M, N, E, e32, e16, nullptr));
  return finish_prefill(h, n_images, mem_len);
}

// ------------------------------------------------------------------ one decode step
// Vocabulary projection logits[M, kVPad] = x[M,512] (fp32 residual stream) * fc_w^T + b.  fp32 engine: CUDA cores (strict
// parity).  bf16 / tf32 engines with `tx` (fp32 TMA view of x): tcgen05 kind::tf32 - the operands keep 10 mantissa bits,
// which moves a logit by ~5e-4 (K = 512, |w| ~ 0.05), far inside either mode's tolerance.
static int vocab_logits(capdec_handle h, cudaStream_t s, const float* x, const CUtensorMap* tx, float* logits, int M, const DecState* st) {
  EpiLinear<float, false> ev{logits, h->fc_b, kVPad};
  set_tag(h, "gemm_vocab", gemm_bytes(M, kVPad, kD, 4, 4, 4));
  if (tx != nullptr && (h->fast || h->umma_f32) && !h->no_tf32_head)
    return launch_umma<EpiLinear<float, false>, float>(h, s, *tx, h->m_fc, M, kVPad, kD, ev, st);
  return launch_simt<float, float, EpiLinear<float, false>>(h, s, x, kD, h->fc_w, kD, M, kVPad, kD, ev, st);
}

template <typename Elem>
static int launch_self_attn(capdec_handle h, cudaStream_t s, int l, int M) {
  const Elem* kc = (const Elem*)h->self_k + (size_t)l * h->self_layer_elems;
  const Elem* vc = (const Elem*)h->self_v + (size_t)l * h->self_layer_elems;
  const int blocks = (M * kH + kAttnWarps - 1) / kAttnWarps;
  ProfScope ps(h, s);
  if (h->beam_mode) {
    if (h->t_pad == 64)
      self_attn_kernel<Elem, 16, true><<<blocks, kAttnWarps * 32, 0, s>>>((const Elem*)h->q, kc, vc, (Elem*)h->a, M, h->t_pad, h->st, h->beam.anc);
    else
      self_attn_kernel<Elem, 32, true><<<blocks, kAttnWarps * 32, 0, s>>>((const Elem*)h->q, kc, vc, (Elem*)h->a, M, h->t_pad, h->st, h->beam.anc);
  } else if (h->t_pad == 64) {
    self_attn_kernel<Elem, 16, false><<<blocks, kAttnWarps * 32, 0, s>>>((const Elem*)h->q, kc, vc, (Elem*)h->a, M, h->t_pad, h->st, nullptr);
  } else {
    self_attn_kernel<Elem, 32, false><<<blocks, kAttnWarps * 32, 0, s>>>((const Elem*)h->q, kc, vc, (Elem*)h->a, M, h->t_pad, h->st, nullptr);
  }
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  return 0;
}

template <typename Elem, int GT>
static int launch_cross_gt(capdec_handle h, cudaStream_t s, const Elem* ck, const Elem* cv, int n_images, int group, int S) {
  const int warps = n_images * kH * (group / GT);
  ProfScope ps(h, s);
  cross_attn_kernel<Elem, GT><<<(warps + kAttnWarps - 1) / kAttnWarps, kAttnWarps * 32, 0, s>>>(
      (const Elem*)h->q, ck, cv, (Elem*)h->a, n_images, group, S, h->st);
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  return 0;
}

template <typename Elem>
static int launch_cross_attn(capdec_handle h, cudaStream_t s, int l, int n_images, int group, int S) {
  const size_t tile = (size_t)n_images * kH * S * kHD;  // one (layer, kv) block
  const Elem* ck = (const Elem*)h->cross_kv + (size_t)(2 * l) * tile;
  const Elem* cv = ck + tile;
  if (group % 6 == 0) return launch_cross_gt<Elem, 6>(h, s, ck, cv, n_images, group, S);
  if (group % 4 == 0) return launch_cross_gt<Elem, 4>(h, s, ck, cv, n_images, group, S);
  if (group % 2 == 0) return launch_cross_gt<Elem, 2>(h, s, ck, cv, n_images, group, S);
  return launch_cross_gt<Elem, 1>(h, s, ck, cv, n_images, group, S);
}

template <typename Elem>
static int launch_add_ln(capdec_handle h, cudaStream_t s, int l, int which, int M) {
  ProfScope ps(h, s);
  add_ln_kernel<Elem><<<(M + 3) / 4, 128, 0, s>>>(h->x, h->t, h->layer[l].ln_w[which], h->layer[l].ln_b[which],
                                                   (Elem*)h->xb, M, h->st);
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  return 0;
}

template <typename Elem>
static int enqueue_step_t(capdec_handle h, cudaStream_t s, int n_images, int S, int group) {
  const int M = n_images * group, L = h->cfg.n_layer, F = h->cfg.d_ff;
  const size_t es = h->esz;
  const DecState* st = h->st;
  const void* xin = h->bf16 ? h->xb : (void*)h->x;  // GEMM A operand view of the residual stream
  const double ln_bytes = (double)M * kD * (4 + 4 + 4 + (h->bf16 ? 2 : 0)) + 2.0 * kD * 4;
  {
    set_tag(h, "embed_pe", (double)M * kD * (4 + 4 + (h->bf16 ? 2 : 0)));
    ProfScope ps(h, s);
    embed_pe_kernel<Elem><<<(M * (kD / 4) + 255) / 256, 256, 0, s>>>(h->st, h->cp, h->emb, h->pe, h->x, (Elem*)h->xb, M);
  }
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  for (int l = 0; l < L; ++l) {
    LayerParams& lp = h->layer[l];
    {  // self-attention block
      float* kc32 = (float*)h->self_k + (size_t)l * h->self_layer_elems;
      float* vc32 = (float*)h->self_v + (size_t)l * h->self_layer_elems;
      bf16* kc16 = (bf16*)h->self_k + (size_t)l * h->self_layer_elems;
      bf16* vc16 = (bf16*)h->self_v + (size_t)l * h->self_layer_elems;
      EpiQKV<float> e32{(float*)h->q, kc32, vc32, lp.b_qkv, st, h->t_pad};
      EpiQKV<bf16> e16{(bf16*)h->q, kc16, vc16, lp.b_qkv, st, h->t_pad};
      set_tag(h, "gemm_qkv", gemm_bytes(M, 3 * kD, kD, es, es, es));
      TRY(gemm(h, s, xin, &h->tm_xb, lp.w_qkv, &lp.m_qkv, M, 3 * kD, kD, e32, e16, st));
      // bytes: mean cached length over a rollout is filled in by the reporter (per-step n varies);
      // here the launch is tagged with the bytes at the CURRENT host-side step estimate
      set_tag(h, "self_attn", (double)M * kD * es * 2 + (double)M * 2 * kD * es * h->prof_step_n);
      TRY(launch_self_attn<Elem>(h, s, l, M));
      EpiLinear<float, false> eo{h->t, lp.b_so, kD};
      set_tag(h, "gemm_self_out", gemm_bytes(M, kD, kD, es, es, 4));
      TRY(gemm(h, s, h->a, &h->tm_a, lp.w_so, &lp.m_so, M, kD, kD, eo, eo, st));
      set_tag(h, "add_ln", ln_bytes);
      TRY(launch_add_ln<Elem>(h, s, l, 0, M));
    }
    {  // cross-attention block
      EpiLinear<float, false> q32{(float*)h->q, lp.b_cq, kD};
      EpiLinear<bf16, false> q16{(bf16*)h->q, lp.b_cq, kD};
      set_tag(h, "gemm_cross_q", gemm_bytes(M, kD, kD, es, es, es));
      TRY(gemm(h, s, xin, &h->tm_xb, lp.w_cq, &lp.m_cq, M, kD, kD, q32, q16, st));
      set_tag(h, "cross_attn", (double)n_images * S * 2 * kD * es + (double)M * kD * es * 2);
      TRY(launch_cross_attn<Elem>(h, s, l, n_images, group, S));
      EpiLinear<float, false> eo{h->t, lp.b_co, kD};
      set_tag(h, "gemm_cross_out", gemm_bytes(M, kD, kD, es, es, 4));
      TRY(gemm(h, s, h->a, &h->tm_a, lp.w_co, &lp.m_co, M, kD, kD, eo, eo, st));
      set_tag(h, "add_ln", ln_bytes);
      TRY(launch_add_ln<Elem>(h, s, l, 1, M));
    }
    {  // feed-forward block
      EpiLinear<float, true> h32{(float*)h->hbuf, lp.b_1, F};
      EpiLinear<bf16, true> h16{(bf16*)h->hbuf, lp.b_1, F};
      set_tag(h, "gemm_ffn1", gemm_bytes(M, F, kD, es, es, es));
      TRY(gemm(h, s, xin, &h->tm_xb, lp.w_1, &lp.m_1, M, F, kD, h32, h16, st));
      EpiLinear<float, false> eo{h->t, lp.b_2, kD};
      set_tag(h, "gemm_ffn2", gemm_bytes(M, kD, F, es, es, 4));
      TRY(gemm(h, s, h->hbuf, &h->tm_h, lp.w_2, &lp.m_2, M, kD, F, eo, eo, st));
      set_tag(h, "add_ln", ln_bytes);
      TRY(launch_add_ln<Elem>(h, s, l, 2, M));
    }
  }
  // vocabulary projection in fp32 on the fp32 residual stream, then pick/sample
  TRY(vocab_logits(h, s, h->x, h->umma_f32 ? &h->tm_xb : nullptr, h->logits, M, st));  // (tf32 mode: tm_xb views x)
  if (h->beam_mode) {
    set_tag(h, "beam_head", (double)M * (kVPad * 4 + 8 + 4 + 2 * h->t_pad * 4));
    ProfScope ps(h, s);
    beam_head_kernel<<<n_images, group * 32, 0, s>>>(h->st, h->cp, h->beam, h->logits, n_images, group, h->cfg.vocab, h->t_pad, h->beam_grid_rule);
  } else {
    set_tag(h, "head", (double)M * (kVPad * 4 + 8 + 4));
    ProfScope ps(h, s);
    head_kernel<<<(M + 3) / 4, 128, 0, s>>>(h->st, h->cp, h->logits, h->fin, M, h->cfg.vocab);
  }
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  return 0;
}

// ---- v2 step (bf16 + tcgen05)
static int fast_self_attn(capdec_handle h, cudaStream_t s, int l, int M) {
  const DecState* st = h->st;
  bf16* kc = (bf16*)h->self_k + (size_t)l * h->self_layer_elems;
  bf16* vc = (bf16*)h->self_v + (size_t)l * h->self_layer_elems;
  set_tag(h, "self_attn", (double)M * kD * 2 * 2 + (double)M * 2 * kD * 2 * h->prof_step_n);
  const int blocks = (M * kH + kAttnWarps - 1) / kAttnWarps;
  cudaError_t e;
  {
    ProfScope ps(h, s);
    if (h->beam_mode)
      e = launch_ex(self_attn_stream_kernel<true>, dim3(blocks), dim3(kAttnWarps * 32), 0, s, 1, h->use_pdl, (const bf16*)h->q,
                    (const bf16*)kc, (const bf16*)vc, (bf16*)h->a, M, h->t_pad, st, (const int*)h->beam.anc);
    else if (h->self_attn_its == 4)
      e = launch_ex(self_attn_stream_kernel<false, 4>, dim3(blocks), dim3(kAttnWarps * 32), 0, s, 1, h->use_pdl, (const bf16*)h->q,
                    (const bf16*)kc, (const bf16*)vc, (bf16*)h->a, M, h->t_pad, st, (const int*)nullptr);
    else
      e = launch_ex(self_attn_stream_kernel<false>, dim3(blocks), dim3(kAttnWarps * 32), 0, s, 1, h->use_pdl, (const bf16*)h->q,
                    (const bf16*)kc, (const bf16*)vc, (bf16*)h->a, M, h->t_pad, st, (const int*)nullptr);
  }
  CUDA_TRY(h, e);
  h->launches_this_enqueue++;
  return 0;
}

static int fast_cross_attn(capdec_handle h, cudaStream_t s, int l, int n_images, int group, int S) {
  const DecState* st = h->st;
  const int M = n_images * group;
  set_tag(h, "cross_attn", (double)n_images * S * 2 * kD * 2 + (double)M * kD * 2 * 2);
  const int KT = ca_key_rows(S) / 16;
  const int per_warp = (KT + 3) / 4;
  const int ca_compact = (!h->no_ca_compact && ca_compact_ok(S, group)) ? 1 : 0;
  const size_t ca_bytes = ca_compact ? ca_smem_bytes_compact(S, h->dyn_smem_aligned) : ca_smem_bytes(S, group);
  cudaError_t e;
  {
    ProfScope ps(h, s);
    // one query per (image, head) and a short memory: streaming warps, all resident at once (B200: cfg2 335 -> 321
    // us/step, cfg5 592 -> 532; at S = 196 the tensor-core kernel is level, 561 vs 565)
    if (group == 1 && S <= 128 && !h->no_cross_stream)
      e = launch_ex(cross_attn_stream_kernel, dim3((n_images * kH + kAttnWarps - 1) / kAttnWarps), dim3(kAttnWarps * 32), 0,
                    s, 1, h->use_pdl, (const bf16*)h->q, (const bf16*)h->cross_kv, (bf16*)h->a, l, n_images, S, st);
    else if (per_warp <= 1)
      e = launch_ex(cross_attn_tma_kernel<1>, dim3(n_images * kH), dim3(kCaThreads), ca_bytes, s, 1, h->use_pdl,
                    h->tm_ckv, (const bf16*)h->q, (bf16*)h->a, l, n_images, group, S, st, ca_compact);
    else if (per_warp <= 2)
      e = launch_ex(cross_attn_tma_kernel<2>, dim3(n_images * kH), dim3(kCaThreads), ca_bytes, s, 1, h->use_pdl,
                    h->tm_ckv, (const bf16*)h->q, (bf16*)h->a, l, n_images, group, S, st, ca_compact);
    else
      e = launch_ex(cross_attn_tma_kernel<4>, dim3(n_images * kH), dim3(kCaThreads), ca_bytes, s, 1, h->use_pdl,
                    h->tm_ckv, (const bf16*)h->q, (bf16*)h->a, l, n_images, group, S, st, ca_compact);
  }
  CUDA_TRY(h, e);
  h->launches_this_enqueue++;
  return 0;
}

// vocabulary projection + pick + next step's embedding (plain rollouts), or logits + per-image top-k + embedding (beam search)
static int fast_head(capdec_handle h, cudaStream_t s, int n_images, int group) {
  const int M = n_images * group;
  const DecState* st = h->st;
  cudaError_t e;
  if (h->beam_mode) {
    // beam search: logits to memory (fp32, as in the v1 step), per-image top-k head, then the next step's embedding
    TRY(vocab_logits(h, s, h->x, &h->tm_x32, h->logits, M, st));
    {
      set_tag(h, "beam_head", (double)M * (kVPad * 4 + 8 + 4 + 2 * h->t_pad * 4));
      ProfScope ps(h, s);
      beam_head_kernel<<<n_images, group * 32, 0, s>>>(h->st, h->cp, h->beam, h->logits, n_images, group, h->cfg.vocab, h->t_pad, h->beam_grid_rule);
    }
    CUDA_TRY(h, cudaGetLastError());
    h->launches_this_enqueue++;
    {
      set_tag(h, "embed_pe", (double)M * kD * (4 + 4 + 2));
      ProfScope ps(h, s);
      embed_pe_kernel<bf16><<<(M * (kD / 4) + 255) / 256, 256, 0, s>>>(h->st, h->cp, h->emb, h->pe, h->x, (bf16*)h->xb, M);
    }
    CUDA_TRY(h, cudaGetLastError());
    h->launches_this_enqueue++;
    return 0;
  }
  if (M >= h->tf32_head_min_m && !h->no_tf32_head) {
    // large batches: logits on tcgen05 kind::tf32, then the v1 head (pick / sample / bookkeeping + the next step's
    // embedding) - two short launches instead of the fused CUDA-core head, whose 0.4 GFLOP of fp32 FMAs take ~33 us
    TRY(vocab_logits(h, s, h->x, &h->tm_x32, h->logits, M, st));
    {
      set_tag(h, "head", (double)M * (kVPad * 4 + 8 + 4) + (double)M * kD * (4 + 4 + 2));
      ProfScope ps(h, s);
      head_kernel<<<(M + 3) / 4, 128, 0, s>>>(h->st, h->cp, h->logits, h->fin, M, h->cfg.vocab, h->emb, h->pe, h->x, (bf16*)h->xb);
    }
    CUDA_TRY(h, cudaGetLastError());
    h->launches_this_enqueue++;
    return 0;
  }
  {
    set_tag(h, "vocab_head", (double)kD * kVPad * 4 + (double)M * (kD * 4 + 8 + 4) + (double)M * kD * (4 + 4 + 2));
    ProfScope ps(h, s);
    // rows per warp: the smallest tile that still covers the batch in one wave of CTAs (148 SMs, one CTA each);
    // env CAPDEC_HEAD_RW overrides (1, 2, 4 or 6)
    int rw = M <= 4 * 148 ? 1 : (M <= 8 * 148 ? 2 : (M <= 16 * 148 ? 4 : 6));
    if (h->force_head_rw > 0) rw = h->force_head_rw;
    auto launch_head = [&](auto kern, int rows, int bufs) {
      return launch_ex(kern, dim3((M + rows - 1) / rows), dim3(kVhWarps * 32), vh_smem_bytes(rows, bufs), s, 1, h->use_pdl, h->st,
                       (const CallParams*)h->cp, h->x, (const float*)h->fc_wT, (const float*)h->fc_b, h->fin, M, h->cfg.vocab,
                       (const float*)h->emb, (const float*)h->pe, (bf16*)h->xb);
    };
    if (rw == 1) e = launch_head(vocab_head_kernel<1, 6>, 4, 6);
    else if (rw == 2) e = launch_head(vocab_head_kernel<2, 6>, 8, 6);
    else if (rw == 4) e = launch_head(vocab_head_kernel<4, 6>, 16, 6);
    else e = launch_head(vocab_head_kernel<6, 5>, 24, 5);
  }
  CUDA_TRY(h, e);
  h->launches_this_enqueue++;
  return 0;
}

// The step as projection chains: in_proj(0), then per layer self-attention | out_proj+LN1+cross-q | cross-attention |
// out_proj+LN2+FFN+LN3[+in_proj of the next layer], then the head: 2 + 4 L kernels instead of 1 + 9 L.
template <int CS>
static int enqueue_step_chain(capdec_handle h, cudaStream_t s, int n_images, int S, int group) {
  using SH = ChainShape<CS>;
  const int M = n_images * group, L = h->cfg.n_layer;
  const long long rank_stride = (long long)L * SH::UL * kChUnit;
  auto base = [&](int l, int kind) { return (const uint8_t*)h->chain_w + chain_unit_offset<CS>(l, kind) * kChUnit; };
  auto caches = [&](ChainParams& P, int l) {
    P.kcache = (bf16*)h->self_k + (size_t)l * h->self_layer_elems;
    P.vcache = (bf16*)h->self_v + (size_t)l * h->self_layer_elems;
    P.b_qkv = h->layer[l].b_qkv;
  };
  ChainParams P0{};
  P0.rank_stride = rank_stride; P0.st = h->st; P0.M = M; P0.t_pad = h->t_pad;
  P0.x = h->x; P0.q = (bf16*)h->q; P0.part = h->chain_part;
  {
    ChainParams P = P0;
    P.wstream = base(0, kChainQKV);
    caches(P, 0);
    set_tag(h, "chain_in_proj", gemm_bytes(M, 3 * kD, kD, 2, 2, 2));
    TRY(launch_chain<kChainQKV>(h, s, h->tc_xb, P, M));
  }
  for (int l = 0; l < L; ++l) {
    LayerParams& lp = h->layer[l];
    TRY(fast_self_attn(h, s, l, M));
    {
      ChainParams P = P0;
      P.wstream = base(l, kChainAttnOut);
      P.b_proj = lp.b_so; P.ln_g = lp.ln_w[0]; P.ln_b = lp.ln_b[0]; P.b_cq = lp.b_cq;
      set_tag(h, "chain_attn_out", gemm_bytes(M, kD, kD, 2, 2, 0) + gemm_bytes(M, kD, kD, 2, 0, 2) + (double)M * kD * 8);
      TRY(launch_chain<kChainAttnOut>(h, s, h->tc_a, P, M));
    }
    TRY(fast_cross_attn(h, s, l, n_images, group, S));
    {
      ChainParams P = P0;
      P.wstream = base(l, kChainFFN);  // the next layer's in_proj units follow in the stream
      P.b_proj = lp.b_co; P.ln_g = lp.ln_w[1]; P.ln_b = lp.ln_b[1];
      P.b_1 = lp.b_1; P.b_2 = lp.b_2; P.ln3_g = lp.ln_w[2]; P.ln3_b = lp.ln_b[2];
      P.has_qkv = l + 1 < L ? 1 : 0;
      if (P.has_qkv) caches(P, l + 1);
      set_tag(h, "chain_ffn", gemm_bytes(M, kD, kD, 2, 2, 0) + 2.0 * kD * kChFF * 2 + (double)M * kD * 8 +
                                  (P.has_qkv ? gemm_bytes(M, 3 * kD, kD, 2, 0, 2) : 0.0));
      TRY(launch_chain<kChainFFN>(h, s, h->tc_a, P, M));
    }
  }
  return fast_head(h, s, n_images, group);
}

// One kernel per projection (9 per layer); in use when the chains are off (CAPDEC_CHAIN=0) or d_ff != 2048.
static int enqueue_step_fast(capdec_handle h, cudaStream_t s, int n_images, int S, int group) {
  if (h->chain_now) return h->chain_cs == 8 ? enqueue_step_chain<8>(h, s, n_images, S, group) : enqueue_step_chain<4>(h, s, n_images, S, group);
  const int M = n_images * group, L = h->cfg.n_layer, F = h->cfg.d_ff;
  const DecState* st = h->st;
  const SwapLNParams no_ln{};
  cudaError_t e;
  const double ln_extra = (double)M * kD * (4 + 4 + 2);  // residual read, x and xb written
  for (int l = 0; l < L; ++l) {
    LayerParams& lp = h->layer[l];
    bf16* kc = (bf16*)h->self_k + (size_t)l * h->self_layer_elems;
    bf16* vc = (bf16*)h->self_v + (size_t)l * h->self_layer_elems;
    {  // self-attention block
      SwapQKV<bf16> eq{(bf16*)h->q, kc, vc, lp.b_qkv, st, h->t_pad};
      set_tag(h, "gemm_qkv", gemm_bytes(M, 3 * kD, kD, 2, 2, 2));
      if (use_ws(h, 3 * kD, M)) TRY(launch_ws(h, s, lp.m_qkv, h->tx_ws, M, 3 * kD, WsQKV{h->ts_q, lp.ts_k, lp.ts_v, lp.b_qkv, st}, st));
      else TRY((launch_swap<false>(h, s, lp.m_qkv, h->tx_xb, M, 3 * kD, kD, eq, no_ln, st)));
      TRY(fast_self_attn(h, s, l, M));
      SwapLNParams ln{h->x, (bf16*)h->xb, lp.b_so, lp.ln_w[0], lp.ln_b[0], h->ts_x, h->ts_xb, h->ln_tma_store};
      set_tag(h, "gemm_self_out_ln", gemm_bytes(M, kD, kD, 2, 2, 0) + ln_extra);
      TRY((launch_swap<true>(h, s, lp.m_so, h->tx_a, M, kD, kD, SwapNoLN{}, ln, st)));
    }
    {  // cross-attention block
      SwapLinear<bf16, false> eq{(bf16*)h->q, lp.b_cq, kD};
      set_tag(h, "gemm_cross_q", gemm_bytes(M, kD, kD, 2, 2, 2));
      if (use_ws(h, kD, M)) TRY(launch_ws(h, s, lp.m_cq, h->tx_ws, M, kD, WsLinear<false>{h->ts_q, lp.b_cq}, st));
      else TRY((launch_swap<false>(h, s, lp.m_cq, h->tx_xb, M, kD, kD, eq, no_ln, st)));
      TRY(fast_cross_attn(h, s, l, n_images, group, S));
      SwapLNParams ln{h->x, (bf16*)h->xb, lp.b_co, lp.ln_w[1], lp.ln_b[1], h->ts_x, h->ts_xb, h->ln_tma_store};
      set_tag(h, "gemm_cross_out_ln", gemm_bytes(M, kD, kD, 2, 2, 0) + ln_extra);
      TRY((launch_swap<true>(h, s, lp.m_co, h->tx_a, M, kD, kD, SwapNoLN{}, ln, st)));
    }
    {  // feed-forward block
      SwapLinear<bf16, true> e1{(bf16*)h->hbuf, lp.b_1, F};
      set_tag(h, "gemm_ffn1", gemm_bytes(M, F, kD, 2, 2, 2));
      if (use_ws(h, F, M)) TRY(launch_ws(h, s, lp.m_1, h->tx_ws, M, F, WsLinear<true>{h->ts_h, lp.b_1}, st));
      else TRY((launch_swap<false>(h, s, lp.m_1, h->tx_xb, M, F, kD, e1, no_ln, st)));
      const int row_tiles = (M + h->bnb - 1) / h->bnb;
      if (M <= h->part_rows && 16 * row_tiles <= 148 && F % (4 * kSwapBK) == 0 && !h->no_ffn2_split) {
        // small batch: K = d_ff divided over 4 CTAs per tile (fp32 partial products), closed by a separate sum_ln launch.
        // CAPDEC_FFN2_FUSE=1: closed by the last CTA of each row tile instead (SwapPartialLN) -- one launch fewer, but
        // measured slower on B200 (cfg2 320 vs 293 us/step): 8 CTAs x 7 warps walk 256 rows five at a time behind two
        // fences and an atomic, where sum_ln's 256 warps take one row each inside a launch that is mostly floor.
        if (!h->no_ffn2_fuse) {
          SwapPartialLN ep{h->part, (long long)M * kD, h->x, (bf16*)h->xb, lp.b_2, lp.ln_w[2], lp.ln_b[2], h->tickets};
          set_tag(h, "gemm_ffn2_sumln", gemm_bytes(M, kD, F, 2, 2, 0) + 2 * 4.0 * M * kD * 4 + (double)M * kD * (4 + 4 + 2));
          TRY((launch_swap<false>(h, s, lp.m_2, h->tx_h, M, kD, F, ep, no_ln, st, 4)));
          continue;
        }
        SwapLinear<float, false> ep{h->part, nullptr, kD, (long long)M * kD};
        set_tag(h, "gemm_ffn2_part", gemm_bytes(M, kD, F, 2, 2, 0) + 4.0 * M * kD * 4);
        TRY((launch_swap<false>(h, s, lp.m_2, h->tx_h, M, kD, F, ep, no_ln, st, 4)));
        set_tag(h, "sum_ln", (double)M * kD * (4 * 4 + 4 + 4 + 2));
        ProfScope ps(h, s);
        e = launch_ex(sum_ln_kernel<4>, dim3((M + 3) / 4), dim3(128), 0, s, 1, h->use_pdl, h->x, (const float*)h->part,
                      (const float*)lp.b_2, (const float*)lp.ln_w[2], (const float*)lp.ln_b[2], (bf16*)h->xb, M, st);
        CUDA_TRY(h, e);
        h->launches_this_enqueue++;
      } else {
        SwapLNParams ln{h->x, (bf16*)h->xb, lp.b_2, lp.ln_w[2], lp.ln_b[2], h->ts_x, h->ts_xb, h->ln_tma_store};
        set_tag(h, "gemm_ffn2_ln", gemm_bytes(M, kD, F, 2, 2, 0) + ln_extra);
        TRY((launch_swap<true>(h, s, lp.m_2, h->tx_h, M, kD, F, SwapNoLN{}, ln, st)));
      }
    }
  }
  return fast_head(h, s, n_images, group);
}

static int enqueue_step(capdec_handle h, cudaStream_t s, int n_images, int S, int group) {
  h->launches_this_enqueue = 0;
  if (h->fast && !(h->beam_mode && h->beam_v1)) return enqueue_step_fast(h, s, n_images, S, group);
  return h->bf16 ? enqueue_step_t<bf16>(h, s, n_images, S, group) : enqueue_step_t<float>(h, s, n_images, S, group);
}

static int ensure_activation_maps(capdec_handle h, int M) {
  if (h->umma_f32) {  // TF32 route: the fp32 step buffers themselves are the TMA operands (x stands where xb does in bf16 mode)
    if (h->maps_for_m == M) return 0;
    TRY(make_map(h, &h->tm_xb, h->x, M, kD, kUmmaBM, true));
    TRY(make_map(h, &h->tm_a, h->a, M, kD, kUmmaBM, true));
    TRY(make_map(h, &h->tm_h, h->hbuf, M, h->cfg.d_ff, kUmmaBM, true));
    h->maps_for_m = M;
    return 0;
  }
  if (!h->use_umma || h->maps_for_m == M) return 0;
  TRY(make_map(h, &h->tm_xb, h->xb, M, kD, kUmmaBM));
  TRY(make_map(h, &h->tm_a, h->a, M, kD, kUmmaBM));
  TRY(make_map(h, &h->tm_h, h->hbuf, M, h->cfg.d_ff, kUmmaBM));
  if (h->fast) {
    // sequences per GEMM CTA: small enough that (n_out/128) x (M/bnb) CTAs cover the SMs
    h->bnb = h->force_bnb ? h->force_bnb : (M <= 768 ? 32 : (M <= 2048 ? 64 : 128));
    if (h->bnb != 32 && h->bnb != 64 && h->bnb != 128) return fail(h, CAPDEC_ERR_INVALID, "CAPDEC_BNB must be 32, 64 or 128");
    TRY(make_map(h, &h->tx_xb.tile, h->xb, M, kD, h->bnb));
    TRY(make_map(h, &h->tx_a.tile, h->a, M, kD, h->bnb));
    TRY(make_map(h, &h->tx_h.tile, h->hbuf, M, h->cfg.d_ff, h->bnb));
    TRY(make_map3(h, &h->tx_xb.whole, h->xb, M, kD, h->bnb, kSwapWholeKB));
    TRY(make_map3(h, &h->tx_a.whole, h->a, M, kD, h->bnb, kSwapWholeKB));
    TRY(make_map3(h, &h->tx_h.whole, h->hbuf, M, h->cfg.d_ff, h->bnb, kSwapWholeKB));
    // measured (B200): staging + bulk stores win for 32-row tiles (cfg2 372.6 -> 368.6 us/step), per-thread stores
    // for wider ones (cfg3 974 -> 958, cfg5 627 -> 620): the 96 KB tile takes ~3.3k cycles to drain
    h->ln_tma_store = h->force_ln_tma_store >= 0 ? h->force_ln_tma_store : (h->bnb == 32 ? 1 : 0);
    TRY(make_map(h, &h->tm_x32, h->x, M, kD, kUmmaBM, true));  // fp32 view of the residual stream: the kind::tf32 vocabulary projection
    TRY(make_store_map(h, &h->ts_x, h->x, M, kD, h->bnb, true));
    TRY(make_store_map(h, &h->ts_xb, h->xb, M, kD, h->bnb, false));
    if (M >= h->ws_min_m) {  // output views of the weight-stationary GEMM's bulk stores; its X tiles are always 128 rows
      TRY(make_store_map(h, &h->ts_q, h->q, M, kD, 128, false));
      TRY(make_store_map(h, &h->ts_h, h->hbuf, M, h->cfg.d_ff, 128, false));
      TRY(make_map(h, &h->tx_ws, h->xb, M, kD, 128));
    }
    // Measured on B200 (tools/chain_ab.sh, same box): the chains win only while every cluster of a launch is
    // co-resident (8 CTAs x <= 15 clusters): cfg2 (256 rows) 269.6 -> 266.3 us/step.  Beyond that each CTA still streams
    // its whole 768 KB weight share per 32-row tile through one TMA ring (~60 B/clk) and the row tiles queue up:
    // cfg5 (1024 rows) 489 -> 505 us/step with 4-CTA clusters, cfg3 (3072 rows) 860 -> 1290.  So larger batches keep one
    // kernel per projection; CAPDEC_CHAIN=2 forces the chains for any batch (tests, A/B), CAPDEC_CHAIN=0 turns them off.
    h->chain_now = h->chain && (h->chain_always || (M + 31) / 32 <= h->chain_clusters[0]);
    if (h->chain_now) {
      // 32 rows per cluster while all clusters of a launch are co-resident, else 64 (half the weight traffic per row)
      h->chain_bnb = h->force_chain_bnb ? h->force_chain_bnb : ((M + 31) / 32 <= std::max(1, h->chain_clusters[0]) ? 32 : 64);
      TRY(make_map3(h, &h->tc_xb, h->xb, M, kD, h->chain_bnb, 8));
      TRY(make_map3(h, &h->tc_a, h->a, M, kD, h->chain_bnb, 8));
    }
  }
  h->maps_for_m = M;
  return 0;
}

// Capture one step into a graph (on a private stream: the caller's may be the legacy stream).
// `nsteps` consecutive steps in one graph: the kernels read the step index from DecState, so a step's launches are
// the same at every position, and the programmatic-launch edges continue across the step boundary.
static int get_step_graph(capdec_handle h, int n_images, int S, int group, int nsteps, StepGraph** out) {
  // beam steps are different kernels (and the stop rule is a launch argument)
  auto key = std::make_tuple(n_images, S, h->beam_mode ? -(group + 100 * h->beam_grid_rule) : group, nsteps);
  auto it = h->graphs.find(key);
  if (it != h->graphs.end()) { *out = &it->second; return 0; }
  cudaStream_t cs;
  CUDA_TRY(h, cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
  StepGraph g;
  cudaError_t e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
  int rc = 0;
  if (e == cudaSuccess) {
    int kernels = 0;
    for (int i = 0; i < nsteps && rc == 0; ++i) {
      rc = enqueue_step(h, cs, n_images, S, group);
      kernels += h->launches_this_enqueue;
    }
    h->launches_this_enqueue = kernels;
    e = cudaStreamEndCapture(cs, &g.graph);
  }
  if (rc == 0 && e == cudaSuccess) e = cudaGraphInstantiate(&g.exec, g.graph, 0);
  cudaStreamDestroy(cs);
  if (rc) return rc;
  if (e != cudaSuccess) {
    if (g.graph) cudaGraphDestroy(g.graph);
    return fail(h, CAPDEC_ERR_CUDA, "step graph capture failed: %s", cudaGetErrorString(e));
  }
  g.kernels = h->launches_this_enqueue;
  auto ins = h->graphs.emplace(key, g);
  *out = &ins.first->second;
  return 0;
}

// ------------------------------------------------------------------ decode
static int validate_decode(capdec_handle h, const capdec_decode_args* a) {
  if (!h->prefilled) return fail(h, CAPDEC_ERR_STATE, "capdec_decode before capdec_prefill");
  if (a->group < 1 || a->n_sample < 0 || a->n_sample > a->group) return fail(h, CAPDEC_ERR_INVALID, "bad group/n_sample (%d/%d)", a->group, a->n_sample);
  if (a->group > h->cfg.max_group) return fail(h, CAPDEC_ERR_CAPACITY, "group %d exceeds capacity %d", a->group, h->cfg.max_group);
  if (a->max_len < 2 || a->max_len > h->cfg.max_len) return fail(h, CAPDEC_ERR_CAPACITY, "max_len %d outside [2,%d]", a->max_len, h->cfg.max_len);
  if (a->max_len - 1 > h->cfg.pe_rows)  // the reference fails at vit:32 (prefix longer than the pe table)
    return fail(h, CAPDEC_ERR_INVALID, "max_len %d needs %d positional rows, table has %d", a->max_len, a->max_len - 1, h->cfg.pe_rows);
  const int V = h->cfg.vocab;
  if (a->start_token < 0 || a->start_token >= V || a->end_token < 0 || a->end_token >= V || a->pad_token < 0 || a->pad_token >= V)
    return fail(h, CAPDEC_ERR_INVALID, "start/end/pad token outside the vocabulary");
  if (!a->ids_out) return fail(h, CAPDEC_ERR_INVALID, "ids_out is null");
  return 0;
}

extern "C" int capdec_decode(capdec_handle h, const capdec_decode_args* a, void* stream) {
  if (!h || !a) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  TRY(validate_decode(h, a));
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  const int n_images = h->n_images, S = h->mem_len, group = a->group, M = n_images * group;
  TRY(ensure_activation_maps(h, M));

  CallParams cp;
  cp.ids = (long long*)a->ids_out;
  cp.logp = a->logp_out;
  cp.logits_dump = a->logits_out;
  cp.seed = a->seed;
  cp.seq_id_base = a->seq_id_base;
  cp.max_len = a->max_len;
  cp.group = group;
  cp.n_sample = a->n_sample;
  cp.start_token = a->start_token;
  cp.end_token = a->end_token;
  cp.pad_token = a->pad_token;
  cp.no_early_stop = (a->flags & CAPDEC_DECODE_NO_EARLY_STOP) ? 1 : 0;
  cp.vocab = h->cfg.vocab;
  cp.forced = (a->flags & CAPDEC_DECODE_TEACHER_FORCED) ? 1 : 0;
  cp.n_sample_rows = n_images * a->n_sample;
  cp.n_greedy_rows = M - cp.n_sample_rows;
  // by value through a one-thread kernel, not a host -> device copy: a copy would queue on the H2D copy engine behind a
  // capdec_prefetch_host transfer in flight (205 MB at cfg3) and hold the whole rollout back until it has landed
  set_call_params_kernel<<<1, 1, 0, s>>>(h->cp, cp);
  CUDA_TRY(h, cudaGetLastError());
  init_rollout_kernel<<<std::min((M * a->max_len + 255) / 256, 1024), 256, 0, s>>>(h->st, h->cp, h->fin, M);
  CUDA_TRY(h, cudaGetLastError());
  h->launches++;

  const int T = a->max_len - 1;
  if (h->fast) {
    // input of step 0; later steps get theirs from the head kernel of the step before
    set_tag(h, "embed_pe", (double)M * kD * (4 + 4 + 2));
    ProfScope ps(h, s);
    embed_pe_kernel<bf16><<<(M * (kD / 4) + 255) / 256, 256, 0, s>>>(h->st, h->cp, h->emb, h->pe, h->x, (bf16*)h->xb, M);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
  }
  if (h->use_graph && !h->profiling) {
    StepGraph* g = nullptr;
    const int per = std::max(1, std::min(h->graph_steps, T));
    if (T / per > 0) {
      TRY(get_step_graph(h, n_images, S, group, per, &g));
      for (int t = 0; t + per <= T; t += per) CUDA_TRY(h, cudaGraphLaunch(g->exec, s));
      h->launches += (long long)g->kernels * (T / per);
    }
    if (T % per) {
      TRY(get_step_graph(h, n_images, S, group, T % per, &g));
      CUDA_TRY(h, cudaGraphLaunch(g->exec, s));
      h->launches += g->kernels;
    }
  } else {
    for (int t = 0; t < T; ++t) {
      h->prof_step_n = t + 1;
      TRY(enqueue_step(h, s, n_images, S, group));
      h->launches += h->launches_this_enqueue;
    }
  }
  if (a->steps_out) {
    int steps = 0;
    CUDA_TRY(h, cudaMemcpyAsync(&steps, &h->st->steps_executed, sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    *a->steps_out = steps;
  }
  return CAPDEC_OK;
}

// ------------------------------------------------------------------ full-sequence teacher-forced pass
// The step kernels' activation buffers and tensor maps live in the handle; the full pass runs the same GEMM launchers over
// (sequences x positions) rows, so it swaps its own workspace in for the duration of the call and puts the step state back.
namespace {
struct ActState {
  float *x, *logits;
  void *xb, *q, *a, *hbuf;
  int maps_for_m, bnb, ln_tma_store;
  bool chain, chain_now;
  CUtensorMap tm_xb, tm_a, tm_h, ts_q, ts_h, ts_x, ts_xb, tx_ws, tc_xb, tc_a;
  ActMaps tx_xb, tx_a, tx_h;
};
void save_act(capdec_handle h, ActState* o) {
  o->x = h->x; o->logits = h->logits; o->xb = h->xb; o->q = h->q; o->a = h->a; o->hbuf = h->hbuf;
  o->maps_for_m = h->maps_for_m; o->bnb = h->bnb; o->ln_tma_store = h->ln_tma_store; o->chain = h->chain; o->chain_now = h->chain_now;
  o->tm_xb = h->tm_xb; o->tm_a = h->tm_a; o->tm_h = h->tm_h; o->ts_q = h->ts_q; o->ts_h = h->ts_h; o->ts_x = h->ts_x;
  o->ts_xb = h->ts_xb; o->tx_ws = h->tx_ws; o->tc_xb = h->tc_xb; o->tc_a = h->tc_a; o->tx_xb = h->tx_xb; o->tx_a = h->tx_a; o->tx_h = h->tx_h;
}
void restore_act(capdec_handle h, const ActState& o) {
  h->x = o.x; h->logits = o.logits; h->xb = o.xb; h->q = o.q; h->a = o.a; h->hbuf = o.hbuf;
  h->maps_for_m = o.maps_for_m; h->bnb = o.bnb; h->ln_tma_store = o.ln_tma_store; h->chain = o.chain; h->chain_now = o.chain_now;
  h->tm_xb = o.tm_xb; h->tm_a = o.tm_a; h->tm_h = o.tm_h; h->ts_q = o.ts_q; h->ts_h = o.ts_h; h->ts_x = o.ts_x;
  h->ts_xb = o.ts_xb; h->tx_ws = o.tx_ws; h->tc_xb = o.tc_xb; h->tc_a = o.tc_a; h->tx_xb = o.tx_xb; h->tx_a = o.tx_a; h->tx_h = o.tx_h;
}
}  // namespace

static int ensure_full_workspace(capdec_handle h) {
  if (h->full_x) return 0;
  size_t cap = 65536;  // rows (sequence positions) per chunk of images: 9.7 KB of workspace each
  if (const char* e = getenv("CAPDEC_FULL_ROWS")) cap = (size_t)std::max(256, atoi(e));
  TRY(dev_alloc(h, &h->full_x, cap * kD));
  TRY(dev_alloc(h, &h->full_logits, cap * kVPad));
  TRY(dev_alloc_bytes(h, &h->full_xb, cap * kD * 2));
  TRY(dev_alloc_bytes(h, &h->full_q, cap * kD * 2));
  TRY(dev_alloc_bytes(h, &h->full_a, cap * kD * 2));
  TRY(dev_alloc_bytes(h, &h->full_h, cap * (size_t)h->cfg.d_ff * 2));
  h->full_rows_cap = cap;
  return 0;
}

// env CAPDEC_FULL_SYNC=1 (debugging aid): synchronise after every stage of the full pass and name the one that failed
static int full_sync(capdec_handle h, cudaStream_t s, const char* what, int l) {
  static const bool on = getenv("CAPDEC_FULL_SYNC") != nullptr && atoi(getenv("CAPDEC_FULL_SYNC")) != 0;
  if (!on) return 0;
  const cudaError_t e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return fail(h, CAPDEC_ERR_CUDA, "full pass: %s (layer %d) failed: %s", what, l, cudaGetErrorString(e));
  return 0;
}

static int launch_dense_attn(capdec_handle h, cudaStream_t s, const DenseAttnArgs& A, int units, const char* tag, double bytes) {
  set_tag(h, tag, bytes);
  ProfScope ps(h, s);
  const dim3 grid(units, kH), block(kAdWarps * 32);
  if (A.nk <= 64) dense_attn_kernel<8><<<grid, block, dense_attn_smem(8), s>>>(A);
  else if (A.nk <= 128) dense_attn_kernel<16><<<grid, block, dense_attn_smem(16), s>>>(A);
  else if (A.nk <= 208) dense_attn_kernel<26><<<grid, block, dense_attn_smem(26), s>>>(A);
  else dense_attn_kernel<32><<<grid, block, dense_attn_smem(32), s>>>(A);
  CUDA_TRY(h, cudaGetLastError());
  h->launches_this_enqueue++;
  return 0;
}

// One chunk of images [img0, img0 + nc): every decoder layer over Mc = nc * group * T rows, then the vocabulary head.
static int full_pass_chunk(capdec_handle h, cudaStream_t s, const capdec_forward_args* a, int img0, int nc) {
  const int G = a->group, T = a->max_len - 1, S = h->mem_len, n_images = h->n_images, L = h->cfg.n_layer, F = h->cfg.d_ff;
  const int seq0 = img0 * G, n_seq = nc * G, Mc = n_seq * T;
  h->maps_for_m = -1;
  TRY(ensure_activation_maps(h, Mc));
  // the weight-stationary GEMM's views, for any Mc (the step path builds them from 512 rows up)
  TRY(make_store_map(h, &h->ts_q, h->q, Mc, kD, 128, false));
  TRY(make_store_map(h, &h->ts_h, h->hbuf, Mc, F, 128, false));
  TRY(make_map(h, &h->tx_ws, h->xb, Mc, kD, 128));
  const SwapLNParams no_ln{};
  (void)no_ln;
  {
    set_tag(h, "full_embed", (double)Mc * kD * (4 + 4 + 4 + 2));
    ProfScope ps(h, s);
    const size_t n4 = (size_t)Mc * (kD / 4);
    embed_full_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>((const long long*)a->ids, a->max_len, h->cfg.vocab, seq0, n_seq, T,
                                                                 h->emb, h->pe, h->x, (bf16*)h->xb);
    CUDA_TRY(h, cudaGetLastError());
    h->launches_this_enqueue++;
  }
  const double ln_extra = (double)Mc * kD * (4 + 4 + 2);
  const size_t plane = (size_t)n_images * kH * S * kHD;
  for (int l = 0; l < L; ++l) {
    LayerParams& lp = h->layer[l];
    bf16* kc = (bf16*)h->self_k + (size_t)l * h->self_layer_elems;
    bf16* vc = (bf16*)h->self_v + (size_t)l * h->self_layer_elems;
    {  // self-attention block
      WsQKVFull eq{};
      eq.ts_q = h->ts_q; eq.kc = kc; eq.vc = vc; eq.bias = lp.b_qkv; eq.T = T; eq.t_pad = h->t_pad; eq.seq0 = seq0; eq.rows = Mc;
      set_tag(h, "full_gemm_qkv", gemm_bytes(Mc, 3 * kD, kD, 2, 2, 2));
      TRY(launch_ws(h, s, lp.m_qkv, h->tx_ws, Mc, 3 * kD, eq, nullptr));
      TRY(full_sync(h, s, "qkv", l));
      DenseAttnArgs A{(const bf16*)h->q, (bf16*)h->a, kc, vc, (long long)h->t_pad * kHD, kD, kD, seq0, T, T, T};
      TRY(launch_dense_attn(h, s, A, n_seq, "full_self_attn", (double)Mc * kD * 2 * 4));
      TRY(full_sync(h, s, "self_attn", l));
      SwapLNParams ln{h->x, (bf16*)h->xb, lp.b_so, lp.ln_w[0], lp.ln_b[0], h->ts_x, h->ts_xb, h->ln_tma_store};
      set_tag(h, "full_gemm_self_out_ln", gemm_bytes(Mc, kD, kD, 2, 2, 0) + ln_extra);
      TRY((launch_swap<true>(h, s, lp.m_so, h->tx_a, Mc, kD, kD, SwapNoLN{}, ln, nullptr)));
      TRY(full_sync(h, s, "self_out_ln", l));
    }
    {  // cross-attention block
      set_tag(h, "full_gemm_cross_q", gemm_bytes(Mc, kD, kD, 2, 2, 2));
      TRY(launch_ws(h, s, lp.m_cq, h->tx_ws, Mc, kD, WsLinear<false>{h->ts_q, lp.b_cq}, nullptr));
      TRY(full_sync(h, s, "cross_q", l));
      const bf16* ck = (const bf16*)h->cross_kv + (size_t)(2 * l) * plane;
      DenseAttnArgs A{(const bf16*)h->q, (bf16*)h->a, ck, ck + plane, (long long)S * kHD, kD, kD, img0, G * T, S, 0};
      TRY(launch_dense_attn(h, s, A, nc, "full_cross_attn", (double)nc * S * 2 * kD * 2 + (double)Mc * kD * 2 * 2));
      TRY(full_sync(h, s, "cross_attn", l));
      SwapLNParams ln{h->x, (bf16*)h->xb, lp.b_co, lp.ln_w[1], lp.ln_b[1], h->ts_x, h->ts_xb, h->ln_tma_store};
      set_tag(h, "full_gemm_cross_out_ln", gemm_bytes(Mc, kD, kD, 2, 2, 0) + ln_extra);
      TRY((launch_swap<true>(h, s, lp.m_co, h->tx_a, Mc, kD, kD, SwapNoLN{}, ln, nullptr)));
      TRY(full_sync(h, s, "cross_out_ln", l));
    }
    {  // feed-forward block
      set_tag(h, "full_gemm_ffn1", gemm_bytes(Mc, F, kD, 2, 2, 2));
      TRY(launch_ws(h, s, lp.m_1, h->tx_ws, Mc, F, WsLinear<true>{h->ts_h, lp.b_1}, nullptr));
      TRY(full_sync(h, s, "ffn1", l));
      SwapLNParams ln{h->x, (bf16*)h->xb, lp.b_2, lp.ln_w[2], lp.ln_b[2], h->ts_x, h->ts_xb, h->ln_tma_store};
      set_tag(h, "full_gemm_ffn2_ln", gemm_bytes(Mc, kD, F, 2, 2, 0) + ln_extra);
      TRY((launch_swap<true>(h, s, lp.m_2, h->tx_h, Mc, kD, F, SwapNoLN{}, ln, nullptr)));
      TRY(full_sync(h, s, "ffn2_ln", l));
    }
  }
  // vocabulary head on the fp32 residual stream (as the step head), then log-softmax + gather
  CUtensorMap tm_x32;
  TRY(make_map(h, &tm_x32, h->x, Mc, kD, kUmmaBM, true));
  TRY(vocab_logits(h, s, h->x, &tm_x32, h->logits, Mc, nullptr));
  {
    set_tag(h, "full_logp", (double)Mc * (kVPad * 4 + 8 + 4) + (a->logits_out ? (double)Mc * h->cfg.vocab * 4 : 0.0));
    ProfScope ps(h, s);
    logp_full_kernel<<<(Mc + 3) / 4, 128, 0, s>>>(h->logits, (const long long*)a->ids, a->max_len, h->cfg.vocab, seq0, n_seq, T,
                                                  a->end_token, a->logp_out, a->logits_out);
    CUDA_TRY(h, cudaGetLastError());
    h->launches_this_enqueue++;
  }
  return 0;
}

extern "C" int capdec_forward_full(capdec_handle h, const capdec_forward_args* a, void* stream) {
  if (!h || !a || !a->ids) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  if (!a->logp_out && !a->logits_out) return fail(h, CAPDEC_ERR_INVALID, "neither logp_out nor logits_out given");
  if (!h->prefilled) return fail(h, CAPDEC_ERR_STATE, "capdec_forward_full before capdec_prefill");
  if (!h->fast) return fail(h, CAPDEC_ERR_STATE, "the full-sequence pass needs the bf16 tcgen05 engine (CAPDEC_PRECISION_BF16 on sm_100, v2 step)");
  if (h->cfg.d_ff % 128 != 0) return fail(h, CAPDEC_ERR_STATE, "the full-sequence pass needs d_ff %% 128 == 0 (have %d)", h->cfg.d_ff);
  if (a->group < 1 || a->group > h->cfg.max_group) return fail(h, CAPDEC_ERR_CAPACITY, "group %d outside [1,%d]", a->group, h->cfg.max_group);
  if (a->max_len < 2 || a->max_len > h->cfg.max_len) return fail(h, CAPDEC_ERR_CAPACITY, "max_len %d outside [2,%d]", a->max_len, h->cfg.max_len);
  const int T = a->max_len - 1;
  if (T > h->cfg.pe_rows) return fail(h, CAPDEC_ERR_INVALID, "max_len %d needs %d positional rows, table has %d", a->max_len, T, h->cfg.pe_rows);
  if (T > h->t_pad || T > 256) return fail(h, CAPDEC_ERR_CAPACITY, "%d positions exceed the self cache (%d) or the attention kernel (256)", T, h->t_pad);
  if (h->mem_len > 256) return fail(h, CAPDEC_ERR_CAPACITY, "memory length %d > 256", h->mem_len);
  if (a->end_token < 0 || a->end_token >= h->cfg.vocab) return fail(h, CAPDEC_ERR_INVALID, "end token outside the vocabulary");
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  TRY(ensure_full_workspace(h));
  const size_t per_image = (size_t)a->group * T;
  if (per_image > h->full_rows_cap) return fail(h, CAPDEC_ERR_CAPACITY, "group * positions = %zu exceeds the workspace (%zu rows; CAPDEC_FULL_ROWS)", per_image, h->full_rows_cap);
  const int chunk = (int)std::min<size_t>((size_t)h->n_images, h->full_rows_cap / per_image);
  ActState saved;
  save_act(h, &saved);
  h->x = h->full_x; h->logits = h->full_logits; h->xb = h->full_xb; h->q = h->full_q; h->a = h->full_a; h->hbuf = h->full_h;
  h->chain = false;
  h->launches_this_enqueue = 0;
  int rc = 0;
  for (int i0 = 0; i0 < h->n_images && rc == 0; i0 += chunk) rc = full_pass_chunk(h, s, a, i0, std::min(chunk, h->n_images - i0));
  restore_act(h, saved);
  h->launches += h->launches_this_enqueue;
  return rc;
}

// ------------------------------------------------------------------ beam search (beam.cuh)
static int ensure_beam_buffers(capdec_handle h) {
  if (h->beam.score) return 0;
  const size_t Mmax = (size_t)h->cfg.max_images * h->cfg.max_group;
  TRY(dev_alloc(h, &h->beam.score, Mmax));
  TRY(dev_alloc(h, &h->beam.comp_score, Mmax));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.live, Mmax * sizeof(int)));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.k, (size_t)h->cfg.max_images * sizeof(int)));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.anc, 2 * Mmax * h->t_pad * sizeof(int)));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.par, Mmax * h->cfg.max_len * sizeof(int)));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.comp_slot, Mmax * sizeof(int)));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.comp_pos, Mmax * sizeof(int)));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.comp_n, (size_t)h->cfg.max_images * sizeof(int)));
  TRY(dev_alloc_bytes(h, (void**)&h->beam.images_done, sizeof(int)));
  TRY(dev_alloc(h, &h->beam_tok, Mmax * h->cfg.max_len));
  return 0;
}

extern "C" int capdec_beam_search(capdec_handle h, const capdec_beam_args* a, void* stream) {
  if (!h || !a) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  if (!h->prefilled) return fail(h, CAPDEC_ERR_STATE, "capdec_beam_search before capdec_prefill");
  const int V = h->cfg.vocab;
  if (a->beam_size < 1 || a->beam_size > kBeamMax || a->beam_size > h->cfg.max_group || a->beam_size > V)
    return fail(h, CAPDEC_ERR_CAPACITY, "beam_size %d outside 1..min(%d, max_group %d)", a->beam_size, kBeamMax, h->cfg.max_group);
  if (a->max_len < 2 || a->max_len > h->cfg.max_len) return fail(h, CAPDEC_ERR_CAPACITY, "max_len %d outside 2..%d", a->max_len, h->cfg.max_len);
  if (a->start_token < 0 || a->start_token >= V || a->end_token < 0 || a->end_token >= V || a->pad_token < 0 || a->pad_token >= V)
    return fail(h, CAPDEC_ERR_INVALID, "start/end/pad token outside the vocabulary");
  if (!a->ids_out || !a->len_out) return fail(h, CAPDEC_ERR_INVALID, "ids_out / len_out is null");
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  TRY(ensure_beam_buffers(h));
  const int n_images = h->n_images, S = h->mem_len, beam = a->beam_size, M = n_images * beam;
  TRY(ensure_activation_maps(h, M));

  CallParams cp{};
  cp.ids = h->beam_tok;
  cp.max_len = a->max_len;
  cp.group = beam;
  cp.start_token = a->start_token;
  cp.end_token = a->end_token;
  cp.pad_token = a->pad_token;
  cp.n_greedy_rows = M;
  cp.vocab = V;
  set_call_params_kernel<<<1, 1, 0, s>>>(h->cp, cp);
  CUDA_TRY(h, cudaGetLastError());
  init_rollout_kernel<<<std::min((M * a->max_len + 255) / 256, 1024), 256, 0, s>>>(h->st, h->cp, h->fin, M);
  beam_init_kernel<<<std::min((2 * M * h->t_pad + 255) / 256, 1024), 256, 0, s>>>(h->beam, h->cp, n_images, beam, h->t_pad);
  CUDA_TRY(h, cudaGetLastError());
  h->launches += 2;

  struct ModeGuard {  // the step enqueue / graph cache read the mode from the handle
    capdec_handle h;
    ~ModeGuard() { h->beam_mode = false; }
  } guard{h};
  h->beam_mode = true;
  h->beam_grid_rule = (a->flags & CAPDEC_BEAM_STOP_GRID) ? 1 : 0;
  const int T = a->max_len - 1;
  if (h->fast && !h->beam_v1) {  // input of step 0 (the v2 step embeds at its end, for the step after)
    embed_pe_kernel<bf16><<<(M * (kD / 4) + 255) / 256, 256, 0, s>>>(h->st, h->cp, h->emb, h->pe, h->x, (bf16*)h->xb, M);
    CUDA_TRY(h, cudaGetLastError());
    h->launches++;
  }
  if (h->use_graph && !h->profiling) {
    StepGraph* g = nullptr;
    const int per = std::max(1, std::min(h->graph_steps, T));
    if (T / per > 0) {
      TRY(get_step_graph(h, n_images, S, beam, per, &g));
      for (int t = 0; t + per <= T; t += per) CUDA_TRY(h, cudaGraphLaunch(g->exec, s));
      h->launches += (long long)g->kernels * (T / per);
    }
    if (T % per) {
      TRY(get_step_graph(h, n_images, S, beam, T % per, &g));
      CUDA_TRY(h, cudaGraphLaunch(g->exec, s));
      h->launches += g->kernels;
    }
  } else {
    for (int t = 0; t < T; ++t) {
      h->prof_step_n = t + 1;
      TRY(enqueue_step(h, s, n_images, S, beam));
      h->launches += h->launches_this_enqueue;
    }
  }
  beam_finalize_kernel<<<(n_images + 3) / 4, 128, 0, s>>>(h->st, h->cp, h->beam, n_images, beam, (long long*)a->ids_out, a->len_out, a->score_out);
  CUDA_TRY(h, cudaGetLastError());
  h->launches++;
  if (a->steps_out) {
    int steps = 0;
    CUDA_TRY(h, cudaMemcpyAsync(&steps, &h->st->steps_executed, sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    *a->steps_out = steps;
  }
  return CAPDEC_OK;
}

// Host buffers in, host buffers out (what generate() / SCSTLoss see when their tensors live on the CPU).
// The memory is copied in up to 8 image chunks on a private copy stream; each chunk's K/V projection starts as soon
// as its copy has landed, so the host link and the prefill overlap instead of running back to back.
static int ensure_host_staging(capdec_handle h) {
  if (h->stage_mem[0]) return 0;
  const size_t Mmax = (size_t)h->cfg.max_images * h->cfg.max_group;
  for (int i = 0; i < 2; ++i) TRY(dev_alloc(h, &h->stage_mem[i], (size_t)h->cfg.max_images * h->cfg.max_mem * kD));
  TRY(dev_alloc(h, &h->stage_ids, Mmax * h->cfg.max_len));
  TRY(dev_alloc(h, &h->stage_logp, Mmax * h->cfg.max_len));
  CUDA_TRY(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
  for (auto& slot : h->copy_ev)
    for (auto& e : slot) CUDA_TRY(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  CUDA_TRY(h, cudaEventCreateWithFlags(&h->copy_start, cudaEventDisableTiming));
  return 0;
}

// The memory of one call, host -> staging slot, in image chunks on the copy stream; one event per chunk.
static int enqueue_host_copy(capdec_handle h, int slot, const float* memory_host, int n_images, int mem_len) {
  const int n_chunks = std::min<int>(kHostCopyChunks, n_images);
  const int per = (n_images + n_chunks - 1) / n_chunks;
  for (int c = 0, img0 = 0; img0 < n_images; ++c, img0 += per) {
    const int n = std::min(per, n_images - img0);
    const size_t off = (size_t)img0 * mem_len * kD;
    CUDA_TRY(h, cudaMemcpyAsync(h->stage_mem[slot] + off, memory_host + off, (size_t)n * mem_len * kD * sizeof(float),
                                cudaMemcpyHostToDevice, h->copy_stream));
    CUDA_TRY(h, cudaEventRecord(h->copy_ev[slot][c], h->copy_stream));
  }
  return 0;
}

extern "C" int capdec_prefetch_host(capdec_handle h, const float* memory_host, int32_t n_images, int32_t mem_len) {
  TRY(check_prefill_args(h, memory_host, n_images, mem_len));
  CUDA_TRY(h, cudaSetDevice(h->device));
  TRY(ensure_host_staging(h));
  // the slot the newest outstanding prefetch (or, with none, the last capdec_generate_host) did NOT use; its last
  // reader - a prefill two calls back - has completed, because capdec_generate_host returns synchronised
  const int slot = h->last_slot ^ 1;
  if (h->pf[slot].valid) return fail(h, CAPDEC_ERR_STATE, "two prefetches are already outstanding (each capdec_generate_host consumes one)");
  TRY(enqueue_host_copy(h, slot, memory_host, n_images, mem_len));
  h->pf[slot] = {memory_host, n_images, mem_len, true, ++h->pf_seq};
  h->last_slot = slot;
  return CAPDEC_OK;
}

extern "C" int capdec_generate_host(capdec_handle h, const float* memory_host, int32_t n_images, int32_t mem_len,
                                    const capdec_decode_args* ah, void* stream) {
  if (!h || !memory_host || !ah) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  if (!ah->ids_out) return fail(h, CAPDEC_ERR_INVALID, "ids_out is null");
  if (ah->logits_out) return fail(h, CAPDEC_ERR_INVALID, "logits_out is not supported by capdec_generate_host");
  if (ah->flags & CAPDEC_DECODE_TEACHER_FORCED) return fail(h, CAPDEC_ERR_INVALID, "teacher forcing needs device buffers (capdec_decode)");
  TRY(check_prefill_args(h, memory_host, n_images, mem_len));
  if (ah->group < 1 || ah->group > h->cfg.max_group || ah->max_len < 2 || ah->max_len > h->cfg.max_len)
    return fail(h, CAPDEC_ERR_CAPACITY, "group/max_len outside capacity");
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  TRY(ensure_host_staging(h));
  const size_t M = (size_t)n_images * ah->group;
  // the oldest outstanding prefetch of this very buffer, if any: its copy is already under way (or done)
  int slot = -1;
  for (int i = 0; i < 2; ++i)
    if (h->pf[i].valid && h->pf[i].host == memory_host && h->pf[i].n_images == n_images && h->pf[i].mem_len == mem_len &&
        (slot < 0 || h->pf[i].seq < h->pf[slot].seq))
      slot = i;
  if (slot >= 0) {
    h->pf[slot].valid = false;
  } else {
    // no prefetch: copy now, into the slot no outstanding prefetch holds (both held: this call matches neither, drop them)
    slot = !h->pf[0].valid ? 0 : (!h->pf[1].valid ? 1 : 0);
    if (h->pf[0].valid && h->pf[1].valid) h->pf[0].valid = h->pf[1].valid = false;
    // the copy stream starts after whatever the caller's stream still does with the staging buffer
    CUDA_TRY(h, cudaEventRecord(h->copy_start, s));
    CUDA_TRY(h, cudaStreamWaitEvent(h->copy_stream, h->copy_start, 0));
    TRY(enqueue_host_copy(h, slot, memory_host, n_images, mem_len));
    if (!h->pf[slot ^ 1].valid) h->last_slot = slot;
  }
  h->prefilled = false;
  h->launches_this_enqueue = 0;
  const int n_chunks = std::min<int>(kHostCopyChunks, n_images);
  const int per = (n_images + n_chunks - 1) / n_chunks;
  for (int c = 0, img0 = 0; img0 < n_images; ++c, img0 += per) {
    const int n = std::min(per, n_images - img0);
    const size_t off = (size_t)img0 * mem_len * kD;
    CUDA_TRY(h, cudaStreamWaitEvent(s, h->copy_ev[slot][c], 0));
    TRY(prefill_chunk(h, s, h->stage_mem[slot] + off, img0, n, n_images, mem_len));
  }
  TRY(finish_prefill(h, n_images, mem_len));
  capdec_decode_args ad = *ah;
  ad.ids_out = (int64_t*)h->stage_ids;
  ad.logp_out = ah->logp_out ? h->stage_logp : nullptr;
  int steps = 0;
  ad.steps_out = nullptr;
  TRY(capdec_decode(h, &ad, stream));
  CUDA_TRY(h, cudaMemcpyAsync(ah->ids_out, h->stage_ids, M * ah->max_len * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  if (ah->logp_out)
    CUDA_TRY(h, cudaMemcpyAsync(ah->logp_out, h->stage_logp, M * (ah->max_len - 1) * sizeof(float), cudaMemcpyDeviceToHost, s));
  CUDA_TRY(h, cudaMemcpyAsync(&steps, &h->st->steps_executed, sizeof(int), cudaMemcpyDeviceToHost, s));
  CUDA_TRY(h, cudaStreamSynchronize(s));
  if (ah->steps_out) *ah->steps_out = steps;
  return CAPDEC_OK;
}

extern "C" int capdec_sample_logits(capdec_handle h, const float* logits, int32_t n_rows, int32_t mode, uint64_t seed,
                                    uint64_t seq_id_base, int32_t step, int64_t* ids_out, float* logp_out, void* stream) {
  if (!h || !logits || !ids_out || n_rows < 1) return fail(h, CAPDEC_ERR_INVALID, "bad argument");
  CUDA_TRY(h, cudaSetDevice(h->device));
  sample_rows_kernel<<<(n_rows + 3) / 4, 128, 0, (cudaStream_t)stream>>>(logits, n_rows, h->cfg.vocab, mode, seed, seq_id_base,
                                                                         step, (long long*)ids_out, logp_out);
  CUDA_TRY(h, cudaGetLastError());
  h->launches++;
  return CAPDEC_OK;
}

// ------------------------------------------------------------------ per-kernel timing
extern "C" int capdec_profile_enable(capdec_handle h, int32_t on) {
  if (!h) return fail(h, CAPDEC_ERR_INVALID, "null handle");
  CUDA_TRY(h, cudaSetDevice(h->device));
  for (auto& r : h->prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  h->prof.clear();
  h->profiling = on != 0;
  return CAPDEC_OK;
}

// JSON: {"tag": {"launches": n, "ms": total device ms, "bytes": total algorithmic bytes}, ...}
extern "C" int capdec_profile_report(capdec_handle h, char* buf, int32_t buf_len) {
  if (!h || !buf || buf_len < 2) return fail(h, CAPDEC_ERR_INVALID, "bad argument");
  CUDA_TRY(h, cudaSetDevice(h->device));
  CUDA_TRY(h, cudaDeviceSynchronize());
  struct Acc { long long n = 0; double ms = 0, bytes = 0; };
  std::map<std::string, Acc> acc;
  for (auto& r : h->prof) {
    float ms = 0.f;
    CUDA_TRY(h, cudaEventElapsedTime(&ms, r.e0, r.e1));
    Acc& a = acc[r.tag];
    a.n++; a.ms += ms; a.bytes += r.bytes;
  }
  std::string out = "{";
  bool first = true;
  for (auto& kv : acc) {
    char line[256];
    snprintf(line, sizeof line, "%s\"%s\": {\"launches\": %lld, \"ms\": %.6f, \"bytes\": %.0f}", first ? "" : ", ",
             kv.first.c_str(), kv.second.n, kv.second.ms, kv.second.bytes);
    out += line;
    first = false;
  }
  out += "}";
  if ((int)out.size() + 1 > buf_len) return fail(h, CAPDEC_ERR_CAPACITY, "profile report needs %zu bytes", out.size() + 1);
  memcpy(buf, out.c_str(), out.size() + 1);
  return CAPDEC_OK;
}

// ------------------------------------------------------------------ diagnostics
extern "C" int capdec_debug_gemm(capdec_handle h, const float* a, const float* w, const float* bias, float* c, int32_t m,
                                 int32_t n, int32_t k, int32_t relu, void* stream) {
  if (!h || !a || !w || !c || m < 1) return fail(h, CAPDEC_ERR_INVALID, "bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  h->launches_this_enqueue = 0;
  int rc = 0;
  if (h->umma_f32 && k % kUmmaBK == 0 && n % 128 == 0) {  // TF32 route: the caller's fp32 matrices are the TMA operands
    CUtensorMap ta;
    WeightMaps wm;
    rc = make_map(h, &ta, a, m, k, kUmmaBM, true);
    if (!rc) rc = make_weight_maps(h, &wm, w, n, k, true);
    if (!rc) {
      if (relu) { EpiLinear<float, true> e{c, bias, n}; rc = launch_umma<EpiLinear<float, true>, float>(h, s, ta, wm, m, n, k, e, nullptr); }
      else { EpiLinear<float, false> e{c, bias, n}; rc = launch_umma<EpiLinear<float, false>, float>(h, s, ta, wm, m, n, k, e, nullptr); }
    }
  } else if (!h->bf16) {
    if (relu) { EpiLinear<float, true> e{c, bias, n}; rc = launch_simt<float, float, EpiLinear<float, true>>(h, s, a, k, w, k, m, n, k, e, nullptr); }
    else { EpiLinear<float, false> e{c, bias, n}; rc = launch_simt<float, float, EpiLinear<float, false>>(h, s, a, k, w, k, m, n, k, e, nullptr); }
  } else {
    bf16 *ab = nullptr, *wb = nullptr;
    if (cudaMalloc(&ab, (size_t)m * k * 2) != cudaSuccess || cudaMalloc(&wb, (size_t)n * k * 2) != cudaSuccess) {
      cudaFree(ab);  // (cudaFree(nullptr) is a no-op)
      cudaGetLastError();
      return fail(h, CAPDEC_ERR_CUDA, "debug gemm: cudaMalloc of the bf16 operands failed");
    }
    pack_rows_kernel<bf16><<<(unsigned)(((size_t)m * k + 255) / 256), 256, 0, s>>>(a, ab, m, k, m);
    pack_rows_kernel<bf16><<<(unsigned)(((size_t)n * k + 255) / 256), 256, 0, s>>>(w, wb, n, k, n);
    if (h->use_umma) {
      CUtensorMap ta;
      WeightMaps wm;
      rc = make_map(h, &ta, ab, m, k, kUmmaBM);
      if (!rc) rc = make_weight_maps(h, &wm, wb, n, k);
      if (!rc) {
        if (relu) { EpiLinear<float, true> e{c, bias, n}; rc = launch_umma(h, s, ta, wm, m, n, k, e, nullptr); }
        else { EpiLinear<float, false> e{c, bias, n}; rc = launch_umma(h, s, ta, wm, m, n, k, e, nullptr); }
      }
    } else {
      if (relu) { EpiLinear<float, true> e{c, bias, n}; rc = launch_simt<bf16, bf16, EpiLinear<float, true>>(h, s, ab, k, wb, k, m, n, k, e, nullptr); }
      else { EpiLinear<float, false> e{c, bias, n}; rc = launch_simt<bf16, bf16, EpiLinear<float, false>>(h, s, ab, k, wb, k, m, n, k, e, nullptr); }
    }
    cudaError_t e2 = cudaStreamSynchronize(s);
    cudaFree(ab);
    cudaFree(wb);
    if (!rc && e2 != cudaSuccess) rc = fail(h, CAPDEC_ERR_CUDA, "debug gemm failed: %s", cudaGetErrorString(e2));
  }
  h->launches += h->launches_this_enqueue + (h->bf16 ? 2 : 0);
  return rc;
}

extern "C" int capdec_debug_cross_kv(capdec_handle h, int32_t layer, int32_t which, float* out, void* stream) {
  if (!h || !out || layer < 0 || layer >= h->cfg.n_layer || which < 0 || which > 1) return fail(h, CAPDEC_ERR_INVALID, "bad argument");
  if (!h->prefilled) return fail(h, CAPDEC_ERR_STATE, "no prefill");
  CUDA_TRY(h, cudaSetDevice(h->device));
  const size_t tile = (size_t)h->n_images * kH * h->mem_len * kHD;
  const size_t n = (size_t)h->n_images * h->mem_len * kD;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  if (h->bf16)
    gather_cross_kernel<bf16><<<blocks, 256, 0, (cudaStream_t)stream>>>((const bf16*)h->cross_kv + (size_t)(2 * layer + which) * tile, out, h->n_images, h->mem_len);
  else
    gather_cross_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>((const float*)h->cross_kv + (size_t)(2 * layer + which) * tile, out, h->n_images, h->mem_len);
  CUDA_TRY(h, cudaGetLastError());
  h->launches++;
  return CAPDEC_OK;
}

// ------------------------------------------------------------------ CIDEr-D reward (cider.cuh)
extern "C" int capdec_cider_d(capdec_handle h, const capdec_cider_args* a, void* stream) {
  if (!h || !a || !a->hyp_ids || !a->ref_ids || !a->ref_offsets || !a->scores_out) return fail(h, CAPDEC_ERR_INVALID, "null argument");
  if (a->n_images < 1 || a->group < 1 || a->hyp_len < 1 || a->ref_len < 1 || a->n_refs < 0)
    return fail(h, CAPDEC_ERR_INVALID, "cider: sizes must be positive");
  if (a->ref_offsets[0] != 0 || a->ref_offsets[a->n_images] != a->n_refs) return fail(h, CAPDEC_ERR_INVALID, "cider: ref_offsets must run from 0 to n_refs");
  const int words = std::min<int>(a->ref_len, kCdMaxTok);
  const int per_ref = words >= 4 ? 4 * words - 6 : words * (words + 1) / 2;
  for (int i = 0; i < a->n_images; ++i) {
    const int n = a->ref_offsets[i + 1] - a->ref_offsets[i];
    if (n < 0) return fail(h, CAPDEC_ERR_INVALID, "cider: ref_offsets must not decrease");
    if ((long long)n * per_ref > kCdImgKeys)
      return fail(h, CAPDEC_ERR_CAPACITY, "cider: image %d has %d references of up to %d words: more than %d n-grams", i, n, words, kCdImgKeys);
  }
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  size_t want = 1024;
  while (want < (size_t)2 * std::max(1, a->n_refs) * per_ref) want <<= 1;
  if (want > h->cider_tab || a->n_images + 1 > h->cider_off_cap) {  // (grown on demand; earlier buffers stay until destroy)
    h->cider_tab = std::max(want, h->cider_tab);
    h->cider_off_cap = std::max(a->n_images + 1, h->cider_off_cap);
    TRY(dev_alloc(h, &h->cider_keys, h->cider_tab));
    TRY(dev_alloc(h, &h->cider_cnt, h->cider_tab));
    TRY(dev_alloc(h, &h->cider_off, (size_t)h->cider_off_cap));
  }
  CUDA_TRY(h, cudaMemsetAsync(h->cider_keys, 0xff, h->cider_tab * sizeof(unsigned long long), s));
  CUDA_TRY(h, cudaMemsetAsync(h->cider_cnt, 0, h->cider_tab * sizeof(unsigned int), s));
  CUDA_TRY(h, cudaMemcpyAsync(h->cider_off, a->ref_offsets, (size_t)(a->n_images + 1) * sizeof(int), cudaMemcpyHostToDevice, s));
  CiderArgs A{};
  A.hyp = (const long long*)a->hyp_ids; A.refs = (const long long*)a->ref_ids; A.ref_offsets = h->cider_off;
  A.n_images = a->n_images; A.group = a->group; A.hyp_len = a->hyp_len; A.ref_len = a->ref_len;
  A.start_token = a->start_token; A.end_token = a->end_token; A.pad_token = a->pad_token;
  A.tab_keys = h->cider_keys; A.tab_cnt = h->cider_cnt; A.tab_mask = (unsigned int)(h->cider_tab - 1);
  A.scores = a->scores_out;
  cider_df_kernel<<<a->n_images, kCdThreads, 0, s>>>(A);
  cider_score_kernel<<<a->n_images * a->group, kCdThreads, 0, s>>>(A);
  CUDA_TRY(h, cudaGetLastError());
  h->launches += 2;
  return CAPDEC_OK;
}

// ------------------------------------------------------------------ allocation-free projection (capdec_gemm_static)
// capdec_debug_gemm converts its fp32 operands to bf16 in scratch buffers it cudaMalloc's and frees per call.  Callers
// that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid it) get the same
// kernels with that scratch in a file-scope device array: 4 MiB, held by every GPU that loads the library until exit.
// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once
// per device with cudaGetSymbolAddress (that device current) and handed to the kernels.
constexpr size_t kStaticScratchElems = (size_t)2 << 20;  // bf16 elements: operand rows of A plus rows of W, 4 MiB in all
static __device__ __align__(256) unsigned short g_static_scratch[kStaticScratchElems];

static int static_scratch_address(capdec_handle h, bf16** out) {
  static std::mutex mu;
  static void* addr[64] = {};
  if (h->device < 0 || h->device >= 64) return fail(h, CAPDEC_ERR_INVALID, "device index %d outside the scratch table", h->device);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[h->device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);  // caller has made h->device current
    if (e != cudaSuccess || !p)
      return fail(h, CAPDEC_ERR_CUDA, "cudaGetSymbolAddress(g_static_scratch) failed on device %d: %s", h->device, cudaGetErrorString(e));
    addr[h->device] = p;
  }
  *out = static_cast<bf16*>(addr[h->device]);
  return 0;
}

extern "C" int capdec_gemm_static(capdec_handle h, const float* a, const float* w, const float* bias, float* c, int32_t m,
                                  int32_t n, int32_t k, int32_t relu, void* stream) {
  if (!h || !a || !w || !c || m < 1 || n < 1 || k < 1) return fail(h, CAPDEC_ERR_INVALID, "bad argument");
  if (h->profiling) return fail(h, CAPDEC_ERR_STATE, "capdec_gemm_static while per-launch profiling is on (the event pairs are allocations)");
  const size_t a_elems = ((size_t)m * k + 127) / 128 * 128, w_elems = (size_t)n * k;
  if (h->bf16 && a_elems + w_elems > kStaticScratchElems)
    return fail(h, CAPDEC_ERR_CAPACITY, "capdec_gemm_static: (m + n) * k = %zu operand elements exceed the static scratch of %zu (4 MiB); "
                "use capdec_debug_gemm or a smaller tile", (size_t)m * k + w_elems, kStaticScratchElems);
  cudaStream_t s = (cudaStream_t)stream;
  CUDA_TRY(h, cudaSetDevice(h->device));
  h->launches_this_enqueue = 0;
  int rc = 0;
  if (!h->bf16) {  // fp32 engine: the CUDA-core GEMM reads the operands in place, no scratch at all
    if (relu) { EpiLinear<float, true> e{c, bias, n}; rc = launch_simt<float, float, EpiLinear<float, true>>(h, s, a, k, w, k, m, n, k, e, nullptr); }
    else { EpiLinear<float, false> e{c, bias, n}; rc = launch_simt<float, float, EpiLinear<float, false>>(h, s, a, k, w, k, m, n, k, e, nullptr); }
    h->launches += h->launches_this_enqueue;
    return rc;
  }
  bf16* scratch = nullptr;
  TRY(static_scratch_address(h, &scratch));
  bf16 *ab = scratch, *wb = scratch + a_elems;
  // every element the GEMM reads is written here first (rows beyond m / n are filled by the TMA unit, not read from
  // memory), so nothing of an earlier call's contents can reach the result and nothing needs clearing
  pack_rows_kernel<bf16><<<(unsigned)(((size_t)m * k + 255) / 256), 256, 0, s>>>(a, ab, m, k, m);
  pack_rows_kernel<bf16><<<(unsigned)((w_elems + 255) / 256), 256, 0, s>>>(w, wb, n, k, n);
  CUDA_TRY(h, cudaGetLastError());
  if (h->use_umma) {
    CUtensorMap ta;
    WeightMaps wm;  // (tensor maps are encoded on the host: no device memory involved)
    rc = make_map(h, &ta, ab, m, k, kUmmaBM);
    if (!rc) rc = make_weight_maps(h, &wm, wb, n, k);
    if (!rc) {
      if (relu) { EpiLinear<float, true> e{c, bias, n}; rc = launch_umma(h, s, ta, wm, m, n, k, e, nullptr); }
      else { EpiLinear<float, false> e{c, bias, n}; rc = launch_umma(h, s, ta, wm, m, n, k, e, nullptr); }
    }
  } else {
    if (relu) { EpiLinear<float, true> e{c, bias, n}; rc = launch_simt<bf16, bf16, EpiLinear<float, true>>(h, s, ab, k, wb, k, m, n, k, e, nullptr); }
    else { EpiLinear<float, false> e{c, bias, n}; rc = launch_simt<bf16, bf16, EpiLinear<float, false>>(h, s, ab, k, wb, k, m, n, k, e, nullptr); }
  }
  h->launches += h->launches_this_enqueue + 2;
  return rc;
}
