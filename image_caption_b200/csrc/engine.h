// engine.h - host-side state of one caption decode engine (one per process and device).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/capdec.h"
#include "common.cuh"

namespace capdec {

constexpr int kHostCopyChunks = 8;  // capdec_generate_host: image chunks of the host copy

// W operand tensor maps for the three N-tile widths the tcgen05 GEMM is instantiated for.
struct WeightMaps {
  CUtensorMap bn[3];  // 32, 64, 128
  const void* raw = nullptr;  // the packed matrix itself ([rows, K] row-major, engine element type)
  int rows = 0, k = 0;
  CUtensorMap w3;     // [64, rows, K/64] view, box = 64 x 128 rows x 4 k-blocks: half of a K = 512 weight slice per copy
};
// TMA views of one activation matrix for the v2 GEMMs
struct ActMaps {
  CUtensorMap tile;   // box = 64 columns x bnb rows (one k-block)
  CUtensorMap whole;  // [64, rows, K/64] view, box = 64 x bnb rows x 8 k-blocks: all of K = 512 in one copy
};

struct LayerParams {
  // GEMM weights in the engine's element type, [N, K] row-major (nn.Linear layout, K-major)
  void *w_qkv, *w_so, *w_cq, *w_co, *w_1, *w_2;
  float *b_qkv, *b_so, *b_cq, *b_co, *b_1, *b_2;
  float *ln_w[3], *ln_b[3];
  WeightMaps m_qkv, m_so, m_cq, m_co, m_1, m_2;
  CUtensorMap ts_k, ts_v;  // this layer's self K / V cache as a 4-D store view (gemm_ws.cuh, in_proj epilogue)
};

struct StepGraph {
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int kernels = 0;
};

}  // namespace capdec

struct capdec_engine {
  capdec_config cfg;
  int device = 0;
  bool bf16 = false;
  bool use_umma = false;
  bool umma_f32 = false;   // CAPDEC_PREC_TF32: fp32 storage, projections on tcgen05 kind::tf32 (gemm_umma_kernel<.., float>)
  bool use_graph = true;
  size_t esz = 4;
  int t_pad = 64;
  std::string err;

  // ---- parameters (engine-owned copies)
  std::vector<void*> allocs;  // everything to cudaFree at destroy
  capdec::LayerParams layer[CAPDEC_MAX_LAYERS];
  void* w_ckv = nullptr;      // [L*1024, 512] cross K|V projection rows of every layer
  float* b_ckv = nullptr;
  capdec::WeightMaps m_ckv;
  float *emb = nullptr, *pe = nullptr, *fc_w = nullptr, *fc_b = nullptr;  // fc padded to kVPad rows
  bool have_weights = false;

  // ---- encoder hand-off: the encoder's output projection folded into the cross K/V projection
  int proj_dim = 0;              // input width of the folded projection (768 for ViT-B/16); 0 = not set
  void* w_ckv_f = nullptr;       // [L*1024, proj_dim] = w_ckv * proj_w, engine element type
  float* b_ckv_f = nullptr;      // [L*1024] = w_ckv * proj_b + b_ckv
  float* fold_tmp = nullptr;     // fp32 scratch: [proj_dim, 512] transposed projection + [L*1024, proj_dim] product
  void* tok_b = nullptr;         // [max_images*max_mem, proj_dim] gathered tokens (GEMM operand)
  capdec::WeightMaps m_ckv_f;

  // ---- caches
  void *self_k = nullptr, *self_v = nullptr;  // [L][M_max][H][t_pad][64]
  void* cross_kv = nullptr;                   // [L][2][n_images][H][S][64]
  size_t self_layer_elems = 0;

  // ---- step workspace
  float *x = nullptr, *t = nullptr, *logits = nullptr;
  void *xb = nullptr, *q = nullptr, *a = nullptr, *hbuf = nullptr, *mem_b = nullptr;
  unsigned char* fin = nullptr;
  capdec::DecState* st = nullptr;
  capdec::CallParams* cp = nullptr;
  CUtensorMap tm_xb, tm_a, tm_h;  // A-operand maps for the current M
  int maps_for_m = -1;

  // ---- v2 step path (bf16 + tcgen05 only): swap-AB GEMMs, TMA cross-attention, fused head
  bool fast = false;               // v2 kernels in use (off: CAPDEC_FLAG_V1_STEP)
  bool use_pdl = false;            // programmatic dependent launch between the step's kernels
  int bnb = 32;                    // sequences per GEMM CTA (UMMA N) for the current M
  int force_bnb = 0, force_stages = 0, force_ksplit = 0;
  bool no_ffn2_split = false;  // tuning overrides (env CAPDEC_BNB / CAPDEC_STAGES)
  capdec::ActMaps tx_xb, tx_a, tx_h;  // activation maps with box rows = bnb
  CUtensorMap ts_q, ts_h;          // bf16 store views of q [M,512] and the FFN hidden [M,d_ff], box = 128 x 128 (gemm_ws.cuh)
  CUtensorMap ts_x, ts_xb;         // store views of the residual stream (fp32 / bf16), box = 128 x bnb
  int ln_tma_store = 1;            // LayerNorm epilogue leaves through bulk tensor stores
  int force_ln_tma_store = -1;     // env CAPDEC_LN_TMA_STORE
  int force_head_rw = 0;           // env CAPDEC_HEAD_RW: rows per warp of the vocabulary head (1, 2, 4, 6)
  int ws_dbg = 0;
  capdec::WeightMaps m_fc;         // fc_w [128, 512] fp32 as a kind::tf32 operand (vocabulary projection of large batches)
  CUtensorMap tm_x32;              // fp32 view of x [M,512], box = 32 x 128
  int self_attn_its = 4;           // env CAPDEC_SELF_ATTN_ITS: cached positions per chunk / 4 of the streaming self-attention: 4 (16
                                   // positions, 56 registers, 9 CTAs per SM) or 8 (32 positions, 72 registers, 7 CTAs per SM)
  bool no_tf32_head = false;       // env CAPDEC_NO_TF32_HEAD: vocabulary projection on CUDA cores everywhere
  int tf32_head_min_m = 1024;      // env CAPDEC_TF32_HEAD_MIN_M (B200: cfg3, 3072 rows, 807 -> 800 us/step; level at 1024 rows; 1024 keeps the 1536-row shards of cfg4 at N = 2 on the route of the unsharded batch)
  bool no_ws_prefill = false;      // env CAPDEC_NO_WS_PREFILL: prefill through the one-tile-per-CTA GEMM (gemm_umma.cuh)
  int ws_min_ctas = 96;
  int ws_min_m = 512;             // smallest batch that takes the weight-stationary GEMM (env CAPDEC_WS_MIN_M)
  CUtensorMap tx_ws;               // xb as [M,512], box = 64 x 128 rows
  bool no_ws = false;              // env CAPDEC_NO_WS: never use the weight-stationary persistent GEMM (gemm_ws.cuh)
  bool no_whole_k = false;         // env CAPDEC_NO_WHOLE_K: one TMA copy per k-block instead of whole-K copies
  bool dyn_smem_aligned = false;   // dynamic shared memory starts 1 KB-aligned on this device (probed at create)
  bool no_ca_compact = false;      // env CAPDEC_NO_CA_COMPACT: cross-attention with separate merge scratch and alignment slack (3 CTAs per SM at S = 196)
  bool no_cross_stream = false;    // env CAPDEC_NO_CROSS_STREAM: keep the tensor-core cross-attention for group == 1
  CUtensorMap tm_ckv;              // whole cross cache as [rows,64], box = 64 x ca_key_rows(S)
  float* fc_wT = nullptr;          // [512][kVPad] fp32: fc_out.weight transposed + padded
  float* part = nullptr;           // [4][min(M_max,512)][512] fp32 split-K partial products (small batches)
  int part_rows = 0;
  unsigned* tickets = nullptr;     // [row tiles] last-CTA election of the self-closing split-K FFN2 (SwapPartialLN)
  bool no_ffn2_fuse = true;        // env CAPDEC_FFN2_FUSE=1 clears it: the split-K FFN2 closes itself (no sum_ln launch)

  // ---- cluster-resident projection chains (layer_chain.cuh): in_proj | out_proj+LN+cross-q | out_proj+LN+FFN+LN[+in_proj]
  bool chain = false;              // in use (bf16 + tcgen05 v2 step, d_ff == 2048; env CAPDEC_CHAIN=0 turns it off)
  bool chain_now = false;          // chains used for the current batch size (all clusters co-resident, or forced)
  bool chain_always = false;       // env CAPDEC_CHAIN=2
  int chain_cs = 8;                // CTAs per cluster (env CAPDEC_CHAIN_CS: 4 or 8)
  int chain_bnb = 32;              // sequence rows per cluster for the current M (env CAPDEC_CHAIN_BNB)
  int force_chain_bnb = 0;
  int chain_max_clusters = 0;      // co-resident clusters of the FFN chain kernel (cudaOccupancyMaxActiveClusters), per BNB below
  int chain_clusters[2] = {0, 0};  // [BNB 32, BNB 64]
  unsigned char* chain_w = nullptr;  // [CS][L][UL][16 KB] per-CTA weight units in consumption order
  float* chain_part = nullptr;     // [row tiles][CS][BNB][512] fp32 linear2 partial products
  CUtensorMap tc_xb, tc_a;         // [64, M, 8] views of xb / a, box = 64 x chain_bnb x 8

  // ---- full-sequence teacher-forced pass (capdec_forward_full): its own activation workspace, allocated on first use
  float *full_x = nullptr, *full_logits = nullptr;
  void *full_xb = nullptr, *full_q = nullptr, *full_a = nullptr, *full_h = nullptr;
  size_t full_rows_cap = 0;        // rows (sequence positions) per chunk of images (env CAPDEC_FULL_ROWS, default 65536)

  // ---- CIDEr-D reward workspace (cider.cuh), grown on demand
  unsigned long long* cider_keys = nullptr;
  unsigned int* cider_cnt = nullptr;
  int* cider_off = nullptr;
  size_t cider_tab = 0;
  int cider_off_cap = 0;

  // ---- host-buffer entry point staging (device side)
  float* stage_mem[2] = {nullptr, nullptr};   // two slots: capdec_prefetch_host fills one while the other is decoded
  long long* stage_ids = nullptr;
  float* stage_logp = nullptr;
  cudaStream_t copy_stream = nullptr;          // host -> device chunks of the memory, overlapped with the prefill
  cudaEvent_t copy_ev[2][8] = {};              // [slot][chunk]
  struct Prefetch { const float* host; int n_images, mem_len; bool valid; unsigned long long seq; };
  Prefetch pf[2] = {{nullptr, 0, 0, false, 0}, {nullptr, 0, 0, false, 0}};
  unsigned long long pf_seq = 0;
  int last_slot = 1;                           // slot of the newest prefetch / un-prefetched copy
  cudaEvent_t copy_start = nullptr;

  // ---- current rollout
  int n_images = 0, mem_len = 0;
  bool prefilled = false;
  bool prefilled_once = false;
  std::map<std::tuple<int, int, int, int>, capdec::StepGraph> graphs;  // (n_images, mem_len, group, steps in the graph)
  // ---- beam search: the step kernels with an ancestry table for self-attention and a per-image head
  bool beam_mode = false;          // set for the duration of capdec_beam_search
  bool beam_v1 = false;            // env CAPDEC_BEAM_V1: beam search on the v1 step kernels also in bf16 + tcgen05 mode
  int beam_grid_rule = 0;          // stop rule of the Grid model (grid:308) instead of the ViT model's (vit:402)
  capdec::BeamState beam{};        // buffers allocated on first use
  long long* beam_tok = nullptr;   // [M_max][max_len] token of each slot at each position
  int graph_steps = 7;  // decode steps captured per graph (env CAPDEC_GRAPH_STEPS); B200 cfg2: 317.3 us/step with 1, 313.4 with 7, 312.6 with 49
  long long launches = 0;
  int launches_this_enqueue = 0;

  // ---- per-kernel timing (capdec_profile_*): events around every launch of a direct (non-graph) rollout
  struct ProfRec { const char* tag; cudaEvent_t e0, e1; double bytes; };
  bool profiling = false;
  const char* tag = "other";
  int prof_step_n = 1;             // cached self positions of the step being enqueued (direct launches)
  double tag_bytes = 0.0;          // algorithmic bytes of the launch being enqueued
  std::vector<ProfRec> prof;
  std::string prof_report;
};
