/*
 * capdec.h - C ABI of the B200-native caption decode engine (libcapdec.so).
 *
 * Drop-in boundary for the batched autoregressive decoding path of
 * wonder-dream/image_caption.  Plain pointers and sizes only; no torch types, no C++
 * exceptions cross this boundary.  Every entry point returns 0 on success or a negative
 * capdec_status; capdec_last_error() gives the text.
 *
 * What each entry point replaces in the reference (paths relative to the reference root):
 *   capdec_set_weights   the decoder's nn.Module parameters, read by state-dict name
 *                        (models/vit_transformer_model.py:111-139, grid_transformer_model.py:119-134)
 *   capdec_prefill       the memory->K,V projection torch redoes inside every decoder call
 *                        (vit_transformer_model.py:312 -> nn.TransformerDecoderLayer._mha_block)
 *   capdec_decode        the greedy loop  vit_transformer_model.py:307-323 / grid_transformer_model.py:237-249
 *                        and the sampling loop utils/scst_loss.py:216-252 (one call = one rollout)
 *   capdec_generate_host the same two loops entered with HOST buffers (what generate()/SCSTLoss see
 *                        when tensors live on the CPU): H2D, prefill, decode, D2H in one call
 *   capdec_sample_logits softmax + multinomial + log_softmax.gather of utils/scst_loss.py:229-239
 *                        on caller-provided logits (used to test the sampling distribution)
 *
 * Sequences are image-major: sequence m belongs to image m / group; inside a group rows
 * g < n_sample are sampled (scst semantics: cumulative `finished`, log-prob 0 once finished),
 * rows g >= n_sample are greedy (vit semantics).  Plain greedy = {group 1, n_sample 0}; the
 * reference's single-sample rollout = {group 1, n_sample 1}; multi-sample SCST with a greedy
 * baseline (BASELINE.json config 3) = {group 6, n_sample 5}.
 *
 * Threading: one handle per (process, device); calls on one handle must be serialised by the
 * caller.  Work is enqueued on the caller's stream; only capdec_decode's read-back of the
 * step count and capdec_generate_host synchronise.
 */
#ifndef CAPDEC_H_
#define CAPDEC_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CAPDEC_ABI_VERSION 1
#define CAPDEC_MAX_LAYERS 12

typedef struct capdec_engine* capdec_handle;

typedef enum {
  CAPDEC_OK = 0,
  CAPDEC_ERR_INVALID = -1,     /* bad argument / unsupported shape                */
  CAPDEC_ERR_CUDA = -2,        /* a CUDA runtime/driver call failed               */
  CAPDEC_ERR_STATE = -3,       /* call order violated (e.g. decode before prefill) */
  CAPDEC_ERR_CAPACITY = -4,    /* exceeds the sizes given to capdec_create         */
  CAPDEC_ERR_NO_DEVICE = -5    /* no usable sm_100 device                          */
} capdec_status;

typedef enum {
  CAPDEC_PREC_FP32 = 0, /* fp32 weights / KV / CUDA-core GEMMs: the strict-parity mode         */
  CAPDEC_PREC_BF16 = 1, /* bf16 weights / KV, tcgen05 GEMMs with fp32 accumulate; fp32 residual */
  CAPDEC_PREC_TF32 = 2  /* fp32 weights / KV / step kernels as FP32, but every projection on tcgen05 kind::tf32 (operands read
                           with a 10-bit mantissa, fp32 accumulate): the fp32-class throughput mode                       */
} capdec_precision;

typedef struct {
  /* Hard shape limits of this build (capdec_create returns CAPDEC_ERR_INVALID outside them):
   * vocab in [2,128], d_model == 512, n_head == 8, d_ff a multiple of 128, max_len in [2,128],
   * memory length <= 256 on the bf16 tensor-core path.  Every shipped reference config
   * (d_model 512 / nhead 8 / ffn 2048 / vocab 109 / max_len <= 64, S = 196 or 49) is inside them. */
  int32_t vocab;       /* V   (data/vocab.json: 109); 2..128                  */
  int32_t d_model;     /* must be 512                                          */
  int32_t n_head;      /* must be 8 (head dim 64)                              */
  int32_t n_layer;     /* 1..CAPDEC_MAX_LAYERS                                 */
  int32_t d_ff;        /* multiple of 128                                      */
  int32_t pe_rows;     /* rows of pos_encoder.pe (the model's max_len)         */
  int32_t max_mem;     /* largest memory length S (196 ViT, 49 grid)           */
  int32_t max_images;  /* largest number of images per call                    */
  int32_t max_group;   /* largest sequences-per-image                          */
  int32_t max_len;     /* largest generate(max_len=...); 2..128               */
  int32_t precision;   /* capdec_precision                                     */
  int32_t flags;       /* CAPDEC_FLAG_*                                        */
} capdec_config;

#define CAPDEC_FLAG_NO_GRAPH 1   /* launch step kernels directly instead of replaying a CUDA graph */
#define CAPDEC_FLAG_NO_UMMA 2    /* bf16 mode: route GEMMs through the CUDA-core kernel (debug)     */
#define CAPDEC_FLAG_V1_STEP 4    /* bf16 mode: the first-generation step kernels (one GEMM tile per CTA,
                                    separate LayerNorm / vocabulary / head kernels); kept for A/B runs */
#define CAPDEC_FLAG_NO_PDL 8     /* do not chain the step's kernels with programmatic dependent launch */
#define CAPDEC_FLAG_BATCH_INVARIANT 32 /* bf16 mode: ONE kernel route whatever the batch size (64-row swap-AB tiles, no
                                    projection chains, no weight-stationary GEMM, no split-K FFN2), so that a sequence's
                                    tokens and log-probs are bit-identical however the batch is sharded or split.  The
                                    default picks the fastest route per batch size; routes differ in fp32 summation
                                    order, so results across DIFFERENT batch sizes then agree only within the bf16
                                    tolerance (and a near-tie may flip a token). */

/* fp32 DEVICE pointers, borrowed for the duration of capdec_set_weights only. */
typedef struct {
  const float* self_in_w;   /* [3D, D]  self_attn.in_proj_weight       */
  const float* self_in_b;   /* [3D]                                      */
  const float* self_out_w;  /* [D, D]   self_attn.out_proj.weight      */
  const float* self_out_b;  /* [D]                                       */
  const float* cross_in_w;  /* [3D, D]  multihead_attn.in_proj_weight  */
  const float* cross_in_b;  /* [3D]                                      */
  const float* cross_out_w; /* [D, D]   multihead_attn.out_proj.weight */
  const float* cross_out_b; /* [D]                                       */
  const float* lin1_w;      /* [F, D]   linear1.weight                  */
  const float* lin1_b;      /* [F]                                       */
  const float* lin2_w;      /* [D, F]   linear2.weight                  */
  const float* lin2_b;      /* [D]                                       */
  const float* norm_w[3];   /* [D]      norm1/2/3.weight                */
  const float* norm_b[3];   /* [D]      norm1/2/3.bias                  */
} capdec_layer_weights;

typedef struct {
  const float* embedding;   /* [V, D]        embedding.weight  */
  const float* pe;          /* [pe_rows, D]  pos_encoder.pe    */
  const float* fc_out_w;    /* [V, D]        fc_out.weight     */
  const float* fc_out_b;    /* [V]           fc_out.bias       */
  capdec_layer_weights layer[CAPDEC_MAX_LAYERS];
} capdec_weights;

typedef struct {
  int32_t group;        /* sequences per image (>= 1)                                          */
  int32_t n_sample;     /* rows g < n_sample of each group are sampled, the rest greedy        */
  int32_t start_token;
  int32_t end_token;
  int32_t pad_token;    /* fills ids_out beyond the last executed step                         */
  int32_t max_len;      /* as generate(max_len=...): at most max_len-1 steps                   */
  int32_t flags;        /* CAPDEC_DECODE_*                                                      */
  int32_t reserved;     /* must be 0                                                            */
  uint64_t seed;        /* Philox key                                                           */
  uint64_t seq_id_base; /* global index of this call's first sequence (sharding invariance)    */
  int64_t* ids_out;     /* [n_seq, max_len]                                                     */
  float* logp_out;      /* [n_seq, max_len-1] or NULL; 0 for greedy rows and finished rows      */
  float* logits_out;    /* NULL, or [n_seq, max_len-1, vocab]: every step's logits (parity)     */
  int32_t* steps_out;   /* HOST int: number of steps executed (L = steps+1); may be NULL       */
} capdec_decode_args;

/* Run all max_len-1 steps even if the stop condition is met (sharded greedy rollouts: the
 * global "all rows emit <end> on one step" test of vit:322 needs every rank's later tokens). */
#define CAPDEC_DECODE_NO_EARLY_STOP 1
/* Teacher forcing: ids_out holds the whole sequences on entry ([M, max_len], column 0 = <start>) and is not written;
 * step n feeds ids[:, n] and scores ids[:, n+1] (logp_out, rows g < n_sample; 0 after a row's <end>, scst:239) instead
 * of picking a token; logits_out receives every step's logits.  This is the reference's per-step call
 * `model.decoder(prefix, memory, tgt_mask=causal)` (utils/scst_loss.py:223, vit_transformer_model.py:312) evaluated for
 * a given sequence: row k of its output = the logits of step k.  Stop rules apply to the forced tokens. */
#define CAPDEC_DECODE_TEACHER_FORCED 2

int capdec_abi_version(void);
const char* capdec_last_error(capdec_handle h); /* h may be NULL: error of a failed create */

int capdec_create(const capdec_config* cfg, int device, capdec_handle* out);
int capdec_destroy(capdec_handle h);

/* Re-pack (and in bf16 mode round) the module's fp32 parameters into engine-owned buffers. */
int capdec_set_weights(capdec_handle h, const capdec_weights* w, void* cuda_stream);

/* memory: DEVICE fp32 [n_images, mem_len, D] contiguous.  Projects it once to every layer's
 * cross-attention K and V. */
int capdec_prefill(capdec_handle h, const float* memory, int32_t n_images, int32_t mem_len, void* cuda_stream);

/* Encoder hand-off.  The ViT encoder ends in `features = self.projection(patch_features)`
 * (models/vit_transformer_model.py:61,95-98) and the decoder only ever multiplies that memory by the cross-attention
 * K/V weights, so the projection can be folded into them:  K|V = tokens (Wkv P)^T + (Wkv pb + bkv).
 * capdec_set_memory_projection forms the folded weights on the device (fp32 product, then the engine's element type);
 * call it after capdec_set_weights and again whenever either set of parameters changes.
 *   proj_w [D, in_dim], proj_b [D]: DEVICE fp32 (nn.Linear layout); in_dim a multiple of 64.
 *   cross_in_w / cross_in_b: HOST arrays of n_layer DEVICE pointers to multihead_attn.in_proj_weight [3D, D] / in_proj_bias.
 * capdec_prefill_tokens then replaces encoder.projection + capdec_prefill: tokens is DEVICE fp32
 * [n_images, tokens_per_image, in_dim]; rows first_token .. first_token+mem_len-1 of every image are the memory
 * positions (first_token = 1 drops the class token, vit:95). */
int capdec_set_memory_projection(capdec_handle h, const float* proj_w, const float* proj_b, int32_t in_dim,
                                 const float* const* cross_in_w, const float* const* cross_in_b, void* cuda_stream);
int capdec_prefill_tokens(capdec_handle h, const float* tokens, int32_t n_images, int32_t tokens_per_image,
                          int32_t first_token, int32_t mem_len, void* cuda_stream);

/* One rollout over the images of the last prefill.  ids_out/logp_out/logits_out are DEVICE
 * pointers.  Synchronises the stream only if steps_out != NULL. */
int capdec_decode(capdec_handle h, const capdec_decode_args* args, void* cuda_stream);

/* Same with HOST pointers for memory, ids_out, logp_out (logits_out must be NULL): copies in,
 * prefills, decodes, copies out and synchronises. */
int capdec_generate_host(capdec_handle h, const float* memory_host, int32_t n_images, int32_t mem_len,
                         const capdec_decode_args* args_host, void* cuda_stream);

/* Optional overlap for back-to-back capdec_generate_host calls (no counterpart in the reference, whose memory never leaves
 * the device; it exists for the host-buffer boundary above): start the host -> device copy of the memory that a LATER
 * capdec_generate_host call will be given, on the engine's copy stream, so that it runs underneath the decode of the call
 * in between.  That later call (same pointer, n_images and mem_len) then waits for this copy instead of issuing its own.
 * At most two prefetches may be outstanding (CAPDEC_ERR_STATE beyond); the host buffer must stay unchanged (and should be
 * pinned) until the call that consumes it returns; a capdec_generate_host that matches no prefetch copies as before. */
int capdec_prefetch_host(capdec_handle h, const float* memory_host, int32_t n_images, int32_t mem_len);

/* Batched beam search over the images of the last prefill: the third `generate` method of the reference,
 * `_beam_search` (models/vit_transformer_model.py:327-420, models/grid_transformer_model.py:253-322), which the
 * reference runs one image at a time and without a KV cache.  Same rules: scores are sums of log-softmax values
 * (no length normalisation), step 0 expands one hypothesis, the beam shrinks as hypotheses emit <end>, an image
 * stops when all of a step's hypotheses ended (ViT model, vit:402; the Grid model's rule is CAPDEC_BEAM_STOP_GRID)
 * or after max_len-1 steps, the result is the first completed
 * hypothesis with the highest score, else the best live one.  beam_size <= min(8, max_group of the handle).
 * ids_out DEVICE int64 [n_images, max_len]: row i holds its sequence (with <start>, and <end> if it completed)
 * followed by pad_token; len_out DEVICE int32 [n_images]; score_out DEVICE fp32 [n_images] or NULL.
 * steps_out as in capdec_decode. */
#define CAPDEC_BEAM_STOP_GRID 1 /* Grid model's stop rule (grid:308): stop once len(completed) >= current beam width */
typedef struct capdec_beam_args {
  int32_t beam_size;
  int32_t flags;
  int32_t start_token, end_token, pad_token;
  int32_t max_len;
  int64_t* ids_out;
  int32_t* len_out;
  float* score_out;
  int32_t* steps_out;
} capdec_beam_args;
int capdec_beam_search(capdec_handle h, const capdec_beam_args* args, void* cuda_stream);

/* Sampler alone: logits DEVICE fp32 [n_rows, vocab]; row r uses sequence id seq_id_base + r.
 * mode 0 = argmax, 1 = sample.  ids_out DEVICE int64 [n_rows], logp_out DEVICE fp32 [n_rows] or NULL. */
int capdec_sample_logits(capdec_handle h, const float* logits, int32_t n_rows, int32_t mode, uint64_t seed,
                         uint64_t seq_id_base, int32_t step, int64_t* ids_out, float* logp_out, void* cuda_stream);

/* Number of kernels this handle has launched (graph replays count their kernel nodes). */
int64_t capdec_launch_count(capdec_handle h);

/* Per-kernel device timing for bench.py's roofline: while enabled, rollouts launch their kernels
 * directly (no graph) with a CUDA-event pair around every launch on the launching stream.
 * capdec_profile_report synchronises and writes JSON {"tag": {"launches","ms","bytes"}} where
 * bytes are the ALGORITHMIC bytes of those launches (DESIGN.md lists the formula per tag). */
int capdec_profile_enable(capdec_handle h, int32_t on);
int capdec_profile_report(capdec_handle h, char* buf, int32_t buf_len);

/* ---- Full-sequence teacher-forced pass (forward only).
 * Replaces the reference's whole-caption decoder call `self.decoder(captions[:, :-1], memory, tgt_mask=causal)`
 * (models/vit_transformer_model.py:216-255 training forward; utils/scst_loss.py:223-226 is the same call on a growing
 * prefix) plus `log_softmax(...).gather(...)` with finished rows zeroed (utils/scst_loss.py:226-239), for GIVEN token
 * sequences: one pass over (sequences x positions) rows instead of max_len-1 decode steps.  Same results as
 * capdec_decode with CAPDEC_DECODE_TEACHER_FORCED (same weights, caches and head arithmetic; the attention and GEMM
 * summation orders differ, so logits agree to the bf16 tolerance, not bit for bit).  Needs capdec_prefill first and a bf16
 * tcgen05 engine (CAPDEC_ERR_STATE otherwise: there is no fallback).  Leaves the self K/V cache holding positions
 * [0, max_len-1) of every sequence.  Images are processed in chunks of <= CAPDEC_FULL_ROWS (default 65536) rows.
 *   ids        DEVICE int64 [n_images*group, max_len], column 0 = <start> (values outside the vocabulary are clamped)
 *   logp_out   DEVICE fp32 [n_images*group, max_len-1] or NULL: log p(ids[:, k+1] | ids[:, :k+1]), 0 once a row's
 *              ids[:, 1..k] contains end_token
 *   logits_out DEVICE fp32 [n_images*group, max_len-1, vocab] or NULL                                             */
typedef struct capdec_forward_args {
  int32_t group;
  int32_t max_len;
  int32_t end_token;
  int32_t reserved;
  const int64_t* ids;
  float* logp_out;
  float* logits_out;
} capdec_forward_args;
int capdec_forward_full(capdec_handle h, const capdec_forward_args* a, void* cuda_stream);

/* ---- diagnostics (used by the parity tests; not part of the reference-facing surface) ---- */

/* C[M,N] = A[M,K] * W[N,K]^T + bias through the engine's GEMM of the handle's precision
 * (bf16 mode: operands are rounded to bf16, tcgen05 path).  All DEVICE fp32. relu: 0/1. */
int capdec_debug_gemm(capdec_handle h, const float* a, const float* w, const float* bias, float* c,
                      int32_t m, int32_t n, int32_t k, int32_t relu, void* cuda_stream);

/* Copy layer `layer`'s cross K (which=0) or V (which=1) of the last prefill to out as DEVICE fp32
 * [n_images, mem_len, D] (the reference's memory @ W^T + b layout). */
int capdec_debug_cross_kv(capdec_handle h, int32_t layer, int32_t which, float* out, void* cuda_stream);

/* CIDEr-D reward of an SCST step on the device: what CiderRewardCalculator.compute_reward
 * (utils/scst_loss.py:20-54, called at scst:179-180 on the strings of scst:256-269) returns per prediction, computed from
 * token ids.  The arithmetic lives in pycocoevalcap (not vendored by the reference, not installable here): this follows the
 * published CIDEr-D algorithm as restated in oracle/cider_d.py; parity is against that restatement ("unpinned").
 * Sequences are read as the reference reads them: stop at the first end_token, skip start_token and pad_token.
 * Token ids must be < 32768; at most 128 words per sequence are used; all references of one image together may hold at
 * most 4096 n-grams (about 8 references of 128 words), else CAPDEC_ERR_CAPACITY.  The scorer's corpus is THIS call's
 * references (document frequencies and log N from the n_images reference sets), as in the reference's per-batch call. */
typedef struct {
  const int64_t* hyp_ids;      /* DEVICE [n_images * group, hyp_len]: hypotheses, image-major             */
  const int64_t* ref_ids;      /* DEVICE [n_refs, ref_len]: reference captions as token ids                 */
  const int32_t* ref_offsets;  /* HOST   [n_images + 1]: references of image i = rows ref_offsets[i] .. ref_offsets[i+1]-1 */
  int32_t n_images, group, hyp_len, n_refs, ref_len;
  int32_t start_token, end_token, pad_token;
  double* scores_out;          /* DEVICE [n_images * group] fp64, as the reference's Python floats          */
} capdec_cider_args;
int capdec_cider_d(capdec_handle h, const capdec_cider_args* a, void* cuda_stream);

/* Allocation-free projection: C[m,n] = act(A[m,k] W[n,k]^T + bias), fp32 DEVICE pointers, the computation of
 * capdec_debug_gemm (same kernels, same rounding: bf16 operands / fp32 accumulate on a bf16 engine) with the bf16 operand
 * scratch in a file-scope device array instead of cudaMalloc'd buffers.  It has NO counterpart in the reference: it was
 * added for callers that cannot call cudaMalloc (CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  It calls no cudaMalloc / cudaFree / allocating library routine and does not
 * synchronise: the work is enqueued on `cuda_stream`.
 *   limit   bf16 engine: (m rounded so that m*k is a multiple of 128) * k + n * k <= 2 Mi elements (4 MiB of scratch, held
 *           by every GPU that loads the library until exit); beyond it CAPDEC_ERR_CAPACITY.  Shapes as capdec_debug_gemm:
 *           k % 64 == 0 and n % 128 == 0 on the tensor-core route.  fp32 engine: no scratch, no limit.
 *   sharing every call on a device uses the SAME array, which keeps the previous call's contents; each call overwrites
 *           all it reads, so nothing needs clearing -- but two calls that overlap on the GPU (different streams, or a
 *           graph replayed concurrently) RACE on it.  Serialise them on one stream.
 *   state   CAPDEC_ERR_STATE while capdec_profile_enable(h, 1) is in effect (its event pairs are allocations). */
int capdec_gemm_static(capdec_handle h, const float* a, const float* w, const float* bias, float* c, int32_t m, int32_t n,
                       int32_t k, int32_t relu, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* CAPDEC_H_ */
