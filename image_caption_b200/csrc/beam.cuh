// beam.cuh - batched beam search on the decode engine.
// Restates `_beam_search` (models/vit_transformer_model.py:327-420, grid_transformer_model.py:253-322) for every
// image of the batch at once: the reference handles one image per call and re-runs the whole decoder over the
// prefix each step; here an image's `beam` hypotheses are `beam` rows of the step batch (they share the image's
// cross K/V like the rollouts of an SCST group) and the self-attention cache is never copied when hypotheses are
// re-ranked -- each slot keeps, per position, the cache row that holds its ancestor's K/V (`anc`).
//
// Semantics kept from the reference:
//   * scores are sums of log_softmax values, no length normalisation (vit:364, 378);
//   * step 0 expands hypothesis 0 only (vit:367-375); later steps take the top k of the k x V candidates,
//     k = number of hypotheses still alive (vit:381, 405-409): the beam shrinks as hypotheses end;
//   * a hypothesis that emits <end> moves to the completed list with its score (vit:395-399), in slot order;
//   * an image stops after max_len-1 steps, or -- ViT model -- when every hypothesis of a step ended (vit:402),
//     -- Grid model -- when the completed list is at least as long as the beam is wide at that moment (grid:308);
//   * result = first completed hypothesis with the highest score, else the best live one (vit:412-416).
// torch.topk leaves the order of exactly equal candidates unspecified; here the lower flat index (hypothesis * V +
// word) wins.
#pragma once
#include "common.cuh"
#include "pointwise.cuh"

namespace capdec {

constexpr int kBeamMax = 8;  // hypotheses per image (one warp each in beam_head_kernel)

__global__ void beam_init_kernel(BeamState b, const CallParams* cp, int n_images, int beam, int t_pad) {
  const CallParams p = *cp;
  const int M = n_images * beam;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
  for (int m = tid; m < M; m += nt) {
    b.score[m] = 0.f;
    b.live[m] = (m % beam) == 0;  // step 0: all hypotheses are the same, only the first is expanded (vit:367)
    b.comp_score[m] = 0.f; b.comp_slot[m] = 0; b.comp_pos[m] = 0;
  }
  for (int i = tid; i < n_images; i += nt) { b.k[i] = beam; b.comp_n[i] = 0; }
  for (int i = tid; i < 2 * M * t_pad; i += nt) b.anc[i] = (i / t_pad) % M;  // every slot reads its own row
  for (int i = tid; i < M * p.max_len; i += nt) b.par[i] = i / p.max_len;
  if (tid == 0) *b.images_done = 0;
}

// logits: [M, kVPad] fp32 (bias added).  One CTA per image, one warp per hypothesis slot.
__global__ void __launch_bounds__(kBeamMax * 32)
beam_head_kernel(DecState* st, const CallParams* cp, BeamState b, const float* __restrict__ logits, int n_images,
                 int beam, int V, int t_pad, int grid_rule) {
  if (st->done) return;
  __shared__ float cand[kBeamMax][kVPad];
  __shared__ int sel_idx[kBeamMax];
  __shared__ float sel_score[kBeamMax];
  __shared__ int s_ended[kBeamMax], s_filled[kBeamMax];
  const CallParams p = *cp;
  const int n = st->step;
  const int img = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = img * beam;
  const int k = b.k[img];
  const int* anc_old = b.anc + (size_t)(n & 1) * n_images * beam * t_pad;
  int* anc_new = b.anc + (size_t)((n + 1) & 1) * n_images * beam * t_pad;

  if (warp < beam) {  // candidate scores of this slot: score + log_softmax(logits)  (vit:364, 378)
    const int m = m0 + warp;
    float l[4];
    load_logits_row(logits + (size_t)m * kVPad, lane, V, l);
    float mx = fmaxf(fmaxf(l[0], l[1]), fmaxf(l[2], l[3]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float se = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) se += (lane * 4 + i < V) ? expf(l[i] - mx) : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
    const float lse = logf(se);
    const bool alive = k > 0 && b.live[m] != 0;
    const float sc = b.score[m];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int v = lane * 4 + i;
      cand[warp][v] = (alive && v < V) ? sc + ((l[i] - mx) - lse) : -INFINITY;
    }
  }
  __syncthreads();

  if (warp == 0) {  // top k of the beam x V candidates, ties to the lower flat index
    for (int r = 0; r < k; ++r) {
      float bv = -INFINITY;
      int bi = 0x7fffffff;
      for (int f = lane; f < beam * kVPad; f += 32) {
        const float c = cand[f / kVPad][f % kVPad];
        const int flat = (f / kVPad) * V + (f % kVPad);  // the reference's index: hypothesis * vocab_size + word
        if (c > bv || (c == bv && c != -INFINITY && flat < bi)) { bv = c; bi = flat; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
      }
      if (lane == 0) {
        sel_idx[r] = bi;
        sel_score[r] = bv;
        if (bi != 0x7fffffff) cand[bi / V][bi % V] = -INFINITY;
      }
      __syncwarp();
    }
  }
  __syncthreads();

  // new slot j (< k) continues hypothesis `parent` with `word`; slots >= k are empty from now on
  int parent = warp, word = p.pad_token;
  bool filled = false;
  if (warp < beam) {
    filled = warp < k && sel_idx[warp] != 0x7fffffff;
    if (filled) { parent = sel_idx[warp] / V; word = sel_idx[warp] % V; }
    // ancestry of the new slot: the parent's rows for positions <= n, its own row from n + 1 on
    const int* src = anc_old + (size_t)(m0 + parent) * t_pad;
    int* dst = anc_new + (size_t)(m0 + warp) * t_pad;
    for (int q = lane; q < t_pad; q += 32) dst[q] = (q <= n) ? src[q] : m0 + warp;
  }
  __syncthreads();  // every read of the old per-slot state is done
  if (warp < beam && lane == 0) {
    const int m = m0 + warp;
    const bool ended = filled && word == p.end_token;
    p.ids[(size_t)m * p.max_len + n + 1] = word;  // token of slot m at position n + 1 (input of the next step)
    b.par[(size_t)m * p.max_len + n + 1] = m0 + parent;
    b.score[m] = filled ? sel_score[warp] : 0.f;
    b.live[m] = filled && !ended;
    s_ended[warp] = ended ? 1 : 0;
    s_filled[warp] = filled ? 1 : 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int n_ended = 0;
    if (k > 0) {
      int cn = b.comp_n[img];
      for (int j = 0; j < beam; ++j)  // completion order = slot order (vit:396)
        if (s_filled[j] && s_ended[j]) {
          b.comp_score[m0 + cn] = sel_score[j];
          b.comp_slot[m0 + cn] = m0 + j;
          b.comp_pos[m0 + cn] = n + 1;
          ++cn; ++n_ended;
        }
      b.comp_n[img] = cn;
      int k_new = k - n_ended;                                  // vit:405-409, grid:310-316
      if (grid_rule && n_ended > 0 && cn >= k) k_new = 0;       // grid:308 (k = beam width before this step's removals)
      b.k[img] = k_new;
      if (k_new == 0) atomicAdd(b.images_done, 1);              // vit:402 / grid:308,315
    }
    __threadfence();
    const unsigned t = atomicAdd(&st->ticket, 1u);
    if (t == gridDim.x - 1) {  // last image: advance the step, stop when every beam is empty
      __threadfence();
      st->ticket = 0u;
      st->steps_executed = n + 1;
      st->step = n + 1;
      if (atomicAdd(b.images_done, 0) == n_images) st->done = 1;
      __threadfence();
    }
  }
}

// One warp per image: pick the result (vit:412-416) and walk the parent links back to <start>.
__global__ void beam_finalize_kernel(const DecState* st, const CallParams* cp, BeamState b, int n_images, int beam,
                                     long long* __restrict__ out_ids, int* __restrict__ out_len, float* __restrict__ out_score) {
  const CallParams p = *cp;
  const int img = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (img >= n_images) return;
  const int m0 = img * beam;
  int slot = m0, pos = st->steps_executed;
  float best = -INFINITY;
  const int cn = b.comp_n[img];
  if (cn > 0) {  // first maximum of the completed scores (torch.argmax)
    for (int i = 0; i < cn; ++i)
      if (b.comp_score[m0 + i] > best) { best = b.comp_score[m0 + i]; slot = b.comp_slot[m0 + i]; pos = b.comp_pos[m0 + i]; }
  } else {       // nothing ended within max_len: the best live hypothesis
    for (int j = 0; j < beam; ++j)
      if (b.live[m0 + j] && b.score[m0 + j] > best) { best = b.score[m0 + j]; slot = m0 + j; }
  }
  long long* o = out_ids + (size_t)img * p.max_len;
  for (int q = pos + 1 + lane; q < p.max_len; q += 32) o[q] = p.pad_token;
  if (lane == 0) {
    for (int q = pos; q >= 0; --q) {
      o[q] = p.ids[(size_t)slot * p.max_len + q];
      slot = b.par[(size_t)slot * p.max_len + q];
    }
    out_len[img] = pos + 1;
    if (out_score != nullptr) out_score[img] = best;
  }
}

}  // namespace capdec
