// full_pass.cuh - pointwise ends of the full-sequence (teacher-forced) pass: the reference's
// `model.decoder(ids[:, :-1], memory, tgt_mask=causal)` over whole captions (vit:155-182 called as at scst:223 / the
// training forward vit:216-255) followed by `log_softmax(...).gather(ids[:, 1:])` with finished rows zeroed
// (scst:226-239).  Rows of every activation matrix are (sequence, position), sequence-major: row = seq * T + pos.
#pragma once
#include "common.cuh"
#include "pointwise.cuh"

namespace capdec {

// x[row] = emb[ids[seq, pos]] * sqrt(512) + pe[pos]  (vit:165-167), fp32 residual stream + bf16 GEMM operand.
// ids: [n_seq_total, max_len] int64; this launch covers sequences [seq0, seq0 + n_seq), T = max_len - 1 positions each.
__global__ void __launch_bounds__(256)
embed_full_kernel(const long long* __restrict__ ids, int max_len, int vocab, int seq0, int n_seq, int T,
                  const float* __restrict__ emb, const float* __restrict__ pe, float* __restrict__ x, bf16* __restrict__ xb) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // one float4 each
  if (idx >= (size_t)n_seq * T * (kD / 4)) return;
  const int row = (int)(idx / (kD / 4)), c = (int)(idx % (kD / 4)) * 4;
  const int sq = row / T, pos = row - sq * T;
  const int tok = clamp_token(ids[(size_t)(seq0 + sq) * max_len + pos], vocab);
  const float4 e = __ldg(reinterpret_cast<const float4*>(emb + (size_t)tok * kD + c));
  const float4 p = __ldg(reinterpret_cast<const float4*>(pe + (size_t)pos * kD + c));
  const float sc = 22.627416997969522f;  // sqrt(512) rounded to fp32, as torch does for the scalar (same as embed_pe_kernel)
  float4 r;
  r.x = __fadd_rn(__fmul_rn(e.x, sc), p.x); r.y = __fadd_rn(__fmul_rn(e.y, sc), p.y);
  r.z = __fadd_rn(__fmul_rn(e.z, sc), p.z); r.w = __fadd_rn(__fmul_rn(e.w, sc), p.w);
  *reinterpret_cast<float4*>(x + (size_t)row * kD + c) = r;
  store4(xb + (size_t)row * kD + c, r.x, r.y, r.z, r.w);
}

// One warp per row.  logits: [n_seq * T, kVPad] fp32 (bias added).  logp_out[seq, pos] = log_softmax(logits)[ids[seq,
// pos + 1]], or 0 when one of ids[seq, 1 .. pos] is <end> (scst:236-239: finished rows contribute nothing);
// logits_out[seq, pos, :V] = the logits (optional).  Same arithmetic as the step head (warp_pick with a forced token).
__global__ void __launch_bounds__(128)
logp_full_kernel(const float* __restrict__ logits, const long long* __restrict__ ids, int max_len, int V, int seq0, int n_seq,
                 int T, int end_token, float* __restrict__ logp_out, float* __restrict__ logits_out) {
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= n_seq * T) return;
  const int sq = row / T, pos = row - sq * T;
  const long long* seq_ids = ids + (size_t)(seq0 + sq) * max_len;
  float l[4];
  load_logits_row(logits + (size_t)row * kVPad, lane, V, l);
  if (logits_out != nullptr) {
    float* d = logits_out + ((size_t)(seq0 + sq) * T + pos) * V;
#pragma unroll
    for (int i = 0; i < 4; ++i) if (lane * 4 + i < V) d[lane * 4 + i] = l[i];
  }
  if (logp_out == nullptr) return;
  bool was = false;
  for (int j = 1 + lane; j <= pos; j += 32) was |= seq_ids[j] == (long long)end_token;
  was = __any_sync(0xffffffffu, was);
  float lp;
  warp_pick(l, lane, false, 0.f, V, &lp, clamp_token(seq_ids[pos + 1], V));
  if (lane == 0) logp_out[(size_t)(seq0 + sq) * T + pos] = was ? 0.f : lp;
}

}  // namespace capdec
