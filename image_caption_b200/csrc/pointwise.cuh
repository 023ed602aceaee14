// pointwise.cuh - the non-GEMM, non-attention kernels of the decode step:
//   init_rollout_kernel  rollout state, ids[:,0] = <start>          (vit:307, scst:216-218)
//   embed_pe_kernel      embedding * sqrt(d) + pe[n]                 (vit:166-169)
//   add_ln_kernel        x = LayerNorm(x + t), eps 1e-5              (torch post-LN blocks, K5/K9/K10)
//   head_kernel          argmax | softmax-sample + log-prob, finished bookkeeping,
//                        stop condition and step advance               (vit:315-323, scst:226-249, K12/K13)
//   convert / gather helpers for prefill and diagnostics
#pragma once
#include "common.cuh"

namespace capdec {

// teacher forcing reads caller-supplied token ids: an id outside the vocabulary must not become an out-of-bounds read
__device__ __forceinline__ int clamp_token(long long t, int vocab) { return t < 0 ? 0 : (t >= vocab ? vocab - 1 : (int)t); }

__global__ void init_rollout_kernel(DecState* st, const CallParams* cp, unsigned char* fin, int M) {
  const CallParams p = *cp;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid == 0) {
    st->step = 0; st->done = 0; st->steps_executed = 0; st->ticket = 0u;
    st->cnt_end_now = 0; st->cnt_finished = 0;
    st->greedy_done = (p.n_greedy_rows == 0);
    st->sample_done = (p.n_sample_rows == 0);
  }
  const int total = M * p.max_len;
  for (int i = tid; i < total; i += gridDim.x * blockDim.x) {
    if (!p.forced) p.ids[i] = (i % p.max_len == 0) ? p.start_token : p.pad_token;
    if (p.logp != nullptr && i < M * (p.max_len - 1)) p.logp[i] = 0.f;
  }
  if (tid < M) fin[tid] = 0;
}

// x[m,:] = emb[tok]*sqrt(512) + pe[n,:]; separate multiply and add (no FMA contraction) so the
// fp32 mode reproduces the reference's two-op rounding.
// The per-call parameters reach the device as a kernel argument (see capdec_decode).
__global__ void set_call_params_kernel(CallParams* dst, const CallParams src) { *dst = src; }

template <typename Elem>
__global__ void embed_pe_kernel(const DecState* st, const CallParams* cp, const float* __restrict__ emb,
                                const float* __restrict__ pe, float* __restrict__ x, Elem* __restrict__ xb, int M) {
  asm volatile("griddepcontrol.wait;" ::: "memory");  // ids / DecState of the previous step (PDL launches)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (st->done) return;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;  // one float4 each
  if (idx >= M * (kD / 4)) return;
  const int m = idx / (kD / 4), c = (idx % (kD / 4)) * 4;
  const int n = st->step;
  if (n >= cp->max_len - 1) return;  // launched after the last step's head (v2 beam search): no next position
  long long tok = cp->ids[(size_t)m * cp->max_len + n];
  if (cp->forced) tok = clamp_token(tok, cp->vocab);  // caller-supplied sequence (teacher forcing): never index outside the table
  const float4 e = __ldg(reinterpret_cast<const float4*>(emb + (size_t)tok * kD + c));
  const float4 p = __ldg(reinterpret_cast<const float4*>(pe + (size_t)n * kD + c));
  const float sc = 22.627416997969522f;  // sqrt(512) rounded to fp32, as torch does for the scalar
  float4 r;
  r.x = __fadd_rn(__fmul_rn(e.x, sc), p.x); r.y = __fadd_rn(__fmul_rn(e.y, sc), p.y);
  r.z = __fadd_rn(__fmul_rn(e.z, sc), p.z); r.w = __fadd_rn(__fmul_rn(e.w, sc), p.w);
  *reinterpret_cast<float4*>(x + (size_t)m * kD + c) = r;
  if (xb != nullptr) store4(xb + (size_t)m * kD + c, r.x, r.y, r.z, r.w);
}

// One warp per row of 512: x <- LN(x + t) * gamma + beta (biased variance, eps inside sqrt).
template <typename Elem>
__global__ void __launch_bounds__(128)
add_ln_kernel(float* __restrict__ x, const float* __restrict__ t, const float* __restrict__ gamma,
              const float* __restrict__ beta, Elem* __restrict__ xb, int M, const DecState* st) {
  if (st->done) return;
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= M) return;
  const int lane = threadIdx.x & 31;
  float v[16];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 128 + lane * 4;
    const float4 a = *reinterpret_cast<const float4*>(x + (size_t)row * kD + c);
    const float4 b = *reinterpret_cast<const float4*>(t + (size_t)row * kD + c);
    v[4 * i] = a.x + b.x; v[4 * i + 1] = a.y + b.y; v[4 * i + 2] = a.z + b.z; v[4 * i + 3] = a.w + b.w;
    s += (v[4 * i] + v[4 * i + 1]) + (v[4 * i + 2] + v[4 * i + 3]);
  }
  const float mean = warp_sum(s) * (1.f / kD);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { v[i] -= mean; q = fmaf(v[i], v[i], q); }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / kD) + 1e-5f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 128 + lane * 4;
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c));
    const float4 b = __ldg(reinterpret_cast<const float4*>(beta + c));
    const float r0 = v[4 * i] * rstd * g.x + b.x, r1 = v[4 * i + 1] * rstd * g.y + b.y;
    const float r2 = v[4 * i + 2] * rstd * g.z + b.z, r3 = v[4 * i + 3] * rstd * g.w + b.w;
    *reinterpret_cast<float4*>(x + (size_t)row * kD + c) = make_float4(r0, r1, r2, r3);
    if (xb != nullptr) store4(xb + (size_t)row * kD + c, r0, r1, r2, r3);
  }
}

// x <- LN(x + bias + sum_z part[z]) : closes a split-K projection whose k-slices wrote fp32 partial
// products part[z][M][512] (small batches: FFN2's K = 2048 divided over 4 CTAs per tile).  One warp per row.
// NZ is a compile-time constant so that all 4 + 4*NZ row loads are in flight together (with a run-time slice count
// they were issued one slice at a time: ~2.9 us of the kernel's in-chain time at cfg2, measured by skipping it);
// bias / gamma / beta do not depend on the step and are read before the programmatic-launch wait.
template <int NZ>
__global__ void __launch_bounds__(128)
sum_ln_kernel(float* __restrict__ x, const float* __restrict__ part, const float* __restrict__ bias,
              const float* __restrict__ gamma, const float* __restrict__ beta, bf16* __restrict__ xb, int M, const DecState* st) {
  const int lane = threadIdx.x & 31;
  float4 bi[4], ga[4], be[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 128 + lane * 4;
    bi[i] = __ldg(reinterpret_cast<const float4*>(bias + c));
    ga[i] = __ldg(reinterpret_cast<const float4*>(gamma + c));
    be[i] = __ldg(reinterpret_cast<const float4*>(beta + c));
  }
  const bool done = st->done != 0;  // written by the previous step's head kernel, long complete
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (done) return;
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= M) return;
  sum_ln_row<NZ>(x, part, M, row, lane, bi, ga, be, xb);
}

// Warp-level argmax / categorical draw over one row of kVPad logits (entries >= V are -inf).
// Lane l owns vocabulary entries 4l .. 4l+3.  Returns token; *logp_out = log_softmax[token].
// forced >= 0 (teacher forcing): the returned token is `forced` and *logp_out its log-probability.
__device__ __forceinline__ int warp_pick(const float (&l)[4], int lane, bool sample, float u, int V, float* logp_out,
                                         int forced = -1) {
  float mx = fmaxf(fmaxf(l[0], l[1]), fmaxf(l[2], l[3]));
  mx = warp_max(mx);
  float e[4];
  float ls = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) { e[i] = expf(l[i] - mx); ls += e[i]; }  // exp(-inf) = 0 for padding
  // inclusive scan of lane sums
  float inc = ls;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  const float total = __shfl_sync(0xffffffffu, inc, 31);
  int tok;
  if (!sample) {
    // first index attaining the maximum (torch.argmax tie rule)
    int best = 0x7fffffff;
#pragma unroll
    for (int i = 3; i >= 0; --i) if (l[i] == mx) best = lane * 4 + i;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    tok = best;
  } else {
    const float target = u * total;
    float run = inc - ls;  // exclusive prefix
    int cand = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      run += e[i];
      if (cand == 0x7fffffff && run > target && lane * 4 + i < V) cand = lane * 4 + i;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, o));
    tok = (cand == 0x7fffffff) ? V - 1 : cand;
  }
  if (forced >= 0) tok = forced;
  float lt = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) if (lane * 4 + i == tok) lt = l[i];
  lt = __shfl_sync(0xffffffffu, lt, tok >> 2);
  *logp_out = (lt - mx) - logf(total);
  return tok;
}

__device__ __forceinline__ void load_logits_row(const float* __restrict__ row, int lane, int V, float (&l)[4]) {
  const float4 r = *reinterpret_cast<const float4*>(row + lane * 4);
  l[0] = r.x; l[1] = r.y; l[2] = r.z; l[3] = r.w;
#pragma unroll
  for (int i = 0; i < 4; ++i) if (lane * 4 + i >= V) l[i] = -INFINITY;
}

// logits: [M, kVPad] fp32 (bias already added).  One warp per sequence.
__global__ void __launch_bounds__(128)
head_kernel(DecState* st, const CallParams* cp, const float* __restrict__ logits, unsigned char* fin, int M, int V,
            const float* __restrict__ emb = nullptr, const float* __restrict__ pe = nullptr, float* __restrict__ x = nullptr,
            bf16* __restrict__ xb = nullptr) {
  if (st->done) return;
  const CallParams p = *cp;
  const int n = st->step;
  const int m = blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (m < M) {
    float l[4];
    load_logits_row(logits + (size_t)m * kVPad, lane, V, l);
    if (p.logits_dump != nullptr) {
      float* d = p.logits_dump + ((size_t)m * (p.max_len - 1) + n) * V;
#pragma unroll
      for (int i = 0; i < 4; ++i) if (lane * 4 + i < V) d[lane * 4 + i] = l[i];
    }
    const bool sample = (m % p.group) < p.n_sample;
    const float u = sample ? philox_uniform(p.seed, p.seq_id_base + (unsigned long long)m, n) : 0.f;
    float lp;
    const int forced = p.forced ? clamp_token(p.ids[(size_t)m * p.max_len + n + 1], V) : -1;
    const int tok = warp_pick(l, lane, sample, u, V, &lp, forced);
    if (lane == 0) {
      if (!p.forced) p.ids[(size_t)m * p.max_len + n + 1] = tok;
      const bool ended = tok == p.end_token;
      if (sample) {
        const bool was = fin[m] != 0;
        if (p.logp != nullptr) p.logp[(size_t)m * (p.max_len - 1) + n] = was ? 0.f : lp;  // scst:239
        if (ended) fin[m] = 1;
        if (was || ended) atomicAdd(&st->cnt_finished, 1);
      } else if (ended) {
        atomicAdd(&st->cnt_end_now, 1);
      }
    }
    // with `emb`: the embedding of the token just picked = input of the next step (as vocab_head_kernel does; the
    // bf16 step's kind::tf32 vocabulary route), so that no separate embed_pe launch follows
    if (emb != nullptr && n + 1 < p.max_len - 1) {
      const float sc = 22.627416997969522f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = i * 128 + lane * 4;
        const float4 e4 = __ldg(reinterpret_cast<const float4*>(emb + (size_t)tok * kD + c));
        const float4 p4 = __ldg(reinterpret_cast<const float4*>(pe + (size_t)(n + 1) * kD + c));
        float4 o;
        o.x = __fadd_rn(__fmul_rn(e4.x, sc), p4.x); o.y = __fadd_rn(__fmul_rn(e4.y, sc), p4.y);
        o.z = __fadd_rn(__fmul_rn(e4.z, sc), p4.z); o.w = __fadd_rn(__fmul_rn(e4.w, sc), p4.w);
        *reinterpret_cast<float4*>(x + (size_t)m * kD + c) = o;
        store4(xb + (size_t)m * kD + c, o.x, o.y, o.z, o.w);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned t = atomicAdd(&st->ticket, 1u);
    if (t == gridDim.x - 1) {  // last block: evaluate the stop condition, advance the step
      __threadfence();
      const int ce = atomicAdd(&st->cnt_end_now, 0), cf = atomicAdd(&st->cnt_finished, 0);
      if (p.n_greedy_rows > 0 && ce == p.n_greedy_rows) st->greedy_done = 1;  // vit:322
      if (p.n_sample_rows > 0 && cf == p.n_sample_rows) st->sample_done = 1;  // scst:248
      st->cnt_end_now = 0; st->cnt_finished = 0; st->ticket = 0u;
      st->steps_executed = n + 1;
      st->step = n + 1;
      if (st->greedy_done && st->sample_done && !p.no_early_stop) st->done = 1;
      __threadfence();
    }
  }
}

// Fused vocabulary projection + pick + next-step embedding (SURVEY K11+K12+K13+K2): logits never
// round-trip to HBM and the step needs no separate embedding kernel.
// wT: fc_out.weight transposed and padded, [512][kVPad] fp32; x: [M,512] fp32 residual stream.
// The 256 KB of wT stream through shared memory in 64-row slabs (cp.async); BUFS of the eight slabs are requested
// before the dependency wait (the weights do not depend on the step), so that after the wait a warp only waits
// for its x rows.  Lane l accumulates vocabulary entries 4l..4l+3 of each of its warp's RW rows.
constexpr int kVhSlab = 64, kVhWarps = 4;
__host__ __device__ constexpr int vh_smem_bytes(int rows, int bufs) { return rows * kD * 4 + bufs * kVhSlab * kVPad * 4; }

// 4 warps per CTA, RW rows per warp (R = 4 RW rows per CTA).  Every k step reads one 512-byte weight row per warp
// from shared memory (4 wavefronts) whatever RW is, so one row per warp is bound by shared-memory bandwidth once
// many rows share an SM (measured: 16 warps x 1 row, 57 us at 3072 rows); RW rows per warp reuse the weight vector
// RW times (4 + RW/4 wavefronts per k for 4 RW FMA instructions).  At small batch the per-row chain of 512
// dependent steps is the critical path instead, so small batches keep RW = 1 and many CTAs.
// The accumulation order per logit (k ascending, one fmaf chain) does not depend on RW.
template <int RW, int BUFS>
__global__ void __launch_bounds__(kVhWarps * 32)
vocab_head_kernel(DecState* st, const CallParams* cp, float* __restrict__ x, const float* __restrict__ wT,
                  const float* __restrict__ bias, unsigned char* fin, int M, int V, const float* __restrict__ emb,
                  const float* __restrict__ pe, bf16* __restrict__ xb) {
  constexpr int kThreads = kVhWarps * 32;
  constexpr int R = kVhWarps * RW;
  static_assert(BUFS >= 2 && BUFS <= 6, "BUFS");
  extern __shared__ __align__(16) uint8_t vh_smem[];
  float* xs = reinterpret_cast<float*>(vh_smem);   // [R][512]
  float* ws = xs + R * kD;                         // [BUFS][64][128]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  auto load_slab = [&](int slab, int buf) {  // 64 x 128 floats = 2048 float4
    const float4* src = reinterpret_cast<const float4*>(wT + (size_t)slab * kVhSlab * kVPad);
    float4* dst = reinterpret_cast<float4*>(ws + buf * kVhSlab * kVPad);
#pragma unroll
    for (int i = 0; i < 2048 / kThreads; ++i) {
      const int idx = tid + i * kThreads;
      const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(dst + idx));
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src + idx) : "memory");
    }
  };
  // the weights do not depend on the previous kernel: start streaming them before the dependency wait
#pragma unroll
  for (int b = 0; b < BUFS; ++b) {
    load_slab(b, b);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // the call parameters are constant for the rollout and DecState was written by the previous step's head kernel
  // (a whole step upstream): read both, and the next position's encoding row, before the wait
  const CallParams p = *cp;
  const int n = st->step;
  const bool done = st->done != 0;
  const float4 bv = __ldg(reinterpret_cast<const float4*>(bias) + lane);
  float4 pe4[4];
  {
    const int pn = (n + 1 < p.max_len - 1) ? n + 1 : 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) pe4[i] = __ldg(reinterpret_cast<const float4*>(pe + (size_t)pn * kD + i * 128 + lane * 4));
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (done) { asm volatile("cp.async.wait_group 0;" ::: "memory"); return; }
  const int r0 = blockIdx.x * R;
  // x rows of this CTA -> smem (rows >= M read as zero)
  for (int i = tid; i < R * kD / 4; i += kThreads) {
    const int r = i / (kD / 4), c = i % (kD / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r0 + r < M) v = *reinterpret_cast<const float4*>(x + (size_t)(r0 + r) * kD + c * 4);
    reinterpret_cast<float4*>(xs)[i] = v;
  }
  float acc[RW][4];
#pragma unroll
  for (int j = 0; j < RW; ++j) { acc[j][0] = 0.f; acc[j][1] = 0.f; acc[j][2] = 0.f; acc[j][3] = 0.f; }
  const float* xr = xs + warp * RW * kD;
  constexpr int kSlabs = kD / kVhSlab;
  for (int sl = 0; sl < kSlabs; ++sl) {
    // groups committed so far: BUFS + sl (one per iteration, possibly empty); slabs 0..sl are needed
    asm volatile("cp.async.wait_group %0;" ::"n"(BUFS - 1) : "memory");
    __syncthreads();
    const float* wb = ws + (sl % BUFS) * kVhSlab * kVPad + lane * 4;
#pragma unroll 2
    for (int k = 0; k < kVhSlab; k += 4) {
      float av[RW][4];
#pragma unroll
      for (int j = 0; j < RW; ++j) {
        const float4 a = *reinterpret_cast<const float4*>(xr + j * kD + sl * kVhSlab + k);
        av[j][0] = a.x; av[j][1] = a.y; av[j][2] = a.z; av[j][3] = a.w;
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const float4 w = *reinterpret_cast<const float4*>(wb + (k + kk) * kVPad);
#pragma unroll
        for (int j = 0; j < RW; ++j) {
          acc[j][0] = fmaf(av[j][kk], w.x, acc[j][0]); acc[j][1] = fmaf(av[j][kk], w.y, acc[j][1]);
          acc[j][2] = fmaf(av[j][kk], w.z, acc[j][2]); acc[j][3] = fmaf(av[j][kk], w.w, acc[j][3]);
        }
      }
    }
    __syncthreads();  // every warp is done with buffer (sl % BUFS): refill it
    if (sl + BUFS < kSlabs) load_slab(sl + BUFS, sl % BUFS);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
#pragma unroll
  for (int j = 0; j < RW; ++j) {
    const int m = r0 + warp * RW + j;
    if (m < M) {  // warp-uniform
      float l[4] = {acc[j][0] + bv.x, acc[j][1] + bv.y, acc[j][2] + bv.z, acc[j][3] + bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) if (lane * 4 + i >= V) l[i] = -INFINITY;
      if (p.logits_dump != nullptr) {
        float* d = p.logits_dump + ((size_t)m * (p.max_len - 1) + n) * V;
#pragma unroll
        for (int i = 0; i < 4; ++i) if (lane * 4 + i < V) d[lane * 4 + i] = l[i];
      }
      const bool sample = (m % p.group) < p.n_sample;
      const float u = sample ? philox_uniform(p.seed, p.seq_id_base + (unsigned long long)m, n) : 0.f;
      float lp;
      const int forced = p.forced ? clamp_token(p.ids[(size_t)m * p.max_len + n + 1], V) : -1;
      const int tok = warp_pick(l, lane, sample, u, V, &lp, forced);
      if (lane == 0) {
        if (!p.forced) p.ids[(size_t)m * p.max_len + n + 1] = tok;
        const bool ended = tok == p.end_token;
        if (sample) {
          const bool was = fin[m] != 0;
          if (p.logp != nullptr) p.logp[(size_t)m * (p.max_len - 1) + n] = was ? 0.f : lp;  // scst:239
          if (ended) fin[m] = 1;
          if (was || ended) atomicAdd(&st->cnt_finished, 1);
        } else if (ended) {
          atomicAdd(&st->cnt_end_now, 1);
        }
      }
      // embedding of the token just picked = input of the next step (vit:166-169); pe row n+1 exists iff a next step does
      if (n + 1 < p.max_len - 1) {
        const float sc = 22.627416997969522f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int c = i * 128 + lane * 4;
          const float4 e4 = __ldg(reinterpret_cast<const float4*>(emb + (size_t)tok * kD + c));
          const float4 p4 = pe4[i];
          float4 o;
          o.x = __fadd_rn(__fmul_rn(e4.x, sc), p4.x); o.y = __fadd_rn(__fmul_rn(e4.y, sc), p4.y);
          o.z = __fadd_rn(__fmul_rn(e4.z, sc), p4.z); o.w = __fadd_rn(__fmul_rn(e4.w, sc), p4.w);
          *reinterpret_cast<float4*>(x + (size_t)m * kD + c) = o;
          store4(xb + (size_t)m * kD + c, o.x, o.y, o.z, o.w);
        }
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned tk = atomicAdd(&st->ticket, 1u);
    if (tk == gridDim.x - 1) {  // last block: evaluate the stop condition, advance the step
      __threadfence();
      const int ce = atomicAdd(&st->cnt_end_now, 0), cf = atomicAdd(&st->cnt_finished, 0);
      if (p.n_greedy_rows > 0 && ce == p.n_greedy_rows) st->greedy_done = 1;  // vit:322
      if (p.n_sample_rows > 0 && cf == p.n_sample_rows) st->sample_done = 1;  // scst:248
      st->cnt_end_now = 0; st->cnt_finished = 0; st->ticket = 0u;
      st->steps_executed = n + 1;
      st->step = n + 1;
      if (st->greedy_done && st->sample_done && !p.no_early_stop) st->done = 1;
      __threadfence();
    }
  }
}

// Stand-alone sampler for capdec_sample_logits: logits [rows, V] (unpadded).
__global__ void __launch_bounds__(128)
sample_rows_kernel(const float* __restrict__ logits, int rows, int V, int mode, unsigned long long seed,
                   unsigned long long seq_id_base, int step, long long* __restrict__ ids, float* __restrict__ logp) {
  const int r = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  float l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int v = lane * 4 + i;
    l[i] = v < V ? __ldg(logits + (size_t)r * V + v) : -INFINITY;
  }
  const float u = mode ? philox_uniform(seed, seq_id_base + (unsigned long long)r, step) : 0.f;
  float lp;
  const int tok = warp_pick(l, lane, mode != 0, u, V, &lp);
  if (lane == 0) { ids[r] = tok; if (logp) logp[r] = lp; }
}

template <typename Out>
__global__ void convert_kernel(const float* __restrict__ in, Out* __restrict__ out, size_t n4) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const float4 v = __ldg(reinterpret_cast<const float4*>(in) + i);
  store4(out + i * 4, v.x, v.y, v.z, v.w);
}

// Encoder hand-off: rows first..first+mem_len-1 of every image's [tokens_per_image, dim] fp32 token block
// (the ViT encoder output minus its class token, vit:95) -> contiguous [n_images*mem_len, dim] GEMM operand.
template <typename Out>
__global__ void gather_tokens_kernel(const float* __restrict__ in, Out* __restrict__ out, int n_images, int tokens_per_image,
                                     int first, int mem_len, int dim4) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_images * mem_len * dim4) return;
  const int c = (int)(i % dim4);
  const size_t r = i / dim4;
  const int img = (int)(r / mem_len), s = (int)(r % mem_len);
  const float4 v = __ldg(reinterpret_cast<const float4*>(in) + ((size_t)img * tokens_per_image + first + s) * dim4 + c);
  store4(out + i * 4, v.x, v.y, v.z, v.w);
}

// pack rows [r0, r0+rows) of a [*, cols] fp32 matrix into out (Out), optional zero row padding
template <typename Out>
__global__ void pack_rows_kernel(const float* __restrict__ in, Out* __restrict__ out, int rows, int cols, int rows_out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)rows_out * cols) return;
  const int r = (int)(i / cols);
  out[i] = from_f32<Out>(r < rows ? in[i] : 0.f);
}

// out[c][r] = in[r][c] for r < rows (else 0): fc_out.weight [V,512] -> [512][kVPad]
__global__ void transpose_pad_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols, int rows_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cols * rows_out) return;
  const int c = i / rows_out, r = i % rows_out;
  out[i] = r < rows ? in[(size_t)r * cols + c] : 0.f;
}

// diagnostics: cache tile [n_images][H][S][64] -> fp32 [n_images, S, 512]
template <typename Elem>
__global__ void gather_cross_kernel(const Elem* __restrict__ src, float* __restrict__ out, int n_images, int S) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_images * S * kD) return;
  const int d = (int)(i % kD);
  const size_t r = i / kD;
  const int s = (int)(r % S), img = (int)(r / S);
  out[i] = to_f32(src[(((size_t)img * kH + (d >> 6)) * S + s) * kHD + (d & 63)]);
}

}  // namespace capdec
