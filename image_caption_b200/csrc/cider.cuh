// cider.cuh - CIDEr-D reward of an SCST step on the device.
//
// The reference computes the reward on the CPU with pycocoevalcap's Cider() on strings (utils/scst_loss.py:20-54, called at
// scst:179-180 after ids -> words at scst:256-269); once the rollout takes milliseconds that Python loop is the SCST
// step's bottleneck.  pycocoevalcap is NOT vendored in the reference and not installable here, so this follows the
// published CIDEr-D algorithm (Vedantam et al. 2015; pycocoevalcap/cider/cider_scorer.py: n = 4, sigma = 6, clipped
// tf-idf cosine, x10) as restated in oracle/cider_d.py -- parity is against that restatement only ("unpinned").
//
// Everything is integer n-gram bookkeeping on token ids (a word <-> id map is injective, so n-grams of ids == n-grams of
// words), HBM/latency-bound and tiny: no tensor cores.  An n-gram of n <= 4 tokens (< 32768 each) is ONE 64-bit key
// {n-1 : 2 bits | 4 x 15 bits}; a sequence's n-grams are sorted in shared memory (bitonic), run-length counted, and
//   cider_df_kernel     one CTA per image: distinct n-grams over all its references -> +1 in a global open-addressing table
//   cider_score_kernel  one CTA per hypothesis: tf-idf vectors of the hypothesis and of each reference of its image,
//                       clipped dot products by binary search, Gaussian length penalty, mean over n, /#refs, x10
// Arithmetic in fp64 like the reference's Python floats (a few thousand operations per caption).
#pragma once
#include "common.cuh"

namespace capdec {

constexpr int kCdMaxTok = 128;                 // tokens of one sequence that are looked at
constexpr int kCdSeqKeys = 512;                // >= 4 * kCdMaxTok - 6: n-grams of one sequence (bitonic size)
constexpr int kCdImgKeys = 4096;               // n-grams of all references of one image (bitonic size)
constexpr int kCdThreads = 256;
constexpr unsigned long long kCdEmpty = ~0ull;

struct CiderArgs {
  const long long* hyp;      // [n_images * group, hyp_len] hypotheses, image-major
  const long long* refs;     // [n_refs, ref_len] reference captions as token ids
  const int* ref_offsets;    // [n_images + 1]: references of image i are rows ref_offsets[i] .. ref_offsets[i+1]-1
  int n_images, group, hyp_len, ref_len;
  int start_token, end_token, pad_token;
  unsigned long long* tab_keys;  // [tab_size] open-addressing table of n-gram keys (kCdEmpty = free)
  unsigned int* tab_cnt;         // [tab_size] document frequency
  unsigned int tab_mask;         // tab_size - 1 (power of two)
  double* scores;            // [n_images * group]
};

// scst:256-269: stop at the first <end>, skip <start> and <pad>.  One thread; returns the number of words kept.
__device__ inline int cd_tokens(const long long* seq, int len, int start, int end, int pad, int* tok) {
  int n = 0;
  for (int i = 0; i < len && n < kCdMaxTok; ++i) {
    const long long t = seq[i];
    if (t == end) break;
    if (t != start && t != pad) tok[n++] = (int)(t & 0x7fff);
  }
  return n;
}
__device__ __forceinline__ unsigned long long cd_key(const int* tok, int i, int n) {  // n-gram of n tokens starting at i
  unsigned long long k = (unsigned long long)(n - 1) << 60;
  for (int j = 0; j < n; ++j) k |= (unsigned long long)tok[i + j] << (45 - 15 * j);
  return k;
}
// all n-grams (n = 1..4) of tok[0..cnt) appended to keys[base..]; every thread of the CTA takes part; returns their number
__device__ inline int cd_ngrams(const int* tok, int cnt, unsigned long long* keys, int base) {
  int total = 0;
  for (int n = 1; n <= 4; ++n) {
    const int m = cnt - n + 1;
    if (m <= 0) break;
    for (int i = threadIdx.x; i < m; i += blockDim.x) keys[base + total + i] = cd_key(tok, i, n);
    total += m;
  }
  return total;
}
// ascending bitonic sort of keys[0..size) (size a power of two, padded with kCdEmpty); whole CTA
__device__ inline void cd_sort(unsigned long long* keys, int size) {
  for (int k = 2; k <= size; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < size; i += blockDim.x) {
        const int p = i ^ j;
        if (p > i) {
          const unsigned long long a = keys[i], b = keys[p];
          if (((i & k) == 0) == (a > b)) { keys[i] = b; keys[p] = a; }
        }
      }
    }
  __syncthreads();
}
// sorted keys[0..n) -> distinct keys in ukeys[0..nu) with their term frequencies; returns nu.  Thread 0 walks the sorted
// list once (n <= 506: a microsecond, off the rollout's critical path); every thread of the CTA must call it.
__device__ inline int cd_unique(const unsigned long long* keys, int n, unsigned long long* ukeys, int* tf, int* counter) {
  if (threadIdx.x == 0) {
    int u = -1;
    for (int i = 0; i < n; ++i) {
      if (i == 0 || keys[i] != keys[i - 1]) { ++u; ukeys[u] = keys[i]; tf[u] = 1; }
      else tf[u] += 1;
    }
    *counter = u + 1;
  }
  __syncthreads();
  return *counter;
}
__device__ __forceinline__ unsigned int cd_hash(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return (unsigned int)k;
}

// grid = n_images.  Document frequency (cider_scorer.compute_doc_freq): +1 per image for every n-gram that occurs in
// any of its references.
__global__ void __launch_bounds__(kCdThreads)
cider_df_kernel(const __grid_constant__ CiderArgs A) {
  __shared__ unsigned long long keys[kCdImgKeys];
  __shared__ int tok[kCdMaxTok];
  __shared__ int s_cnt, s_total;
  const int img = blockIdx.x;
  for (int i = threadIdx.x; i < kCdImgKeys; i += blockDim.x) keys[i] = kCdEmpty;
  if (threadIdx.x == 0) s_total = 0;
  __syncthreads();
  for (int r = A.ref_offsets[img]; r < A.ref_offsets[img + 1]; ++r) {
    if (threadIdx.x == 0) s_cnt = cd_tokens(A.refs + (size_t)r * A.ref_len, A.ref_len, A.start_token, A.end_token, A.pad_token, tok);
    __syncthreads();
    const int base = s_total;
    const int m = s_cnt >= 4 ? 4 * s_cnt - 6 : s_cnt * (s_cnt + 1) / 2;  // this reference's n-grams
    if (base + m <= kCdImgKeys) {  // (the host checks the capacity; a reference that would not fit is skipped, never overrun)
      const int got = cd_ngrams(tok, s_cnt, keys, base);
      __syncthreads();
      if (threadIdx.x == 0) s_total = base + got;
    }
    __syncthreads();
  }
  cd_sort(keys, kCdImgKeys);
  const int n = s_total;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (i == 0 || keys[i] != keys[i - 1]) {  // one insert per distinct n-gram of this image
      const unsigned long long k = keys[i];
      unsigned int slot = cd_hash(k) & A.tab_mask;
      for (;;) {
        const unsigned long long prev = atomicCAS(&A.tab_keys[slot], kCdEmpty, k);
        if (prev == kCdEmpty || prev == k) { atomicAdd(&A.tab_cnt[slot], 1u); break; }
        slot = (slot + 1) & A.tab_mask;
      }
    }
  }
}

// tf-idf vector of one sorted, run-length-counted sequence: vec = tf * (log N - log max(1, df)) (cider_scorer.counts2vec)
__device__ inline void cd_vector(const CiderArgs& A, const unsigned long long* ukeys, const int* tf, int nu, double log_n, double* vec,
                                 double* norm2 /*[4]*/, int* length /*bigram count*/) {
  for (int i = threadIdx.x; i < nu; i += blockDim.x) {
    const unsigned long long k = ukeys[i];
    unsigned int slot = cd_hash(k) & A.tab_mask, df = 0;
    for (;;) {
      const unsigned long long cur = A.tab_keys[slot];
      if (cur == k) { df = A.tab_cnt[slot]; break; }
      if (cur == kCdEmpty) break;
      slot = (slot + 1) & A.tab_mask;
    }
    const double v = (double)tf[i] * (log_n - log(df > 1u ? (double)df : 1.0));
    vec[i] = v;
    const int n = (int)(k >> 60);
    atomicAdd(&norm2[n], v * v);
    if (n == 1) atomicAdd(length, tf[i]);
  }
}

// grid = n_images * group.  cider_scorer.compute_cider for one hypothesis.
__global__ void __launch_bounds__(kCdThreads)
cider_score_kernel(const __grid_constant__ CiderArgs A) {
  __shared__ unsigned long long keys[kCdSeqKeys], hk[kCdSeqKeys], rk[kCdSeqKeys];
  __shared__ int htf[kCdSeqKeys], rtf[kCdSeqKeys], tok[kCdMaxTok];
  __shared__ double hv[kCdSeqKeys], rv[kCdSeqKeys];
  __shared__ double hnorm2[4], rnorm2[4], val[4], score[4];
  __shared__ int s_cnt, s_counter, hlen, rlen;
  const int m = blockIdx.x, img = m / A.group;
  const double log_n = log((double)A.n_images);  // ref_len = log(number of images in the call)
  auto load = [&](const long long* seq, int len, unsigned long long* uk, int* tf) -> int {
    for (int i = threadIdx.x; i < kCdSeqKeys; i += blockDim.x) keys[i] = kCdEmpty;
    if (threadIdx.x == 0) s_cnt = cd_tokens(seq, len, A.start_token, A.end_token, A.pad_token, tok);
    __syncthreads();
    const int n = cd_ngrams(tok, s_cnt, keys, 0);
    cd_sort(keys, kCdSeqKeys);
    return cd_unique(keys, n, uk, tf, &s_counter);
  };
  const int nh = load(A.hyp + (size_t)m * A.hyp_len, A.hyp_len, hk, htf);
  if (threadIdx.x < 4) { hnorm2[threadIdx.x] = 0.0; score[threadIdx.x] = 0.0; }
  if (threadIdx.x == 0) hlen = 0;
  __syncthreads();
  cd_vector(A, hk, htf, nh, log_n, hv, hnorm2, &hlen);
  __syncthreads();
  const int r0 = A.ref_offsets[img], r1 = A.ref_offsets[img + 1];
  for (int r = r0; r < r1; ++r) {
    const int nr = load(A.refs + (size_t)r * A.ref_len, A.ref_len, rk, rtf);
    if (threadIdx.x < 4) { rnorm2[threadIdx.x] = 0.0; val[threadIdx.x] = 0.0; }
    if (threadIdx.x == 0) rlen = 0;
    __syncthreads();
    cd_vector(A, rk, rtf, nr, log_n, rv, rnorm2, &rlen);
    __syncthreads();
    for (int i = threadIdx.x; i < nh; i += blockDim.x) {  // cider_scorer.sim: sum over the hypothesis' n-grams found in the reference
      const unsigned long long k = hk[i];
      int lo = 0, hi = nr;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (rk[mid] < k) lo = mid + 1; else hi = mid; }
      if (lo < nr && rk[lo] == k) atomicAdd(&val[(int)(k >> 60)], fmin(hv[i], rv[lo]) * rv[lo]);
    }
    __syncthreads();
    if (threadIdx.x < 4) {
      const int n = threadIdx.x;
      double v = val[n];
      const double nh_ = sqrt(hnorm2[n]), nr_ = sqrt(rnorm2[n]);
      if (nh_ != 0.0 && nr_ != 0.0) v /= nh_ * nr_;
      const double delta = (double)(hlen - rlen);
      v *= exp(-(delta * delta) / (2.0 * 6.0 * 6.0));
      score[n] += v;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const int nrefs = r1 - r0;
    double s = (score[0] + score[1] + score[2] + score[3]) / 4.0;
    s = nrefs > 0 ? s / (double)nrefs : 0.0;
    A.scores[m] = s * 10.0;
  }
}

}  // namespace capdec
