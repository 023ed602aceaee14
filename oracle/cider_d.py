"""CIDEr-D restated (oracle - test infrastructure only; PARITY UNPINNED).

The reference computes SCST rewards with ``pycocoevalcap.cider.cider.Cider().compute_score(gts, res)``
(utils/scst_loss.py:27, :50) on strings produced by ``_decode_captions`` (scst:256-269).  pycocoevalcap is a
third-party dependency that the reference neither vendors nor pins (no requirements file names a version) and that
cannot be installed here (no network), and the reference ships no golden reward values.  This file therefore
restates the published algorithm - Vedantam, Zitnick, Parikh, "CIDEr: Consensus-based Image Description Evaluation"
(CVPR 2015), in the CIDEr-D form of pycocoevalcap/cider/cider_scorer.py - and the device kernel (csrc/cider.cuh) is
checked against THIS restatement only:

  * n-grams of n = 1..4 words, term frequency per sentence (``precook``);
  * document frequency = number of images whose reference set contains the n-gram (``compute_doc_freq``);
  * tf-idf weight  tf * (log N - log max(1, df)),  N = number of images in the call (``counts2vec``; the scorer's
    corpus is the call's own references, which is how scst:50 uses it: one call per batch);
  * per n: sum over the hypothesis' n-grams of min(w_hyp, w_ref) * w_ref, divided by the two vector norms when both are
    non-zero, times exp(-(l_hyp - l_ref)^2 / (2 sigma^2)), sigma = 6, where l is the number of BIGRAMS of the
    sentence (the scorer accumulates ``length`` only for n == 2) (``sim``);
  * score = 10 * mean over n of the sum over references, divided by the number of references (``compute_cider``).
Python floats (fp64) throughout, as in the scorer.
"""
from __future__ import annotations

import math
from collections import defaultdict
from typing import Dict, List, Sequence, Tuple

N_MAX, SIGMA = 4, 6.0


def words_from_ids(ids: Sequence[int], start: int, end: int, pad: int) -> List[int]:
    """scst:256-269 without the id -> word lookup: stop at the first <end>, skip <start> and <pad>."""
    out = []
    for t in ids:
        t = int(t)
        if t == end:
            break
        if t != start and t != pad:
            out.append(t)
    return out


def precook(words: Sequence) -> Dict[Tuple, int]:
    counts: Dict[Tuple, int] = defaultdict(int)
    for n in range(1, N_MAX + 1):
        for i in range(len(words) - n + 1):
            counts[tuple(words[i:i + n])] += 1
    return counts


def cider_d(hyps: List[Sequence], refs: List[List[Sequence]]) -> List[float]:
    """``hyps[i]`` = word list of prediction i, ``refs[i]`` = list of word lists of its references.
    Returns what ``Cider().compute_score`` returns per prediction (its second value)."""
    ctest = [precook(h) for h in hyps]
    crefs = [[precook(r) for r in rs] for rs in refs]
    df: Dict[Tuple, float] = defaultdict(float)
    for rs in crefs:
        for ng in set(ng for r in rs for ng in r):
            df[ng] += 1
    log_n = math.log(float(len(crefs)))

    def counts2vec(cnts):
        vec = [defaultdict(float) for _ in range(N_MAX)]
        norm = [0.0] * N_MAX
        length = 0
        for ng, tf in cnts.items():
            d = math.log(max(1.0, df[ng]))
            n = len(ng) - 1
            vec[n][ng] = float(tf) * (log_n - d)
            norm[n] += vec[n][ng] ** 2
            if n == 1:
                length += tf
        return vec, [math.sqrt(x) for x in norm], length

    def sim(vh, vr, nh, nr, lh, lr):
        delta = float(lh - lr)
        val = [0.0] * N_MAX
        for n in range(N_MAX):
            for ng in vh[n]:
                val[n] += min(vh[n][ng], vr[n][ng]) * vr[n][ng]
            if nh[n] != 0 and nr[n] != 0:
                val[n] /= nh[n] * nr[n]
            val[n] *= math.e ** (-(delta ** 2) / (2 * SIGMA ** 2))
        return val

    scores = []
    for test, rs in zip(ctest, crefs):
        vec, norm, length = counts2vec(test)
        score = [0.0] * N_MAX
        for r in rs:
            vr, nr, lr = counts2vec(r)
            s = sim(vec, vr, norm, nr, length, lr)
            score = [a + b for a, b in zip(score, s)]
        avg = sum(score) / N_MAX
        avg = avg / len(rs) if rs else 0.0
        scores.append(avg * 10.0)
    return scores


def cider_d_from_ids(hyp_ids, ref_ids, ref_offsets, group: int, start: int, end: int, pad: int) -> List[float]:
    """Same call shape as the C ABI: ``hyp_ids`` (n_images*group, L) image-major, the references of image i are rows
    ``ref_offsets[i]:ref_offsets[i+1]`` of ``ref_ids``.  Every hypothesis of an image is scored against that image's
    references with the document frequencies of the call's n_images reference sets."""
    n_images = len(ref_offsets) - 1
    ref_sets = [[words_from_ids(ref_ids[r], start, end, pad) for r in range(ref_offsets[i], ref_offsets[i + 1])]
                for i in range(n_images)]
    out = []
    for g in range(group):  # one scorer call per rollout slot, as SCSTLoss calls it once per rollout (scst:179-180)
        hyps = [words_from_ids(hyp_ids[i * group + g], start, end, pad) for i in range(n_images)]
        out.append(cider_d(hyps, ref_sets))
    return [out[m % group][m // group] for m in range(n_images * group)]
