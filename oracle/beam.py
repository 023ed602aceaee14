"""CPU restatement of the reference's beam search (TEST INFRASTRUCTURE ONLY: imported by tests/ and smoke()).

Follows ``ViTTransformerCaptioning._beam_search`` (models/vit_transformer_model.py:327-420) and
``GridTransformerCaptioning._beam_search`` (models/grid_transformer_model.py:253-322): one image at a time, the
decoder re-run over the whole prefix every step, scores = sums of log-softmax values without length normalisation,
the beam shrinking as hypotheses emit <end>.  The two models differ in one stop rule (``stop_rule``):

* ``"vit"``:  stop when every hypothesis selected at a step ended (vit:402);
* ``"grid"``: stop when ``len(completed) >= beam_size`` with the beam width of that moment (grid:308).

Pinned by ``tests/golden/beam_*.npz`` (outputs of the real reference, ``tests/golden/make_golden_beam.py``).
Unlike the reference it also returns the winning score, and it keeps the per-step candidate gap so that parity
tests can tell a genuine disagreement from a near-tie flipped by reduced precision.
"""
from typing import List, Tuple

import torch


def beam_search_one(dec: torch.nn.Module, memory: torch.Tensor, start_token: int, end_token: int, max_len: int,
                    beam_size: int = 5, stop_rule: str = "vit") -> Tuple[torch.Tensor, float, float]:
    """``memory``: (1, S, D).  Returns (sequence 1-D int64, score, smallest gap between the last selected and the
    first rejected candidate over all steps)."""
    assert memory.shape[0] == 1 and stop_rule in ("vit", "grid")
    vocab = dec.fc_out.out_features
    sequences = torch.full((beam_size, 1), start_token, dtype=torch.long)
    scores = torch.zeros(beam_size)
    mem = memory.expand(beam_size, -1, -1)
    completed: List[torch.Tensor] = []
    completed_scores: List[float] = []
    min_gap = float("inf")
    with torch.no_grad():
        for step in range(max_len - 1):
            if sequences.size(0) == 0:  # grid:277
                break
            t = sequences.size(1)
            mask = torch.triu(torch.full((t, t), float("-inf")), diagonal=1)  # vit:149-153
            logp = torch.log_softmax(dec(sequences, mem, tgt_mask=mask)[:, -1, :], dim=-1)
            if step == 0:  # vit:367-375: all hypotheses are identical, expand the first
                cand = logp[0]
                parents = torch.zeros(beam_size, dtype=torch.long)
            else:          # vit:377-392
                cand = (scores.unsqueeze(1) + logp).view(-1)
                parents = None
            k = beam_size
            srt, order = torch.sort(cand, descending=True, stable=True)  # ties: lower flat index first
            top_scores, top_idx = srt[:k], order[:k]
            if cand.numel() > k:
                min_gap = min(min_gap, float(srt[k - 1] - srt[k]))
            if parents is None:
                parents = top_idx // vocab
                words = top_idx % vocab
            else:
                words = top_idx
            sequences = torch.cat([sequences[parents], words.unsqueeze(1)], dim=1)
            scores = top_scores
            ended = sequences[:, -1] == end_token
            if ended.any():
                for idx in ended.nonzero(as_tuple=True)[0]:  # vit:396-399
                    completed.append(sequences[idx])
                    completed_scores.append(float(scores[idx]))
                if stop_rule == "vit" and bool(ended.all()):                 # vit:402
                    break
                if stop_rule == "grid" and len(completed) >= beam_size:       # grid:308
                    break
                keep = ~ended
                sequences, scores, mem = sequences[keep], scores[keep], mem[keep]
                if sequences.size(0) == 0:                                    # grid:315
                    break
                beam_size = sequences.size(0)
    if completed:  # vit:412-416
        best = int(torch.tensor(completed_scores).argmax())
        return completed[best], completed_scores[best], min_gap
    best = int(scores.argmax())
    return sequences[best], float(scores[best]), min_gap


def beam_search(dec: torch.nn.Module, memory: torch.Tensor, start_token: int, end_token: int, max_len: int,
                beam_size: int = 5, stop_rule: str = "vit", pad_token: int = 0):
    """Every image of ``memory`` (B, S, D) in turn (vit:333-341).  Returns (ids (B, max_len) padded, lengths,
    scores, gaps)."""
    n = memory.shape[0]
    ids = torch.full((n, max_len), pad_token, dtype=torch.long)
    lens = torch.zeros(n, dtype=torch.int32)
    scores = torch.zeros(n)
    gaps = torch.zeros(n)
    for i in range(n):
        seq, sc, gap = beam_search_one(dec, memory[i:i + 1], start_token, end_token, max_len, beam_size, stop_rule)
        ids[i, :seq.numel()] = seq
        lens[i] = seq.numel()
        scores[i] = sc
        gaps[i] = gap
    return ids, lens, scores, gaps
