"""Device-side stand-in for the reference's ``CiderRewardCalculator`` (utils/scst_loss.py:20-54).

``DeviceCiderReward(engine).compute_reward(predictions, references)`` takes what the reference's calculator takes -
``predictions``: list of str, ``references``: list of (list of str | str) - and returns a list of floats, computed by
``capdec_cider_d`` on token ids.  Words are mapped to integers by a dictionary built per call (any injective map gives
the same n-gram statistics).  The arithmetic follows the published CIDEr-D algorithm (oracle/cider_d.py); the
reference's own scorer, pycocoevalcap, is not vendored and cannot be installed here, so parity with it is unpinned.
``compute_reward_from_ids`` skips the strings: SCST can hand over the rollout's ids and the dataset's caption ids.
"""
from __future__ import annotations

from typing import List, Sequence

import torch


class DeviceCiderReward:
    PAD, END = 0, 1  # ids of the per-call dictionary: 0 pads the rows, 1 ends a sentence; words start at 2

    def __init__(self, engine):
        self.engine = engine

    def compute_reward(self, predictions: Sequence[str], references: Sequence) -> List[float]:
        vocab = {}

        def ids(sentence: str) -> List[int]:
            return [vocab.setdefault(w, len(vocab) + 2) for w in sentence.split()] + [self.END]

        hyp = [ids(p) for p in predictions]
        ref_sets = [[ids(r) for r in (rs if isinstance(rs, list) else [rs])] for rs in references]
        if len(vocab) + 2 >= 32768:
            raise ValueError("more than 32765 distinct words in one reward call")
        offsets = [0]
        for rs in ref_sets:
            offsets.append(offsets[-1] + len(rs))
        flat = [r for rs in ref_sets for r in rs]

        def pack(rows: List[List[int]]) -> torch.Tensor:
            width = max(1, max((len(r) for r in rows), default=1))
            t = torch.full((max(1, len(rows)), width), self.PAD, dtype=torch.int64)
            for i, r in enumerate(rows):
                t[i, :len(r)] = torch.tensor(r, dtype=torch.int64)
            return t[:len(rows)] if rows else t[:0]

        scores = self.engine.cider_d(pack(hyp), pack(flat), offsets, group=1, start_token=-1, end_token=self.END,
                                     pad_token=self.PAD)
        return scores.cpu().tolist()

    def compute_reward_from_ids(self, hyp_ids: torch.Tensor, ref_ids: torch.Tensor, ref_offsets, group: int,
                                start_token: int, end_token: int, pad_token: int) -> torch.Tensor:
        return self.engine.cider_d(hyp_ids, ref_ids, ref_offsets, group, start_token, end_token, pad_token)
