"""The reference decoder re-assembled from torch.nn building blocks (oracle - test
infrastructure only; also the timed CPU baseline of bench.py).

``/root/reference`` does not exist on the GPU box, and reference sources may not be copied,
so this port constructs the SAME third-party modules the reference constructs, in the same
order (vit:111-147: Embedding, sinusoid buffer, one nn.TransformerDecoderLayer cloned
``num_layers`` times by nn.TransformerDecoder, Linear, then the U(-0.1,0.1) re-init of
vit:142-147), so that under the same ``torch.manual_seed`` it yields bit-identical
parameters with identical state-dict keys.  ``tests/golden/make_golden.py`` asserts both claims
against the real reference when it generates the fixtures (it needs ``/root/reference``), and
``tests/test_oracle_golden.py`` re-checks them everywhere through the per-tensor checksums and
the reference outputs stored in ``tests/golden/*.npz``.

The loops below restate vit:296-325 (greedy) and scst:202-254 (sampling) in the reference's
own no-KV-cache form: the whole prefix is re-decoded at every step.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np
import torch
from torch import nn


class _SinusoidTable(nn.Module):
    """Holds the ``pe`` buffer (1, max_len, d) of vit:18-26 under the key ``pos_encoder.pe``."""

    def __init__(self, d_model: int, max_len: int):
        super().__init__()
        pos = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
        freq = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
        table = torch.zeros(max_len, d_model)
        table[:, 0::2] = torch.sin(pos * freq)
        table[:, 1::2] = torch.cos(pos * freq)
        self.register_buffer("pe", table.unsqueeze(0))


class PortDecoder(nn.Module):
    """Same parameters, same keys, same forward function as the reference TransformerDecoder
    (vit:103-182, grid:113-158).  Dropout is not applied: the port is an eval()-mode oracle."""

    def __init__(self, vocab_size: int, d_model: int = 512, nhead: int = 8, num_layers: int = 6,
                 dim_feedforward: int = 2048, dropout: float = 0.1, max_len: int = 100):
        super().__init__()
        self.d_model, self.vocab_size, self.nhead = d_model, vocab_size, nhead
        self.embedding = nn.Embedding(vocab_size, d_model)
        self.pos_encoder = _SinusoidTable(d_model, max_len)
        proto = nn.TransformerDecoderLayer(d_model=d_model, nhead=nhead, dim_feedforward=dim_feedforward,
                                           dropout=dropout, batch_first=True)
        self.transformer_decoder = nn.TransformerDecoder(proto, num_layers=num_layers)
        self.fc_out = nn.Linear(d_model, vocab_size)
        with torch.no_grad():  # vit:142-147
            self.embedding.weight.uniform_(-0.1, 0.1)
            self.fc_out.bias.zero_()
            self.fc_out.weight.uniform_(-0.1, 0.1)
        self.eval()

    @staticmethod
    def causal_mask(t: int, device=None) -> torch.Tensor:
        """vit:149-153."""
        return torch.full((t, t), float("-inf"), device=device).triu(1)

    def forward(self, tgt, memory, tgt_mask=None, tgt_key_padding_mask=None, memory_key_padding_mask=None):
        t = tgt.size(1)
        x = self.embedding(tgt) * math.sqrt(self.d_model) + self.pos_encoder.pe[:, :t]
        x = self.transformer_decoder(x, memory, tgt_mask=tgt_mask,
                                     tgt_key_padding_mask=tgt_key_padding_mask,
                                     memory_key_padding_mask=memory_key_padding_mask)
        return self.fc_out(x)


def build_decoder(seed: int = 0, vocab_size: int = 109, d_model: int = 512, nhead: int = 8,
                  num_layers: int = 6, dim_feedforward: int = 2048, max_len: int = 52,
                  perturb: float = 0.0) -> PortDecoder:
    """Random-init decoder as SURVEY section 8d specifies (``torch.manual_seed(seed)`` then the
    reference constructor order).  nn.TransformerDecoder deep-copies one layer, so at init all
    layers are identical; ``perturb > 0`` adds N(0, perturb^2) noise to every parameter from a
    second seeded generator so that tests can tell layers (and LayerNorm gains/biases) apart.
    """
    torch.manual_seed(seed)
    dec = PortDecoder(vocab_size, d_model, nhead, num_layers, dim_feedforward, 0.1, max_len)
    if perturb > 0:
        g = torch.Generator().manual_seed(seed + 7919)
        with torch.no_grad():
            for p in dec.parameters():
                p.add_(torch.randn(p.shape, generator=g) * perturb)
    return dec


def synthetic_memory(n_images: int, mem_len: int, d_model: int = 512, seed: int = 1234) -> torch.Tensor:
    """``torch.Generator().manual_seed(1234)``; randn(B,S,512) float32."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n_images, mem_len, d_model, generator=g)


def weights_to_numpy(dec: nn.Module) -> Dict[str, np.ndarray]:
    return {k: v.detach().cpu().numpy() for k, v in dec.state_dict().items()}


@torch.no_grad()
def greedy_recompute(dec: nn.Module, memory: torch.Tensor, start_token: int, end_token: int,
                     max_len: int = 50) -> torch.Tensor:
    """vit:307-323 on an already-encoded memory: no KV cache, one full decoder call per step."""
    seq = torch.full((memory.size(0), 1), start_token, dtype=torch.long, device=memory.device)
    for _ in range(max_len - 1):
        mask = PortDecoder.causal_mask(seq.size(1), memory.device)
        word = dec(seq, memory, tgt_mask=mask)[:, -1].argmax(-1)
        seq = torch.cat([seq, word[:, None]], 1)
        if bool((word == end_token).all()):
            break
    return seq


@torch.no_grad()
def sample_recompute(dec: nn.Module, memory: torch.Tensor, start_token: int, end_token: int,
                     max_len: int = 50, generator=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """scst:216-252 on an already-encoded memory (eval mode, torch.multinomial draws)."""
    b = memory.size(0)
    seq = torch.full((b, 1), start_token, dtype=torch.long, device=memory.device)
    done = torch.zeros(b, dtype=torch.bool, device=memory.device)
    lps = []
    for _ in range(max_len - 1):
        mask = PortDecoder.causal_mask(seq.size(1), memory.device)
        logits = dec(seq, memory, tgt_mask=mask)[:, -1]
        word = torch.multinomial(logits.softmax(-1), 1, generator=generator).squeeze(1)
        lp = logits.log_softmax(-1).gather(1, word[:, None]).squeeze(1).masked_fill(done, 0.0)
        lps.append(lp)
        seq = torch.cat([seq, word[:, None]], 1)
        done = done | (word == end_token)
        if bool(done.all()):
            break
    return seq, torch.stack(lps, 1)


@torch.no_grad()
def teacher_forced_logits(dec: nn.Module, ids: torch.Tensor, memory: torch.Tensor) -> torch.Tensor:
    """One causal pass over ``ids[:, :-1]`` (the training forward, vit:233-253): row k equals the
    step-k logits of the incremental loops."""
    tgt = ids[:, :-1]
    return dec(tgt, memory, tgt_mask=PortDecoder.causal_mask(tgt.size(1), memory.device))
