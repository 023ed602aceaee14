"""Random-init decoder weights and synthetic encoder memory of the reference's shapes, for the
benchmarks and examples (there are no trained checkpoints: the shipped ``.pth`` files are LFS stubs).

``random_decoder`` returns an ``nn.Module`` that is ONLY a weight container with the reference
decoder's state-dict keys (models/vit_transformer_model.py:111-139: ``embedding``,
``pos_encoder.pe``, ``transformer_decoder.layers.{i}.*``, ``fc_out``) - exactly what
``CaptionDecodeEngine`` reads from a live reference model.  It has no forward(): decoding is
the engine's job.
"""
from __future__ import annotations

import math

import torch
from torch import nn


class _PE(nn.Module):
    def __init__(self, d_model: int, rows: int):
        super().__init__()
        pos = torch.arange(rows, dtype=torch.float32)[:, None]
        w = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32) * (-math.log(10000.0) / d_model))
        t = torch.zeros(rows, d_model)
        t[:, 0::2], t[:, 1::2] = torch.sin(pos * w), torch.cos(pos * w)
        self.register_buffer("pe", t[None])


class DecoderWeights(nn.Module):
    def __init__(self, vocab: int = 109, d_model: int = 512, nhead: int = 8, layers: int = 6, d_ff: int = 2048,
                 pe_rows: int = 52):
        super().__init__()
        self.embedding = nn.Embedding(vocab, d_model)
        self.pos_encoder = _PE(d_model, pe_rows)
        layer = nn.TransformerDecoderLayer(d_model, nhead, d_ff, 0.1, batch_first=True)
        self.transformer_decoder = nn.TransformerDecoder(layer, layers)
        self.fc_out = nn.Linear(d_model, vocab)
        with torch.no_grad():
            self.embedding.weight.uniform_(-0.1, 0.1)
            self.fc_out.weight.uniform_(-0.1, 0.1)
            self.fc_out.bias.zero_()
        self.eval()


def random_decoder(seed: int = 0, pe_rows: int = 52, layers: int = 6, distinct_layers: bool = True) -> DecoderWeights:
    """``torch.manual_seed(seed)`` then the reference constructor order.  torch clones
    one layer ``layers`` times, so ``distinct_layers`` adds small seeded noise to tell them apart."""
    torch.manual_seed(seed)
    dec = DecoderWeights(pe_rows=pe_rows, layers=layers)
    if distinct_layers:
        g = torch.Generator().manual_seed(seed + 7919)
        with torch.no_grad():
            for p in dec.parameters():
                p.add_(torch.randn(p.shape, generator=g) * 0.02)
    return dec


def random_memory(n_images: int, mem_len: int, seed: int = 1234, d_model: int = 512) -> torch.Tensor:
    """``randn(B, S, 512)`` float32 from ``torch.Generator().manual_seed(1234)``."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n_images, mem_len, d_model, generator=g)
