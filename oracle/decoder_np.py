"""numpy restatement of the reference caption decoder (oracle - test infrastructure only).

Every function cites the reference site it restates.  ``vit:`` means
``models/vit_transformer_model.py``, ``grid:`` ``models/grid_transformer_model.py``
and ``scst:`` ``utils/scst_loss.py`` of wonder-dream/image_caption.  The layer
arithmetic itself lives in third-party PyTorch (``torch.nn.TransformerDecoderLayer``
with ``batch_first=True, norm_first=False, activation=relu``, LayerNorm eps 1e-5,
``F.multi_head_attention_forward`` with packed in_proj), de-facto pinned to the
installed torch 2.11.0; it is restated here from its documented definition and
pinned by ``tests/golden`` (outputs of the real reference).

All arithmetic is float32, as in the reference.

Weights are a plain ``dict[str, np.ndarray]`` keyed by the decoder's state-dict
names (``embedding.weight``, ``pos_encoder.pe``,
``transformer_decoder.layers.{i}.self_attn.in_proj_weight`` ... ``fc_out.bias``).
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Optional, Tuple

import numpy as np

Weights = Dict[str, np.ndarray]
F32 = np.float32


# --------------------------------------------------------------------------- helpers
def n_layers(w: Weights) -> int:
    n = 0
    while f"transformer_decoder.layers.{n}.linear1.weight" in w:
        n += 1
    return n


def positional_table(max_len: int, d_model: int) -> np.ndarray:
    """Sinusoidal table, vit:18-25 (grid:18-25): pe[p,2i]=sin(p*w_i), pe[p,2i+1]=cos(p*w_i),
    w_i = exp(2i * -ln(1e4)/d).  Returns (max_len, d_model) float32.

    The engine and the tests read the module's own buffer; this restatement only
    documents it and is checked against that buffer.
    """
    pos = np.arange(max_len, dtype=F32)[:, None]
    div = np.exp(np.arange(0, d_model, 2, dtype=F32) * F32(-math.log(10000.0) / d_model)).astype(F32)
    pe = np.zeros((max_len, d_model), F32)
    pe[:, 0::2] = np.sin(pos * div)
    pe[:, 1::2] = np.cos(pos * div)
    return pe


def square_subsequent_mask(t: int) -> np.ndarray:
    """vit:149-153 (grid:142-145): (t,t) float mask, -inf strictly above the diagonal."""
    m = np.zeros((t, t), F32)
    m[np.triu_indices(t, k=1)] = -np.inf
    return m


def _layer_norm(x: np.ndarray, g: np.ndarray, b: np.ndarray, eps: float = 1e-5) -> np.ndarray:
    """torch.nn.LayerNorm over the last dim (biased variance, eps inside the sqrt)."""
    mu = x.mean(-1, keepdims=True, dtype=F32)
    xc = x - mu
    var = (xc * xc).mean(-1, keepdims=True, dtype=F32)
    return (xc / np.sqrt(var + F32(eps))) * g + b


def _softmax(s: np.ndarray) -> np.ndarray:
    m = s.max(-1, keepdims=True)
    e = np.exp(s - m)
    return e / e.sum(-1, keepdims=True, dtype=F32)


def log_softmax(s: np.ndarray) -> np.ndarray:
    """F.log_softmax over the last dim (scst:230)."""
    m = s.max(-1, keepdims=True)
    z = s - m
    return z - np.log(np.exp(z).sum(-1, keepdims=True, dtype=F32))


def _split_heads(x: np.ndarray, h: int) -> np.ndarray:
    b, t, d = x.shape
    return x.reshape(b, t, h, d // h).transpose(0, 2, 1, 3)  # (B,H,t,hd)


def _merge_heads(x: np.ndarray) -> np.ndarray:
    b, h, t, hd = x.shape
    return x.transpose(0, 2, 1, 3).reshape(b, t, h * hd)


def _attention(q, k, v, n_head: int, mask: Optional[np.ndarray]) -> np.ndarray:
    """softmax(q k^T / sqrt(hd) + mask) v per head; q (B,t,D), k/v (B,s,D)."""
    qh, kh, vh = _split_heads(q, n_head), _split_heads(k, n_head), _split_heads(v, n_head)
    hd = qh.shape[-1]
    s = (qh @ kh.transpose(0, 1, 3, 2)) * F32(1.0 / math.sqrt(hd))
    if mask is not None:
        s = s + mask
    return _merge_heads(_softmax(s.astype(F32)) @ vh)


def _lw(w: Weights, i: int, name: str) -> np.ndarray:
    return w[f"transformer_decoder.layers.{i}.{name}"]


# --------------------------------------------------------------------------- full-prefix form
def decoder_forward(w: Weights, tgt: np.ndarray, memory: np.ndarray, n_head: int = 8,
                    causal: bool = True) -> np.ndarray:
    """TransformerDecoder.forward, vit:155-182 / grid:147-158.

    tgt (B,t) integer ids, memory (B,S,D) float32 -> logits (B,t,V) float32.
    embedding * sqrt(d_model) (vit:166) + pe[:, :t] (vit:32, dropout = identity in
    eval) -> L post-LN decoder layers (vit:123-134; torch _sa_block/_mha_block/_ff_block)
    -> fc_out (vit:180).  ``causal`` applies the vit:149-153 mask as every masked call
    site does (vit:311-312, scst:222-223).
    """
    d = w["embedding.weight"].shape[1]
    t = tgt.shape[1]
    pe = w["pos_encoder.pe"].reshape(-1, d)
    if t > pe.shape[0]:
        raise ValueError(f"prefix length {t} exceeds positional table ({pe.shape[0]} rows)")  # vit:32
    x = w["embedding.weight"][tgt].astype(F32) * F32(math.sqrt(d)) + pe[None, :t]
    mask = square_subsequent_mask(t) if causal else None
    memory = memory.astype(F32, copy=False)
    for i in range(n_layers(w)):
        # --- self-attention block: x = norm1(x + out_proj(attn(in_proj(x))))
        wi, bi = _lw(w, i, "self_attn.in_proj_weight"), _lw(w, i, "self_attn.in_proj_bias")
        qkv = x @ wi.T + bi
        q, k, v = qkv[..., :d], qkv[..., d:2 * d], qkv[..., 2 * d:]
        a = _attention(q, k, v, n_head, mask)
        a = a @ _lw(w, i, "self_attn.out_proj.weight").T + _lw(w, i, "self_attn.out_proj.bias")
        x = _layer_norm(x + a, _lw(w, i, "norm1.weight"), _lw(w, i, "norm1.bias"))
        # --- cross-attention block: memory is re-projected to K,V on every call
        wi, bi = _lw(w, i, "multihead_attn.in_proj_weight"), _lw(w, i, "multihead_attn.in_proj_bias")
        q = x @ wi[:d].T + bi[:d]
        k = memory @ wi[d:2 * d].T + bi[d:2 * d]
        v = memory @ wi[2 * d:].T + bi[2 * d:]
        a = _attention(q, k, v, n_head, None)
        a = a @ _lw(w, i, "multihead_attn.out_proj.weight").T + _lw(w, i, "multihead_attn.out_proj.bias")
        x = _layer_norm(x + a, _lw(w, i, "norm2.weight"), _lw(w, i, "norm2.bias"))
        # --- feed-forward block
        h = np.maximum(x @ _lw(w, i, "linear1.weight").T + _lw(w, i, "linear1.bias"), F32(0))
        f = h @ _lw(w, i, "linear2.weight").T + _lw(w, i, "linear2.bias")
        x = _layer_norm(x + f, _lw(w, i, "norm3.weight"), _lw(w, i, "norm3.bias"))
    return (x @ w["fc_out.weight"].T + w["fc_out.bias"]).astype(F32)


def greedy_search(w: Weights, memory: np.ndarray, start_token: int, end_token: int,
                  max_len: int = 50, n_head: int = 8) -> np.ndarray:
    """_greedy_search, vit:296-325 / grid:230-251, from the encoder output onwards.

    Re-runs the whole decoder over the whole prefix at every step (no KV cache), takes
    argmax of the last row, appends, and stops only when EVERY row emitted ``end_token``
    on this very step (vit:322).  Returns int64 (B, L), L <= max_len, column 0 = start.
    """
    b = memory.shape[0]
    gen = np.full((b, 1), start_token, np.int64)
    for _ in range(max_len - 1):
        logits = decoder_forward(w, gen, memory, n_head)
        nxt = logits[:, -1].argmax(-1).astype(np.int64)
        gen = np.concatenate([gen, nxt[:, None]], 1)
        if (nxt == end_token).all():
            break
    return gen


def sample_with_log_probs(w: Weights, memory: np.ndarray, start_token: int, end_token: int,
                          max_len: int, draw: Callable[[np.ndarray, int], np.ndarray],
                          n_head: int = 8) -> Tuple[np.ndarray, np.ndarray]:
    """SCSTLoss._sample_with_log_probs, scst:202-254, from the encoder output onwards,
    with the model in eval() (dropout caveat).

    ``draw(logits (B,V), step) -> ids (B,)`` stands in for ``torch.multinomial(softmax, 1)``
    (scst:229-233).  Log-prob of the drawn id from log_softmax (scst:230,236), forced to 0
    for rows already finished (scst:239); ``finished`` is cumulative (scst:246); stop when
    all rows finished (scst:248).  Returns (ids (B,L) int64, log_probs (B,L-1) float32).
    """
    b = memory.shape[0]
    gen = np.full((b, 1), start_token, np.int64)
    finished = np.zeros(b, bool)
    lps = []
    for step in range(max_len - 1):
        logits = decoder_forward(w, gen, memory, n_head)[:, -1]
        lsm = log_softmax(logits)
        nxt = np.asarray(draw(logits, step)).astype(np.int64)
        lp = lsm[np.arange(b), nxt]
        lp = np.where(finished, F32(0), lp).astype(F32)
        lps.append(lp)
        gen = np.concatenate([gen, nxt[:, None]], 1)
        finished |= nxt == end_token
        if finished.all():
            break
    return gen, np.stack(lps, 1)


def trim_caption(ids, start_token: int, end_token: int, pad_token: int = 0):
    """What the callers keep of one row (scst:256-269, inference_vit_transformer.py:113-124):
    tokens before the first <end>, without <start>/<pad>."""
    out = []
    for t in ids:
        t = int(t)
        if t == end_token:
            break
        if t not in (start_token, pad_token):
            out.append(t)
    return out


# --------------------------------------------------------------------------- KV-cached form
class KVCachedDecoder:
    """The same function as ``decoder_forward(...)[:, -1]`` evaluated incrementally - the form
    the CUDA engine implements.  ``prefill`` projects the memory to cross K/V
    once per image (the reference redoes it every call, vit:312 -> torch _mha_block); ``step``
    consumes one token per sequence and appends its self-attention K/V.

    Sequences are image-major: sequence ``m`` uses image ``m // group``.
    """

    def __init__(self, w: Weights, n_head: int = 8):
        self.w = w
        self.h = n_head
        self.d = w["embedding.weight"].shape[1]
        self.L = n_layers(w)
        self.pe = w["pos_encoder.pe"].reshape(-1, self.d)

    def prefill(self, memory: np.ndarray, group: int = 1) -> None:
        d = self.d
        self.group = group
        self.ck, self.cv = [], []
        memory = memory.astype(F32, copy=False)
        for i in range(self.L):
            wi, bi = _lw(self.w, i, "multihead_attn.in_proj_weight"), _lw(self.w, i, "multihead_attn.in_proj_bias")
            self.ck.append(memory @ wi[d:2 * d].T + bi[d:2 * d])
            self.cv.append(memory @ wi[2 * d:].T + bi[2 * d:])
        self.sk = [None] * self.L
        self.sv = [None] * self.L
        self.n = 0

    def step(self, tokens: np.ndarray) -> np.ndarray:
        """tokens (M,) -> logits (M,V) for position ``self.n``; advances the cache."""
        w, d = self.w, self.d
        if self.n >= self.pe.shape[0]:
            raise ValueError("position exceeds positional table")
        x = (w["embedding.weight"][tokens].astype(F32) * F32(math.sqrt(d)) + self.pe[self.n])[:, None]
        g = self.group
        for i in range(self.L):
            qkv = x @ _lw(w, i, "self_attn.in_proj_weight").T + _lw(w, i, "self_attn.in_proj_bias")
            q, k, v = qkv[..., :d], qkv[..., d:2 * d], qkv[..., 2 * d:]
            self.sk[i] = k if self.sk[i] is None else np.concatenate([self.sk[i], k], 1)
            self.sv[i] = v if self.sv[i] is None else np.concatenate([self.sv[i], v], 1)
            a = _attention(q, self.sk[i], self.sv[i], self.h, None)
            a = a @ _lw(w, i, "self_attn.out_proj.weight").T + _lw(w, i, "self_attn.out_proj.bias")
            x = _layer_norm(x + a, _lw(w, i, "norm1.weight"), _lw(w, i, "norm1.bias"))
            wi, bi = _lw(w, i, "multihead_attn.in_proj_weight"), _lw(w, i, "multihead_attn.in_proj_bias")
            q = x @ wi[:d].T + bi[:d]
            ck = np.repeat(self.ck[i], g, 0) if g > 1 else self.ck[i]
            cv = np.repeat(self.cv[i], g, 0) if g > 1 else self.cv[i]
            a = _attention(q, ck, cv, self.h, None)
            a = a @ _lw(w, i, "multihead_attn.out_proj.weight").T + _lw(w, i, "multihead_attn.out_proj.bias")
            x = _layer_norm(x + a, _lw(w, i, "norm2.weight"), _lw(w, i, "norm2.bias"))
            hdn = np.maximum(x @ _lw(w, i, "linear1.weight").T + _lw(w, i, "linear1.bias"), F32(0))
            f = hdn @ _lw(w, i, "linear2.weight").T + _lw(w, i, "linear2.bias")
            x = _layer_norm(x + f, _lw(w, i, "norm3.weight"), _lw(w, i, "norm3.bias"))
        self.n += 1
        return (x[:, 0] @ w["fc_out.weight"].T + w["fc_out.bias"]).astype(F32)


def rollout_cached(w: Weights, memory: np.ndarray, start_token: int, end_token: int, max_len: int,
                   group: int = 1, n_sample: int = 0,
                   draw: Optional[Callable[[np.ndarray, int, np.ndarray], np.ndarray]] = None,
                   n_head: int = 8, pad_token: int = 0):
    """The engine's rollout restated on the CPU: image-major batch of ``B_img*group``
    sequences; within a group rows ``g < n_sample`` are sampled (scst:202-254 semantics),
    rows ``g >= n_sample`` are greedy (vit:296-325 semantics).

    Stops after the first step at which (all greedy rows emitted <end> simultaneously at some
    step so far - latched) and (all sampled rows have finished).  Returns
    (ids (M,max_len) int64 padded with ``pad_token``, logp (M,max_len-1) float32, steps,
    logits (M,steps,V)).  ``draw(logits_rows, step, row_index) -> ids``.
    """
    b_img = memory.shape[0]
    m = b_img * group
    dec = KVCachedDecoder(w, n_head)
    dec.prefill(memory, group)
    is_sample = (np.arange(m) % group) < n_sample
    ids = np.full((m, max_len), pad_token, np.int64)
    ids[:, 0] = start_token
    logp = np.zeros((m, max_len - 1), F32)
    finished = np.zeros(m, bool)
    greedy_done = not (~is_sample).any()
    sample_done = not is_sample.any()
    all_logits = []
    steps = 0
    for n in range(max_len - 1):
        logits = dec.step(ids[:, n])
        all_logits.append(logits)
        nxt = logits.argmax(-1).astype(np.int64)
        if is_sample.any():
            rows = np.nonzero(is_sample)[0]
            nxt[rows] = np.asarray(draw(logits[rows], n, rows)).astype(np.int64)
            lsm = log_softmax(logits[rows])
            lp = lsm[np.arange(len(rows)), nxt[rows]]
            logp[rows, n] = np.where(finished[rows], F32(0), lp)
        ids[:, n + 1] = nxt
        steps = n + 1
        ended = nxt == end_token
        finished |= ended & is_sample
        if (~is_sample).any() and ended[~is_sample].all():
            greedy_done = True
        if is_sample.any() and finished[is_sample].all():
            sample_done = True
        if greedy_done and sample_done:
            break
    return ids, logp, steps, np.stack(all_logits, 1)
