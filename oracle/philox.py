"""Philox4x32-10 and the engine's categorical sampler, restated in numpy (oracle - test
infrastructure only).

The reference draws with ``torch.multinomial(softmax(logits), 1)`` (utils/scst_loss.py:229-233)
on torch's global generator.  Reproducing torch's CUDA draw order is neither possible with a
KV cache nor required ("RNG parity"): parity is distributional.  The engine
therefore defines its own counter-based scheme, documented here and in DESIGN.md:

    key     = (seed & 0xffffffff, seed >> 32)
    counter = (seq_id & 0xffffffff, seq_id >> 32, step, 0)
    r       = Philox4x32-10(counter, key)
    u       = (r[0] >> 8) * 2**-24                      in [0, 1)
    token   = first v with  sum_{j<=v} exp(l_j - max l)  >  u * sum_j exp(l_j - max l)

``seq_id`` is the GLOBAL sequence index (image-major: image*group + g), so a rollout is
bit-reproducible and independent of how the batch is sharded over GPUs.
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32 with 10 rounds (Salmon et al., SC'11).  Inputs broadcastable
    unsigned 32-bit arrays; returns four uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, np.uint64) & _MASK for c in (c0, c1, c2, c3))
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def uniform(seed: int, seq_id, step) -> np.ndarray:
    """u in [0,1) for (seed, global sequence id, step) - float64 holding an exact 24-bit value."""
    seq_id = np.asarray(seq_id, np.uint64)
    r0, _, _, _ = philox4x32_10(seq_id & _MASK, seq_id >> np.uint64(32), np.asarray(step, np.uint64),
                                np.uint64(0), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return (r0 >> np.uint32(8)).astype(np.float64) * (2.0 ** -24)


def _cdf(logits: np.ndarray) -> np.ndarray:
    z = logits.astype(np.float64)
    e = np.exp(z - z.max(-1, keepdims=True))
    return np.cumsum(e, -1)


def categorical_from_uniform(logits: np.ndarray, u: np.ndarray) -> np.ndarray:
    """Inverse-CDF draw: first index whose inclusive cumulative weight exceeds u * total."""
    c = _cdf(logits)
    target = (np.asarray(u, np.float64) * c[..., -1])[..., None]
    idx = (c > target).argmax(-1)
    none = ~(c > target).any(-1)
    return np.where(none, logits.shape[-1] - 1, idx).astype(np.int64)


def draw_consistent(logits: np.ndarray, u: np.ndarray, ids: np.ndarray, eps: float = 2e-6) -> np.ndarray:
    """True where ``ids`` is an admissible inverse-CDF draw for ``u`` once float32 rounding of
    the device-side cumulative sums (relative ``eps``) is allowed for."""
    c = _cdf(logits)
    tot = c[..., -1]
    rows = np.arange(logits.shape[0])
    hi = c[rows, ids] / tot
    lo = np.where(ids > 0, c[rows, np.maximum(ids - 1, 0)] / tot, 0.0)
    u = np.asarray(u, np.float64)
    return (u >= lo - eps) & (u <= hi + eps)


def make_draw(seed: int, seq_id_base: int = 0):
    """``draw(logits_rows, step, row_index)`` callback for ``decoder_np.rollout_cached`` that
    applies the engine's scheme."""
    def draw(logits, step, rows):
        u = uniform(seed, seq_id_base + np.asarray(rows, np.uint64), step)
        return categorical_from_uniform(logits, u)
    return draw
