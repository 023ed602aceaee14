"""Data-parallel rollouts: images sharded over ranks, one all-gather of ids / log-probs per rollout.

The reference is single-device; BASELINE.json shards the image batch over
1/2/4/8 GPUs with weights replicated.  Sequences are independent given their image, so the
data path needs no collective; the only exchange is the gather of results for the reward
computation that follows (utils/scst_loss.py:175-180).

* contiguous blocks of images per rank, all ``group`` rollouts of an image on one rank;
* Philox counters use the GLOBAL sequence id (``seq_id_base = first_image * group``), so the
  draws do not depend on the number of ranks.  The TOKENS are bit-identical for every N when all
  shards take the same kernel route: always in fp32 mode and with ``batch_invariant=True``
  (CAPDEC_FLAG_BATCH_INVARIANT); in the default bf16 mode the engine picks the fastest route per
  batch size (projection chains / swap-AB / weight-stationary GEMMs differ in fp32 summation
  order), so shards of a different size class agree with the unsharded rollout within the bf16
  tolerance only and a near-tie may flip a token (bench.py's ``verify`` reports both);
* a rank cannot know the global stop step without a per-step collective, and the greedy rule
  (vit:322: ALL rows emit <end> on the SAME step) depends on what finished rows keep emitting,
  so ranks decode all ``max_len - 1`` steps (``early_stop=False``) and the gathered result is
  trimmed to the width the reference would have produced on the global batch (vit:322 for
  greedy rows, scst:248 for sampled rows).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_images: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous split; the first ``n_images % world_size`` ranks take one extra image."""
    base, extra = divmod(n_images, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def reference_width(ids: torch.Tensor, is_sample_row: torch.Tensor, end_token: int) -> int:
    """Number of columns (including <start>) the reference loops would have produced for this
    global batch: greedy rows stop at the first step where ALL of them emit <end> at once
    (vit:322); sampled rows when every one has emitted <end> at least once (scst:246-249)."""
    steps_total = ids.size(1) - 1
    tok = ids[:, 1:]
    need = 0
    greedy = tok[~is_sample_row]
    if greedy.numel():
        hit = (greedy == end_token).all(0).nonzero()
        need = max(need, int(hit[0]) + 1 if hit.numel() else steps_total)
    sampled = tok[is_sample_row]
    if sampled.numel():
        fin = ((sampled == end_token).cumsum(1) > 0).all(0).nonzero()
        need = max(need, int(fin[0]) + 1 if fin.numel() else steps_total)
    return need + 1


def gather_rollout(ids_local: torch.Tensor, logp_local: Optional[torch.Tensor], counts, group: int,
                   process_group=None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """All-gather padded (M_local, max_len) ids and (M_local, max_len-1) log-probs.  ``counts``
    lists the images per rank (ragged shards are padded to the largest and cut after the gather)."""
    world = dist.get_world_size(process_group)
    m_max = max(counts) * group

    def pad(t: torch.Tensor) -> torch.Tensor:
        if t.size(0) == m_max:
            return t.contiguous()
        out = t.new_zeros((m_max,) + tuple(t.shape[1:]))
        out[: t.size(0)] = t
        return out

    # NCCL moves device memory only: results that came back through the host-buffer entry point are staged
    # on this rank's GPU for the gather and returned on the host again
    stage = None
    if not ids_local.is_cuda and dist.get_backend(process_group) == "nccl":
        stage = torch.device("cuda", torch.cuda.current_device())

    def gather(t: torch.Tensor) -> torch.Tensor:
        src = pad(t).to(stage, non_blocking=True) if stage is not None else pad(t)
        buf = src.new_empty((world * m_max,) + tuple(t.shape[1:]))
        dist.all_gather_into_tensor(buf, src, group=process_group)
        parts = [buf[r * m_max: r * m_max + counts[r] * group] for r in range(world)]
        out = torch.cat(parts, 0)
        return out.cpu() if stage is not None else out

    # ids travel as int32 (half the bytes; the vocabulary is tiny) and widen after the gather
    ids = gather(ids_local.to(torch.int32)).to(torch.int64)
    logp = gather(logp_local) if logp_local is not None else None
    return ids, logp


def sharded_rollout(local_decode: Callable[..., Tuple[torch.Tensor, Optional[torch.Tensor]]],
                    n_images_global: int, start_token: int, end_token: int, max_len: int, group: int = 1,
                    n_sample: int = 0, seed: int = 0, pad_token: int = 0, process_group=None):
    """Run ``local_decode(lo, hi, seq_id_base) -> (ids (M_local, max_len) padded, logp or None)`` on
    this rank's images ``[lo, hi)`` and return the gathered, reference-width result on every rank.

    ``local_decode`` is normally ``lambda lo, hi, base: engine.decode_from_memory(memory[lo:hi], ...,
    seq_id_base=base, trim=False, early_stop=False)`` unpacked to ``(ids, logp)``; tests inject a CPU stand-in."""
    world = dist.get_world_size(process_group)
    rank = dist.get_rank(process_group)
    bounds = [shard_bounds(n_images_global, world, r) for r in range(world)]
    counts = [hi - lo for lo, hi in bounds]
    lo, hi = bounds[rank]
    ids_local, logp_local = local_decode(lo, hi, lo * group)
    ids, logp = gather_rollout(ids_local, logp_local, counts, group, process_group)
    is_sample = (torch.arange(ids.size(0), device=ids.device) % group) < n_sample
    width = reference_width(ids, is_sample, end_token)
    ids = ids[:, :width]
    logp = logp[:, : width - 1] if logp is not None else None
    return ids, logp


class HostFeeder:
    """Double-buffered host -> device feed for back-to-back sharded rollouts: ``prefetch(t)`` starts the copy of a pinned
    CPU tensor on a side stream, ``take()`` returns the oldest prefetched copy once the current stream may read it.  Calling
    ``prefetch(next)`` before decoding ``take()``'s result puts the next rollout's copy underneath the current decode (the
    single-GPU host entry point does the same inside the C ABI: capdec_prefetch_host).  Plumbing only: no kernels."""

    def __init__(self, device: torch.device):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self._queue = []

    def prefetch(self, host_tensor: torch.Tensor) -> None:
        with torch.cuda.stream(self.stream):
            dev_tensor = host_tensor.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self._queue.append((dev_tensor, ev))

    def take(self) -> torch.Tensor:
        dev_tensor, ev = self._queue.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        dev_tensor.record_stream(cur)
        return dev_tensor
