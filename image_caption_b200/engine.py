"""Host-side mirror of the reference's decode entry points over libcapdec.so.

``CaptionDecodeEngine`` wraps a live reference decoder module (``model.decoder`` of
``ViTTransformerCaptioning`` / ``GridTransformerCaptioning``; models/vit_transformer_model.py:103-182,
models/grid_transformer_model.py:113-158) and serves

* ``generate(model, images, start_token, end_token, max_len=50, method='greedy')``
  - models/vit_transformer_model.py:276-325, models/grid_transformer_model.py:222-251
* ``sample_with_log_probs(model, images, start_token, end_token, max_len, device)``
  - utils/scst_loss.py:202-254
* ``decode_from_memory(memory, ...)`` - the same loops entered after the encoder, used by the
  benchmarks and by multi-sample SCST (``group`` rollouts per image sharing one cross K/V).

PyTorch is used for device memory, streams and the (untouched) encoders only; every decode
kernel lives in the C-ABI library.  There is no CPU or eager fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import _lib

_LAYER_FIELDS = (
    ("self_in_w", "self_attn.in_proj_weight"), ("self_in_b", "self_attn.in_proj_bias"),
    ("self_out_w", "self_attn.out_proj.weight"), ("self_out_b", "self_attn.out_proj.bias"),
    ("cross_in_w", "multihead_attn.in_proj_weight"), ("cross_in_b", "multihead_attn.in_proj_bias"),
    ("cross_out_w", "multihead_attn.out_proj.weight"), ("cross_out_b", "multihead_attn.out_proj.bias"),
    ("lin1_w", "linear1.weight"), ("lin1_b", "linear1.bias"), ("lin2_w", "linear2.weight"), ("lin2_b", "linear2.bias"),
)


@dataclass
class Rollout:
    """Result of one rollout.  ``ids`` (M, L) int64 with column 0 = <start>; ``logp`` (M, L-1)
    float32 (0 for greedy rows and after a sampled row finished) or None; ``logits`` (M, L-1, V) or
    None; ``steps`` = L-1."""
    ids: torch.Tensor
    logp: Optional[torch.Tensor]
    steps: int
    logits: Optional[torch.Tensor] = None


def _stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def vit_encoder_tokens(encoder: torch.nn.Module, images: torch.Tensor) -> torch.Tensor:
    """``VisionTransformerEncoder.forward`` (models/vit_transformer_model.py:70-100) up to, not including, the
    class-token drop and the output projection: (B, 1 + patches, 768)."""
    vit = encoder.vit
    x = vit._process_input(images)                                          # vit:84
    x = torch.cat([vit.class_token.expand(x.shape[0], -1, -1), x], dim=1)   # vit:88-89
    return vit.encoder(x)                                                   # vit:92


class CaptionDecodeEngine:
    def __init__(self, decoder: torch.nn.Module, precision: str = "bf16", max_images: int = 64,
                 max_group: int = 1, max_mem: int = 196, max_len: int = 50, use_graph: bool = True,
                 use_umma: bool = True, step_impl: str = "v2", pdl: bool = True, batch_invariant: bool = False):
        self._lib = _lib.load()
        self.decoder = decoder
        sd = dict(decoder.state_dict())
        emb = sd["embedding.weight"]
        if not emb.is_cuda:
            raise RuntimeError("CaptionDecodeEngine needs the decoder on a CUDA device (no CPU path exists)")
        self.device = emb.device
        self.vocab, self.d_model = emb.shape
        self.n_layer = len(decoder.transformer_decoder.layers)
        first = decoder.transformer_decoder.layers[0]
        self.n_head = first.self_attn.num_heads
        self.d_ff = first.linear1.out_features
        self.pe_rows = sd["pos_encoder.pe"].shape[1]
        if getattr(first, "norm_first", False):
            raise ValueError("only post-LN decoder layers (norm_first=False) are supported, as in the reference")
        if precision not in ("bf16", "fp32", "tf32"):
            raise ValueError(f"precision must be 'bf16', 'fp32' or 'tf32', got {precision!r}")
        self.precision = precision
        if step_impl not in ("v1", "v2"):
            raise ValueError(f"step_impl must be 'v1' or 'v2', got {step_impl!r}")
        self._flags = ((0 if use_graph else _lib.FLAG_NO_GRAPH) | (0 if use_umma else _lib.FLAG_NO_UMMA)
                       | (_lib.FLAG_V1_STEP if step_impl == "v1" else 0) | (0 if pdl else _lib.FLAG_NO_PDL)
                       | (_lib.FLAG_BATCH_INVARIANT if batch_invariant else 0))
        self._cap = dict(max_images=max_images, max_group=max_group, max_mem=max_mem, max_len=max_len)
        self._h = C.c_void_p()
        self._weights_sig = None
        self._keep = []  # fp32 contiguous views handed to set_weights
        self._create()

    # ------------------------------------------------------------------ handle management
    def _create(self) -> None:
        cfg = _lib.Config(vocab=self.vocab, d_model=self.d_model, n_head=self.n_head, n_layer=self.n_layer,
                          d_ff=self.d_ff, pe_rows=self.pe_rows, max_mem=self._cap["max_mem"],
                          max_images=self._cap["max_images"], max_group=self._cap["max_group"],
                          max_len=self._cap["max_len"],
                          precision={"bf16": _lib.PREC_BF16, "fp32": _lib.PREC_FP32, "tf32": _lib.PREC_TF32}[self.precision],
                          flags=self._flags)
        h = C.c_void_p()
        rc = self._lib.capdec_create(C.byref(cfg), self.device.index or 0, C.byref(h))
        _lib.check(None, rc)
        self._h = h
        self._weights_sig = None
        self._proj_sig = None

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.capdec_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ensure_capacity(self, n_images: int, group: int, mem_len: int, max_len: int) -> None:
        need = dict(max_images=n_images, max_group=group, max_mem=mem_len, max_len=max_len)
        if all(need[k] <= self._cap[k] for k in need):
            return
        torch.cuda.synchronize(self.device)
        self.close()
        self._cap = {k: max(self._cap[k], need[k]) for k in need}
        self._create()

    # ------------------------------------------------------------------ weights
    def _signature(self):
        return tuple((p.data_ptr(), p._version) for p in self.decoder.state_dict().values())

    def sync_weights(self, force: bool = False) -> None:
        """Re-pack the module's parameters if any was updated in place (optimiser step) or rebound
        (EMA ``apply_shadow``/``restore``, utils/optimizations.py:444-456)."""
        sig = self._signature()
        if not force and sig == self._weights_sig:
            return
        sd = self.decoder.state_dict()
        keep = []

        def ptr(name: str) -> int:
            t = sd[name].detach()
            if t.device != self.device:
                raise RuntimeError(f"parameter {name} is on {t.device}, engine on {self.device}")
            t = t.to(torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        w = _lib.Weights()
        w.embedding, w.pe = ptr("embedding.weight"), ptr("pos_encoder.pe")
        w.fc_out_w, w.fc_out_b = ptr("fc_out.weight"), ptr("fc_out.bias")
        for i in range(self.n_layer):
            pre = f"transformer_decoder.layers.{i}."
            lw = w.layer[i]
            for field, key in _LAYER_FIELDS:
                setattr(lw, field, ptr(pre + key))
            for j in range(3):
                lw.norm_w[j] = ptr(f"{pre}norm{j + 1}.weight")
                lw.norm_b[j] = ptr(f"{pre}norm{j + 1}.bias")
        _lib.check(self._h, self._lib.capdec_set_weights(self._h, C.byref(w), _stream_ptr(self.device)))
        self._keep = keep  # released on the next repack; the pack kernels are stream-ordered before that
        self._weights_sig = sig

    def sync_projection(self, projection: torch.nn.Linear) -> None:
        """Fold an encoder output projection (``VisionTransformerEncoder.projection``, vit:61) into the cross K/V
        weights (capdec_set_memory_projection).  Re-done when the projection or the decoder parameters changed."""
        self.sync_weights()
        sig = (projection.weight.data_ptr(), projection.weight._version, projection.bias.data_ptr(),
               projection.bias._version, self._weights_sig)
        if sig == getattr(self, "_proj_sig", None):
            return
        if projection.out_features != self.d_model:
            raise ValueError(f"projection must map to d_model={self.d_model}, got {projection.out_features}")
        sd = self.decoder.state_dict()
        keep = [projection.weight.detach().to(self.device, torch.float32).contiguous(),
                projection.bias.detach().to(self.device, torch.float32).contiguous()]
        ws = (C.c_void_p * self.n_layer)()
        bs = (C.c_void_p * self.n_layer)()
        for i in range(self.n_layer):
            w = sd[f"transformer_decoder.layers.{i}.multihead_attn.in_proj_weight"].detach().to(torch.float32).contiguous()
            b = sd[f"transformer_decoder.layers.{i}.multihead_attn.in_proj_bias"].detach().to(torch.float32).contiguous()
            keep += [w, b]
            ws[i], bs[i] = w.data_ptr(), b.data_ptr()
        _lib.check(self._h, self._lib.capdec_set_memory_projection(
            self._h, keep[0].data_ptr(), keep[1].data_ptr(), projection.in_features, ws, bs, _stream_ptr(self.device)))
        self._keep_proj = keep  # released on the next fold; the fold kernels are stream-ordered before that
        self._proj_sig = sig

    # ------------------------------------------------------------------ decode
    def _check_lengths(self, max_len: int) -> None:
        if max_len < 2:
            raise ValueError("max_len must be >= 2")
        if max_len - 1 > self.pe_rows:  # the reference fails with a shape error at vit:32
            raise ValueError(f"max_len={max_len} needs {max_len - 1} positions but pos_encoder.pe has {self.pe_rows}")
        if max_len > 128:
            raise ValueError("max_len > 128 is not built")

    def decode_from_memory(self, memory: torch.Tensor, start_token: int, end_token: int, max_len: int = 50,
                           group: int = 1, n_sample: int = 0, seed: int = 0, seq_id_base: int = 0,
                           pad_token: int = 0, return_logits: bool = False, trim: bool = True,
                           early_stop: bool = True, projection: Optional[torch.nn.Linear] = None,
                           first_token: int = 0) -> Rollout:
        """Decode ``group`` sequences per image of ``memory`` (B, S, 512) float32: rows ``g < n_sample``
        of each group are sampled, the rest greedy.  A CUDA ``memory`` runs on the current stream and
        returns CUDA tensors; a CPU ``memory`` takes the host-buffer entry point
        (capdec_generate_host) and returns CPU tensors.

        Encoder hand-off: with ``projection`` (the encoder's final ``nn.Linear``, vit:61) ``memory``
        is the encoder's *un-projected* CUDA token tensor (B, T, in_features); tokens ``first_token ..`` of every
        image are the memory positions (``first_token=1`` drops the ViT class token, vit:95) and the projection is
        applied inside the K/V prefill GEMM (capdec_prefill_tokens)."""
        if projection is not None:
            if memory.dim() != 3 or memory.shape[2] != projection.in_features or not memory.is_cuda:
                raise ValueError(f"tokens must be CUDA (B, T, {projection.in_features}), got {tuple(memory.shape)}")
            if not 0 <= first_token < memory.shape[1]:
                raise ValueError("first_token outside the token axis")
        elif memory.dim() != 3 or memory.shape[2] != self.d_model:
            raise ValueError(f"memory must be (B, S, {self.d_model}), got {tuple(memory.shape)}")
        self._check_lengths(max_len)
        n_images, mem_len = memory.shape[0], memory.shape[1] - (first_token if projection is not None else 0)
        m = n_images * group
        self._ensure_capacity(n_images, group, mem_len, max_len)
        if projection is not None:
            self.sync_projection(projection)
        else:
            self.sync_weights()
        memory = memory.detach().to(torch.float32).contiguous()
        steps = C.c_int32(0)
        args = _lib.DecodeArgs(group=group, n_sample=n_sample, start_token=start_token, end_token=end_token,
                               pad_token=pad_token, max_len=max_len, seed=seed, seq_id_base=seq_id_base,
                               flags=0 if early_stop else _lib.DECODE_NO_EARLY_STOP,
                               steps_out=C.pointer(steps))
        want_logp = n_sample > 0
        stream = _stream_ptr(self.device)
        if memory.is_cuda:
            if memory.device != self.device:
                raise RuntimeError("memory and decoder are on different devices")
            ids = torch.empty((m, max_len), dtype=torch.int64, device=self.device)
            logp = torch.empty((m, max_len - 1), dtype=torch.float32, device=self.device) if want_logp else None
            logits = (torch.empty((m, max_len - 1, self.vocab), dtype=torch.float32, device=self.device)
                      if return_logits else None)
            args.ids_out = ids.data_ptr()
            args.logp_out = logp.data_ptr() if want_logp else None
            args.logits_out = logits.data_ptr() if return_logits else None
            if projection is not None:
                _lib.check(self._h, self._lib.capdec_prefill_tokens(self._h, memory.data_ptr(), n_images, memory.shape[1],
                                                                    first_token, mem_len, stream))
            else:
                _lib.check(self._h, self._lib.capdec_prefill(self._h, memory.data_ptr(), n_images, mem_len, stream))
            _lib.check(self._h, self._lib.capdec_decode(self._h, C.byref(args), stream))
        else:
            if return_logits:
                raise ValueError("return_logits needs a CUDA memory tensor")
            ids = torch.empty((m, max_len), dtype=torch.int64, pin_memory=True)
            logp = torch.empty((m, max_len - 1), dtype=torch.float32, pin_memory=True) if want_logp else None
            logits = None
            args.ids_out = ids.data_ptr()
            args.logp_out = logp.data_ptr() if want_logp else None
            _lib.check(self._h, self._lib.capdec_generate_host(self._h, memory.data_ptr(), n_images, mem_len,
                                                               C.byref(args), stream))
        n = int(steps.value)
        if trim:
            ids = ids[:, :n + 1]
            logp = logp[:, :n] if logp is not None else None
            logits = logits[:, :n] if logits is not None else None
        return Rollout(ids=ids, logp=logp, steps=n, logits=logits)

    def prefetch_memory(self, memory: torch.Tensor) -> None:
        """Start the host -> device copy of a CPU ``memory`` (B, S, 512) float32 that a LATER ``decode_from_memory`` call
        will be given (capdec_prefetch_host): the copy runs underneath whatever the engine decodes in between.  The tensor
        must be contiguous float32 (pin it for a truly asynchronous copy) and stay unchanged until that call returns."""
        if memory.is_cuda or memory.dim() != 3 or memory.shape[2] != self.d_model or memory.dtype != torch.float32 \
                or not memory.is_contiguous():
            raise ValueError(f"memory must be a contiguous CPU float32 (B, S, {self.d_model}) tensor")
        self._ensure_capacity(memory.shape[0], 1, memory.shape[1], 2)
        if self._weights_sig is None:  # (a handle without weights refuses prefill-type calls; the consuming call re-checks them)
            self.sync_weights()
        _lib.check(self._h, self._lib.capdec_prefetch_host(self._h, memory.data_ptr(), memory.shape[0], memory.shape[1]))
        self._keep_prefetch = getattr(self, "_keep_prefetch", [])[-1:] + [memory]

    def score(self, memory: torch.Tensor, ids: torch.Tensor, end_token: int, group: int = 1,
              return_logits: bool = True, pad_token: int = 0, mode: str = "steps"):
        """Teacher-forced pass over given sequences (CAPDEC_DECODE_TEACHER_FORCED): ``ids`` (B*group, L) int64 with
        column 0 = <start>; returns ``(logits (M, L-1, V) or None, logp (M, L-1))`` where ``logits[:, k]`` is what the
        reference's ``model.decoder(ids[:, :k+1], memory, tgt_mask=causal)[:, -1]`` returns (scst:223-226, vit:312-315)
        and ``logp[:, k] = log_softmax(logits[:, k])[ids[:, k+1]]``, zero after a row's <end> (scst:236-239).
        Inference only: no autograd graph is built.

        ``mode="steps"`` replays the L-1 decode steps with forced tokens (any engine).  ``mode="full"`` is the
        full-sequence pass (capdec_forward_full): the whole-caption decoder call of the training forward
        (vit:216-255) as ONE pass over (sequences x positions) rows - every projection a single large tcgen05 GEMM, K/V
        read once per sequence instead of once per step; bf16 tensor-core engine only (CapdecError otherwise)."""
        if mode not in ("steps", "full"):
            raise ValueError(f"mode must be 'steps' or 'full', got {mode!r}")
        if memory.dim() != 3 or memory.shape[2] != self.d_model or not memory.is_cuda:
            raise ValueError(f"memory must be CUDA (B, S, {self.d_model}), got {tuple(memory.shape)}")
        n_images, mem_len = memory.shape[0], memory.shape[1]
        if ids.dim() != 2 or ids.shape[0] != n_images * group:
            raise ValueError(f"ids must be ({n_images * group}, L), got {tuple(ids.shape)}")
        max_len = ids.shape[1]
        self._check_lengths(max_len)
        ids = ids.to(self.device, torch.int64).contiguous()
        if int(ids.min()) < 0 or int(ids.max()) >= self.vocab:
            raise ValueError("token id outside the vocabulary")
        self._ensure_capacity(n_images, group, mem_len, max_len)
        self.sync_weights()
        memory = memory.detach().to(torch.float32).contiguous()
        m = n_images * group
        logp = torch.empty((m, max_len - 1), dtype=torch.float32, device=self.device)
        logits = (torch.empty((m, max_len - 1, self.vocab), dtype=torch.float32, device=self.device)
                  if return_logits else None)
        args = _lib.DecodeArgs(group=group, n_sample=group, start_token=int(ids[0, 0]), end_token=end_token,
                               pad_token=pad_token, max_len=max_len, seed=0, seq_id_base=0,
                               flags=_lib.DECODE_NO_EARLY_STOP | _lib.DECODE_TEACHER_FORCED)
        args.ids_out = ids.data_ptr()
        args.logp_out = logp.data_ptr()
        args.logits_out = logits.data_ptr() if return_logits else None
        stream = _stream_ptr(self.device)
        _lib.check(self._h, self._lib.capdec_prefill(self._h, memory.data_ptr(), n_images, mem_len, stream))
        if mode == "full":
            fa = _lib.ForwardArgs(group=group, max_len=max_len, end_token=end_token)
            fa.ids = ids.data_ptr()
            fa.logp_out = logp.data_ptr()
            fa.logits_out = logits.data_ptr() if return_logits else None
            _lib.check(self._h, self._lib.capdec_forward_full(self._h, C.byref(fa), stream))
        else:
            _lib.check(self._h, self._lib.capdec_decode(self._h, C.byref(args), stream))
        self._keep_ids = ids
        return logits, logp

    @torch.no_grad()
    def decoder_forward(self, tgt: torch.Tensor, memory: torch.Tensor, mode: str = "steps") -> torch.Tensor:
        """``model.decoder(tgt, memory, tgt_mask=causal)`` (vit:155-182 as called at vit:312 / scst:223) on the
        engine: (B, t) token ids -> (B, t, V) logits.  Causal mask only (the hot path's call shape); no autograd."""
        b, t = tgt.shape
        ids = torch.cat([tgt.to(self.device, torch.int64),
                         torch.zeros((b, 1), dtype=torch.int64, device=self.device)], dim=1)  # step t-1 scores a dummy token
        logits, _ = self.score(memory, ids, end_token=0, mode=mode)  # the end token only shapes the (discarded) log-probs
        return logits

    def beam_search_from_memory(self, memory: torch.Tensor, start_token: int, end_token: int, max_len: int = 50,
                                beam_size: int = 5, pad_token: int = 0, stop_rule: str = "vit"):
        """Batched beam search (C ABI ``capdec_beam_search``) over ``memory`` (B, S, 512) CUDA float32.
        ``stop_rule``: ``"vit"`` = an image stops when every hypothesis of a step ended
        (models/vit_transformer_model.py:402), ``"grid"`` = when the completed list is at least as long as
        the beam is wide (models/grid_transformer_model.py:308).  Returns ``(ids, lengths, scores)``:
        ids (B, max_len) int64 padded with ``pad_token``, lengths (B,) int32, scores (B,) float32."""
        if memory.dim() != 3 or memory.shape[2] != self.d_model:
            raise ValueError(f"memory must be (B, S, {self.d_model}), got {tuple(memory.shape)}")
        if not memory.is_cuda:
            raise ValueError("beam search needs a CUDA memory tensor")
        if stop_rule not in ("vit", "grid"):
            raise ValueError("stop_rule must be 'vit' or 'grid'")
        if not 1 <= beam_size <= 8:
            raise ValueError("beam_size must be in 1..8")
        self._check_lengths(max_len)
        n_images, mem_len = memory.shape[0], memory.shape[1]
        self._ensure_capacity(n_images, beam_size, mem_len, max_len)
        self.sync_weights()
        memory = memory.detach().to(torch.float32).contiguous()
        ids = torch.empty((n_images, max_len), dtype=torch.int64, device=self.device)
        lens = torch.empty((n_images,), dtype=torch.int32, device=self.device)
        scores = torch.empty((n_images,), dtype=torch.float32, device=self.device)
        steps = C.c_int32(0)
        args = _lib.BeamArgs(beam_size=beam_size, flags=_lib.BEAM_STOP_GRID if stop_rule == "grid" else 0,
                             start_token=start_token, end_token=end_token, pad_token=pad_token, max_len=max_len,
                             ids_out=ids.data_ptr(), len_out=lens.data_ptr(), score_out=scores.data_ptr(),
                             steps_out=C.pointer(steps))
        stream = _stream_ptr(self.device)
        _lib.check(self._h, self._lib.capdec_prefill(self._h, memory.data_ptr(), n_images, mem_len, stream))
        _lib.check(self._h, self._lib.capdec_beam_search(self._h, C.byref(args), stream))
        return ids, lens, scores

    # ------------------------------------------------------------------ encoder hand-off
    fuse_encoder_projection = True

    def _encode(self, model: torch.nn.Module, images: torch.Tensor):
        """``model.encoder(images)`` -- or, for the ViT encoder, its token tensor *before* the class-token drop and
        the 768 -> 512 projection (vit:80-98), which the engine folds into the K/V prefill.  Returns the keyword
        arguments of ``decode_from_memory``'s memory input."""
        enc = model.encoder
        vit, proj = getattr(enc, "vit", None), getattr(enc, "projection", None)
        if (self.fuse_encoder_projection and vit is not None and isinstance(proj, torch.nn.Linear)
                and not torch.is_grad_enabled() and proj.in_features % 64 == 0 and images.is_cuda):
            # vit:95-98 (class-token drop + projection) happen in the prefill GEMM
            return dict(memory=vit_encoder_tokens(enc, images), projection=proj, first_token=1)
        return dict(memory=enc(images))

    def scst_rollouts(self, model: torch.nn.Module, images: torch.Tensor, start_token: int, end_token: int,
                      max_len: int = 50, n_samples: int = 1, seed: Optional[int] = None):
        """The two rollouts of one SCST step from ONE encoder pass: ``SCSTLoss.forward`` encodes the batch twice
        (scst:213 for the sampled rollout, vit:304 inside ``model.generate`` for the greedy baseline).  Returns
        ``(sample_ids (B, n_samples, L), sample_logp (B, n_samples, L-1), greedy_ids (B, L))``; the sampled and greedy
        sequences of an image share its cross K/V on the device.  Eval-mode encoder and decoder, no autograd (the
        caller recomputes the sampled log-probs teacher-forced when it needs gradients, as ``sample_with_log_probs``
        does)."""
        with torch.no_grad():
            was_training = model.training
            model.eval()
            try:
                kw = self._encode(model, images)
            finally:
                model.train(was_training)
            if seed is None:
                seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            g = n_samples + 1
            out = self.decode_from_memory(start_token=start_token, end_token=end_token, max_len=max_len, group=g,
                                          n_sample=n_samples, seed=seed, **kw)
        b, width = images.shape[0], out.ids.shape[1]
        ids = out.ids.view(b, g, width)
        logp = out.logp.view(b, g, width - 1)
        return ids[:, :n_samples], logp[:, :n_samples], ids[:, n_samples]

    # ------------------------------------------------------------------ reference-facing entry points
    @torch.no_grad()
    def generate(self, model: torch.nn.Module, images: torch.Tensor, start_token: int, end_token: int,
                 max_len: int = 50, method: str = "greedy", beam_size: int = 5) -> torch.Tensor:
        """Drop-in for ``model.generate`` (vit:276-294 / grid:222-228); both methods run on the engine.

        ``beam_search``: the reference decodes one image per call and concatenates the results, which only
        works when all captions have the same length (``torch.cat`` at vit:341 / grid:263); here the batch is
        decoded at once and returned as (B, longest) padded with 0 -- identical for one image or equal lengths.
        The ViT model always uses a beam of 5 (vit:291); the Grid model's extra stop rule (grid:308) is selected
        by the class name."""
        if method == "greedy":
            model.eval()  # vit:298 - a side effect the reference never undoes
            return self.decode_from_memory(start_token=start_token, end_token=end_token, max_len=max_len,
                                           **self._encode(model, images)).ids
        if method == "beam_search":
            model.eval()  # vit:329
            is_grid = "grid" in type(model).__name__.lower()
            if not is_grid:
                beam_size = 5
            memory = model.encoder(images)
            ids, lens, _ = self.beam_search_from_memory(memory, start_token, end_token, max_len, beam_size=beam_size,
                                                        stop_rule="grid" if is_grid else "vit")
            return ids[:, :int(lens.max().item())]
        raise ValueError(f"Unknown generation method: {method}")

    def sample_with_log_probs(self, model: torch.nn.Module, images: torch.Tensor, start_token: int,
                              end_token: int, max_len: int, device=None, seed: Optional[int] = None
                              ) -> Tuple[torch.Tensor, torch.Tensor]:
        """Drop-in for ``SCSTLoss._sample_with_log_probs`` (scst:202-254).

        The rollout samples from the eval-mode distribution on the engine.  When autograd is
        enabled the returned log-probs are recomputed by ONE teacher-forced pass of the reference
        decoder over the sampled ids (row k of that pass is the step-k distribution), so
        ``loss.backward()`` reaches decoder and encoder exactly as scst:183-191 expects.

        Stated deviation: the reference runs this loop under ``model.train()``
        (scst:161), i.e. it samples from and scores with a distribution whose prefix activations are
        re-dropped at every step.  A KV-cached engine cannot reproduce that; tokens come from the
        eval-mode distribution, and the differentiable scoring pass runs in whatever mode the caller
        left the model in (one dropout draw if it is in train mode; in ``eval()`` it reproduces the
        engine's own log-probs to the precision tolerance, which the GPU tests assert).  ``device`` is
        where the reference creates its token tensor (scst:216); results are returned there."""
        memory = model.encoder(images)  # scst:213
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())  # follows torch's global generator
        out = self.decode_from_memory(memory, start_token, end_token, max_len, group=1, n_sample=1, seed=seed)
        dst = torch.device(device) if device is not None else out.ids.device
        if not torch.is_grad_enabled():
            return out.ids.to(dst), out.logp.to(dst)
        ids = out.ids
        t = ids.size(1) - 1
        mask = model.decoder.generate_square_subsequent_mask(t, ids.device)
        logits = model.decoder(ids[:, :-1], memory, tgt_mask=mask)
        lp = torch.log_softmax(logits, dim=-1).gather(2, ids[:, 1:, None]).squeeze(2)
        ended = ids[:, 1:] == end_token
        finished_before = (ended.cumsum(1) - ended.long()) > 0  # finished strictly before this step
        return ids.to(dst), lp.masked_fill(finished_before, 0.0).to(dst)

    # ------------------------------------------------------------------ reward
    def cider_d(self, hyp_ids: torch.Tensor, ref_ids: torch.Tensor, ref_offsets, group: int = 1, start_token: int = -1,
                end_token: int = -1, pad_token: int = -1) -> torch.Tensor:
        """CIDEr-D of every hypothesis against its image's references on the device (C ABI ``capdec_cider_d``):
        ``hyp_ids`` (n_images*group, L) int64 image-major, ``ref_ids`` (n_refs, Lr) int64, ``ref_offsets`` n_images+1
        ints (host).  Sequences are read like scst:256-269 (stop at ``end_token``, skip ``start_token`` / ``pad_token``).
        Returns float64 (n_images*group,), what ``CiderRewardCalculator.compute_reward`` (scst:29-54) returns."""
        hyp = hyp_ids.to(self.device, torch.int64).contiguous()
        refs = ref_ids.to(self.device, torch.int64).contiguous()
        offs = [int(x) for x in ref_offsets]
        n_images = len(offs) - 1
        if hyp.dim() != 2 or hyp.shape[0] != n_images * group:
            raise ValueError(f"hyp_ids must be ({n_images * group}, L), got {tuple(hyp.shape)}")
        if refs.dim() != 2 or refs.shape[0] != offs[-1]:
            raise ValueError(f"ref_ids must be ({offs[-1]}, Lr), got {tuple(refs.shape)}")
        out = torch.empty(hyp.shape[0], dtype=torch.float64, device=self.device)
        off_arr = (C.c_int32 * len(offs))(*offs)
        args = _lib.CiderArgs(hyp_ids=hyp.data_ptr(), ref_ids=refs.data_ptr(), ref_offsets=off_arr, n_images=n_images,
                              group=group, hyp_len=hyp.shape[1], n_refs=refs.shape[0], ref_len=refs.shape[1],
                              start_token=start_token, end_token=end_token, pad_token=pad_token, scores_out=out.data_ptr())
        _lib.check(self._h, self._lib.capdec_cider_d(self._h, C.byref(args), _stream_ptr(self.device)))
        self._keep_cider = (hyp, refs)  # alive until the kernels have run (stream-ordered; replaced by the next call)
        return out

    # ------------------------------------------------------------------ diagnostics
    def profile(self, on: bool) -> None:
        """Per-kernel CUDA-event timing of the following rollouts (direct launches, no graph)."""
        _lib.check(self._h, self._lib.capdec_profile_enable(self._h, int(on)))

    def profile_report(self) -> dict:
        import json
        buf = C.create_string_buffer(1 << 16)
        _lib.check(self._h, self._lib.capdec_profile_report(self._h, buf, len(buf)))
        return json.loads(buf.value.decode())

    def launch_count(self) -> int:
        return int(self._lib.capdec_launch_count(self._h))

    def sample_logits(self, logits: torch.Tensor, mode: int, seed: int = 0, seq_id_base: int = 0, step: int = 0):
        logits = logits.to(self.device, torch.float32).contiguous()
        if logits.dim() != 2 or logits.shape[1] != self.vocab:
            raise ValueError("logits must be (rows, vocab)")
        ids = torch.empty(logits.shape[0], dtype=torch.int64, device=self.device)
        lp = torch.empty(logits.shape[0], dtype=torch.float32, device=self.device)
        _lib.check(self._h, self._lib.capdec_sample_logits(self._h, logits.data_ptr(), logits.shape[0], mode, seed,
                                                           seq_id_base, step, ids.data_ptr(), lp.data_ptr(),
                                                           _stream_ptr(self.device)))
        return ids, lp

    def debug_gemm(self, a: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False):
        a = a.to(self.device, torch.float32).contiguous()
        w = w.to(self.device, torch.float32).contiguous()
        bias = bias.to(self.device, torch.float32).contiguous() if bias is not None else None
        c = torch.empty((a.shape[0], w.shape[0]), dtype=torch.float32, device=self.device)
        _lib.check(self._h, self._lib.capdec_debug_gemm(self._h, a.data_ptr(), w.data_ptr(),
                                                        bias.data_ptr() if bias is not None else None, c.data_ptr(),
                                                        a.shape[0], w.shape[0], a.shape[1], int(relu),
                                                        _stream_ptr(self.device)))
        return c

    def debug_cross_kv(self, layer: int, which: int, n_images: int, mem_len: int) -> torch.Tensor:
        out = torch.empty((n_images, mem_len, self.d_model), dtype=torch.float32, device=self.device)
        _lib.check(self._h, self._lib.capdec_debug_cross_kv(self._h, layer, which, out.data_ptr(),
                                                            _stream_ptr(self.device)))
        return out

    def prefill_tokens(self, tokens: torch.Tensor, projection: torch.nn.Linear, first_token: int = 1) -> None:
        tokens = tokens.detach().to(self.device, torch.float32).contiguous()
        mem_len = tokens.shape[1] - first_token
        self._ensure_capacity(tokens.shape[0], 1, mem_len, 2)
        self.sync_projection(projection)
        self._keep_mem = tokens
        _lib.check(self._h, self._lib.capdec_prefill_tokens(self._h, tokens.data_ptr(), tokens.shape[0], tokens.shape[1],
                                                            first_token, mem_len, _stream_ptr(self.device)))

    def prefill(self, memory: torch.Tensor) -> None:
        memory = memory.detach().to(self.device, torch.float32).contiguous()
        self._ensure_capacity(memory.shape[0], 1, memory.shape[1], 2)
        self.sync_weights()
        self._keep_mem = memory
        _lib.check(self._h, self._lib.capdec_prefill(self._h, memory.data_ptr(), memory.shape[0], memory.shape[1],
                                                     _stream_ptr(self.device)))
