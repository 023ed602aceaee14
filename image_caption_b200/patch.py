"""Install the engine behind the reference's own call surface.

``install(model)`` rebinds ``model.generate`` (models/vit_transformer_model.py:276,
models/grid_transformer_model.py:222) so that the reference's scripts, validation loops and
``SCSTLoss.forward`` (utils/scst_loss.py:169-172) reach the B200 path unchanged;
``install_scst(scst_loss)`` rebinds ``SCSTLoss._sample_with_log_probs`` (utils/scst_loss.py:202).
``uninstall`` restores the originals.  The reference files are not edited.
"""
from __future__ import annotations

import types
from typing import Optional

import torch

from .engine import CaptionDecodeEngine


def engine_for(model: torch.nn.Module, **engine_kwargs) -> CaptionDecodeEngine:
    """One engine per model instance, created on first use from ``model.decoder`` (which must then be on a CUDA
    device: there is no CPU path, ``CaptionDecodeEngine`` raises)."""
    eng: Optional[CaptionDecodeEngine] = getattr(model, "_capdec_engine", None)
    if eng is None:
        kw = dict(getattr(model, "_capdec_engine_kwargs", {}))
        kw.update(engine_kwargs)
        eng = CaptionDecodeEngine(model.decoder, **kw)
        object.__setattr__(model, "_capdec_engine", eng)  # not a submodule / parameter
    return eng


def install(model: torch.nn.Module, **engine_kwargs) -> Optional[CaptionDecodeEngine]:
    """Rebind ``model.generate`` (vit:276 / grid:222; same positional order, same defaults; the ViT model's
    signature has no ``beam_size`` and ignores it, vit:291).  The engine is created now when the decoder already
    lives on a CUDA device, else on the first call (scripts move the model with ``.to(device)`` after building it)."""
    if getattr(model, "_capdec_reference_generate", None) is None:
        object.__setattr__(model, "_capdec_engine_kwargs", dict(engine_kwargs))
        object.__setattr__(model, "_capdec_reference_generate", model.generate)

        def generate(self, images, start_token, end_token, max_len=50, method="greedy", beam_size=5):
            return engine_for(self).generate(self, images, start_token, end_token, max_len, method, beam_size)

        object.__setattr__(model, "generate", types.MethodType(generate, model))
    on_cuda = any(p.is_cuda for p in model.decoder.parameters())
    return engine_for(model) if on_cuda else None


def uninstall(model: torch.nn.Module) -> None:
    if getattr(model, "_capdec_reference_generate", None) is not None:
        object.__delattr__(model, "generate")
        object.__delattr__(model, "_capdec_reference_generate")
        object.__delattr__(model, "_capdec_engine_kwargs")
    eng = getattr(model, "_capdec_engine", None)
    if eng is not None:
        eng.close()
        object.__delattr__(model, "_capdec_engine")


def install_scst(scst_loss, **engine_kwargs) -> None:
    """Rebind ``scst_loss._sample_with_log_probs(model, images, start, end, max_len, device)``."""
    if getattr(scst_loss, "_capdec_reference_sample", None) is not None:
        return
    object.__setattr__(scst_loss, "_capdec_reference_sample", scst_loss._sample_with_log_probs)

    def _sample_with_log_probs(self, model, images, start_token, end_token, max_len, device):
        return engine_for(model, **engine_kwargs).sample_with_log_probs(model, images, start_token, end_token,
                                                                         max_len, device)

    object.__setattr__(scst_loss, "_sample_with_log_probs", types.MethodType(_sample_with_log_probs, scst_loss))


def uninstall_scst(scst_loss) -> None:
    if getattr(scst_loss, "_capdec_reference_sample", None) is not None:
        object.__delattr__(scst_loss, "_sample_with_log_probs")
        object.__delattr__(scst_loss, "_capdec_reference_sample")
