"""image_caption_b200 - B200-native batched caption decoding for wonder-dream/image_caption.

The package holds only what the hot path needs: the CUDA kernels and C ABI (``csrc/``,
``libcapdec.so``), the host-side mirror of the reference's decode entry points (``engine``),
the monkey-patch installer (``patch``) and the data-parallel gather (``dist``).
Importing the package does not load the native library; constructing an engine does, and
fails loudly if it is missing.
"""
from .engine import CaptionDecodeEngine, Rollout  # noqa: F401
from .patch import install, install_scst, uninstall, uninstall_scst, engine_for  # noqa: F401

__all__ = ["CaptionDecodeEngine", "Rollout", "install", "install_scst", "uninstall", "uninstall_scst", "engine_for"]
