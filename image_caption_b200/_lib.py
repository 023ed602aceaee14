"""ctypes binding of libcapdec.so (include/capdec.h).  No fallback: if the library is missing
or a call fails, an exception is raised."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CAPDEC_LIB") or os.path.join(_HERE, "libcapdec.so")  # CAPDEC_LIB: A/B builds (tuning aid)

ABI_VERSION = 1
MAX_LAYERS = 12
PREC_FP32, PREC_BF16, PREC_TF32 = 0, 1, 2
FLAG_NO_GRAPH, FLAG_NO_UMMA, FLAG_V1_STEP, FLAG_NO_PDL = 1, 2, 4, 8
FLAG_BATCH_INVARIANT = 32
DECODE_NO_EARLY_STOP = 1
DECODE_TEACHER_FORCED = 2
BEAM_STOP_GRID = 1

EXPORTS = (
    "capdec_abi_version", "capdec_last_error", "capdec_create", "capdec_destroy", "capdec_set_weights",
    "capdec_prefill", "capdec_decode", "capdec_generate_host", "capdec_sample_logits", "capdec_launch_count",
    "capdec_debug_gemm", "capdec_debug_cross_kv", "capdec_profile_enable", "capdec_profile_report",
    "capdec_beam_search", "capdec_set_memory_projection", "capdec_prefill_tokens", "capdec_gemm_static", "capdec_cider_d",
    "capdec_forward_full", "capdec_prefetch_host",
)


class CapdecError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"capdec error {code}: {message}")
        self.code = code


class Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "vocab", "d_model", "n_head", "n_layer", "d_ff", "pe_rows", "max_mem", "max_images", "max_group",
        "max_len", "precision", "flags")]


_fp = C.c_void_p  # device pointers travel as integers


class LayerWeights(C.Structure):
    _fields_ = [(n, _fp) for n in (
        "self_in_w", "self_in_b", "self_out_w", "self_out_b", "cross_in_w", "cross_in_b", "cross_out_w",
        "cross_out_b", "lin1_w", "lin1_b", "lin2_w", "lin2_b")] + [("norm_w", _fp * 3), ("norm_b", _fp * 3)]


class Weights(C.Structure):
    _fields_ = [("embedding", _fp), ("pe", _fp), ("fc_out_w", _fp), ("fc_out_b", _fp),
                ("layer", LayerWeights * MAX_LAYERS)]


class DecodeArgs(C.Structure):
    _fields_ = [("group", C.c_int32), ("n_sample", C.c_int32), ("start_token", C.c_int32),
                ("end_token", C.c_int32), ("pad_token", C.c_int32), ("max_len", C.c_int32),
                ("flags", C.c_int32), ("reserved", C.c_int32),
                ("seed", C.c_uint64), ("seq_id_base", C.c_uint64), ("ids_out", _fp), ("logp_out", _fp),
                ("logits_out", _fp), ("steps_out", C.POINTER(C.c_int32))]


class BeamArgs(C.Structure):
    _fields_ = [("beam_size", C.c_int32), ("flags", C.c_int32), ("start_token", C.c_int32), ("end_token", C.c_int32),
                ("pad_token", C.c_int32), ("max_len", C.c_int32), ("ids_out", _fp), ("len_out", _fp),
                ("score_out", _fp), ("steps_out", C.POINTER(C.c_int32))]


class ForwardArgs(C.Structure):
    _fields_ = [("group", C.c_int32), ("max_len", C.c_int32), ("end_token", C.c_int32), ("reserved", C.c_int32),
                ("ids", _fp), ("logp_out", _fp), ("logits_out", _fp)]


class CiderArgs(C.Structure):
    _fields_ = [("hyp_ids", _fp), ("ref_ids", _fp), ("ref_offsets", C.POINTER(C.c_int32)),
                ("n_images", C.c_int32), ("group", C.c_int32), ("hyp_len", C.c_int32), ("n_refs", C.c_int32),
                ("ref_len", C.c_int32), ("start_token", C.c_int32), ("end_token", C.c_int32), ("pad_token", C.c_int32),
                ("scores_out", _fp)]


_lib = None


def load() -> C.CDLL:
    """Load libcapdec.so from the package directory (built in-tree by __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python __graft_entry__.py build` "
            "(nvcc, sm_100a). There is no CPU or PyTorch fallback for the decode path.")
    lib = C.CDLL(LIB_PATH)
    ab_build = bool(os.environ.get("CAPDEC_LIB"))  # an older A/B build may lack the newest entry points
    for name in EXPORTS:
        if not hasattr(lib, name) and not ab_build:
            raise ImportError(f"{LIB_PATH} does not export {name}")
    H = C.c_void_p
    lib.capdec_abi_version.restype = C.c_int
    lib.capdec_last_error.restype = C.c_char_p
    lib.capdec_last_error.argtypes = [H]
    lib.capdec_create.argtypes = [C.POINTER(Config), C.c_int, C.POINTER(H)]
    lib.capdec_destroy.argtypes = [H]
    lib.capdec_set_weights.argtypes = [H, C.POINTER(Weights), C.c_void_p]
    lib.capdec_prefill.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    lib.capdec_decode.argtypes = [H, C.POINTER(DecodeArgs), C.c_void_p]
    lib.capdec_generate_host.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(DecodeArgs), C.c_void_p]
    lib.capdec_sample_logits.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32, C.c_uint64, C.c_uint64, C.c_int32,
                                         C.c_void_p, C.c_void_p, C.c_void_p]
    lib.capdec_beam_search.argtypes = [H, C.POINTER(BeamArgs), C.c_void_p]
    if hasattr(lib, "capdec_prefill_tokens"):
        lib.capdec_set_memory_projection.argtypes = [H, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p),
                                                     C.POINTER(C.c_void_p), C.c_void_p]
        lib.capdec_prefill_tokens.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    if hasattr(lib, "capdec_prefetch_host"):
        lib.capdec_prefetch_host.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32]
    if hasattr(lib, "capdec_forward_full"):
        lib.capdec_forward_full.argtypes = [H, C.POINTER(ForwardArgs), C.c_void_p]
    lib.capdec_launch_count.restype = C.c_int64
    lib.capdec_launch_count.argtypes = [H]
    lib.capdec_debug_gemm.argtypes = [H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_int32, C.c_void_p]
    if hasattr(lib, "capdec_gemm_static"):
        lib.capdec_gemm_static.argtypes = [H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                           C.c_int32, C.c_int32, C.c_void_p]
    if hasattr(lib, "capdec_cider_d"):
        lib.capdec_cider_d.argtypes = [H, C.POINTER(CiderArgs), C.c_void_p]
    lib.capdec_debug_cross_kv.argtypes = [H, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
    lib.capdec_profile_enable.argtypes = [H, C.c_int32]
    lib.capdec_profile_report.argtypes = [H, C.c_char_p, C.c_int32]
    for name in EXPORTS:
        if name not in ("capdec_last_error", "capdec_launch_count") and hasattr(lib, name):
            getattr(lib, name).restype = C.c_int
    if lib.capdec_abi_version() != ABI_VERSION:
        raise ImportError(f"libcapdec ABI {lib.capdec_abi_version()} != binding {ABI_VERSION}")
    _lib = lib
    return lib


def check(handle, rc: int) -> None:
    if rc != 0:
        msg = load().capdec_last_error(handle)
        raise CapdecError(rc, msg.decode() if msg else "unknown")
