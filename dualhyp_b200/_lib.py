"""ctypes binding of libdualhyp.so (include/dualhyp.h) -- the thin stub a `ger` maintainer would add.

There is no CPU fallback anywhere in this package: if the shared library is missing, or a compute entry is
called without an sm_100 device, the call raises.  Loading the library (and listing its symbols) works on
a machine without a GPU; `dh_init` is what requires the device.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, os.environ.get("DUALHYP_LIB", "libdualhyp.so"))   # DUALHYP_LIB: developer override (trace builds)

EPI_BF16, EPI_F32, EPI_RESID_F32, EPI_RESID_BF16, EPI_QKV_ROPE, EPI_SWIGLU, EPI_SWIGLU_BWD = range(7)


class DualHypError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    """Mirror of `dh_gemm_args` (include/dualhyp.h)."""
    _fields_ = [
        ("A", C.c_void_p), ("lda", C.c_int64),
        ("B", C.c_void_p), ("ldb", C.c_int64),
        ("A2", C.c_void_p), ("lda2", C.c_int64),
        ("B2", C.c_void_p), ("ldb2", C.c_int64),
        ("M", C.c_int32), ("N", C.c_int32), ("K1", C.c_int32), ("K2", C.c_int32),
        ("a_layout", C.c_int32), ("b_layout", C.c_int32), ("a2_layout", C.c_int32), ("b2_layout", C.c_int32),
        ("epilogue", C.c_int32), ("accumulate", C.c_int32),
        ("alpha", C.c_float),
        ("seq_len", C.c_int32), ("q_per_kv", C.c_int32), ("split_k", C.c_int32),
        ("out", C.c_void_p), ("ldo", C.c_int64),
        ("out2", C.c_void_p), ("ldo2", C.c_int64),
        ("aux", C.c_void_p), ("ldaux", C.c_int64),
        ("col_scale", C.c_void_p), ("col_bias", C.c_void_p),
        ("rope_cos", C.c_void_p), ("rope_sin", C.c_void_p),
        ("positions", C.c_void_p),
        ("m_limit", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
    ]


class DecodeLayer(C.Structure):
    """Mirror of `dh_decode_layer` (include/dualhyp.h)."""
    _fields_ = [("wqkv", C.c_void_p), ("wproj", C.c_void_p), ("w12", C.c_void_p), ("wmlp", C.c_void_p),
                ("norm2", C.c_void_p), ("norm_next", C.c_void_p), ("cache_k", C.c_void_p), ("cache_v", C.c_void_p),
                ("prefix_kv", C.c_void_p), ("gate", C.c_void_p)]


class DecodeDesc(C.Structure):
    """Mirror of `dh_decode_desc` (include/dualhyp.h)."""
    _fields_ = [("n_layer", C.c_int32), ("n_seq", C.c_int32), ("C", C.c_int32), ("I", C.c_int32), ("n_groups", C.c_int32),
                ("q_per_kv", C.c_int32), ("max_len", C.c_int32), ("prefix_len", C.c_int32),
                ("eps", C.c_float),
                ("x", C.c_void_p), ("x2", C.c_void_p), ("n", C.c_void_p), ("qkv", C.c_void_p), ("y", C.c_void_p), ("act", C.c_void_p),
                ("hidden", C.c_void_p), ("ws", C.c_void_p), ("ws_bytes", C.c_size_t),
                ("positions", C.c_void_p), ("rope_cos", C.c_void_p), ("rope_sin", C.c_void_p),
                ("barrier", C.c_void_p), ("layers", C.POINTER(DecodeLayer)), ("trace", C.c_void_p)]


_P, _I, _F, _L, _Z = C.c_void_p, C.c_int, C.c_float, C.c_int64, C.c_size_t

# name -> (restype, argtypes); every symbol include/dualhyp.h declares
SIGNATURES = {
    "dh_abi_version": (_I, []),
    "dh_last_error": (C.c_char_p, []),
    "dh_init": (_I, [_I]),
    "dh_sm_count": (_I, []),
    "dh_gemm": (_I, [C.POINTER(GemmArgs), _P]),
    "dh_gemm_streamk_workspace_bytes": (_Z, [_I, _I, _I]),
    "dh_gemm_swiglu_streamk_noalloc": (_I, [C.POINTER(GemmArgs), _P]),
    "dh_rmsnorm_fwd": (_I, [_P, _I, _P, _P, _P, _I, _I, _F, _P]),
    "dh_splitk_resid_rmsnorm": (_I, [_P, _I, _L, _P, _P, _P, _P, _I, _I, _F, _P]),
    "dh_splitk_qkv_rope_cache": (_I, [_P, _I, _L, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "dh_rmsnorm_bwd": (_I, [_P, _I, _P, _P, _P, _P, _P, _P, _I, _I, _P]),
    "dh_embedding_fwd": (_I, [_P, _P, _I, _P, _I, _I, _I, _I, _P]),
    "dh_rope_apply": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dh_rope_qkv_inplace": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dh_attention_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dh_attention_fwd_varlen": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _L, _I, _I, _I, _P]),
    "dh_attention_bwd_workspace": (_Z, [_I, _I, _I, _I]),
    "dh_attention_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _P, _Z, _P]),
    "dh_attention_bwd_varlen": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _P, _P, _I, _I, _L, _I, _I, _I, _P, _Z, _P]),
    "dh_attention_decode": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dh_attention_decode_stream": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dh_attention_decode_rows": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dh_decode_plan_bytes": (_Z, [_I]),
    "dh_decode_workspace_bytes": (_Z, [C.POINTER(DecodeDesc)]),
    "dh_decode_plan_build": (_I, [C.POINTER(DecodeDesc), _P]),
    "dh_decode_layers": (_I, [_P, _P, _P]),
    "dh_prefix_project_fwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dh_prefix_project_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dh_kv_cache_write": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "dh_cross_entropy": (_I, [_P, _L, _P, _P, _I, _I, _F, _I, _P]),
    "dh_compact_targets": (_I, [_P, _I, _P, _P, _P, _P]),
    "dh_gather_rows": (_I, [_P, _P, _P, _P, _I, _I, _P]),
    "dh_scatter_rows": (_I, [_P, _P, _P, _P, _I, _I, _P]),
    "dh_loss_reduce": (_I, [_P, _P, _P, _I, _P]),
    "dh_argmax": (_I, [_P, _L, _P, _I, _I, _P]),
    "dh_relpred_fwd": (_I, [_P] * 11 + [_I, _I, _I, _I, _P]),
    "dh_relpred_bwd": (_I, [_P] * 15 + [_I, _I, _I, _I, _P]),
    "dh_cast_f32_to_bf16": (_I, [_P, _L, _P, _L, _L, _L, _P]),
    "dh_relpred_head_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dh_relpred_head_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dh_dropout_bf16": (_I, [_P, _P, _L, _F, C.c_uint64, C.c_uint64, _I, _P]),
    "dh_adamw_step": (_I, [_P, _P, _P, _P, _L, _F, _F, _F, _F, _F, _I, _F, _P]),
}

_lib: Optional[C.CDLL] = None
_initialised_device: Optional[int] = None


def load() -> C.CDLL:
    """dlopen the library and type every entry point.  Raises if the .so has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise DualHypError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C dualhyp_b200/csrc`). dualhyp_b200 has no CPU or PyTorch fallback path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().dh_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise DualHypError(f"{what} failed ({rc}): {last_error()}")


def init(device: int = 0) -> None:
    """Per-device init; raises on machines without an sm_100 GPU (no fallback)."""
    global _initialised_device
    if _initialised_device == device:
        return
    check(load().dh_init(device), "dh_init")
    _initialised_device = device


def source_id() -> str:
    """Identity of the device code for evidence files (profiles/traffic.json): sha256 over the kernel sources with comments and
    white space removed.  The library file itself is not hashed: nvcc does not write byte-identical files from identical sources, and
    build() recompiles on every fresh checkout."""
    import glob
    import hashlib
    import re
    root = os.path.dirname(_HERE)
    files = sorted(glob.glob(os.path.join(_HERE, "csrc", "*.cu")) + glob.glob(os.path.join(_HERE, "csrc", "*.cuh"))) + \
        [os.path.join(_HERE, "csrc", "Makefile"), os.path.join(root, "include", "dualhyp.h")]
    h = hashlib.sha256()
    for f in files:
        text = open(f, "r", encoding="utf-8", errors="replace").read()
        if not f.endswith("Makefile"):
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            text = re.sub(r"//[^\n]*", "", text)
        h.update(os.path.basename(f).encode())
        h.update(re.sub(r"\s+", "", text).encode())
    return h.hexdigest()[:16]
