"""ctypes binding of libcram_b200.so (the C ABI declared in include/cram_b200.h).

There is no CPU fallback: if the library is missing and cannot be built, importing the ops fails
loudly. The library is built in-tree by ``cram_b200.build`` / ``__graft_entry__.build()``.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

_lock = threading.Lock()
_lib = None

CRAM_F32, CRAM_F16, CRAM_BF16 = 0, 1, 2
ABI_VERSION = 1

c_i32p = C.POINTER(C.c_int32)
vp = C.c_void_p


class AttnPrefillParams(C.Structure):
    """struct cram_attn_prefill_params"""
    _fields_ = [
        ("q", vp), ("k", vp), ("v", vp), ("out", vp), ("lse", vp),
        ("q_row_stride", C.c_int64), ("k_row_stride", C.c_int64), ("v_row_stride", C.c_int64),
        ("out_row_stride", C.c_int64),
        ("cu_seqlens_q", vp), ("cu_seqlens_k", vp),
        ("n_seq", C.c_int32), ("total_q", C.c_int32), ("total_k", C.c_int32), ("max_seqlen_q", C.c_int32),
        ("max_seqlen_k", C.c_int32),
        ("n_heads", C.c_int32), ("n_kv_heads", C.c_int32), ("head_dim", C.c_int32),
        ("dtype", C.c_int32), ("scale", C.c_float),
        ("logcred", vp), ("logcred_dtype", C.c_int32),
        ("head_mask", C.c_uint64),
        ("bias_len", vp),
        ("causal", C.c_int32), ("kernel", C.c_int32),
    ]


class AttnDecodeParams(C.Structure):
    """struct cram_attn_decode_params"""
    _fields_ = [
        ("q", vp), ("out", vp),
        ("q_row_stride", C.c_int64), ("out_row_stride", C.c_int64),
        ("k_cache", vp), ("v_cache", vp),
        ("page_stride", C.c_int64), ("token_stride", C.c_int64),
        ("block_table", vp), ("block_table_stride", C.c_int32),
        ("seqlens", vp),
        ("n_seq", C.c_int32), ("page_size", C.c_int32), ("max_seqlen", C.c_int32),
        ("n_heads", C.c_int32), ("n_kv_heads", C.c_int32), ("head_dim", C.c_int32),
        ("dtype", C.c_int32), ("scale", C.c_float),
        ("logcred", vp), ("logcred_stride", C.c_int64), ("logcred_dtype", C.c_int32),
        ("head_mask", C.c_uint64),
        ("bias_len", vp),
        ("n_splits", C.c_int32),
        ("workspace", vp), ("workspace_bytes", C.c_size_t),
    ]


# name -> (restype, argtypes); must list every symbol include/cram_b200.h declares.
SIGNATURES = {
    "cram_abi_version": (C.c_int, []),
    "cram_last_error": (C.c_char_p, []),
    "cram_span_index": (C.c_int, [vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, vp, vp, vp, vp, vp]),
    "cram_log_cred": (C.c_int, [vp, C.c_int64, C.c_int, vp, vp]),
    "cram_attn_prefill": (C.c_int, [C.POINTER(AttnPrefillParams), vp]),
    "cram_attn_decode_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "cram_attn_decode": (C.c_int, [C.POINTER(AttnDecodeParams), vp]),
    "cram_attn_decode_static": (C.c_int, [C.POINTER(AttnDecodeParams), vp]),
    "cram_kv_append": (C.c_int, [vp, vp, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, vp, vp, vp,
                                 C.c_int64, C.c_int64, C.c_int, vp]),
    "cram_sweep_splice_heads": (C.c_int, [vp, vp, vp, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, vp, vp]),
    "cram_sweep_answer_ie": (C.c_int, [vp, C.c_int, C.c_int64, C.c_int, vp, C.c_float, vp, vp, C.c_int, C.c_int,
                                       vp, vp, vp, vp]),
    "cram_rmsnorm": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_int, C.c_float, C.c_int, vp]),
    "cram_rope": (C.c_int, [vp, C.c_int64, C.c_int64, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp]),
    "cram_silu_mul": (C.c_int, [vp, vp, C.c_int64, C.c_int, C.c_int, vp]),
    "cram_linear_decode": (C.c_int, [vp, C.c_int64, C.c_int, vp, C.c_int64, C.c_int, C.c_int, vp, vp, C.c_int64, C.c_int,
                                     C.c_int, vp]),
    "cram_rope_kv_append": (C.c_int, [vp, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp,
                                      C.c_int64, C.c_int64, C.c_int, C.c_int, vp]),
    "cram_rank_heads": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp]),
}


# A/B-timing switches exported by the library but deliberately NOT part of include/cram_b200.h (engine.py reads them
# through the environment variables CRAM_B200_PDL / CRAM_B200_LINEAR_RING only).
DEBUG_SIGNATURES = {
    "cram_debug_pdl": (None, [C.c_int]),
    "cram_debug_linear_ring": (None, [C.c_int]),
}


class CramError(RuntimeError):
    pass


def library_path() -> str:
    return os.environ.get("CRAM_B200_LIB") or _build.LIB_PATH


def library_fingerprint() -> str:
    """sha256 of csrc/ + the header + the nvcc flags the loaded in-tree library was built from (bench.py records it)."""
    return _build.built_fingerprint()


def load(build_if_missing: bool = True):
    """Load (building first if needed) libcram_b200.so and set the ctypes signatures."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = os.environ.get("CRAM_B200_LIB")  # override: variant / debug builds (scripts/trace_attn.py)
        if not path:
            # The in-tree library must be the build of the sources next to it: a stale .so (sources edited, *.so is
            # git-ignored but travels with the snapshot) would silently run old kernels. build_library() is a cheap
            # fingerprint compare when nothing changed, a rebuild when nvcc is there, and an error otherwise.
            path = _build.LIB_PATH
            if os.path.exists(path) and not build_if_missing:
                if not _build.is_current():
                    raise CramError(f"{path} does not match cram_b200/csrc (stale build); run `python -m cram_b200.build`")
            elif not os.path.exists(path) and not build_if_missing:
                raise CramError(f"{path} is missing; run `python -m cram_b200.build` (no CPU fallback exists)")
            else:
                path = _build.build_library()
        elif not os.path.exists(path):
            raise CramError(f"CRAM_B200_LIB={path} does not exist")
        lib = C.CDLL(path)
        for name, (res, args) in list(SIGNATURES.items()) + list(DEBUG_SIGNATURES.items()):
            fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        got = lib.cram_abi_version()
        if got != ABI_VERSION:
            raise CramError(f"libcram_b200.so ABI version {got} != binding version {ABI_VERSION}; rebuild")
        _lib = lib
        return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().cram_last_error()
        raise CramError(f"{what} failed (status {rc}): {msg.decode() if msg else '?'}")
