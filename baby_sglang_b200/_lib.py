"""ctypes binding of libbsgl_b200.so (the C ABI in include/bsgl_b200.h).

ctypes, not pybind: the library stays independent of torch's C++ ABI, and `CDLL` calls drop
the GIL, which the reference's overlap worker thread relies on (overlap_worker.py:104-106).
Loading fails loudly: there is no fallback implementation.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int64, c_void_p

from . import LIB_PATH

BF16, F16, F32 = 0, 1, 2
TILE_TOKENS = 16
PLAN_HDR_INTS = 4
MIN_TILES_PER_CTA = 16     # BSGL_MIN_TILES_PER_CTA in the header (bsgl_min_tiles_per_cta() gives the effective value)

_i = c_int
_l = c_int64
_p = c_void_p

# name -> (restype, argtypes); mirrors include/bsgl_b200.h one to one
SIGNATURES = {
    "bsgl_version": (c_int, []),
    "bsgl_last_error": (c_char_p, []),
    "bsgl_sm_count": (c_int, [POINTER(c_int)]),
    "bsgl_heads_per_group": (c_int, [_i]),
    "bsgl_min_tiles_per_cta": (c_int, []),
    "bsgl_kv_store": (c_int, [_p, _p, _l, _p, _p, _p, _l, _i, _i, _l, _l, _l, _l, _i, _i, _p]),
    "bsgl_build_decode_metadata": (c_int, [_l, _p, _p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "bsgl_build_extend_metadata": (c_int, [_l, _p, _p, _p, _i, _l, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "bsgl_gather_kv_indices": (c_int, [_p, _l, _p, _p, _i, _p, _p]),
    "bsgl_partial_sizes": (c_int, [_i, _i, _i, _i, POINTER(c_int64), POINTER(c_int64)]),
    "bsgl_paged_attention": (c_int, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, c_float, _i, _p, _p, _p, _i, _p]),
    "bsgl_merge_partials": (c_int, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p]),
    "bsgl_static_scratch_bytes": (c_int, []),
    "bsgl_paged_attention_static": (c_int, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, c_float, _i, _p, _i, _p]),
    "bsgl_rope_kv_store": (c_int, [_p, _l, _p, _p, _p, _p, _p, _l, _p, _l, _i, _i, _i, _i, _p]),
    "bsgl_greedy_sample": (c_int, [_p, _l, _i, _i, _p, _i, _p, _p, _p, _p, _p]),
    "bsgl_extend_attention": (c_int, [_p, _p, _p, _p, _l, _p, _p, _p, _p, _p, _i, _l, _i, _i, _i, c_float, _i, _p, _p, _p]),
    "bsgl_page_table_write_rows": (c_int, [_p, _l, _l, _i, _p, _i, _p, _p, _p]),
    "bsgl_page_table_append": (c_int, [_p, _l, _l, _i, _p, _p, _p, _i, _p]),
    "bsgl_slot_ring_pop": (c_int, [_p, _l, _l, _l, _p, _p]),
    "bsgl_slot_ring_push": (c_int, [_p, _l, _l, _l, _p, _l, _i, _p]),
    "bsgl_comm_alloc": (c_int, [_l, POINTER(c_void_p), _p]),
    "bsgl_comm_open": (c_int, [_p, POINTER(c_void_p)]),
    "bsgl_comm_close": (c_int, [_p]),
    "bsgl_comm_free": (c_int, [_p]),
    "bsgl_oproj_flag_bytes": (c_int, [_i, _i, POINTER(c_int64)]),
    "bsgl_oproj_allreduce": (c_int, [_p, _l, _p, _l, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _i, _i, _l,
                                      ctypes.c_double, _i, _p]),
}

_lib = None


class BsglError(RuntimeError):
    pass


def load():
    """Load (once) and return the CDLL.  Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise BsglError(
            f"{LIB_PATH} is missing: build it with `python -m baby_sglang_b200.build` "
            "(nvcc, sm_100a).  baby_sglang_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the .so is stale w.r.t. the header
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().bsgl_last_error().decode("utf-8", "replace")
        raise BsglError(f"{what} failed (code {rc}): {msg}")
