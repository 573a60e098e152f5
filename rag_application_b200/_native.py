"""ctypes binding of libbrag.so -- one prototype per symbol declared in include/brag.h.

The library is the only compute path: if it is missing or fails to load this module raises;
there is no CPU fallback (BASELINE.json north_star)."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# BRAG_LIB: load another build of the same ABI (A/B timing of kernel variants in tools/); default = the in-tree build
LIB_PATH = os.environ.get("BRAG_LIB") or os.path.join(_HERE, "csrc", "libbrag.so")

c_p = C.c_void_p
i32, i64 = C.c_int32, C.c_int64

# name -> (restype, argtypes); must list every symbol of include/brag.h
PROTOTYPES = {
    "brag_create": (C.c_int, [C.c_int, C.POINTER(c_p)]),
    "brag_destroy": (C.c_int, [c_p]),
    "brag_last_error": (C.c_char_p, []),
    "brag_version": (C.c_int, []),
    "brag_set_id_offset": (C.c_int, [c_p, i64]),
    "brag_dense_set": (C.c_int, [c_p, C.c_char_p, c_p, i64, i32, i32, i32]),
    "brag_dense_append": (C.c_int, [c_p, C.c_char_p, c_p, i64, i32, i32, i32]),
    "brag_dense_update": (C.c_int, [c_p, C.c_char_p, c_p, c_p, i64]),
    "brag_dense_drop": (C.c_int, [c_p, C.c_char_p]),
    "brag_sparse_set_csr": (C.c_int, [c_p, C.c_char_p, c_p, c_p, c_p, i64, i64, i32]),
    "brag_sparse_drop": (C.c_int, [c_p, C.c_char_p]),
    "brag_sparse_df": (C.c_int, [c_p, C.c_char_p, c_p, i64]),
    "brag_mask_set": (C.c_int, [c_p, i32, c_p, i64, i32]),
    "brag_dense_topk": (C.c_int, [c_p, C.c_char_p, c_p, i32, i32, i32, c_p, c_p, c_p, i32, c_p]),
    "brag_dense_rescore": (C.c_int, [c_p, C.c_char_p, c_p, i32, c_p, i32, i32, i32, c_p, c_p, c_p, i32, c_p]),
    "brag_dense_rescore_noalloc": (C.c_int, [c_p, C.c_char_p, c_p, i32, c_p, i32, i32, i32, c_p, c_p, c_p, c_p]),
    "brag_sparse_topk": (C.c_int, [c_p, C.c_char_p, c_p, c_p, c_p, i32, i32, i32, i32, c_p, c_p, c_p, i32, c_p]),
    "brag_rrf": (C.c_int, [c_p, c_p, c_p, i32, i32, i32, i32, i32, c_p, c_p, c_p, i32, c_p]),
    "brag_merge_topk": (C.c_int, [c_p, c_p, c_p, c_p, i32, i32, i32, c_p, c_p, c_p, i32, c_p]),
    "brag_hybrid_topk": (C.c_int, [c_p, C.c_char_p, C.c_char_p, c_p, c_p, c_p, c_p, i32, i32, i32, i32, c_p, c_p, c_p, i32, c_p]),
    "brag_shard_local_lists": (C.c_int, [c_p, C.c_char_p, C.c_char_p, c_p, c_p, c_p, c_p, i32, i32, i32, i32, c_p, i32, c_p]),
    "brag_shard_fuse": (C.c_int, [c_p, c_p, i32, i32, i32, i32, i32, c_p, c_p, c_p, c_p, c_p, c_p, i32, c_p]),
    "brag_shard_window_create": (C.c_int, [c_p, i32, i32, i32, c_p]),
    "brag_shard_connect": (C.c_int, [c_p, i32, i32, c_p, c_p]),
    "brag_shard_hybrid_topk": (C.c_int, [c_p, C.c_char_p, C.c_char_p, c_p, c_p, c_p, c_p, i32, i32, i32, i32, c_p, c_p, c_p,
                                         c_p, c_p, c_p, i32, c_p]),
    "brag_shard_close": (C.c_int, [c_p]),
    "brag_save": (C.c_int, [c_p, C.c_char_p]),
    "brag_load": (C.c_int, [c_p, C.c_char_p]),
    "brag_profile_enable": (C.c_int, [c_p, i32]),
    "brag_profile_read": (C.c_int, [c_p, c_p, c_p, c_p]),
    "brag_profile_reset": (C.c_int, [c_p]),
    "brag_dense_path": (C.c_int, [c_p, C.c_char_p, i32]),
    "brag_force_dense_path": (C.c_int, [c_p, i32]),
    "brag_force_sparse_path": (C.c_int, [c_p, i32]),
    "brag_sparse_compact_after": (C.c_int, [c_p, i64]),
    "brag_sparse_offsets_min_df": (C.c_int, [c_p, i64]),
    "brag_sparse_compact_info": (C.c_int, [c_p, C.c_char_p, c_p, c_p]),
    "brag_rerank_meanpool": (C.c_int, [c_p, c_p, i32, c_p, i32, i32, i32, c_p, c_p, i32, c_p]),
    "brag_sparse_append": (C.c_int, [c_p, C.c_char_p, c_p, c_p, c_p, c_p, i64, i64]),
    "brag_bm25_encode": (C.c_int, [C.c_char_p, c_p, i32, C.c_double, C.c_double, C.c_double, i32, i32, c_p, c_p, c_p]),
    "brag_bm25_stem": (C.c_int, [C.c_char_p, C.c_char_p, i32]),
    "brag_murmur3_32": (C.c_uint32, [C.c_char_p, i32, C.c_uint32]),
}

_lib = None


class NativeLibraryMissing(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load libbrag.so and attach prototypes.  Raises (never falls back) when absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeLibraryMissing(
            f"{LIB_PATH} not built. Run `python -m rag_application_b200.build` "
            "(needs nvcc; targets sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class BragError(RuntimeError):
    pass


def check(rc: int) -> None:
    if rc != 0:
        msg = load().brag_last_error()
        raise BragError(msg.decode("utf-8", "replace") if msg else f"libbrag error {rc}")
