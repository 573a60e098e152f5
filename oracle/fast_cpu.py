"""Fast CPU implementation of the three-stage path, for TIMING the CPU baseline at sizes the pure
numpy oracle cannot reach.  TEST INFRASTRUCTURE ONLY (oracle/__init__.py); PARITY UNPINNED.

Same results as oracle.hybrid_simple, engineered fairly ("fast CPU path"):
multithreaded BLAS sgemm + argpartition for the dense stage, the C inverted index
(oracle/bm25_cpu.c, pthreads) for BM25, and the oracle's RRF."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np

from . import oracle as O

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libbm25_cpu.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.run(["make", "-C", _HERE, "-s"], check=True)
        _LIB = C.CDLL(path)
        _LIB.bm25_topk_batch.restype = C.c_int
        _LIB.bm25_topk_batch.argtypes = [C.c_void_p] * 3 + [C.c_int64, C.c_int64] + [C.c_void_p] * 3 + \
            [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
    return _LIB


def dense_topk_batch(x_stored: np.ndarray, q_stored: np.ndarray, k: int, mask: Optional[np.ndarray] = None,
                     block: int = 1 << 18) -> Tuple[np.ndarray, np.ndarray]:
    """float32 sgemm over row blocks + argpartition, then exact (score desc, id asc) ordering of the
    survivors with float64 re-scoring (so the result equals oracle.dense_topk)."""
    n = x_stored.shape[0]
    B = q_stored.shape[0]
    kk = min(n, k + 32)
    best_s = np.full((B, 0), -np.inf, np.float32)
    best_i = np.zeros((B, 0), np.int64)
    for s in range(0, n, block):
        e = min(n, s + block)
        sc = q_stored @ x_stored[s:e].T                         # [B, e-s] BLAS
        if mask is not None:
            sc[:, ~mask[s:e]] = -np.inf
        m = min(kk, e - s)
        part = np.argpartition(-sc, m - 1, axis=1)[:, :m]
        best_s = np.concatenate([best_s, np.take_along_axis(sc, part, 1)], 1)
        best_i = np.concatenate([best_i, part + s], 1)
        if best_s.shape[1] > 4 * kk:
            keep = np.argpartition(-best_s, kk - 1, axis=1)[:, :kk]
            best_s, best_i = np.take_along_axis(best_s, keep, 1), np.take_along_axis(best_i, keep, 1)
    out_s = np.full((B, k), -np.inf, np.float32)
    out_i = np.full((B, k), -1, np.int64)
    for b in range(B):
        ok = np.isfinite(best_s[b])
        ids = best_i[b][ok]
        keep = np.argsort(-best_s[b][ok], kind="stable")[:kk]
        ids = ids[keep]
        ex = O.dense_scores(x_stored[ids], q_stored[b])
        o = O.topk_order(ex, ids, k)
        out_s[b, :len(o)], out_i[b, :len(o)] = ex[o], ids[o]
    return out_s, out_i


def sparse_topk_batch(term_ptr, doc, w, n_docs, q_ptr, q_terms, q_vals, k, mask=None, threads: int = 0):
    term_ptr = np.ascontiguousarray(term_ptr, np.int64)
    doc = np.ascontiguousarray(doc, np.uint32)
    w = np.ascontiguousarray(w, np.float32)
    q_ptr = np.ascontiguousarray(q_ptr, np.int32)
    q_terms = np.ascontiguousarray(q_terms, np.int32)
    q_vals = np.ascontiguousarray(q_vals, np.float64)
    B = len(q_ptr) - 1
    s = np.empty((B, k), np.float32)
    i = np.empty((B, k), np.int64)
    c = np.empty((B,), np.int32)
    m8 = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    rc = _lib().bm25_topk_batch(term_ptr.ctypes.data, doc.ctypes.data, w.ctypes.data, len(term_ptr) - 1, n_docs,
                                q_ptr.ctypes.data, q_terms.ctypes.data, q_vals.ctypes.data, B, k,
                                None if m8 is None else m8.ctypes.data, s.ctypes.data, i.ctypes.data, c.ctypes.data,
                                threads or (os.cpu_count() or 1))
    if rc != 0:
        raise MemoryError("bm25_topk_batch: allocation failed")
    return s, i, c


def hybrid_batch(x_stored, q_stored, term_ptr, doc, w, n_docs, q_ptr, q_terms, q_vals, k, threads: int = 0):
    ds, di = dense_topk_batch(x_stored, q_stored, k)
    ss, si, sc = sparse_topk_batch(term_ptr, doc, w, n_docs, q_ptr, q_terms, q_vals, k, threads=threads)
    out = []
    for b in range(q_stored.shape[0]):
        out.append(O.rrf([di[b][di[b] >= 0], si[b][:sc[b]]], limit=k))
    return out
