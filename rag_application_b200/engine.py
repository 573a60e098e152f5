"""Python handle over the C ABI (include/brag.h).  Thin: validates shapes, picks host vs device
pointers, and returns numpy arrays (host calls) or torch CUDA tensors (device calls).

Host calls (numpy in / numpy out) are the reference-facing path: host->device and
device->host copies happen inside the call.  Device calls (torch CUDA tensors in/out) are
stream-ordered on torch's current stream and never synchronise."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _native

DTYPES = {"f32": 0, "f16": 1, "bf16": 2, "i8": 3}
STAGES = ("dense_scan", "dense_select", "sparse_scan", "sparse_select", "rrf", "merge")
TIE_ID, TIE_FIRST_SEEN = 0, 1


def to_storage(x: np.ndarray, dtype: str) -> np.ndarray:
    """float32 rows -> storage bits (f32: float32, f16: float16, bf16: uint16 RNE bits, i8: int8 -- the rows of an
    int8 space are the integer values themselves, e.g. the reference's `quantized` vector, qdrant_handler.py:144-146)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    if dtype == "i8":
        r = np.rint(x)
        if not (np.array_equal(r, x) and r.min(initial=0) >= -128 and r.max(initial=0) <= 127):
            raise ValueError("int8 storage takes integer values in [-128, 127]")
        return r.astype(np.int8)
    if dtype == "f32":
        return x
    if dtype == "f16":
        return x.astype(np.float16)
    if dtype == "bf16":
        u = x.view(np.uint32).astype(np.uint64)
        return (((u + 0x7FFF + ((u >> 16) & 1)) >> 16) & 0xFFFF).astype(np.uint16)
    raise ValueError(f"unknown storage dtype {dtype!r}")


def _is_torch_cuda(t) -> bool:
    return hasattr(t, "data_ptr") and hasattr(t, "is_cuda") and t.is_cuda


def _ptr(a):
    if a is None:
        return None
    if _is_torch_cuda(a):
        return a.data_ptr()
    return a.ctypes.data


def _stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


class Engine:
    """One GPU, one corpus shard.  Fails loudly when libbrag.so or a B200 is missing."""

    def __init__(self, device: int = 0):
        self._lib = _native.load()
        h = C.c_void_p()
        _native.check(self._lib.brag_create(device, C.byref(h)))
        self._h = h
        self.device = device
        self._spaces = {}

    def close(self):
        if getattr(self, "_h", None):
            self._lib.brag_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ corpus
    def set_id_offset(self, offset: int):
        _native.check(self._lib.brag_set_id_offset(self._h, int(offset)))

    def dense_set(self, name: str, rows, dtype: str = "f32"):
        """rows: numpy float32 [n,d] (converted to `dtype` on the host) or a torch CUDA tensor
        already in the storage dtype (float32 / float16 / bfloat16)."""
        if _is_torch_cuda(rows):
            import torch
            tmap = {torch.float32: "f32", torch.float16: "f16", torch.bfloat16: "bf16", torch.int8: "i8"}
            dtype = tmap[rows.dtype]
            rows = rows.contiguous()
            n, d = rows.shape
            torch.cuda.current_stream().synchronize()
            _native.check(self._lib.brag_dense_set(self._h, name.encode(), rows.data_ptr(), n, d, DTYPES[dtype], 1))
        else:
            rows = np.asarray(rows)
            n, d = rows.shape
            st = to_storage(rows, dtype)
            _native.check(self._lib.brag_dense_set(self._h, name.encode(), st.ctypes.data, n, d, DTYPES[dtype], 0))
        self._spaces[name] = (n, d, dtype)

    def dense_append(self, name: str, rows):
        """Append float32 [m,d] rows (host) to an existing space, converted to its storage dtype."""
        n, d, dtype = self._spaces[name]
        rows = np.asarray(rows)
        assert rows.ndim == 2 and rows.shape[1] == d
        st = to_storage(rows, dtype)
        _native.check(self._lib.brag_dense_append(self._h, name.encode(), st.ctypes.data, rows.shape[0], d, DTYPES[dtype], 0))
        self._spaces[name] = (n + rows.shape[0], d, dtype)

    def dense_update(self, name: str, row_ids, rows):
        """Overwrite rows `row_ids` (local) with float32 [m,d] rows (host)."""
        n, d, dtype = self._spaces[name]
        ids = np.ascontiguousarray(row_ids, dtype=np.int64)
        rows = np.asarray(rows)
        assert rows.shape == (len(ids), d)
        st = to_storage(rows, dtype)
        _native.check(self._lib.brag_dense_update(self._h, name.encode(), ids.ctypes.data, st.ctypes.data, len(ids)))

    def dense_drop(self, name: str):
        _native.check(self._lib.brag_dense_drop(self._h, name.encode()))
        self._spaces.pop(name, None)

    def sparse_set_csr(self, name: str, term_ptr, doc, w, n_docs: int):
        if _is_torch_cuda(term_ptr):
            import torch
            assert term_ptr.dtype == torch.int64 and doc.dtype in (torch.int32, torch.uint32) and w.dtype == torch.float32
            term_ptr, doc, w = term_ptr.contiguous(), doc.contiguous(), w.contiguous()
            torch.cuda.current_stream().synchronize()
            _native.check(self._lib.brag_sparse_set_csr(self._h, name.encode(), term_ptr.data_ptr(), doc.data_ptr(),
                                                        w.data_ptr(), term_ptr.numel() - 1, int(n_docs), 1))
        else:
            term_ptr = np.ascontiguousarray(term_ptr, dtype=np.int64)
            doc = np.ascontiguousarray(doc, dtype=np.uint32)
            w = np.ascontiguousarray(w, dtype=np.float32)
            _native.check(self._lib.brag_sparse_set_csr(self._h, name.encode(), term_ptr.ctypes.data, doc.ctypes.data,
                                                        w.ctypes.data, len(term_ptr) - 1, int(n_docs), 0))

    def sparse_append(self, name: str, delta_term_ptr, doc, w, row_map, n_docs: int):
        """Splice the postings of new documents (a CSR over the grown vocabulary) into a space already in HBM."""
        delta_term_ptr = np.ascontiguousarray(delta_term_ptr, dtype=np.int64)
        doc = np.ascontiguousarray(doc, dtype=np.uint32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        row_map = np.ascontiguousarray(row_map, dtype=np.int32)
        _native.check(self._lib.brag_sparse_append(self._h, name.encode(), delta_term_ptr.ctypes.data, doc.ctypes.data,
                                                   w.ctypes.data, row_map.ctypes.data, len(delta_term_ptr) - 1,
                                                   int(n_docs)))

    def sparse_drop(self, name: str):
        _native.check(self._lib.brag_sparse_drop(self._h, name.encode()))

    def sparse_df(self, name: str, n_terms: int) -> np.ndarray:
        df = np.zeros(n_terms, dtype=np.int64)
        _native.check(self._lib.brag_sparse_df(self._h, name.encode(), df.ctypes.data, n_terms))
        return df

    def mask_set(self, mask_id: int, mask):
        """mask: numpy bool [n] (packed here) / uint32 words, or torch CUDA int32 words."""
        if _is_torch_cuda(mask):
            import torch
            torch.cuda.current_stream().synchronize()
            _native.check(self._lib.brag_mask_set(self._h, mask_id, mask.data_ptr(), mask.numel() * 32, 1))
            return
        mask = np.asarray(mask)
        if mask.dtype == np.bool_:
            nbits = len(mask)
            words = np.packbits(mask, bitorder="little")
            pad = (-len(words)) % 4
            if pad:
                words = np.concatenate([words, np.zeros(pad, np.uint8)])
            words = words.view(np.uint32)
        else:
            words = np.ascontiguousarray(mask, dtype=np.uint32)
            nbits = len(words) * 32
        _native.check(self._lib.brag_mask_set(self._h, mask_id, words.ctypes.data, nbits, 0))

    # ------------------------------------------------------------------ queries
    def _outs(self, B, k, dev, score_dtype="f32"):
        if dev:
            import torch
            d = torch.device("cuda", self.device)
            sdt = torch.float32 if score_dtype == "f32" else torch.float64
            return (torch.empty((B, k), dtype=sdt, device=d), torch.empty((B, k), dtype=torch.int64, device=d),
                    torch.empty((B,), dtype=torch.int32, device=d))
        sdt = np.float32 if score_dtype == "f32" else np.float64
        return np.empty((B, k), sdt), np.empty((B, k), np.int64), np.empty((B,), np.int32)

    def dense_topk(self, name: str, Q, k: int, mask_id: int = 0):
        dev = _is_torch_cuda(Q)
        Q = Q.contiguous() if dev else np.ascontiguousarray(Q, dtype=np.float32)
        B = Q.shape[0]
        s, i, c = self._outs(B, k, dev)
        _native.check(self._lib.brag_dense_topk(self._h, name.encode(), _ptr(Q), B, k, mask_id, _ptr(s), _ptr(i), _ptr(c),
                                                int(dev), _stream() if dev else None))
        return s, i, c

    def dense_rescore(self, name: str, Q, cand, k: int, mask_id: int = 0):
        dev = _is_torch_cuda(Q)
        Q = Q.contiguous() if dev else np.ascontiguousarray(Q, dtype=np.float32)
        cand = cand.contiguous() if dev else np.ascontiguousarray(cand, dtype=np.int64)
        B, m = cand.shape
        s, i, c = self._outs(B, k, dev)
        _native.check(self._lib.brag_dense_rescore(self._h, name.encode(), _ptr(Q), B, _ptr(cand), m, k, mask_id, _ptr(s),
                                                   _ptr(i), _ptr(c), int(dev), _stream() if dev else None))
        return s, i, c

    def dense_rescore_noalloc(self, name: str, Q, cand, k: int, out, mask_id: int = 0):
        """brag_dense_rescore_noalloc: device tensors only, the caller owns the outputs `out` = (float32 [B,k],
        int64 [B,k], int32 [B]); nothing is allocated here or in the library (static scratch, B * m <= 65536)."""
        if not (_is_torch_cuda(Q) and _is_torch_cuda(cand) and all(_is_torch_cuda(t) for t in out)):
            raise ValueError("dense_rescore_noalloc takes CUDA tensors only")
        if not (Q.is_contiguous() and cand.is_contiguous() and all(t.is_contiguous() for t in out)):
            raise ValueError("dense_rescore_noalloc takes contiguous tensors (it will not copy)")
        B, m = cand.shape
        s, i, c = out
        _native.check(self._lib.brag_dense_rescore_noalloc(self._h, name.encode(), _ptr(Q), B, _ptr(cand), m, k, mask_id,
                                                           _ptr(s), _ptr(i), _ptr(c), _stream()))
        return s, i, c

    def sparse_topk(self, name: str, q_ptr, q_terms, q_vals, k: int, mask_id: int = 0, max_terms: int = -1):
        """max_terms: upper bound of the longest query (terms).  Device callers should pass it: without it the engine
        copies q_ptr back and synchronises the stream in the middle of the call."""
        dev = _is_torch_cuda(q_ptr)
        if not dev:
            q_ptr = np.ascontiguousarray(q_ptr, dtype=np.int32)
            q_terms = np.ascontiguousarray(q_terms, dtype=np.int32)
            q_vals = np.ascontiguousarray(q_vals, dtype=np.float64)
        B = q_ptr.shape[0] - 1
        s, i, c = self._outs(B, k, dev)
        _native.check(self._lib.brag_sparse_topk(self._h, name.encode(), _ptr(q_ptr), _ptr(q_terms), _ptr(q_vals), B, k,
                                                 mask_id, int(max_terms), _ptr(s), _ptr(i), _ptr(c), int(dev),
                                                 _stream() if dev else None))
        return s, i, c

    def rrf(self, ids, counts, limit: int, tie: int = TIE_ID):
        """ids [L,B,kmax] int64, counts [L,B] int32 -> (scores f64 [B,limit], ids, counts)."""
        dev = _is_torch_cuda(ids)
        if not dev:
            ids = np.ascontiguousarray(ids, dtype=np.int64)
            counts = np.ascontiguousarray(counts, dtype=np.int32)
        L, B, kmax = ids.shape
        s, i, c = self._outs(B, limit, dev, "f64")
        _native.check(self._lib.brag_rrf(self._h, _ptr(ids), _ptr(counts), L, B, kmax, limit, tie, _ptr(s), _ptr(i), _ptr(c),
                                         int(dev), _stream() if dev else None))
        return s, i, c

    def rerank_meanpool(self, query_hidden, doc_hidden):
        """query_hidden [Tq, d], doc_hidden [n, T, d] float32 -> (order int32 [n], scores float32 [n])."""
        dev = _is_torch_cuda(doc_hidden)
        if dev:
            import torch
            query_hidden = query_hidden.to(torch.float32).contiguous()
            doc_hidden = doc_hidden.to(torch.float32).contiguous()
            n, T, d = doc_hidden.shape
            order = torch.empty(n, dtype=torch.int32, device=doc_hidden.device)
            scores = torch.empty(n, dtype=torch.float32, device=doc_hidden.device)
        else:
            query_hidden = np.ascontiguousarray(query_hidden, dtype=np.float32)
            doc_hidden = np.ascontiguousarray(doc_hidden, dtype=np.float32)
            n, T, d = doc_hidden.shape
            order, scores = np.empty(n, np.int32), np.empty(n, np.float32)
        assert query_hidden.ndim == 2 and query_hidden.shape[1] == d
        _native.check(self._lib.brag_rerank_meanpool(self._h, _ptr(query_hidden), query_hidden.shape[0], _ptr(doc_hidden),
                                                     n, T, d, _ptr(order), _ptr(scores), int(dev),
                                                     _stream() if dev else None))
        return order, scores

    def merge_topk(self, scores, ids, counts):
        """[G,B,k] per-shard lists -> [B,k]."""
        dev = _is_torch_cuda(ids)
        if not dev:
            scores = np.ascontiguousarray(scores, dtype=np.float32)
            ids = np.ascontiguousarray(ids, dtype=np.int64)
            counts = np.ascontiguousarray(counts, dtype=np.int32)
        G, B, k = ids.shape
        s, i, c = self._outs(B, k, dev)
        _native.check(self._lib.brag_merge_topk(self._h, _ptr(scores), _ptr(ids), _ptr(counts), G, B, k, _ptr(s), _ptr(i),
                                                _ptr(c), int(dev), _stream() if dev else None))
        return s, i, c

    def _hybrid_inputs(self, Q, q_ptr, q_terms, q_vals):
        dev = _is_torch_cuda(Q)
        if dev:
            return dev, Q.contiguous(), q_ptr, q_terms, q_vals
        return (dev, np.ascontiguousarray(Q, dtype=np.float32), np.ascontiguousarray(q_ptr, dtype=np.int32),
                np.ascontiguousarray(q_terms, dtype=np.int32), np.ascontiguousarray(q_vals, dtype=np.float64))

    @staticmethod
    def _check_out(out, B, k):
        s, i, c = out
        ok = (s.shape == (B, k) and i.shape == (B, k) and c.shape == (B,) and str(s.dtype).endswith("float64") and
              str(i.dtype).endswith("int64") and str(c.dtype).endswith("int32"))
        if not ok:
            raise ValueError("out must be (float64 [B,k], int64 [B,k], int32 [B]) buffers, C-contiguous")
        return s, i, c

    def hybrid_topk(self, dense_name: str, sparse_name: str, Q, q_ptr, q_terms, q_vals, k: int, mask_id: int = 0,
                    max_terms: int = -1, out=None):
        """out: optional (scores, ids, counts) buffers to write into and return -- a serving loop reuses (pinned) result
        buffers instead of paying for fresh pages on every call."""
        dev, Q, q_ptr, q_terms, q_vals = self._hybrid_inputs(Q, q_ptr, q_terms, q_vals)
        B = Q.shape[0]
        s, i, c = self._check_out(out, B, k) if out is not None else self._outs(B, k, dev, "f64")
        _native.check(self._lib.brag_hybrid_topk(self._h, dense_name.encode(), sparse_name.encode(), _ptr(Q), _ptr(q_ptr),
                                                 _ptr(q_terms), _ptr(q_vals), B, k, mask_id, int(max_terms), _ptr(s),
                                                 _ptr(i), _ptr(c), int(dev), _stream() if dev else None))
        return s, i, c

    # ------------------------------------------------------------------ row sharding (include/brag.h, shard.cu)
    def _stage_outs(self, B, k, dev, want):
        if not want:
            return None, None, None
        if dev:
            import torch
            d = torch.device("cuda", self.device)
            return (torch.empty((2, B, k), dtype=torch.float32, device=d), torch.empty((2, B, k), dtype=torch.int64, device=d),
                    torch.empty((2, B), dtype=torch.int32, device=d))
        return np.empty((2, B, k), np.float32), np.empty((2, B, k), np.int64), np.empty((2, B), np.int32)

    def shard_local_lists(self, dense_name: str, sparse_name: str, Q, q_ptr, q_terms, q_vals, k: int, mask_id: int = 0,
                          max_terms: int = -1):
        """This shard's dense and sparse top-k as packed keys [2, B, k] (uint64 bits carried as int64)."""
        dev, Q, q_ptr, q_terms, q_vals = self._hybrid_inputs(Q, q_ptr, q_terms, q_vals)
        B = Q.shape[0]
        if dev:
            import torch
            keys = torch.empty((2, B, k), dtype=torch.int64, device=Q.device)
        else:
            keys = np.empty((2, B, k), np.int64)
        _native.check(self._lib.brag_shard_local_lists(self._h, dense_name.encode(), sparse_name.encode(), _ptr(Q),
                                                       _ptr(q_ptr), _ptr(q_terms), _ptr(q_vals), B, k, mask_id,
                                                       int(max_terms), _ptr(keys), int(dev), _stream() if dev else None))
        return keys

    def shard_fuse(self, keys, limit: int, tie: int = TIE_ID, stages: bool = False):
        """keys [G, 2, B, k] gathered from every shard -> (rrf scores f64 [B, limit], ids, counts)
        (+ the two merged lists (scores [2,B,k], ids [2,B,k], counts [2,B]) when `stages`)."""
        dev = _is_torch_cuda(keys)
        keys = keys.contiguous() if dev else np.ascontiguousarray(keys, dtype=np.int64)
        G, two, B, k = keys.shape
        assert two == 2
        s, i, c = self._outs(B, limit, dev, "f64")
        ts, ti, tc = self._stage_outs(B, k, dev, stages)
        _native.check(self._lib.brag_shard_fuse(self._h, _ptr(keys), G, B, k, limit, tie, _ptr(s), _ptr(i), _ptr(c),
                                                _ptr(ts), _ptr(ti), _ptr(tc), int(dev), _stream() if dev else None))
        return (s, i, c, (ts, ti, tc)) if stages else (s, i, c)

    def shard_window_create(self, world: int, max_batch: int, max_k: int) -> bytes:
        """Allocate this engine's exchange window; returns its 64-byte CUDA IPC handle."""
        buf = (C.c_uint8 * 64)()
        _native.check(self._lib.brag_shard_window_create(self._h, world, max_batch, max_k, buf))
        return bytes(buf)

    def shard_connect(self, rank: int, world: int, handles: Optional[Sequence[bytes]] = None,
                      local_peers: Optional[Sequence["Engine"]] = None):
        """handles: every rank's window handle (one process per GPU); local_peers: the world's engines, in rank order,
        when they all live in this process."""
        if local_peers is not None:
            arr = (C.c_void_p * world)(*[p._h for p in local_peers])
            _native.check(self._lib.brag_shard_connect(self._h, rank, world, None, arr))
        else:
            blob = b"".join(handles)
            assert len(blob) == 64 * world
            _native.check(self._lib.brag_shard_connect(self._h, rank, world, blob, None))

    def shard_hybrid_topk(self, dense_name: str, sparse_name: str, Q, q_ptr, q_terms, q_vals, k: int, mask_id: int = 0,
                          max_terms: int = -1, stages: bool = False, out=None):
        """Every rank calls this with the same batch, in the same order; identical result on every rank."""
        dev, Q, q_ptr, q_terms, q_vals = self._hybrid_inputs(Q, q_ptr, q_terms, q_vals)
        B = Q.shape[0]
        s, i, c = self._check_out(out, B, k) if out is not None else self._outs(B, k, dev, "f64")
        ts, ti, tc = self._stage_outs(B, k, dev, stages)
        _native.check(self._lib.brag_shard_hybrid_topk(self._h, dense_name.encode(), sparse_name.encode(), _ptr(Q),
                                                       _ptr(q_ptr), _ptr(q_terms), _ptr(q_vals), B, k, mask_id,
                                                       int(max_terms), _ptr(s), _ptr(i), _ptr(c), _ptr(ts), _ptr(ti),
                                                       _ptr(tc), int(dev), _stream() if dev else None))
        return (s, i, c, (ts, ti, tc)) if stages else (s, i, c)

    def shard_close(self):
        _native.check(self._lib.brag_shard_close(self._h))

    # ------------------------------------------------------------------ persistence
    def save(self, path: str):
        """Every space of this engine (rows, postings, built scan indexes) -> one flat file."""
        _native.check(self._lib.brag_save(self._h, str(path).encode()))

    def load(self, path: str, host_state: Optional[dict] = None):
        """Spaces of a file written by save() back into HBM (nothing is rebuilt).  `host_state`: what host_state()
        returned when the file was written (the {dense name: (n, d, dtype)} table dense_append / dense_update need)."""
        _native.check(self._lib.brag_load(self._h, str(path).encode()))
        if host_state:
            self._spaces.update({k: tuple(v) for k, v in host_state["spaces"].items()})

    def host_state(self) -> dict:
        return {"spaces": dict(self._spaces)}

    # ------------------------------------------------------------------ introspection
    def profile_enable(self, on: bool = True):
        _native.check(self._lib.brag_profile_enable(self._h, int(on)))

    def profile_reset(self):
        _native.check(self._lib.brag_profile_reset(self._h))

    def profile_read(self):
        ms = (C.c_double * len(STAGES))()
        ln = (C.c_int64 * len(STAGES))()
        tot = C.c_int64()
        _native.check(self._lib.brag_profile_read(self._h, ms, ln, C.byref(tot)))
        return ({s: ms[i] for i, s in enumerate(STAGES)}, {s: ln[i] for i, s in enumerate(STAGES)}, tot.value)

    def dense_path(self, name: str, B: int) -> str:
        p = self._lib.brag_dense_path(self._h, name.encode(), B)
        return {0: "gemv", 1: "tcgen05"}.get(p, "?")

    def force_sparse_path(self, path: Optional[str]):
        _native.check(self._lib.brag_force_sparse_path(self._h, {None: -1, "scatter": 0, "bitmap": 1}[path]))

    def sparse_compact_after(self, queries: int):
        """Queries one filter serves before the index of its passing documents is built (0: first use, < 0: never)."""
        _native.check(self._lib.brag_sparse_compact_after(self._h, int(queries)))

    def sparse_offsets_min_df(self, min_df: int):
        """Terms with fewer postings get no per-sub-tile offset row (their lists are searched); 0 = automatic."""
        _native.check(self._lib.brag_sparse_offsets_min_df(self._h, int(min_df)))

    def sparse_compact_info(self, name: str):
        """(mask_id the compacted index of this space was built for or 0, documents it holds)."""
        mid, nd = C.c_int32(0), C.c_int64(0)
        _native.check(self._lib.brag_sparse_compact_info(self._h, name.encode(), C.byref(mid), C.byref(nd)))
        return int(mid.value), int(nd.value)

    def force_dense_path(self, path: Optional[str]):
        _native.check(self._lib.brag_force_dense_path(self._h, {None: -1, "gemv": 0, "tcgen05": 1, "tcgen05_full": 2}[path]))
