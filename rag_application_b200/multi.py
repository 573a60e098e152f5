"""Several GPUs behind ONE engine object (SURVEY 8 row e behind the boundary of row b): `MultiEngine` has the method
contract of `Engine` (host arrays in, host arrays out) and keeps every space row-sharded over the engines it owns, so the
Qdrant-shaped client (client.py) -- plan executor, filters, paging, upsert -- works unchanged on a collection that
lives on N GPUs: `B200QdrantClient(devices=[0, 1, ...])`.

Rows are dealt to the shards in blocks: shard(row) = (row // block) % G, a pure function of the row number.  Every
space of every collection, every filter mask and every append therefore agree on where a row lives without any
per-collection table, and appended rows keep the shards balanced.  Inside a shard local rows keep the global order, so
(score desc, id asc) per shard merges into (score desc, id asc) globally (brag_merge_topk / brag_shard_fuse on the
first engine).  Per-shard calls run concurrently, one host thread per GPU (the C ABI releases the GIL).

The one-process-per-GPU throughput path of bench.py is sharded.py (contiguous ranges, lists pushed over NVLink peer
memory); this class is the drop-in form of the same idea for the reference-facing client."""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Optional, Sequence

import numpy as np

from .sharded import lists_to_keys


class MultiEngine:
    def __init__(self, devices: Sequence[int], engine_factory: Optional[Callable] = None, block_rows: int = 1 << 16):
        if engine_factory is None:
            from .engine import Engine          # fails loudly if libbrag.so / a B200 is missing
            engine_factory = Engine
        if len(devices) < 1:
            raise ValueError("MultiEngine needs at least one device")
        self.engines = [engine_factory(d) for d in devices]
        self.G = len(self.engines)
        self.block = int(block_rows)
        self.device = devices[0]
        self._pool = ThreadPoolExecutor(max_workers=self.G)
        self._n = {}             # space name -> global rows
        self._dtype = {}

    # ------------------------------------------------------------------ row <-> (shard, local row)
    def _shard_of(self, rows):
        rows = np.asarray(rows, dtype=np.int64)
        return (rows // self.block) % self.G

    def _local_of(self, rows):
        rows = np.asarray(rows, dtype=np.int64)
        return (rows // (self.block * self.G)) * self.block + rows % self.block

    def _global_of(self, g: int, local):
        local = np.asarray(local, dtype=np.int64)
        return ((local // self.block) * self.G + g) * self.block + local % self.block

    def _local_n(self, g: int, n: int) -> int:
        full, rem = divmod(int(n), self.block * self.G)
        return full * self.block + int(min(max(rem - g * self.block, 0), self.block))

    def _rows_of(self, g: int, lo: int, hi: int) -> np.ndarray:
        """global rows in [lo, hi) that live on shard g, ascending"""
        r = np.arange(lo, hi, dtype=np.int64)
        return r[self._shard_of(r) == g]

    def _map(self, fn, shards=None):
        shards = range(self.G) if shards is None else shards
        return list(self._pool.map(fn, shards))

    def close(self):
        for e in self.engines:
            e.close()
        self._pool.shutdown(wait=True)

    # ------------------------------------------------------------------ persistence: one flat file per shard
    def save(self, path: str):
        self._map(lambda g: self.engines[g].save(f"{path}.shard{g}"))

    def load(self, path: str, host_state: dict):
        if host_state["G"] != self.G or host_state["block"] != self.block:
            raise ValueError("the index was saved with another number of shards / block size")
        self._map(lambda g: self.engines[g].load(f"{path}.shard{g}", host_state["engines"][g]))
        self._n.update(host_state["n"])
        self._dtype.update(host_state["dtype"])

    def host_state(self) -> dict:
        return {"G": self.G, "block": self.block, "n": dict(self._n), "dtype": dict(self._dtype),
                "engines": [e.host_state() for e in self.engines]}

    # ------------------------------------------------------------------ corpus
    def dense_set(self, name: str, rows, dtype: str = "f32"):
        rows = np.asarray(rows, dtype=np.float32)
        n = rows.shape[0]
        self._n[name], self._dtype[name] = n, dtype
        idx = [self._rows_of(g, 0, n) for g in range(self.G)]

        def one(g):
            if len(idx[g]):
                self.engines[g].dense_set(name, rows[idx[g]], dtype)
            else:
                self.engines[g].dense_drop(name)
        self._map(one)

    def dense_append(self, name: str, rows):
        rows = np.asarray(rows, dtype=np.float32)
        n0 = self._n[name]
        n1 = n0 + rows.shape[0]

        def one(g):
            idx = self._rows_of(g, n0, n1)
            if not len(idx):
                return
            if self._local_n(g, n0) == 0:
                self.engines[g].dense_set(name, rows[idx - n0], self._dtype[name])
            else:
                self.engines[g].dense_append(name, rows[idx - n0])
        self._map(one)
        self._n[name] = n1

    def dense_update(self, name: str, row_ids, rows):
        ids = np.asarray(row_ids, dtype=np.int64)
        rows = np.asarray(rows, dtype=np.float32)
        sh = self._shard_of(ids)

        def one(g):
            sel = sh == g
            if sel.any():
                self.engines[g].dense_update(name, self._local_of(ids[sel]), rows[sel])
        self._map(one)

    def dense_drop(self, name: str):
        self._map(lambda g: self.engines[g].dense_drop(name))
        self._n.pop(name, None)

    def _split_csr(self, term_ptr, doc, w, lo_docs: int = 0):
        """term-major CSR over global documents -> per shard (term_ptr, local docs, weights); order inside a term kept"""
        term_ptr = np.asarray(term_ptr, dtype=np.int64)
        doc = np.asarray(doc, dtype=np.int64)
        w = np.asarray(w, dtype=np.float32)
        V = len(term_ptr) - 1
        term_of = np.repeat(np.arange(V, dtype=np.int64), np.diff(term_ptr))
        sh = self._shard_of(doc)
        out = []
        for g in range(self.G):
            sel = sh == g
            p = np.zeros(V + 1, dtype=np.int64)
            np.add.at(p, term_of[sel] + 1, 1)
            out.append((np.cumsum(p), self._local_of(doc[sel]).astype(np.uint32), w[sel]))
        return out

    def sparse_set_csr(self, name: str, term_ptr, doc, w, n_docs: int):
        parts = self._split_csr(term_ptr, doc, w)
        self._n[name] = int(n_docs)

        def one(g):
            ln = self._local_n(g, n_docs)
            if ln:
                self.engines[g].sparse_set_csr(name, parts[g][0], parts[g][1], parts[g][2], ln)
            else:
                self.engines[g].sparse_drop(name)
        self._map(one)

    def sparse_append(self, name: str, delta_term_ptr, doc, w, row_map, n_docs: int):
        parts = self._split_csr(delta_term_ptr, doc, w)
        n0 = self._n[name]

        def one(g):
            ln0, ln1 = self._local_n(g, n0), self._local_n(g, n_docs)
            if ln1 == 0:
                return
            if ln0 == 0:
                self.engines[g].sparse_set_csr(name, parts[g][0], parts[g][1], parts[g][2], ln1)
            else:       # every shard takes the grown vocabulary, also the ones that receive no new posting
                self.engines[g].sparse_append(name, parts[g][0], parts[g][1], parts[g][2], row_map, ln1)
        self._map(one)
        self._n[name] = int(n_docs)

    def sparse_drop(self, name: str):
        self._map(lambda g: self.engines[g].sparse_drop(name))
        self._n.pop(name, None)

    def mask_set(self, mask_id: int, mask):
        mask = np.asarray(mask)
        if mask.dtype != np.bool_:
            mask = np.unpackbits(np.ascontiguousarray(mask, dtype=np.uint32).view(np.uint8), bitorder="little").astype(bool)
        n = len(mask)

        def one(g):
            idx = self._rows_of(g, 0, n)
            if len(idx):
                self.engines[g].mask_set(mask_id, mask[idx])
        self._map(one)

    # ------------------------------------------------------------------ queries
    def _active(self, name: str) -> List[int]:
        return [g for g in range(self.G) if self._local_n(g, self._n[name]) > 0]

    def _merge(self, lists, shards, B: int, k: int):
        """per-shard (scores, local ids, counts) -> global (scores [B,k], ids, counts) under (score desc, id asc)"""
        S = np.full((len(shards), B, k), -np.inf, np.float32)
        I = np.full((len(shards), B, k), -1, np.int64)
        C = np.zeros((len(shards), B), np.int32)
        for j, (g, (s, i, c)) in enumerate(zip(shards, lists)):
            s, i, c = np.asarray(s), np.asarray(i), np.asarray(c)
            kk = s.shape[1]
            S[j, :, :kk] = s
            I[j, :, :kk] = np.where(i >= 0, self._global_of(g, np.maximum(i, 0)), -1)
            C[j] = c
        if len(shards) == 1:
            return S[0], I[0], C[0]
        return self.engines[0].merge_topk(S, I, C)

    def dense_topk(self, name: str, Q, k: int, mask_id: int = 0):
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        shards = self._active(name)
        lists = self._map(lambda g: self.engines[g].dense_topk(name, Q, min(k, 512), mask_id), shards)
        return self._merge(lists, shards, Q.shape[0], k)

    def dense_rescore(self, name: str, Q, cand, k: int, mask_id: int = 0):
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        cand = np.asarray(cand, dtype=np.int64)
        shards = self._active(name)
        sh = np.where(cand >= 0, self._shard_of(np.maximum(cand, 0)), -1)

        def one(g):
            loc = np.where(sh == g, self._local_of(np.maximum(cand, 0)), -1)
            return self.engines[g].dense_rescore(name, Q, loc, k, mask_id)
        lists = self._map(one, shards)
        B = Q.shape[0]
        if len(shards) == 1:
            return self._merge(lists, shards, B, k)
        # up to k entries per shard: merge in slices the merge kernel takes (G * k <= 8192), exact under (score desc, id)
        S = np.concatenate([np.asarray(l[0]) for l in lists], axis=1)
        I = np.concatenate([np.where(np.asarray(l[1]) >= 0, self._global_of(g, np.maximum(np.asarray(l[1]), 0)), -1)
                            for g, l in zip(shards, lists)], axis=1)
        out_s = np.full((B, k), -np.inf, np.float32)
        out_i = np.full((B, k), -1, np.int64)
        out_c = np.zeros(B, np.int32)
        for b in range(B):
            ok = I[b] >= 0
            o = np.lexsort((I[b][ok], -S[b][ok].astype(np.float64)))[:k]
            out_s[b, :len(o)], out_i[b, :len(o)], out_c[b] = S[b][ok][o], I[b][ok][o], len(o)
        return out_s, out_i, out_c

    def sparse_topk(self, name: str, q_ptr, q_terms, q_vals, k: int, mask_id: int = 0, max_terms: int = -1):
        shards = self._active(name)
        lists = self._map(lambda g: self.engines[g].sparse_topk(name, q_ptr, q_terms, q_vals, k, mask_id, max_terms), shards)
        return self._merge(lists, shards, len(q_ptr) - 1, k)

    def rrf(self, ids, counts, limit: int, tie: int = 0):
        return self.engines[0].rrf(ids, counts, limit, tie)

    def rerank_meanpool(self, query_hidden, doc_hidden):
        return self.engines[0].rerank_meanpool(query_hidden, doc_hidden)

    def merge_topk(self, scores, ids, counts):
        return self.engines[0].merge_topk(scores, ids, counts)

    def hybrid_topk(self, dense_name: str, sparse_name: str, Q, q_ptr, q_terms, q_vals, k: int, mask_id: int = 0,
                    max_terms: int = -1):
        """Per-shard dense and sparse lists as packed keys with global rows, fused by ONE kernel on the first engine
        (brag_shard_fuse: merge of the dense lists, merge of the sparse lists, RRF)."""
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        B = Q.shape[0]
        shards = sorted(set(self._active(dense_name)) | set(self._active(sparse_name)))

        def one(g):
            keys = np.zeros((2, B, k), np.int64)
            if self._local_n(g, self._n[dense_name]) > 0:
                s, i, _ = self.engines[g].dense_topk(dense_name, Q, k, mask_id)
                i = np.asarray(i)
                keys[0] = lists_to_keys(np.asarray(s), np.where(i >= 0, self._global_of(g, np.maximum(i, 0)), -1))
            if self._local_n(g, self._n[sparse_name]) > 0:
                s, i, _ = self.engines[g].sparse_topk(sparse_name, q_ptr, q_terms, q_vals, k, mask_id, max_terms)
                i = np.asarray(i)
                keys[1] = lists_to_keys(np.asarray(s), np.where(i >= 0, self._global_of(g, np.maximum(i, 0)), -1))
            return keys
        keys = np.stack(self._map(one, shards))
        return self.engines[0].shard_fuse(keys, k)
