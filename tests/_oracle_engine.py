"""TEST-ONLY stand-in for rag_application_b200.Engine backed by the CPU oracle, so the host-side
logic above the C ABI (sharded exchange, Qdrant-shaped client, plan executor) can be exercised on
a machine without a GPU.  Same method contract as Engine: numpy in -> numpy out, torch CPU
tensors in -> torch CPU tensors out; lists padded with id -1 / score -inf, true length in counts.
Never imported by the product."""
from __future__ import annotations

import numpy as np

from oracle import oracle as O


def _np(a):
    return a.numpy() if hasattr(a, "numpy") and not isinstance(a, np.ndarray) else np.asarray(a)


class OracleEngine:
    def __init__(self, device: int = 0):
        self.dense, self.dtype, self.sparse, self.masks = {}, {}, {}, {}
        self.id_offset = 0
        self.calls = []

    def close(self):
        pass

    def save(self, path):
        """CPU mirror of brag_save / brag_load (a pickle instead of the flat file)."""
        import pickle
        with open(path, "wb") as f:
            pickle.dump({"dense": self.dense, "dtype": self.dtype, "sparse": self.sparse, "id_offset": self.id_offset}, f)

    def load(self, path, host_state=None):
        import pickle
        with open(path, "rb") as f:
            st = pickle.load(f)
        self.dense.update(st["dense"]); self.dtype.update(st["dtype"]); self.sparse.update(st["sparse"])
        self.id_offset = st["id_offset"]

    def host_state(self):
        return {"spaces": {}}

    def set_id_offset(self, off):
        self.id_offset = int(off)

    def dense_set(self, name, rows, dtype="f32"):
        self.dense[name] = O.round_storage(_np(rows).astype(np.float32), dtype)
        self.dtype[name] = dtype

    def dense_append(self, name, rows):
        new = O.round_storage(_np(rows).astype(np.float32), self.dtype[name])
        self.dense[name] = np.concatenate([self.dense[name], new])

    def dense_update(self, name, row_ids, rows):
        self.dense[name][np.asarray(row_ids, dtype=np.int64)] = O.round_storage(_np(rows).astype(np.float32), self.dtype[name])

    def dense_drop(self, name):
        self.dense.pop(name, None)

    def sparse_set_csr(self, name, term_ptr, doc, w, n_docs):
        self.sparse[name] = O.CSR(_np(term_ptr).astype(np.int64), _np(doc).astype(np.uint32),
                                  _np(w).astype(np.float32), int(n_docs))

    def sparse_append(self, name, delta_term_ptr, doc, w, row_map, n_docs):
        """CPU mirror of brag_sparse_append: old lists keep their order, the delta follows them inside each term."""
        old = self.sparse[name]
        dptr = _np(delta_term_ptr).astype(np.int64)
        ddoc, dw, rm = _np(doc).astype(np.uint32), _np(w).astype(np.float32), _np(row_map).astype(np.int64)
        v_new = len(dptr) - 1
        assert len(rm) == len(old.term_ptr) - 1 and (np.diff(rm) > 0).all() and n_docs >= old.n_docs
        assert len(ddoc) == 0 or ddoc.min() >= old.n_docs
        inv = np.full(v_new, -1, np.int64)
        inv[rm] = np.arange(len(rm))
        docs, ws, ptr = [], [], [0]
        for t in range(v_new):
            n = int(dptr[t + 1] - dptr[t])
            if inv[t] >= 0:
                a, b = old.term_ptr[inv[t]], old.term_ptr[inv[t] + 1]
                docs.append(old.doc[a:b]); ws.append(old.w[a:b])
                n += int(b - a)
            docs.append(ddoc[dptr[t]:dptr[t + 1]]); ws.append(dw[dptr[t]:dptr[t + 1]])
            ptr.append(ptr[-1] + n)
        self.sparse[name] = O.CSR(np.asarray(ptr, np.int64), np.concatenate(docs).astype(np.uint32) if docs else np.zeros(0, np.uint32),
                                  np.concatenate(ws).astype(np.float32) if ws else np.zeros(0, np.float32), int(n_docs))

    def sparse_drop(self, name):
        self.sparse.pop(name, None)

    def sparse_df(self, name, n_terms):
        return np.diff(self.sparse[name].term_ptr)

    def mask_set(self, mask_id, mask):
        m = _np(mask)
        if m.dtype != np.bool_:
            m = np.unpackbits(m.astype(np.uint32).view(np.uint8), bitorder="little").astype(bool)
        self.masks[mask_id] = m

    def _mask(self, mask_id, n):
        if not mask_id:
            return None
        m = self.masks[mask_id]
        return m[:n]

    def _wrap(self, like, s, i, c):
        if hasattr(like, "numpy") and not isinstance(like, np.ndarray):
            import torch
            return torch.from_numpy(s), torch.from_numpy(i), torch.from_numpy(c)
        return s, i, c

    def _pad(self, lists, k, sdtype=np.float32):
        B = len(lists)
        s = np.full((B, k), -np.inf, sdtype)
        i = np.full((B, k), -1, np.int64)
        c = np.zeros((B,), np.int32)
        for b, (ls, li) in enumerate(lists):
            n = len(li)
            s[b, :n], i[b, :n], c[b] = ls, li, n
        return s, i, c

    def dense_topk(self, name, Q, k, mask_id=0):
        self.calls.append(("dense_topk", name, k))
        x = self.dense[name]
        q = _np(Q).astype(np.float32)      # the query stays float32 (oracle.prep_query)
        out = []
        for b in range(q.shape[0]):
            s, i = O.dense_topk(x, q[b], k, mask=self._mask(mask_id, x.shape[0]), int8=self.dtype[name] == "i8")
            out.append((s, i + self.id_offset))
        return self._wrap(Q, *self._pad(out, k))

    def dense_rescore(self, name, Q, cand, k, mask_id=0):
        self.calls.append(("dense_rescore", name, k))
        x = self.dense[name]
        q = _np(Q).astype(np.float32)      # the query stays float32 (oracle.prep_query)
        cand = _np(cand)
        out = []
        for b in range(q.shape[0]):
            cb = cand[b][cand[b] >= 0] - self.id_offset
            cb = cb[(cb >= 0) & (cb < x.shape[0])]
            s, i = O.dense_topk(x, q[b], k, mask=self._mask(mask_id, x.shape[0]), cand=cb, int8=self.dtype[name] == "i8")
            out.append((s, i + self.id_offset))
        return self._wrap(Q, *self._pad(out, k))

    def sparse_topk(self, name, q_ptr, q_terms, q_vals, k, mask_id=0, max_terms=-1):
        self.calls.append(("sparse_topk", name, k))
        csr = self.sparse[name]
        qp, qt, qv = _np(q_ptr), _np(q_terms), _np(q_vals)
        out = []
        for b in range(len(qp) - 1):
            s, i = O.sparse_topk(csr, qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]], k,
                                 mask=self._mask(mask_id, csr.n_docs))
            out.append((s, i + self.id_offset))
        return self._wrap(q_ptr, *self._pad(out, k))

    def rrf(self, ids, counts, limit, tie=0):
        self.calls.append(("rrf", limit))
        I, Cn = _np(ids), _np(counts)
        L, B, _ = I.shape
        out = []
        for b in range(B):
            out.append(O.rrf([I[l, b, :Cn[l, b]] for l in range(L)], limit=limit, tie="id" if tie == 0 else "first_seen"))
        return self._wrap(ids, *self._pad(out, limit, np.float64))

    def rerank_meanpool(self, query_hidden, doc_hidden):
        return O.rerank_meanpool(_np(query_hidden), _np(doc_hidden))

    def merge_topk(self, scores, ids, counts):
        self.calls.append(("merge_topk",))
        S, I, Cn = _np(scores), _np(ids), _np(counts)
        G, B, k = I.shape
        out = []
        for b in range(B):
            fs = np.concatenate([S[g, b, :Cn[g, b]] for g in range(G)])
            fi = np.concatenate([I[g, b, :Cn[g, b]] for g in range(G)])
            o = O.topk_order(fs, fi, k)
            out.append((fs[o], fi[o]))
        return self._wrap(ids, *self._pad(out, k))

    def shard_local_lists(self, dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id=0, max_terms=-1):
        """CPU mirror of brag_shard_local_lists: [2, B, k] packed keys with global ids."""
        from rag_application_b200.sharded import lists_to_keys
        d = self.dense_topk(dense_name, Q, k, mask_id)
        s = self.sparse_topk(sparse_name, q_ptr, q_terms, q_vals, k, mask_id)
        keys = np.stack([lists_to_keys(_np(d[0]), _np(d[1])), lists_to_keys(_np(s[0]), _np(s[1]))])
        if hasattr(Q, "numpy") and not isinstance(Q, np.ndarray):
            import torch
            return torch.from_numpy(keys)
        return keys

    def shard_fuse(self, keys, limit, tie=0, stages=False):
        """CPU mirror of brag_shard_fuse: merge the G dense lists, the G sparse lists, RRF of the two."""
        from rag_application_b200.sharded import keys_to_lists
        K = _np(keys)
        G, _, B, k = K.shape
        S, I = keys_to_lists(K)
        merged = []
        for l in range(2):
            C = (I[:, l] >= 0).sum(axis=2).astype(np.int32)
            merged.append([_np(t) for t in self.merge_topk(S[:, l], I[:, l], C)])
        ids = np.stack([merged[0][1], merged[1][1]])
        cnt = np.stack([merged[0][2], merged[1][2]])
        r = self._wrap(keys, *[_np(t) for t in self.rrf(ids, cnt, limit, tie)])
        if stages:
            return r + ((np.stack([merged[0][0], merged[1][0]]), ids, cnt),)
        return r

    def hybrid_topk(self, dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id=0, max_terms=-1):
        d = self.dense_topk(dense_name, Q, k, mask_id)
        s = self.sparse_topk(sparse_name, q_ptr, q_terms, q_vals, k, mask_id)
        ids = np.stack([_np(d[1]), _np(s[1])])
        cnt = np.stack([_np(d[2]), _np(s[2])])
        r = self.rrf(ids, cnt, k)
        return self._wrap(Q, *[_np(t) for t in r])
