"""CPU oracle (numpy) for dense top-k -> BM25 sparse top-k -> RRF.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED (see oracle/__init__.py): restates the published algorithm of
qdrant-client local mode (un-vendored, un-pinned dependency, reference requirements.txt:3)
for exactly the requests the reference issues:

* schema / derived vectors ......... app/core/vector_store/qdrant/qdrant_handler.py:57-86, :144-150
* nested prefetch tree + RRF ....... app/core/vector_store/qdrant/qdrant_handler.py:296-372
* sparse vector format ............. app/core/embedding/embedding_handler.py:123-138
* default limits ................... app/api/v1/endpoints/mcp/qdrant_search_mcp_endpoint.py:19-28

Semantics restated:
 1 COSINE = L2-normalise rows at insert and the query at search, score = dot; zero norm
   divided by float32 eps.  DOT = raw dot.
 2 ranking order is canonical (score desc, id asc) -- upstream's argsort tie order is
   implementation defined; north_star fixes "ties broken by id".
 3 sparse score = sum over shared indices, ascending index order, accumulated in Python
   float (float64), returned as float32; documents sharing no index are NOT candidates.
   Modifier.IDF multiplies query values by ln(1 + (N - df + 0.5)/(df + 0.5)).
 4 RRF score(id) = sum_lists 1/(2 + pos), pos 0-based, float64; default limit 10.
 5 nested prefetch: evaluate children, then fuse (FusionQuery) or re-score the union of
   child ids with this level's query/using/filter/limit.
 6 exact / hnsw_ef are no-ops (always brute force).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

EPS32 = float(np.finfo(np.float32).eps)
RRF_CONST = 2
DEFAULT_FUSION_LIMIT = 10


# ----------------------------------------------------------------------------- storage
def round_storage(x: np.ndarray, dtype: str) -> np.ndarray:
    """Round float32 values to the storage dtype and return them as float32.

    bf16 is round-to-nearest-even on the upper 16 bits (what the device converter does)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    if dtype in ("f32", "i8"):              # int8 spaces hold the integer values the caller sent
        return x
    if dtype == "f16":
        return x.astype(np.float16).astype(np.float32)
    if dtype == "bf16":
        u = x.view(np.uint32).astype(np.uint64)
        r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint32) << 16
        return r.astype(np.uint32).view(np.float32).reshape(x.shape)
    raise ValueError(dtype)


def to_bf16_bits(x: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    return (((u + 0x7FFF + ((u >> 16) & 1)) >> 16) & 0xFFFF).astype(np.uint16).reshape(x.shape)


def normalize_rows(x: np.ndarray) -> np.ndarray:
    """Cosine preprocessing in float32 (item 1)."""
    x = np.array(x, dtype=np.float32, copy=True)
    n = np.linalg.norm(x, axis=-1, keepdims=True).astype(np.float32)
    return (x / np.where(n != 0.0, n, np.float32(EPS32))).astype(np.float32)


def quantize_int8(x: Sequence[float]) -> np.ndarray:
    """qdrant_handler.py:144-146 / :300-302 -- (x*127).astype(int8) truncates toward zero;
    the trailing clip is a no-op."""
    return np.clip((np.array(x) * 127).astype(np.int8), -128, 127)


# ----------------------------------------------------------------------------- dense
def dense_scores(x_stored: np.ndarray, q_stored: np.ndarray) -> np.ndarray:
    """Exact dot of stored values: float64 accumulate, float32 result.  [N,d]x[d] -> [N]."""
    return (x_stored.astype(np.float64) @ q_stored.astype(np.float64)).astype(np.float32)


def int8_cosine_scores(x_int: np.ndarray, q_int: np.ndarray) -> np.ndarray:
    """Row a9 with int8 STORAGE (include/brag.h BRAG_I8): cosine of the integer vectors the reference builds
    (qdrant_handler.py:144-146 rows, :300-302 query).  Exact integer dot and squared norms, then one float64
    expression dot / (sqrt(ssx) * sqrt(ssq)) rounded to float32; a zero row or query scores 0.  (Qdrant holding the
    same values in a float32 COSINE space normalises rows and query in float32 first; the two differ by float32
    rounding of the norms, ~1e-7 relative -- tests/test_oracle_kat.py pins that bound.)"""
    xi = np.asarray(x_int).astype(np.int64)
    qi = np.asarray(q_int).astype(np.int64)
    dot = xi @ qi
    ssx = (xi * xi).sum(axis=-1)
    ssq = int((qi * qi).sum())
    den = np.sqrt(ssx.astype(np.float64)) * np.sqrt(np.float64(ssq))
    out = np.zeros(dot.shape, np.float64)
    ok = den != 0.0
    out[ok] = dot[ok].astype(np.float64) / den[ok]
    return out.astype(np.float32)


def topk_order(scores: np.ndarray, ids: np.ndarray, k: int) -> np.ndarray:
    """Indices of the k best entries under (score desc, id asc)."""
    order = np.lexsort((ids, -scores.astype(np.float64)))
    return order[:k]


def dense_topk(x_stored: np.ndarray, q_stored: np.ndarray, k: int,
               mask: Optional[np.ndarray] = None, cand: Optional[np.ndarray] = None, int8: bool = False
               ) -> Tuple[np.ndarray, np.ndarray]:
    """Rows a1/a2/a5: full or candidate-restricted scan, bitmask predicate, top-k.
    mask: bool [N] (True = keep).  cand: int row ids (union semantics: duplicates dropped).
    int8: the space is stored as int8 (scores = int8_cosine_scores)."""
    n = x_stored.shape[0]
    ids = np.arange(n, dtype=np.int64) if cand is None else np.unique(np.asarray(cand, dtype=np.int64))
    if mask is not None:
        ids = ids[mask[ids]]
    score = int8_cosine_scores if int8 else dense_scores
    s = score(x_stored[ids], q_stored) if len(ids) else np.zeros(0, np.float32)
    o = topk_order(s, ids, k)
    return s[o], ids[o]


# ----------------------------------------------------------------------------- sparse
def bm25_tf_weight(tf: np.ndarray, doc_len: np.ndarray, k1: float = 1.2, b: float = 0.75,
                   avgdl: float = 256.0) -> np.ndarray:
    """fastembed Qdrant/bm25 document-side weight: float64 formula, float32 result."""
    tf = np.asarray(tf, dtype=np.float64)
    dl = np.asarray(doc_len, dtype=np.float64)
    return (tf * (k1 + 1.0) / (tf + k1 * (1.0 - b + b * dl / avgdl))).astype(np.float32)


def idf(n_docs: int, df: int) -> float:
    return math.log(1.0 + (n_docs - df + 0.5) / (df + 0.5))


def sparse_dot(q_idx: Sequence[int], q_val: Sequence[float], d_idx: Sequence[int],
               d_val: Sequence[float]) -> Optional[np.float32]:
    """Two-pointer dot over sorted indices, Python-float accumulation (item 3)."""
    i = j = 0
    res = 0.0
    overlap = False
    while i < len(q_idx) and j < len(d_idx):
        a, b_ = q_idx[i], d_idx[j]
        if a == b_:
            overlap = True
            res += float(q_val[i]) * float(d_val[j])
            i += 1
            j += 1
        elif a < b_:
            i += 1
        else:
            j += 1
    return np.float32(res) if overlap else None


@dataclass
class CSR:
    """Term-major inverted index: postings of term t are doc[term_ptr[t]:term_ptr[t+1]] (doc asc)."""
    term_ptr: np.ndarray  # int64 [V+1]
    doc: np.ndarray       # uint32 [nnz]
    w: np.ndarray         # float32 [nnz]
    n_docs: int

    @property
    def n_terms(self) -> int:
        return len(self.term_ptr) - 1


def build_csr(docs: Sequence[Tuple[Sequence[int], Sequence[float]]], n_terms: int) -> CSR:
    """docs[i] = (term rows, weights).  Rows inside a doc need not be sorted."""
    rows, dd, ww = [], [], []
    for i, (ti, tv) in enumerate(docs):
        rows.append(np.asarray(ti, dtype=np.int64))
        dd.append(np.full(len(ti), i, dtype=np.int64))
        ww.append(np.asarray(tv, dtype=np.float32))
    if rows:
        rows, dd, ww = np.concatenate(rows), np.concatenate(dd), np.concatenate(ww)
    else:
        rows, dd, ww = np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0, np.float32)
    o = np.lexsort((dd, rows))
    rows, dd, ww = rows[o], dd[o], ww[o]
    ptr = np.zeros(n_terms + 1, dtype=np.int64)
    np.add.at(ptr, rows + 1, 1)
    return CSR(np.cumsum(ptr), dd.astype(np.uint32), ww, len(docs))


def sparse_scores_csr(csr: CSR, q_terms: Sequence[int], q_vals: Sequence[float]
                      ) -> Tuple[np.ndarray, np.ndarray]:
    """Inverted-index evaluation of item 3.  q_terms ascending.  Returns (acc float64 [N], touched bool [N]).
    Accumulation order per doc = ascending term row = upstream's ascending-index order."""
    acc = np.zeros(csr.n_docs, dtype=np.float64)
    touched = np.zeros(csr.n_docs, dtype=bool)
    last = -1
    for t, v in zip(q_terms, q_vals):
        assert t > last, "query terms must be strictly ascending"
        last = t
        if t < 0 or t >= csr.n_terms:
            continue
        lo, hi = int(csr.term_ptr[t]), int(csr.term_ptr[t + 1])
        d = csr.doc[lo:hi]
        acc[d] += float(v) * csr.w[lo:hi].astype(np.float64)
        touched[d] = True
    return acc, touched


def sparse_topk(csr: CSR, q_terms: Sequence[int], q_vals: Sequence[float], k: int,
                mask: Optional[np.ndarray] = None, cand: Optional[np.ndarray] = None
                ) -> Tuple[np.ndarray, np.ndarray]:
    """Row a4: overlap-only candidates, float32 score, (score desc, id asc) top-k."""
    acc, touched = sparse_scores_csr(csr, q_terms, q_vals)
    if mask is not None:
        touched &= mask
    if cand is not None:
        sel = np.zeros_like(touched)
        sel[np.asarray(cand, dtype=np.int64)] = True
        touched &= sel
    ids = np.nonzero(touched)[0].astype(np.int64)
    s = acc[ids].astype(np.float32)
    o = topk_order(s, ids, k)
    return s[o], ids[o]


def apply_idf(csr: CSR, q_terms: Sequence[int], q_vals: Sequence[float]) -> List[float]:
    """Modifier.IDF: query values (as Python floats) times idf(N, df_t)."""
    out = []
    for t, v in zip(q_terms, q_vals):
        df = int(csr.term_ptr[t + 1] - csr.term_ptr[t]) if 0 <= t < csr.n_terms else 0
        out.append(float(v) * idf(csr.n_docs, df))
    return out


# ----------------------------------------------------------------------------- rerank (row f4)
def rerank_meanpool(query_hidden: np.ndarray, doc_hidden: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """The tensor code of rerank_documents (app/core/models/huggingface/huggingface.py:183-189): hidden states
    mean-pooled over every position (padding included), dot products, `np.argsort(scores)[::-1]`.  The reference's
    argsort is numpy's default (unstable); ties are pinned here as the reversed STABLE ascending order, i.e. equal
    scores put the higher index first.  Returns (order, scores)."""
    q = np.asarray(query_hidden, dtype=np.float32).mean(axis=0, dtype=np.float32)
    dm = np.asarray(doc_hidden, dtype=np.float32).mean(axis=1, dtype=np.float32)
    scores = (dm @ q).astype(np.float32)
    return np.argsort(scores, kind="stable")[::-1].astype(np.int32), scores


# ----------------------------------------------------------------------------- fusion
def rrf(lists: Sequence[Sequence[int]], limit: int = DEFAULT_FUSION_LIMIT, tie: str = "id"
        ) -> Tuple[np.ndarray, np.ndarray]:
    """Row a6 (item 4).  tie='id' -> canonical (score desc, id asc); tie='first_seen' ->
    upstream's stable sort (list 0 order, then new ids of list 1, ...)."""
    scores: Dict[int, float] = {}
    first: Dict[int, int] = {}
    for lst in lists:
        for pos, pid in enumerate(lst):
            pid = int(pid)
            if pid in scores:
                scores[pid] += 1.0 / (RRF_CONST + pos)
            else:
                first[pid] = len(first)
                scores[pid] = 1.0 / (RRF_CONST + pos)
    if tie == "id":
        items = sorted(scores.items(), key=lambda kv: (-kv[1], kv[0]))
    else:
        items = sorted(scores.items(), key=lambda kv: (-kv[1], first[kv[0]]))
    items = items[:limit]
    return (np.array([s for _, s in items], dtype=np.float64),
            np.array([i for i, _ in items], dtype=np.int64))


# ----------------------------------------------------------------------------- collection + plan
@dataclass
class OracleCollection:
    """In-memory collection with named dense spaces and one or more sparse spaces."""
    dense: Dict[str, np.ndarray] = field(default_factory=dict)       # name -> stored float32 [N,d]
    distance: Dict[str, str] = field(default_factory=dict)          # name -> "cosine" | "dot"
    storage: Dict[str, str] = field(default_factory=dict)           # name -> "f32" | "f16" | "bf16"
    sparse: Dict[str, CSR] = field(default_factory=dict)
    sparse_modifier: Dict[str, str] = field(default_factory=dict)   # name -> "none" | "idf"
    n: int = 0

    def add_dense(self, name: str, x: np.ndarray, distance: str = "cosine", storage: str = "f32"):
        x = np.asarray(x, dtype=np.float32)
        if storage == "i8":
            assert distance == "cosine" and np.array_equal(np.rint(x), x) and np.abs(x).max(initial=0) <= 128
            self.dense[name] = x                          # the integer values; scored by int8_cosine_scores
        else:
            if distance == "cosine":
                x = normalize_rows(x)
            self.dense[name] = round_storage(x, storage)
        self.distance[name] = distance
        self.storage[name] = storage
        self.n = x.shape[0]

    def add_sparse(self, name: str, csr: CSR, modifier: str = "none"):
        self.sparse[name] = csr
        self.sparse_modifier[name] = modifier
        self.n = max(self.n, csr.n_docs)

    def prep_query(self, name: str, q: Sequence[float]) -> np.ndarray:
        """The query stays float32 whatever the storage dtype of the space: qdrant holds the query vector the
        reference sends (qdrant_handler.py:363-369, a Python float list) in float32 and only the STORED vectors
        carry a datatype.  (Round 1 rounded it to the storage dtype on both sides; the engine now only feeds the
        rounded copy to its 16-bit tensor-core scan and re-scores the candidates with the float32 query.)"""
        q = np.asarray(q, dtype=np.float32)
        if self.distance[name] == "cosine" and self.storage[name] != "i8":
            q = normalize_rows(q[None, :])[0]
        return q


@dataclass
class Node:
    """One Prefetch / top-level request.  query: dense list, ("sparse", terms, vals) or "rrf"."""
    query: Any
    using: Optional[str] = None
    limit: Optional[int] = None
    prefetch: List["Node"] = field(default_factory=list)
    mask: Optional[np.ndarray] = None


def run_node(col: OracleCollection, node: Node, rrf_tie: str = "id", propagate: bool = False,
             _inherited: Optional[np.ndarray] = None) -> Tuple[np.ndarray, np.ndarray]:
    """Item 5: bottom-up evaluation of the prefetch tree.  Returns (scores, ids).
    propagate=False is qdrant-client LOCAL mode: a filter acts on the node that carries it.  propagate=True is the
    Qdrant SERVER's planner, which merges a request's filter into every prefetch below it (the reference talks to a
    server and passes query_filter at the root, qdrant_handler.py:363-372) -- restated from upstream knowledge like the
    rest of this file, pinned by tests/golden/kat.json: tree.final_masked_server."""
    own = node.mask
    if _inherited is not None:
        own = _inherited if own is None else (own & _inherited)
    child = [run_node(col, c, rrf_tie, propagate, own if propagate else None) for c in node.prefetch]
    if _inherited is not None:
        node = Node(node.query, node.using, node.limit, node.prefetch, own)
    if isinstance(node.query, str) and node.query == "rrf":
        lim = node.limit if node.limit is not None else DEFAULT_FUSION_LIMIT
        s, i = rrf([c[1] for c in child], limit=lim, tie=rrf_tie)
        if node.mask is not None and len(i):
            keep = node.mask[i]
            s, i = s[keep], i[keep]
        return s, i
    cand = None
    if child:
        cand = np.unique(np.concatenate([c[1] for c in child])) if any(len(c[1]) for c in child) \
            else np.zeros(0, np.int64)
    lim = node.limit if node.limit is not None else DEFAULT_FUSION_LIMIT
    if isinstance(node.query, tuple) and node.query[0] == "sparse":
        _, terms, vals = node.query
        csr = col.sparse[node.using]
        if col.sparse_modifier.get(node.using, "none") == "idf":
            vals = apply_idf(csr, terms, vals)
        return sparse_topk(csr, terms, vals, lim, mask=node.mask, cand=cand)
    q = col.prep_query(node.using, node.query)
    return dense_topk(col.dense[node.using], q, lim, mask=node.mask, cand=cand, int8=col.storage[node.using] == "i8")


def reference_tree(dense_vector: Sequence[float], sparse_terms: Sequence[int], sparse_vals: Sequence[float],
                   sp: Dict[str, int], mask: Optional[np.ndarray] = None) -> Node:
    """The exact request QdrantHandler.hybrid_search builds (qdrant_handler.py:305-372)."""
    dv = list(dense_vector)
    qq = quantize_int8(dv).astype(np.float32).tolist()
    mat = Node(dv, "dense", sp["dense_limit"], [
        Node(dv[:256], "matryoshka_256", sp["matryoshka_256_limit"], [
            Node(dv[:128], "matryoshka_128", sp["matryoshka_128_limit"], [
                Node(dv[:64], "matryoshka_64", sp["matryoshka_64_limit"])])])])
    qd = Node(dv, "dense", sp["dense_limit"], [Node(qq, "quantized", sp["quantized_limit"])])
    s1 = Node(("sparse", list(sparse_terms), list(sparse_vals)), "sparse", sp["sparse_limit"])
    f1 = Node("rrf", None, None, [qd, s1])
    return Node(dv, "dense", sp["final_limit"], [mat, f1], mask=mask)


def hybrid_simple(col: OracleCollection, dense_vector, sparse_terms, sparse_vals, k: int,
                  dense_name="dense", sparse_name="sparse", mask=None, rrf_tie="id"):
    """north_star three-stage shape: dense top-k, sparse top-k, RRF(limit=k)."""
    d = Node(list(dense_vector), dense_name, k, mask=mask)
    s = Node(("sparse", list(sparse_terms), list(sparse_vals)), sparse_name, k, mask=mask)
    return run_node(col, Node("rrf", None, k, [d, s]), rrf_tie)
