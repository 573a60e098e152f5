"""Synthetic corpora / queries for tests and benchmarks.

numpy generators (seeded PCG64) are used by the parity tests; the torch generators build the
full-size benchmark corpora directly in HBM (same distributions, device RNG)."""
from __future__ import annotations

from typing import Tuple

import numpy as np

K1, B_PARAM, AVGDL = 1.2, 0.75, 256.0
SEED_CORPUS, SEED_QUERY, SEED_MASK = 1234, 5678, 4242


def bm25_weight_table(doc_len: int, max_tf: int = 256) -> np.ndarray:
    """w[tf] for a fixed document length: float64 formula, float32 result (fastembed Qdrant/bm25)."""
    tf = np.arange(max_tf + 1, dtype=np.float64)
    w = tf * (K1 + 1.0) / (tf + K1 * (1.0 - B_PARAM + B_PARAM * doc_len / AVGDL))
    return w.astype(np.float32)


def zipf_cdf(V: int) -> np.ndarray:
    p = 1.0 / np.arange(1, V + 1, dtype=np.float64)
    p /= p.sum()
    return np.cumsum(p)


# ------------------------------------------------------------------ numpy (tests)
def dense_corpus(n: int, d: int, seed: int = SEED_CORPUS, clustered: bool = False) -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(seed))
    if clustered:
        c = rng.standard_normal((min(4096, max(1, n // 16)), d)).astype(np.float32)
        x = c[rng.integers(0, len(c), n)] + 0.3 * rng.standard_normal((n, d)).astype(np.float32)
    else:
        x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x.astype(np.float32)


def dense_queries(x: np.ndarray, B: int, seed: int = SEED_QUERY) -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(seed))
    r = rng.integers(0, x.shape[0], B)
    q = x[r] + 0.1 * rng.standard_normal((B, x.shape[1])).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q.astype(np.float32)


def zipf_corpus(n_docs: int, V: int = 50000, L: int = 285, seed: int = SEED_CORPUS
                ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """Term-major CSR of a Zipf(1.0) corpus: returns (term_ptr int64 [V+1], doc uint32, w float32, tf)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    cdf = zipf_cdf(V)
    terms = np.minimum(np.searchsorted(cdf, rng.random((n_docs, L))), V - 1).astype(np.int64)
    key = (np.arange(n_docs, dtype=np.int64)[:, None] * V + terms).ravel()
    uk, tf = np.unique(key, return_counts=True)
    doc, term = uk // V, uk % V
    o = np.lexsort((doc, term))
    doc, term, tf = doc[o], term[o], tf[o]
    ptr = np.zeros(V + 1, dtype=np.int64)
    np.add.at(ptr, term + 1, 1)
    ptr = np.cumsum(ptr)
    w = bm25_weight_table(L, int(tf.max()))[tf]
    return ptr, doc.astype(np.uint32), w.astype(np.float32), tf.astype(np.int32)


def sparse_queries(term_ptr: np.ndarray, doc: np.ndarray, n_docs: int, B: int, n_terms: int = 8,
                   regime: str = "zipf", seed: int = SEED_QUERY, skip_top: int = 64
                   ) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """(q_ptr int32 [B+1], q_terms int32 ascending per query, q_vals float64).  Values follow the
    reference: query encoded with the document-side formula, tf=1, |q|=n_terms
    (embedding_handler.py:123)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    V = len(term_ptr) - 1
    df = np.diff(term_ptr).astype(np.float64)
    qv = float(bm25_weight_table(n_terms, 1)[1])
    q_terms = []
    if regime == "zipf":
        p = df.copy()
        p[np.argsort(-df)[:skip_top]] = 0.0
        p /= p.sum()
        for _ in range(B):
            q_terms.append(np.sort(rng.choice(V, size=n_terms, replace=False, p=p)))
    else:  # "doc": terms of a random document
        order = np.argsort(doc, kind="stable")
        dptr = np.searchsorted(doc[order], np.arange(n_docs + 1))
        term_of = np.repeat(np.arange(V), np.diff(term_ptr))
        for _ in range(B):
            dd = rng.integers(0, n_docs)
            cand = term_of[order[dptr[dd]:dptr[dd + 1]]]
            q_terms.append(np.sort(rng.choice(cand, size=min(n_terms, len(cand)), replace=False)))
    q_ptr = np.zeros(B + 1, dtype=np.int32)
    q_ptr[1:] = np.cumsum([len(t) for t in q_terms])
    flat = np.concatenate(q_terms).astype(np.int32)
    return q_ptr, flat, np.full(len(flat), qv, dtype=np.float64)


def random_mask(n: int, sel: float = 0.10, seed: int = SEED_MASK) -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.random(n) < sel


# ------------------------------------------------------------------ torch (benchmarks, on device)
def dense_corpus_torch(n: int, d: int, dtype, device, seed: int = SEED_CORPUS, block: int = 1 << 20):
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty((n, d), dtype=dtype, device=device)
    for s in range(0, n, block):
        e = min(n, s + block)
        x = torch.randn((e - s, d), generator=g, device=device, dtype=torch.float32)
        x /= x.norm(dim=1, keepdim=True)
        out[s:e] = x.to(dtype)
    return out


def dense_queries_torch(x, B: int, seed: int = SEED_QUERY):
    import torch
    g = torch.Generator(device=x.device)
    g.manual_seed(seed)
    r = torch.randint(0, x.shape[0], (B,), generator=g, device=x.device)
    q = x[r].float() + 0.1 * torch.randn((B, x.shape[1]), generator=g, device=x.device)
    q /= q.norm(dim=1, keepdim=True)
    return q.contiguous()


def zipf_corpus_torch(n_docs: int, V: int, L: int, device, seed: int = SEED_CORPUS, block_docs: int = 1 << 18):
    """Device-side build of the term-major CSR: (term_ptr int64 [V+1], doc int32 [nnz], w float32 [nnz])."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    cdf = torch.from_numpy(zipf_cdf(V)).to(device=device, dtype=torch.float32)
    wt = None
    blocks = []
    counts = []
    for s in range(0, n_docs, block_docs):
        e = min(n_docs, s + block_docs)
        u = torch.rand((e - s, L), generator=g, device=device)
        terms = torch.searchsorted(cdf, u).clamp_(max=V - 1)
        docs = torch.arange(s, e, device=device, dtype=torch.int64)[:, None]
        key = (terms * n_docs + docs).reshape(-1)      # term-major key inside the block
        del u, terms
        key, _ = torch.sort(key)
        uk, tf = torch.unique_consecutive(key, return_counts=True)
        del key
        term = uk // n_docs
        doc = (uk - term * n_docs).to(torch.int32)
        cnt = torch.bincount(term, minlength=V)
        blocks.append((term.to(torch.int32), doc, tf.to(torch.int16)))
        del uk, term, tf
        counts.append(cnt)
    counts = torch.stack(counts)                        # [nb, V]
    total = counts.sum(0)
    term_ptr = torch.zeros(V + 1, dtype=torch.int64, device=device)
    term_ptr[1:] = torch.cumsum(total, 0)
    nnz = int(term_ptr[-1].item())
    doc_out = torch.empty(nnz, dtype=torch.int32, device=device)
    w_out = torch.empty(nnz, dtype=torch.float32, device=device)
    max_tf = max(int(b[2].max().item()) for b in blocks)
    wt = torch.from_numpy(bm25_weight_table(L, max_tf)).to(device)
    base = term_ptr[:-1].clone()
    for (term, doc, tf), cnt in zip(blocks, counts):
        term = term.long()
        start = torch.cumsum(cnt, 0) - cnt                    # first index of each term inside the block
        pos = torch.arange(term.numel(), device=device) - start[term]
        dest = base[term] + pos
        doc_out[dest] = doc
        w_out[dest] = wt[tf.long()]
        base += cnt
        del term, pos, dest
    return term_ptr, doc_out, w_out


def sparse_queries_torch(term_ptr, B: int, n_terms: int = 8, seed: int = SEED_QUERY, skip_top: int = 64):
    """zipf regime on device; returns numpy-compatible CPU tensors (q_ptr, q_terms, q_vals)."""
    import torch
    df = (term_ptr[1:] - term_ptr[:-1]).double().cpu()
    p = df.clone()
    p[torch.argsort(-df)[:skip_top]] = 0
    g = torch.Generator()
    g.manual_seed(seed)
    terms = torch.multinomial((p / p.sum()).expand(B, -1), n_terms, replacement=False, generator=g)
    terms, _ = torch.sort(terms, dim=1)
    q_ptr = torch.arange(0, (B + 1) * n_terms, n_terms, dtype=torch.int32)
    qv = float(bm25_weight_table(n_terms, 1)[1])
    return q_ptr, terms.reshape(-1).to(torch.int32), torch.full((B * n_terms,), qv, dtype=torch.float64)
