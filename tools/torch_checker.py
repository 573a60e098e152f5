"""Independent checker for config-sized answers: plain torch on the device, written from the definitions.

TEST / BENCH INFRASTRUCTURE.  It shares no code with libbrag.so and none with oracle/: scores are recomputed from the
raw corpus tensors (dense: chunked float64 `X[i:j] @ q`; sparse: float64 index_add_ over the raw term-major CSR, terms
in ascending order, which is the reference's accumulation order), ranked under (score desc, id asc) and
fused with RRF (constant 2).  Used by tests/test_config_parity.py at the BASELINE config sizes and by bench.py's
`parity` object on the timed step's answers (outside the timed region)."""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

TIE_EPS = 4e-7          # |relative score gap| below which two dense rows count as a near-tie (fp32 rounding of a f64 sum)


def _rank(scores: np.ndarray, ids: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    o = np.lexsort((ids, -scores.astype(np.float64)))[:k]
    return scores[o], ids[o]


def round_like_storage(q, dtype):
    """A float32 query rounded to the corpus storage dtype (what the engine scores with)."""
    import torch
    return q.to(dtype).to(torch.float32) if dtype != torch.float32 else q


def dense_scores(X, q, chunk: int = 1 << 18):
    """float32(float64 dot) of every stored row with the float32 query `q` [d]; X [n, d] in its storage dtype."""
    import torch
    out = torch.empty(X.shape[0], dtype=torch.float32, device=X.device)
    qd = q.to(torch.float64)
    for s in range(0, X.shape[0], chunk):
        e = min(X.shape[0], s + chunk)
        out[s:e] = (X[s:e].to(torch.float64) @ qd).to(torch.float32)
    return out


def dense_topk(X, q, k: int, mask=None, id_offset: int = 0, margin: int = 64):
    """(scores f32, global ids) of the k best rows under (score desc, id asc); mask = bool tensor [n] or None.
    Also returns the full score vector (device) so a caller can explain near-tie swaps."""
    import torch
    sc = dense_scores(X, q)
    use = sc if mask is None else torch.where(mask, sc, torch.full_like(sc, -float("inf")))
    kk = min(use.numel(), k + margin)
    top_s, top_i = torch.topk(use, kk)
    top_s, top_i = top_s.cpu().numpy(), top_i.cpu().numpy().astype(np.int64)
    keep = np.isfinite(top_s) if mask is not None else np.ones(kk, bool)
    s, i = _rank(top_s[keep], top_i[keep], k)
    if kk > k and len(s) == k and keep.sum() == kk:
        assert top_s.min() < s[-1], "checker: more than `margin` rows tie with the k-th score"
    return s, i + id_offset, sc


def sparse_topk(term_ptr, doc, w, n_docs: int, q_terms: Sequence[int], q_vals: Sequence[float], k: int, mask=None,
                id_offset: int = 0, margin: int = 64):
    """Overlap-only sparse dot: float64 accumulation in ascending term order, float32 score, top-k (score desc, id asc).
    term_ptr [V+1] (host or device), doc int32 [nnz], w float32 [nnz] on the device."""
    import torch
    dev = doc.device
    acc = torch.zeros(n_docs, dtype=torch.float64, device=dev)
    touched = torch.zeros(n_docs, dtype=torch.bool, device=dev)
    tp = term_ptr.cpu() if hasattr(term_ptr, "cpu") else torch.as_tensor(term_ptr)
    last = -1
    for t, v in zip(q_terms, q_vals):
        t = int(t)
        assert t > last
        last = t
        if t < 0 or t >= tp.numel() - 1:
            continue
        lo, hi = int(tp[t]), int(tp[t + 1])
        d = doc[lo:hi].long()
        acc.index_add_(0, d, w[lo:hi].to(torch.float64) * float(v))      # a term's documents are distinct
        touched[d] = True
    if mask is not None:
        touched &= mask
    sc = torch.where(touched, acc.to(torch.float32), torch.full((n_docs,), -float("inf"), device=dev))
    kk = min(n_docs, k + margin)
    top_s, top_i = torch.topk(sc, kk)
    top_s, top_i = top_s.cpu().numpy(), top_i.cpu().numpy().astype(np.int64)
    keep = np.isfinite(top_s)
    s, i = _rank(top_s[keep], top_i[keep], k)
    if keep.sum() == kk and kk > k:
        # ties at the boundary beyond the margin would make the id set ambiguous: widen until resolved
        if top_s.min() == s[-1]:
            return sparse_topk(term_ptr, doc, w, n_docs, q_terms, q_vals, k, mask, id_offset, margin * 8)
    return s, i + id_offset


def rrf(lists: Sequence[Sequence[int]], limit: int) -> Tuple[np.ndarray, np.ndarray]:
    """score(id) = sum over lists of 1 / (2 + rank), rank 0-based, float64, accumulated in list order;
    order (score desc, id asc)."""
    score = {}
    for lst in lists:
        for pos, pid in enumerate(lst):
            pid = int(pid)
            score[pid] = score.get(pid, 0.0) + 1.0 / (2 + pos)
    items = sorted(score.items(), key=lambda kv: (-kv[1], kv[0]))[:limit]
    return np.array([s for _, s in items], np.float64), np.array([i for i, _ in items], np.int64)


def dense_lists_agree(got_s, got_i, exp_s, exp_i, all_scores=None, id_offset: int = 0, rtol: float = 1e-3) -> dict:
    """Compare one query's dense list with the checker's: ids equal, or every disagreement a near-tie in the checker's
    own scores; scores within rtol.  Returns {"ids_equal", "explained", "max_rel_err"}."""
    got_s, got_i = np.asarray(got_s), np.asarray(got_i)
    n = len(exp_i)
    res = {"ids_equal": bool(len(got_i) >= n and np.array_equal(got_i[:n], exp_i)), "explained": True, "max_rel_err": 0.0}
    if len(got_i) < n:
        res["explained"] = False
        return res
    if all_scores is not None and hasattr(all_scores, "cpu"):
        import torch
        idx = torch.as_tensor(np.asarray(got_i[:n] - id_offset, dtype=np.int64), device=all_scores.device)
        ref = all_scores[idx].cpu().numpy()
    elif all_scores is not None:
        ref = np.asarray(all_scores)[got_i[:n] - id_offset]
    else:
        ref = exp_s
    den = np.maximum(np.abs(ref), 1e-6)
    res["max_rel_err"] = float(np.max(np.abs(got_s[:n] - ref) / den)) if n else 0.0
    if res["max_rel_err"] > rtol:
        res["explained"] = False
    if not res["ids_equal"]:
        if all_scores is None:
            res["explained"] = False
        else:
            for pos in np.nonzero(got_i[:n] != exp_i)[0]:
                if abs(float(ref[pos]) - float(exp_s[pos])) > TIE_EPS * max(1.0, abs(float(exp_s[pos]))):
                    res["explained"] = False
    return res
