#!/usr/bin/env python
"""Generates tests/golden/*.json -- small known-answer vectors for the hybrid-retrieval path.

PARITY UNPINNED: the reference's engine (qdrant-client local mode, un-vendored and un-pinned,
reference requirements.txt:3) is not importable in the build container, so these vectors are NOT
outputs of the reference.  They come from an INDEPENDENT brute-force restatement written here with
plain Python loops and Python floats (float64) -- deliberately sharing no code with oracle/ -- of
the semantics SURVEY.md section 8c lists for the requests the reference issues
(app/core/vector_store/qdrant/qdrant_handler.py:296-372).  The oracle (CPU tests) and the CUDA
path (GPU tests) are both checked against them.  On any machine where `import qdrant_client`
works, `python tests/golden/make_golden.py --qdrant` additionally replays the same points through
QdrantClient(":memory:") and reports disagreements; nothing in the test-suite depends on that.

Run from the repo root:  python tests/golden/make_golden.py
"""
from __future__ import annotations

import json
import math
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def f32(x: float) -> float:
    return struct.unpack("f", struct.pack("f", x))[0]


# ----------------------------------------------------------------------------- brute force
def bf_normalize(v):
    """float32 L2 normalisation the way numpy does it on float32 arrays (qdrant local mode)."""
    a = np.asarray(v, dtype=np.float32)
    n = np.float32(np.linalg.norm(a))
    if n == 0:
        n = np.float32(np.finfo(np.float32).eps)
    return [float(t) for t in (a / n).astype(np.float32)]


def bf_dot(a, b):
    s = 0.0
    for x, y in zip(a, b):
        s += float(x) * float(y)
    return f32(s)


def bf_rank(scored):
    """[(score, id)] -> sorted by (score desc, id asc)."""
    return sorted(scored, key=lambda t: (-t[0], t[1]))


def bf_dense(rows, q, k, cosine, mask=None, cand=None):
    if cosine:
        rows = [bf_normalize(r) for r in rows]
        q = bf_normalize(q)
    ids = range(len(rows)) if cand is None else sorted(set(cand))
    sc = [(bf_dot(rows[i], q), i) for i in ids if mask is None or mask[i]]
    return bf_rank(sc)[:k]


def bf_sparse(docs, q_idx, q_val, k, mask=None):
    """docs: list of (indices, values); per-doc two-pointer over sorted indices; only docs that
    share an index are candidates; Python-float accumulation, float32 result."""
    out = []
    order = sorted(range(len(q_idx)), key=lambda i: q_idx[i])
    qi = [q_idx[i] for i in order]
    qv = [q_val[i] for i in order]
    for d, (di, dv) in enumerate(docs):
        if mask is not None and not mask[d]:
            continue
        o = sorted(range(len(di)), key=lambda i: di[i])
        di2, dv2 = [di[i] for i in o], [dv[i] for i in o]
        i = j = 0
        s, hit = 0.0, False
        while i < len(qi) and j < len(di2):
            if qi[i] == di2[j]:
                s += float(qv[i]) * float(f32(dv2[j]))
                hit = True
                i += 1
                j += 1
            elif qi[i] < di2[j]:
                i += 1
            else:
                j += 1
        if hit:
            out.append((f32(s), d))
    return bf_rank(out)[:k]


def bf_rrf(lists, limit):
    sc = {}
    for lst in lists:
        for pos, pid in enumerate(lst):
            sc[pid] = sc.get(pid, 0.0) + 1.0 / (2 + pos)
    return bf_rank([(s, i) for i, s in sc.items()])[:limit]


def bf_idf(n, df):
    return math.log(1.0 + (n - df + 0.5) / (df + 0.5))


def bf_quant(v):
    """(x*127).astype(int8): truncation toward zero (qdrant_handler.py:144-146, :300-302)."""
    return [float(int(float(x) * 127)) for x in v]


# ----------------------------------------------------------------------------- cases
def case_dense_small():
    rows = [[1, 0, 0, 0], [0.6, 0.8, 0, 0], [0, 0, 0, 0], [0.6, 0.8, 0, 0], [-1, 0, 0, 0]]
    q = [2, 0, 0, 0]
    return {"rows": rows, "q": q,
            "cosine_top3": bf_dense(rows, q, 3, True), "dot_top3": bf_dense(rows, q, 3, False),
            "cosine_top3_mask": bf_dense(rows, q, 3, True, mask=[0, 1, 1, 1, 1]),
            "note": "row 2 is the zero vector (eps guard); rows 1 and 3 tie exactly (lower id first)"}


def case_sparse_small():
    docs = [([1, 5, 9], [1.0, 0.5, 2.0]), ([2, 5], [1.5, 1.0]), ([7], [3.0]), ([1, 2, 9], [0.25, 0.5, 0.75]),
            ([9], [1.0]), ([3, 4], [1.0, 1.0])]
    q_idx, q_val = [9, 1, 5], [1.0, 2.0, 0.5]
    n = len(docs)
    df = {t: sum(1 for d in docs if t in d[0]) for t in q_idx}
    q_idf = [v * bf_idf(n, df[t]) for t, v in zip(q_idx, q_val)]
    return {"docs": docs, "q_idx": q_idx, "q_val": q_val,
            "top10": bf_sparse(docs, q_idx, q_val, 10), "top2": bf_sparse(docs, q_idx, q_val, 2),
            "q_val_idf": q_idf, "top10_idf": bf_sparse(docs, q_idx, q_idf, 10),
            "note": "docs 2 and 5 share no index with the query and must be absent"}


def case_rrf():
    a, b, c, d = 10, 11, 12, 13
    lists = [[a, b, c], [c, a, d]]
    long = [list(range(100, 130)), list(range(115, 145))]
    return {"lists": lists, "fused": bf_rrf(lists, 10), "long_lists": long, "long_default_limit": bf_rrf(long, 10),
            "long_limit_100": bf_rrf(long, 100),
            "note": "a:1/2+1/3, c:1/4+1/2, b:1/3, d:1/4 -> a,c,b,d ; default fusion limit is 10"}


def case_quant():
    xs = [0.999, -0.5, 1.0, -1.0, 0.0039, -0.0079, 0.5, 0.0078]
    return {"x": xs, "q": bf_quant(xs), "note": "0.999->126, -0.5->-63, 1.0->127 (trunc toward zero)"}


def case_tree(seed=7, n=200, d=768, V=60):
    """The exact prefetch tree of QdrantHandler.hybrid_search on a seeded corpus."""
    rng = np.random.Generator(np.random.PCG64(seed))
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    qv = (x[17] + 0.2 * rng.standard_normal(d)).astype(np.float32)
    qv /= np.linalg.norm(qv)
    docs = []
    for _ in range(n):
        m = int(rng.integers(0, 9))
        idx = sorted(int(t) for t in rng.choice(V, m, replace=False))
        docs.append((idx, [f32(float(v)) for v in rng.random(m) + 0.1]))
    q_idx = sorted(int(t) for t in rng.choice(V, 5, replace=False))
    q_val = [float(v) for v in rng.random(5) + 0.5]
    sp = {"matryoshka_64_limit": 100, "matryoshka_128_limit": 80, "matryoshka_256_limit": 60, "dense_limit": 40,
          "quantized_limit": 100, "sparse_limit": 50, "final_limit": 30}
    rows = [[float(t) for t in r] for r in x]
    q = [float(t) for t in qv]
    mask = [bool(b) for b in (rng.random(n) < 0.6)]

    def pre(k):
        return [r[:k] for r in rows]

    quant_rows = [bf_quant(r) for r in rows]
    p1 = bf_dense(pre(64), q[:64], sp["matryoshka_64_limit"], True)
    p2 = bf_dense(pre(128), q[:128], sp["matryoshka_128_limit"], True, cand=[i for _, i in p1])
    p3 = bf_dense(pre(256), q[:256], sp["matryoshka_256_limit"], True, cand=[i for _, i in p2])
    p4 = bf_dense(rows, q, sp["dense_limit"], True, cand=[i for _, i in p3])
    q1 = bf_dense(quant_rows, bf_quant(q), sp["quantized_limit"], True)
    q2 = bf_dense(rows, q, sp["dense_limit"], True, cand=[i for _, i in q1])
    s1 = bf_sparse(docs, q_idx, q_val, sp["sparse_limit"])
    f1 = bf_rrf([[i for _, i in q2], [i for _, i in s1]], 10)
    cand = [i for _, i in p4] + [i for _, i in f1]
    r = bf_dense(rows, q, sp["final_limit"], True, cand=cand)
    rm = bf_dense(rows, q, sp["final_limit"], True, cand=cand, mask=mask)
    # the same request answered with SERVER semantics: Qdrant merges the root query_filter into every prefetch
    # (the reference talks to a server, qdrant_session.py:21), so every stage ranks only the passing rows
    m1 = bf_dense(pre(64), q[:64], sp["matryoshka_64_limit"], True, mask=mask)
    m2 = bf_dense(pre(128), q[:128], sp["matryoshka_128_limit"], True, cand=[i for _, i in m1], mask=mask)
    m3 = bf_dense(pre(256), q[:256], sp["matryoshka_256_limit"], True, cand=[i for _, i in m2], mask=mask)
    m4 = bf_dense(rows, q, sp["dense_limit"], True, cand=[i for _, i in m3], mask=mask)
    n1 = bf_dense(quant_rows, bf_quant(q), sp["quantized_limit"], True, mask=mask)
    n2 = bf_dense(rows, q, sp["dense_limit"], True, cand=[i for _, i in n1], mask=mask)
    t1 = bf_sparse(docs, q_idx, q_val, sp["sparse_limit"], mask=mask)
    g1 = bf_rrf([[i for _, i in n2], [i for _, i in t1]], 10)
    rs = bf_dense(rows, q, sp["final_limit"], True, cand=[i for _, i in m4] + [i for _, i in g1], mask=mask)
    return {"seed": seed, "n": n, "d": d, "V": V, "search_params": sp,
            "docs": docs, "q_idx": q_idx, "q_val": q_val, "mask": mask,
            "stages": {"P1": p1, "P4": p4, "Q1": q1, "Q2": q2, "S1": s1, "F1": f1},
            "final": r, "final_masked": rm, "final_masked_server": rs,
            "note": "dense rows/query are regenerated from the seed by the test (PCG64, same draws)"}


def main():
    out = {"dense_small": case_dense_small(), "sparse_small": case_sparse_small(), "rrf": case_rrf(),
           "quant": case_quant(), "tree": case_tree()}
    path = os.path.join(HERE, "kat.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path, os.path.getsize(path), "bytes")
    if "--qdrant" in sys.argv:
        replay_qdrant(out)


def replay_qdrant(out):  # pragma: no cover - needs the absent dependency
    from qdrant_client import QdrantClient, models
    c = out["dense_small"]
    cl = QdrantClient(":memory:")
    cl.create_collection("t", vectors_config={"v": models.VectorParams(size=4, distance=models.Distance.COSINE)})
    cl.upsert("t", [models.PointStruct(id=i, vector={"v": r}, payload={}) for i, r in enumerate(c["rows"])])
    got = cl.query_points("t", query=c["q"], using="v", limit=3).points
    print("qdrant-client cosine top3:", [(p.score, p.id) for p in got], "golden:", c["cosine_top3"])


if __name__ == "__main__":
    main()
