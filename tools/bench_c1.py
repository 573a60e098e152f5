#!/usr/bin/env python
"""BASELINE.json configs[0] (C1), both sides: 10 000 chunks, 384-d float32 COSINE + BM25 sparse, RRF top-10, 100 queries.

  * the FAITHFUL CPU path (what qdrant-client local mode executes): float32 `X @ q` + full argsort per
    query, and one two-pointer sparse dot per document per query (pure Python; timed on a few queries, the rest scaled);
  * the fast CPU path (oracle/fast_cpu.py: BLAS + C inverted index), all host cores;
  * the GPU through the C ABI with host buffers (float32 storage -> dense_gemv_kernel), asserted equal to the oracle on
    all 100 queries before anything is timed;
  * the reference's whole prefetch tree (qdrant_handler.py:305-372) through B200QdrantClient with the default limits of
    mcp/qdrant_search_mcp_endpoint.py:19-28, checked against the oracle's plan executor;
  * the float32 GEMV scan's HBM fraction on a corpus large enough to measure (1 M x 384 float32).
Run on the GPU box: python tools/bench_c1.py > gpurun_out/c1.json"""
import asyncio
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
cores = os.cpu_count() or 1
for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
    os.environ[var] = str(cores)

import torch  # noqa: E402
from oracle import fast_cpu, oracle as O  # noqa: E402
from rag_application_b200 import Engine, models as M, synth  # noqa: E402
from rag_application_b200.client import B200QdrantClient  # noqa: E402

peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
n, d, k, B, V = 10_000, 384, 10, 100, 50_000
x = O.normalize_rows(synth.dense_corpus(n, d, seed=synth.SEED_CORPUS))
q = O.normalize_rows(synth.dense_queries(x, B, seed=synth.SEED_QUERY))
ptr, doc, w, _ = synth.zipf_corpus(n, V, 285, seed=synth.SEED_CORPUS)
qp, qt, qv = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=synth.SEED_QUERY)
out = {"config": "C1: 10000 x 384 f32 COSINE + BM25 (V=50000, L=285), RRF top-10, 100 queries", "host_cores": cores}

# ---- oracle answers (the checker of everything below)
col = O.OracleCollection()
col.add_dense("dense", x, "cosine", "f32")
col.add_sparse("sparse", O.CSR(ptr, doc, w, n))
want = [O.hybrid_simple(col, q[b], qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]], k) for b in range(B)]

# ---- faithful CPU path
order = np.argsort(doc, kind="stable")
term_of = np.repeat(np.arange(V), np.diff(ptr))
dptr = np.searchsorted(doc[order], np.arange(n + 1))
docs = [(term_of[order[dptr[i]:dptr[i + 1]]].tolist(), w[order[dptr[i]:dptr[i + 1]]].tolist()) for i in range(n)]
t0 = time.perf_counter()
dense_lists = []
for b in range(B):
    sc = x @ q[b]
    dense_lists.append(np.argsort(sc)[::-1][:k])
t_dense = time.perf_counter() - t0
NS = 4
t0 = time.perf_counter()
sparse_lists = []
for b in range(NS):
    t, v = qt[qp[b]:qp[b + 1]].tolist(), qv[qp[b]:qp[b + 1]].tolist()
    scored = []
    for i, (di, dv) in enumerate(docs):
        r = O.sparse_dot(t, v, di, dv)
        if r is not None:
            scored.append((-float(r), i))
    scored.sort()
    sparse_lists.append(np.array([i for _, i in scored[:k]], np.int64))
t_sparse = (time.perf_counter() - t0) / NS
faithful_ok = True
for b in range(NS):
    _, ei = O.rrf([dense_lists[b], sparse_lists[b]], limit=k)
    faithful_ok &= bool(np.array_equal(ei, want[b][1]))
per_query = t_dense / B + t_sparse
out["cpu_faithful"] = {"dense_ms_per_query": t_dense / B * 1e3, "sparse_ms_per_query": t_sparse * 1e3,
                       "qps": 1.0 / per_query, "cores_used": "numpy BLAS threads for X@q; the sparse loop is one Python thread",
                       "sparse_sample": f"{NS} of {B} queries timed (pure-Python two-pointer loop over {n} documents each)",
                       "fused_ids_equal_oracle_on_sample": faithful_ok}

# ---- fast CPU path
fast_cpu.hybrid_batch(x, q, ptr, doc, w, n, qp, qt, qv, k, threads=cores)
t0 = time.perf_counter()
for _ in range(5):
    res = fast_cpu.hybrid_batch(x, q, ptr, doc, w, n, qp, qt, qv, k, threads=cores)
dt = (time.perf_counter() - t0) / 5
out["cpu_fast"] = {"ms_per_100_queries": dt * 1e3, "qps": B / dt, "cores": cores,
                   "ids_equal_oracle": all(np.array_equal(res[b][1], want[b][1]) for b in range(B))}

# ---- GPU, float32 storage (dense_gemv_kernel), host buffers through the C ABI
eng = Engine(0)
eng.dense_set("dense", x, "f32")
eng.sparse_set_csr("sparse", ptr, doc, w, n)
s, i, c = eng.hybrid_topk("dense", "sparse", q, qp, qt, qv, k)
ok = all(int(c[b]) == len(want[b][1]) and np.array_equal(i[b][:len(want[b][1])], want[b][1]) and
         np.array_equal(s[b][:len(want[b][1])], want[b][0]) for b in range(B))
assert ok, "GPU C1 answers differ from the oracle"
lat = []
for _ in range(20):
    t0 = time.perf_counter()
    eng.hybrid_topk("dense", "sparse", q, qp, qt, qv, k)
    lat.append(time.perf_counter() - t0)
one = []
for b in range(B):
    t0 = time.perf_counter()
    eng.hybrid_topk("dense", "sparse", q[b:b + 1], np.array([0, qp[b + 1] - qp[b]], np.int32), qt[qp[b]:qp[b + 1]],
                    qv[qp[b]:qp[b + 1]], k)
    one.append(time.perf_counter() - t0)
out["gpu"] = {"dense_path": eng.dense_path("dense", B), "ids_and_scores_equal_oracle": ok,
              "batch100_p50_ms": float(np.median(lat)) * 1e3, "batch100_qps": B / float(np.median(lat)),
              "single_query_p50_ms": float(np.median(one)) * 1e3, "single_query_qps": 1.0 / float(np.median(one)),
              "speedup_vs_faithful_cpu_batch100": (B / float(np.median(lat))) / (1.0 / per_query),
              "speedup_vs_fast_cpu_batch100": (B / float(np.median(lat))) / (B / dt)}
eng.close()

# ---- the reference's prefetch tree through the drop-in client at C1, default limits of the MCP endpoint
sp = dict(matryoshka_64_limit=100, matryoshka_128_limit=80, matryoshka_256_limit=60, dense_limit=40, quantized_limit=40,
          sparse_limit=50, final_limit=30)
from _client_cases import reference_create_collection, reference_hybrid_search, reference_points  # noqa: E402
doc_major = [(np.asarray(t, np.int64).tolist(), np.asarray(v, np.float32).tolist()) for t, v in docs]


async def tree():
    client = B200QdrantClient(device=0)
    await reference_create_collection(client, "u", dense=d)
    pts = reference_points(x, doc_major)
    for a in range(0, n, 2000):
        await client.upsert(collection_name="u", points=pts[a:a + 2000])
    lat, eq = [], 0
    ocol = O.OracleCollection()
    ocol.add_dense("dense", x, "cosine", "f32")
    xq = np.stack([np.asarray(p.vector["quantized"], np.float32) for p in pts])
    ocol.add_dense("quantized", xq, "cosine", "f32")
    for m in (64, 128, 256):
        ocol.add_dense(f"matryoshka_{m}", x[:, :m], "cosine", "f32")
    ocol.add_sparse("sparse", O.CSR(ptr, doc, w, n))
    for b in range(20):
        sv = {"indices": qt[qp[b]:qp[b + 1]].tolist(), "values": qv[qp[b]:qp[b + 1]].tolist()}
        t0 = time.perf_counter()
        got = await reference_hybrid_search(client, "u", q[b].tolist(), sv, sp)
        lat.append(time.perf_counter() - t0)
        if b < 5:
            es, ei = O.run_node(ocol, O.reference_tree(q[b].tolist(), sv["indices"], sv["values"], sp))
            eq += int([p.payload["row"] for p in got] == ei.tolist())
    await client.close()
    return lat, eq

lat, eq = asyncio.run(tree())
out["reference_tree_through_client"] = {"limits": sp, "p50_ms": float(np.median(lat[1:])) * 1e3, "first_query_ms_incl_upload": lat[0] * 1e3,
                                         "rows_equal_oracle_plan": f"{eq}/5"}

# ---- float32 GEMV scan on a corpus large enough to time: 1 M x 384 float32
dev = torch.device("cuda", 0)
eng = Engine(0)
eng.profile_enable(True)
n2 = 1 << 20
x2 = synth.dense_corpus_torch(n2, d, torch.float32, dev)
q2 = synth.dense_queries_torch(x2, 8)
eng.dense_set("f32", x2)
for Bq in (1, 8):
    qb = q2[:Bq].contiguous()
    for _ in range(3):
        eng.dense_topk("f32", qb, 100)
    eng.profile_reset()
    for _ in range(20):
        eng.dense_topk("f32", qb, 100)
    ms, calls, _ = eng.profile_read()
    scan = ms["dense_scan"] / calls["dense_scan"]
    byt = n2 * d * 4
    out[f"gemv_f32_1Mx384_B{Bq}"] = {"scan_ms": scan, "achieved_gbs": byt / (scan * 1e-3) / 1e9,
                                     "frac_of_measured_hbm": byt / (scan * 1e-3) / 1e9 / peaks["hbm_gbs"]}
eng.close()
print(json.dumps(out, indent=1))
