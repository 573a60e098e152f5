#!/usr/bin/env python
"""The reference's own request (qdrant_handler.py:296-372: matryoshka 64 -> 128 -> 256 -> dense cascade, quantized ->
dense, sparse, RRF, final dense re-score) through B200QdrantClient on a collection built with the reference's upsert
shape (six named vectors per point), one query at a time as the reference issues them.  Reports p50 latency per
query and the engine's share of it.  usage: python tools/reference_plan_probe.py [--n 100000] [--queries 60]"""
import argparse
import asyncio
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from rag_application_b200 import models as M, synth  # noqa: E402
from rag_application_b200.client import B200QdrantClient  # noqa: E402
from _client_cases import reference_create_collection, reference_hybrid_search  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=100_000)
ap.add_argument("--queries", type=int, default=60)
ap.add_argument("--storage", default="bf16")
a = ap.parse_args()
n, d = a.n, 768
x = synth.dense_corpus(n, d, seed=1, clustered=True)
ptr, doc, w, _ = synth.zipf_corpus(n, 20000, 120, seed=2)
# doc-major view of the sparse corpus for the upsert
order = np.argsort(doc, kind="stable")
term = np.repeat(np.arange(len(ptr) - 1), np.diff(ptr))[order]
dw, dd = w[order], doc[order]
dptr = np.searchsorted(dd, np.arange(n + 1))
SP = {"matryoshka_64_limit": 2000, "matryoshka_128_limit": 1000, "matryoshka_256_limit": 500, "dense_limit": 100,
      "quantized_limit": 200, "sparse_limit": 200, "final_limit": 10}


async def main():
    client = B200QdrantClient(storage=a.storage)
    await reference_create_collection(client, "u")
    t0 = time.perf_counter()
    for lo in range(0, n, 20000):
        pts = []
        for i in range(lo, min(n, lo + 20000)):
            e = x[i]
            q8 = np.clip((e * 127).astype(np.int8), -128, 127).astype(np.float32)
            pts.append(M.PointStruct(id=i, vector={"dense": e, "matryoshka_64": e[:64], "matryoshka_128": e[:128],
                                                   "matryoshka_256": e[:256], "quantized": q8,
                                                   "sparse": {"indices": term[dptr[i]:dptr[i + 1]], "values": dw[dptr[i]:dptr[i + 1]]}},
                                     payload={"row": i}))
        await client.upsert(collection_name="u", points=pts)
    t_up = time.perf_counter() - t0
    rng = np.random.default_rng(3)
    qs = synth.dense_queries(x, a.queries, seed=4)
    lat = []
    eng = client.core.engine
    for j in range(a.queries + 5):
        dv = qs[j % a.queries].tolist()
        r = int(rng.integers(0, n))
        sv = {"indices": term[dptr[r]:dptr[r + 1]][:8].tolist(), "values": dw[dptr[r]:dptr[r + 1]][:8].tolist()}
        if j == 5:
            eng.profile_enable(True)
            eng.profile_reset()
        t0 = time.perf_counter()
        pts = await reference_hybrid_search(client, "u", dv, sv, SP)
        if j >= 5:
            lat.append((time.perf_counter() - t0) * 1e3)
        assert len(pts) == SP["final_limit"]
    ms, calls, launches = eng.profile_read()
    print(json.dumps({"points": n, "d": d, "storage": a.storage, "limits": SP, "upsert_and_first_commit_s": t_up,
                      "query_p50_ms": float(np.median(lat)), "query_p95_ms": float(np.percentile(lat, 95)),
                      "engine_ms_per_query": sum(ms.values()) / a.queries,
                      "engine_calls_per_query": sum(calls.values()) / a.queries}))
    await client.close()


asyncio.run(main())
