#!/usr/bin/env python
"""Single-query latency of one 1/8 shard of C4 (1.25 M x 1024 bf16 + 1.25 M BM25 docs) on one GPU: the fused hybrid
call, and the sharded host path (local lists -> pack -> [all-gather skipped at world 1] -> unpack -> merge x2 -> RRF),
to see what the exchange's host side costs beside the scans.  usage: python tools/shard_b1_probe.py"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rag_application_b200 import Engine, synth  # noqa: E402
from rag_application_b200.sharded import ShardedHybrid  # noqa: E402

dev = torch.device("cuda", 0)
n, d, k = 1_250_000, 1024, 100
eng = Engine(0)
x = synth.dense_corpus_torch(n, d, torch.bfloat16, dev)
q = synth.dense_queries_torch(x, 8)
eng.dense_set("dense", x)
del x
tp, doc, w = synth.zipf_corpus_torch(n, 50000, 285, dev)
qp, qt, qv = synth.sparse_queries_torch(tp, 8, 8)
eng.sparse_set_csr("sparse", tp, doc, w, n)
del doc, w
torch.cuda.empty_cache()
q1 = q[:1].contiguous()
qp1, qt1, qv1 = qp.to(dev)[:2].contiguous(), qt.to(dev)[:8].contiguous(), qv.to(dev)[:8].contiguous()
sh = ShardedHybrid(eng, 0, 1, 0)


def p50(fn, iters=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    lat = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        lat.append(a.elapsed_time(b))
    return float(np.median(lat))


out = {"rows": n, "hbm_floor_ms": n * d * 2 / 6.5488e12 * 1e3}
out["fused_hybrid_ms"] = p50(lambda: eng.hybrid_topk("dense", "sparse", q1, qp1, qt1, qv1, k))
out["sharded_path_ms"] = p50(lambda: sh.hybrid_topk("dense", "sparse", q1, qp1, qt1, qv1, k))
out["dense_topk_ms"] = p50(lambda: eng.dense_topk("dense", q1, k))
out["sparse_topk_ms"] = p50(lambda: eng.sparse_topk("sparse", qp1, qt1, qv1, k))
print(json.dumps(out))
