#!/usr/bin/env python
"""Experiment: can a few BM25 worker warps per SM run beside the resident tcgen05 dense scan (different pipes, the
dense CTA leaves ~27 KB of shared memory), and what does that do to both under the power cap?  Two engine handles on
one GPU (separate scratch), two streams.  Run with BRAG_BX_WORKERS=2 (or 1, 3) so the sparse CTA fits beside the
dense one.  usage: BRAG_BX_WORKERS=2 python tools/overlap_probe.py"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rag_application_b200 import Engine, synth  # noqa: E402

dev = torch.device("cuda", 0)
n, d, B, k = 10_000_000, 1024, 1024, 100
ed, es = Engine(0), Engine(0)
x = synth.dense_corpus_torch(n, d, torch.bfloat16, dev)
q = synth.dense_queries_torch(x, B)
ed.dense_set("dense", x)
del x
tp, doc, w = synth.zipf_corpus_torch(n, 50000, 285, dev)
qp, qt, qv = synth.sparse_queries_torch(tp, B, 8)
es.sparse_set_csr("sparse", tp, doc, w, n)
del doc, w
torch.cuda.empty_cache()
qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
es.force_sparse_path("bitmap")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(dense, sparse, iters=5):
    def once():
        if dense:
            with torch.cuda.stream(s1):
                ed.dense_topk("dense", q, k)
        if sparse:
            with torch.cuda.stream(s2):
                es.sparse_topk("sparse", qp_d, qt_d, qv_d, k)
    for _ in range(2):
        once()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    s1.wait_event(a); s2.wait_event(a)
    for _ in range(iters):
        once()
    e1, e2 = torch.cuda.Event(), torch.cuda.Event()
    e1.record(s1); e2.record(s2)
    torch.cuda.current_stream().wait_event(e1); torch.cuda.current_stream().wait_event(e2)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


out = {"workers": os.environ.get("BRAG_BX_WORKERS", "all")}
out["dense_alone_ms"] = run(True, False)
out["sparse_alone_ms"] = run(False, True)
out["both_streams_ms"] = run(True, True)
print(json.dumps(out))
