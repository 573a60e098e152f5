#!/usr/bin/env python
"""Sparse stage alone at benchmark scale: both scan kernels (scatter-accumulate / bitmap filter) timed on the same
synthetic Zipf corpus and query batch, results compared bit for bit (a full-size parity property: the two kernels
share no scoring code).  usage: python tools/bm25_probe.py [--n 10000000] [--batch 1024] [--paths bitmap,scatter]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rag_application_b200 import Engine, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=10_000_000)
ap.add_argument("--vocab", type=int, default=50_000)
ap.add_argument("--doc-len", type=int, default=285)
ap.add_argument("--batch", type=int, default=1024)
ap.add_argument("--q-terms", type=int, default=8)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--paths", default="bitmap,scatter")
ap.add_argument("--mask", type=float, default=0.0, help="payload-filter selectivity (0 = no filter)")
a = ap.parse_args()

dev = torch.device("cuda", 0)
eng = Engine(0)
tp, doc, w = synth.zipf_corpus_torch(a.n, a.vocab, a.doc_len, dev)
qp, qt, qv = synth.sparse_queries_torch(tp, a.batch, a.q_terms)
ldf = tp[1:] - tp[:-1]
alg_bytes = float(ldf[qt.to(dev).long()].sum().item()) * 8.0
import time
torch.cuda.synchronize()
_t0 = time.perf_counter()
eng.sparse_set_csr("s", tp, doc, w, a.n)
torch.cuda.synchronize()
index_build_ms = (time.perf_counter() - _t0) * 1e3      # device-resident CSR -> interleave + every scan index
del doc, w
torch.cuda.empty_cache()
mask_id = 0
if a.mask > 0:
    g = torch.Generator(device=dev)
    g.manual_seed(synth.SEED_MASK)
    eng.mask_set(1, (torch.rand(a.n, generator=g, device=dev) < a.mask).cpu().numpy())
    mask_id = 1
qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
eng.profile_enable(True)
out = {"n": a.n, "batch": a.batch, "k": a.k, "algorithmic_bytes": alg_bytes, "mask": a.mask,
       "index_build_ms": index_build_ms}
res = {}
for path in a.paths.split(","):
    eng.force_sparse_path(path)
    for _ in range(2):
        r = eng.sparse_topk("s", qp_d, qt_d, qv_d, a.k, mask_id=mask_id)
    eng.profile_reset()
    for _ in range(a.iters):
        r = eng.sparse_topk("s", qp_d, qt_d, qv_d, a.k, mask_id=mask_id)
    ms, calls, _ = eng.profile_read()
    scan = ms["sparse_scan"] / max(1, calls["sparse_scan"])
    out[path] = {"scan_ms": scan, "select_ms": ms["sparse_select"] / max(1, calls["sparse_select"]),
                 "algorithmic_gbs": alg_bytes / (scan * 1e-3) / 1e9}
    res[path] = [t.cpu().numpy() for t in r]
eng.force_sparse_path(None)
if len(res) == 2:
    (s0, i0, c0), (s1, i1, c1) = res.values()
    out["kernels_agree_bit_exact"] = bool(np.array_equal(i0, i1) and np.array_equal(c0, c1)
                                          and np.array_equal(s0.view(np.uint32), s1.view(np.uint32)))
print(json.dumps(out))
