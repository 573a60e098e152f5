#!/usr/bin/env python
"""Dense stage alone: scan time per (n, d, batch) with per-stage timers; run under ncu for a launch list.
usage: python tools/dense_probe.py [--n 1048576] [--d 768] [--batch 1024] [--dtype f16] [--mask 0.1]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rag_application_b200 import Engine, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1 << 20)
ap.add_argument("--d", type=int, default=768)
ap.add_argument("--batch", default="1024")
ap.add_argument("--dtype", default="f16")
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--mask", type=float, default=0.0)
ap.add_argument("--dump", default="", help="write the last batch's (scores, ids) to this .npz (to compare runs under different knobs)")
a = ap.parse_args()
dev = torch.device("cuda", 0)
eng = Engine(0)
x = synth.dense_corpus_torch(a.n, a.d, torch.float16 if a.dtype == "f16" else torch.bfloat16, dev)
batches = [int(b) for b in a.batch.split(",")]
q = synth.dense_queries_torch(x, max(batches))
eng.dense_set("x", x)
del x
mask_id = 0
if a.mask > 0:
    g = torch.Generator(device=dev)
    g.manual_seed(synth.SEED_MASK)
    eng.mask_set(1, (torch.rand(a.n, generator=g, device=dev) < a.mask).cpu().numpy())
    mask_id = 1
eng.profile_enable(True)
out = {"n": a.n, "d": a.d, "mask": a.mask}
for B in batches:
    qb = q[:B].contiguous()
    for _ in range(2):
        eng.dense_topk("x", qb, a.k, mask_id=mask_id)
    eng.profile_reset()
    for _ in range(a.iters):
        r = eng.dense_topk("x", qb, a.k, mask_id=mask_id)
    ms, calls, _ = eng.profile_read()
    scan = ms["dense_scan"] / a.iters
    out[f"B{B}"] = {"scan_ms": scan, "select_ms": ms["dense_select"] / a.iters,
                    "tflops": 2.0 * a.n * a.d * B / (scan * 1e-3) / 1e12, "gbs": a.n * a.d * 2 / (scan * 1e-3) / 1e9}
if a.dump:
    import numpy as np
    np.savez(a.dump, **{f"r{i}": t.cpu().numpy() for i, t in enumerate(r)})
print(json.dumps(out))
