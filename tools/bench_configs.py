#!/usr/bin/env python
"""Per-stage numbers for the BASELINE.json configs that are not the headline bench line
(C2 dense 1M x 768 fp16 at B=1 and B=1024, C3 BM25 1M docs B=1024, C5-shard filtered dense).
Writes one JSON object; run on the GPU box: python tools/bench_configs.py > gpurun_out/configs.json"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rag_application_b200 import Engine, synth  # noqa: E402

dev = torch.device("cuda", 0)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(np.min(ts))


out = {"peaks": {k: peaks.get(k) for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained")}}
eng = Engine(0)
eng.profile_enable(True)


def stage_ms(fn, stage, iters=10):
    fn(); fn()
    eng.profile_reset()
    for _ in range(iters):
        fn()
    ms, calls, _ = eng.profile_read()
    return ms[stage] / max(1, calls[stage])


# ---- C2: dense exact top-100, 1M x 768 fp16
n, d, k = 1 << 20, 768, 100
x = synth.dense_corpus_torch(n, d, torch.float16, dev)
q = synth.dense_queries_torch(x, 1024)
eng.dense_set("c2", x)
del x
for B in (1, 8, 1024):
    qb = q[:B].contiguous()
    p50, best = timeit(lambda: eng.dense_topk("c2", qb, k))
    scan = stage_ms(lambda: eng.dense_topk("c2", qb, k), "dense_scan")
    rec = {"call_p50_ms": p50, "call_min_ms": best, "scan_kernel_ms": scan, "path": eng.dense_path("c2", B)}
    if B <= 8:
        byt = n * d * 2
        rec.update(bound="hbm", achieved_gbs=byt / (scan * 1e-3) / 1e9, frac=byt / (scan * 1e-3) / 1e9 / peaks["hbm_gbs"],
                   call_gbs=byt / (p50 * 1e-3) / 1e9)
    else:
        fl = 2.0 * n * d * B
        rec.update(bound="tensor", achieved_tflops=fl / (scan * 1e-3) / 1e12,
                   frac_burst=fl / (scan * 1e-3) / 1e12 / peaks["bf16_tflops"],
                   frac_sustained=fl / (scan * 1e-3) / 1e12 / peaks["bf16_tflops_sustained"])
    out[f"C2_dense_1Mx768_f16_B{B}"] = rec
# C5-like shard: 10 % filter bitmask at batch 1 (row skipping) and batch 1024
mask = torch.rand(n, device=dev) < 0.10
words = torch.zeros((n + 31) // 32, dtype=torch.int64, device=dev)
idx = torch.nonzero(mask).squeeze(1)
words.scatter_add_(0, idx // 32, (1 << (idx % 32)).to(torch.int64))
eng.mask_set(1, words.to(torch.int32))
for B in (1, 1024):
    qb = q[:B].contiguous()
    scan = stage_ms(lambda: eng.dense_topk("c2", qb, k, mask_id=1), "dense_scan")
    out[f"C5like_dense_1Mx768_f16_mask10pct_B{B}"] = {"scan_kernel_ms": scan, "rows_kept": int(mask.sum().item())}
eng.dense_drop("c2")
torch.cuda.empty_cache()

# ---- C3: BM25 top-100 over 1M docs, 50k-term Zipf vocab, 8-term queries, batch 1024
nd, V = 1 << 20, 50000
tp, doc, w = synth.zipf_corpus_torch(nd, V, 285, dev)
qp, qt, qv = synth.sparse_queries_torch(tp, 1024, 8)
eng.sparse_set_csr("c3", tp, doc, w, nd)
df = (tp[1:] - tp[:-1])
byt = float(df[qt.to(dev).long()].sum().item()) * 8.0
qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
p50, best = timeit(lambda: eng.sparse_topk("c3", qp_d, qt_d, qv_d, k), iters=10)
scan = stage_ms(lambda: eng.sparse_topk("c3", qp_d, qt_d, qv_d, k), "sparse_scan")
out["C3_bm25_1Mdocs_B1024"] = {"call_p50_ms": p50, "scan_kernel_ms": scan, "algorithmic_bytes": byt, "nnz": int(tp[-1].item()),
                               "achieved_gbs": byt / (scan * 1e-3) / 1e9, "frac": byt / (scan * 1e-3) / 1e9 / peaks["hbm_gbs"],
                               "qps": 1024 / (p50 * 1e-3)}
qp1, qt1, qv1 = qp_d[:2].contiguous(), qt_d[:8].contiguous(), qv_d[:8].contiguous()
p50, best = timeit(lambda: eng.sparse_topk("c3", qp1, qt1, qv1, k), iters=20)
out["C3_bm25_1Mdocs_B1"] = {"call_p50_ms": p50, "call_min_ms": best}
print(json.dumps(out, indent=1))
