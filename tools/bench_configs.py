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
    rec = {"rows_kept": int(mask.sum().item())}
    for label, path in (("full_space_masked", "tcgen05_full"), ("compacted_copy", "tcgen05")):
        eng.force_dense_path(path)
        rec[label + "_scan_ms"] = stage_ms(lambda: eng.dense_topk("c2", qb, k, mask_id=1), "dense_scan")
    eng.force_dense_path(None)
    out[f"C5like_dense_1Mx768_f16_mask10pct_B{B}"] = rec
eng.dense_drop("c2")
torch.cuda.empty_cache()


def pack_mask(m):
    w = torch.zeros((m.numel() + 31) // 32, dtype=torch.int64, device=dev)
    ix = torch.nonzero(m).squeeze(1)
    w.scatter_add_(0, ix // 32, (1 << (ix % 32)).to(torch.int64))
    return w.to(torch.int32)


# ---- C5 shard (1 of 8 GPUs): 12.5M x 384 fp16 + BM25 over the same 12.5M docs, 10 % payload filter
if "--c5" in sys.argv:
    n5, d5 = 12_500_000, 384
    x = synth.dense_corpus_torch(n5, d5, torch.float16, dev)
    q5 = synth.dense_queries_torch(x, 4096)
    eng.dense_set("c5", x)
    del x
    torch.cuda.empty_cache()
    tp, doc, w = synth.zipf_corpus_torch(n5, 50000, 285, dev)
    qp, qt, qv = synth.sparse_queries_torch(tp, 4096, 8)
    eng.sparse_set_csr("c5s", tp, doc, w, n5)
    del doc, w
    torch.cuda.empty_cache()
    g = torch.Generator(device=dev)
    g.manual_seed(synth.SEED_MASK)
    m5 = torch.rand(n5, generator=g, device=dev) < 0.10
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eng.mask_set(2, pack_mask(m5))
    a.record()
    eng.dense_topk("c5", q5[:1].contiguous(), k, mask_id=2)       # first call under this filter: builds the copy
    b.record()
    torch.cuda.synchronize()
    rec = {"rows_kept": int(m5.sum().item()), "first_call_with_copy_build_ms": a.elapsed_time(b)}
    qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
    eng.sparse_compact_after(0)
    a.record()
    eng.sparse_topk("c5s", qp_d[:2].contiguous(), qt_d[:8].contiguous(), qv_d[:8].contiguous(), k, mask_id=2)
    b.record()
    torch.cuda.synchronize()
    rec["sparse_first_call_with_index_build_ms"] = a.elapsed_time(b)
    for B in (1, 4096):
        qb = q5[:B].contiguous()
        qpb, qtb, qvb = qp_d[:B + 1].contiguous(), qt_d[:8 * B].contiguous(), qv_d[:8 * B].contiguous()
        for mid, label in ((2, "filtered"), (2, "filtered_sparse_full_index"), (0, "unfiltered")):
            eng.sparse_compact_after(-1 if label == "filtered_sparse_full_index" else 0)
            p50, best = timeit(lambda: eng.hybrid_topk("c5", "c5s", qb, qpb, qtb, qvb, k, mask_id=mid), iters=10)
            fn = lambda: eng.hybrid_topk("c5", "c5s", qb, qpb, qtb, qvb, k, mask_id=mid)
            fn(); eng.profile_reset()
            for _ in range(5):
                fn()
            ms, calls, _ = eng.profile_read()
            rec[f"hybrid_B{B}_{label}"] = {"call_p50_ms": p50, "qps_per_gpu": B / (p50 * 1e-3),
                                           "dense_scan_ms": ms["dense_scan"] / 5, "sparse_scan_ms": ms["sparse_scan"] / 5}
    rec["hbm_floor_ms_B1_filtered"] = (n5 / 8 + 0.10 * n5 * (d5 * 2 + 4)) / (peaks["hbm_gbs"] * 1e9) * 1e3
    rec["tensor_floor_ms_B4096_filtered"] = 2.0 * 0.10 * n5 * d5 * 4096 / (peaks["bf16_tflops_sustained"] * 1e12) * 1e3
    out["C5_shard_12.5Mx384_f16_mask10pct"] = rec
    eng.dense_drop("c5")
    eng.sparse_drop("c5s")
    torch.cuda.empty_cache()
    if "--c5-only" in sys.argv:
        print(json.dumps(out, indent=1))
        sys.exit(0)

# ---- the three full scans of the reference's tree at 1 M points (qdrant_handler.py:305-372): P1 = matryoshka_64
#      (float32, kind::tf32), Q1 = `quantized` (int8 rows, kind::i8; and the float32 space it replaces), S1 = C3 below
def leaf_scan(label, name, nbytes, flop_per_query, qs, tensor_scale=1.0):
    for B in (1, 1024):
        qb = qs[:B].contiguous()
        p50, best = timeit(lambda: eng.dense_topk(name, qb, k))
        scan = stage_ms(lambda: eng.dense_topk(name, qb, k), "dense_scan")
        rec = {"call_p50_ms": p50, "scan_kernel_ms": scan, "path": eng.dense_path(name, B), "bytes": nbytes}
        if B == 1:
            rec.update(bound="hbm", achieved_gbs=nbytes / (scan * 1e-3) / 1e9, frac=nbytes / (scan * 1e-3) / 1e9 / peaks["hbm_gbs"])
        else:
            ach = flop_per_query * B / (scan * 1e-3) / 1e12
            rec.update(bound="tensor", achieved_tops=ach, peak_note=f"measured sustained bf16 x {tensor_scale} (nominal ratio of this MMA kind)",
                       frac_sustained=ach / (peaks["bf16_tflops_sustained"] * tensor_scale),
                       hbm_gbs=nbytes / (scan * 1e-3) / 1e9)
        out[f"{label}_B{B}"] = rec


n1 = 1 << 20
x = synth.dense_corpus_torch(n1, 768, torch.float32, dev)
q = synth.dense_queries_torch(x, 1024)
xm = x[:, :64].contiguous()
xm = xm / xm.norm(dim=1, keepdim=True)
eng.dense_set("p1", xm)
qm = q[:, :64].contiguous()
leaf_scan("P1_matryoshka64_1Mx64_f32_tf32", "p1", n1 * 64 * 4, 2.0 * n1 * 64, qm / qm.norm(dim=1, keepdim=True), tensor_scale=0.5)
eng.dense_drop("p1")
del xm
xi = (x * 127).to(torch.int8)                       # truncation toward zero, like numpy's astype(int8)
qi = (q * 127).to(torch.int8).to(torch.float32)
eng.dense_set("q1", xi)
leaf_scan("Q1_quantized_1Mx768_int8_i8", "q1", n1 * 768, 2.0 * n1 * 768, qi, tensor_scale=2.0)
eng.dense_drop("q1")
xf = xi.to(torch.float32)
xf = xf / xf.norm(dim=1, keepdim=True)
del xi
qf = (qi / qi.norm(dim=1, keepdim=True)).contiguous()
eng.dense_set("q1f", xf)
leaf_scan("Q1_quantized_1Mx768_as_float32_tf32", "q1f", n1 * 768 * 4, 2.0 * n1 * 768, qf, tensor_scale=0.5)
eng.force_dense_path("gemv")
for B in (1, 8):
    qb = qf[:B].contiguous()
    scan = stage_ms(lambda: eng.dense_topk("q1f", qb, k), "dense_scan")
    out[f"Q1_quantized_1Mx768_as_float32_gemv_B{B}"] = {"scan_kernel_ms": scan, "achieved_gbs": n1 * 768 * 4 / (scan * 1e-3) / 1e9,
                                                       "frac": n1 * 768 * 4 / (scan * 1e-3) / 1e9 / peaks["hbm_gbs"]}
eng.force_dense_path(None)
eng.dense_drop("q1f")
del x, xf, q, qi, qf
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

# ---- f2: 8192 new documents appended to the C3 index in HBM vs re-uploading the grown CSR from host memory
import time
n_new = 8192
d_ptr, d_doc, d_w, _ = synth.zipf_corpus(n_new, V, 285, seed=777)
d_doc = (d_doc.astype(np.int64) + nd).astype(np.uint32)
h_tp, h_doc, h_w = tp.cpu().numpy(), doc.cpu().numpy().astype(np.uint32), w.cpu().numpy()
torch.cuda.synchronize()
t0 = time.perf_counter()
eng.sparse_append("c3", d_ptr, d_doc, d_w, np.arange(V, dtype=np.int32), nd + n_new)
torch.cuda.synchronize()
t_app = (time.perf_counter() - t0) * 1e3
after = eng.sparse_topk("c3", qp_d, qt_d, qv_d, k)
# the same grown corpus built on the host (what a commit did before brag_sparse_append) and uploaded
t0 = time.perf_counter()
term_h = np.repeat(np.arange(V), np.diff(h_tp))
term_d = np.repeat(np.arange(V), np.diff(d_ptr))
o = np.argsort(np.concatenate([term_h, term_d]), kind="stable")
g_doc, g_w = np.concatenate([h_doc, d_doc])[o], np.concatenate([h_w, d_w])[o]
g_ptr = h_tp + d_ptr
t_host = (time.perf_counter() - t0) * 1e3
t0 = time.perf_counter()
eng.sparse_set_csr("c3f", g_ptr, g_doc, g_w, nd + n_new)
torch.cuda.synchronize()
t_up = (time.perf_counter() - t0) * 1e3
fresh = eng.sparse_topk("c3f", qp_d, qt_d, qv_d, k)
same = all(bool((a == b).all().item()) for a, b in zip(after, fresh))
out["C3_append_8192_docs"] = {"sparse_append_ms": t_app, "host_merge_ms": t_host, "fresh_upload_ms": t_up,
                              "delta_postings": int(d_ptr[-1]), "answers_identical": same}
print(json.dumps(out, indent=1))
