#!/usr/bin/env python
"""Headline benchmark: hybrid (dense + BM25 + RRF) top-100 QPS at 10M x 1024 bf16, batch 1024
(BASELINE.json metric / configs[3]), row-sharded over --gpus B200s of one box.

One "step" = one batch of B synthetic hybrid queries through the whole hot path.
  value  : QPS with the query batch already resident in HBM (device-pointer C-ABI calls)
  e2e    : QPS through the host-buffer C-ABI call (H2D of the batch + D2H of the results inside)
  parity : the timed step's answers on sampled queries against tools/torch_checker.py (torch on the device, no code
           shared with libbrag.so), outside the timed region; at N > 1 also "every rank holds the same lists"
  --workload c5 : configs[4] (100 M x 384 f16 + 10 % payload filter, batch 4096) sharded over N >= 4 GPUs
  --impl reference : the CPU restatement (oracle/fast_cpu.py) timed on the host cores.
Prints ONE JSON line on rank 0."""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "hybrid top-100 QPS @10Mx1024 batch 1024"
UNIT = "queries/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=["c4", "c5"],
                    help="c4 = BASELINE configs[3] (the metric's config, default); c5 = configs[4]: 100M x 384 f16 + 10%% filter")
    ap.add_argument("--n", "--rows", dest="n", type=int, default=None, help="corpus rows / documents (whole job); --rows under torchrun, whose parser takes --n for itself")
    ap.add_argument("--d", type=int, default=None)
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--dtype", default=None, choices=["bf16", "f16"])
    ap.add_argument("--selectivity", type=float, default=None, help="payload-filter pass rate (c5: 0.10; c4: none)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "collective"],
                    help="N > 1: lists pushed over NVLink peer memory by the tail kernels, or an NCCL all-gather")
    ap.add_argument("--parity-queries", type=int, default=8, help="sampled queries checked after the timed loop (0 = skip)")
    ap.add_argument("--vocab", type=int, default=50_000)
    ap.add_argument("--doc-len", type=int, default=285)
    ap.add_argument("--q-terms", type=int, default=8)
    ap.add_argument("--cpu-sample-rows", type=int, default=1 << 19, help="rows/docs of the CPU-baseline sample")
    ap.add_argument("--cpu-sample-queries", type=int, default=256)
    ap.add_argument("--b1-iters", type=int, default=30, help="batch-1 latency iterations (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    c5 = a.workload == "c5"
    a.n = a.n if a.n is not None else (100_000_000 if c5 else 10_000_000)
    a.d = a.d if a.d is not None else (384 if c5 else 1024)
    a.batch = a.batch if a.batch is not None else (4096 if c5 else 1024)
    a.dtype = a.dtype or ("f16" if c5 else "bf16")
    a.selectivity = a.selectivity if a.selectivity is not None else (0.10 if c5 else None)
    return a


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return {"tflops": j.get("bf16_tflops_sustained", 1400.0), "tflops_burst": j.get("bf16_tflops"),
                "hbm_gbs": j.get("hbm_gbs", 6650.0), "source": "measured"}
    return {"tflops": 1400.0, "tflops_burst": 1590.0, "hbm_gbs": 6650.0, "source": "fallback"}


def ncu_traffic(kernel, world):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed
    `ncu --set full` capture of this same command at N=1 (profiles/ncu_traffic.json); None otherwise."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if world != 1 or not os.path.exists(p):
        return None
    with open(p) as f:
        return json.load(f).get(kernel)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(len(r) >= 6 and r[2 + j] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------ reference arm
def run_reference(args, rank, world):
    """CPU restatement timed on the host cores: a bounded sample (rows and queries) of the same
    workload, scaled linearly in corpus size to the full config."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    # torchrun exports OMP_NUM_THREADS=1, which throttled this arm 3x at N > 1 in round 1: the CPU baseline always gets
    # every host core (BLAS reads these when it is first loaded, i.e. below)
    for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = str(cores)
    import torch
    from oracle import fast_cpu
    from rag_application_b200 import synth
    torch.set_num_threads(cores)
    # same sample as the in-run cpu_baseline leg of the GPU arm (CPU_SAMPLE_ROWS x CPU_SAMPLE_QUERIES)
    ns, bs = min(args.cpu_sample_rows, args.n), min(args.cpu_sample_queries, args.batch)
    x = synth.dense_corpus_torch(ns, args.d, torch.bfloat16 if args.dtype == "bf16" else torch.float16, "cpu")
    q = synth.dense_queries_torch(x, bs).numpy()
    xs = x.float().numpy()
    tp, doc, w = synth.zipf_corpus_torch(ns, args.vocab, args.doc_len, "cpu", block_docs=1 << 16)
    qp, qt, qv = synth.sparse_queries_torch(tp, bs, args.q_terms)
    tp, doc, w = tp.numpy(), doc.numpy().astype(np.uint32), w.numpy()
    qp, qt, qv = qp.numpy(), qt.numpy(), qv.numpy()

    def step():
        fast_cpu.hybrid_batch(xs, q, tp, doc, w, ns, qp, qt, qv, args.k, threads=cores)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    scale = ns / args.n
    val = bs * args.steps / dt * scale
    sample = (f"{bs} queries x {ns} rows/docs per step (1/{args.n // ns if ns else 1} of the corpus); "
              f"QPS scaled by rows ratio {scale:.5f} (cost linear in N)")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args, args.gpus),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def config_dict(args, world):
    flt = f", payload filter Bernoulli({args.selectivity}) seed 4242" if args.selectivity else ""
    return {"workload": f"{args.workload.upper()} hybrid dense+BM25+RRF top-{args.k}: {args.n} chunks x {args.d}-d {args.dtype}, "
                        f"V={args.vocab} Zipf(1.0) L={args.doc_len}, {args.q_terms}-term queries, batch {args.batch}{flt}"
                        + (" (strong scaling over N >= 4: the postings of 100 M documents do not fit fewer GPUs)"
                           if args.workload == "c5" else ""),
            "n": args.n, "d": args.d, "batch": args.batch, "k": args.k, "shards": world,
            "l2": "inputs larger than L2 (corpus shard >> 126 MB)", "parallelism": f"row-shard x{world}"}




# ------------------------------------------------------------------------------------ parity of the timed step
def parity_check(args, rank, world, dev, raw, queries, answer, lo, mask_t):
    """The step's answers on sampled queries against the torch checker (tools/torch_checker.py: float64 on the raw
    corpus tensors, no code shared with libbrag.so).  answer = (rrf scores, ids, counts, (stage scores, ids, counts))
    of one extra, untimed call with the stage lists exposed.  At N > 1 every rank ranks its own shard with the checker,
    the (score, id) lists are all-gathered and globally sorted on every rank -- a torch global sort of the gathered
    payload -- and the ranks' fused lists are compared with each other."""
    import torch
    import torch.distributed as dist
    from tools import torch_checker as T
    x, tp, doc, w = raw
    q_dev, qp, qt, qv = queries
    fs, fi, fc = [t.cpu().numpy() for t in answer[:3]]
    ts, ti, tc = [t.cpu().numpy() for t in answer[3]]
    k, n_loc = args.k, x.shape[0]
    rows = sorted(set(np.linspace(0, args.batch - 1, args.parity_queries).astype(int).tolist()))
    res = {"queries": len(rows), "checker": "tools/torch_checker.py (torch float64 on device)", "dense_ids_equal": 0,
           "dense_explained_by_near_ties": 0, "sparse_ids_equal": 0, "sparse_scores_bit_equal": 0, "fused_ids_equal": 0,
           "max_rel_score_err": 0.0}
    for b in rows:
        ds, di, all_s = T.dense_topk(x, q_dev[b], k, mask=mask_t, id_offset=lo)
        ss, si = T.sparse_topk(tp, doc, w, n_loc, qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]], k, mask=mask_t, id_offset=lo)
        if world > 1:
            def gather(s, i):
                pad_s = torch.full((k,), -float("inf"), device=dev)
                pad_i = torch.full((k,), -1, dtype=torch.int64, device=dev)
                pad_s[:len(s)] = torch.from_numpy(s).to(dev)
                pad_i[:len(i)] = torch.from_numpy(i).to(dev)
                gs = torch.empty(world * k, device=dev)
                gi = torch.empty(world * k, dtype=torch.int64, device=dev)
                dist.all_gather_into_tensor(gs, pad_s)
                dist.all_gather_into_tensor(gi, pad_i)
                gs, gi = gs.cpu().numpy(), gi.cpu().numpy()
                keep = gi >= 0
                o = np.lexsort((gi[keep], -gs[keep].astype(np.float64)))[:k]
                return gs[keep][o], gi[keep][o]
            ds, di = gather(ds, di)
            ss, si = gather(ss, si)
            all_s = None
        r = T.dense_lists_agree(ts[0][b], ti[0][b], ds, di, all_scores=all_s, id_offset=lo)
        if all_s is None and not r["ids_equal"]:
            # sharded: near-ties are judged on the gathered checker scores of the same ids
            ref = dict(zip(di.tolist(), ds.tolist()))
            r["explained"] = all(abs(ref.get(int(g), 9e9) - float(e)) <= T.TIE_EPS * max(1.0, abs(float(e))) or int(g) == int(h)
                                 for g, e, h in zip(ti[0][b][:len(di)], ds, di))
        res["dense_ids_equal"] += int(r["ids_equal"])
        res["dense_explained_by_near_ties"] += int(r["explained"] and not r["ids_equal"])
        res["max_rel_score_err"] = max(res["max_rel_score_err"], r["max_rel_err"])
        c = len(si)
        s_ok = int(tc[1][b]) == c and np.array_equal(ti[1][b][:c], si)
        res["sparse_ids_equal"] += int(s_ok)
        res["sparse_scores_bit_equal"] += int(s_ok and np.array_equal(ts[1][b][:c].view(np.uint32), ss.view(np.uint32)))
        # fusion: RRF of the checker's lists (of the engine's dense list where it differs by an explained near-tie)
        dl = di if r["ids_equal"] else ti[0][b][:int(tc[0][b])]
        es, ei = T.rrf([dl, si], k)
        res["fused_ids_equal"] += int(int(fc[b]) == len(ei) and np.array_equal(fi[b][:len(ei)], ei)
                                      and np.array_equal(fs[b][:len(ei)], es))
    res["ids_equal"] = bool(res["dense_ids_equal"] + res["dense_explained_by_near_ties"] == len(rows)
                            and res["sparse_ids_equal"] == len(rows) and res["fused_ids_equal"] == len(rows))
    if world > 1:
        # every rank holds the same fused lists
        h = torch.from_numpy(fi.astype(np.int64)).to(dev)
        hs = torch.stack([h.sum(), (h * torch.arange(1, h.numel() + 1, device=dev).view_as(h)).sum()])
        allh = torch.empty(world * 2, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allh, hs)
        res["ranks_agree"] = bool((allh.view(world, 2) == hs[None]).all().item())
        res["ids_equal"] = res["ids_equal"] and res["ranks_agree"]
    return res


# ------------------------------------------------------------------------------------ our arm
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from rag_application_b200 import Engine, synth
    from rag_application_b200.sharded import ShardedHybrid, shard_bounds

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ.pop("NCCL_DEBUG")        # its banner goes to stdout, which carries the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()
    tdt = torch.bfloat16 if args.dtype == "bf16" else torch.float16
    lo, hi = shard_bounds(args.n, world, rank)
    n_loc = hi - lo
    k = args.k
    # raw corpus tensors are kept for the parity check when they fit beside the engine's copies
    est_gb = (2 * n_loc * args.d * 2 + 2.4 * n_loc * 256 * 8) / 1e9
    keep_raw = args.parity_queries > 0 and est_gb < 125

    # ---- corpus shard, built in HBM
    eng = Engine(local)
    x = synth.dense_corpus_torch(n_loc, args.d, tdt, dev, seed=synth.SEED_CORPUS + rank)
    q_dev = synth.dense_queries_torch(x, args.batch)
    if world > 1:
        dist.broadcast(q_dev, 0)
    cpu_rows = min(args.cpu_sample_rows, n_loc)
    cpu_x = x[:cpu_rows].float().cpu().numpy() if rank == 0 and world == 1 and not args.no_cpu_baseline else None
    eng.dense_set("dense", x)
    if not keep_raw:
        del x
        x = None
    torch.cuda.empty_cache()
    tp, doc, w = synth.zipf_corpus_torch(n_loc, args.vocab, args.doc_len, dev, seed=synth.SEED_CORPUS + rank)
    df = (tp[1:] - tp[:-1]).clone()
    if world > 1:
        dist.all_reduce(df)
    gptr = torch.zeros_like(tp)
    gptr[1:] = torch.cumsum(df, 0)
    qp, qt, qv = synth.sparse_queries_torch(gptr, args.batch, args.q_terms)
    t0 = time.perf_counter()
    eng.sparse_set_csr("sparse", tp, doc, w, n_loc)
    torch.cuda.synchronize()
    sparse_index_s = time.perf_counter() - t0
    # algorithmic BM25 bytes of this batch on this shard: sum_t df_t * 8 B
    ldf = (tp[1:] - tp[:-1])
    bm25_bytes = float(ldf[qt.to(dev).long()].sum().item()) * 8.0
    cpu_csr = None
    if cpu_x is not None:
        keep = doc < cpu_rows
        term_of = torch.repeat_interleave(torch.arange(args.vocab, device=dev), ldf)
        sp = torch.zeros(args.vocab + 1, dtype=torch.int64, device=dev)
        sp[1:] = torch.cumsum(torch.bincount(term_of[keep], minlength=args.vocab), 0)
        cpu_csr = (sp.cpu().numpy(), doc[keep].cpu().numpy().astype(np.uint32), w[keep].cpu().numpy(), cpu_rows)
        del term_of, keep
    if not keep_raw:
        del tp, doc, w
        tp = doc = w = None
    torch.cuda.empty_cache()

    # ---- payload filter (c5): Bernoulli(selectivity), seed 4242 + rank, the same bits for both spaces
    mask_id, mask_t, sel_real = 0, None, None
    if args.selectivity:
        g = torch.Generator(device=dev)
        g.manual_seed(synth.SEED_MASK + rank)
        mask_t = torch.rand(n_loc, generator=g, device=dev) < args.selectivity
        eng.mask_set(1, mask_t.cpu().numpy())
        mask_id = 1
        sel_real = float(mask_t.float().mean().item())
        eng.sparse_compact_after(0)            # the filter is fixed for the whole run: build its index on first use

    max_batch = max(args.batch, 1)
    sh = ShardedHybrid(eng, rank, world, lo, exchange=args.exchange if world > 1 else "collective",
                       max_batch=max_batch, max_k=max(k, 16))
    qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
    mt = args.q_terms

    def step_dev():
        if world == 1:
            return eng.hybrid_topk("dense", "sparse", q_dev, qp_d, qt_d, qv_d, k, mask_id=mask_id, max_terms=mt)
        return sh.hybrid_topk("dense", "sparse", q_dev, qp_d, qt_d, qv_d, k, mask_id=mask_id, max_terms=mt)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    # the first call under a filter builds the filter-compacted dense copy and sparse index (c5): timed apart
    first_call_ms = None
    if mask_id:
        first_call_ms, _ = timed(step_dev, 1)
    eng.profile_enable(True)
    for _ in range(args.warmup):
        step_dev()
    barrier()
    eng.profile_reset()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms_total, last = timed(step_dev, args.steps)
    stage_ms, stage_calls, launches = eng.profile_read()
    clk = clocks.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    value = args.batch / (ms_step * 1e-3)
    last = [t.clone() for t in last[:3]]

    # ---- parity of the timed step's answers (outside the timed region)
    parity = None
    if args.parity_queries > 0:
        if not keep_raw:
            parity = {"skipped": f"raw corpus tensors not kept beside the engine's copies (~{est_gb:.0f} GB per GPU)"}
        else:
            eng.profile_enable(False)
            if world == 1:
                keys = eng.shard_local_lists("dense", "sparse", q_dev, qp_d, qt_d, qv_d, k, mask_id=mask_id, max_terms=mt)
                ans = eng.shard_fuse(keys[None].contiguous(), k, stages=True)
            else:
                ans = sh.hybrid_topk("dense", "sparse", q_dev, qp_d, qt_d, qv_d, k, mask_id=mask_id, max_terms=mt, stages=True)
            same = all(bool(torch.equal(a, b)) for a, b in zip(last, ans[:3]))
            parity = parity_check(args, rank, world, dev, (x, tp, doc, w), (q_dev, qp.numpy(), qt.numpy(), qv.numpy()),
                                  ans, lo, mask_t)
            parity["staged_call_equals_timed_call"] = same
            if world > 1 and args.exchange == "peer":
                # the peer exchange against the collective one (NCCL all-gather of the same keys)
                col = sh.hybrid_topk("dense", "sparse", q_dev, qp_d, qt_d, qv_d, k, mask_id=mask_id, max_terms=mt,
                                     exchange="collective")
                parity["peer_equals_collective"] = all(bool(torch.equal(a, b)) for a, b in zip(last, col[:3]))
                parity["ids_equal"] = parity["ids_equal"] and parity["peer_equals_collective"]
            parity["ids_equal"] = bool(parity["ids_equal"] and same)
            eng.profile_enable(True)
    x = tp = doc = w = None
    torch.cuda.empty_cache()

    # ---- end to end: host buffers through the public call, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        q_host = torch.empty((args.batch, args.d), dtype=torch.float32).pin_memory()
        q_host.copy_(q_dev)
        qp_h, qt_h, qv_h = qp.pin_memory(), qt.pin_memory(), qv.pin_memory()
        h2d = q_host.numel() * 4 + qp_h.numel() * 4 + qt_h.numel() * 4 + qv_h.numel() * 8
        d2h = args.batch * k * (8 + 8) + args.batch * 4
        qn, qpn, qtn, qvn = q_host.numpy(), qp_h.numpy(), qt_h.numpy(), qv_h.numpy()
        # result buffers of a serving loop: pinned, reused by every call
        out_pin = (torch.empty((args.batch, k), dtype=torch.float64).pin_memory().numpy(),
                   torch.empty((args.batch, k), dtype=torch.int64).pin_memory().numpy(),
                   torch.empty((args.batch,), dtype=torch.int32).pin_memory().numpy())
        if world == 1:
            def step_host():
                return eng.hybrid_topk("dense", "sparse", qn, qpn, qtn, qvn, k, mask_id=mask_id, max_terms=mt, out=out_pin)
        elif args.exchange == "peer":
            def step_host():        # every rank: H2D of the batch, sharded call, D2H of the (identical) result
                return sh.hybrid_topk("dense", "sparse", qn, qpn, qtn, qvn, k, mask_id=mask_id, max_terms=mt, out=out_pin)
        else:
            def step_host():
                s, i, c = sh.hybrid_topk("dense", "sparse", q_host.to(dev, non_blocking=True),
                                         qp_h.to(dev, non_blocking=True), qt_h.to(dev, non_blocking=True),
                                         qv_h.to(dev, non_blocking=True), k, mask_id=mask_id, max_terms=mt)
                return s.cpu(), i.cpu(), c.cpu()
        for _ in range(max(1, args.warmup)):
            step_host()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            out_h = step_host()
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": args.batch * args.steps / float(dt.item()), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "copies": "per rank (every rank uploads the batch and downloads the result)",
               "equals_device_call": bool(np.array_equal(np.asarray(out_h[1]), last[1].cpu().numpy()))}

    # ---- batch-1 latency (the metric's second half): p50 over single-query hybrid calls
    b1 = None
    if args.b1_iters > 0:
        q1, qp1 = q_dev[:1].contiguous(), qp_d[:2].contiguous()
        nt1 = int(qp[1])
        qt1, qv1 = qt_d[:nt1].contiguous(), qv_d[:nt1].contiguous()

        def step_b1():
            if world == 1:
                return eng.hybrid_topk("dense", "sparse", q1, qp1, qt1, qv1, k, mask_id=mask_id, max_terms=mt)
            return sh.hybrid_topk("dense", "sparse", q1, qp1, qt1, qv1, k, mask_id=mask_id, max_terms=mt)
        for _ in range(3):
            step_b1()
        lat = []
        for _ in range(args.b1_iters):
            ms1, _ = timed(step_b1, 1)
            lat.append(ms1)
        rows_scanned = n_loc * (sel_real if sel_real else 1.0)
        floor_bytes = rows_scanned * args.d * 2 + (n_loc / 8 if sel_real else 0)
        b1 = {"p50_ms": float(np.median(lat)), "p95_ms": float(np.percentile(lat, 95)), "iters": args.b1_iters,
              "hbm_floor_ms": floor_bytes / (peaks["hbm_gbs"] * 1e9) * 1e3,
              "hbm_floor": "dense rows that pass the filter x d x 2 B (+ N/8 mask bytes) at the measured copy bandwidth"}
        if world == 1 or args.exchange == "peer":
            # the same call captured once in a CUDA graph and replayed (stream-ordered and allocation-free after warm-up;
            # the peer exchange has no collective inside, so the sharded call captures too): what a serving loop with
            # fixed query buffers pays per query, launch overhead removed
            try:
                eng.profile_enable(False)        # stage events cannot be read back from a captured call
                eager = [t.clone() for t in step_b1()]
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    for _ in range(3):
                        step_b1()
                torch.cuda.current_stream().wait_stream(side)
                barrier()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    captured = step_b1()
                graph.replay()
                torch.cuda.synchronize()
                same = all(bool(torch.equal(a, b)) for a, b in zip(eager, captured))
                glat = []
                for _ in range(args.b1_iters):
                    ms1, _ = timed(graph.replay, 1)
                    glat.append(ms1)
                b1["graph_replay_p50_ms"] = float(np.median(glat))
                b1["graph_replay_equals_eager"] = same
                barrier()
                del graph
            except Exception as ex:      # noqa: BLE001 -- an extra; the eager numbers above stand
                b1["graph_replay_p50_ms"] = None
                b1["graph_replay_error"] = str(ex)[:200]

    # ---- c5: the same batch without the filter, for the filtered / unfiltered comparison
    unfiltered = None
    if mask_id:
        def step_nomask():
            if world == 1:
                return eng.hybrid_topk("dense", "sparse", q_dev, qp_d, qt_d, qv_d, k, max_terms=mt)
            return sh.hybrid_topk("dense", "sparse", q_dev, qp_d, qt_d, qv_d, k, max_terms=mt)
        step_nomask()
        ms_u, _ = timed(step_nomask, max(2, args.steps // 4))
        ms_u /= max(2, args.steps // 4)
        unfiltered = {"ms_per_step": ms_u, "value": args.batch / (ms_u * 1e-3), "unit": UNIT}

    # ---- CPU baseline (rank 0, N=1): bounded sample of the same workload on the host cores
    cpu = None
    if rank == 0 and cpu_csr is not None:
        from oracle import fast_cpu
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        bs = min(args.cpu_sample_queries, args.batch)
        sp, sdoc, sw, ns = cpu_csr
        qs = q_dev[:bs].cpu().numpy()
        qpn, qtn, qvn = qp[: bs + 1].numpy(), qt[: int(qp[bs])].numpy(), qv[: int(qp[bs])].numpy()
        fast_cpu.hybrid_batch(cpu_x[:ns], qs[:8], sp, sdoc, sw, ns, qpn[:9], qtn, qvn, k, threads=cores)  # warm
        t0 = time.perf_counter()
        reps = 0
        while reps < 3 or time.perf_counter() - t0 < 10.0:
            fast_cpu.hybrid_batch(cpu_x[:ns], qs, sp, sdoc, sw, ns, qpn, qtn, qvn, k, threads=cores)
            reps += 1
            if reps >= 40:
                break
        dt = (time.perf_counter() - t0) / reps
        cpu = {"value": bs / dt * (ns / args.n), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{bs} queries x {ns} rows/docs, {reps} passes ({dt * reps:.1f} s CPU wall), QPS scaled by "
                         f"{ns}/{args.n} (cost linear in N); the --impl reference arm times the same sample"}

    if rank == 0:
        # the main launch runs once per group of <= 1024 queries (4 at C5's batch 4096): per-launch figures
        groups = (args.batch + 1023) // 1024
        dense_ms = stage_ms["dense_scan"] / args.steps / groups
        n_scan = n_loc * (sel_real if sel_real else 1.0)          # rows the tensor-core scan really multiplies
        flops = 2.0 * n_scan * args.d * args.batch / groups
        ach = flops / (dense_ms * 1e-3) / 1e12 if dense_ms > 0 else 0.0
        bm_ms = stage_ms["sparse_scan"] / max(1, stage_calls["sparse_scan"])
        bm_bytes = bm25_bytes * (sel_real if sel_real else 1.0)   # postings of the documents that pass the filter
        line = {
            "metric": METRIC if args.workload == "c4" else "hybrid top-100 QPS @100Mx384 batch 4096, 10% payload filter (BASELINE configs[4])", "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic", "config": config_dict(args, world),
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clk, "parity": parity,
            "roofline": {"kernel": "dense_umma_kernel (tcgen05 scan + fused top-k)", "bound": "tensor", "achieved": ach,
                         "peak": peaks["tflops"], "unit": "TFLOP/s", "frac": ach / peaks["tflops"],
                         "peak_kind": f"{peaks['source']} sustained cuBLAS bf16", "traffic": ncu_traffic("dense_umma_kernel", world),
                         "flops_per_launch": flops, "avg_launch_ms": dense_ms,
                         # frac can exceed 1: `peak` is what cuBLAS sustains under the box's power cap, and this scan
                         # moves less HBM data per flop than a GEMM that writes its output; the other two denominators:
                         "frac_of_measured_burst": (ach / peaks["tflops_burst"]) if peaks.get("tflops_burst") else None,
                         "frac_of_nominal_2250": ach / 2250.0},
            "cpu_baseline": cpu, "batch1": b1,
            "stages_ms_per_step": {s: stage_ms[s] / args.steps for s in stage_ms},
            # N > 1: signal + fuse kernels of the sharded call, i.e. the exchange as this rank sees it -- it includes
            # the wait for the slowest peer's lists
            "exchange_ms": (stage_ms["merge"] / args.steps) if world > 1 else 0.0,
            "exchange": (args.exchange if world > 1 else None),
            # what one rank stores into its peers' windows per step: both ranked lists (B x k keys of 8 B) to each of the
            # world - 1 peers, plus one 8-byte flag each; the cross-shard bound words (one atomicMax per improved bound
            # and peer, at most a few per query and list) are not counted
            "nvlink_bytes_per_step_per_rank": (int((world - 1) * (2 * args.batch * k * 8 + 8))
                                               if world > 1 and args.exchange == "peer" else 0),
            "sparse_index_build_s": sparse_index_s,
            # second kernel of the step: the sparse scan against the HBM roofline, on the ALGORITHMIC bytes an
            # exhaustive posting scan reads (sum_t df_t * 8 B); `traffic` = DRAM bytes the pruning kernel really moves
            "bm25": {"kernel": "bm25x_scan_kernel (bitmap-filter BM25 scan + fused top-k)", "bound": "hbm",
                     "algorithmic_bytes_per_step": bm_bytes, "ms": bm_ms,
                     "achieved_gbs": bm_bytes / (bm_ms * 1e-3) / 1e9 if bm_ms > 0 else None,
                     "peak_gbs": peaks["hbm_gbs"],
                     "frac": bm_bytes / (bm_ms * 1e-3) / 1e9 / peaks["hbm_gbs"] if bm_ms > 0 else None,
                     "traffic": ncu_traffic("bm25x_scan_kernel", world)},
        }
        if mask_id:
            line["filter"] = {"selectivity": sel_real, "first_call_ms_incl_compacted_copy_and_index_build": first_call_ms,
                              "unfiltered": unfiltered}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
