#!/usr/bin/env python
"""Turn ncu output into the committed profile artefacts.
  launches:  python tools/ncu_summarize.py launches <launches.csv> <tag>
             (from `ncu --metrics gpu__time_duration.sum --clock-control none -k regex:... --csv --log-file launches.csv ...`)
             -> profiles/r2_launches_c4_<tag>.csv, profiles/r2_launch_shares_c4_<tag>.csv
  full:      python tools/ncu_summarize.py full <rep.ncu-rep> <tag>
             (from `ncu --set full --clock-control none --import-source on ...`)
             -> profiles/r2_ncu_summary_c4_<tag>.txt, profiles/ncu_traffic.json"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "profiles")
RND = os.environ.get("BRAG_ROUND", "r2")          # prefix of the artefacts written


def to_ms(v, u):
    v = float(v.replace(",", ""))
    return v / 1e6 if u in ("ns", "nsecond") else (v / 1e3 if u in ("us", "usecond") else (v * 1e3 if u in ("s", "second") else v))


def launches(path, tag):
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h = rows[hi]
    c = {n: i for i, n in enumerate(h)}
    agg, seq = collections.OrderedDict(), []
    for r in rows[hi + 1:]:
        if len(r) < len(h) or r[c["Metric Name"]] != "gpu__time_duration.sum":
            continue
        k = r[c["Kernel Name"]].split("(")[0].replace("void ", "")
        g = r[c["Grid Size"]]
        b = r[c["Block Size"]]
        ms = to_ms(r[c["Metric Value"]], r[c["Metric Unit"]])
        seq.append((r[c["ID"]], k, g, b, ms))
    # a kernel launched with two grid shapes per step (seed + main launch) is listed per shape
    shapes = collections.defaultdict(set)
    for _, k, g, _, _ in seq:
        shapes[k].add(g)
    longest = collections.defaultdict(float)
    for _, k, _, _, ms in seq:
        longest[k] = max(longest[k], ms)
    for _, k, g, _, ms in seq:
        key = k if len(shapes[k]) == 1 else f"{k} grid={g}"
        if k.split("<")[0] in ("dense_umma_kernel", "bm25x_scan_kernel") and ms < 0.1 * longest[k]:
            key = k + " [seed launch]"
        agg.setdefault(key, []).append(ms)
    tot = sum(sum(v) for v in agg.values())
    with open(os.path.join(PROF, f"{RND}_launch_shares_c4_{tag}.csv"), "w") as f:
        f.write("kernel,launches,total_ms,avg_ms,share_pct  (ncu --metrics gpu__time_duration.sum --clock-control none; "
                "bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --b1-iters 0; includes the one-time index "
                "build kernels; cold-cache "
                "serialised: compare shares)\n")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            line = f"\"{k}\",{len(v)},{sum(v):.3f},{sum(v)/len(v):.4f},{100*sum(v)/tot:.2f}"
            f.write(line + "\n")
            print(line)
    with open(os.path.join(PROF, f"{RND}_launches_c4_{tag}.csv"), "w") as f:
        f.write("id,kernel,grid,block,ms\n")
        for s in seq:
            f.write(f"{s[0]},{s[1]},\"{s[2]}\",\"{s[3]}\",{s[4]:.4f}\n")


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct"]


def full(rep, tag):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h = rows[0]
    ci = {n: i for i, n in enumerate(h)}
    out = ["ncu --set full --clock-control none --import-source on, bench.py --steps 1 --warmup 1 (C4: 10M x 1024 bf16, "
           "batch 1024), launches of the timed step",
           "(times under ncu are cold-cache/serialised and taken at whatever clock the replay ran at; bench.py reports "
           "the live numbers)", ""]
    sc = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}
    traffic, best = {}, {}
    for r in rows[2:]:
        name = r[ci["Kernel Name"]].split("(")[0].replace("void ", "")
        out.append(f"== {name}  grid {r[ci['launch__grid_size']]}")
        for w in WANT:
            if w in ci:
                out.append(f"  {w:85s} {r[ci[w]]:>18s} {rows[1][ci[w]]}")
        t = (float(r[ci["dram__bytes_read.sum"]].replace(",", "")) * sc[rows[1][ci["dram__bytes_read.sum"]]] +
             float(r[ci["dram__bytes_write.sum"]].replace(",", "")) * sc[rows[1][ci["dram__bytes_write.sum"]]])
        out.append(f"  traffic (dram read+write per launch) = {t/1e9:.2f} GB")
        out.append("")
        key = name.split("<")[0]
        dur = to_ms(r[ci["gpu__time_duration.sum"]], rows[1][ci["gpu__time_duration.sum"]])
        if dur > best.get(key, 0):          # the main launch of a kernel that also has a short seed launch
            best[key], traffic[key] = dur, t
    open(os.path.join(PROF, f"{RND}_ncu_summary_c4_{tag}.txt"), "w").write("\n".join(out))
    json.dump(traffic, open(os.path.join(PROF, "ncu_traffic.json"), "w"), indent=1)
    print("\n".join(out))
    print(traffic)


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
