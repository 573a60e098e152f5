timeout 600 python -m pytest tests -m gpu -x -q -k "dense or config or hybrid" 2>&1 | tail -4
P=$PWD/rag_application_b200/csrc/libbrag_prev.so
S="--no-cpu-baseline --no-e2e --b1-iters 0 --parity-queries 2"
show() { python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(j['ms_per_step'],3), {k:round(v,3) for k,v in j['stages_ms_per_step'].items()}, j['parity']['ids_equal'], j['clocks'].get('sm_mhz'))"; }
for r in 1 2; do
echo "== shard old"; BRAG_LIB=$P timeout 200 python bench.py --n 1250000 $S 2>/dev/null | show
echo "== shard new"; timeout 200 python bench.py --n 1250000 $S 2>/dev/null | show
done
echo "== c4 old"; BRAG_LIB=$P timeout 300 python bench.py $S 2>/dev/null | show
echo "== c4 new"; timeout 300 python bench.py $S 2>/dev/null | show
echo "== c4 old"; BRAG_LIB=$P timeout 300 python bench.py $S 2>/dev/null | show
echo "== c4 new"; timeout 300 python bench.py $S 2>/dev/null | show
echo "== configs old"; BRAG_LIB=$P timeout 500 python tools/bench_configs.py 2>/dev/null > gpurun_out/r2_t11_configs_old.json
echo "== configs new"; timeout 500 python tools/bench_configs.py 2>/dev/null > gpurun_out/r2_t11_configs_new.json
python - <<'PY'
import json
a=json.load(open("gpurun_out/r2_t11_configs_old.json")); b=json.load(open("gpurun_out/r2_t11_configs_new.json"))
for k in b:
    if isinstance(b[k],dict) and "scan_kernel_ms" in b[k] and k in a:
        print(k, round(a[k]["scan_kernel_ms"],4), "->", round(b[k]["scan_kernel_ms"],4), "call", round(a[k].get("call_p50_ms",0),4), "->", round(b[k].get("call_p50_ms",0),4))
PY
