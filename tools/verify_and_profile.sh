#!/bin/bash
# final single-GPU verification + profile artefacts (tag v26)
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s40_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/s40_tests.log
tail -3 gpurun_out/s40_tests.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/s40_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s40_smoke.log
timeout 900 python bench.py > gpurun_out/s40_bench.json 2> gpurun_out/s40_bench.err; echo bench_rc=$?
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s40_bench_ref.json 2> gpurun_out/s40_bench_ref.err; echo ref_rc=$?
KF='regex:bm25|dense_|um_bound|um_seed|merge_|rescore_|finalize_|rrf_|round_queries|remap_|gather_rows|mask_|publish_'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" --csv --log-file gpurun_out/s40_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --b1-iters 0 > gpurun_out/s40_ncu_l.log 2>&1; echo launches_rc=$?
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:bm25x_scan|dense_umma_kernel' --launch-skip 4 --launch-count 4 -o gpurun_out/s40_full -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --b1-iters 0 > gpurun_out/s40_ncu_f.log 2>&1; echo full_rc=$?
timeout 900 python tools/bench_configs.py --c5 > gpurun_out/s40_configs.json 2> gpurun_out/s40_configs.err; echo cfg_rc=$?
