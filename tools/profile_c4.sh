python bench.py > gpurun_out/s26_bench.json 2> gpurun_out/s26_bench.err; echo bench_rc=$?
KF='regex:bm25|dense_|um_bound|um_seed|merge_|rescore_|finalize_|rrf_|round_queries|remap_|gather_rows|mask_'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" --csv --log-file gpurun_out/s26_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --b1-iters 0 > gpurun_out/s26_ncu_l.log 2>&1; echo launches_rc=$?
ncu --set full --clock-control none --import-source on -k 'regex:bm25x_scan|dense_umma_kernel' --launch-skip 4 --launch-count 4 -o gpurun_out/s26_full -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --b1-iters 0 > gpurun_out/s26_ncu_f.log 2>&1; echo full_rc=$?
python tools/bench_configs.py --c5 > gpurun_out/s26_configs.json 2> gpurun_out/s26_configs.err; echo cfg_rc=$?
