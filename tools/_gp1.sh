python -m pytest tests -m gpu -x -q > gpurun_out/s15_tests.log 2>&1; echo tests_rc=$? >> gpurun_out/s15_tests.log; tail -3 gpurun_out/s15_tests.log
python bench.py > gpurun_out/s15_bench.json 2> gpurun_out/s15_bench.err; echo bench_rc=$?; cat gpurun_out/s15_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s15_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --b1-iters 0 > gpurun_out/s15_ncu_l.log 2>&1; echo launches_rc=$?
ncu --set full --clock-control none --import-source on -k regex:'bm25x_scan|dense_umma_kernel' --launch-skip 2 --launch-count 2 -o gpurun_out/s15_full -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --b1-iters 0 > gpurun_out/s15_ncu_f.log 2>&1; echo full_rc=$?
