#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "compacted or sparse" > gpurun_out/s29_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/s29_tests.log
tail -3 gpurun_out/s29_tests.log
timeout 900 python tools/bench_configs.py --c5 --c5-only > gpurun_out/s29_c5.json 2>gpurun_out/s29_c5.err
echo "c5 rc=$?"; grep -n "build_ms\|sparse_scan_ms\|call_p50" gpurun_out/s29_c5.json
