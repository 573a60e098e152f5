#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_rerank.py tests/test_graph_index.py -m gpu -x -q > gpurun_out/s31_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/s31_tests.log
tail -15 gpurun_out/s31_tests.log
timeout 900 python tools/bench_configs.py > gpurun_out/s31_configs.json 2> gpurun_out/s31_configs.err
echo "configs rc=$?"; grep -n "C3_append" -A7 gpurun_out/s31_configs.json; tail -3 gpurun_out/s31_configs.err
