#!/bin/bash
# sparse filter compaction: parity + C5 shard numbers
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sparse or hybrid or filter" > gpurun_out/s27_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/s27_tests.log
tail -5 gpurun_out/s27_tests.log
timeout 900 python tools/bench_configs.py --c5 > gpurun_out/s27_c5.log 2>&1
echo "c5 rc=$?"; tail -40 gpurun_out/s27_c5.log
