#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "sparse or hybrid or client or bm25 or logical" > gpurun_out/s33_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/s33_tests.log
tail -5 gpurun_out/s33_tests.log
timeout 600 python tools/bm25_probe.py --iters 10 > gpurun_out/s33_probe_c4.json 2> gpurun_out/s33_probe.err
echo "probe rc=$?"; cat gpurun_out/s33_probe_c4.json; tail -3 gpurun_out/s33_probe.err
