#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s30_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/s30_tests.log
tail -15 gpurun_out/s30_tests.log
