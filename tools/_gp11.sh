#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/s35_bench.json 2> gpurun_out/s35_bench.err
echo "bench rc=$?"; tail -c 3000 gpurun_out/s35_bench.json; tail -5 gpurun_out/s35_bench.err
