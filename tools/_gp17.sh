#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python tools/bench_configs.py > gpurun_out/s41_configs.json 2> gpurun_out/s41_configs.err; echo cfg_rc=$?
grep -n "C3_append" -A7 gpurun_out/s41_configs.json
