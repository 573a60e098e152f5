#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "dense or hybrid or graph or logical" > gpurun_out/s38_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/s38_tests.log
tail -4 gpurun_out/s38_tests.log
timeout 600 python tools/dense_probe.py --n 12500000 --d 384 --batch 128,1024,4096 --dtype f16 2>/dev/null | tee gpurun_out/s38_dense_d384.json
timeout 600 python tools/dense_probe.py --n 1048576 --d 768 --batch 1,1024 --dtype f16 2>/dev/null | tee gpurun_out/s38_dense_d768.json
timeout 600 python tools/dense_probe.py --n 10000000 --d 1024 --batch 1,1024 --dtype bf16 2>/dev/null | tee gpurun_out/s38_dense_d1024.json
