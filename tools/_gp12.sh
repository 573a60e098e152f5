#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/dense_probe.py --n 12500000 --d 384 --batch 128,256,512,1024,2048,4096 --dtype f16 2>/dev/null | tee gpurun_out/s36_dense_d384.json
timeout 600 python tools/dense_probe.py --n 4000000 --d 1024 --batch 1024,4096 --dtype bf16 2>/dev/null | tee gpurun_out/s36_dense_d1024.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'um_|dense_umma|selk|round_q|rescore|finalize' --csv --log-file gpurun_out/s36_d384_launches.csv python tools/dense_probe.py --n 12500000 --d 384 --batch 4096 --dtype f16 --iters 1 > /dev/null 2>&1
echo "ncu rc=$?"
