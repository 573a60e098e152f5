#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'dense_umma_kernel' --launch-skip 5 --launch-count 1 \
  -o gpurun_out/s37_umma_d384 -f python tools/dense_probe.py --n 12500000 --d 384 --batch 1024 --dtype f16 --iters 2 > gpurun_out/s37_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/s37_ncu.log; ls -la gpurun_out/s37_umma_d384.ncu-rep
