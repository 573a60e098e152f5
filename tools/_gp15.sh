#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
: > gpurun_out/s39_ab.txt
for rep in 1 2 3; do
  for lib in old new; do
    if [ $lib = old ]; then export BRAG_LIB=/root/repo/tools/_ab/libbrag_old.so; else unset BRAG_LIB; fi
    echo "$lib $rep" >> gpurun_out/s39_ab.txt
    timeout 300 python tools/dense_probe.py --n 10000000 --d 1024 --batch 1024 --dtype bf16 --iters 10 2>/dev/null >> gpurun_out/s39_ab.txt
  done
done
cat gpurun_out/s39_ab.txt
