#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
for n in 10000000 1048576; do
  BRAG_BX_WARM=1 timeout 600 python tools/bm25_probe.py --n $n --iters 10 --paths bitmap 2>/dev/null | tee gpurun_out/s34_warm_$n.json
done
