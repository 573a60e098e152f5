#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
: > gpurun_out/s32_c3_knobs.txt
for subs in 1 2 4; do for seed in 0 2 4; do
  echo "SUBS=$subs SEED=$seed" >> gpurun_out/s32_c3_knobs.txt
  BRAG_BX_SUBS=$subs BRAG_BX_SEED=$seed timeout 300 python tools/bm25_probe.py --n 1048576 --batch 1024 --paths bitmap --iters 10 2>&1 | grep -o '"bitmap": {[^}]*}' >> gpurun_out/s32_c3_knobs.txt
done; done
cat gpurun_out/s32_c3_knobs.txt
