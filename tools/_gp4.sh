#!/bin/bash
# sparse filter compaction: parity re-check + launch list of the index build at the C5 shard
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "compacted" > gpurun_out/s28_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/s28_tests.log
tail -3 gpurun_out/s28_tests.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none \
  -k regex:'sparse_keep|sparse_renumber|bm25x_wrange|bm25x_nan|bm25x_tau|bm25x_bitmap|suboff|DeviceSelect|DeviceScan|mask_wordcount|mask_rowmap|interleave|DeviceCompact|Sweep' \
  --csv --log-file gpurun_out/s28_build_launches.csv python tools/bench_configs.py --c5 --c5-only > gpurun_out/s28_ncu.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/s28_ncu.log
