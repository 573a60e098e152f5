#!/bin/bash
# compute-sanitizer over the CI-sized parity tests (race / memory checking of the scans' lock-free protocols:
# bound-manager warps, atomicMax bounds, mbarrier pipelines, TMEM alloc/dealloc).
# ONE tool per gpurun call (B200_PROFILING.md): tools/sanitize.sh memcheck | racecheck | synccheck | initcheck
# The full-size tests are left out (the tools slow kernels 10-100x); logs go to gpurun_out/ and are copied to profiles/.
set -u
tool=${1:-memcheck}
sel='hybrid_three_stage or rrf_ranks or merge_topk or (sparse_topk_bit_exact and 300-500-20) or (sparse_topk_bit_exact and 5000-2000-60 and zipf) or (dense_tcgen05 and 300-64 and 16) or (dense_tcgen05 and 20000-768 and 128 and bf16) or (dense_gemv and 37-64 and f32) or (logical_shards and 2) or dense_rescore_union or sparse_overlap_only'
mkdir -p gpurun_out
compute-sanitizer --tool "$tool" --error-exitcode 77 --print-limit 20 \
    python -m pytest tests/test_gpu_parity.py -m gpu -x -q -p no:cacheprovider -k "$sel" \
    > "gpurun_out/r2_sanitize_${tool}.log" 2>&1
rc=$?
echo "compute-sanitizer --tool $tool exit code: $rc" >> "gpurun_out/r2_sanitize_${tool}.log"
tail -15 "gpurun_out/r2_sanitize_${tool}.log"
exit $rc
