// Device helpers shared by the two sparse scan kernels (bm25.cu: scatter-accumulate, bm25x.cu: bitmap filter).
#pragma once
#include "common.cuh"

namespace brag {

__device__ __forceinline__ void warp_compact(uint64_t* list, int* cnt, int lcap, int k_sel, uint64_t& thr, int lane) {
    const int c = min(*cnt, lcap);
    for (int i = c + lane; i < lcap; i += 32) list[i] = 0ull;
    __syncwarp();
    bitonic_desc(list, lcap, lane, 32, [] { __syncwarp(); });
    const int nc = min(c, k_sel);
    if (nc == k_sel) thr = max(thr, list[k_sel - 1]);
    __syncwarp();
    if (lane == 0) *cnt = nc;
    __syncwarp();
}

// double prefilter for "float32(acc) can reach score t": any acc whose float rounding is >= t is > pre
__device__ __forceinline__ double prefilter_of_score(float t) { return (double)nextafterf(t, -INFINITY); }

__device__ __forceinline__ void worker_bar(int nthreads) { asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory"); }

}  // namespace brag
