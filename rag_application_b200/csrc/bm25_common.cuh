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

// Per-(term, 2048-document sub-tile) posting offsets are stored only for terms whose postings outweigh such a row
// (sub_row[t] = row of term t in the table, or -1): a hashed vocabulary of millions of rare stems would otherwise cost
// V * (N / 2048) words, more than the postings themselves.  The short list of a row-less term is searched instead:
// first posting at or after document `target`, looking forward from `from` (gallop, then binary search) -- a scan
// visits a term's sub-tiles in ascending order, so the search resumes where the last one ended.
__device__ __forceinline__ uint32_t list_lower_bound(const uint2* __restrict__ list, uint32_t from, uint32_t n,
                                                     uint32_t target) {
    uint32_t lo = from, hi = from, step = 1;
    while (hi < n && __ldg(&list[hi].x) < target) { lo = hi + 1; hi += step; step <<= 1; }
    hi = min(hi, n);
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(&list[mid].x) < target) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ void worker_bar(int nthreads) { asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory"); }

}  // namespace brag
