// Reciprocal rank fusion of one query's ranked lists by one CTA (reference request
// qdrant_handler.py:357-360).  Shared by rrf_kernel (select.cu) and the sharded fuse kernel (shard.cu).
#pragma once
#include "common.cuh"

namespace brag {

struct RrfEntry {
    double s;
    uint32_t tb;   // tie-break, ascending
    int32_t src;   // flattened source index
    __device__ bool operator<(const RrfEntry& o) const { return s < o.s || (s == o.s && tb > o.tb); }
};

constexpr int RRF_THREADS = 256;
constexpr int RRF_MAX = 8192;          // 8192 * (16 + 12) bytes = 224 KB of the 227 KB a CTA may use

// fid[E] / fpos[E]: the lists' ids flattened in list order with their 0-based ranks (shared memory, filled and
// synchronised by the caller); ent: n2 >= max(E, limit) entries (a power of two); *n_unique = 0 on entry.
// score(id) = sum over its occurrences, in list order, of 1 / (2 + rank) in float64 -- upstream's accumulation order;
// order = (score desc, id asc) for tie_mode 0, (score desc, first seen) otherwise.  Writes `limit` padded slots.
__device__ __forceinline__ void rrf_cta(RrfEntry* ent, const int64_t* fid, const int32_t* fpos, int E, int n2, int limit,
                                        int tie_mode, int* n_unique, double* __restrict__ out_scores,
                                        int64_t* __restrict__ out_ids, int32_t* __restrict__ out_count, int tid,
                                        int nthreads) {
    for (int e = tid; e < n2; e += nthreads) {
        RrfEntry r;
        r.s = -1.0; r.tb = 0xFFFFFFFFu; r.src = -1;
        if (e < E) {
            const int64_t id = fid[e];
            double s = 0.0;
            bool first = true;
            for (int f = 0; f < E; ++f) {
                if (fid[f] == id) {
                    if (f < e) first = false;
                    s += 1.0 / (double)(2 + fpos[f]);   // list order == upstream accumulation order
                }
            }
            if (first) {
                r.s = s;
                r.tb = (tie_mode == 0) ? (uint32_t)id : (uint32_t)e;
                r.src = e;
                atomicAdd(n_unique, 1);
            }
        }
        ent[e] = r;
    }
    __syncthreads();
    bitonic_desc(ent, n2, tid, nthreads, [] { __syncthreads(); });
    const int nu = min(*n_unique, limit);
    for (int i = tid; i < limit; i += nthreads) {
        const bool ok = i < nu;
        out_scores[i] = ok ? ent[i].s : -INFINITY;
        out_ids[i] = ok ? fid[ent[i].src] : -1;
    }
    if (tid == 0) *out_count = nu;
}

}  // namespace brag
