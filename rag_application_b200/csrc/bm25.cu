// BM25 / sparse-dot scoring over a term-major CSR inverted index with fused top-k
// (reference request qdrant_handler.py:347-354).
//
// Semantics restated from qdrant-client local mode: score(d) = sum over shared terms, in
// ascending term order, accumulated in float64, returned as float32; documents sharing no
// term are not candidates.
//
// Roofline: HBM.  Algorithmic bytes per query = sum_t df_t * 8 B (u32 doc + f32 weight).
// Design: warp-autonomous.  A worker warp owns a 2048-document sub-tile whose f64 accumulators live
// in its private shared-memory slab (initialised to the sentinel -0.0 = "not touched"; -0.0 + p == p
// bit-for-bit, so the first add equals the oracle's 0.0 + p).  A term's postings hit unique docs,
// so they update the slab without atomics, and terms are applied in ascending order with only
// __syncwarp between them -- bit-reproducible.  A per-(term, sub-tile) offset table makes every
// segment lookup O(1); the first 128 postings of term t+1 are in flight while term t is applied.
// Selection is fused into the update: every update compares the new partial sum with the query's
// global score bound (common.cuh: per-query score histograms + a bound-manager warp per CTA), and
// the rare documents that exceed it are recorded; since the LAST update of a document produces its
// final score, re-reading the recorded slots after the last term finds every candidate without
// sweeping the slab.  The slab is then reset blindly (32 STS.128 per lane).  If the record buffer
// overflows (no bound yet) the warp falls back to a full sweep of its slab.  Candidates go to the
// warp's shared-memory list (bitonic compaction if it ever fills) and are emitted per work item.
// Work items (doc split, query) are claimed split-major so CTAs that run together read the same
// doc range and share posting segments in L2.
// Known edge: a document whose every matching product is exactly -0.0 (zero weight times a
// negative query value) is indistinguishable from "not touched" and is not returned.
#include "common.cuh"
#include "kernels.h"

namespace brag {

constexpr int BM_WARPS = 12;                     // worker warps (one CTA per SM); one more warp manages the bounds
constexpr int BM_THREADS = (BM_WARPS + 1) * 32;
constexpr int BM_HEAD = 4;       // posting chunks (x32) prefetched per term
constexpr int BM_TG = 16;        // query terms per group
constexpr int BM_REC = 96;       // recorded over-the-bound slots per sub-tile visit before falling back
constexpr uint32_t BM_INVALID = 0xFFFFFFFFu;
constexpr unsigned long long BM_SENTINEL = 0x8000000000000000ull;  // -0.0: "document not touched"

// per-warp shared-memory slab: acc[BM25_SUB] f64 | list[lcap] u64 | rec[BM_REC] u32 | cnt, nrec
__host__ __device__ inline size_t bm_warp_bytes(int lcap) {
    return (size_t)BM25_SUB * 8 + (size_t)lcap * 8 + (size_t)BM_REC * 4 + 16;
}

__global__ void bm25_suboff_kernel(const int64_t* __restrict__ term_ptr, const uint32_t* __restrict__ doc,
                                   int64_t V, int ns1, uint32_t* __restrict__ sub_off) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V * ns1) return;
    const int64_t t = i / ns1;
    const int j = (int)(i - t * ns1);
    const int64_t lo0 = term_ptr[t], hi0 = term_ptr[t + 1];
    const uint64_t target = (uint64_t)j * BM25_SUB;
    int64_t lo = lo0, hi = hi0;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if ((uint64_t)doc[mid] < target) lo = mid + 1; else hi = mid;
    }
    sub_off[i] = (uint32_t)(lo - lo0);
}

cudaError_t bm25_build_suboffsets(const int64_t* term_ptr, const uint32_t* doc, int64_t V, int64_t n_docs,
                                  uint32_t* sub_off, cudaStream_t st) {
    const int ns1 = (int)((n_docs + BM25_SUB - 1) / BM25_SUB) + 1;
    const int64_t total = V * ns1;
    if (total == 0) return cudaSuccess;
    const int64_t blocks = (total + 255) / 256;
    bm25_suboff_kernel<<<(unsigned)blocks, 256, 0, st>>>(term_ptr, doc, V, ns1, sub_off);
    return cudaGetLastError();
}

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

__device__ __forceinline__ void worker_bar() { asm volatile("bar.sync 1, %0;" ::"n"(BM_WARPS * 32) : "memory"); }

__global__ void __launch_bounds__(BM_THREADS, 1)
bm25_scan_kernel(const int64_t* __restrict__ term_ptr, const uint32_t* __restrict__ doc,
                 const float* __restrict__ w, const uint32_t* __restrict__ sub_off, int64_t V,
                 int64_t n_docs, const int32_t* __restrict__ q_ptr, const int32_t* __restrict__ q_terms,
                 const double* __restrict__ q_vals, int B, const uint32_t* __restrict__ mask, int k_sel,
                 int lcap, uint64_t* __restrict__ part, int splits, int* __restrict__ work_counter,
                 uint32_t* __restrict__ gthr, uint32_t* __restrict__ coarse, uint32_t* __restrict__ fine) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ int s_item;
    __shared__ int s_done;
    if (threadIdx.x == 0) s_done = 0;
    __syncthreads();

    if (warp == BM_WARPS) {
        // ===== bound manager: queries blockIdx.x, blockIdx.x + gridDim.x, ...
        while (true) {
            const int done = *reinterpret_cast<volatile int*>(&s_done);
            for (int q = (int)blockIdx.x; q < B; q += (int)gridDim.x) {
                const uint32_t bound = hist_bound(coarse, fine, q, (uint32_t)k_sel, lane);
                if (lane == 0 && bound != 0u) atomicMax(gthr + q, bound);
            }
            if (done) break;
            __nanosleep(20000);
        }
        return;
    }

    uint8_t* slab = smem_raw + (size_t)warp * bm_warp_bytes(lcap);
    double* acc = reinterpret_cast<double*>(slab);
    uint64_t* list = reinterpret_cast<uint64_t*>(slab + (size_t)BM25_SUB * 8);
    uint32_t* rec = reinterpret_cast<uint32_t*>(slab + (size_t)BM25_SUB * 8 + (size_t)lcap * 8);
    int* cntp = reinterpret_cast<int*>(slab + (size_t)BM25_SUB * 8 + (size_t)lcap * 8 + (size_t)BM_REC * 4);
    int* nrecp = cntp + 1;
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int ns1 = ns + 1;
    const int ns_per_split = (ns + splits - 1) / splits;
    const int n_items = splits * B;
    const double sentinel = __longlong_as_double((long long)BM_SENTINEL);
    const double2 sent2 = make_double2(sentinel, sentinel);

    for (int i = lane; i < BM25_SUB; i += 32) acc[i] = sentinel;
    if (lane == 0) *nrecp = 0;
    __syncwarp();

    while (true) {
        worker_bar();
        if (threadIdx.x == 0) s_item = atomicAdd(work_counter, 1);
        worker_bar();
        const int item = s_item;
        if (item >= n_items) break;
        const int split = item / B, q = item - split * B;
        const int qb = q_ptr[q], qe = q_ptr[q + 1];
        const int j_begin = split * ns_per_split, j_end = min(ns, j_begin + ns_per_split);
        uint64_t thr = 0ull;              // local: k_sel-th best key of this warp's list after a compaction
        double pre_local = -INFINITY;
        if (lane == 0) *cntp = 0;
        __syncwarp();

        for (int j = j_begin + warp; j < j_end; j += BM_WARPS) {
            const uint32_t sub_base = (uint32_t)j * BM25_SUB;
            // the query's global bound: k_sel-th best score over everything any warp has pushed so far
            const uint32_t g = __ldcg(gthr + q);
            const double pre = (g != 0u) ? fmax(pre_local, prefilter_of_score(ordered_f32(g))) : pre_local;
            bool any = false;
            for (int tg = qb; tg < qe; tg += BM_TG) {
                const int nq = min(BM_TG, qe - tg);
                // ---- segment setup: lane l < nq owns term tg + l
                uint32_t n = 0;
                int64_t base = 0;
                double qv = 0.0;
                if (lane < nq) {
                    const int t = q_terms[tg + lane];
                    if (t >= 0 && t < V) {
                        const uint32_t lo = __ldg(sub_off + (size_t)t * ns1 + j);
                        const uint32_t hi = __ldg(sub_off + (size_t)t * ns1 + j + 1);
                        n = hi - lo;
                        base = term_ptr[t] + lo;
                        qv = q_vals[tg + lane];
                    }
                }
                if (!__any_sync(0xFFFFFFFFu, n != 0)) continue;
                any = true;

                // ---- the group's postings as one stream of 128-posting pieces, terms in ascending order;
                //      two pieces are always in flight ahead of the one being applied
                struct Piece { uint32_t d[BM_HEAD]; float w[BM_HEAD]; double qv; bool last; };
                int ft = 0;            // fetch cursor: term index inside the group, posting offset in its segment
                uint32_t fi = 0;
                auto fetch = [&](Piece& P) -> bool {
                    uint32_t n_t = 0;
                    while (ft < nq) {                       // warp-uniform
                        n_t = __shfl_sync(0xFFFFFFFFu, n, ft);
                        if (fi < n_t) break;
                        ++ft;
                        fi = 0;
                    }
                    if (ft >= nq) return false;
                    const int64_t b_t = __shfl_sync(0xFFFFFFFFu, base, ft);
                    P.qv = __shfl_sync(0xFFFFFFFFu, qv, ft);
#pragma unroll
                    for (int u = 0; u < BM_HEAD; ++u) {
                        const uint32_t idx = fi + (uint32_t)u * 32u + lane;
                        P.d[u] = BM_INVALID;
                        P.w[u] = 0.0f;
                        if (idx < n_t) { P.d[u] = __ldg(doc + b_t + idx); P.w[u] = __ldg(w + b_t + idx); }
                    }
                    fi += BM_HEAD * 32u;
                    P.last = fi >= n_t;                     // next piece belongs to another term
                    if (P.last) { ++ft; fi = 0; }
                    return true;
                };
                auto apply = [&](uint32_t d, float ww, double qv_t) {
                    if (d != BM_INVALID) {
                        const uint32_t dl = d - sub_base;
                        const double nv = __dadd_rn(acc[dl], __dmul_rn(qv_t, (double)ww));   // -0.0 + p == p
                        acc[dl] = nv;
                        if (nv > pre) {                       // rare once the bound is warm
                            const int pos = atomicAdd(nrecp, 1);
                            if (pos < BM_REC) rec[pos] = dl;
                        }
                    }
                };
                Piece p0, p1, p2;
                bool v0 = fetch(p0);
                bool v1 = v0 && fetch(p1);
                bool v2 = v1 && fetch(p2);
#pragma unroll 1
                while (v0) {
#pragma unroll
                    for (int u = 0; u < BM_HEAD; ++u) apply(p0.d[u], p0.w[u], p0.qv);
                    if (p0.last) __syncwarp();   // a term's updates are visible before the next term's
                    p0 = p1; v0 = v1;
                    p1 = p2; v1 = v2;
                    v2 = v1 && fetch(p2);
                }
            }
            if (!any) continue;
            // ---- selection: re-read the recorded slots (their last update left the final score there)
            const int nrec = *nrecp;
            double pre_eff = pre;
            auto push = [&](double a, uint32_t dl) {
                const uint32_t dg = sub_base + dl;
                if (mask_test(mask, dg)) {
                    const uint64_t key = make_key((float)a, dg);
                    if (key > thr && (uint32_t)(key >> 32) >= g) {
                        list[atomicAdd(cntp, 1)] = key;
                        hist_count(coarse, fine, q, (uint32_t)(key >> 32));
                    }
                }
            };
            if (nrec > BM_REC) {
                // no useful bound yet (or a flood of ties): sweep the whole slab once, resetting as we go
#pragma unroll 1
                for (int it = 0; it < BM25_SUB / 64; ++it) {
                    if (*cntp > lcap - 64) {
                        warp_compact(list, cntp, lcap, k_sel, thr, lane);
                        if (thr != 0ull) { pre_local = prefilter_of_score(key_score(thr)); pre_eff = fmax(pre_eff, pre_local); }
                    }
                    const int d0 = it * 64 + lane * 2;
                    const double2 v = *reinterpret_cast<const double2*>(acc + d0);
                    *reinterpret_cast<double2*>(acc + d0) = sent2;
                    if (v.x > pre_eff || v.y > pre_eff) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const double a = h ? v.y : v.x;
                            if (a > pre_eff && (unsigned long long)__double_as_longlong(a) != BM_SENTINEL) push(a, d0 + h);
                        }
                    }
                    __syncwarp();
                }
            } else {
                if (nrec > 0) {
                    if (*cntp > lcap - BM_REC) {
                        warp_compact(list, cntp, lcap, k_sel, thr, lane);
                        if (thr != 0ull) { pre_local = prefilter_of_score(key_score(thr)); pre_eff = fmax(pre_eff, pre_local); }
                    }
                    for (int i = lane; i < nrec; i += 32) {
                        const uint32_t dl = rec[i];
                        // a slot may be recorded several times: whoever swaps the sentinel in owns it
                        const unsigned long long old =
                            atomicExch(reinterpret_cast<unsigned long long*>(acc + dl), BM_SENTINEL);
                        if (old != BM_SENTINEL) {
                            const double a = __longlong_as_double((long long)old);
                            if (a > pre_eff) push(a, dl);
                        }
                    }
                    __syncwarp();
                }
                double2* acc2 = reinterpret_cast<double2*>(acc);
#pragma unroll 8
                for (int i = lane; i < BM25_SUB / 2; i += 32) acc2[i] = sent2;
            }
            if (lane == 0) *nrecp = 0;
            __syncwarp();
        }
        // ---- emit this warp's list for (split, warp, q)
        warp_compact(list, cntp, lcap, k_sel, thr, lane);
        uint64_t* out = part + ((size_t)(split * BM_WARPS + warp) * B + q) * k_sel;
        for (int i = lane; i < k_sel; i += 32) out[i] = list[i];
        __syncwarp();
    }
    if (threadIdx.x == 0) *reinterpret_cast<volatile int*>(&s_done) = 1;
}

Bm25Plan bm25_plan(int64_t n_docs, int B, int num_sms) {
    Bm25Plan p;
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int resident = num_sms;
    // enough items for ~8 claims per resident CTA, but never more splits than sub-tile groups
    int splits = (resident * 8 + B - 1) / B;
    int max_splits = (ns + BM_WARPS - 1) / BM_WARPS;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    p.splits = splits;
    p.lists_per_query = splits * BM_WARPS;
    int64_t items = (int64_t)splits * B;
    p.grid = (int)(items < resident ? items : resident);
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    p.gthr_bytes = up((size_t)B * sizeof(uint32_t));
    p.coarse_bytes = up((size_t)B * HIST_COARSE_BINS * sizeof(uint32_t));
    p.fine_bytes = up((size_t)B * HIST_FINE_BINS * sizeof(uint32_t));
    p.scratch_bytes = 256 + p.gthr_bytes + p.coarse_bytes + p.fine_bytes;   // work counter | bounds | histograms
    return p;
}

cudaError_t bm25_scan(const int64_t* term_ptr, const uint32_t* doc, const float* w, const uint32_t* sub_off,
                      int64_t V, int64_t n_docs, const int32_t* q_ptr, const int32_t* q_terms,
                      const double* q_vals, int B, const uint32_t* mask, int k_sel, uint64_t* part,
                      const Bm25Plan& plan, void* scratch, cudaStream_t st) {
    if (k_sel > 512) return cudaErrorInvalidValue;
    int lcap = 256;
    while (lcap < k_sel + BM_REC) lcap <<= 1;   // k_sel kept after a compaction + one visit's pushes (<= BM_REC)
    const size_t smem = bm_warp_bytes(lcap) * BM_WARPS;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(bm25_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) return e;
        // two CTAs per SM need the largest shared-memory carve-out; the default heuristic sizes it for one
        e = cudaFuncSetAttribute(bm25_scan_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
        attr_smem = smem;
    }
    uint8_t* sp = reinterpret_cast<uint8_t*>(scratch);
    int* work_counter = reinterpret_cast<int*>(sp);
    uint32_t* gthr = reinterpret_cast<uint32_t*>(sp + 256);
    uint32_t* coarse = reinterpret_cast<uint32_t*>(sp + 256 + plan.gthr_bytes);
    uint32_t* fine = reinterpret_cast<uint32_t*>(sp + 256 + plan.gthr_bytes + plan.coarse_bytes);
    cudaError_t e = cudaMemsetAsync(scratch, 0, plan.scratch_bytes, st);
    if (e != cudaSuccess) return e;
    bm25_scan_kernel<<<plan.grid, BM_THREADS, smem, st>>>(term_ptr, doc, w, sub_off, V, n_docs, q_ptr, q_terms,
                                                          q_vals, B, mask, k_sel, lcap, part, plan.splits,
                                                          work_counter, gthr, coarse, fine);
    return cudaGetLastError();
}

}  // namespace brag
