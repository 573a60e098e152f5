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
#include <algorithm>

#include "common.cuh"
#include "kernels.h"
#include "bm25_common.cuh"

namespace brag {

constexpr int BM_WARPS = 12;                     // max worker warps (one CTA per SM); one more warp manages the bounds
constexpr int BM_THREADS = (BM_WARPS + 1) * 32;
constexpr size_t BM_SMEM_MAX = 232448 - 64;      // opt-in shared memory per CTA on sm_100
constexpr int BM_HEAD = 4;       // posting chunks (x32) prefetched per term
constexpr int BM_TG = 16;        // query terms per group
constexpr int BM_REC = 96;       // recorded over-the-bound slots per sub-tile visit before falling back
constexpr uint32_t BM_INVALID = 0xFFFFFFFFu;
constexpr unsigned long long BM_SENTINEL = 0x8000000000000000ull;  // -0.0: "document not touched"

constexpr int BM_ACC = BM25_SUB + 32;   // + one dummy slot per lane: padding lanes of a chunk update it
// per-warp shared-memory slab: acc[BM_ACC] f64 | list[lcap] u64 | rec[BM_REC] u32 | cnt, nrec
__host__ __device__ inline size_t bm_warp_bytes(int lcap) {
    return (size_t)BM_ACC * 8 + (size_t)lcap * 8 + (size_t)BM_REC * 4 + 16;
}

__global__ void bm25_interleave_kernel(const uint32_t* __restrict__ doc, const float* __restrict__ w, int64_t nnz,
                                       uint2* __restrict__ post) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nnz) post[i] = make_uint2(doc[i], __float_as_uint(w[i]));
}

cudaError_t bm25_interleave(const uint32_t* doc, const float* w, int64_t nnz, uint2* post, cudaStream_t st) {
    if (nnz == 0) return cudaSuccess;
    bm25_interleave_kernel<<<(unsigned)((nnz + 255) / 256), 256, 0, st>>>(doc, w, nnz, post);
    return cudaGetLastError();
}

__global__ void bm25_suboff_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post,
                                   const int32_t* __restrict__ row_terms, int64_t n_rows, int ns1,
                                   uint32_t* __restrict__ sub_off) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows * ns1) return;
    const int64_t t = row_terms[i / ns1];
    const int j = (int)(i % ns1);
    const int64_t lo0 = term_ptr[t], hi0 = term_ptr[t + 1];
    const uint64_t target = (uint64_t)j * BM25_SUB;
    int64_t lo = lo0, hi = hi0;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if ((uint64_t)post[mid].x < target) lo = mid + 1; else hi = mid;
    }
    sub_off[i] = (uint32_t)(lo - lo0);
}

cudaError_t bm25_build_suboffsets(const int64_t* term_ptr, const uint2* post, const int32_t* row_terms, int64_t n_rows,
                                  int64_t n_docs, uint32_t* sub_off, cudaStream_t st) {
    const int ns1 = (int)((n_docs + BM25_SUB - 1) / BM25_SUB) + 1;
    const int64_t total = n_rows * ns1;
    if (total == 0) return cudaSuccess;
    const int64_t blocks = (total + 255) / 256;
    bm25_suboff_kernel<<<(unsigned)blocks, 256, 0, st>>>(term_ptr, post, row_terms, n_rows, ns1, sub_off);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(BM_THREADS, 1)
bm25_scan_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post,
                 const uint32_t* __restrict__ sub_off, const int32_t* __restrict__ sub_row, int64_t V,
                 int64_t n_docs, const int32_t* __restrict__ q_ptr, const int32_t* __restrict__ q_terms,
                 const double* __restrict__ q_vals, int B, const uint32_t* __restrict__ mask, int k_sel,
                 int lcap, int nw, uint64_t* __restrict__ part, int splits, int* __restrict__ work_counter,
                 uint32_t* __restrict__ gthr, uint32_t* __restrict__ coarse, uint32_t* __restrict__ fine,
                 const ShardBounds xb) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ int s_item;
    __shared__ int s_done;
    if (threadIdx.x == 0) s_done = 0;
    __syncthreads();

    if (warp == nw) {
        // ===== bound manager (last warp): queries blockIdx.x, blockIdx.x + gridDim.x, ...
        while (true) {
            const int done = *reinterpret_cast<volatile int*>(&s_done);
            for (int q = (int)blockIdx.x; q < B; q += (int)gridDim.x) {
                publish_bound(gthr, q, hist_bound(coarse, fine, q, (uint32_t)k_sel, lane), xb, lane);
            }
            if (done) break;
            __nanosleep(20000);
        }
        return;
    }

    uint8_t* slab = smem_raw + (size_t)warp * bm_warp_bytes(lcap);
    double* acc = reinterpret_cast<double*>(slab);
    uint64_t* list = reinterpret_cast<uint64_t*>(slab + (size_t)BM_ACC * 8);
    uint32_t* rec = reinterpret_cast<uint32_t*>(slab + (size_t)BM_ACC * 8 + (size_t)lcap * 8);
    int* cntp = reinterpret_cast<int*>(slab + (size_t)BM_ACC * 8 + (size_t)lcap * 8 + (size_t)BM_REC * 4);
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
        worker_bar(nw * 32);
        if (threadIdx.x == 0) s_item = atomicAdd(work_counter, 1);
        worker_bar(nw * 32);
        const int item = s_item;
        if (item >= n_items) break;
        const int split = item / B, q = item - split * B;
        const int qb = q_ptr[q], qe = q_ptr[q + 1];
        const int j_begin = split * ns_per_split, j_end = min(ns, j_begin + ns_per_split);
        uint64_t thr = 0ull;              // local: k_sel-th best key of this warp's list after a compaction
        double pre_local = -INFINITY;
        if (lane == 0) *cntp = 0;
        __syncwarp();

        // first term group of the query: term row, query value and posting base are the same for every
        // sub-tile, and the (lo, hi) offsets of the NEXT sub-tile are fetched while this one is processed
        int t0 = -1, srow0 = -1;
        double qv0 = 0.0;
        int64_t tp0 = 0;
        uint32_t df0 = 0, cur0 = 0;       // row-less term: list length, where the last search ended
        if (lane < min(BM_TG, qe - qb)) {
            const int t = q_terms[qb + lane];
            if (t >= 0 && t < V) {
                t0 = t; qv0 = q_vals[qb + lane]; tp0 = __ldg(term_ptr + t);
                srow0 = __ldg(sub_row + t);
                df0 = (uint32_t)(__ldg(term_ptr + t + 1) - tp0);
            }
        }
        // posting range of this lane's term inside sub-tile s: table row, or a forward search of the short list
        auto range_of = [&](int s, uint32_t& lo, uint32_t& hi) {
            if (srow0 >= 0) {
                lo = __ldg(sub_off + (size_t)srow0 * ns1 + s);
                hi = __ldg(sub_off + (size_t)srow0 * ns1 + s + 1);
            } else {
                lo = list_lower_bound(post + tp0, cur0, df0, (uint32_t)s * BM25_SUB);
                hi = list_lower_bound(post + tp0, lo, df0, (uint32_t)(s + 1) * BM25_SUB);
                cur0 = hi;
            }
        };
        uint32_t lo_nx = 0, hi_nx = 0;
        if (t0 >= 0 && j_begin + warp < j_end) range_of(j_begin + warp, lo_nx, hi_nx);
        for (int j = j_begin + warp; j < j_end; j += nw) {
            const uint32_t sub_base = (uint32_t)j * BM25_SUB;
            // the query's global bound: k_sel-th best score over everything any warp has pushed so far
            const uint32_t g = __ldcg(gthr + q);
            const double pre = (g != 0u) ? fmax(pre_local, prefilter_of_score(ordered_f32(g))) : pre_local;
            bool any = false;
            const uint32_t lo_cur = lo_nx, hi_cur = hi_nx;
            if (t0 >= 0 && j + nw < j_end) range_of(j + nw, lo_nx, hi_nx);
            for (int tg = qb; tg < qe; tg += BM_TG) {
                const int nq = min(BM_TG, qe - tg);
                // ---- segment setup: lane l < nq owns term tg + l
                uint32_t n = 0;
                int64_t base = 0;
                double qv = 0.0;
                if (tg == qb) {
                    if (t0 >= 0) { n = hi_cur - lo_cur; base = tp0 + lo_cur; qv = qv0; }
                } else if (lane < nq) {
                    const int t = q_terms[tg + lane];
                    if (t >= 0 && t < V) {
                        const int srow = __ldg(sub_row + t);
                        const int64_t tp = __ldg(term_ptr + t);
                        uint32_t lo, hi;
                        if (srow >= 0) {
                            lo = __ldg(sub_off + (size_t)srow * ns1 + j);
                            hi = __ldg(sub_off + (size_t)srow * ns1 + j + 1);
                        } else {
                            const uint32_t df = (uint32_t)(__ldg(term_ptr + t + 1) - tp);
                            lo = list_lower_bound(post + tp, 0u, df, sub_base);
                            hi = list_lower_bound(post + tp, lo, df, sub_base + BM25_SUB);
                        }
                        n = hi - lo;
                        base = tp + lo;
                        qv = q_vals[tg + lane];
                    }
                }
                unsigned active = __ballot_sync(0xFFFFFFFFu, n != 0);   // terms with postings in this sub-tile
                if (active == 0u) continue;
                any = true;

                // ---- the group's postings as one stream of <=128-posting pieces, terms in ascending order;
                //      two pieces are always in flight ahead of the one being applied
                struct Piece { uint2 v[BM_HEAD]; double qv; uint32_t cnt; bool last; };
                uint32_t c_n = 0, c_i = 0;       // current term's segment length / fetch offset
                int64_t c_base = 0;
                double c_qv = 0.0;
                bool more = true;
                auto fetch = [&](Piece& P) -> bool {
                    if (c_i >= c_n) {                       // next term with postings (warp-uniform)
                        if (active == 0u) { more = false; return false; }
                        const int t = __ffs(active) - 1;
                        active &= active - 1;
                        c_n = __shfl_sync(0xFFFFFFFFu, n, t);
                        c_base = __shfl_sync(0xFFFFFFFFu, base, t);
                        c_qv = __shfl_sync(0xFFFFFFFFu, qv, t);
                        c_i = 0;
                    }
                    const uint32_t rem = c_n - c_i;
                    const uint2* src = post + c_base + c_i;
                    P.qv = c_qv;
                    P.cnt = rem;
#pragma unroll
                    for (int u = 0; u < BM_HEAD; ++u) {
                        const uint32_t idx = (uint32_t)u * 32u + lane;
                        P.v[u] = make_uint2(BM_INVALID, 0u);
                        if (idx < rem) P.v[u] = __ldg(src + idx);
                    }
                    c_i += BM_HEAD * 32u;
                    P.last = c_i >= c_n;                    // the next piece belongs to another term
                    return true;
                };
                // two chunks at a time: the documents of one term are distinct, so the accumulator reads
                // of a pair are issued together, then the adds, then the writes
                auto apply2 = [&](const uint2& x0, const uint2& x1, double qv_t) {
                    const uint32_t dl0 = (x0.x != BM_INVALID) ? x0.x - sub_base : (uint32_t)BM25_SUB + lane;
                    const uint32_t dl1 = (x1.x != BM_INVALID) ? x1.x - sub_base : (uint32_t)BM25_SUB + lane;
                    const double a0 = acc[dl0], a1 = acc[dl1];
                    const double n0 = __dadd_rn(a0, __dmul_rn(qv_t, (double)__uint_as_float(x0.y)));   // -0.0 + p == p
                    const double n1 = __dadd_rn(a1, __dmul_rn(qv_t, (double)__uint_as_float(x1.y)));
                    acc[dl0] = n0;
                    acc[dl1] = n1;
                    if (n0 > pre || n1 > pre) {               // rare once the bound is warm
                        if (n0 > pre && x0.x != BM_INVALID) { const int pos = atomicAdd(nrecp, 1); if (pos < BM_REC) rec[pos] = dl0; }
                        if (n1 > pre && x1.x != BM_INVALID) { const int pos = atomicAdd(nrecp, 1); if (pos < BM_REC) rec[pos] = dl1; }
                    }
                };
                auto apply = [&](const Piece& P) {
                    apply2(P.v[0], P.v[1], P.qv);
                    if (P.cnt > 64u) apply2(P.v[2], P.v[3], P.qv);   // warp-uniform
                    if (P.last) __syncwarp();     // a term's updates are visible before the next term's
                };
                Piece p0, p1, p2, p3, p4;           // four pieces in flight ahead of the one being applied
                bool v0 = fetch(p0);
                bool v1 = v0 && fetch(p1);
                bool v2 = v1 && fetch(p2);
                bool v3 = v2 && fetch(p3);
                bool v4 = v3 && fetch(p4);
#pragma unroll 1
                while (true) {
                    if (!v0) break;
                    apply(p0);
                    v0 = more && fetch(p0);
                    if (!v1) break;
                    apply(p1);
                    v1 = more && fetch(p1);
                    if (!v2) break;
                    apply(p2);
                    v2 = more && fetch(p2);
                    if (!v3) break;
                    apply(p3);
                    v3 = more && fetch(p3);
                    if (!v4) break;
                    apply(p4);
                    v4 = more && fetch(p4);
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
        uint64_t* out = part + ((size_t)(split * nw + warp) * B + q) * k_sel;
        for (int i = lane; i < k_sel; i += 32) out[i] = list[i];
        __syncwarp();
    }
    if (threadIdx.x == 0) *reinterpret_cast<volatile int*>(&s_done) = 1;
}

static int bm25_lcap(int k_sel) {
    int lcap = 256;
    while (lcap < k_sel + BM_REC) lcap <<= 1;   // k_sel kept after a compaction + one visit's pushes (<= BM_REC)
    return lcap;
}

Bm25Plan bm25_plan(int64_t n_docs, int B, int k_sel, int num_sms) {
    Bm25Plan p;
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    p.workers = (int)std::min<size_t>(BM_WARPS, BM_SMEM_MAX / bm_warp_bytes(bm25_lcap(k_sel)));
    const int resident = num_sms;
    // enough items for ~8 claims per resident CTA, but never more splits than sub-tile groups
    int splits = (resident * 8 + B - 1) / B;
    int max_splits = (ns + p.workers - 1) / p.workers;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    p.splits = splits;
    p.lists_per_query = splits * p.workers;
    p.visits_per_list = ns / p.lists_per_query;
    int64_t items = (int64_t)splits * B;
    p.grid = (int)(items < resident ? items : resident);
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    p.gthr_bytes = up((size_t)B * sizeof(uint32_t));
    p.coarse_bytes = up((size_t)B * HIST_COARSE_BINS * sizeof(uint32_t));
    p.fine_bytes = up((size_t)B * HIST_FINE_BINS * sizeof(uint32_t));
    p.scratch_bytes = 256 + p.gthr_bytes + p.coarse_bytes + p.fine_bytes;   // work counter | bounds | histograms
    return p;
}

// kernel attributes are per device: brag_create calls this with the engine's device current
cudaError_t bm25_init_device() {
    cudaError_t e = cudaFuncSetAttribute(bm25_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BM_SMEM_MAX);
    if (e != cudaSuccess) return e;
    // the largest shared-memory carve-out; the default heuristic sizes it from the first launch
    return cudaFuncSetAttribute(bm25_scan_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

cudaError_t bm25_scan(const int64_t* term_ptr, const uint2* post, const uint32_t* sub_off, const int32_t* sub_row,
                      int64_t V, int64_t n_docs, const int32_t* q_ptr, const int32_t* q_terms,
                      const double* q_vals, int B, const uint32_t* mask, int k_sel, uint64_t* part,
                      const Bm25Plan& plan, void* scratch, cudaStream_t st, const ShardBounds& xb) {
    if (k_sel > 512) return cudaErrorInvalidValue;
    const int lcap = bm25_lcap(k_sel);
    const size_t smem = bm_warp_bytes(lcap) * plan.workers;
    if (smem > BM_SMEM_MAX) return cudaErrorInvalidValue;
    uint8_t* sp = reinterpret_cast<uint8_t*>(scratch);
    int* work_counter = reinterpret_cast<int*>(sp);
    uint32_t* gthr = reinterpret_cast<uint32_t*>(sp + 256);
    uint32_t* coarse = reinterpret_cast<uint32_t*>(sp + 256 + plan.gthr_bytes);
    uint32_t* fine = reinterpret_cast<uint32_t*>(sp + 256 + plan.gthr_bytes + plan.coarse_bytes);
    cudaError_t e = cudaMemsetAsync(scratch, 0, plan.scratch_bytes, st);
    if (e != cudaSuccess) return e;
    bm25_scan_kernel<<<plan.grid, (plan.workers + 1) * 32, smem, st>>>(term_ptr, post, sub_off, sub_row, V, n_docs, q_ptr, q_terms,
                                                          q_vals, B, mask, k_sel, lcap, plan.workers, part, plan.splits,
                                                          work_counter, gthr, coarse, fine, xb);
    return cudaGetLastError();
}

}  // namespace brag
