// BM25 / sparse-dot scoring over a term-major CSR inverted index with fused top-k
// (reference request qdrant_handler.py:347-354).
//
// Semantics restated from qdrant-client local mode: score(d) = sum over shared terms, in
// ascending term order, accumulated in float64, returned as float32; documents sharing no
// term are not candidates.
//
// Roofline: HBM.  Algorithmic bytes per query = sum_t df_t * 8 B (u32 doc + f32 weight).
// Design: warp-autonomous.  A warp owns a 2048-document sub-tile whose f64 accumulators live in
// its private shared-memory slab (initialised to the sentinel -0.0 = "not touched"; -0.0 + p == p
// bit-for-bit, so the first add equals the oracle's 0.0 + p).  A term's postings hit unique docs,
// so they update the slab without atomics, and terms are applied in ascending order with only
// __syncwarp between them -- bit-reproducible.  A per-(term, sub-tile) offset table makes every
// segment lookup O(1); the first 128 postings of term t+1 are in flight while term t is applied.
// After the terms, the warp sweeps its slab once (LDS.128, conflict-free): a double-precision
// prefilter against the running k-th-best threshold rejects almost everything with one compare;
// survivors go to the warp's candidate list (bitonic compaction in shared memory when it fills).
// Thresholds are shared between the warps of a CTA and, through global memory, between CTAs
// working on the same query.  Work items (doc split, query) are claimed split-major so CTAs that
// run together read the same doc range and share posting segments in L2.
// Known edge: a document whose every matching product is exactly -0.0 (zero weight times a
// negative query value) is indistinguishable from "not touched" and is not returned.
#include "common.cuh"
#include "kernels.h"

namespace brag {

constexpr int BM_WARPS = 5;
constexpr int BM_THREADS = BM_WARPS * 32;
constexpr int BM_HEAD = 4;       // posting chunks (x32) prefetched per term
constexpr int BM_TG = 16;        // query terms per group
constexpr uint32_t BM_INVALID = 0xFFFFFFFFu;
constexpr unsigned long long BM_SENTINEL = 0x8000000000000000ull;  // -0.0: "document not touched"

// per-warp shared-memory slab: acc[BM25_SUB] f64 | list[lcap] u64 | cnt
__host__ __device__ inline size_t bm_warp_bytes(int lcap) { return (size_t)BM25_SUB * 8 + (size_t)lcap * 8 + 16; }

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

// float threshold -> double prefilter: any acc whose float rounding is >= thr_score is > pre
__device__ __forceinline__ double prefilter_of(uint64_t thr) {
    if (thr == 0ull) return -INFINITY;
    const float t = key_score(thr);
    return (double)nextafterf(t, -INFINITY);
}

__global__ void __launch_bounds__(BM_THREADS, 2)
bm25_scan_kernel(const int64_t* __restrict__ term_ptr, const uint32_t* __restrict__ doc,
                 const float* __restrict__ w, const uint32_t* __restrict__ sub_off, int64_t V,
                 int64_t n_docs, const int32_t* __restrict__ q_ptr, const int32_t* __restrict__ q_terms,
                 const double* __restrict__ q_vals, int B, const uint32_t* __restrict__ mask, int k_sel,
                 int lcap, uint64_t* __restrict__ part, int splits, int* __restrict__ work_counter,
                 unsigned long long* __restrict__ g_thr) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* slab = smem_raw + (size_t)warp * bm_warp_bytes(lcap);
    double* acc = reinterpret_cast<double*>(slab);
    uint64_t* list = reinterpret_cast<uint64_t*>(slab + (size_t)BM25_SUB * 8);
    int* cntp = reinterpret_cast<int*>(slab + (size_t)BM25_SUB * 8 + (size_t)lcap * 8);
    __shared__ int s_item;
    __shared__ unsigned long long s_cta_thr;
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int ns1 = ns + 1;
    const int ns_per_split = (ns + splits - 1) / splits;
    const int n_items = splits * B;
    const double sentinel = __longlong_as_double((long long)BM_SENTINEL);

    for (int i = lane; i < BM25_SUB; i += 32) acc[i] = sentinel;
    __syncwarp();

    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) { s_item = atomicAdd(work_counter, 1); s_cta_thr = 0ull; }
        __syncthreads();
        const int item = s_item;
        if (item >= n_items) break;
        const int split = item / B, q = item - split * B;
        const int qb = q_ptr[q], qe = q_ptr[q + 1];
        const int j_begin = split * ns_per_split, j_end = min(ns, j_begin + ns_per_split);
        uint64_t thr = __ldcg(g_thr + q);
        double pre = prefilter_of(thr);
        if (lane == 0) *cntp = 0;
        __syncwarp();

        int iter = 0;
        for (int j = j_begin + warp; j < j_end; j += BM_WARPS, ++iter) {
            const uint32_t sub_base = (uint32_t)j * BM25_SUB;
            {   // adopt a tighter bound found by sibling warps (same query) or other CTAs
                uint64_t t = s_cta_thr;
                if ((iter & 15) == 15) t = max(t, (uint64_t)__ldcg(g_thr + q));
                if (t > thr) { thr = t; pre = prefilter_of(thr); }
            }
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

                // ---- terms in ascending order; head of term t+1 is in flight while term t is applied
                uint32_t hd[BM_HEAD], nd[BM_HEAD];
                float hw[BM_HEAD], nw[BM_HEAD];
                auto load_head = [&](int64_t b_t, uint32_t n_t, uint32_t (&d)[BM_HEAD], float (&ww)[BM_HEAD]) {
#pragma unroll
                    for (int u = 0; u < BM_HEAD; ++u) {
                        const uint32_t idx = (uint32_t)u * 32u + lane;
                        d[u] = BM_INVALID;
                        ww[u] = 0.0f;
                        if (idx < n_t) { d[u] = __ldg(doc + b_t + idx); ww[u] = __ldg(w + b_t + idx); }
                    }
                };
                auto apply = [&](uint32_t d, float ww, double qv_t) {
                    if (d != BM_INVALID) {
                        const uint32_t dl = d - sub_base;
                        acc[dl] = __dadd_rn(acc[dl], __dmul_rn(qv_t, (double)ww));   // -0.0 + p == p
                    }
                };
                int64_t b_t = __shfl_sync(0xFFFFFFFFu, base, 0);
                uint32_t n_t = __shfl_sync(0xFFFFFFFFu, n, 0);
                load_head(b_t, n_t, hd, hw);
#pragma unroll 1
                for (int t = 0; t < nq; ++t) {
                    const double qv_t = __shfl_sync(0xFFFFFFFFu, qv, t);
                    const int64_t b_n = __shfl_sync(0xFFFFFFFFu, base, (t + 1) & 31);
                    const uint32_t n_n = (t + 1 < nq) ? __shfl_sync(0xFFFFFFFFu, n, (t + 1) & 31) : 0u;
                    load_head(b_n, n_n, nd, nw);
#pragma unroll
                    for (int u = 0; u < BM_HEAD; ++u) apply(hd[u], hw[u], qv_t);
                    // tail of a long segment: BM_HEAD chunks per trip, loads batched ahead of the updates
                    for (uint32_t i0 = BM_HEAD * 32u; i0 < n_t; i0 += BM_HEAD * 32u) {
                        uint32_t td[BM_HEAD];
                        float tw[BM_HEAD];
#pragma unroll
                        for (int u = 0; u < BM_HEAD; ++u) {
                            const uint32_t idx = i0 + (uint32_t)u * 32u + lane;
                            td[u] = BM_INVALID;
                            tw[u] = 0.0f;
                            if (idx < n_t) { td[u] = __ldg(doc + b_t + idx); tw[u] = __ldg(w + b_t + idx); }
                        }
#pragma unroll
                        for (int u = 0; u < BM_HEAD; ++u) apply(td[u], tw[u], qv_t);
                    }
                    __syncwarp();   // this term's updates are visible before the next term's
#pragma unroll
                    for (int u = 0; u < BM_HEAD; ++u) { hd[u] = nd[u]; hw[u] = nw[u]; }
                    b_t = b_n;
                    n_t = n_n;
                }
            }
            if (!any) continue;

            // ---- scan the sub-tile: prefilter against the running threshold, push survivors, reset
#pragma unroll 1
            for (int it = 0; it < BM25_SUB / 64; ++it) {
                if (*cntp > lcap - 64) { warp_compact(list, cntp, lcap, k_sel, thr, lane); pre = prefilter_of(thr);
                                         if (lane == 0) atomicMax(&s_cta_thr, (unsigned long long)thr); }
                const int d0 = it * 64 + lane * 2;
                const double2 v = *reinterpret_cast<const double2*>(acc + d0);
                *reinterpret_cast<double2*>(acc + d0) = make_double2(sentinel, sentinel);
                if (v.x > pre || v.y > pre) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const double a = h ? v.y : v.x;
                        if (a > pre && (unsigned long long)__double_as_longlong(a) != BM_SENTINEL) {
                            const uint32_t dg = sub_base + d0 + h;
                            if (mask_test(mask, dg)) {
                                const uint64_t key = make_key((float)a, dg);
                                if (key > thr) list[atomicAdd(cntp, 1)] = key;
                            }
                        }
                    }
                }
                __syncwarp();
            }
        }
        // ---- emit this warp's list for (split, warp, q)
        warp_compact(list, cntp, lcap, k_sel, thr, lane);
        if (lane == 0 && thr != 0ull) atomicMax(g_thr + q, (unsigned long long)thr);
        uint64_t* out = part + ((size_t)(split * BM_WARPS + warp) * B + q) * k_sel;
        for (int i = lane; i < k_sel; i += 32) out[i] = list[i];
        __syncwarp();
    }
}

Bm25Plan bm25_plan(int64_t n_docs, int B, int num_sms) {
    Bm25Plan p;
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int resident = num_sms * 2;
    // enough items for ~8 claims per resident CTA, but never more splits than sub-tile quads
    int splits = (resident * 8 + B - 1) / B;
    int max_splits = (ns + BM_WARPS - 1) / BM_WARPS;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    p.splits = splits;
    p.lists_per_query = splits * BM_WARPS;
    int64_t items = (int64_t)splits * B;
    p.grid = (int)(items < resident ? items : resident);
    return p;
}

cudaError_t bm25_scan(const int64_t* term_ptr, const uint32_t* doc, const float* w, const uint32_t* sub_off,
                      int64_t V, int64_t n_docs, const int32_t* q_ptr, const int32_t* q_terms,
                      const double* q_vals, int B, const uint32_t* mask, int k_sel, uint64_t* part,
                      const Bm25Plan& plan, int* work_counter, unsigned long long* g_thr, cudaStream_t st) {
    if (k_sel > 512) return cudaErrorInvalidValue;
    int lcap = 256;
    while (lcap < 2 * k_sel || lcap - 64 < k_sel + 64) lcap <<= 1;   // room for k_sel kept + 64 pushes per scan trip
    const size_t smem = bm_warp_bytes(lcap) * BM_WARPS;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(bm25_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) return e;
        attr_smem = smem;
    }
    cudaError_t e = cudaMemsetAsync(work_counter, 0, sizeof(int), st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(g_thr, 0, (size_t)B * sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    bm25_scan_kernel<<<plan.grid, BM_THREADS, smem, st>>>(term_ptr, doc, w, sub_off, V, n_docs, q_ptr, q_terms,
                                                          q_vals, B, mask, k_sel, lcap, part, plan.splits,
                                                          work_counter, g_thr);
    return cudaGetLastError();
}

}  // namespace brag
