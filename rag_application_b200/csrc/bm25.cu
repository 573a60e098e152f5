// BM25 / sparse-dot scoring over a term-major CSR inverted index with fused top-k
// (reference request qdrant_handler.py:347-354).
//
// Semantics restated from qdrant-client local mode: score(d) = sum over shared terms, in
// ascending term order, accumulated in float64, returned as float32; documents sharing no
// term are not candidates.
//
// Roofline: HBM.  Algorithmic bytes per query = sum_t df_t * 8 B (u32 doc + f32 weight).
// Design: warp-autonomous.  A warp owns a 2048-document sub-tile: f64 accumulators and byte
// "touched" flags live in its private shared-memory slab, so a term's postings (unique docs)
// update it without atomics and terms are applied in order with only __syncwarp between
// them -- bit-reproducible.  A per-(term, sub-tile) offset table makes every segment lookup
// O(1); posting chunks stream through a software pipeline DEPTH chunks deep across term
// boundaries.  Each warp keeps a private candidate list with a running k-th-best threshold.
// Work items (doc split, query) are claimed split-major so concurrently running CTAs read
// the same doc range and share posting segments in L2.
#include "common.cuh"
#include "kernels.h"

namespace brag {

constexpr int BM_WARPS = 4;
constexpr int BM_THREADS = BM_WARPS * 32;
constexpr int BM_LCAP = 1024;    // candidate slots per warp
constexpr int BM_DEPTH = 8;      // posting chunks in flight per lane
constexpr int BM_TG = 16;        // query terms per group
constexpr uint32_t BM_INVALID = 0xFFFFFFFFu;

struct BmWarpSmem {
    double acc[BM25_SUB];
    uint64_t list[BM_LCAP];
    uint32_t flags[BM25_SUB / 4];
    double seg_qv[BM_TG];
    int64_t seg_base[BM_TG];
    uint32_t seg_n[BM_TG];
    int seg_cend[BM_TG];
    int cnt;
    int pad;
};

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

__device__ __forceinline__ void warp_compact(BmWarpSmem& S, int k_sel, uint64_t& thr, int lane) {
    const int c = min(S.cnt, BM_LCAP);
    for (int i = c + lane; i < BM_LCAP; i += 32) S.list[i] = 0ull;
    __syncwarp();
    bitonic_desc(S.list, BM_LCAP, lane, 32, [] { __syncwarp(); });
    const int nc = min(c, k_sel);
    thr = (nc == k_sel) ? S.list[k_sel - 1] : 0ull;
    __syncwarp();
    if (lane == 0) S.cnt = nc;
    __syncwarp();
}

__global__ void __launch_bounds__(BM_THREADS, 2)
bm25_scan_kernel(const int64_t* __restrict__ term_ptr, const uint32_t* __restrict__ doc,
                 const float* __restrict__ w, const uint32_t* __restrict__ sub_off, int64_t V,
                 int64_t n_docs, const int32_t* __restrict__ q_ptr, const int32_t* __restrict__ q_terms,
                 const double* __restrict__ q_vals, int B, const uint32_t* __restrict__ mask, int k_sel,
                 uint64_t* __restrict__ part, int splits, int* __restrict__ work_counter) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    BmWarpSmem& S = reinterpret_cast<BmWarpSmem*>(smem_raw)[threadIdx.x >> 5];
    __shared__ int s_item;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int ns1 = ns + 1;
    const int ns_per_split = (ns + splits - 1) / splits;
    const int n_items = splits * B;

    for (int i = lane; i < BM25_SUB / 4; i += 32) S.flags[i] = 0u;
    __syncwarp();

    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) s_item = atomicAdd(work_counter, 1);
        __syncthreads();
        const int item = s_item;
        if (item >= n_items) break;
        const int split = item / B, q = item - split * B;
        const int qb = q_ptr[q], qe = q_ptr[q + 1];
        const int j_begin = split * ns_per_split, j_end = min(ns, j_begin + ns_per_split);
        uint64_t thr = 0ull;
        if (lane == 0) S.cnt = 0;
        __syncwarp();

        for (int j = j_begin + warp; j < j_end; j += BM_WARPS) {
            const uint32_t sub_base = (uint32_t)j * BM25_SUB;
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
                int nch = (int)((n + 31) >> 5);
                int cend = nch;  // inclusive scan over lanes
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    int v = __shfl_up_sync(0xFFFFFFFFu, cend, off);
                    if (lane >= off) cend += v;
                }
                const int total = __shfl_sync(0xFFFFFFFFu, cend, 31);
                if (total == 0) continue;
                any = true;
                if (lane < BM_TG) {
                    S.seg_n[lane] = n; S.seg_base[lane] = base; S.seg_qv[lane] = qv; S.seg_cend[lane] = cend;
                }
                __syncwarp();

                // ---- software-pipelined posting stream
                uint32_t pd[BM_DEPTH];
                float pw[BM_DEPTH];
                int pc = 0, pt = 0, pstart = 0;  // producer chunk, term, first chunk of term
                auto issue = [&](int s) {
                    pd[s] = BM_INVALID;
                    pw[s] = 0.0f;
                    if (pc < total) {
                        while (pc >= S.seg_cend[pt]) { pstart = S.seg_cend[pt]; ++pt; }
                        const uint32_t idx = (uint32_t)(pc - pstart) * 32u + lane;
                        if (idx < S.seg_n[pt]) {
                            const int64_t a = S.seg_base[pt] + idx;
                            pd[s] = __ldg(doc + a);
                            pw[s] = __ldg(w + a);
                        }
                    }
                    ++pc;
                };
#pragma unroll
                for (int s = 0; s < BM_DEPTH; ++s) issue(s);
                int cc = 0, ct = 0;
                double cqv = S.seg_qv[0];
                while (cc < total) {
#pragma unroll
                    for (int s = 0; s < BM_DEPTH; ++s) {
                        if (cc < total) {
                            if (cc >= S.seg_cend[ct]) {
                                while (cc >= S.seg_cend[ct]) ++ct;
                                cqv = S.seg_qv[ct];
                                __syncwarp();  // previous term's updates visible before this term's
                            }
                            if (pd[s] != BM_INVALID) {
                                const uint32_t dl = pd[s] - sub_base;
                                const double prod = __dmul_rn(cqv, (double)pw[s]);
                                uint8_t* fl = reinterpret_cast<uint8_t*>(S.flags) + dl;
                                if (*fl) {
                                    S.acc[dl] = __dadd_rn(S.acc[dl], prod);
                                } else {
                                    S.acc[dl] = __dadd_rn(0.0, prod);
                                    *fl = 1;
                                }
                            }
                            ++cc;
                            issue(s);
                        }
                    }
                }
                __syncwarp();
            }
            if (!any) continue;

            // ---- scan touched docs of this sub-tile, push survivors, clear flags
#pragma unroll 1
            for (int r = 0; r < BM25_SUB / 128; ++r) {
                if (S.cnt > BM_LCAP - 128) warp_compact(S, k_sel, thr, lane);
                const int wi = r * 32 + lane;
                const uint32_t fw = S.flags[wi];
                if (fw) {
                    S.flags[wi] = 0u;
#pragma unroll
                    for (int bb = 0; bb < 4; ++bb) {
                        if ((fw >> (8 * bb)) & 0xFFu) {
                            const uint32_t dl = (uint32_t)wi * 4 + bb;
                            const uint32_t dg = sub_base + dl;
                            if (mask_test(mask, dg)) {
                                const uint64_t key = make_key((float)S.acc[dl], dg);
                                if (key > thr) {
                                    const int pos = atomicAdd(&S.cnt, 1);
                                    S.list[pos] = key;
                                }
                            }
                        }
                    }
                }
                __syncwarp();
            }
        }
        // ---- emit this warp's list for (split, warp, q)
        warp_compact(S, k_sel, thr, lane);
        uint64_t* out = part + ((size_t)(split * BM_WARPS + warp) * B + q) * k_sel;
        for (int i = lane; i < k_sel; i += 32) out[i] = S.list[i];
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
                      const Bm25Plan& plan, int* work_counter, cudaStream_t st) {
    if (k_sel > BM_LCAP - 128 - 128) return cudaErrorInvalidValue;
    static bool attr_set = false;
    const size_t smem = sizeof(BmWarpSmem) * BM_WARPS;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(bm25_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    cudaError_t e = cudaMemsetAsync(work_counter, 0, sizeof(int), st);
    if (e != cudaSuccess) return e;
    bm25_scan_kernel<<<plan.grid, BM_THREADS, smem, st>>>(term_ptr, doc, w, sub_off, V, n_docs, q_ptr, q_terms,
                                                          q_vals, B, mask, k_sel, part, plan.splits, work_counter);
    return cudaGetLastError();
}

}  // namespace brag
