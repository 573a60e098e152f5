// BM25 / sparse-dot scoring, bitmap-filter variant (reference request
// qdrant_handler.py:347-354).  Same results as bm25.cu (scatter-accumulate), bit for bit: the score of a
// document is the float64 sum of its matching products in ascending term order, rounded to float32.
//
// Idea: the scatter kernel spends its time adding postings of documents that cannot reach the top k.
// Here a document's score is only evaluated when it CAN exceed the query's running global bound:
//   * every term has an upper bound of its contribution, ub_t = max(0, qv_t*maxw_t, qv_t*minw_t); with
//     the bounds sorted descending and prefix-summed (S_c), a document that contains c of the query's
//     terms scores at most S_c, so it needs c >= m where m = 1 + #{c : S_c <= bound};
//   * per (term, 2048-document sub-tile) the set of documents containing the term is a 2048-bit map:
//     stored at upload for dense terms (df >= N/64: the map is smaller than the postings it stands for),
//     built in shared memory from the few postings otherwise;
//   * the per-document term count is a bit-sliced carry-save sum of the <= MAXT maps (each lane owns
//     64 documents = 2 words per map), compared with m bit-slice-wise, and-ed with the payload filter;
//   * only the surviving documents are scored: weight of term t for document d = posting number
//     sub_off[t][tile] + rank_t(d), rank = prefix popcount of the map (one 8-byte load, no search).
// With a warm bound <1 % of the documents survive, so the kernel reads 1 bit instead of 8 bytes for
// most postings.  Selection (global histograms, bound-manager warp, per-warp lists) is bm25.cu's.
// Roofline: HBM against the ALGORITHMIC bytes sum_t df_t * 8 B (what an exhaustive scan must read);
// the bytes this kernel really moves are reported separately from ncu (profiles/).
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "kernels.h"
#include "bm25_common.cuh"

namespace brag {

constexpr int BX_MAX_WORKERS = 23;                  // + the bound-manager warp = 24 warps (<= 80 registers/thread)
constexpr int BX_THREADS = (BX_MAX_WORKERS + 1) * 32;
constexpr size_t BX_SMEM_MAX = 232448 - 64;
constexpr int BX_WORDS = BM25_SUB / 32;             // 64 map words per sub-tile, 2 per lane
constexpr unsigned long long BX_SENTINEL = 0x8000000000000000ull;   // -0.0 + p == p: first add equals 0.0 + p

constexpr int BX_STAGE = 64;                        // postings of a map-less term staged in shared memory per sub-tile

// per-warp shared memory: bm[MAXT][64] u32 | pre[MAXT][64] u16 | stage[MAXT][64] (doc, weight) | list[lcap] u64 |
//                         sbase[MAXT] i64 | sqv[MAXT] f64 | clist[32] u16 | cnt
__host__ __device__ inline size_t bx_warp_bytes(int maxt, int lcap) {
    return (size_t)maxt * BX_WORDS * 4 + (size_t)maxt * BX_WORDS * 2 + (size_t)maxt * BX_STAGE * 8 + (size_t)lcap * 8 +
           (size_t)maxt * 16 + 64 + 16;
}

// ------------------------------------------------------------------------------------ index build
__global__ void bm25x_wrange_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post, int64_t V,
                                    float2* __restrict__ wrange) {
    __shared__ float s_mn[8], s_mx[8];
    for (int64_t t = blockIdx.x; t < V; t += gridDim.x) {
        const int64_t lo = term_ptr[t], hi = term_ptr[t + 1];
        float mn = INFINITY, mx = -INFINITY;
        for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
            const float w = __uint_as_float(__ldg(&post[i].y));
            mn = fminf(mn, w);
            mx = fmaxf(mx, w);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            mn = fminf(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, off));
            mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, off));
        }
        if ((threadIdx.x & 31) == 0) { s_mn[threadIdx.x >> 5] = mn; s_mx[threadIdx.x >> 5] = mx; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int i = 1; i < (int)(blockDim.x >> 5); ++i) { mn = fminf(mn, s_mn[i]); mx = fmaxf(mx, s_mx[i]); }
            if (hi == lo) { mn = 0.f; mx = 0.f; }
            // NaN weights: fminf/fmaxf drop them; such a term gets an infinite bound (never pruned on)
            wrange[t] = make_float2(mn, mx);
        }
        __syncthreads();
    }
}

__global__ void bm25x_nan_kernel(const uint2* __restrict__ post, int64_t nnz, int* __restrict__ flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nnz) {
        const float w = __uint_as_float(post[i].y);
        if (!(fabsf(w) <= 3.0e38f)) *flag = 1;      // NaN or Inf weight: pruning bounds are meaningless
    }
}

__global__ void bm25x_bitmap_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post,
                                    const int32_t* __restrict__ dense_terms, int64_t words_per_row,
                                    uint32_t* __restrict__ bits) {
    const int row = blockIdx.y;
    const int64_t t = dense_terms[row];
    const int64_t lo = term_ptr[t], hi = term_ptr[t + 1];
    uint32_t* out = bits + (size_t)row * words_per_row;
    for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t d = __ldg(&post[i].x);
        atomicOr(out + (d >> 5), 1u << (d & 31));
    }
}

cudaError_t bm25x_build_wrange(const int64_t* term_ptr, const uint2* post, int64_t V, int64_t nnz, float2* wrange,
                               int* nonfinite_flag, cudaStream_t st) {
    if (V == 0) return cudaSuccess;
    bm25x_wrange_kernel<<<(unsigned)std::min<int64_t>(V, 148 * 16), 256, 0, st>>>(term_ptr, post, V, wrange);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(nonfinite_flag, 0, sizeof(int), st);
    if (e != cudaSuccess || nnz == 0) return e;
    bm25x_nan_kernel<<<(unsigned)((nnz + 255) / 256), 256, 0, st>>>(post, nnz, nonfinite_flag);
    return cudaGetLastError();
}

cudaError_t bm25x_build_bitmaps(const int64_t* term_ptr, const uint2* post, const int32_t* dense_terms, int n_dense,
                                int64_t n_docs, uint32_t* bits, cudaStream_t st) {
    if (n_dense == 0) return cudaSuccess;
    const int64_t ns = (n_docs + BM25_SUB - 1) / BM25_SUB;
    const int64_t words_per_row = ns * BX_WORDS;
    cudaError_t e = cudaMemsetAsync(bits, 0, (size_t)n_dense * words_per_row * 4, st);
    if (e != cudaSuccess) return e;
    for (int r0 = 0; r0 < n_dense; r0 += 32768) {
        const int rows = std::min(32768, n_dense - r0);
        bm25x_bitmap_kernel<<<dim3(64, rows), 256, 0, st>>>(term_ptr, post, dense_terms + r0, words_per_row,
                                                           bits + (size_t)r0 * words_per_row);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ------------------------------------------------------------------------------------ scan
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int MAXT>
__global__ void __launch_bounds__(BX_THREADS, 1)
bm25x_scan_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post,
                  const uint32_t* __restrict__ sub_off, const int32_t* __restrict__ bm_row,
                  const uint32_t* __restrict__ bits, const float2* __restrict__ wrange, int64_t V, int64_t n_docs,
                  const int32_t* __restrict__ q_ptr, const int32_t* __restrict__ q_terms,
                  const double* __restrict__ q_vals, int B, const uint32_t* __restrict__ mask, int k_sel, int lcap,
                  int nw, uint64_t* __restrict__ part, int splits, int* __restrict__ work_counter,
                  uint32_t* __restrict__ gthr, uint32_t* __restrict__ coarse, uint32_t* __restrict__ fine) {
    constexpr int PLANES = (MAXT == 8) ? 4 : (MAXT == 16 ? 5 : 6);     // counts 0..MAXT
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
                const uint32_t bound = hist_bound(coarse, fine, q, (uint32_t)k_sel, lane);
                if (lane == 0 && bound != 0u) atomicMax(gthr + q, bound);
            }
            if (done) break;
            __nanosleep(10000);
        }
        return;
    }

    uint8_t* slab = smem_raw + (size_t)warp * bx_warp_bytes(MAXT, lcap);
    uint32_t* bm = reinterpret_cast<uint32_t*>(slab);                                   // [MAXT][64]
    uint16_t* pre = reinterpret_cast<uint16_t*>(slab + (size_t)MAXT * BX_WORDS * 4);    // [MAXT][64]
    uint2* stage = reinterpret_cast<uint2*>(slab + (size_t)MAXT * BX_WORDS * 6);        // [MAXT][BX_STAGE]
    uint64_t* list = reinterpret_cast<uint64_t*>(stage + MAXT * BX_STAGE);
    int64_t* sbase = reinterpret_cast<int64_t*>(list + lcap);
    double* sqv = reinterpret_cast<double*>(sbase + MAXT);
    uint16_t* clist = reinterpret_cast<uint16_t*>(sqv + MAXT);
    int* cntp = reinterpret_cast<int*>(clist + 32);
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int ns1 = ns + 1;
    const int ns_per_split = (ns + splits - 1) / splits;
    const int n_items = splits * B;
    const double sentinel = __longlong_as_double((long long)BX_SENTINEL);

    while (true) {
        worker_bar(nw * 32);
        if (threadIdx.x == 0) s_item = atomicAdd(work_counter, 1);
        worker_bar(nw * 32);
        const int item = s_item;
        if (item >= n_items) break;
        const int split = item / B, q = item - split * B;
        const int qb = q_ptr[q];
        const int nq = min(q_ptr[q + 1] - qb, MAXT);      // the launcher guarantees nq <= MAXT
        const int j_begin = split * ns_per_split, j_end = min(ns, j_begin + ns_per_split);
        uint64_t thr = 0ull;
        double pre_local = -INFINITY;
        if (lane == 0) *cntp = 0;

        // ---- per-query term setup: lane l < nq owns term l
        int t0 = -1, row0 = -1;
        double qv0 = 0.0, ub = 0.0;
        int64_t tp0 = 0;
        if (lane < nq) {
            const int t = q_terms[qb + lane];
            if (t >= 0 && t < V) {
                t0 = t;
                qv0 = q_vals[qb + lane];
                tp0 = __ldg(term_ptr + t);
                row0 = __ldg(bm_row + t);
                const float2 wr = __ldg(wrange + t);
                ub = fmax(0.0, fmax(qv0 * (double)wr.x, qv0 * (double)wr.y));
                if (!(ub == ub)) ub = INFINITY;           // NaN query value: never prune on this term
                ub *= 1.0 + 1e-12;                         // slack for the rounding of the float64 sums
            }
        }
        if (lane < MAXT) sqv[lane] = qv0;
        // S_c for c = rank + 1: sum of the bounds that sort at or before this lane's (descending, ties by lane)
        double su = 0.0;
#pragma unroll 1
        for (int o = 0; o < nq; ++o) {
            const double uo = __shfl_sync(0xFFFFFFFFu, ub, o);
            if (uo > ub || (uo == ub && o <= lane)) su += uo;
        }
        su *= 1.0 + 1e-12;
        const unsigned sparse_terms = __ballot_sync(0xFFFFFFFFu, t0 >= 0 && row0 < 0);
        __syncwarp();

        uint32_t lo_nx = 0, hi_nx = 0;
        if (t0 >= 0 && j_begin + warp < j_end) {
            lo_nx = __ldg(sub_off + (size_t)t0 * ns1 + (j_begin + warp));
            hi_nx = __ldg(sub_off + (size_t)t0 * ns1 + (j_begin + warp) + 1);
        }
#pragma unroll 1
        for (int j = j_begin + warp; j < j_end; j += nw) {
            const uint32_t sub_base = (uint32_t)j * BM25_SUB;
            const uint32_t g = __ldcg(gthr + q);
            double pre_eff = (g != 0u) ? fmax(pre_local, prefilter_of_score(ordered_f32(g))) : pre_local;
            const uint32_t lo_cur = lo_nx, hi_cur = hi_nx;
            if (t0 >= 0 && j + nw < j_end) {
                lo_nx = __ldg(sub_off + (size_t)t0 * ns1 + (j + nw));
                hi_nx = __ldg(sub_off + (size_t)t0 * ns1 + (j + nw) + 1);
            }
            const uint32_t n = (t0 >= 0) ? hi_cur - lo_cur : 0u;
            const unsigned active = __ballot_sync(0xFFFFFFFFu, n != 0u);
            // a document needs at least m of the query's terms to beat the bound
            const int m = 1 + __popc(__ballot_sync(0xFFFFFFFFu, lane < nq && su <= pre_eff));
            if (__popc(active) < m) continue;

            // ---- phase 1: the (term, sub-tile) document maps into shared memory.  Stored maps and the postings of
            //      map-less terms are all fetched with cp.async before anything waits: one memory round trip
            const int64_t base_cur = tp0 + lo_cur;
            if (n != 0u) sbase[lane] = base_cur;
            for (unsigned a = active; a; a &= a - 1) {
                const int l = __ffs(a) - 1;
                const int row = __shfl_sync(0xFFFFFFFFu, row0, l);
                uint32_t* dst = bm + l * BX_WORDS + 2 * lane;
                if (row >= 0) {
                    cp_async8(dst, bits + ((size_t)row * ns + j) * BX_WORDS + 2 * lane);
                } else {
                    *reinterpret_cast<uint2*>(dst) = make_uint2(0u, 0u);
                    const uint32_t n_l = __shfl_sync(0xFFFFFFFFu, n, l);
                    const uint2* src = post + __shfl_sync(0xFFFFFFFFu, base_cur, l);
                    if ((uint32_t)lane < n_l) cp_async8(stage + l * BX_STAGE + lane, src + lane);
                    if ((uint32_t)lane + 32u < n_l) cp_async8(stage + l * BX_STAGE + lane + 32, src + lane + 32);
                }
            }
            cp_async_wait_all();
            __syncwarp();
            for (unsigned a = active & sparse_terms; a; a &= a - 1) {
                const int l = __ffs(a) - 1;
                const uint32_t n_l = __shfl_sync(0xFFFFFFFFu, n, l);
                for (uint32_t i = lane; i < min(n_l, (uint32_t)BX_STAGE); i += 32) {
                    const uint32_t dl = stage[l * BX_STAGE + i].x - sub_base;
                    atomicOr(bm + l * BX_WORDS + (dl >> 5), 1u << (dl & 31));
                }
                if (n_l > (uint32_t)BX_STAGE) {            // rare: a locally dense stretch of a map-less term
                    const uint2* src = post + __shfl_sync(0xFFFFFFFFu, base_cur, l);
                    for (uint32_t i = BX_STAGE + lane; i < n_l; i += 32) {
                        const uint32_t dl = __ldg(&src[i].x) - sub_base;
                        atomicOr(bm + l * BX_WORDS + (dl >> 5), 1u << (dl & 31));
                    }
                }
            }
            __syncwarp();

            // ---- phase 2: per-word ranks + bit-sliced count of the maps (lane owns words 2*lane, 2*lane+1);
            //      two terms share one warp scan (12-bit counts packed in the halves of a register)
            uint32_t c0[PLANES], c1[PLANES];
#pragma unroll
            for (int p = 0; p < PLANES; ++p) { c0[p] = 0u; c1[p] = 0u; }
            for (unsigned a = active; a;) {
                const int la = __ffs(a) - 1;
                a &= a - 1;
                const bool two = a != 0u;
                const int lb = two ? __ffs(a) - 1 : la;
                a &= a - 1;
                const uint2 wa = *reinterpret_cast<const uint2*>(bm + la * BX_WORDS + 2 * lane);
                uint2 wb = *reinterpret_cast<const uint2*>(bm + lb * BX_WORDS + 2 * lane);
                if (!two) wb = make_uint2(0u, 0u);
                const uint32_t pa0 = __popc(wa.x), pb0 = __popc(wb.x);
                const uint32_t own = (pa0 + __popc(wa.y)) | ((pb0 + __popc(wb.y)) << 16);
                uint32_t inc = own;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, off);
                    if (lane >= off) inc += o;
                }
                const uint32_t ex = inc - own;
                const uint32_t exa = ex & 0xFFFFu, exb = ex >> 16;
                *reinterpret_cast<uint32_t*>(pre + la * BX_WORDS + 2 * lane) = exa | ((exa + pa0) << 16);
                if (two) *reinterpret_cast<uint32_t*>(pre + lb * BX_WORDS + 2 * lane) = exb | ((exb + pb0) << 16);
                uint32_t k0 = wa.x, k1 = wa.y;
#pragma unroll
                for (int p = 0; p < PLANES; ++p) {
                    const uint32_t t0_ = c0[p] & k0, t1_ = c1[p] & k1;
                    c0[p] ^= k0; c1[p] ^= k1;
                    k0 = t0_; k1 = t1_;
                }
                k0 = wb.x; k1 = wb.y;
#pragma unroll
                for (int p = 0; p < PLANES; ++p) {
                    const uint32_t t0_ = c0[p] & k0, t1_ = c1[p] & k1;
                    c0[p] ^= k0; c1[p] ^= k1;
                    k0 = t0_; k1 = t1_;
                }
            }
            // count >= m, bit-slice compare from the top plane down
            uint32_t gt0 = 0u, gt1 = 0u, eq0 = 0xFFFFFFFFu, eq1 = 0xFFFFFFFFu;
#pragma unroll
            for (int p = PLANES - 1; p >= 0; --p) {
                if ((m >> p) & 1) { eq0 &= c0[p]; eq1 &= c1[p]; }
                else { gt0 |= eq0 & c0[p]; gt1 |= eq1 & c1[p]; eq0 &= ~c0[p]; eq1 &= ~c1[p]; }
            }
            uint32_t cand0 = gt0 | eq0, cand1 = gt1 | eq1;
            if (m >> PLANES) { cand0 = 0u; cand1 = 0u; }
            if (mask != nullptr) {
                const uint2 mw = __ldg(reinterpret_cast<const uint2*>(mask + (sub_base >> 5)) + lane);
                cand0 &= mw.x; cand1 &= mw.y;
            }

            // ---- phase 3: score the survivors, 32 at a time (compacted through clist so every lane has one)
            const int own_c = __popc(cand0) + __popc(cand1);
            int pos = own_c;                                  // -> exclusive prefix of own_c
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_up_sync(0xFFFFFFFFu, pos, off);
                if (lane >= off) pos += o;
            }
            const int total = __shfl_sync(0xFFFFFFFFu, pos, 31);
            pos -= own_c;
#pragma unroll 1
            for (int r0 = 0; r0 < total; r0 += 32) {
                if (*cntp > lcap - 32) {
                    warp_compact(list, cntp, lcap, k_sel, thr, lane);
                    if (thr != 0ull) { pre_local = prefilter_of_score(key_score(thr)); pre_eff = fmax(pre_eff, pre_local); }
                }
                while ((cand0 | cand1) != 0u && pos < r0 + 32) {      // this lane's candidates that fall in the window
                    int dl;
                    if (cand0) { dl = __ffs(cand0) - 1 + 64 * lane; cand0 &= cand0 - 1; }
                    else { dl = __ffs(cand1) - 1 + 64 * lane + 32; cand1 &= cand1 - 1; }
                    clist[pos - r0] = (uint16_t)dl;
                    ++pos;
                }
                __syncwarp();
                const bool mine = r0 + lane < total;
                const int dl = mine ? (int)clist[lane] : 0;
                const int word = dl >> 5, bit = dl & 31;
                const uint32_t below = (1u << bit) - 1u;
                float wt[MAXT];
                unsigned hit = 0u, from_stage = 0u;
                uint32_t rk[MAXT];
#pragma unroll
                for (int l = 0; l < MAXT; ++l) {
                    const uint32_t wv = bm[l * BX_WORDS + word];
                    const bool h = mine && ((active >> l) & 1u) && ((wv >> bit) & 1u);
                    rk[l] = (uint32_t)pre[l * BX_WORDS + word] + __popc(wv & below);
                    const bool st = h && ((sparse_terms >> l) & 1u) && rk[l] < (uint32_t)BX_STAGE;
                    const bool gl = h && !st;
                    // unconditional load (a dummy, L1-resident address when this term misses): the MAXT loads issue
                    // back to back instead of one dependent branch per term
                    const uint2* ap = gl ? post + sbase[l] + rk[l] : post;
                    wt[l] = __uint_as_float(__ldg(&ap->y));
                    hit |= (h ? 1u : 0u) << l;
                    from_stage |= (st ? 1u : 0u) << l;
                }
#pragma unroll
                for (int l = 0; l < MAXT; ++l)
                    if ((from_stage >> l) & 1u) wt[l] = __uint_as_float(stage[l * BX_STAGE + rk[l]].y);
                double acc = sentinel;
#pragma unroll
                for (int l = 0; l < MAXT; ++l)
                    if ((hit >> l) & 1u) acc = __dadd_rn(acc, __dmul_rn(sqv[l], (double)wt[l]));
                if (mine && acc > pre_eff) {
                    const uint32_t dg = sub_base + (uint32_t)dl;
                    const uint64_t key = make_key((float)acc, dg);
                    if (key > thr && (uint32_t)(key >> 32) >= g) {
                        list[atomicAdd(cntp, 1)] = key;
                        hist_count(coarse, fine, q, (uint32_t)(key >> 32));
                    }
                }
                __syncwarp();
            }
        }
        // ---- emit this warp's list for (split, warp, q)
        warp_compact(list, cntp, lcap, k_sel, thr, lane);
        uint64_t* out = part + ((size_t)(split * nw + warp) * B + q) * k_sel;
        for (int i = lane; i < k_sel; i += 32) out[i] = list[i];
        __syncwarp();
    }
    if (threadIdx.x == 0) *reinterpret_cast<volatile int*>(&s_done) = 1;
}

static int bx_lcap(int k_sel) {
    int lcap = 128;
    while (lcap < k_sel + 32) lcap <<= 1;     // k_sel kept after a compaction + one round of pushes (<= 32)
    return lcap;
}

static int bx_maxt(int max_terms) { return max_terms <= 8 ? 8 : (max_terms <= 16 ? 16 : 32); }

bool bm25x_supported(int max_terms, int k_sel) { return max_terms >= 0 && max_terms <= 32 && k_sel <= 512; }

Bm25Plan bm25x_plan(int64_t n_docs, int B, int k_sel, int max_terms, int num_sms) {
    Bm25Plan p;
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    p.workers = (int)std::min<size_t>(BX_MAX_WORKERS, BX_SMEM_MAX / bx_warp_bytes(bx_maxt(max_terms), bx_lcap(k_sel)));
    p.workers = std::max(1, std::min(p.workers, ns));
    const int resident = num_sms;
    int splits = (resident * 8 + B - 1) / B;
    int max_splits = (ns + p.workers - 1) / p.workers;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    p.splits = splits;
    p.lists_per_query = splits * p.workers;
    int64_t items = (int64_t)splits * B;
    p.grid = (int)(items < resident ? items : resident);
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    p.gthr_bytes = up((size_t)B * sizeof(uint32_t));
    p.coarse_bytes = up((size_t)B * HIST_COARSE_BINS * sizeof(uint32_t));
    p.fine_bytes = up((size_t)B * HIST_FINE_BINS * sizeof(uint32_t));
    p.scratch_bytes = 256 + p.gthr_bytes + p.coarse_bytes + p.fine_bytes;
    return p;
}

template <int MAXT>
static cudaError_t bx_launch(const int64_t* term_ptr, const uint2* post, const uint32_t* sub_off, const int32_t* bm_row,
                             const uint32_t* bits, const float2* wrange, int64_t V, int64_t n_docs,
                             const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals, int B,
                             const uint32_t* mask, int k_sel, uint64_t* part, const Bm25Plan& plan, void* scratch,
                             cudaStream_t st) {
    const int lcap = bx_lcap(k_sel);
    const size_t smem = bx_warp_bytes(MAXT, lcap) * plan.workers;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(bm25x_scan_kernel<MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
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
    bm25x_scan_kernel<MAXT><<<plan.grid, (plan.workers + 1) * 32, smem, st>>>(
        term_ptr, post, sub_off, bm_row, bits, wrange, V, n_docs, q_ptr, q_terms, q_vals, B, mask, k_sel, lcap,
        plan.workers, part, plan.splits, work_counter, gthr, coarse, fine);
    return cudaGetLastError();
}

cudaError_t bm25x_scan(const int64_t* term_ptr, const uint2* post, const uint32_t* sub_off, const int32_t* bm_row,
                       const uint32_t* bits, const float2* wrange, int64_t V, int64_t n_docs, const int32_t* q_ptr,
                       const int32_t* q_terms, const double* q_vals, int B, int max_terms, const uint32_t* mask,
                       int k_sel, uint64_t* part, const Bm25Plan& plan, void* scratch, cudaStream_t st) {
    if (!bm25x_supported(max_terms, k_sel)) return cudaErrorInvalidValue;
    switch (bx_maxt(max_terms)) {
        case 8:
            return bx_launch<8>(term_ptr, post, sub_off, bm_row, bits, wrange, V, n_docs, q_ptr, q_terms, q_vals, B, mask,
                                k_sel, part, plan, scratch, st);
        case 16:
            return bx_launch<16>(term_ptr, post, sub_off, bm_row, bits, wrange, V, n_docs, q_ptr, q_terms, q_vals, B, mask,
                                 k_sel, part, plan, scratch, st);
        default:
            return bx_launch<32>(term_ptr, post, sub_off, bm_row, bits, wrange, V, n_docs, q_ptr, q_terms, q_vals, B, mask,
                                 k_sel, part, plan, scratch, st);
    }
}

}  // namespace brag
