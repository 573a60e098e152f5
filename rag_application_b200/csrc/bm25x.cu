// BM25 / sparse-dot scoring, bitmap-filter variant (reference request
// qdrant_handler.py:347-354).  Same results as bm25.cu (scatter-accumulate), bit for bit: the score of a
// document is the float64 sum of its matching products in ascending term order, rounded to float32.
//
// Idea: the scatter kernel spends its time adding postings of documents that cannot reach the top k.
// Here a document's score is only evaluated when it CAN exceed the query's running global bound:
//   * every term t splits its postings at a weight threshold tau_t into a low class (w <= tau_t, >= ~70 % of
//     them; for BM25: tf = 1) and a high class.  A low posting contributes at most ua_t = max(0, qv*min_t,
//     qv*tau_t), a high one at most ua_t + ue_t with ue_t = max(ua_t, qv*max_t) - ua_t.  With the 2*nq values
//     {ua_t} U {ue_t} sorted descending and prefix-summed (S_c), a document that contains c1 of the query's
//     terms, c2 of them in their high class, scores at most S_(c1+c2): it needs c1 + c2 >= m, where
//     m = 1 + #{c : S_c <= bound}.  Valid for any sign of weights and query values;
//   * per (term, 2048-document sub-tile) the documents containing the term (and those in its high class) are
//     2048-bit maps: stored at upload for dense terms (df >= N/256, within as many bytes as the postings take: the
//     scan is issue-bound, and building a map from postings costs ~10 instructions per posting where fetching a
//     stored one costs two cp.async per lane), built in shared memory from the few postings (staged with cp.async)
//     otherwise;
//   * c1 + c2 per document is a bit-sliced carry-save sum of the <= 2*MAXT maps (each lane owns 64*SUBS
//     documents), compared with m bit-slice-wise, and-ed with the payload filter;
//   * only the surviving documents are scored: weight of term t for document d = posting number
//     sub_off[t][tile] + rank_t(d), rank = prefix popcount of the map (one 8-byte load, no search); survivors
//     are compacted so that every lane of a scoring round holds one; the float64 sum runs in ascending term order.
// With a warm bound ~0.3 % of the documents survive, so the kernel reads 2 bits instead of 8 bytes for most
// postings.  Selection (global histograms, bound-manager warp, per-warp lists) is bm25.cu's.
// Roofline: HBM against the ALGORITHMIC bytes sum_t df_t * 8 B (what an exhaustive scan must read);
// the bytes this kernel really moves are reported separately from ncu (profiles/).
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "kernels.h"
#include "bm25_common.cuh"

namespace brag {

constexpr int BX_MAX_WORKERS = 23;                  // + the bound-manager warp = 24 warps; bx_worker_cap() per instantiation
constexpr size_t BX_SMEM_MAX = 232448 - 64;
constexpr int BX_WORDS = BM25_SUB / 32;             // 64 map words per sub-tile, 2 per lane
constexpr unsigned long long BX_SENTINEL = 0x8000000000000000ull;   // -0.0 + p == p: first add equals 0.0 + p

// One visit of a worker warp covers SUBS consecutive 2048-document sub-tiles (the granularity of sub_off): more
// documents per visit amortise the per-visit and per-term overheads, fewer leave room for more resident warps.
template <int MAXT, int SUBS>
struct BxCfg {
    static constexpr int WPL = 2 * SUBS;          // map words per lane
    static constexpr int WORDS = 64 * SUBS;       // map words per tile
    static constexpr int TILE = BM25_SUB * SUBS;  // documents per tile
    static constexpr int POOL = 128 * SUBS;       // postings of map-less terms staged in shared memory per tile
    static constexpr int PLANES = (MAXT == 8) ? 5 : (MAXT == 16 ? 6 : 7);     // bit planes of the counts 0..2*MAXT
};

// per-warp shared memory: bm[MAXT][WORDS] u32 | bmh[MAXT][WORDS] u32 (the rank table pre[MAXT][WORDS] u16 overwrites
//                         each term's high map once it has been counted) | pool[POOL] (doc, weight) | list[lcap] u64 |
//                         sbase[MAXT] i64 | sqv[MAXT] f64 | spoff[MAXT] i32 | spcnt[MAXT] i32 | clist[32] u16 | cnt
__host__ __device__ constexpr size_t bx_warp_bytes(int maxt, int subs, int lcap) {
    return (size_t)maxt * 64 * subs * 8 + (size_t)128 * subs * 8 + (size_t)lcap * 8 + (size_t)maxt * 24 + 64 + 16;
}
// Worker warps per CTA of an instantiation: what shared memory allows with the shortest list (lcap = 128), at most
// BX_MAX_WORKERS; registers are allocated per 4 warps, so a count of 16 or more is rounded down to 4k - 1 (+ the manager
// warp = 4k warps): <8, 2> gets 19 workers = 640 threads -> 96 registers instead of the 80 that 21+ warps allow.
constexpr int bx_worker_cap(int maxt, int subs) {
    const size_t f = BX_SMEM_MAX / bx_warp_bytes(maxt, subs, 128);
    const int w = (int)(f < (size_t)BX_MAX_WORKERS ? f : (size_t)BX_MAX_WORKERS);
    return w >= 16 ? (w + 1) / 4 * 4 - 1 : w;
}
constexpr int bx_max_threads(int maxt, int subs) { return (bx_worker_cap(maxt, subs) + 1) * 32; }

// words of one stored document map: the sub-tile count is rounded up to 4 so a tile of any SUBS stays inside its row
int64_t bm25x_row_words(int64_t n_docs) {
    const int64_t ns = (n_docs + BM25_SUB - 1) / BM25_SUB;
    return (ns + 3) / 4 * 4 * BX_WORDS;
}

// ------------------------------------------------------------------------------------ index build
// Weight range of every term, and whether any weight is NaN / Inf, in one pass over the postings: a warp walks chunks of
// WR_CHUNK consecutive postings (four coalesced loads per lane in flight), follows the term boundaries from one search
// per chunk, keeps min / max in registers and folds them into the term's slot (order-preserving uint32 images, so
// atomicMin / atomicMax work) once per (chunk, term).  `slots` aliases wrange: {min image, max image} per term,
// converted in place by bm25x_wrange_fin_kernel.
constexpr int WR_CHUNK = 8192;

__global__ void bm25x_wrange_init_kernel(uint2* __restrict__ slots, int64_t V) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < V) slots[t] = make_uint2(0xFFFFFFFFu, 0u);
}

__device__ __forceinline__ void wr_fold(uint2* __restrict__ slots, int64_t t, float mn, float mx) {
    if (mn <= mx) {          // false when nothing finite-or-infinite was seen (empty segment, or only NaNs)
        atomicMin(&slots[t].x, f32_ordered(mn));
        atomicMax(&slots[t].y, f32_ordered(mx));
    }
}

__global__ void bm25x_wrange_walk_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post, int64_t V,
                                         int64_t nnz, uint2* __restrict__ slots, int* __restrict__ flag) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t nchunks = (nnz + WR_CHUNK - 1) / WR_CHUNK;
    bool bad = false;
    for (int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < nchunks; c += nwarps) {
        const int64_t i0 = c * WR_CHUNK, i1 = min(nnz, i0 + (int64_t)WR_CHUNK);
        int64_t t = term_of_posting(term_ptr, V, i0);
        int64_t end = __ldg(term_ptr + t + 1);
        float mn = INFINITY, mx = -INFINITY;
        for (int64_t base = i0; base < i1; base += 128) {
            float w[4];
            bool have[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t i = base + u * 32 + lane;
                have[u] = i < i1;
                w[u] = have[u] ? __uint_as_float(__ldcs(post + i).y) : 0.f;
                bad |= !(fabsf(w[u]) <= 3.0e38f);      // NaN or Inf weight: pruning bounds are meaningless
            }
            if (base + 128 <= end) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (have[u]) { mn = fminf(mn, w[u]); mx = fmaxf(mx, w[u]); }
                continue;
            }
            // a term ends inside this block: fold what the warp holds, then one search per posting
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                mn = fminf(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, o));
                mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
            }
            if (lane == 0) wr_fold(slots, t, mn, mx);
            mn = INFINITY;
            mx = -INFINITY;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (have[u] && w[u] == w[u])
                    wr_fold(slots, term_of_posting(term_ptr, V, base + u * 32 + lane), w[u], w[u]);
            if (base + 128 < i1) {
                t = term_of_posting(term_ptr, V, base + 128);
                end = __ldg(term_ptr + t + 1);
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            mn = fminf(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, o));
            mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
        }
        if (lane == 0) wr_fold(slots, t, mn, mx);
    }
    if (bad) *flag = 1;
}

__global__ void bm25x_wrange_fin_kernel(const int64_t* __restrict__ term_ptr, float2* __restrict__ wrange, int64_t V) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= V) return;
    const uint2 s = reinterpret_cast<const uint2*>(wrange)[t];
    float2 r;
    if (term_ptr[t + 1] == term_ptr[t]) r = make_float2(0.f, 0.f);
    else if (s.x > s.y) r = make_float2(INFINITY, -INFINITY);      // only NaN weights: never pruned on
    else r = make_float2(ordered_f32(s.x), ordered_f32(s.y));
    wrange[t] = r;
}

// tau[t]: weight threshold that splits the postings of term t into a "low" class (w <= tau, at least ~70 % of them)
// and a "high" class: 64-bin histogram over [min, max], tau = upper edge of the first bin whose cumulative count
// reaches 70 % (tau = max when the term has a single weight).  For BM25 the low class is the tf = 1 postings.
constexpr int TAU_SAMPLES = 4096;

__global__ void bm25x_tau_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post, int64_t V,
                                 const float2* __restrict__ wrange, float* __restrict__ tau) {
    __shared__ unsigned int hist[64];
    for (int64_t t = blockIdx.x; t < V; t += gridDim.x) {
        const int64_t lo = term_ptr[t], hi = term_ptr[t + 1];
        const float2 wr = wrange[t];
        const float span = wr.y - wr.x;
        // tau only steers how well the two classes prune (the scan and the maps use the same value, any value is
        // safe), so long lists are sampled: TAU_SAMPLES postings at an even stride
        const int64_t stride = max((int64_t)1, (hi - lo) / TAU_SAMPLES);
        const int64_t ns = (hi - lo + stride - 1) / stride;
        if (threadIdx.x < 64) hist[threadIdx.x] = 0u;
        __syncthreads();
        if (span > 0.f && span < 3.0e38f) {
            const float inv = 64.f / span;
            for (int64_t j = threadIdx.x; j < ns; j += blockDim.x) {
                const float w = __uint_as_float(__ldg(&post[lo + j * stride].y));
                int b = (int)((w - wr.x) * inv);
                b = min(63, max(0, b));
                atomicAdd(&hist[b], 1u);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float tv = wr.y;
            if (span > 0.f && span < 3.0e38f) {
                const unsigned long long need = (unsigned long long)(ns * 7 / 10);
                unsigned long long cum = 0;
                for (int b = 0; b < 64; ++b) {
                    cum += hist[b];
                    if (cum >= need) {
                        // upper edge of the bin; the class test is w > tau everywhere, so rounding here is harmless
                        tv = (b >= 63) ? wr.y : wr.x + span * (float)(b + 1) / 64.f;
                        break;
                    }
                }
            }
            tau[t] = fminf(tv, wr.y);
        }
        __syncthreads();
    }
}

// present map (every posting) and high map (postings with w > tau) of the dense terms
__global__ void bm25x_bitmap_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post,
                                    const int32_t* __restrict__ dense_terms, const float* __restrict__ tau,
                                    int64_t words_per_row, uint32_t* __restrict__ bits, uint32_t* __restrict__ bits_hi) {
    const int row = blockIdx.y;
    const int64_t t = dense_terms[row];
    const int64_t lo = term_ptr[t], hi = term_ptr[t + 1];
    const float tv = tau[t];
    uint32_t* out = bits + (size_t)row * words_per_row;
    uint32_t* out_hi = bits_hi + (size_t)row * words_per_row;
    for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t)gridDim.x * blockDim.x) {
        const uint2 p = __ldg(&post[i]);
        atomicOr(out + (p.x >> 5), 1u << (p.x & 31));
        if (__uint_as_float(p.y) > tv) atomicOr(out_hi + (p.x >> 5), 1u << (p.x & 31));
    }
}

cudaError_t bm25x_build_wrange(const int64_t* term_ptr, const uint2* post, int64_t V, int64_t nnz, float2* wrange,
                               float* tau, int* nonfinite_flag, cudaStream_t st) {
    if (V == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(nonfinite_flag, 0, sizeof(int), st);
    if (e != cudaSuccess) return e;
    bm25x_wrange_init_kernel<<<(unsigned)((V + 255) / 256), 256, 0, st>>>(reinterpret_cast<uint2*>(wrange), V);
    if (nnz) {
        const int64_t blocks = std::min<int64_t>((nnz + WR_CHUNK * 8 - 1) / (WR_CHUNK * 8), 148 * 8);
        bm25x_wrange_walk_kernel<<<(unsigned)blocks, 256, 0, st>>>(term_ptr, post, V, nnz,
                                                                  reinterpret_cast<uint2*>(wrange), nonfinite_flag);
    }
    bm25x_wrange_fin_kernel<<<(unsigned)((V + 255) / 256), 256, 0, st>>>(term_ptr, wrange, V);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    bm25x_tau_kernel<<<(unsigned)std::min<int64_t>(V, 148 * 16), 256, 0, st>>>(term_ptr, post, V, wrange, tau);
    return cudaGetLastError();
}

cudaError_t bm25x_build_bitmaps(const int64_t* term_ptr, const uint2* post, const int32_t* dense_terms, int n_dense,
                                const float* tau, int64_t n_docs, uint32_t* bits, uint32_t* bits_hi, cudaStream_t st) {
    if (n_dense == 0) return cudaSuccess;
    const int64_t words_per_row = bm25x_row_words(n_docs);
    cudaError_t e = cudaMemsetAsync(bits, 0, (size_t)n_dense * words_per_row * 4, st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(bits_hi, 0, (size_t)n_dense * words_per_row * 4, st);
    if (e != cudaSuccess) return e;
    for (int r0 = 0; r0 < n_dense; r0 += 32768) {
        const int rows = std::min(32768, n_dense - r0);
        bm25x_bitmap_kernel<<<dim3(64, rows), 256, 0, st>>>(term_ptr, post, dense_terms + r0, tau, words_per_row,
                                                           bits + (size_t)r0 * words_per_row,
                                                           bits_hi + (size_t)r0 * words_per_row);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ------------------------------------------------------------------------------------ scan
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gsrc) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(d), "l"(gsrc), "n"(BYTES) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// A lane's WPL consecutive map words (16-byte aligned when WPL is a multiple of 4): one 128-bit shared access per
// four words.  Two 64-bit accesses at a 16-byte lane stride are 2-way bank-conflicted (ncu: 4 wavefronts each
// instead of 2), so the wide form halves the shared-memory wavefronts of the map reads.
template <int WPL>
__device__ __forceinline__ void lds_words(const uint32_t* p, uint32_t (&v)[WPL]) {
    if constexpr (WPL % 4 == 0) {
#pragma unroll
        for (int w = 0; w < WPL; w += 4) {
            const uint4 t = *reinterpret_cast<const uint4*>(p + w);
            v[w] = t.x; v[w + 1] = t.y; v[w + 2] = t.z; v[w + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int w = 0; w < WPL; w += 2) {
            const uint2 t = *reinterpret_cast<const uint2*>(p + w);
            v[w] = t.x; v[w + 1] = t.y;
        }
    }
}
template <int WPL>
__device__ __forceinline__ void sts_zero_words(uint32_t* p) {
    if constexpr (WPL % 4 == 0) {
#pragma unroll
        for (int w = 0; w < WPL; w += 4) *reinterpret_cast<uint4*>(p + w) = make_uint4(0u, 0u, 0u, 0u);
    } else {
#pragma unroll
        for (int w = 0; w < WPL; w += 2) *reinterpret_cast<uint2*>(p + w) = make_uint2(0u, 0u);
    }
}

template <int MAXT, int SUBS>
__global__ void __launch_bounds__(bx_max_threads(MAXT, SUBS), 1)
bm25x_scan_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post,
                  const uint32_t* __restrict__ sub_off, const int32_t* __restrict__ sub_row,
                  const int32_t* __restrict__ bm_row,
                  const uint32_t* __restrict__ bits, const uint32_t* __restrict__ bits_hi,
                  const float2* __restrict__ wrange, const float* __restrict__ tau, int64_t V, int64_t n_docs,
                  const int32_t* __restrict__ q_ptr, const int32_t* __restrict__ q_terms,
                  const double* __restrict__ q_vals, int B, const uint32_t* __restrict__ mask, int k_sel, int lcap,
                  int nw, uint64_t* __restrict__ part, int splits, int seed_tiles, int seed_qpi, int seed_mode, unsigned mgr_sleep_ns,
                  int* __restrict__ work_counter,
                  uint32_t* __restrict__ gthr, uint32_t* __restrict__ coarse, uint32_t* __restrict__ fine,
                  const ShardBounds xb) {
    using C = BxCfg<MAXT, SUBS>;
    constexpr int WPL = C::WPL, WORDS = C::WORDS, POOL = C::POOL, PLANES = C::PLANES;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ int s_item;
    __shared__ int s_done;
    if (threadIdx.x == 0) s_done = 0;
    __syncthreads();

    if (warp == nw) {
        // ===== bound manager (last warp): queries blockIdx.x, blockIdx.x + gridDim.x, ...
        if (seed_mode) return;                             // a seed launch publishes its bounds itself
        while (true) {
            const int done = *reinterpret_cast<volatile int*>(&s_done);
            for (int q = (int)blockIdx.x; q < B; q += (int)gridDim.x) {
                publish_bound(gthr, q, hist_bound(coarse, fine, q, (uint32_t)k_sel, lane), xb, lane);
            }
            if (done >= nw) break;
            __nanosleep(mgr_sleep_ns);
        }
        return;
    }

    uint8_t* slab = smem_raw + (size_t)warp * bx_warp_bytes(MAXT, SUBS, lcap);
    uint32_t* bm = reinterpret_cast<uint32_t*>(slab);                                 // [MAXT][WORDS] present maps
    uint32_t* bmh = bm + MAXT * WORDS;                                                // [MAXT][WORDS] high maps
    // rank table: term l's ranks (u16) live in the first half of bmh[l], written after bmh[l] has been counted
    uint16_t* pre = reinterpret_cast<uint16_t*>(bmh);                                 // stride 2 * WORDS u16 per term
    uint2* pool = reinterpret_cast<uint2*>(slab + (size_t)MAXT * WORDS * 8);          // [POOL]
    uint64_t* list = reinterpret_cast<uint64_t*>(pool + POOL);
    int64_t* sbase = reinterpret_cast<int64_t*>(list + lcap);
    double* sqv = reinterpret_cast<double*>(sbase + MAXT);
    int* spoff = reinterpret_cast<int*>(sqv + MAXT);
    int* spcnt = spoff + MAXT;
    uint16_t* clist = reinterpret_cast<uint16_t*>(spcnt + MAXT);
    int* cntp = reinterpret_cast<int*>(clist + 32);
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int ns1 = ns + 1;
    const int nt = (ns + SUBS - 1) / SUBS;                    // tiles of SUBS sub-tiles
    // Seed launch (seed_mode): a SAMPLE that only produces a starting bound.  Work item = seed_qpi queries; seed_tiles
    // consecutive warps share a query and visit one tile each (tiles spread over the corpus), scoring just the
    // ~k_sel documents with the highest (terms present + terms in their high class) count -- the likely best of the
    // tile.  The k_sel-th best score of the group's lists together is a valid lower bound of the query's k_sel-th best
    // overall (those documents exist), published straight into gthr; no list is written and no histogram is touched, because the
    // main launch scans every tile, the sampled ones included.  Main launch: tiles [0, nt) in `splits` ranges, work
    // item = (range, query, slot w) claimed by one warp, which takes every nw-th tile of the range starting at the
    // w-th and writes list range * nw + w.  The seed's bound keeps the main launch's first visits from scoring every
    // document that shares a term with the query.
    const int nt_per_split = (nt + splits - 1) / splits;
    const size_t row_words = (size_t)((ns + 3) / 4 * 4) * BX_WORDS;
    const int n_items = seed_mode ? (B + seed_qpi - 1) / seed_qpi : splits * B;
    const double sentinel = __longlong_as_double((long long)BX_SENTINEL);

    while (true) {
        // Seed launch: the CTA's warps share one item (seed_qpi queries).  Main launch: every warp claims its own
        // (range, query, warp slot) -- the same decomposition as a CTA-wide (range, query) item, without the two
        // barriers per item and without the warps of an item waiting for its slowest one.
        int item, wslot;
        if (seed_mode) {
            worker_bar(nw * 32);
            if (threadIdx.x == 0) s_item = atomicAdd(work_counter, 1);
            worker_bar(nw * 32);
            item = s_item;
            wslot = warp;
            if (item >= n_items) break;
        } else {
            int wi = 0;
            if (lane == 0) wi = atomicAdd(work_counter, 1);
            wi = __shfl_sync(0xFFFFFFFFu, wi, 0);
            item = wi / nw;
            wslot = wi - item * nw;
            if (item >= n_items) break;
        }
        int q, j_first, j_end, j_stride, slot;
        if (seed_mode) {
            // seed_tiles consecutive warps share a query, one sampled tile each (the tiles are evenly spaced and
            // shifted per query so a batch does not pile on one tile); the first of them combines the lists below
            const int qi = warp / seed_tiles, ti = warp - qi * seed_tiles;
            q = item * seed_qpi + qi;
            j_stride = max(1, nt / seed_tiles);
            slot = 0;
            if (qi >= seed_qpi || q >= B) continue;        // idle warps of a seed item (whole groups: no partner waits)
            j_first = (q * 7) % j_stride + ti * j_stride;
            j_end = min(nt, j_first + 1);
        } else {
            const int split = item / B;
            q = item - split * B;
            const int j_begin = split * nt_per_split;
            j_first = j_begin + wslot; j_stride = nw; slot = split * nw + wslot;
            j_end = min(nt, j_begin + nt_per_split);
        }
        const int qb = q_ptr[q];
        const int nq = min(q_ptr[q + 1] - qb, MAXT);      // the launcher guarantees nq <= MAXT
        uint64_t thr = 0ull;
        double pre_local = -INFINITY;
        if (lane == 0) *cntp = 0;

        // ---- per-query term setup: lane l < nq owns term l
        int t0 = -1, row0 = -1, srow0 = -1;
        double qv0 = 0.0, ua = 0.0, ue = 0.0;     // bound of a low-class posting; extra of a high-class one
        float tau0 = 0.f;
        int64_t tp0 = 0;
        uint32_t df0 = 0, cur0 = 0;               // term without an offset row: list length, end of the last search
        if (lane < nq) {
            const int t = q_terms[qb + lane];
            if (t >= 0 && t < V) {
                t0 = t;
                qv0 = q_vals[qb + lane];
                tp0 = __ldg(term_ptr + t);
                row0 = __ldg(bm_row + t);
                srow0 = __ldg(sub_row + t);
                df0 = (uint32_t)(__ldg(term_ptr + t + 1) - tp0);
                tau0 = __ldg(tau + t);
                const float2 wr = __ldg(wrange + t);
                // low class: w in [min, tau]; high class: w in (tau, max]
                ua = fmax(0.0, fmax(qv0 * (double)wr.x, qv0 * (double)tau0));
                const double uh = fmax(ua, qv0 * (double)wr.y);
                ue = uh - ua;
                if (!(ua == ua) || !(ue == ue)) { ua = INFINITY; ue = 0.0; }   // NaN query value: never prune on this term
                ua *= 1.0 + 1e-12;                         // slack for the rounding of the float64 sums
                ue *= 1.0 + 1e-12;
            }
        }
        if (lane < MAXT) sqv[lane] = qv0;
        // The 2*nq bounds {ua_l} U {ue_l} sorted descending and prefix-summed give S_c; a document with c1 terms
        // present, c2 of them high, scores at most S_(c1+c2).  sa / se = S at the rank of this lane's ua / ue
        // (ties broken by kind, then lane).
        double sa = 0.0, se = 0.0;
#pragma unroll 1
        for (int o = 0; o < nq; ++o) {
            const double ao = __shfl_sync(0xFFFFFFFFu, ua, o);
            const double eo = __shfl_sync(0xFFFFFFFFu, ue, o);
            if (ao > ua || (ao == ua && o <= lane)) sa += ao;
            if (eo > ua) sa += eo;
            if (ao >= ue) se += ao;
            if (eo > ue || (eo == ue && o <= lane)) se += eo;
        }
        sa *= 1.0 + 1e-12;
        se *= 1.0 + 1e-12;
        const unsigned sparse_terms = __ballot_sync(0xFFFFFFFFu, t0 >= 0 && row0 < 0);
        const uint32_t* so = sub_off + (size_t)max(srow0, 0) * ns1;     // this lane's row of the offset table
        // posting range of this lane's term inside sub-tiles [s_lo, s_hi): table row, or (rare terms, bm25_common.cuh)
        // a forward search of the short list
        auto range_of = [&](int s_lo, int s_hi, uint32_t& lo, uint32_t& hi) {
            if (srow0 >= 0) {
                lo = __ldg(so + s_lo);
                hi = __ldg(so + s_hi);
            } else {
                lo = list_lower_bound(post + tp0, cur0, df0, (uint32_t)s_lo * BM25_SUB);
                hi = list_lower_bound(post + tp0, lo, df0, (uint32_t)min((int64_t)s_hi * BM25_SUB, (int64_t)0xFFFFFFFFll));
                cur0 = hi;
            }
        };
        __syncwarp();

        uint32_t lo_nx = 0, hi_nx = 0;
        if (t0 >= 0 && j_first < j_end) range_of(j_first * SUBS, min(ns, j_first * SUBS + SUBS), lo_nx, hi_nx);
        uint32_t g_seen = 0u;
        double pre_g = -INFINITY;
        // one scored-but-not-yet-summed candidate per lane, carried from a scoring round to the next one
        float pwt[MAXT];
        unsigned phit = 0u;
        uint32_t pdg = 0u;
#pragma unroll
        for (int l = 0; l < MAXT; ++l) pwt[l] = 0.f;
        auto flush_pending = [&](double pre_now, uint32_t g_now) {
            if (phit != 0u) {
                double acc = sentinel;
#pragma unroll
                for (int l = 0; l < MAXT; ++l)
                    if ((phit >> l) & 1u) acc = __dadd_rn(acc, __dmul_rn(sqv[l], (double)pwt[l]));
                if (acc > pre_now) {
                    const uint64_t key = make_key((float)acc, pdg);
                    if (key > thr && (uint32_t)(key >> 32) >= g_now) {
                        list[atomicAdd(cntp, 1)] = key;
                        if (!seed_mode) hist_count(coarse, fine, q, (uint32_t)(key >> 32));
                    }
                }
                phit = 0u;
            }
            __syncwarp();
        };
#pragma unroll 1
        for (int j = j_first; j < j_end; j += j_stride) {
            const uint32_t tile_base = (uint32_t)j * C::TILE;
            const uint32_t g = __ldcg(gthr + q);
            if (g != g_seen) { g_seen = g; pre_g = prefilter_of_score(ordered_f32(g)); }
            double pre_eff = fmax(pre_local, pre_g);
            const uint32_t lo_cur = lo_nx, hi_cur = hi_nx;
            if (t0 >= 0 && j + j_stride < j_end)
                range_of((j + j_stride) * SUBS, min(ns, (j + j_stride) * SUBS + SUBS), lo_nx, hi_nx);
            const uint32_t n = (t0 >= 0) ? hi_cur - lo_cur : 0u;
            const unsigned active = __ballot_sync(0xFFFFFFFFu, n != 0u);
            // a document needs (terms present) + (terms in their high class) >= m to beat the bound
            const int m = 1 + __popc(__ballot_sync(0xFFFFFFFFu, lane < nq && sa <= pre_eff)) +
                          __popc(__ballot_sync(0xFFFFFFFFu, lane < nq && se <= pre_eff));
            if (2 * __popc(active) < m) continue;

            // ---- phase 1: the (term, tile) document maps into shared memory.  Stored maps and the postings of
            //      map-less terms are all fetched with cp.async before anything waits: one memory round trip
            const int64_t base_cur = tp0 + lo_cur;
            const bool is_sp = n != 0u && row0 < 0;
            int poff = is_sp ? (int)n : 0;                       // -> exclusive prefix: slot of this term in the pool
#pragma unroll
            for (int off = 1; off < MAXT; off <<= 1) {
                const int o = __shfl_up_sync(0xFFFFFFFFu, poff, off);
                if (lane >= off) poff += o;
            }
            poff -= is_sp ? (int)n : 0;
            const int pcnt = is_sp ? max(0, min((int)n, POOL - poff)) : 0;
            if (n != 0u) { sbase[lane] = base_cur; spoff[lane] = poff; spcnt[lane] = pcnt; }
            for (unsigned a = active; a; a &= a - 1) {
                const int l = __ffs(a) - 1;
                const int row = __shfl_sync(0xFFFFFFFFu, row0, l);
                uint32_t* dst = bm + l * WORDS + WPL * lane;
                uint32_t* dsth = bmh + l * WORDS + WPL * lane;
                if (row >= 0) {
                    const size_t at = (size_t)row * row_words + (size_t)j * WORDS + WPL * lane;
                    if (SUBS == 1) { cp_async<8>(dst, bits + at); cp_async<8>(dsth, bits_hi + at); }
                    else {
#pragma unroll
                        for (int w = 0; w < WPL; w += 4) { cp_async<16>(dst + w, bits + at + w); cp_async<16>(dsth + w, bits_hi + at + w); }
                    }
                } else {
                    sts_zero_words<WPL>(dst);
                    sts_zero_words<WPL>(dsth);
                    const int c_l = __shfl_sync(0xFFFFFFFFu, pcnt, l);
                    uint2* pd = pool + __shfl_sync(0xFFFFFFFFu, poff, l);
                    const uint2* src = post + __shfl_sync(0xFFFFFFFFu, base_cur, l);
                    for (int i = lane; i < c_l; i += 32) cp_async<8>(pd + i, src + i);
                }
            }
            cp_async_wait_all();
            __syncwarp();
            for (unsigned a = active & sparse_terms; a; a &= a - 1) {
                const int l = __ffs(a) - 1;
                const int n_l = (int)__shfl_sync(0xFFFFFFFFu, n, l);
                const int c_l = __shfl_sync(0xFFFFFFFFu, pcnt, l);
                const uint2* pd = pool + __shfl_sync(0xFFFFFFFFu, poff, l);
                const float tau_l = __shfl_sync(0xFFFFFFFFu, tau0, l);
                for (int i = lane; i < c_l; i += 32) {
                    const uint2 pp = pd[i];
                    const uint32_t dl = pp.x - tile_base;
                    atomicOr(bm + l * WORDS + (dl >> 5), 1u << (dl & 31));
                    if (__uint_as_float(pp.y) > tau_l) atomicOr(bmh + l * WORDS + (dl >> 5), 1u << (dl & 31));
                }
                if (n_l > c_l) {                              // rare: the pool is full, read the rest from global
                    const uint2* src = post + __shfl_sync(0xFFFFFFFFu, base_cur, l);
                    for (int i = c_l + lane; i < n_l; i += 32) {
                        const uint2 pp = __ldg(&src[i]);
                        const uint32_t dl = pp.x - tile_base;
                        atomicOr(bm + l * WORDS + (dl >> 5), 1u << (dl & 31));
                        if (__uint_as_float(pp.y) > tau_l) atomicOr(bmh + l * WORDS + (dl >> 5), 1u << (dl & 31));
                    }
                }
            }
            __syncwarp();

            // ---- phase 2: bit-sliced count of the maps (lane owns words WPL*lane ..), two terms per step:
            //      count += present_a + high_a + present_b + high_b -- two full adders on the ones plane, one on the
            //      twos plane, then the single carry ripples up through half adders
            uint32_t cn[PLANES][WPL];
#pragma unroll
            for (int p = 0; p < PLANES; ++p)
#pragma unroll
                for (int w = 0; w < WPL; ++w) cn[p][w] = 0u;
            for (unsigned a = active; a;) {
                const int la = __ffs(a) - 1;
                a &= a - 1;
                const bool two = a != 0u;
                const int lb = two ? __ffs(a) - 1 : la;
                a &= a - 1;
                uint32_t wa[WPL], ha[WPL], wb[WPL], hb[WPL];
                lds_words<WPL>(bm + la * WORDS + WPL * lane, wa);
                lds_words<WPL>(bm + lb * WORDS + WPL * lane, wb);
                lds_words<WPL>(bmh + la * WORDS + WPL * lane, ha);
                lds_words<WPL>(bmh + lb * WORDS + WPL * lane, hb);
#pragma unroll
                for (int w = 0; w < WPL; ++w) {
                    const uint32_t wbw = two ? wb[w] : 0u, hbw = two ? hb[w] : 0u;
                    uint32_t c = cn[0][w];
                    const uint32_t ta = (c & wa[w]) | (c & ha[w]) | (wa[w] & ha[w]);
                    c = c ^ wa[w] ^ ha[w];
                    const uint32_t tb = (c & wbw) | (c & hbw) | (wbw & hbw);
                    cn[0][w] = c ^ wbw ^ hbw;
                    const uint32_t d = cn[1][w];
                    uint32_t k = (d & ta) | (d & tb) | (ta & tb);
                    cn[1][w] = d ^ ta ^ tb;
#pragma unroll
                    for (int p = 2; p < PLANES; ++p) { const uint32_t t_ = cn[p][w] & k; cn[p][w] ^= k; k = t_; }
                }
            }
            // count >= m, bit-slice compare from the top plane down, under the payload filter
            uint32_t keep[WPL];
#pragma unroll
            for (int w = 0; w < WPL; ++w) keep[w] = 0xFFFFFFFFu;
            if (mask != nullptr) {
#pragma unroll
                for (int w = 0; w < WPL; w += 2) {
                    const uint2 mw = __ldg(reinterpret_cast<const uint2*>(mask + (tile_base >> 5) + WPL * lane + w));
                    keep[w] = mw.x; keep[w + 1] = mw.y;
                }
            }
            uint32_t cand[WPL];
            auto count_ge = [&](int mm) {
                int c = 0;
#pragma unroll
                for (int w = 0; w < WPL; ++w) {
                    uint32_t gt = 0u, eq = 0xFFFFFFFFu;
#pragma unroll
                    for (int p = PLANES - 1; p >= 0; --p) {
                        if ((mm >> p) & 1) eq &= cn[p][w];
                        else { gt |= eq & cn[p][w]; eq &= ~cn[p][w]; }
                    }
                    cand[w] = ((mm >> PLANES) ? 0u : (gt | eq)) & keep[w];
                    c += __popc(cand[w]);
                }
                return c;
            };
            int own_c = count_ge(m);
            if (seed_mode && __reduce_add_sync(0xFFFFFFFFu, own_c) > k_sel) {
                // sample: the largest count threshold that still leaves k_sel documents of this tile
                int lo = m, hi = 2 * nq;
                while (lo < hi) {
                    const int mid = (lo + hi + 1) >> 1;
                    if (__reduce_add_sync(0xFFFFFFFFu, count_ge(mid)) >= k_sel) lo = mid; else hi = mid - 1;
                }
                own_c = count_ge(lo);
            }

            // ---- phase 3: score the survivors, 32 at a time (compacted through clist so every lane has one)
            int pos = own_c;                                  // -> exclusive prefix of own_c
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_up_sync(0xFFFFFFFFu, pos, off);
                if (lane >= off) pos += o;
            }
            const int total = __shfl_sync(0xFFFFFFFFu, pos, 31);
            pos -= own_c;
            if (total == 0) continue;                          // nothing of this tile can reach the bound
            // per-word ranks of the present maps (a posting's slot in its term's list = sub-tile offset + rank), only
            // for tiles with survivors; two terms share one warp scan (13-bit counts packed in the halves of a register).
            // The rank table overwrites the high maps, which every lane has finished reading.
            __syncwarp();
            for (unsigned a = active; a;) {
                const int la = __ffs(a) - 1;
                a &= a - 1;
                const bool two = a != 0u;
                const int lb = two ? __ffs(a) - 1 : la;
                a &= a - 1;
                uint32_t loc[WPL];                       // packed exclusive prefix inside the lane
                uint32_t own = 0u;
                uint32_t va[WPL], vb[WPL];
                lds_words<WPL>(bm + la * WORDS + WPL * lane, va);
                lds_words<WPL>(bm + lb * WORDS + WPL * lane, vb);
#pragma unroll
                for (int w = 0; w < WPL; ++w) {
                    loc[w] = own;
                    own += (uint32_t)__popc(va[w]) | (two ? (uint32_t)__popc(vb[w]) << 16 : 0u);
                }
                uint32_t inc = own;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, off);
                    if (lane >= off) inc += o;
                }
                const uint32_t ex = inc - own;
                uint32_t pa[WPL / 2], pb[WPL / 2];
#pragma unroll
                for (int w = 0; w < WPL; w += 2) {
                    const uint32_t r0 = ex + loc[w], r1 = ex + loc[w + 1];
                    pa[w / 2] = (r0 & 0xFFFFu) | (r1 << 16);
                    pb[w / 2] = (r0 >> 16) | (r1 & 0xFFFF0000u);
                }
                if constexpr (WPL % 4 == 0) {           // a lane's packed ranks are 8-byte aligned: one 64-bit store per 4 words
#pragma unroll
                    for (int w = 0; w < WPL; w += 4) {
                        *reinterpret_cast<uint2*>(pre + la * 2 * WORDS + WPL * lane + w) = make_uint2(pa[w / 2], pa[w / 2 + 1]);
                        if (two) *reinterpret_cast<uint2*>(pre + lb * 2 * WORDS + WPL * lane + w) = make_uint2(pb[w / 2], pb[w / 2 + 1]);
                    }
                } else {
#pragma unroll
                    for (int w = 0; w < WPL; w += 2) {
                        *reinterpret_cast<uint32_t*>(pre + la * 2 * WORDS + WPL * lane + w) = pa[w / 2];
                        if (two) *reinterpret_cast<uint32_t*>(pre + lb * 2 * WORDS + WPL * lane + w) = pb[w / 2];
                    }
                }
            }
#pragma unroll 1
            for (int r0 = 0; r0 < total; r0 += 32) {
                if (*cntp > lcap - 32) {
                    warp_compact(list, cntp, lcap, k_sel, thr, lane);
                    if (thr != 0ull) { pre_local = prefilter_of_score(key_score(thr)); pre_eff = fmax(pre_eff, pre_local); }
                }
                // this lane's candidates that fall in the window [r0, r0 + 32), in document order
#pragma unroll
                for (int w = 0; w < WPL; ++w) {
                    while (cand[w] != 0u && pos < r0 + 32) {
                        clist[pos - r0] = (uint16_t)(__ffs(cand[w]) - 1 + 32 * (WPL * lane + w));
                        cand[w] &= cand[w] - 1;
                        ++pos;
                    }
                }
                __syncwarp();
                const bool mine = r0 + lane < total;
                const int dl = mine ? (int)clist[lane] : 0;
                const int word = dl >> 5, bit = dl & 31;
                const uint32_t below = (1u << bit) - 1u;
                // the previous round's candidate (usually the previous visit's): its weight loads have had a whole
                // visit to land; sum and push it now, then start this round's loads into the same registers
                flush_pending(pre_eff, g);
                uint32_t rk[MAXT];
                unsigned hit = 0u;
#pragma unroll
                for (int l = 0; l < MAXT; ++l) {
                    const uint32_t wv = bm[l * WORDS + word];
                    const bool h = mine && ((active >> l) & 1u) && ((wv >> bit) & 1u);
                    rk[l] = (uint32_t)pre[l * 2 * WORDS + word] + __popc(wv & below);
                    const bool st = h && ((sparse_terms >> l) & 1u) && (int)rk[l] < spcnt[l];
                    const bool gl = h && !st;
                    // unconditional load (a dummy, L1-resident address when this term misses): the MAXT loads issue
                    // back to back instead of one dependent branch per term.  Generic loads: the posting is either in
                    // the global array or in this visit's shared-memory pool (read now, before the pool is refilled).
                    const uint2* ap = st ? pool + spoff[l] + rk[l] : (gl ? post + sbase[l] + rk[l] : post);
                    pwt[l] = __uint_as_float(ap->y);
                    hit |= (h ? 1u : 0u) << l;
                }
                phit = hit;
                pdg = tile_base + (uint32_t)dl;
                __syncwarp();
            }
        }
        // ---- the last round's candidates, then emit this warp's list for (split, warp, q)
        if (*cntp > lcap - 32) warp_compact(list, cntp, lcap, k_sel, thr, lane);
        flush_pending(fmax(pre_local, pre_g), g_seen);
        if (seed_mode) {
            warp_compact(list, cntp, lcap, k_sel, thr, lane);          // sorted, best first; *cntp <= k_sel entries
            if (seed_tiles == 1) {
                publish_bound(gthr, q, (uint32_t)(thr >> 32), xb, lane);   // thr: the k_sel-th best key, 0 if fewer were found
                continue;
            }
            // the k_sel-th best of the group's lists together: a key's rank in the union is its index in its own list
            // plus the number of greater keys in the others (binary searches; keys are distinct documents)
            const int qi = warp / seed_tiles, ti = warp - qi * seed_tiles;
            asm volatile("bar.sync %0, %1;" ::"r"(2 + qi), "r"(32 * seed_tiles) : "memory");
            if (ti == 0) {
                const size_t wb = bx_warp_bytes(MAXT, SUBS, lcap);
                int tot = 0;
                for (int t = 0; t < seed_tiles; ++t) tot += *reinterpret_cast<const int*>(reinterpret_cast<const uint8_t*>(cntp) + t * wb);
                uint32_t found = 0u;
                if (tot >= k_sel) {
                    for (int t = 0; t < seed_tiles; ++t) {
                        const uint64_t* L = reinterpret_cast<const uint64_t*>(reinterpret_cast<const uint8_t*>(list) + t * wb);
                        const int c = *reinterpret_cast<const int*>(reinterpret_cast<const uint8_t*>(cntp) + t * wb);
                        for (int i = lane; i < c; i += 32) {
                            const uint64_t x = L[i];
                            int r = i;
                            for (int t2 = 0; t2 < seed_tiles; ++t2) {
                                if (t2 == t) continue;
                                const uint64_t* L2 = reinterpret_cast<const uint64_t*>(reinterpret_cast<const uint8_t*>(list) + t2 * wb);
                                int lo = 0, hi = *reinterpret_cast<const int*>(reinterpret_cast<const uint8_t*>(cntp) + t2 * wb);
                                while (lo < hi) { const int mid = (lo + hi) >> 1; if (L2[mid] > x) lo = mid + 1; else hi = mid; }
                                r += lo;
                            }
                            if (r == k_sel - 1) found = (uint32_t)(x >> 32);
                        }
                    }
                }
                found = __reduce_max_sync(0xFFFFFFFFu, found);
                publish_bound(gthr, q, found, xb, lane);
            }
            continue;
        }
        // the merge that follows sorts whatever it gathers, so a list of <= k_sel keys leaves unsorted (under a warm
        // global bound an item keeps a dozen keys: the 256-key bitonic sort was most of an item's fixed cost)
        __syncwarp();
        if (*cntp > k_sel) warp_compact(list, cntp, lcap, k_sel, thr, lane);
        const int n_keep = *cntp;
        uint64_t* out = part + ((size_t)slot * B + q) * k_sel;
        for (int i = lane; i < k_sel; i += 32) out[i] = i < n_keep ? list[i] : 0ull;
        __syncwarp();
    }
    if (lane == 0) atomicAdd(&s_done, 1);             // the manager leaves once all nw worker warps are done
}

static int bx_lcap(int k_sel) {
    int lcap = 128;
    while (lcap < k_sel + 32) lcap <<= 1;     // k_sel kept after a compaction + one round of pushes (<= 32)
    return lcap;
}

static int bx_maxt(int max_terms) { return max_terms <= 8 ? 8 : (max_terms <= 16 ? 16 : 32); }

// pause of a CTA's bound-manager warp between sweeps of its queries' histograms (BRAG_BX_MGR_NS, tuning knob)
static unsigned bx_mgr_sleep() {
    static const unsigned ns = [] { const char* e = getenv("BRAG_BX_MGR_NS"); const int v = e ? atoi(e) : 4000; return (unsigned)(v < 0 ? 0 : v); }();
    return ns;
}

// sub-tiles per visit: 2 unless BRAG_BX_SUBS says otherwise (tuning knob, read once)
static int bx_subs() {
    static int subs = [] {
        const char* e = getenv("BRAG_BX_SUBS");
        const int v = e ? atoi(e) : 2;
        return (v == 1 || v == 2 || v == 4) ? v : 2;
    }();
    return subs;
}

bool bm25x_supported(int max_terms, int k_sel) { return max_terms >= 0 && max_terms <= 32 && k_sel <= 512; }

Bm25Plan bm25x_plan(int64_t n_docs, int B, int k_sel, int max_terms, int num_sms) {
    Bm25Plan p;
    const int subs = bx_subs();
    const int ns = (int)((n_docs + BM25_SUB - 1) / BM25_SUB);
    const int nt = (ns + subs - 1) / subs;
    p.workers = (int)std::min<size_t>(bx_worker_cap(bx_maxt(max_terms), subs), BX_SMEM_MAX / bx_warp_bytes(bx_maxt(max_terms), subs, bx_lcap(k_sel)));
    p.workers = std::max(1, std::min(p.workers, nt));
    // BRAG_BX_WORKERS (experiments): fewer worker warps per CTA, e.g. to run beside another resident kernel
    static const int wcap = [] { const char* e = getenv("BRAG_BX_WORKERS"); return e ? atoi(e) : 0; }();
    if (wcap > 0) p.workers = std::min(p.workers, wcap);
    const int resident = num_sms;
    // seed launch: BRAG_BX_SEED sampled tiles per query (default 2), when the corpus has tiles to spare
    static const int seed_cfg = [] { const char* e = getenv("BRAG_BX_SEED"); const int v = e ? atoi(e) : 2; return v < 0 ? 0 : (v > 4 ? 4 : v); }();
    p.seed_tiles = (nt >= 32) ? seed_cfg : 0;
    // queries per seed item: seed_tiles warps each
    p.seed_qpi = std::max(1, std::min(p.workers / std::max(1, p.seed_tiles), (B + resident - 1) / resident));
    if (p.seed_tiles > 1 && (p.workers < p.seed_tiles || p.seed_qpi > 13)) p.seed_tiles = 1;   // named barriers 2 .. 15
    const int nt_main = nt;
    // ranges per query: enough items to keep every SM busy ~8 times over (BRAG_BX_ITEMS_PER_SM overrides; tuning knob)
    static const int items_per_sm = [] { const char* e = getenv("BRAG_BX_ITEMS_PER_SM"); const int v = e ? atoi(e) : 8; return v < 1 ? 1 : (v > 64 ? 64 : v); }();
    int splits = (resident * items_per_sm + B - 1) / B;
    int max_splits = (nt_main + p.workers - 1) / p.workers;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > resident) splits = resident;     // batch 1: one split per SM
    p.splits = splits;
    p.visits_per_list = nt_main / (splits * p.workers);
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

struct BxArgs {
    const int64_t* term_ptr; const uint2* post; const uint32_t* sub_off; const int32_t* sub_row; const int32_t* bm_row; const uint32_t* bits;
    const uint32_t* bits_hi; const float2* wrange; const float* tau; int64_t V, n_docs; const int32_t *q_ptr, *q_terms; const double* q_vals; int B;
    const uint32_t* mask; int k_sel; uint64_t* part; const Bm25Plan* plan; void* scratch; cudaStream_t st;
    ShardBounds xb;
};

template <int MAXT, int SUBS>
static cudaError_t bx_launch(const BxArgs& a) {
    const Bm25Plan& plan = *a.plan;
    const int lcap = bx_lcap(a.k_sel);
    const size_t smem = bx_warp_bytes(MAXT, SUBS, lcap) * plan.workers;
    if (smem > BX_SMEM_MAX) return cudaErrorInvalidValue;
    uint8_t* sp = reinterpret_cast<uint8_t*>(a.scratch);
    int* work_counter = reinterpret_cast<int*>(sp);
    uint32_t* gthr = reinterpret_cast<uint32_t*>(sp + 256);
    uint32_t* coarse = reinterpret_cast<uint32_t*>(sp + 256 + plan.gthr_bytes);
    uint32_t* fine = reinterpret_cast<uint32_t*>(sp + 256 + plan.gthr_bytes + plan.coarse_bytes);
    cudaError_t e = cudaMemsetAsync(a.scratch, 0, plan.scratch_bytes, a.st);
    if (e != cudaSuccess) return e;
    if (plan.seed_tiles > 0) {
        const int seed_items = (a.B + plan.seed_qpi - 1) / plan.seed_qpi;
        bm25x_scan_kernel<MAXT, SUBS><<<std::min(seed_items, plan.grid), (plan.workers + 1) * 32, smem, a.st>>>(
            a.term_ptr, a.post, a.sub_off, a.sub_row, a.bm_row, a.bits, a.bits_hi, a.wrange, a.tau, a.V, a.n_docs, a.q_ptr, a.q_terms,
            a.q_vals, a.B, a.mask, a.k_sel, lcap, plan.workers, a.part, plan.splits, plan.seed_tiles, plan.seed_qpi, 1, bx_mgr_sleep(),
            work_counter + 1, gthr, coarse, fine, ShardBounds());
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    bm25x_scan_kernel<MAXT, SUBS><<<plan.grid, (plan.workers + 1) * 32, smem, a.st>>>(
        a.term_ptr, a.post, a.sub_off, a.sub_row, a.bm_row, a.bits, a.bits_hi, a.wrange, a.tau, a.V, a.n_docs, a.q_ptr, a.q_terms,
        a.q_vals, a.B, a.mask, a.k_sel, lcap, plan.workers, a.part, plan.splits, plan.seed_tiles, plan.seed_qpi, 0, bx_mgr_sleep(),
        work_counter, gthr, coarse, fine, a.xb);
    return cudaGetLastError();
}

template <int MAXT>
static cudaError_t bx_launch_subs(const BxArgs& a) {
    switch (bx_subs()) {
        case 1: return bx_launch<MAXT, 1>(a);
        case 4: return bx_launch<MAXT, 4>(a);
        default: return bx_launch<MAXT, 2>(a);
    }
}

// kernel attributes are per device: brag_create calls this with the engine's device current
template <int MAXT, int SUBS>
static cudaError_t bx_init_one() {
    return cudaFuncSetAttribute(bm25x_scan_kernel<MAXT, SUBS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BX_SMEM_MAX);
}
cudaError_t bm25x_init_device() {
    cudaError_t e = bx_init_one<8, 1>();
    if (e == cudaSuccess) e = bx_init_one<8, 2>();
    if (e == cudaSuccess) e = bx_init_one<8, 4>();
    if (e == cudaSuccess) e = bx_init_one<16, 1>();
    if (e == cudaSuccess) e = bx_init_one<16, 2>();
    if (e == cudaSuccess) e = bx_init_one<16, 4>();
    if (e == cudaSuccess) e = bx_init_one<32, 1>();
    if (e == cudaSuccess) e = bx_init_one<32, 2>();
    if (e == cudaSuccess) e = bx_init_one<32, 4>();
    return e;
}

cudaError_t bm25x_scan(const int64_t* term_ptr, const uint2* post, const uint32_t* sub_off, const int32_t* sub_row,
                       const int32_t* bm_row,
                       const uint32_t* bits, const uint32_t* bits_hi, const float2* wrange, const float* tau, int64_t V,
                       int64_t n_docs, const int32_t* q_ptr,
                       const int32_t* q_terms, const double* q_vals, int B, int max_terms, const uint32_t* mask,
                       int k_sel, uint64_t* part, const Bm25Plan& plan, void* scratch, cudaStream_t st,
                       const ShardBounds& xb) {
    if (!bm25x_supported(max_terms, k_sel)) return cudaErrorInvalidValue;
    const BxArgs a{term_ptr, post, sub_off, sub_row, bm_row, bits, bits_hi, wrange, tau, V, n_docs, q_ptr, q_terms, q_vals, B,
                   mask, k_sel, part, &plan, scratch, st, xb};
    switch (bx_maxt(max_terms)) {
        case 8: return bx_launch_subs<8>(a);
        case 16: return bx_launch_subs<16>(a);
        default: return bx_launch_subs<32>(a);
    }
}

}  // namespace brag
