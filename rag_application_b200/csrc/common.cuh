// Shared device helpers: sortable (score,id) keys, smem bitonic sort, vector loads.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace brag {

// ---------------------------------------------------------------- sortable keys
// key = ordered(score) << 32 | (0xFFFFFFFF - row).  Bigger key == better under
// (score desc, row asc).  key 0 is the empty slot.
__device__ __forceinline__ uint32_t f32_ordered(float f) {
    f += 0.0f;  // -0 -> +0 so equal scores compare equal
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_f32(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(u);
}
__device__ __forceinline__ uint64_t make_key(float s, uint32_t row) {
    return ((uint64_t)f32_ordered(s) << 32) | (uint64_t)(0xFFFFFFFFu - row);
}
__device__ __forceinline__ float key_score(uint64_t k) { return ordered_f32((uint32_t)(k >> 32)); }
__device__ __forceinline__ uint32_t key_row(uint64_t k) { return 0xFFFFFFFFu - (uint32_t)k; }

// ---------------------------------------------------------------- global score bounds
// Every scan keeps, per query, a histogram of the scores it has pushed as candidates, over the
// order-preserving uint32 image of the float score: coarse bin = top 7 bits, fine bin = top 14
// bits (sign + exponent + 5 mantissa bits, ~3 % wide).  Each counted entry is a distinct corpus
// row, so "at least k rows counted in bins >= e" proves that the k-th best score is >= edge(e):
// that edge is published as the query's bound (f32_ordered units, 0 = none) and candidates
// below it are dropped by every CTA.  The bound only ever rises.
constexpr int HIST_FINE_SHIFT = 18;
constexpr int HIST_FINE_BINS = 1 << (32 - HIST_FINE_SHIFT);        // 16384
constexpr int HIST_COARSE_SHIFT = 25;
constexpr int HIST_COARSE_BINS = 1 << (32 - HIST_COARSE_SHIFT);    // 128
constexpr int HIST_FINE_PER_COARSE = HIST_FINE_BINS / HIST_COARSE_BINS;  // 128

__device__ __forceinline__ void hist_count(uint32_t* __restrict__ coarse, uint32_t* __restrict__ fine, int q,
                                           uint32_t ord) {
    atomicAdd(fine + (size_t)q * HIST_FINE_BINS + (ord >> HIST_FINE_SHIFT), 1u);
    atomicAdd(coarse + (size_t)q * HIST_COARSE_BINS + (ord >> HIST_COARSE_SHIFT), 1u);
}

// inclusive suffix sum over lanes >= lane
__device__ __forceinline__ uint32_t warp_suffix_sum(uint32_t v, int lane) {
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_down_sync(0xFFFFFFFFu, v, off);
        if (lane + off < 32) v += o;
    }
    return v;
}

// Whole warp: highest histogram edge that at least k counted rows reach (0 = fewer than k counted).
// Counts are read with ld.cg while other CTAs keep adding, so the result is a valid, possibly
// slightly stale bound.
__device__ __forceinline__ uint32_t hist_bound(const uint32_t* __restrict__ coarse, const uint32_t* __restrict__ fine,
                                               int q, uint32_t k, int lane) {
    const uint4 c4 = __ldcg(reinterpret_cast<const uint4*>(coarse + (size_t)q * HIST_COARSE_BINS) + lane);
    const uint32_t own = c4.x + c4.y + c4.z + c4.w;     // lane l owns coarse bins 4l .. 4l+3
    const uint32_t suf = warp_suffix_sum(own, lane);
    const unsigned has = __ballot_sync(0xFFFFFFFFu, suf >= k);
    if (has == 0u) return 0u;
    const int L = 31 - __clz(has);
    const uint32_t al = __shfl_sync(0xFFFFFFFFu, suf - own, L);     // rows in lanes above L
    const uint32_t y = __shfl_sync(0xFFFFFFFFu, c4.y, L), z = __shfl_sync(0xFFFFFFFFu, c4.z, L),
                   w = __shfl_sync(0xFFFFFFFFu, c4.w, L);
    int cb;
    uint32_t above;                                                  // rows in coarse bins above cb
    if (al + w >= k) { cb = 4 * L + 3; above = al; }
    else if (al + w + z >= k) { cb = 4 * L + 2; above = al + w; }
    else if (al + w + z + y >= k) { cb = 4 * L + 1; above = al + w + z; }
    else { cb = 4 * L; above = al + w + z + y; }
    const uint4 f4 = __ldcg(reinterpret_cast<const uint4*>(fine + (size_t)q * HIST_FINE_BINS +
                                                          (size_t)cb * HIST_FINE_PER_COARSE) + lane);
    const uint32_t fown = f4.x + f4.y + f4.z + f4.w;
    const uint32_t fsuf = warp_suffix_sum(fown, lane);
    const unsigned fhas = __ballot_sync(0xFFFFFFFFu, above + fsuf >= k);
    if (fhas == 0u) return (uint32_t)cb << HIST_COARSE_SHIFT;        // fine counts lag: the coarse edge holds
    const int FL = 31 - __clz(fhas);
    const uint32_t fa = above + __shfl_sync(0xFFFFFFFFu, fsuf - fown, FL);
    const uint32_t fy = __shfl_sync(0xFFFFFFFFu, f4.y, FL), fz = __shfl_sync(0xFFFFFFFFu, f4.z, FL),
                   fw = __shfl_sync(0xFFFFFFFFu, f4.w, FL);
    int fb;
    if (fa + fw >= k) fb = 3;
    else if (fa + fw + fz >= k) fb = 2;
    else if (fa + fw + fz + fy >= k) fb = 1;
    else fb = 0;
    return (uint32_t)(cb * HIST_FINE_PER_COARSE + 4 * FL + fb) << HIST_FINE_SHIFT;
}

// ---------------------------------------------------------------- cross-shard bounds (row sharding, shard.cu)
// Any shard's k-th best score is a lower bound of the global k-th best, so the shards of one box tell each other
// their bounds while they scan: rank g owns one word per (list, query) inside its peer-mapped window,
// (epoch << 32 | bound); a bound manager that raises its query's bound above everything it knew pushes it into every
// peer's word with a system-scope atomicMax over NVLink, and folds what the peers pushed into its own gthr.  The epoch
// (one per sharded call, the same on every rank) keeps a word written for another batch from being used: a newer
// epoch replaces an older one under max, an older one never survives, and a reader ignores any epoch but its own.
struct ShardBounds {
    unsigned long long* const* peers = nullptr;   // device table [world]: rank g's words of this list (peers[rank] = own)
    const unsigned* epoch = nullptr;              // device word holding the current sharded call's epoch
    int world = 0, rank = 0;
};

// Whole warp, `bound` warp-uniform: publish a query's bound locally and, when sharded, to the peers.
__device__ __forceinline__ void publish_bound(uint32_t* __restrict__ gthr, int q, uint32_t bound, const ShardBounds& xb,
                                              int lane) {
    if (xb.world <= 1) {
        if (lane == 0 && bound != 0u) atomicMax(gthr + q, bound);
        return;
    }
    const uint32_t ep = *reinterpret_cast<const volatile unsigned*>(xb.epoch);
    uint32_t old = 0u, remote = 0u;
    if (lane == 0) {
        unsigned long long x;
        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(x) : "l"(xb.peers[xb.rank] + q) : "memory");
        if ((uint32_t)(x >> 32) == ep) remote = (uint32_t)x;
        const uint32_t best = max(bound, remote);
        if (best != 0u) old = atomicMax(gthr + q, best);
    }
    old = __shfl_sync(0xFFFFFFFFu, old, 0);
    remote = __shfl_sync(0xFFFFFFFFu, remote, 0);
    // only a bound that beats everything this rank knew is news to a peer
    if (bound > old && bound > remote && lane < xb.world && lane != xb.rank)
        atomicMax_system(xb.peers[lane] + q, ((unsigned long long)ep << 32) | (unsigned long long)bound);
}

// ---------------------------------------------------------------- term-major CSR walks
// term owning posting i of a term-major CSR: the last t with term_ptr[t] <= i (empty terms are skipped)
__device__ __forceinline__ int64_t term_of_posting(const int64_t* __restrict__ term_ptr, int64_t V, int64_t i) {
    int64_t lo = 0, hi = V;              // last t with term_ptr[t] <= i
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(term_ptr + mid) <= i) lo = mid; else hi = mid;
    }
    return lo;
}

// ---------------------------------------------------------------- bitonic sort (descending)
// n must be a power of two; all `nthreads` threads of the group call it; `sync` is
// __syncthreads for a CTA-wide sort or __syncwarp for a warp-private one.
template <typename T, typename SyncF>
__device__ __forceinline__ void bitonic_desc(T* a, int n, int tid, int nthreads, SyncF sync) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < (n >> 1); i += nthreads) {
                // i-th compare-exchange pair of this stage
                int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
                int hi = lo | j;
                bool desc = (lo & k) == 0;
                T x = a[lo], y = a[hi];
                if ((x < y) == desc) { a[lo] = y; a[hi] = x; }
            }
            sync();
        }
    }
}

__device__ __forceinline__ int next_pow2(int x) {
    int p = 1;
    while (p < x) p <<= 1;
    return p;
}

// ---------------------------------------------------------------- loads
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ bool mask_test(const uint32_t* __restrict__ mask, uint32_t row) {
    return mask == nullptr || ((__ldg(mask + (row >> 5)) >> (row & 31)) & 1u);
}

// unpack 8 storage elements (one 16-byte chunk) to float
template <int DT> struct Unpack;
template <> struct Unpack<1> {  // f16
    __device__ static __forceinline__ void run(const uint4& v, float* f) {
        const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
    }
};
template <> struct Unpack<2> {  // bf16
    __device__ static __forceinline__ void run(const uint4& v, float* f) {
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) { f[2 * i] = __uint_as_float(u[i] << 16); f[2 * i + 1] = __uint_as_float(u[i] & 0xFFFF0000u); }
    }
};

}  // namespace brag
