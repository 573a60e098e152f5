// Selection-side kernels: k-way merge of partial (score,row) key lists, candidate re-scoring,
// reciprocal rank fusion, and the per-shard list merge that follows the all-gather.
// (e; reference requests qdrant_handler.py:317-329, :357-372.)
// All of these touch <= a few thousand entries per query: latency-bound, not roofed.
#include "common.cuh"
#include "kernels.h"
#include "rrf.cuh"
#include "rescore.cuh"
#include "shard_push.cuh"

namespace brag {

constexpr int SEL_THREADS = 512;
constexpr int SEL_MAX = 8192;  // keys per CTA sort (64 KB of shared memory)

// ------------------------------------------------------------------ merge of key lists
// bound (optional): per-query lower bound of the kout-th best key's score image (f32_ordered units, 0 = none); keys
// below it cannot be in the result, so they are dropped while loading and the sort runs over a few hundred keys
// instead of P * kk.
__global__ void __launch_bounds__(SEL_THREADS)
merge_keys_kernel(const uint64_t* __restrict__ in, int P, int B, int kk, int GP,
                  uint64_t* __restrict__ out, int kout, const uint32_t* __restrict__ bound) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t* a = reinterpret_cast<uint64_t*>(smem_raw);
    __shared__ int fill;
    const int b = blockIdx.x, pg = blockIdx.y, tid = threadIdx.x;
    const int p0 = pg * GP, np = min(GP, P - p0);
    const int n_in = np * kk;
    int n2;
    if (bound != nullptr) {
        const uint32_t bnd = bound[b];
        if (tid == 0) fill = 0;
        __syncthreads();
        for (int i = tid; i < n_in; i += SEL_THREADS) {
            const int p = p0 + i / kk, j = i % kk;
            const uint64_t v = in[((size_t)p * B + b) * kk + j];
            if (v != 0ull && (uint32_t)(v >> 32) >= bnd) a[atomicAdd(&fill, 1)] = v;
        }
        __syncthreads();
        const int f = fill;
        n2 = next_pow2(max(f, kout));
        for (int i = f + tid; i < n2; i += SEL_THREADS) a[i] = 0ull;
    } else {
        n2 = next_pow2(max(n_in, kout));
        for (int i = tid; i < n2; i += SEL_THREADS) {
            uint64_t v = 0ull;
            if (i < n_in) {
                int p = p0 + i / kk, j = i % kk;
                v = in[((size_t)p * B + b) * kk + j];
            }
            a[i] = v;
        }
    }
    __syncthreads();
    bitonic_desc(a, n2, tid, SEL_THREADS, [] { __syncthreads(); });
    uint64_t* o = out + ((size_t)pg * B + b) * kout;
    for (int i = tid; i < kout; i += SEL_THREADS) o[i] = a[i];
}

// ------------------------------------------------------------------ tail of a sparse scan, one CTA per query
// One launch where there were three (bound, merge, finalize): the bound of the query's complete histograms (published
// like every bound, also to the peers of a sharded call), the keys of the P per-warp lists that reach it -- a few
// hundred of the P * kk slots -- ordered by rank (keys are distinct: count the greater ones, no barriers), and either
// the best k as keys with local rows (keys_out: filter-compacted indexes and sharded calls go on from there) or the
// final list itself.  The lists' scores are already exact (float64 sums rounded once), so there is no re-score.
constexpr int STAIL_CAP = 2048;           // keys gathered per query (16 KB: four CTAs per SM)
constexpr int STAIL_RANK_MAX = 1024;

__global__ void __launch_bounds__(1024)
sparse_tail_kernel(const uint64_t* __restrict__ part, int P, int B, int kk, const uint32_t* __restrict__ coarse,
                   const uint32_t* __restrict__ fine, uint32_t* __restrict__ gthr, const ShardBounds xb, int k,
                   uint64_t* __restrict__ keys_out, int64_t id_offset, float* __restrict__ out_scores,
                   int64_t* __restrict__ out_ids, int32_t* __restrict__ out_counts, const ShardSink sink) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t* a = reinterpret_cast<uint64_t*>(smem_raw);
    __shared__ int fill, count;
    __shared__ uint32_t s_bound;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NT = (int)blockDim.x;
    if (tid == 0) { fill = 0; count = 0; }
    if (warp == 0) {
        publish_bound(gthr, b, hist_bound(coarse, fine, b, (uint32_t)k, lane), xb, lane);
        __syncwarp();
        if (lane == 0) s_bound = *reinterpret_cast<volatile uint32_t*>(gthr + b);
    }
    __syncthreads();
    const uint32_t bound = s_bound;
    const int total = P * kk;
    for (int i = tid; i < total; i += NT) {
        const int p = i / kk, j = i - p * kk;
        const uint64_t v = __ldcg(part + ((size_t)p * B + b) * kk + j);
        if (v != 0ull && (uint32_t)(v >> 32) >= bound) {
            const int slot = atomicAdd(&fill, 1);
            if (slot < STAIL_CAP) a[slot] = v;
        }
    }
    __syncthreads();
    if (fill > STAIL_CAP) {
        // more keys reach the bound than the buffer holds (mass ties): rounds of lists, keeping the best k in between
        __syncthreads();
        if (tid == 0) fill = 0;
        __syncthreads();
        const int R = max(1, (STAIL_CAP / 2) / kk);           // lists per round
        for (int p0 = 0; p0 < P; p0 += R) {
            if (fill + R * kk > STAIL_CAP) {                   // uniform: fill is stable here
                const int f = fill;
                const int n2 = next_pow2(max(f, 32));
                for (int i = f + tid; i < n2; i += NT) a[i] = 0ull;
                __syncthreads();
                bitonic_desc(a, n2, tid, NT, [] { __syncthreads(); });
                if (tid == 0) fill = min(f, k);
                __syncthreads();
            }
            const int np = min(R, P - p0);
            for (int i = tid; i < np * kk; i += NT) {
                const int p = p0 + i / kk, j = i % kk;
                const uint64_t v = __ldcg(part + ((size_t)p * B + b) * kk + j);
                if (v != 0ull && (uint32_t)(v >> 32) >= bound) a[atomicAdd(&fill, 1)] = v;
            }
            __syncthreads();
        }
    }
    const int f = fill;
    const bool ranked = f <= STAIL_RANK_MAX;
    if (!ranked) {
        const int n2 = next_pow2(max(f, 32));
        for (int i = f + tid; i < n2; i += NT) a[i] = 0ull;
        __syncthreads();
        bitonic_desc(a, n2, tid, NT, [] { __syncthreads(); });
    }
    const int n_out = min(f, k);
    if (sink.on) {
        for (int i = n_out + tid; i < k; i += NT) shard_sink_store(sink, B, b, k, i, 0ull, id_offset);
    } else if (keys_out != nullptr) {
        for (int i = n_out + tid; i < k; i += NT) keys_out[(size_t)b * k + i] = 0ull;
    } else {
        for (int i = n_out + tid; i < k; i += NT) { out_scores[(size_t)b * k + i] = -INFINITY; out_ids[(size_t)b * k + i] = -1; }
        if (tid == 0) out_counts[b] = n_out;
    }
    for (int i = tid; i < f; i += NT) {
        const uint64_t x = a[i];
        int r = i;
        if (ranked) {
            r = 0;
            for (int j = 0; j < f; ++j) r += a[j] > x ? 1 : 0;
        }
        if (r >= k) continue;
        if (sink.on) shard_sink_store(sink, B, b, k, r, x, id_offset);
        else if (keys_out != nullptr) keys_out[(size_t)b * k + r] = x;
        else {
            out_scores[(size_t)b * k + r] = key_score(x);
            out_ids[(size_t)b * k + r] = (int64_t)key_row(x) + id_offset;
        }
    }
}

cudaError_t sparse_tail(const uint64_t* part, int P, int B, int kk, const uint32_t* coarse, const uint32_t* fine,
                        uint32_t* gthr, const ShardBounds& xb, int k, uint64_t* keys_out, int64_t id_offset,
                        float* out_scores, int64_t* out_ids, int32_t* out_counts, const ShardSink& sink, cudaStream_t st) {
    if (kk > STAIL_CAP / 2 || k > STAIL_CAP / 2 || k > kk) return cudaErrorInvalidValue;
    sparse_tail_kernel<<<B, B <= 64 ? 1024 : 512, STAIL_CAP * 8, st>>>(part, P, B, kk, coarse, fine, gthr, xb, k, keys_out,
                                                                     id_offset, out_scores, out_ids, out_counts, sink);
    return cudaGetLastError();
}

size_t merge_tmp_elems(int P, int B, int kk) {
    int GP = max(2, SEL_MAX / kk);
    int PG = (P + GP - 1) / GP;
    return (size_t)PG * B * kk;
}

cudaError_t merge_key_lists(const uint64_t* in, int P, int B, int kk, uint64_t* tmp_a, uint64_t* tmp_b,
                            uint64_t* out, int kout, cudaStream_t st, int* launches, const uint32_t* bound) {
    if (kk > SEL_MAX / 2 || kout > SEL_MAX) return cudaErrorInvalidValue;
    const int GP = max(2, SEL_MAX / kk);
    const uint64_t* cur = in;
    uint64_t* bufs[2] = {tmp_a, tmp_b};
    int which = 0;
    while (true) {
        int PG = (P + GP - 1) / GP;
        bool last = (PG == 1);
        uint64_t* dst = last ? out : bufs[which];
        int ko = last ? kout : kk;
        int n_in = min(GP, P) * kk;
        int n2 = 1;
        while (n2 < max(n_in, ko)) n2 <<= 1;
        merge_keys_kernel<<<dim3(B, PG), SEL_THREADS, (size_t)n2 * 8, st>>>(cur, P, B, kk, GP, dst, ko, bound);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        if (launches) ++*launches;
        if (last) break;
        cur = dst;
        P = PG;
        which ^= 1;
    }
    return cudaSuccess;
}

// ------------------------------------------------------------------ scan copies of the queries
// The tensor-core scan reads the queries in the element type of the stored rows, zero padded to whole 128-query
// blocks: 16-bit roundings (f16 / bf16 storage), the float32 values as they are (tf32 reads their upper 19 bits), or
// int8 (int8 storage: the caller sends integer-valued queries, qdrant_handler.py:300-302) with the inverse norm of each
// query.  Only the scan sees this copy; candidates are re-scored from the caller's float32 queries.
__global__ void scan_queries_kernel(const float* __restrict__ Q, int B, int d, int dtype, void* __restrict__ out,
                                    int Bpad, float* __restrict__ qinv) {
    const int w = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (w >= Bpad) return;
    const float* q = Q + (size_t)w * d;
    int ss = 0;
    for (int j = lane; j < d; j += 32) {
        const float v = (w < B) ? q[j] : 0.0f;
        const size_t i = (size_t)w * d + j;
        if (dtype == 0) reinterpret_cast<float*>(out)[i] = v;
        else if (dtype == 1) reinterpret_cast<uint16_t*>(out)[i] = __half_as_ushort(__float2half_rn(v));
        else if (dtype == 2) reinterpret_cast<uint16_t*>(out)[i] = __bfloat16_as_ushort(__float2bfloat16_rn(v));
        else {
            const int iv = max(-128, min(127, __float2int_rn(v)));
            reinterpret_cast<int8_t*>(out)[i] = (int8_t)iv;
            ss += iv * iv;
        }
    }
    if (dtype == 3 && qinv != nullptr) {
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) ss += __shfl_xor_sync(0xFFFFFFFFu, ss, off);
        if (lane == 0 && w < B) qinv[w] = ss > 0 ? (float)(1.0 / sqrt((double)ss)) : 0.0f;
    }
}

cudaError_t scan_queries(const float* Q, int B, int d, int dtype, void* out, int Bpad, float* qinv, cudaStream_t st) {
    const int blocks = (Bpad * 32 + 255) / 256;
    scan_queries_kernel<<<blocks, 256, 0, st>>>(Q, B, d, dtype, out, Bpad, qinv);
    return cudaGetLastError();
}

// int8 spaces: inverse norm of every stored row, and the largest of them (positive floats order like their bits)
__global__ void i8_row_inv_norms_kernel(const int8_t* __restrict__ X, int64_t n, int d, float* __restrict__ xinv,
                                        uint32_t* __restrict__ max_bits) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n) return;
    const uint4* row = reinterpret_cast<const uint4*>(X + (size_t)w * d);
    int ss = 0;
    for (int c = lane; c < d / 16; c += 32) {
        const uint4 v = __ldg(row + c);
        ss = __dp4a((int)v.x, (int)v.x, ss); ss = __dp4a((int)v.y, (int)v.y, ss);
        ss = __dp4a((int)v.z, (int)v.z, ss); ss = __dp4a((int)v.w, (int)v.w, ss);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) ss += __shfl_xor_sync(0xFFFFFFFFu, ss, off);
    if (lane == 0) {
        const float inv = ss > 0 ? (float)(1.0 / sqrt((double)ss)) : 0.0f;
        xinv[w] = inv;
        atomicMax(max_bits, __float_as_uint(inv));
    }
}

cudaError_t i8_row_inv_norms(const void* X, int64_t n, int d, float* xinv, uint32_t* max_bits, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    const int blocks = (int)((n * 32 + 255) / 256);
    i8_row_inv_norms_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const int8_t*>(X), n, d, xinv, max_bits);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ re-scoring (f64 accumulate)
// (row_dot_any: rescore.cuh, shared with the dense scan's tail kernel)
__global__ void rescore_keys_kernel(const void* __restrict__ X, int dtype, int64_t n, int d,
                                    const float* __restrict__ Qs, int B, const uint64_t* __restrict__ keys_in,
                                    int m, uint64_t* __restrict__ keys_out) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= (int64_t)B * m) return;
    const int b = (int)(w / m);
    const uint64_t key = keys_in[w];
    uint64_t out = 0ull;
    if (key != 0ull) {
        const uint32_t row = key_row(key);
        if ((int64_t)row < n) {
            double s = row_dot_any(X, dtype, row, d, Qs + (size_t)b * d, lane);
            out = make_key((float)s, row);
        }
    }
    if (lane == 0) keys_out[w] = out;
}

cudaError_t rescore_keys(const void* X, int dtype, int64_t n, int d, const float* Qs, int B,
                         const uint64_t* keys_in, int m, uint64_t* keys_out, cudaStream_t st) {
    int64_t threads = (int64_t)B * m * 32;
    int blocks = (int)((threads + 255) / 256);
    rescore_keys_kernel<<<blocks, 256, 0, st>>>(X, dtype, n, d, Qs, B, keys_in, m, keys_out);
    return cudaGetLastError();
}

__global__ void rescore_cands_kernel(const void* __restrict__ X, int dtype, int64_t n, int d,
                                     const float* __restrict__ Qs, int B, const int64_t* __restrict__ cand,
                                     int m, int64_t id_offset, const uint32_t* __restrict__ mask,
                                     uint64_t* __restrict__ keys_out) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= (int64_t)B * m) return;
    const int b = (int)(w / m);
    const int64_t row = cand[w] < 0 ? -1 : cand[w] - id_offset;
    uint64_t out = 0ull;
    if (row >= 0 && row < n && mask_test(mask, (uint32_t)row)) {
        double s = row_dot_any(X, dtype, row, d, Qs + (size_t)b * d, lane);
        out = make_key((float)s, (uint32_t)row);
    }
    if (lane == 0) keys_out[w] = out;
}

cudaError_t rescore_cands(const void* X, int dtype, int64_t n, int d, const float* Qs, int B,
                          const int64_t* cand, int m, int64_t id_offset, const uint32_t* mask,
                          uint64_t* keys_out, cudaStream_t st) {
    int64_t threads = (int64_t)B * m * 32;
    int blocks = (int)((threads + 255) / 256);
    rescore_cands_kernel<<<blocks, 256, 0, st>>>(X, dtype, n, d, Qs, B, cand, m, id_offset, mask, keys_out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ finalize: sort, dedup, unpack
__global__ void __launch_bounds__(SEL_THREADS)
finalize_keys_kernel(const uint64_t* __restrict__ keys, int m, int k, int dedup, int64_t id_offset,
                     float* __restrict__ out_scores, int64_t* __restrict__ out_ids,
                     int32_t* __restrict__ out_counts) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t* a = reinterpret_cast<uint64_t*>(smem_raw);
    __shared__ int count;
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n2 = next_pow2(max(m, k));
    for (int i = tid; i < n2; i += SEL_THREADS) a[i] = (i < m) ? keys[(size_t)b * m + i] : 0ull;
    if (tid == 0) count = 0;
    __syncthreads();
    bitonic_desc(a, n2, tid, SEL_THREADS, [] { __syncthreads(); });
    if (dedup) {
        // equal keys == same row (and same score): keep the first copy
        uint64_t prev[SEL_MAX / SEL_THREADS];
        int t = 0;
        for (int i = tid; i < n2; i += SEL_THREADS, ++t) prev[t] = (i > 0) ? a[i - 1] : ~0ull;
        __syncthreads();
        t = 0;
        for (int i = tid; i < n2; i += SEL_THREADS, ++t) if (a[i] == prev[t]) a[i] = 0ull;
        __syncthreads();
        bitonic_desc(a, n2, tid, SEL_THREADS, [] { __syncthreads(); });
    }
    for (int i = tid; i < k; i += SEL_THREADS) {
        uint64_t key = a[i];
        bool ok = key != 0ull;
        out_scores[(size_t)b * k + i] = ok ? key_score(key) : -INFINITY;
        out_ids[(size_t)b * k + i] = ok ? (int64_t)key_row(key) + id_offset : -1;
        if (ok) atomicAdd(&count, 1);
    }
    __syncthreads();
    if (tid == 0) out_counts[b] = count;
}

cudaError_t finalize_keys(const uint64_t* keys, int B, int m, int k, bool dedup, int64_t id_offset,
                          float* out_scores, int64_t* out_ids, int32_t* out_counts, cudaStream_t st) {
    int n2 = 1;
    while (n2 < (m > k ? m : k)) n2 <<= 1;
    if (n2 > SEL_MAX) return cudaErrorInvalidValue;
    finalize_keys_kernel<<<B, SEL_THREADS, (size_t)n2 * 8, st>>>(keys, m, k, dedup ? 1 : 0, id_offset,
                                                               out_scores, out_ids, out_counts);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ reciprocal rank fusion
// (the fusion itself is rrf_cta in rrf.cuh, shared with the sharded fuse kernel of shard.cu)
__global__ void __launch_bounds__(RRF_THREADS)
rrf_kernel(const int64_t* __restrict__ ids, const int32_t* __restrict__ counts, int L, int B, int kmax,
           int limit, int tie_mode, double* __restrict__ out_scores, int64_t* __restrict__ out_ids,
           int32_t* __restrict__ out_counts) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    // layout: entries [n2] (16 B each) | flat ids [E] | flat pos [E]
    __shared__ int s_off[17];
    __shared__ int n_unique;
    const int b = blockIdx.x, tid = threadIdx.x;
    if (tid == 0) {
        int o = 0;
        for (int l = 0; l < L; ++l) { s_off[l] = o; o += min(counts[(size_t)l * B + b], kmax); }
        s_off[L] = o;
        n_unique = 0;
    }
    __syncthreads();
    const int E = s_off[L];
    const int n2 = next_pow2(max(E, max(limit, 1)));
    RrfEntry* ent = reinterpret_cast<RrfEntry*>(smem_raw);
    int64_t* fid = reinterpret_cast<int64_t*>(smem_raw + (size_t)n2 * sizeof(RrfEntry));
    int32_t* fpos = reinterpret_cast<int32_t*>(fid + E);
    for (int l = 0; l < L; ++l) {
        const int c = s_off[l + 1] - s_off[l];
        for (int j = tid; j < c; j += RRF_THREADS) {
            fid[s_off[l] + j] = ids[((size_t)l * B + b) * kmax + j];
            fpos[s_off[l] + j] = j;
        }
    }
    __syncthreads();
    rrf_cta(ent, fid, fpos, E, n2, limit, tie_mode, &n_unique, out_scores + (size_t)b * limit,
            out_ids + (size_t)b * limit, out_counts + b, tid, RRF_THREADS);
}

cudaError_t rrf_fuse(const int64_t* ids, const int32_t* counts, int L, int B, int kmax, int limit,
                     int tie_mode, double* out_scores, int64_t* out_ids, int32_t* out_counts,
                     cudaStream_t st) {
    if (L < 1 || L > 16 || (int64_t)L * kmax > RRF_MAX || limit > RRF_MAX) return cudaErrorInvalidValue;
    int E = L * kmax;
    int n2 = 1;
    while (n2 < (E > limit ? E : limit)) n2 <<= 1;
    size_t smem = (size_t)n2 * sizeof(RrfEntry) + (size_t)E * 12;
    rrf_kernel<<<B, RRF_THREADS, smem, st>>>(ids, counts, L, B, kmax, limit, tie_mode, out_scores, out_ids,
                                             out_counts);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ merge of per-shard lists
__global__ void __launch_bounds__(SEL_THREADS)
merge_scored_kernel(const float* __restrict__ scores, const int64_t* __restrict__ ids,
                    const int32_t* __restrict__ counts, int G, int B, int k, float* __restrict__ out_scores,
                    int64_t* __restrict__ out_ids, int32_t* __restrict__ out_counts) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t* a = reinterpret_cast<uint64_t*>(smem_raw);
    __shared__ int count;
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n_in = G * k;
    const int n2 = next_pow2(n_in);
    if (tid == 0) count = 0;
    for (int i = tid; i < n2; i += SEL_THREADS) {
        uint64_t v = 0ull;
        if (i < n_in) {
            const int g = i / k, j = i % k;
            if (j < counts[(size_t)g * B + b]) {
                const size_t src = ((size_t)g * B + b) * k + j;
                v = make_key(scores[src], (uint32_t)ids[src]);   // global ids < 2^32
            }
        }
        a[i] = v;
    }
    __syncthreads();
    bitonic_desc(a, n2, tid, SEL_THREADS, [] { __syncthreads(); });
    for (int i = tid; i < k; i += SEL_THREADS) {
        const uint64_t key = a[i];
        const bool ok = key != 0ull;
        out_scores[(size_t)b * k + i] = ok ? key_score(key) : -INFINITY;
        out_ids[(size_t)b * k + i] = ok ? (int64_t)key_row(key) : -1;
        if (ok) atomicAdd(&count, 1);
    }
    __syncthreads();
    if (tid == 0) out_counts[b] = count;
}

cudaError_t merge_scored_lists(const float* scores, const int64_t* ids, const int32_t* counts, int G,
                               int B, int k, float* out_scores, int64_t* out_ids, int32_t* out_counts,
                               cudaStream_t st) {
    int n2 = 1;
    while (n2 < G * k) n2 <<= 1;
    if (n2 > SEL_MAX) return cudaErrorInvalidValue;
    merge_scored_kernel<<<B, SEL_THREADS, (size_t)n2 * 8, st>>>(scores, ids, counts, G, B, k, out_scores,
                                                             out_ids, out_counts);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ re-score without any allocation
// brag_dense_rescore_noalloc (include/brag.h): the candidate re-score with its [B][m] key scratch in a file-scope array
// instead of the engine's grow-only cudaMalloc buffers, for callers that may not allocate (CUDA graph capture, tight
// serving loops).  512 KB on every GPU that loads the library, until exit.  Every call on a device shares it: calls
// that overlap on the GPU race.  Nothing has to start at zero -- rescore_cands_kernel writes every slot of its
// B * m keys (0 for an empty / filtered / out-of-range candidate) before finalize_keys_kernel reads exactly those.
static __device__ __align__(256) uint64_t g_noalloc_keys[NOALLOC_MAX_KEYS];

// In host code `g_noalloc_keys` is a host-side stand-in, not a device address: the kernels get the address
// cudaGetSymbolAddress returns for the CURRENT device (brag_create calls this once per engine, device current).
cudaError_t rescore_noalloc_init_device(uint64_t** keys) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_keys);
    if (e != cudaSuccess) return e;
    *keys = reinterpret_cast<uint64_t*>(p);
    return cudaSuccess;
}

// keys: what rescore_noalloc_init_device returned on this device.  Launches only; no allocation, no synchronisation.
cudaError_t dense_rescore_noalloc(uint64_t* keys, const void* X, int dtype, int64_t n, int d, const float* Q, int B,
                                  const int64_t* cand, int m, int k, int64_t id_offset, const uint32_t* mask,
                                  float* out_scores, int64_t* out_ids, int32_t* out_counts, cudaStream_t st) {
    if (keys == nullptr || (int64_t)B * m > NOALLOC_MAX_KEYS) return cudaErrorInvalidValue;
    int n2 = 1;
    while (n2 < (m > k ? m : k)) n2 <<= 1;
    if (n2 > SEL_MAX) return cudaErrorInvalidValue;
    const int64_t threads = (int64_t)B * m * 32;
    rescore_cands_kernel<<<(int)((threads + 255) / 256), 256, 0, st>>>(X, dtype, n, d, Q, B, cand, m, id_offset, mask, keys);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    finalize_keys_kernel<<<B, SEL_THREADS, (size_t)n2 * 8, st>>>(keys, m, k, 1, id_offset, out_scores, out_ids, out_counts);
    return cudaGetLastError();
}

// kernel attributes are per device: brag_create calls this with the engine's device current
cudaError_t select_init_device() {
    cudaError_t e = cudaFuncSetAttribute(merge_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SEL_MAX * 8);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(sparse_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, STAIL_CAP * 8);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(finalize_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SEL_MAX * 8);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(merge_scored_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SEL_MAX * 8);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(rrf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 RRF_MAX * (int)sizeof(RrfEntry) + RRF_MAX * 12);
    return e;
}

}  // namespace brag
