// Dense exact scan for small query batches (B <= 8): an HBM-streaming GEMV with the top-k
// selection fused into the scan (reference request
// qdrant_handler.py:311-315,335-339,363-369).
//
// Roofline: HBM.  Algorithmic bytes per query batch = N * d * sizeof(elt) (+ N/8 mask).
// Layout: X [N, d] row-major, 16-byte chunks; one warp owns a 32-row group (one mask word),
// lanes stride over the 16-byte chunks of a row, 4 rows in flight per step.  Scores never
// leave the SM: each CTA keeps a (score,row) candidate list per query in shared memory with
// a running k-th-best threshold and emits k_sel keys at the end.
#include "common.cuh"
#include "kernels.h"

namespace brag {

constexpr int GEMV_THREADS = 256;
constexpr int GEMV_WARPS = GEMV_THREADS / 32;
constexpr int GEMV_CAP = 1024;        // candidate slots per query per CTA
constexpr int GEMV_CHUNK_GROUPS = 2;  // 32-row groups per warp between compaction checks
constexpr int RPW = 4;                // rows in flight per warp step

template <int NV>
__device__ __forceinline__ void reduce_transpose(float (&v)[NV], int lane) {
    // After the call lane l holds the full sum of v[l >> (5 - log2 NV)] in v[0].
#pragma unroll
    for (int s = 0; s < 5; ++s) {
        const int off = 16 >> s;
        const int n = NV >> s;  // live values before this step
        if (n >= 2) {
            const int half = n / 2;
            const bool upper = (lane & off) != 0;
#pragma unroll
            for (int j = 0; j < half; ++j) {
                float send = upper ? v[j] : v[j + half];
                float keep = upper ? v[j + half] : v[j];
                v[j] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, off);
            }
        } else {
            v[0] += __shfl_xor_sync(0xFFFFFFFFu, v[0], off);
        }
    }
}

template <int DT> struct Chunk { static constexpr int EPC = 8; };
template <> struct Chunk<0> { static constexpr int EPC = 4; };

template <int DT>
__device__ __forceinline__ void unpack_chunk(const uint4& v, float* f) {
    if constexpr (DT == 0) {
        f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
        f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
    } else {
        Unpack<DT>::run(v, f);
    }
}

__device__ __forceinline__ void cta_compact(uint64_t* list, int* cnt, uint64_t* thr, int k_sel, int tid) {
    // list has GEMV_CAP slots; entries >= *cnt are stale -> zero them, sort, keep k_sel.
    int c = min(*cnt, GEMV_CAP);
    for (int i = c + tid; i < GEMV_CAP; i += GEMV_THREADS) list[i] = 0ull;
    __syncthreads();
    bitonic_desc(list, GEMV_CAP, tid, GEMV_THREADS, [] { __syncthreads(); });
    if (tid == 0) {
        int nc = min(c, k_sel);
        *cnt = nc;
        *thr = (nc == k_sel) ? list[k_sel - 1] : 0ull;
    }
    __syncthreads();
}

template <int DT, int QB, int CPL, bool QREG>
__global__ void __launch_bounds__(GEMV_THREADS, 2)
dense_gemv_kernel(const uint4* __restrict__ X, int64_t n, int d, const float* __restrict__ Q,
                  const uint32_t* __restrict__ mask, int k_sel, uint64_t* __restrict__ part,
                  int64_t groups_per_cta, int b_total, int b_off) {
    constexpr int EPC = Chunk<DT>::EPC;
    constexpr int NV = RPW * QB;
    constexpr int LOG2NV = (NV == 4) ? 2 : (NV == 8) ? 3 : (NV == 16) ? 4 : 5;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    float* qs = reinterpret_cast<float*>(smem_raw);
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem_raw + (((size_t)QB * d * 4 + 15) & ~(size_t)15));
    __shared__ int cnt[QB];
    __shared__ uint64_t thr[QB];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int C = d / EPC;  // 16-byte chunks per row
    for (int i = tid; i < QB * d; i += GEMV_THREADS) {
        int qi = i / d;
        qs[i] = (b_off + qi < b_total) ? Q[(size_t)(b_off + qi) * d + (i - qi * d)] : 0.0f;
    }
    if (tid < QB) { cnt[tid] = 0; thr[tid] = 0ull; }
    __syncthreads();

    float qreg[QREG ? CPL * EPC : 1];
    if constexpr (QREG) {
#pragma unroll
        for (int cc = 0; cc < CPL; ++cc) {
            int c = lane + cc * 32;
#pragma unroll
            for (int e = 0; e < EPC; ++e) qreg[cc * EPC + e] = (c < C) ? qs[c * EPC + e] : 0.0f;
        }
    }

    const int64_t ngroups = (n + 31) >> 5;
    const int64_t g_begin = (int64_t)blockIdx.x * groups_per_cta;
    const int64_t g_end = min(ngroups, g_begin + groups_per_cta);

    for (int64_t g0 = g_begin; g0 < g_end; g0 += GEMV_WARPS * GEMV_CHUNK_GROUPS) {
#pragma unroll 1
        for (int it = 0; it < GEMV_CHUNK_GROUPS; ++it) {
            const int64_t g = g0 + (int64_t)it * GEMV_WARPS + warp;
            if (g >= g_end) continue;
            uint32_t word = mask ? __ldg(mask + g) : 0xFFFFFFFFu;
            const int64_t rem = n - (g << 5);
            if (rem < 32) word &= (1u << (int)rem) - 1u;
            const int64_t row0 = g << 5;
            while (word) {
                int rb[RPW];
                bool rv[RPW];
#pragma unroll
                for (int r = 0; r < RPW; ++r) {
                    if (word) { rb[r] = __ffs(word) - 1; word &= word - 1; rv[r] = true; }
                    else { rb[r] = rb[0]; rv[r] = false; }
                }
                uint4 xv[CPL][RPW];
#pragma unroll
                for (int cc = 0; cc < CPL; ++cc) {
                    const int c = lane + cc * 32;
#pragma unroll
                    for (int r = 0; r < RPW; ++r) {
                        if (c < C) xv[cc][r] = ldg_stream(X + (size_t)(row0 + rb[r]) * C + c);
                        else xv[cc][r] = make_uint4(0, 0, 0, 0);
                    }
                }
                float acc[NV];
#pragma unroll
                for (int i = 0; i < NV; ++i) acc[i] = 0.0f;
#pragma unroll
                for (int cc = 0; cc < CPL; ++cc) {
                    const int c = lane + cc * 32;
                    float xf[RPW][EPC];
#pragma unroll
                    for (int r = 0; r < RPW; ++r) unpack_chunk<DT>(xv[cc][r], xf[r]);
#pragma unroll
                    for (int qi = 0; qi < QB; ++qi) {
                        float qv[EPC];
                        if constexpr (QREG) {
#pragma unroll
                            for (int e = 0; e < EPC; ++e) qv[e] = qreg[cc * EPC + e];
                        } else {
                            if (c < C) {
                                const float4* qp = reinterpret_cast<const float4*>(qs + (size_t)qi * d + c * EPC);
#pragma unroll
                                for (int e4 = 0; e4 < EPC / 4; ++e4) {
                                    float4 t = qp[e4];
                                    qv[e4 * 4 + 0] = t.x; qv[e4 * 4 + 1] = t.y; qv[e4 * 4 + 2] = t.z; qv[e4 * 4 + 3] = t.w;
                                }
                            } else {
#pragma unroll
                                for (int e = 0; e < EPC; ++e) qv[e] = 0.0f;
                            }
                        }
#pragma unroll
                        for (int r = 0; r < RPW; ++r)
#pragma unroll
                            for (int e = 0; e < EPC; ++e) acc[r * QB + qi] = fmaf(xf[r][e], qv[e], acc[r * QB + qi]);
                    }
                }
                reduce_transpose<NV>(acc, lane);
                const int idx = lane >> (5 - LOG2NV);
                const int r = idx / QB, qi = idx - r * QB;
                const int rbs = (r == 0) ? rb[0] : (r == 1) ? rb[1] : (r == 2) ? rb[2] : rb[3];
                const bool rvs = (r == 0) ? rv[0] : (r == 1) ? rv[1] : (r == 2) ? rv[2] : rv[3];
                if ((lane & ((32 >> LOG2NV) - 1)) == 0 && rvs && (b_off + qi < b_total)) {
                    const uint64_t key = make_key(acc[0], (uint32_t)(row0 + rbs));
                    if (key > thr[qi]) {
                        int pos = atomicAdd(&cnt[qi], 1);
                        if (pos < GEMV_CAP) lists[(size_t)qi * GEMV_CAP + pos] = key;
                    }
                }
            }
        }
        __syncthreads();
#pragma unroll 1
        for (int qi = 0; qi < QB; ++qi) {
            // at most 32 * GEMV_CHUNK_GROUPS * GEMV_WARPS = 512 pushes per query per outer step
            if (cnt[qi] > GEMV_CAP - 32 * GEMV_CHUNK_GROUPS * GEMV_WARPS)
                cta_compact(lists + (size_t)qi * GEMV_CAP, &cnt[qi], &thr[qi], k_sel, tid);
        }
    }
    __syncthreads();
#pragma unroll 1
    for (int qi = 0; qi < QB; ++qi) {
        if (b_off + qi >= b_total) break;
        cta_compact(lists + (size_t)qi * GEMV_CAP, &cnt[qi], &thr[qi], k_sel, tid);
        uint64_t* out = part + ((size_t)blockIdx.x * b_total + (b_off + qi)) * k_sel;
        for (int i = tid; i < k_sel; i += GEMV_THREADS) out[i] = lists[(size_t)qi * GEMV_CAP + i];
        __syncthreads();
    }
}

template <int DT, int QB, int CPL, bool QREG>
static cudaError_t launch_one(const void* X, int64_t n, int d, const float* Q, const uint32_t* mask,
                              int k_sel, uint64_t* part, int grid, int64_t gpc, int b_total, int b_off,
                              cudaStream_t st) {
    size_t smem = (((size_t)QB * d * 4 + 15) & ~(size_t)15) + (size_t)QB * GEMV_CAP * 8;
    auto kern = dense_gemv_kernel<DT, QB, CPL, QREG>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    kern<<<grid, GEMV_THREADS, smem, st>>>(reinterpret_cast<const uint4*>(X), n, d, Q, mask, k_sel, part, gpc,
                                           b_total, b_off);
    return cudaGetLastError();
}

template <int DT, int QB>
static cudaError_t launch_cpl(int cpl, const void* X, int64_t n, int d, const float* Q, const uint32_t* mask,
                              int k_sel, uint64_t* part, int grid, int64_t gpc, int b_total, int b_off,
                              cudaStream_t st) {
    constexpr bool QR = (QB == 1);
    if (cpl <= 1) return launch_one<DT, QB, 1, QR>(X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
    if (cpl <= 2) return launch_one<DT, QB, 2, QR>(X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
    if (cpl <= 3) return launch_one<DT, QB, 3, QR>(X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
    if (cpl <= 4) return launch_one<DT, QB, 4, QR>(X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
    if (cpl <= 8) return launch_one<DT, QB, 8, false>(X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
    return cudaErrorInvalidValue;
}

template <int DT>
static cudaError_t launch_qb(int qb, int cpl, const void* X, int64_t n, int d, const float* Q,
                             const uint32_t* mask, int k_sel, uint64_t* part, int grid, int64_t gpc,
                             int b_total, int b_off, cudaStream_t st) {
    switch (qb) {
        case 1: return launch_cpl<DT, 1>(cpl, X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
        case 2: return launch_cpl<DT, 2>(cpl, X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
        case 4: return launch_cpl<DT, 4>(cpl, X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
        default: return launch_cpl<DT, 8>(cpl, X, n, d, Q, mask, k_sel, part, grid, gpc, b_total, b_off, st);
    }
}

// part: [grid][B][k_sel] keys.  Queries are processed in groups of <= 8 (one launch each).
cudaError_t dense_gemv_scan(const void* X, int dtype, int64_t n, int d, const float* Q, int B,
                            const uint32_t* mask, int k_sel, uint64_t* part, int grid, cudaStream_t st,
                            int* launches) {
    const int epc = (dtype == 0) ? 4 : 8;
    const int C = d / epc;
    const int cpl = (C + 31) / 32;
    const int64_t ngroups = (n + 31) >> 5;
    // contiguous groups per CTA, rounded to the CTA step so compaction chunks stay aligned
    int64_t gpc = (ngroups + grid - 1) / grid;
    const int step = GEMV_WARPS * GEMV_CHUNK_GROUPS;
    gpc = (gpc + step - 1) / step * step;
    for (int b_off = 0; b_off < B; b_off += 8) {
        int rem = B - b_off;
        int qb = rem >= 8 ? 8 : rem > 2 ? (rem > 4 ? 8 : 4) : rem;
        cudaError_t err;
        if (dtype == 0) err = launch_qb<0>(qb, cpl, X, n, d, Q, mask, k_sel, part, grid, gpc, B, b_off, st);
        else if (dtype == 1) err = launch_qb<1>(qb, cpl, X, n, d, Q, mask, k_sel, part, grid, gpc, B, b_off, st);
        else err = launch_qb<2>(qb, cpl, X, n, d, Q, mask, k_sel, part, grid, gpc, B, b_off, st);
        if (err != cudaSuccess) return err;
        if (launches) ++*launches;
    }
    return cudaSuccess;
}

}  // namespace brag
