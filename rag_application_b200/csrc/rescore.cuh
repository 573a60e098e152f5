// Exact re-scoring of candidate rows ("fp32 re-scoring of candidates" of the north_star):
// float64 dot of a stored row with the caller's float32 query, one warp per row.  Shared by select.cu (candidate
// re-score requests) and dense_umma.cu (the tail of the dense scan).
#pragma once
#include "common.cuh"

namespace brag {

template <int DT>
__device__ __forceinline__ double row_dot_f64(const uint4* __restrict__ X, int64_t row, int d,
                                              const float* __restrict__ q, int lane) {
    constexpr int EPC = (DT == 0) ? 4 : 8;
    const int C = d / EPC;
    const uint4* rp = X + (size_t)row * C;
    const float4* q4 = reinterpret_cast<const float4*>(q);      // d * 4 bytes is a multiple of 16 and so is every row of Q
    double acc = 0.0;
    // four 16-byte chunks of the row in flight per lane (a candidate row is a cold, random read); the sum still runs
    // over this lane's chunks in ascending order
    for (int c0 = lane; c0 < C; c0 += 128) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = (c0 + 32 * u < C) ? __ldg(rp + c0 + 32 * u) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int c = c0 + 32 * u;
            if (c >= C) break;
            float f[8];
            if constexpr (DT == 0) {
                f[0] = __uint_as_float(v[u].x); f[1] = __uint_as_float(v[u].y); f[2] = __uint_as_float(v[u].z); f[3] = __uint_as_float(v[u].w);
                const float4 qa = __ldg(q4 + c);
                acc += (double)f[0] * (double)qa.x; acc += (double)f[1] * (double)qa.y;
                acc += (double)f[2] * (double)qa.z; acc += (double)f[3] * (double)qa.w;
            } else {
                Unpack<DT>::run(v[u], f);
                const float4 qa = __ldg(q4 + 2 * c), qb = __ldg(q4 + 2 * c + 1);
                acc += (double)f[0] * (double)qa.x; acc += (double)f[1] * (double)qa.y;
                acc += (double)f[2] * (double)qa.z; acc += (double)f[3] * (double)qa.w;
                acc += (double)f[4] * (double)qb.x; acc += (double)f[5] * (double)qb.y;
                acc += (double)f[6] * (double)qb.z; acc += (double)f[7] * (double)qb.w;
            }
        }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, off);
    return acc;
}

// int8 rows (cosine of integer vectors): exact integer dot and squared norms, then ONE float64 expression
// idot / (sqrt(ssx) * sqrt(ssq)) -- every operation correctly rounded, so the oracle's numpy float64 line gives the
// same bits.  A zero row or query scores 0.
__device__ __forceinline__ double row_dot_i8(const uint4* __restrict__ X, int64_t row, int d,
                                             const float* __restrict__ q, int lane) {
    const int C = d / 16;
    int dot = 0, ssx = 0, ssq = 0;
    for (int c = lane; c < C; c += 32) {
        const uint4 v = __ldg(X + (size_t)row * C + c);
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t qp = 0;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int qi = max(-128, min(127, __float2int_rn(q[c * 16 + j * 4 + e])));
                qp |= ((uint32_t)qi & 0xFFu) << (8 * e);
            }
            dot = __dp4a((int)w4[j], (int)qp, dot);
            ssx = __dp4a((int)w4[j], (int)w4[j], ssx);
            ssq = __dp4a((int)qp, (int)qp, ssq);
        }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        dot += __shfl_xor_sync(0xFFFFFFFFu, dot, off);
        ssx += __shfl_xor_sync(0xFFFFFFFFu, ssx, off);
        ssq += __shfl_xor_sync(0xFFFFFFFFu, ssq, off);
    }
    if (ssx == 0 || ssq == 0) return 0.0;
    return (double)dot / (sqrt((double)ssx) * sqrt((double)ssq));
}

__device__ __forceinline__ double row_dot_any(const void* X, int dtype, int64_t row, int d, const float* q,
                                              int lane) {
    const uint4* X4 = reinterpret_cast<const uint4*>(X);
    if (dtype == 0) return row_dot_f64<0>(X4, row, d, q, lane);
    if (dtype == 1) return row_dot_f64<1>(X4, row, d, q, lane);
    if (dtype == 3) return row_dot_i8(X4, row, d, q, lane);
    return row_dot_f64<2>(X4, row, d, q, lane);
}

}  // namespace brag
