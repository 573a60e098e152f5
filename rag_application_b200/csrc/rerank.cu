// Rerank of the fused hits as the reference computes it: `rerank_documents`
// (app/core/models/huggingface/huggingface.py:172-189, called from qdrant_handler.py:388-412) mean-pools the encoder's
// last hidden states over ALL token positions (padding included) for the query and for every document, takes the dot
// products and returns `np.argsort(scores)[::-1]`.  The encoder itself is outside this path (no weights offline); this
// file is everything after it: hidden states in, document order out.
//
// One CTA per document: a thread owns columns c, c + 256, ...; for each it sums the column over the T positions in
// position order (coalesced across the CTA for every position), scales by 1/T, multiplies with the query's pooled
// column (recomputed per CTA: Tq * d floats, L2-resident) and the per-thread partial products are reduced in a fixed
// order.  A second single-CTA launch sorts (score, index) keys descending, i.e. score descending and, among equal
// scores, the higher index first, which is what reversing a stable ascending argsort gives.
#include "common.cuh"
#include "kernels.h"

namespace brag {

constexpr int RERANK_THREADS = 256;

__global__ void __launch_bounds__(RERANK_THREADS)
rerank_pool_score_kernel(const float* __restrict__ qh, int Tq, const float* __restrict__ dh, int T, int d,
                         float* __restrict__ scores) {
    __shared__ float s_part[RERANK_THREADS / 32];
    const float* doc = dh + (size_t)blockIdx.x * T * d;
    const float inv_tq = 1.0f / (float)Tq, inv_t = 1.0f / (float)T;
    float acc = 0.f;
    for (int c = threadIdx.x; c < d; c += RERANK_THREADS) {
        float qs = 0.f, ds = 0.f;
        for (int t = 0; t < Tq; ++t) qs += __ldg(qh + (size_t)t * d + c);
        for (int t = 0; t < T; ++t) ds += __ldcs(doc + (size_t)t * d + c);
        acc = fmaf(qs * inv_tq, ds * inv_t, acc);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int w = 0; w < RERANK_THREADS / 32; ++w) s += s_part[w];
        scores[blockIdx.x] = s;
    }
}

__global__ void __launch_bounds__(1024)
rerank_order_kernel(const float* __restrict__ scores, int n, int n2, int32_t* __restrict__ order) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t* keys = reinterpret_cast<uint64_t*>(smem_raw);
    for (int i = threadIdx.x; i < n2; i += blockDim.x)
        keys[i] = i < n ? ((((uint64_t)f32_ordered(scores[i])) << 32) | (uint64_t)(uint32_t)(i + 1)) : 0ull;
    __syncthreads();
    bitonic_desc(keys, n2, (int)threadIdx.x, (int)blockDim.x, [] { __syncthreads(); });
    for (int i = threadIdx.x; i < n; i += blockDim.x) order[i] = (int32_t)((uint32_t)keys[i]) - 1;
}

cudaError_t rerank_meanpool(const float* qh, int Tq, const float* dh, int n, int T, int d, float* scores,
                            int32_t* order, cudaStream_t st) {
    rerank_pool_score_kernel<<<n, RERANK_THREADS, 0, st>>>(qh, Tq, dh, T, d, scores);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    int n2 = 1;
    while (n2 < n) n2 <<= 1;
    rerank_order_kernel<<<1, 1024, (size_t)n2 * sizeof(uint64_t), st>>>(scores, n, n2, order);
    return cudaGetLastError();
}

}  // namespace brag
