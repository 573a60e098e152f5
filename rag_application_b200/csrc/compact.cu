// Filter-compacted copy of a dense space (SURVEY 8 row a5 at low selectivity, config C5: 10 % of the rows pass).
// The batched tcgen05 scan is tensor-bound and the batch-1 scan HBM-bound; both cost time proportional to the rows
// they touch, and a TMA tile cannot skip rows.  So for a (space, filter) pair whose filter keeps at most half of
// the rows, the passing rows are gathered once into a contiguous [nc, d] copy (HBM cost: 2 * nc * d * elt bytes)
// and every scan under that filter runs on the copy; candidates are mapped back through rowmap, which is
// increasing, so the (score desc, id asc) order of the keys is unchanged.
#include <algorithm>

#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "kernels.h"

namespace brag {

__global__ void mask_wordcount_kernel(const uint32_t* __restrict__ mask, int64_t n, int64_t nwords,
                                      uint32_t* __restrict__ cnt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > nwords) return;
    uint32_t c = 0u;
    if (i < nwords) {
        uint32_t w = mask[i];
        const int64_t rem = n - i * 32;
        if (rem < 32) w &= (rem <= 0) ? 0u : ((1u << rem) - 1u);
        c = __popc(w);
    }
    cnt[i] = c;          // cnt[nwords] = 0: after the exclusive scan it holds the total
}

__global__ void mask_rowmap_kernel(const uint32_t* __restrict__ mask, int64_t n, int64_t nwords,
                                   const uint32_t* __restrict__ prefix, uint32_t* __restrict__ rowmap) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nwords) return;
    uint32_t w = mask[i];
    const int64_t rem = n - i * 32;
    if (rem < 32) w &= (rem <= 0) ? 0u : ((1u << rem) - 1u);
    uint32_t at = prefix[i];
    while (w) {
        rowmap[at++] = (uint32_t)(i * 32) + (uint32_t)(__ffs(w) - 1);
        w &= w - 1;
    }
}

// one warp per destination row, 16-byte chunks
__global__ void gather_rows_kernel(const uint4* __restrict__ X, int chunks, const uint32_t* __restrict__ rowmap,
                                   int64_t nc, uint4* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < nc; r += nwarps) {
        const uint4* src = X + (size_t)rowmap[r] * chunks;
        uint4* dst = out + (size_t)r * chunks;
        for (int c = lane; c < chunks; c += 32) dst[c] = ldg_stream(src + c);
    }
}

__global__ void remap_keys_kernel(uint64_t* __restrict__ keys, int64_t count, const uint32_t* __restrict__ rowmap) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint64_t k = keys[i];
    if (k == 0ull) return;
    keys[i] = (k & 0xFFFFFFFF00000000ull) | (uint64_t)(0xFFFFFFFFu - rowmap[key_row(k)]);
}

size_t compact_tmp_bytes(int64_t nwords) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)(nwords + 1));
    return bytes;
}

// prefix: [nwords + 1]; prefix[w] = passing rows before word w, prefix[nwords] = passing rows among the first n
cudaError_t compact_prefix(const uint32_t* mask, int64_t n, uint32_t* prefix, void* tmp, size_t tmp_bytes,
                           cudaStream_t st) {
    const int64_t nwords = (n + 31) / 32;
    mask_wordcount_kernel<<<(unsigned)((nwords + 256) / 256), 256, 0, st>>>(mask, n, nwords, prefix);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, prefix, prefix, (int)(nwords + 1), st);
}

cudaError_t compact_rows(const void* X, int row_bytes, const uint32_t* mask, int64_t n, const uint32_t* prefix,
                         int64_t nc, uint32_t* rowmap, void* out, cudaStream_t st) {
    const int64_t nwords = (n + 31) / 32;
    if (nc == 0) return cudaSuccess;
    mask_rowmap_kernel<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(mask, n, nwords, prefix, rowmap);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const int64_t blocks = std::min<int64_t>((nc + 7) / 8, 148 * 32);
    gather_rows_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<const uint4*>(X), row_bytes / 16, rowmap, nc,
                                                       reinterpret_cast<uint4*>(out));
    return cudaGetLastError();
}

cudaError_t remap_keys(uint64_t* keys, int64_t count, const uint32_t* rowmap, cudaStream_t st) {
    if (count == 0) return cudaSuccess;
    remap_keys_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(keys, count, rowmap);
    return cudaGetLastError();
}

}  // namespace brag
