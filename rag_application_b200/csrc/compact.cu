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

// ---- filter-compacted copy of a sparse space ------------------------------------------------------------------
// The BM25 scans cost time proportional to the postings they walk, filtered out or not.  Under a filter that keeps
// at most half of the documents, the postings of the passing documents are selected once (stable, so every term's
// list stays sorted by document), renumbered to compact document ids, and indexed like any other space; scores are
// sums of the same weights in the same term order, so the results are bit-identical after the id remap.

// cnt[t] = postings of term t whose document passes the mask.  A warp walks chunks of KEEP_CHUNK consecutive
// postings, four coalesced loads per lane in flight; the term is searched once per chunk and followed from there, the
// count lives in registers and reaches cnt[] once per (chunk, term).  Only a 128-posting block that straddles a term
// boundary takes the per-posting search.
constexpr int KEEP_CHUNK = 8192;

__global__ void sparse_keep_count_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post,
                                         int64_t V, int64_t nnz, const uint32_t* __restrict__ mask,
                                         unsigned long long* __restrict__ cnt, int64_t* __restrict__ chunk_cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t nchunks = (nnz + KEEP_CHUNK - 1) / KEEP_CHUNK;
    for (int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < nchunks; c += nwarps) {
        const int64_t i0 = c * KEEP_CHUNK, i1 = min(nnz, i0 + (int64_t)KEEP_CHUNK);
        int64_t t = term_of_posting(term_ptr, V, i0);
        int64_t end = __ldg(term_ptr + t + 1);
        uint32_t local = 0u, ctot = 0u;
        for (int64_t base = i0; base < i1; base += 128) {
            uint32_t doc[4];
            bool keep[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t i = base + u * 32 + lane;
                doc[u] = i < i1 ? __ldcs(post + i).x : 0xFFFFFFFFu;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                keep[u] = doc[u] != 0xFFFFFFFFu && ((__ldg(mask + (doc[u] >> 5)) >> (doc[u] & 31)) & 1u);
            ctot += (uint32_t)keep[0] + (uint32_t)keep[1] + (uint32_t)keep[2] + (uint32_t)keep[3];
            if (base + 128 <= end) {
                local += (uint32_t)keep[0] + (uint32_t)keep[1] + (uint32_t)keep[2] + (uint32_t)keep[3];
                continue;
            }
            // a term ends inside this block: flush the running count, then one search per kept posting
            uint32_t tot = local;
#pragma unroll
            for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
            if (lane == 0 && tot) atomicAdd(cnt + t, (unsigned long long)tot);
            local = 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (keep[u]) atomicAdd(cnt + term_of_posting(term_ptr, V, base + u * 32 + lane), 1ull);
            if (base + 128 < i1) {
                t = term_of_posting(term_ptr, V, base + 128);
                end = __ldg(term_ptr + t + 1);
            }
        }
        uint32_t tot = local;
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
            ctot += __shfl_xor_sync(0xFFFFFFFFu, ctot, o);
        }
        if (lane == 0 && tot) atomicAdd(cnt + t, (unsigned long long)tot);
        if (lane == 0) chunk_cnt[c] = (int64_t)ctot;      // kept postings of this chunk: the write pass's offsets
    }
}

// Second pass of the selection: chunk c's kept postings go to out[chunk_off[c] ...] in their original order (a ballot
// and a popcount per 32 postings), with the document id renumbered to its rank among the passing documents.
__global__ void sparse_keep_write_kernel(const uint2* __restrict__ post, int64_t nnz, const uint32_t* __restrict__ mask,
                                         const uint32_t* __restrict__ prefix, const int64_t* __restrict__ chunk_off,
                                         uint2* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const unsigned below = (1u << lane) - 1u;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t nchunks = (nnz + KEEP_CHUNK - 1) / KEEP_CHUNK;
    for (int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < nchunks; c += nwarps) {
        const int64_t i0 = c * KEEP_CHUNK, i1 = min(nnz, i0 + (int64_t)KEEP_CHUNK);
        int64_t at = __ldg(chunk_off + c);
        if (__ldg(chunk_off + c + 1) == at) continue;     // nothing kept in this chunk
        for (int64_t base = i0; base < i1; base += 128) {
            uint2 p[4];
            uint32_t mw[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t i = base + u * 32 + lane;
                p[u] = i < i1 ? __ldcs(post + i) : make_uint2(0xFFFFFFFFu, 0u);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) mw[u] = p[u].x != 0xFFFFFFFFu ? __ldg(mask + (p[u].x >> 5)) : 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const bool keep = (mw[u] >> (p[u].x & 31)) & 1u;
                const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
                if (keep) {
                    const uint32_t nd = __ldg(prefix + (p[u].x >> 5)) + __popc(mw[u] & ((1u << (p[u].x & 31)) - 1u));
                    out[at + __popc(bal & below)] = make_uint2(nd, p[u].y);
                }
                at += __popc(bal);
            }
        }
    }
}

// ---- append to a sparse space (brag_sparse_append): every old posting list moves to its place in the grown index --
// out[dst_start[t] + (i - term_ptr[t])] = post[i] for posting i of old term t; same chunked walk as the count above.
__global__ void sparse_move_kernel(const int64_t* __restrict__ term_ptr, const uint2* __restrict__ post, int64_t V,
                                   int64_t nnz, const int64_t* __restrict__ dst_start, uint2* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t nchunks = (nnz + KEEP_CHUNK - 1) / KEEP_CHUNK;
    for (int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < nchunks; c += nwarps) {
        const int64_t i0 = c * KEEP_CHUNK, i1 = min(nnz, i0 + (int64_t)KEEP_CHUNK);
        int64_t t = term_of_posting(term_ptr, V, i0);
        int64_t end = __ldg(term_ptr + t + 1);
        int64_t shift = __ldg(dst_start + t) - __ldg(term_ptr + t);
        for (int64_t base = i0; base < i1; base += 128) {
            uint2 p[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t i = base + u * 32 + lane;
                p[u] = i < i1 ? __ldcs(post + i) : make_uint2(0u, 0u);
            }
            if (base + 128 <= end) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int64_t i = base + u * 32 + lane;
                    if (i < i1) out[i + shift] = p[u];
                }
                continue;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t i = base + u * 32 + lane;
                if (i < i1) {
                    const int64_t tt = term_of_posting(term_ptr, V, i);
                    out[i + __ldg(dst_start + tt) - __ldg(term_ptr + tt)] = p[u];
                }
            }
            if (base + 128 < i1) {
                t = term_of_posting(term_ptr, V, base + 128);
                end = __ldg(term_ptr + t + 1);
                shift = __ldg(dst_start + t) - __ldg(term_ptr + t);
            }
        }
    }
}

__global__ void sparse_place_kernel(const uint2* __restrict__ dpost, const int64_t* __restrict__ dst, int64_t n,
                                    uint2* __restrict__ out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) out[dst[j]] = dpost[j];
}

cudaError_t sparse_append_move(const int64_t* term_ptr, const uint2* post, int64_t V, int64_t nnz,
                               const int64_t* dst_start, const uint2* dpost, const int64_t* ddst, int64_t dnnz,
                               uint2* out, cudaStream_t st) {
    if (nnz) {
        const int64_t blocks = std::min<int64_t>((nnz + KEEP_CHUNK * 8 - 1) / (KEEP_CHUNK * 8), 148 * 8);
        sparse_move_kernel<<<(unsigned)blocks, 256, 0, st>>>(term_ptr, post, V, nnz, dst_start, out);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (dnnz) {
        sparse_place_kernel<<<(unsigned)((dnnz + 255) / 256), 256, 0, st>>>(dpost, ddst, dnnz, out);
        return cudaGetLastError();
    }
    return cudaSuccess;
}

int64_t sparse_compact_chunks(int64_t nnz) { return (nnz + KEEP_CHUNK - 1) / KEEP_CHUNK; }

size_t sparse_compact_tmp_bytes(int64_t V, int64_t nnz) {
    size_t a = 0, b = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, a, (const int64_t*)nullptr, (int64_t*)nullptr, (int)(V + 1));
    cub::DeviceScan::ExclusiveSum(nullptr, b, (const int64_t*)nullptr, (int64_t*)nullptr,
                                  (int)(sparse_compact_chunks(nnz) + 1));
    return std::max(a, b) + 256;
}

// new_term_ptr: [V + 1] (term_ptr of the passing postings); chunk_off: [chunks + 1] (where each chunk of the old
// posting array starts in the new one)
cudaError_t sparse_compact_count(const int64_t* term_ptr, const uint2* post, int64_t V, int64_t nnz,
                                 const uint32_t* mask, int64_t* new_term_ptr, int64_t* chunk_off, void* tmp,
                                 size_t tmp_bytes, cudaStream_t st) {
    const int64_t nchunks = sparse_compact_chunks(nnz);
    cudaError_t e = cudaMemsetAsync(new_term_ptr, 0, (size_t)(V + 1) * sizeof(int64_t), st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(chunk_off, 0, (size_t)(nchunks + 1) * sizeof(int64_t), st);
    if (e != cudaSuccess) return e;
    if (nnz) {
        const int64_t blocks = std::min<int64_t>((nnz + KEEP_CHUNK * 8 - 1) / (KEEP_CHUNK * 8), 148 * 8);
        sparse_keep_count_kernel<<<(unsigned)blocks, 256, 0, st>>>(term_ptr, post, V, nnz, mask,
                                                                  reinterpret_cast<unsigned long long*>(new_term_ptr), chunk_off);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    e = cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, new_term_ptr, new_term_ptr, (int)(V + 1), st);
    if (e != cudaSuccess) return e;
    return cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, chunk_off, chunk_off, (int)(nchunks + 1), st);
}

// out: [nnz_c] passing postings in their original order, with compact document ids
cudaError_t sparse_compact_select(const uint2* post, int64_t nnz, const uint32_t* mask, const uint32_t* prefix,
                                  const int64_t* chunk_off, uint2* out, int64_t nnz_c, cudaStream_t st) {
    if (nnz == 0 || nnz_c == 0) return cudaSuccess;
    const int64_t blocks = std::min<int64_t>((nnz + KEEP_CHUNK * 8 - 1) / (KEEP_CHUNK * 8), 148 * 8);
    sparse_keep_write_kernel<<<(unsigned)blocks, 256, 0, st>>>(post, nnz, mask, prefix, chunk_off, out);
    return cudaGetLastError();
}

// rowmap only (the dense gather builds its own inside compact_rows)
cudaError_t compact_rowmap(const uint32_t* mask, int64_t n, const uint32_t* prefix, uint32_t* rowmap, cudaStream_t st) {
    const int64_t nwords = (n + 31) / 32;
    mask_rowmap_kernel<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(mask, n, nwords, prefix, rowmap);
    return cudaGetLastError();
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
