// Row sharding across the GPUs of one box (SURVEY 8 row e; the reference's counterpart is one Qdrant collection
// answering query_points, qdrant_handler.py:363-372 -- here the collection's rows live on several GPUs).
//
// Every rank scans its own rows and ends with two ranked lists per query (dense, sparse) as sorted 64-bit keys
// (ordered score << 32 | ~GLOBAL row).  The exchange is fused into those tail kernels instead of being a collective
// of its own:
//   * shard_finalize_push_kernel sorts a query's candidates and stores the best k keys straight into EVERY rank's
//     peer-mapped window (NVLink stores; the dense list travels while the sparse scan is still running);
//   * shard_signal_kernel publishes "rank r's lists of epoch E have arrived" to every peer (fence.sys + st.release.sys);
//   * shard_fuse_kernel (one CTA per query) waits for the world's flags, merges the G dense lists and the G sparse
//     lists under (score desc, id asc) and fuses the two merged lists with RRF -- one launch instead of
//     all-gather + 2 merges + RRF.
// Slots are double-buffered by epoch parity: a rank cannot finish the fuse of epoch E + 1 before every peer has pushed
// E + 1, which a peer only does after its own fuse of E, so nobody overwrites a slot that is still being read.
// The same window carries the cross-shard score bounds of common.cuh (ShardBounds).
// The collective variant (brag_shard_local_lists -> any all-gather -> brag_shard_fuse) uses the same two kernels
// without the window; it is the path of the world_size-2 gloo tests and the checker of the peer path.
#include <algorithm>

#include "common.cuh"
#include "kernels.h"
#include "rrf.cuh"

namespace brag {

constexpr int SHF_THREADS = 512;
constexpr int SHF_MAX = 8192;               // keys per CTA sort (64 KB)
constexpr unsigned long long SHARD_WAIT_NS = 8000000000ull;   // a peer that never arrives becomes a trap, not a hang

ShardWindowLayout shard_window_layout(int world, int max_batch, int max_k) {
    ShardWindowLayout l;
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    l.world = world; l.max_batch = max_batch; l.max_k = max_k;
    l.slot_bytes = up((size_t)2 * max_batch * max_k * sizeof(uint64_t));
    l.slots_off = 0;
    l.flags_off = l.slots_off + (size_t)2 * world * l.slot_bytes;
    l.xb_off = l.flags_off + up((size_t)2 * world * sizeof(uint64_t));
    l.epoch_off = l.xb_off + up((size_t)2 * max_batch * sizeof(uint64_t));
    l.tables_off = l.epoch_off + 256;
    l.total = l.tables_off + up((size_t)3 * world * sizeof(void*));
    return l;
}

__global__ void shard_epoch_kernel(unsigned* epoch) { *epoch += 1u; }

cudaError_t shard_next_epoch(unsigned* epoch, cudaStream_t st) {
    shard_epoch_kernel<<<1, 1, 0, st>>>(epoch);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ sort + push of one list
// keys [B][m]: a scan's candidates with LOCAL rows (0 = empty).  Writes, per query, the best k as keys with GLOBAL rows
// (0-padded) to out[(list * B + b) * k ...] of every destination: the local buffer `local_out` (collective variant), or
// slot (epoch parity, rank) of every rank's window (peer variant).
__global__ void __launch_bounds__(SHF_THREADS)
shard_finalize_push_kernel(const uint64_t* __restrict__ keys, int m, int k, int64_t id_offset, int list, int B,
                           uint64_t* __restrict__ local_out, const ShardPush dst) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t* a = reinterpret_cast<uint64_t*>(smem_raw);
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n2 = next_pow2(max(m, k));
    for (int i = tid; i < n2; i += SHF_THREADS) a[i] = (i < m) ? keys[(size_t)b * m + i] : 0ull;
    __syncthreads();
    bitonic_desc(a, n2, tid, SHF_THREADS, [] { __syncthreads(); });
    const size_t at = ((size_t)list * B + b) * k;
    size_t slot = 0;
    if (local_out == nullptr) {
        const unsigned ep = *dst.epoch;
        slot = dst.slots_off + ((size_t)(ep & 1u) * dst.world + dst.rank) * dst.slot_bytes;
    }
    for (int i = tid; i < k; i += SHF_THREADS) {
        uint64_t key = a[i];
        if (key != 0ull) key = (key & 0xFFFFFFFF00000000ull) | (uint64_t)(0xFFFFFFFFu - (uint32_t)((int64_t)key_row(key) + id_offset));
        if (local_out != nullptr) {
            local_out[at + i] = key;
        } else {
            for (int g = 0; g < dst.world; ++g)
                reinterpret_cast<uint64_t*>(dst.windows[g] + slot)[at + i] = key;
        }
    }
}

cudaError_t shard_finalize_push(const uint64_t* keys, int B, int m, int k, int64_t id_offset, int list,
                                uint64_t* local_out, const ShardPush& dst, cudaStream_t st) {
    int n2 = 1;
    while (n2 < (m > k ? m : k)) n2 <<= 1;
    if (n2 > SHF_MAX) return cudaErrorInvalidValue;
    shard_finalize_push_kernel<<<B, SHF_THREADS, (size_t)n2 * 8, st>>>(keys, m, k, id_offset, list, B, local_out, dst);
    return cudaGetLastError();
}

// Both lists of this rank are in every window (stream order: the push kernels have completed): tell the world.
__global__ void shard_signal_kernel(const ShardPush dst) {
    const int g = threadIdx.x;
    if (g >= dst.world) return;
    const unsigned ep = *dst.epoch;
    __threadfence_system();
    unsigned long long* flag = reinterpret_cast<unsigned long long*>(dst.windows[g] + dst.flags_off) +
                               (size_t)(ep & 1u) * dst.world + dst.rank;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"((unsigned long long)ep) : "memory");
}

cudaError_t shard_signal(const ShardPush& dst, cudaStream_t st) {
    shard_signal_kernel<<<1, 32, 0, st>>>(dst);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ merge x 2 + RRF, one CTA per query
// src: gathered keys [G][2][B][k] (collective variant), or this rank's window (peer variant: slot (parity, g) holds
// rank g's [2][B][k]; the CTA first waits until every rank's flag shows the current epoch).
__global__ void __launch_bounds__(SHF_THREADS)
shard_fuse_kernel(const uint64_t* __restrict__ gathered, const ShardPush win, int G, int B, int k, int limit, int tie_mode,
                  double* __restrict__ out_scores, int64_t* __restrict__ out_ids, int32_t* __restrict__ out_counts,
                  float* __restrict__ stage_scores, int64_t* __restrict__ stage_ids, int32_t* __restrict__ stage_counts) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ int n_unique;
    __shared__ int s_cnt[2];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n_in = G * k;
    const int n2 = next_pow2(max(n_in, 32));
    const int E_max = 2 * k;
    const int n2r = next_pow2(max(E_max, max(limit, 1)));
    // layout: sort buffer / RRF entries (aliased) | fid [2k] | fpos [2k]
    const size_t head = max((size_t)n2 * 8, (size_t)n2r * sizeof(RrfEntry));
    uint64_t* a = reinterpret_cast<uint64_t*>(smem_raw);
    RrfEntry* ent = reinterpret_cast<RrfEntry*>(smem_raw);
    int64_t* fid = reinterpret_cast<int64_t*>(smem_raw + head);
    int32_t* fpos = reinterpret_cast<int32_t*>(fid + E_max);
    const uint8_t* slots = nullptr;
    if (gathered == nullptr) {
        const unsigned ep = *win.epoch;
        const uint8_t* mine = win.windows[win.rank];
        if (tid < G) {
            const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(mine + win.flags_off) +
                                             (size_t)(ep & 1u) * win.world + tid;
            unsigned long long t0 = 0ull;
            for (unsigned spin = 0;; ++spin) {
                unsigned long long v;
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
                if ((unsigned)v == ep) break;
                if ((spin & 63u) == 63u) {
                    unsigned long long now;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                    if (t0 == 0ull) t0 = now;
                    else if (now - t0 > SHARD_WAIT_NS) __trap();
                    __nanosleep(200);
                }
            }
        }
        slots = mine + win.slots_off + (size_t)(ep & 1u) * win.world * win.slot_bytes;
    }
    if (tid == 0) n_unique = 0;
    __syncthreads();
    for (int l = 0; l < 2; ++l) {
        for (int i = tid; i < n_in; i += SHF_THREADS) {
            const int g = i / k, j = i - g * k;
            const uint64_t* src = gathered != nullptr
                ? gathered + (((size_t)g * 2 + l) * B + b) * k
                : reinterpret_cast<const uint64_t*>(slots + (size_t)g * win.slot_bytes) + ((size_t)l * B + b) * k;
            a[i] = gathered != nullptr ? src[j] : __ldcv(src + j);     // peers wrote the window: never a cached copy
        }
        if (tid == 0) s_cnt[l] = 0;
        __syncthreads();
        // Every shard's list arrives sorted (best first, empties last) and rows are disjoint across shards, so a key's
        // place in the merged list is its index in its own list plus the number of greater keys in each other list
        // (a binary search per list): no sort.  (Was a bitonic sort of G * k keys per list.)
        if (tid < G) {
            const uint64_t* L = a + tid * k;
            int lo = 0, hi = k;                              // first empty slot of shard tid's list
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (L[mid] != 0ull) lo = mid + 1; else hi = mid; }
            atomicAdd(&s_cnt[l], lo);
        }
        for (int i = tid; i < k; i += SHF_THREADS) {
            fid[l * k + i] = -1;
            if (stage_scores != nullptr) {
                stage_scores[((size_t)l * B + b) * k + i] = -INFINITY;
                stage_ids[((size_t)l * B + b) * k + i] = -1;
            }
        }
        __syncthreads();
        for (int i = tid; i < n_in; i += SHF_THREADS) {
            const uint64_t key = a[i];
            if (key == 0ull) continue;
            const int g = i / k;
            int r = i - g * k;
            for (int g2 = 0; g2 < G && r < k; ++g2) {
                if (g2 == g) continue;
                const uint64_t* L = a + g2 * k;
                int lo = 0, hi = k;                          // keys of shard g2 greater than this one
                while (lo < hi) { const int mid = (lo + hi) >> 1; if (L[mid] > key) lo = mid + 1; else hi = mid; }
                r += lo;
            }
            if (r < k) {
                fid[l * k + r] = (int64_t)key_row(key);
                if (stage_scores != nullptr) {
                    stage_scores[((size_t)l * B + b) * k + r] = key_score(key);
                    stage_ids[((size_t)l * B + b) * k + r] = (int64_t)key_row(key);
                }
            }
        }
        __syncthreads();
        if (tid == 0) s_cnt[l] = min(s_cnt[l], k);
        __syncthreads();
    }
    // compact the two lists to [c0 | c1] with their ranks, then fuse
    const int c0 = s_cnt[0], c1 = s_cnt[1];
    if (stage_counts != nullptr && tid < 2) stage_counts[(size_t)tid * B + b] = s_cnt[tid];
    int64_t mv[2];
    int nm = 0;
    for (int j = tid; j < c1; j += SHF_THREADS) mv[nm++] = fid[k + j];      // <= ceil(512 / 512) = 1 per thread
    __syncthreads();
    nm = 0;
    for (int j = tid; j < c1; j += SHF_THREADS) fid[c0 + j] = mv[nm++];
    __syncthreads();
    const int E = c0 + c1;
    for (int e = tid; e < E; e += SHF_THREADS) fpos[e] = e < c0 ? e : e - c0;
    __syncthreads();
    const int n2e = next_pow2(max(E, max(limit, 1)));
    rrf_cta(ent, fid, fpos, E, n2e, limit, tie_mode, &n_unique, out_scores + (size_t)b * limit,
            out_ids + (size_t)b * limit, out_counts + b, tid, SHF_THREADS);
}

size_t shard_fuse_smem(int G, int k, int limit) {
    int n2 = 32;
    while (n2 < G * k) n2 <<= 1;
    int n2r = 1;
    while (n2r < std::max(2 * k, std::max(limit, 1))) n2r <<= 1;
    return std::max((size_t)n2 * 8, (size_t)n2r * sizeof(RrfEntry)) + (size_t)2 * k * 12;
}

cudaError_t shard_fuse(const uint64_t* gathered, const ShardPush& win, int G, int B, int k, int limit, int tie_mode,
                       double* out_scores, int64_t* out_ids, int32_t* out_counts, float* stage_scores,
                       int64_t* stage_ids, int32_t* stage_counts, cudaStream_t st) {
    if (G < 1 || (int64_t)G * k > SHF_MAX || k < 1 || k > 512 || limit > 1024 || limit < 1) return cudaErrorInvalidValue;
    shard_fuse_kernel<<<B, SHF_THREADS, shard_fuse_smem(G, k, limit), st>>>(
        gathered, win, G, B, k, limit, tie_mode, out_scores, out_ids, out_counts, stage_scores, stage_ids, stage_counts);
    return cudaGetLastError();
}

cudaError_t shard_init_device() {
    cudaError_t e = cudaFuncSetAttribute(shard_finalize_push_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SHF_MAX * 8);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(shard_fuse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)shard_fuse_smem(16, 512, 1024));
}

}  // namespace brag
