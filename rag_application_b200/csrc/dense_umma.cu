// Batched dense exact scan on the 5th-gen tensor cores with the top-k selection fused into the
// epilogue (reference request qdrant_handler.py:311-315,335-339,363-369
// issued for a batch of queries).  The [B, N] score matrix is never materialised.
//
// S[q, r] = sum_k Q[q, k] * X[r, k]; tile = 128 queries (UMMA M, TMEM lanes) x 256 corpus rows
// (UMMA N, TMEM columns), K streamed 64 elements (one 128-byte swizzle atom) per pipeline stage.
// One persistent CTA per SM owns a contiguous corpus chunk; for every X tile it sweeps ALL query
// blocks before moving on, so a tile comes from HBM once and is re-read from L2.
//   warp 0      : TMA producer  (cp.async.bulk.tensor -> 128B-swizzled smem, mbarrier tx)
//   warp 1      : TMEM allocator + single-thread tcgen05.mma issuer (accumulators in TMEM, x2 buffered)
//   warps 2..5  : epilogue.  Each thread owns ONE query (its TMEM lane): tcgen05.ld 32 columns at a
//                 time, compare against the query's global score bound; the rare survivors are
//                 appended (unsorted) to the (CTA, query) candidate list in global memory and counted
//                 in the query's global score histogram.
//   warp 6      : bound manager.  Sweeps the histograms of the queries this CTA manages and
//                 publishes, per query, the highest score edge that at least k_sel scanned rows reach.
//                 That is the k_sel-th best score over everything ANY CTA has scanned so far, so the
//                 pass rate decays like k/(rows scanned by the whole GPU) and no list is ever sorted
//                 inside the scan (a full list falls back to a local sort, never seen in practice).
// A second kernel gathers, per query, the candidates that reach the final bound and sorts them.
// Roofline: tensor pipe; flops = 2 * N * d * B.
#include <cuda.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "kernels.h"
#include "rescore.cuh"
#include "shard_push.cuh"

namespace brag {

// max of three floats in one instruction (PTX ISA 8.6, sm_100: FMNMX3); NaN operands are ignored like fmaxf
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

constexpr int UM_BM = 128;
constexpr int UM_BN = 256;
constexpr int UM_BK = 64;
constexpr int UM_STAGES = 4;
constexpr int UM_CAP = 512;            // candidate slots per (CTA, query)
constexpr int UM_THREADS = 224;
constexpr int UM_A_BYTES = UM_BM * UM_BK * 2;  // 16 KB
constexpr int UM_B_BYTES = UM_BN * UM_BK * 2;  // 32 KB
constexpr uint32_t UM_TMEM_COLS = 512;
constexpr int UM_MAXQB = 8;            // query blocks (x128 queries) one launch covers

struct UmSmem {
    uint8_t a[UM_STAGES][UM_A_BYTES];
    uint8_t b[UM_STAGES][UM_B_BYTES];
    uint64_t sortbuf[4][UM_CAP];
    uint64_t full[UM_STAGES];
    uint64_t empty[UM_STAGES];
    uint64_t tfull[2];
    uint64_t tempty[2];
    uint32_t tmem_base;
    int epi_done;                 // epilogue warps that have finished (stops the bound manager)
    float thr[UM_MAXQB][UM_BM];   // per (query block, query) local k-th best after an overflow sort
    int cnt[UM_MAXQB][UM_BM];     // candidates currently in that query's list
};

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t done = 0;
    // try_wait suspends the thread in hardware (hint: up to ~1 us) instead of burning issue slots;
    // the wait is bounded (~2 s of SM clocks) so a protocol bug becomes a trap, not a hung GPU
    long long t0 = 0;
    for (uint32_t spin = 0;; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity), "r"(1000u) : "memory");
        if (done) return;
        if ((spin & 255u) == 255u) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000ll) __trap();
        }
    }
}
// L2 eviction-priority policies (the encodings CUTLASS uses for TMA cache hints)
constexpr uint64_t L2_EVICT_NORMAL = 0x1000000000000000ull;
constexpr uint64_t L2_EVICT_FIRST = 0x12F0000000000000ull;
constexpr uint64_t L2_EVICT_LAST = 0x14F0000000000000ull;
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tmap, int c0, int c1, uint64_t* bar,
                                            uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// KIND: 0 = kind::f16 (f16 / bf16 operands, f32 accumulate), 1 = kind::tf32 (float32 storage), 2 = kind::i8 (int8
// operands, exact int32 accumulate).  Every kind consumes 32 bytes of K per instruction and one 128-byte swizzle atom
// per pipeline stage, so the pipeline below is the same for all three.
template <int KIND>
__device__ __forceinline__ void umma_issue(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    if constexpr (KIND == 0) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
    } else if constexpr (KIND == 1) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
    }
}
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
// wait for every outstanding tcgen05.ld of this thread; the "+r" operands tie the loaded registers
// to the wait so no use of them can be scheduled ahead of it
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
        : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
          "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]),
          "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]),
          "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
        :: "memory");
}

// one column for every lane of the warp's TMEM quarter (warp-collective; used on the rare hit path)
// v[i] for a run-time i without local memory: a 5-level tree of selects over statically indexed registers (31 SELs).
// The hit path used to re-read the column from TMEM (tcgen05.ld is warp-collective, so every hit column cost the whole
// warp a dependent TMEM round trip); now each lane pulls its own hits out of the registers it already holds.
__device__ __forceinline__ uint32_t sel32(const uint32_t (&v)[32], int i) {
    uint32_t a[16], b[8], c[4], d[2];
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = (i & 1) ? v[2 * j + 1] : v[2 * j];
#pragma unroll
    for (int j = 0; j < 8; ++j) b[j] = (i & 2) ? a[2 * j + 1] : a[2 * j];
#pragma unroll
    for (int j = 0; j < 4; ++j) c[j] = (i & 4) ? b[2 * j + 1] : b[2 * j];
#pragma unroll
    for (int j = 0; j < 2; ++j) d[j] = (i & 8) ? c[2 * j + 1] : c[2 * j];
    return (i & 16) ? d[1] : d[0];
}

// K-major, 128-byte swizzle: 8-row groups are 1024 B apart (SBO); LBO unused; version 1; layout 2.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// ------------------------------------------------------------------ epilogue list compaction
// Sort lane L's candidate list (global memory) in this warp's shared buffer, keep the best
// k_sel entries (written to dst), and return the new threshold / count to lane L.
__device__ __forceinline__ void um_compact(unsigned bal, uint64_t* my_list, uint64_t* my_dst, int& cnt, float& thr,
                                           uint64_t* sbuf, int k_sel, int lane) {
    while (bal) {
        const int L = __ffs(bal) - 1;
        bal &= bal - 1;
        uint64_t* src = reinterpret_cast<uint64_t*>(__shfl_sync(0xFFFFFFFFu, (unsigned long long)my_list, L));
        uint64_t* dst = reinterpret_cast<uint64_t*>(__shfl_sync(0xFFFFFFFFu, (unsigned long long)my_dst, L));
        const int c = __shfl_sync(0xFFFFFFFFu, cnt, L);
        const int n2 = next_pow2(max(c, 32));           // <= UM_CAP; short lists sort cheaply
        for (int i = lane; i < n2; i += 32) sbuf[i] = (i < c) ? __ldcg(src + i) : 0ull;
        __syncwarp();
        bitonic_desc(sbuf, n2, lane, 32, [] { __syncwarp(); });
        const int nc = min(c, k_sel);
        if (dst != nullptr)
            for (int i = lane; i < k_sel; i += 32) __stcg(dst + i, i < n2 ? sbuf[i] : 0ull);
        const uint64_t kth = (k_sel - 1 < n2) ? sbuf[k_sel - 1] : 0ull;
        if (lane == L) {
            cnt = nc;
            if (nc == k_sel) thr = key_score(kth);
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------ the kernel
// lists  : [grid][n_qblocks][128][UM_CAP] keys (unsorted), counts: [grid][B] entries used
// gthr   : [B] order-preserving uint32 score bounds (0 = none yet)
// coarse : [B][HIST_COARSE_BINS], fine: [B][HIST_FINE_BINS] row counts per score bin
// I8 (KIND 2): scores are cosines of integer vectors, idot * xinv[row] * qinv[query] (xinv / qinv = inverse norms).
// The accumulators stay int32; a 32-column group is first tested in the integer domain against
// Ti = floor(bound / (qinv * xinv_max)) (three-input integer max, one vote), and only the columns that pass are scaled
// by their own row's inverse norm and compared with the bound itself.
template <int KIND>
__global__ void __launch_bounds__(UM_THREADS, 1)
dense_umma_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x,
                  int64_t n, int n_kb, int B, int qb_base, int n_qblocks, float4* __restrict__ seedv, int seed_tiles,
                  int64_t rows_per_chunk, int share,
                  const uint32_t* __restrict__ mask, int k_sel, uint32_t idesc, uint64_t* __restrict__ lists,
                  int* __restrict__ counts, uint32_t* __restrict__ gthr, uint32_t* __restrict__ coarse,
                  uint32_t* __restrict__ fine, const ShardBounds xb, const float* __restrict__ xinv,
                  const float* __restrict__ qinv, float xinv_max) {
    constexpr int KELEMS = KIND == 0 ? 64 : (KIND == 1 ? 32 : 128);      // elements in one 128-byte K slice
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    UmSmem& S = *reinterpret_cast<UmSmem*>(smem_raw);   // keeps the shared address space (LDS/STS in the sort)
    if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();   // 128B-swizzle atoms need 1024-byte alignment
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // Seed launch (seedv != nullptr): CTA i scores ONE tile (tile i * stride, rows_per_chunk = stride in rows) against
    // every query block and only records, per query, the 4 best scores of that tile in seedv[query][i]; no lists, no
    // histograms.  Main launch: CTA i owns rows [i * rows_per_chunk, (i + 1) * rows_per_chunk) and list `chunk`.
    const bool seeding = seedv != nullptr;
    const int chunk = (int)blockIdx.x;
    // share > 1 (main launch only): `share` neighbouring CTAs scan the SAME rows, each against its own 1/share of the
    // query blocks, so the X tiles in flight at any moment are 148/share instead of 148 (they fit L2 and are read from
    // HBM once).  Lists / counts stay per CTA; the query blocks a CTA never sees keep count 0.  An odd number of query
    // blocks under share = 2: the even part is dealt block by block, and the last block (qx) is split by ROWS instead
    // -- the first CTA of the pair takes it for the first half of the tiles, the second for the rest -- so both do
    // the same work and every X tile is still fetched by both at about the same time.
    const int sh = seeding ? 1 : share;
    const int part = (int)blockIdx.x % sh;
    const bool odd = sh == 2 && (n_qblocks & 1);
    const int n_even = odd ? n_qblocks - 1 : n_qblocks;
    const int qb_lo = part * n_even / sh, qb_hi = (part + 1) * n_even / sh;
    const int qx = n_qblocks - 1;
    const int64_t row_begin = (int64_t)((int)blockIdx.x / sh) * rows_per_chunk;
    const int64_t row_end = min(n, row_begin + (seeding ? (int64_t)UM_BN : rows_per_chunk));
    const int n_tiles = row_end > row_begin ? (int)((row_end - row_begin + UM_BN - 1) / UM_BN) : 0;
    const int x_lo = odd ? (part == 0 ? 0 : (n_tiles + 1) / 2) : 0;          // tiles [x_lo, x_hi) also take block qx
    const int x_hi = odd ? (part == 0 ? (n_tiles + 1) / 2 : n_tiles) : 0;
    // the (tile, query block) units of this CTA, in the order all three roles walk them:
    //   for t: blocks qb_lo .. qb_hi-1, then qx if x_lo <= t < x_hi

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < UM_STAGES; ++s) { mbar_init(&S.full[s], 1); mbar_init(&S.empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&S.tfull[s], 1); mbar_init(&S.tempty[s], 4); }
        S.epi_done = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "r"(UM_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = S.tmem_base;

    if (warp == 0) {
        // ===== TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = 0; t < n_tiles; ++t) {
                const int row0 = (int)(row_begin + (int64_t)t * UM_BN);
                const int n_blk = qb_hi - qb_lo + ((t >= x_lo && t < x_hi) ? 1 : 0);
                for (int iq = 0; iq < n_blk; ++iq) {
                    const int qb = (iq < qb_hi - qb_lo) ? qb_lo + iq : qx;
                    for (int kb = 0; kb < n_kb; ++kb) {
                        mbar_wait(&S.empty[stage], phase ^ 1);
                        mbar_expect_tx(&S.full[stage], UM_A_BYTES + UM_B_BYTES);
                        // queries stay hot in L2 for the whole scan; an X tile is wanted until its last
                        // query block has used it, then it should leave first
                        tma_load_2d(S.a[stage], &tmap_q, kb * KELEMS, (qb_base + qb) * UM_BM, &S.full[stage], L2_EVICT_LAST);
                        tma_load_2d(S.b[stage], &tmap_x, kb * KELEMS, row0, &S.full[stage],
                                    (iq + 1 < n_blk || sh > 1) ? L2_EVICT_NORMAL : L2_EVICT_FIRST);
                        if (++stage == UM_STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread)
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            const int n_units = n_tiles * (qb_hi - qb_lo) + max(0, min(x_hi, n_tiles) - x_lo);   // unit = (X tile, query block)
            for (int u = 0; u < n_units; ++u) {
                const int as = u & 1;
                const uint32_t aphase = (uint32_t)(u >> 1) & 1;
                mbar_wait(&S.tempty[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)as * UM_BN;
                for (int kb = 0; kb < n_kb; ++kb) {
                    mbar_wait(&S.full[stage], phase);
                    tc_fence_after();
                    const uint64_t adesc = make_smem_desc(smem_u32(S.a[stage]));
                    const uint64_t bdesc = make_smem_desc(smem_u32(S.b[stage]));
#pragma unroll
                    for (int k = 0; k < UM_BK / 16; ++k) {
                        // +32 bytes per 16-element K step inside the swizzle atom (>>4 => +2)
                        umma_issue<KIND>(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) ? 1u : 0u);
                    }
                    umma_commit(&S.empty[stage]);
                    if (kb == n_kb - 1) umma_commit(&S.tfull[as]);
                    if (++stage == UM_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 6) {
        // ===== bound manager: queries q = blockIdx.x, blockIdx.x + gridDim.x, ... of this launch
        const int q_lo = qb_base * UM_BM, q_hi = min(B, (qb_base + n_qblocks) * UM_BM);
        while (true) {
            const int done = *reinterpret_cast<volatile int*>(&S.epi_done);
            for (int q = q_lo + (int)blockIdx.x; q < q_hi; q += (int)gridDim.x) {
                publish_bound(gthr, q, hist_bound(coarse, fine, q, (uint32_t)k_sel, lane), xb, lane);
            }
            if (done >= 4) break;
            __nanosleep(500);
        }
    } else {
        // ===== epilogue (warps 2..5): thread <-> query lane; the query block changes with every unit
        const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
        const int qidx = quarter * 32 + lane;         // query within the block
        uint64_t* sbuf = S.sortbuf[warp - 2];
        const uint32_t lane_taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
        for (int qb = 0; qb < n_qblocks; ++qb) {
            S.thr[qb][qidx] = ((qb_base + qb) * UM_BM + qidx < B) ? -INFINITY : INFINITY;
            S.cnt[qb][qidx] = 0;
        }
        int u = 0;
        for (int t = 0; t < n_tiles; ++t) {
            const int64_t row0 = row_begin + (int64_t)t * UM_BN;
            const int n_blk = qb_hi - qb_lo + ((t >= x_lo && t < x_hi) ? 1 : 0);
            for (int iq = 0; iq < n_blk; ++iq, ++u) {
                const int qb = (iq < qb_hi - qb_lo) ? qb_lo + iq : qx;
                const int as = u & 1;
                const uint32_t aphase = (uint32_t)(u >> 1) & 1;
                const int b = (qb_base + qb) * UM_BM + qidx;
                const bool q_valid = b < B;
                if (seeding) {
                    // the 4 best scores of this tile for this query (rows that exist and pass the filter)
                    mbar_wait(&S.tfull[as], aphase);
                    tc_fence_after();
                    const uint32_t taddr_s = lane_taddr + (uint32_t)(as * UM_BN);
                    float t0 = -INFINITY, t1 = -INFINITY, t2 = -INFINITY, t3 = -INFINITY;
                    uint32_t v[32];
#pragma unroll 1
                    for (int g2 = 0; g2 < UM_BN / 32; ++g2) {
                        tmem_ld32_issue(taddr_s + (uint32_t)(g2 * 32), v);
                        const int64_t r0 = row0 + g2 * 32;
                        uint32_t cm = (row_end - r0 >= 32) ? 0xFFFFFFFFu : (row_end > r0 ? (1u << (int)(row_end - r0)) - 1u : 0u);
                        if (mask != nullptr) cm &= __ldg(mask + (r0 >> 5));
                        tmem_ld_wait(v);
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            float a;
                            if constexpr (KIND == 2) a = ((cm >> i) & 1u) ? (float)(int)v[i] * __ldg(xinv + min(r0 + i, n - 1)) : -INFINITY;
                            else a = ((cm >> i) & 1u) ? __uint_as_float(v[i]) : -INFINITY;
                            float hi = fmaxf(t0, a); a = fminf(t0, a); t0 = hi;
                            hi = fmaxf(t1, a); a = fminf(t1, a); t1 = hi;
                            hi = fmaxf(t2, a); a = fminf(t2, a); t2 = hi;
                            t3 = fmaxf(t3, a);
                        }
                    }
                    if constexpr (KIND == 2) {
                        if (q_valid) { const float qi = __ldg(qinv + b); t0 *= qi; t1 *= qi; t2 *= qi; t3 *= qi; }
                    }
                    if (q_valid) seedv[(size_t)b * seed_tiles + blockIdx.x] = make_float4(t0, t1, t2, t3);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&S.tempty[as]);
                    continue;
                }
                uint64_t* my_list = lists + (((size_t)chunk * n_qblocks + qb) * UM_BM + qidx) * UM_CAP;
                float thr = S.thr[qb][qidx];
                int cnt = S.cnt[qb][qidx];
                // the query's global bound (k_sel-th best score over all rows scanned so far by any CTA,
                // rounded down to a histogram edge) -- fetched while we wait for the accumulator
                uint32_t g = q_valid ? __ldcg(gthr + b) : 0u;
                mbar_wait(&S.tfull[as], aphase);
                tc_fence_after();
                float eff = thr;
                if (g != 0u) eff = fmaxf(eff, nextafterf(ordered_f32(g), -INFINITY));   // keeps score >= bound
                const uint32_t taddr = lane_taddr + (uint32_t)(as * UM_BN);
                // I8: the same threshold in the integer domain of the accumulators.  score = idot * xinv[row] * qinv
                // <= idot * xinv_max * qinv for idot >= 0, so score > eff needs idot > eff / (qinv * xinv_max); a
                // non-positive threshold lets every column through to the exact test.
                float my_qinv = 1.f, qsc = 1.f;
                int Ti = 0;
                if constexpr (KIND == 2) {
                    my_qinv = q_valid ? __ldg(qinv + b) : 1.f;
                    qsc = 1.f / (my_qinv * xinv_max);
                }
                auto int_threshold = [&](float e) {
                    return e > 0.f ? __float2int_rd(e * qsc * 0.999999f) - 1 : (int)0x80000000;
                };
                if constexpr (KIND == 2) Ti = int_threshold(eff);
                auto process = [&](uint32_t (&v)[32], int col0) {
                    if constexpr (KIND == 2) {
                        int c0 = __vimax3_s32((int)v[0], (int)v[1], (int)v[2]), c1 = __vimax3_s32((int)v[8], (int)v[9], (int)v[10]);
                        int c2 = __vimax3_s32((int)v[16], (int)v[17], (int)v[18]), c3 = __vimax3_s32((int)v[24], (int)v[25], (int)v[26]);
                        c0 = __vimax3_s32(c0, (int)v[3], (int)v[4]);   c1 = __vimax3_s32(c1, (int)v[11], (int)v[12]);
                        c2 = __vimax3_s32(c2, (int)v[19], (int)v[20]); c3 = __vimax3_s32(c3, (int)v[27], (int)v[28]);
                        c0 = __vimax3_s32(c0, (int)v[5], (int)v[6]);   c1 = __vimax3_s32(c1, (int)v[13], (int)v[14]);
                        c2 = __vimax3_s32(c2, (int)v[21], (int)v[22]); c3 = __vimax3_s32(c3, (int)v[29], (int)v[30]);
                        c0 = __vimax3_s32(c0, (int)v[7], (int)v[15]);  c2 = __vimax3_s32(c2, (int)v[23], (int)v[31]);
                        if (!__any_sync(0xFFFFFFFFu, __vimax3_s32(c0, c1, max(c2, c3)) > Ti)) return;
                        const int64_t r0 = row0 + col0;
                        uint32_t cm = (row_end - r0 >= 32) ? 0xFFFFFFFFu : (row_end > r0 ? (1u << (int)(row_end - r0)) - 1u : 0u);
                        if (mask != nullptr) cm &= __ldg(mask + (r0 >> 5));
                        uint32_t m = 0u;
#pragma unroll
                        for (int i = 0; i < 32; ++i) m |= ((int)v[i] > Ti) ? (1u << i) : 0u;
                        m &= cm;
                        while (m) {                                    // this lane's own hits (rare, lanes diverge)
                            const int i = __ffs(m) - 1;
                            m &= m - 1;
                            const int iv = (int)sel32(v, i);
                            const float s = (float)iv * __ldg(xinv + r0 + i) * my_qinv;   // same two roundings as the seeding pass
                            if (s > eff) {
                                const uint64_t key = make_key(s, (uint32_t)(r0 + i));
                                __stcg(my_list + cnt, key);
                                ++cnt;
                                hist_count(coarse, fine, b, (uint32_t)(key >> 32));
                            }
                        }
                        __syncwarp();
                        return;
                    }
                    // rows of this 32-column group that exist and pass the payload filter (tiles start on
                    // multiples of 256 rows, so a group is exactly one filter word): warp-uniform
                    // most groups hold no candidate for any of the warp's 32 queries: the group's maximum (three-input
                    // max, 18 instructions for 32 scores) against the bound settles those with one vote
                    float c0 = fmax3(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]));
                    float c1 = fmax3(__uint_as_float(v[8]), __uint_as_float(v[9]), __uint_as_float(v[10]));
                    float c2 = fmax3(__uint_as_float(v[16]), __uint_as_float(v[17]), __uint_as_float(v[18]));
                    float c3 = fmax3(__uint_as_float(v[24]), __uint_as_float(v[25]), __uint_as_float(v[26]));
                    c0 = fmax3(c0, __uint_as_float(v[3]), __uint_as_float(v[4]));
                    c1 = fmax3(c1, __uint_as_float(v[11]), __uint_as_float(v[12]));
                    c2 = fmax3(c2, __uint_as_float(v[19]), __uint_as_float(v[20]));
                    c3 = fmax3(c3, __uint_as_float(v[27]), __uint_as_float(v[28]));
                    c0 = fmax3(c0, __uint_as_float(v[5]), __uint_as_float(v[6]));
                    c1 = fmax3(c1, __uint_as_float(v[13]), __uint_as_float(v[14]));
                    c2 = fmax3(c2, __uint_as_float(v[21]), __uint_as_float(v[22]));
                    c3 = fmax3(c3, __uint_as_float(v[29]), __uint_as_float(v[30]));
                    c0 = fmax3(c0, __uint_as_float(v[7]), __uint_as_float(v[15]));
                    c2 = fmax3(c2, __uint_as_float(v[23]), __uint_as_float(v[31]));
                    if (!__any_sync(0xFFFFFFFFu, fmax3(c0, c1, fmaxf(c2, c3)) > eff)) return;
                    const int64_t r0 = row0 + col0;
                    uint32_t cm = (row_end - r0 >= 32) ? 0xFFFFFFFFu : (row_end > r0 ? (1u << (int)(row_end - r0)) - 1u : 0u);
                    if (mask != nullptr) cm &= __ldg(mask + (r0 >> 5));
                    uint32_t m = 0u;
#pragma unroll
                    for (int i = 0; i < 32; ++i) m |= (__uint_as_float(v[i]) > eff) ? (1u << i) : 0u;
                    m &= cm;
                    while (m) {                                    // this lane's own hits (rare, lanes diverge)
                        const int i = __ffs(m) - 1;
                        m &= m - 1;
                        const uint64_t key = make_key(__uint_as_float(sel32(v, i)), (uint32_t)(r0 + i));
                        __stcg(my_list + cnt, key);
                        ++cnt;
                        hist_count(coarse, fine, b, (uint32_t)(key >> 32));
                    }
                    __syncwarp();
                };
                uint32_t va[32], vb[32];
                tmem_ld32_issue(taddr, va);
#pragma unroll 1
                for (int g2 = 0; g2 < UM_BN / 64; ++g2) {
                    const unsigned bal = __ballot_sync(0xFFFFFFFFu, cnt > UM_CAP - 64);
                    if (bal) {   // list full (pathological): sort locally, keep k_sel, tighten the local bound
                        um_compact(bal, my_list, my_list, cnt, thr, sbuf, k_sel, lane);
                        eff = fmaxf(eff, thr);
                        if constexpr (KIND == 2) Ti = int_threshold(eff);
                    }
                    if (__any_sync(0xFFFFFFFFu, q_valid && g == 0u)) {
                        // warm-up: no bound yet -- pick it up as soon as the manager publishes one, so the
                        // initial flood of candidates stops after a few dozen rows per query
                        if (q_valid && g == 0u) {
                            g = __ldcg(gthr + b);
                            if (g != 0u) eff = fmaxf(eff, nextafterf(ordered_f32(g), -INFINITY));
                            if constexpr (KIND == 2) Ti = int_threshold(eff);
                        }
                    }
                    tmem_ld_wait(va);
                    tmem_ld32_issue(taddr + (uint32_t)(g2 * 64 + 32), vb);
                    process(va, g2 * 64);
                    tmem_ld_wait(vb);
                    if (g2 + 1 < UM_BN / 64) tmem_ld32_issue(taddr + (uint32_t)(g2 * 64 + 64), va);
                    process(vb, g2 * 64 + 32);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&S.tempty[as]);
                S.thr[qb][qidx] = thr;
                S.cnt[qb][qidx] = cnt;
            }
        }
        for (int qb = 0; qb < n_qblocks && !seeding; ++qb) {
            const int b = (qb_base + qb) * UM_BM + qidx;
            if (b < B) counts[(size_t)chunk * B + b] = S.cnt[qb][qidx];
        }
        __syncwarp();
        if (lane == 0) atomicAdd(&S.epi_done, 1);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(UM_TMEM_COLS) : "memory");
    }
}

// Seed bound: per query, the k_sel-th largest of the seed_tiles * 4 per-tile best scores.  They are scores of
// distinct rows, so at least k_sel rows reach that value: a valid starting bound for the main launch, about the
// k_sel-th best of seed_tiles * 256 rows.  One CTA per query, bitonic sort of <= 1024 values in shared memory.
__global__ void __launch_bounds__(256)
um_seed_bound_kernel(const float4* __restrict__ seedv, int seed_tiles, int q_lo, int q_hi, int k_sel,
                     uint32_t* __restrict__ gthr) {
    // k_sel-th largest of <= 1024 ordered-float images by radix select, 8 bits per pass: histogram of the current
    // digit over the values that still match the prefix, then the digit whose suffix count crosses the rank
    // (4 passes x 3 barriers; the bitonic sort this replaces took 55)
    __shared__ uint32_t a[1024];
    __shared__ uint32_t hist[256];
    __shared__ uint32_t s_prefix, s_rank;
    const int q = q_lo + (int)blockIdx.x;
    if (q >= q_hi) return;
    const int tid = threadIdx.x;
    const int nv = min(seed_tiles * 4, 1024);
    if (nv < k_sel) return;
    const float* src = reinterpret_cast<const float*>(seedv + (size_t)q * seed_tiles);
    for (int i = tid; i < nv; i += 256) a[i] = f32_ordered(src[i]);
    if (tid == 0) { s_prefix = 0u; s_rank = (uint32_t)k_sel; }      // rank counted from the top, 1-based
    for (int shift = 24; shift >= 0; shift -= 8) {
        hist[tid] = 0u;
        __syncthreads();
        const uint32_t prefix = s_prefix, himask = shift == 24 ? 0u : (0xFFFFFFFFu << (shift + 8));
        for (int i = tid; i < nv; i += 256)
            if ((a[i] & himask) == prefix) atomicAdd(&hist[(a[i] >> shift) & 255u], 1u);
        __syncthreads();
        if (tid < 32) {
            // lane l owns digits 8l .. 8l+7; suffix counts from the top digit down
            uint32_t own = 0u, c[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { c[j] = hist[tid * 8 + j]; own += c[j]; }
            const uint32_t suf = warp_suffix_sum(own, tid);          // values in digits >= 8 * tid
            const uint32_t rank = s_rank;
            const unsigned has = __ballot_sync(0xFFFFFFFFu, suf >= rank);
            const int L = 31 - __clz(has);                           // has != 0: the prefix class holds >= rank values
            if (tid == L) {
                uint32_t above = suf - own;                          // values in digits above this lane's
                int dgt = 7;
                for (; dgt > 0; --dgt) { if (above + c[dgt] >= rank) break; above += c[dgt]; }
                s_prefix = prefix | ((uint32_t)(tid * 8 + dgt) << shift);
                s_rank = rank - above;
            }
        }
        __syncthreads();
    }
    if (tid == 0) {
        const uint32_t v = s_prefix;
        if (v > f32_ordered(-INFINITY)) atomicMax(gthr + q, v);
    }
}

// ------------------------------------------------------------------ tail of the scan, one CTA per query
// One launch where there were four (bound, selection, re-score, finalize):
//   1. the bound of the query's complete histograms (a CTA's manager stops with its own CTA, so the bound it left may
//      predate the last pushes of the others) -- published like every bound, also to the peers of a sharded call;
//   2. the candidates of the n_chunks unsorted (CTA, query) lists that reach it (at least k_sel of them do on a single
//      GPU), sorted, the best k_sel kept;
//   3. those re-scored exactly: float64 dot of the stored row with the caller's float32 query (rescore.cuh), one warp
//      per candidate;
//   4. the re-scored keys (local rows) to keys_out [B][k_sel] (filter-compacted copies go on from there: their rows
//      still need the map back), or the final list itself by rank: unpacked with the id offset and the count, or --
//      sharded calls -- its best k keys with global rows straight into every rank's window (sink).
constexpr int SELK_MAX_THREADS = 1024;   // launched with 512 threads, 1024 for batches of <= 64 queries (the re-score is
                                         // one warp per candidate: a small batch has no other CTAs to hide its latency)
constexpr int SELK_CAP = 2048;     // keys gathered per query (16 KB: four CTAs per SM); a few hundred reach the bound
constexpr int SELK_GW = SELK_CAP / 2 / UM_CAP;   // lists added per round of the overflow path
static_assert(SELK_GW >= 1, "a round must take at least one full list");
constexpr int SELK_RES = 512;      // re-scored keys (k_sel <= 512)
constexpr int SELK_MAXCHUNKS = 256;
constexpr int SELK_RANK_MAX = 1024;   // gathered keys up to which the selection ranks instead of sorting
static_assert(SELK_RANK_MAX + SELK_RES <= SELK_CAP, "the ranked keys are placed behind the gathered ones");

__global__ void __launch_bounds__(SELK_MAX_THREADS)
dense_umma_tail_kernel(const uint64_t* __restrict__ lists, const int* __restrict__ counts,
                       const uint32_t* __restrict__ coarse, const uint32_t* __restrict__ fine,
                       uint32_t* __restrict__ gthr, const ShardBounds xb, int n_chunks, int B, int qb_base, int n_qblocks,
                       int k_sel, const void* __restrict__ X, int dtype, int64_t n, int d, const float* __restrict__ Q,
                       uint64_t* __restrict__ keys_out, int k, int64_t id_offset, float* __restrict__ out_scores,
                       int64_t* __restrict__ out_ids, int32_t* __restrict__ out_counts, const ShardSink sink) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint64_t* a = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* res = a + SELK_CAP;
    __shared__ int fill, total, count;
    __shared__ uint32_t s_bound;
    __shared__ int s_cnt[SELK_MAXCHUNKS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int SELK_THREADS = (int)blockDim.x;
    const int b = qb_base * UM_BM + blockIdx.x;
    if (b >= B) return;
    const int qb = blockIdx.x / UM_BM, qidx = blockIdx.x % UM_BM;
    if (tid == 0) { fill = 0; total = 0; count = 0; }
    __syncthreads();
    if (warp == 0) {
        publish_bound(gthr, b, hist_bound(coarse, fine, b, (uint32_t)k_sel, lane), xb, lane);
        __syncwarp();
        if (lane == 0) s_bound = *reinterpret_cast<volatile uint32_t*>(gthr + b);
    } else {
        int t = 0;
        for (int c = tid - 32; c < n_chunks; c += SELK_THREADS - 32) {
            const int v = min(__ldcg(counts + (size_t)c * B + b), UM_CAP);
            if (c < SELK_MAXCHUNKS) s_cnt[c] = v;
            t += v;
        }
        if (t) atomicAdd(&total, t);
    }
    __syncthreads();
    const uint32_t bound = s_bound;
    const uint64_t* my_lists = lists + ((size_t)qb * UM_BM + qidx) * UM_CAP;
    const size_t chunk_stride = (size_t)n_qblocks * UM_BM * UM_CAP;
    // the usual case: few keys reach the final bound, whatever the lists hold (a batch without a seed launch floods
    // its first tiles into them); the warps stream their chunks' lists without meeting in between
    for (int chunk = warp; chunk < n_chunks; chunk += SELK_THREADS / 32) {
        const int c = chunk < SELK_MAXCHUNKS ? s_cnt[chunk] : min(__ldcg(counts + (size_t)chunk * B + b), UM_CAP);
        const uint64_t* src = my_lists + (size_t)chunk * chunk_stride;
        for (int i = lane; i < c; i += 32) {
            const uint64_t key = __ldcg(src + i);
            if ((uint32_t)(key >> 32) >= bound) {
                const int slot = atomicAdd(&fill, 1);
                if (slot < SELK_CAP) a[slot] = key;
            }
        }
    }
    __syncthreads();
    if (fill > SELK_CAP) {
        // more than the buffer holds (mass ties, or no bound at all): rounds of lists, keeping the best k_sel in between
        __syncthreads();
        if (tid == 0) fill = 0;
        __syncthreads();
        for (int c0 = 0; c0 < n_chunks; c0 += SELK_GW) {     // SELK_GW warps gather: SELK_GW * UM_CAP keys at most per round
            if (fill + SELK_GW * UM_CAP > SELK_CAP) {       // uniform: fill is stable here
                const int f = fill;
                const int n2 = next_pow2(max(f, 32));
                for (int i = f + tid; i < n2; i += SELK_THREADS) a[i] = 0ull;
                __syncthreads();
                bitonic_desc(a, n2, tid, SELK_THREADS, [] { __syncthreads(); });
                if (tid == 0) fill = min(f, k_sel);
                __syncthreads();
            }
            const int chunk = c0 + warp;
            if (warp < SELK_GW && chunk < n_chunks) {
                const int c = min(__ldcg(counts + (size_t)chunk * B + b), UM_CAP);
                const uint64_t* src = my_lists + (size_t)chunk * chunk_stride;
                for (int i = lane; i < c; i += 32) {
                    const uint64_t key = __ldcg(src + i);
                    if ((uint32_t)(key >> 32) >= bound) a[atomicAdd(&fill, 1)] = key;
                }
            }
            __syncthreads();
        }
    }
    const int f = fill;
    // the best k_sel of the gathered keys: by rank (keys are distinct: count the greater ones, no barriers) when
    // there are few, else a sort
    const int n_res = min(f, k_sel);
    uint64_t* top = a + SELK_CAP - SELK_RES;          // their slots, past anything the gather filled in the rank case
    if (f <= SELK_RANK_MAX) {
        for (int i = tid; i < f; i += SELK_THREADS) {
            const uint64_t x = a[i];
            int r = 0;
            for (int j = 0; j < f; ++j) r += a[j] > x ? 1 : 0;
            if (r < k_sel) top[r] = x;
        }
    } else {
        const int n2 = next_pow2(max(f, 32));
        for (int i = f + tid; i < n2; i += SELK_THREADS) a[i] = 0ull;
        __syncthreads();
        bitonic_desc(a, n2, tid, SELK_THREADS, [] { __syncthreads(); });
        top = a;
    }
    __syncthreads();
    // exact scores of those candidates
    for (int i = warp; i < n_res; i += SELK_THREADS / 32) {
        const uint64_t key = top[i];
        uint64_t out = 0ull;
        if (key != 0ull) {
            const uint32_t row = key_row(key);
            if ((int64_t)row < n) out = make_key((float)row_dot_any(X, dtype, row, d, Q + (size_t)b * d, lane), row);
        }
        if (lane == 0) res[i] = out;
    }
    __syncthreads();
    if (keys_out != nullptr) {
        for (int i = tid; i < k_sel; i += SELK_THREADS) keys_out[(size_t)b * k_sel + i] = i < n_res ? res[i] : 0ull;
        return;
    }
    // final order by rank again (a re-scored key keeps its row, so keys stay distinct; 0 = a row that did not exist)
    if (sink.on) {
        int nz = 0;
        for (int i = tid; i < n_res; i += SELK_THREADS) nz += res[i] != 0ull ? 1 : 0;
        if (nz) atomicAdd(&count, nz);
        __syncthreads();
        for (int r = count + tid; r < k; r += SELK_THREADS) shard_sink_store(sink, B, b, k, r, 0ull, id_offset);
        for (int i = tid; i < n_res; i += SELK_THREADS) {
            const uint64_t x = res[i];
            if (x == 0ull) continue;
            int r = 0;
            for (int j = 0; j < n_res; ++j) r += res[j] > x ? 1 : 0;
            if (r < k) shard_sink_store(sink, B, b, k, r, x, id_offset);
        }
        return;
    }
    for (int i = tid; i < k; i += SELK_THREADS) {
        out_scores[(size_t)b * k + i] = -INFINITY;
        out_ids[(size_t)b * k + i] = -1;
    }
    __syncthreads();
    for (int i = tid; i < n_res; i += SELK_THREADS) {
        const uint64_t x = res[i];
        if (x == 0ull) continue;
        int r = 0;
        for (int j = 0; j < n_res; ++j) r += res[j] > x ? 1 : 0;
        if (r < k) {
            out_scores[(size_t)b * k + r] = key_score(x);
            out_ids[(size_t)b * k + r] = (int64_t)key_row(x) + id_offset;
            atomicAdd(&count, 1);
        }
    }
    __syncthreads();
    if (tid == 0) out_counts[b] = count;
}

// ------------------------------------------------------------------ host side
// dtype: BRAG_F32 (0, kind::tf32), BRAG_F16 / BRAG_BF16 (1 / 2, kind::f16), BRAG_I8 (3, kind::i8).  A row must be a
// whole number of 128-byte K slices.
static int um_esz(int dtype) { return dtype == 0 ? 4 : (dtype == 3 ? 1 : 2); }
bool dense_umma_supported(int dtype, int d, int64_t n) {
    return dtype >= 0 && dtype <= 3 && d >= 1 && ((int64_t)d * um_esz(dtype)) % 128 == 0 && n >= 1;
}

int dense_umma_max_ksel() { return UM_CAP - 64; }

// CTAs per shared row range in the main launch: 2 (BRAG_UM_SHARE = 1, 2 or 4 overrides; tuning knob, read once).
// Measured at 10 M x 1024 bf16, batch 1024, inside the hybrid step (the scan is power-capped, so HBM traffic saved
// is clock gained): 1 -> 16.65 ms, 2 -> 13.79 ms, 4 -> 13.95 ms (profiles/r1_dense_share_v30.txt).
static int um_share() {
    static const int v = [] {
        const char* e = getenv("BRAG_UM_SHARE");
        const int x = e ? atoi(e) : 2;
        return (x == 1 || x == 2 || x == 4) ? x : 2;
    }();
    return v;
}

cudaError_t dense_umma_plan(int64_t n, int d, int B, int k_sel, int num_sms, UmmaPlan* plan) {
    if (k_sel > UM_CAP - 64) return cudaErrorInvalidValue;
    (void)d;
    plan->n_qblocks = (B + UM_BM - 1) / UM_BM;
    const int64_t n_tiles = (n + UM_BN - 1) / UM_BN;
    // Seed launch: for batches of more than 64 queries over a corpus of >= 64 tiles, n_seed single-tile CTAs (tiles
    // spread over the corpus) record each query's 4 best scores per tile; the k_sel-th largest of those is the
    // query's starting bound.  Without it every CTA of the main launch pushes its whole first tile through the slow
    // per-hit path before a bound exists (measured: +0.6 ms per launch at batch 1024).
    plan->n_seed = 0;
    if (B > 64 && n_tiles >= 64) {
        const int want = (int)std::min<int64_t>(num_sms, n_tiles / 4);
        if (want * 4 >= k_sel && want * 4 <= 1024) plan->n_seed = want;
    }
    plan->n_chunks = (int)std::min<int64_t>(num_sms, std::max<int64_t>(1, n_tiles));   // one persistent CTA per SM
    const int qb_launch = std::min(plan->n_qblocks, UM_MAXQB);   // lists are reused by successive query groups
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    plan->list_bytes = up((size_t)plan->n_chunks * qb_launch * UM_BM * UM_CAP * sizeof(uint64_t));
    plan->count_bytes = up((size_t)plan->n_chunks * B * sizeof(int));
    plan->gthr_bytes = up((size_t)B * sizeof(uint32_t));
    plan->coarse_bytes = up((size_t)B * HIST_COARSE_BINS * sizeof(uint32_t));
    plan->fine_bytes = up((size_t)B * HIST_FINE_BINS * sizeof(uint32_t));
    plan->seed_bytes = up((size_t)B * plan->n_seed * sizeof(float4));
    plan->scratch_bytes = plan->list_bytes + plan->count_bytes + plan->gthr_bytes + plan->coarse_bytes + plan->fine_bytes +
                          plan->seed_bytes;
    return cudaSuccess;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// [rows, d] row-major matrix of `dtype`, box = [box_rows, one 128-byte K slice], 128-byte swizzle.  (16-bit floats
// move as raw bits: the map's element type only matters for the element size and OOB fill.)
cudaError_t make_tmap_2d(void* tmap_out, const void* base, int64_t rows, int d, int box_rows, int dtype) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return cudaErrorNotSupported;
    const int esz = um_esz(dtype);
    const CUtensorMapDataType dt = esz == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                            : (esz == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
    CUtensorMap tm;
    cuuint64_t gdim[2] = {(cuuint64_t)d, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)d * esz};
    cuuint32_t box[2] = {(cuuint32_t)(128 / esz), (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tm, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    memcpy(tmap_out, &tm, sizeof(CUtensorMap));
    return cudaSuccess;
}

// kernel attributes are per device: brag_create calls this with the engine's device current
cudaError_t dense_umma_init_device() {
    cudaError_t e = cudaFuncSetAttribute(dense_umma_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UmSmem));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dense_umma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UmSmem));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(dense_umma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UmSmem));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(dense_umma_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (SELK_CAP + SELK_RES) * 8);
}

template <int KIND>
static void um_launch(int grid, size_t smem, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tx, int64_t n, int n_kb,
                      int B, int qb0, int nqb, float4* seedv, int n_seed, int64_t rows_per_chunk, int share, const uint32_t* mask,
                      int k_sel, uint32_t idesc, uint64_t* lists, int* counts, uint32_t* gthr, uint32_t* coarse,
                      uint32_t* fine, const ShardBounds& xb, const UmmaI8& i8) {
    dense_umma_kernel<KIND><<<grid, UM_THREADS, smem, st>>>(tq, tx, n, n_kb, B, qb0, nqb, seedv, n_seed, rows_per_chunk, share,
                                                           mask, k_sel, idesc, lists, counts, gthr, coarse, fine, xb,
                                                           i8.xinv, i8.qinv, i8.xinv_max);
}

namespace {
struct UmScratch {
    uint64_t* lists; int* counts; uint32_t *gthr, *coarse, *fine; float4* seedv;
};
UmScratch um_scratch(void* scratch, const UmmaPlan& plan) {
    uint8_t* sp = reinterpret_cast<uint8_t*>(scratch);
    UmScratch u;
    u.lists = reinterpret_cast<uint64_t*>(sp);
    u.counts = reinterpret_cast<int*>(sp + plan.list_bytes);
    u.gthr = reinterpret_cast<uint32_t*>(sp + plan.list_bytes + plan.count_bytes);
    u.coarse = reinterpret_cast<uint32_t*>(sp + plan.list_bytes + plan.count_bytes + plan.gthr_bytes);
    u.fine = reinterpret_cast<uint32_t*>(sp + plan.list_bytes + plan.count_bytes + plan.gthr_bytes + plan.coarse_bytes);
    u.seedv = reinterpret_cast<float4*>(sp + plan.list_bytes + plan.count_bytes + plan.gthr_bytes + plan.coarse_bytes + plan.fine_bytes);
    return u;
}
}  // namespace

// query groups of <= UM_MAXQB blocks (1024 queries): each group is one pass over the corpus (scan, then tail; the
// candidate lists are reused by the next group)
int dense_umma_groups(const UmmaPlan& plan) { return (plan.n_qblocks + UM_MAXQB - 1) / UM_MAXQB; }

// bounds and histograms of every query of the batch: once per call, before the first group
cudaError_t dense_umma_clear(void* scratch, const UmmaPlan& plan, cudaStream_t st) {
    const UmScratch u = um_scratch(scratch, plan);
    return cudaMemsetAsync(u.gthr, 0, plan.gthr_bytes + plan.coarse_bytes + plan.fine_bytes, st);
}

cudaError_t dense_umma_scan(int dtype, int64_t n, int d, int B, const uint32_t* mask, int k_sel, void* scratch,
                            const UmmaPlan& plan, void* tmap_x, void* tmap_q, int group, cudaStream_t st, int* launches,
                            const ShardBounds& xb, const UmmaI8& i8) {
    const size_t smem = sizeof(UmSmem);
    CUtensorMap tq, tx;
    memcpy(&tq, tmap_q, sizeof tq);
    memcpy(&tx, tmap_x, sizeof tx);
    // instruction descriptor: accumulator format (f32 / s32), operand format, N, M (cute/arch/mma_sm100_desc.hpp layout)
    const int kind = dtype == 0 ? 1 : (dtype == 3 ? 2 : 0);
    const uint32_t cfmt = kind == 2 ? 2u : 1u;
    const uint32_t fmt = kind == 1 ? 2u : (kind == 2 ? 1u : (dtype == 2 ? 1u : 0u));   // TF32 | signed 8-bit | BF16 : F16
    const uint32_t idesc = (cfmt << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(UM_BN >> 3) << 17) | ((uint32_t)(UM_BM >> 4) << 24);
    const int n_kb = (int)((int64_t)d * um_esz(dtype) / 128);
    const UmScratch u = um_scratch(scratch, plan);
    auto launch = [&](int grid, int qb0, int nqb, float4* seedv, int n_seed, int64_t rows_per_chunk, int share, const ShardBounds& b) {
        if (kind == 0) um_launch<0>(grid, smem, st, tq, tx, n, n_kb, B, qb0, nqb, seedv, n_seed, rows_per_chunk, share, mask, k_sel, idesc, u.lists, u.counts, u.gthr, u.coarse, u.fine, b, i8);
        else if (kind == 1) um_launch<1>(grid, smem, st, tq, tx, n, n_kb, B, qb0, nqb, seedv, n_seed, rows_per_chunk, share, mask, k_sel, idesc, u.lists, u.counts, u.gthr, u.coarse, u.fine, b, i8);
        else um_launch<2>(grid, smem, st, tq, tx, n, n_kb, B, qb0, nqb, seedv, n_seed, rows_per_chunk, share, mask, k_sel, idesc, u.lists, u.counts, u.gthr, u.coarse, u.fine, b, i8);
    };
    cudaError_t err;
    const int64_t n_tiles = (n + UM_BN - 1) / UM_BN;
    const int n_seed = plan.n_seed, n_main = plan.n_chunks;
    const int64_t seed_stride = n_seed > 0 ? (n_tiles / n_seed) * UM_BN : 0;      // rows between seed tiles
    int nl = 0;
    const int qb0 = group * UM_MAXQB;
    const int nqb = std::min(UM_MAXQB, plan.n_qblocks - qb0);
    const int nq = std::min(B - qb0 * UM_BM, nqb * UM_BM);
    if (n_seed > 0) {
        launch(n_seed, qb0, nqb, u.seedv, n_seed, seed_stride, 1, ShardBounds());
        if ((err = cudaGetLastError()) != cudaSuccess) return err;
        um_seed_bound_kernel<<<nq, 256, 0, st>>>(u.seedv, n_seed, qb0 * UM_BM, qb0 * UM_BM + nq, k_sel, u.gthr);
        if ((err = cudaGetLastError()) != cudaSuccess) return err;
        nl += 2;
    }
    // CTAs sharing a row range (see the kernel): the query blocks split evenly, or (share = 2) the odd block by rows
    const bool splits = nqb % um_share() == 0 || (um_share() == 2 && nqb >= 3);
    const int share = (um_share() > 1 && splits && n_main % um_share() == 0 && n_tiles >= n_main) ? um_share() : 1;
    const int64_t rows_per_chunk = (n_tiles + n_main / share - 1) / (n_main / share) * UM_BN;
    launch(n_main, qb0, nqb, nullptr, 0, rows_per_chunk, share, xb);
    if ((err = cudaGetLastError()) != cudaSuccess) return err;
    ++nl;
    if (launches) *launches = nl;
    return cudaSuccess;
}

cudaError_t dense_umma_tail(int B, int k_sel, void* scratch, const UmmaPlan& plan, int group, const UmmaTail& t,
                            cudaStream_t st, const ShardBounds& xb) {
    if (k_sel > SELK_RES || t.k > SELK_RES) return cudaErrorInvalidValue;
    const UmScratch u = um_scratch(scratch, plan);
    const int qb0 = group * UM_MAXQB;
    const int nqb = std::min(UM_MAXQB, plan.n_qblocks - qb0);
    const int nq = std::min(B - qb0 * UM_BM, nqb * UM_BM);
    dense_umma_tail_kernel<<<nq, B <= 64 ? 1024 : 512, (SELK_CAP + SELK_RES) * 8, st>>>(
        u.lists, u.counts, u.coarse, u.fine, u.gthr, xb, plan.n_chunks, B, qb0, nqb, k_sel, t.X, t.dtype, t.n, t.d, t.Q,
        t.keys_out, t.k, t.id_offset, t.out_scores, t.out_ids, t.out_counts, t.sink);
    return cudaGetLastError();
}

}  // namespace brag
