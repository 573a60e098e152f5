// Internal launcher declarations (host side of each kernel family).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace brag {

struct ShardBounds;
// Destination of a sharded call's ranked lists (shard.cu): every rank's peer-mapped window.
struct ShardPush {
    uint8_t* const* windows = nullptr;    // device table [world] of window bases (windows[rank] = own)
    const unsigned* epoch = nullptr;
    int world = 0, rank = 0;
    size_t slots_off = 0, slot_bytes = 0, flags_off = 0;
};
// What a scan's tail kernel does with a query's final list in a sharded call: the best k keys, rows made global, into
// local_out [2][B][k] (collective variant) or into slot (epoch parity, rank) of every rank's window (shard_push.cuh).
struct ShardSink {
    int on = 0;
    int list = 0;                          // 0 dense, 1 sparse
    uint64_t* local_out = nullptr;
    ShardPush push;
};     // common.cuh: cross-shard bound exchange of a sharded call (default = none)

// Per-device kernel attributes (opt-in shared memory): brag_create calls these with the engine's device current, so a
// second engine on another GPU of the same process gets them too.
cudaError_t dense_umma_init_device();
cudaError_t select_init_device();
// allocation-free candidate re-score (select.cu): key scratch in a static device array of NOALLOC_MAX_KEYS keys
constexpr int NOALLOC_MAX_KEYS = 65536;     // B * m of one call; 512 KB per GPU
cudaError_t rescore_noalloc_init_device(uint64_t** keys);
cudaError_t dense_rescore_noalloc(uint64_t* keys, const void* X, int dtype, int64_t n, int d, const float* Q, int B,
                                  const int64_t* cand, int m, int k, int64_t id_offset, const uint32_t* mask,
                                  float* out_scores, int64_t* out_ids, int32_t* out_counts, cudaStream_t st);
cudaError_t bm25_init_device();
cudaError_t bm25x_init_device();
cudaError_t shard_init_device();

// ---- dense_gemv.cu
cudaError_t dense_gemv_scan(const void* X, int dtype, int64_t n, int d, const float* Q, int B,
                            const uint32_t* mask, int k_sel, uint64_t* part, int grid, cudaStream_t st,
                            int* launches);

// ---- dense_umma.cu
struct UmmaPlan {
    int n_qblocks;          // ceil(B / 128)
    int n_chunks;           // persistent CTAs = corpus chunks = candidate lists per query
    int n_seed;             // single-tile CTAs of the seed launch (0 = no seeding)
    size_t list_bytes, count_bytes, gthr_bytes, coarse_bytes, fine_bytes, seed_bytes;
    size_t scratch_bytes;   // lists | counts | per-query bounds | coarse + fine score histograms
};
// dtype 0 float32 (kind::tf32), 1 / 2 float16 / bfloat16 (kind::f16), 3 int8 (kind::i8, exact int32 dots)
bool dense_umma_supported(int dtype, int d, int64_t n);
// int8 spaces: scores are cosines of integer vectors, idot * xinv[row] * qinv[query]
struct UmmaI8 {
    const float* xinv = nullptr;    // [n] inverse norms of the stored rows
    const float* qinv = nullptr;    // [B] inverse norms of the (quantised) queries
    float xinv_max = 0.f;           // largest xinv over the space
};
// longest candidate list (k + selection margin) the tcgen05 scan keeps per (CTA, query); longer ones take the GEMV scan
int dense_umma_max_ksel();
cudaError_t dense_umma_plan(int64_t n, int d, int B, int k_sel, int num_sms, UmmaPlan* plan);
// A call = dense_umma_clear once, then for every group of <= 1024 queries (dense_umma_groups): dense_umma_scan (seed
// launch + main launch over the whole space; tmap_x / tmap_q: tensor maps of the [n, d] rows and of the [Bpad, d] scan
// copy of the queries, Bpad = n_qblocks * 128, zero padded) and dense_umma_tail (final bound, selection, exact re-score,
// and either the re-scored keys or the final list).
int dense_umma_groups(const UmmaPlan& plan);
cudaError_t dense_umma_clear(void* scratch, const UmmaPlan& plan, cudaStream_t st);
cudaError_t dense_umma_scan(int dtype, int64_t n, int d, int B, const uint32_t* mask, int k_sel, void* scratch,
                            const UmmaPlan& plan, void* tmap_x, void* tmap_q, int group, cudaStream_t st, int* launches,
                            const ShardBounds& xb, const UmmaI8& i8);
struct UmmaTail {
    const void* X = nullptr;        // the rows that were scanned, for the exact re-score
    int dtype = 0;
    int64_t n = 0;
    int d = 0;
    const float* Q = nullptr;       // [B, d] the caller's float32 queries
    uint64_t* keys_out = nullptr;   // [B][k_sel] re-scored keys with local rows; NULL = write the final list below
    ShardSink sink;                 // sharded call (sink.on): the final list goes to the windows instead
    int k = 0;
    int64_t id_offset = 0;
    float* out_scores = nullptr;    // [B][k]
    int64_t* out_ids = nullptr;     // [B][k]
    int32_t* out_counts = nullptr;  // [B]
};
cudaError_t dense_umma_tail(int B, int k_sel, void* scratch, const UmmaPlan& plan, int group, const UmmaTail& t,
                            cudaStream_t st, const ShardBounds& xb);
// builds a CUtensorMap (128 bytes) for a [rows, d] row-major matrix of `dtype`, box = [box_rows, one 128-byte K slice]
cudaError_t make_tmap_2d(void* tmap_out, const void* base, int64_t rows, int d, int box_rows, int dtype);

// ---- select.cu
// multi-level merge of P key lists per query: in [P][B][kk] -> out [B][kout] sorted desc
// bound (optional, [B], f32_ordered units): keys whose score image is below it are dropped before sorting
cudaError_t merge_key_lists(const uint64_t* in, int P, int B, int kk, uint64_t* tmp_a, uint64_t* tmp_b,
                            uint64_t* out, int kout, cudaStream_t st, int* launches, const uint32_t* bound = nullptr);
// tail of a sparse scan (select.cu): final bound + merge of the P lists [P][B][kk] + the final list (keys_out NULL) or
// its best k keys with local rows (keys_out [B][k])
cudaError_t sparse_tail(const uint64_t* part, int P, int B, int kk, const uint32_t* coarse, const uint32_t* fine,
                        uint32_t* gthr, const ShardBounds& xb, int k, uint64_t* keys_out, int64_t id_offset,
                        float* out_scores, int64_t* out_ids, int32_t* out_counts, const ShardSink& sink, cudaStream_t st);
size_t merge_tmp_elems(int P, int B, int kk);
// the tensor-core scan's copy of the queries in the stored element type, [Bpad, d] with zero padding rows
// (dtype 3 also writes qinv[B], the inverse norms of the int8 queries)
cudaError_t scan_queries(const float* Q, int B, int d, int dtype, void* out, int Bpad, float* qinv, cudaStream_t st);
// int8 spaces: xinv[n] inverse row norms; *max_bits = atomicMax of their bit patterns (caller zeroes it first)
cudaError_t i8_row_inv_norms(const void* X, int64_t n, int d, float* xinv, uint32_t* max_bits, cudaStream_t st);
// keys_in [B][m] (rows from a scan) -> re-scored keys (f64 accumulate, f32 score), 0 stays 0
cudaError_t rescore_keys(const void* X, int dtype, int64_t n, int d, const float* Qs, int B,
                         const uint64_t* keys_in, int m, uint64_t* keys_out, cudaStream_t st);
// cand [B][m] global ids (-1 empty) -> scored keys with mask filter; duplicates keep one copy
cudaError_t rescore_cands(const void* X, int dtype, int64_t n, int d, const float* Qs, int B,
                          const int64_t* cand, int m, int64_t id_offset, const uint32_t* mask,
                          uint64_t* keys_out, cudaStream_t st);
// per query: sort m keys desc (optionally dropping duplicates), write first k to outputs
cudaError_t finalize_keys(const uint64_t* keys, int B, int m, int k, bool dedup, int64_t id_offset,
                          float* out_scores, int64_t* out_ids, int32_t* out_counts, cudaStream_t st);
cudaError_t rrf_fuse(const int64_t* ids, const int32_t* counts, int L, int B, int kmax, int limit,
                     int tie_mode, double* out_scores, int64_t* out_ids, int32_t* out_counts,
                     cudaStream_t st);
cudaError_t merge_scored_lists(const float* scores, const int64_t* ids, const int32_t* counts, int G,
                               int B, int k, float* out_scores, int64_t* out_ids, int32_t* out_counts,
                               cudaStream_t st);

// ---- shard.cu: row sharding across the GPUs of one box (exchange fused into the tail kernels)
// One window per engine (a single cudaMalloc, peer-mapped by every other rank), same layout on every rank.
struct ShardWindowLayout {
    int world, max_batch, max_k;
    size_t slot_bytes;      // one rank's lists of one call: [2 lists][max_batch][max_k] keys
    size_t slots_off;       // [2 epoch parities][world] slots
    size_t flags_off;       // [2 epoch parities][world] u64: epoch whose lists of that rank have arrived
    size_t xb_off;          // [2 lists][max_batch] u64 cross-shard bound words (common.cuh: ShardBounds)
    size_t epoch_off;       // u32: the current epoch of this rank (bumped on the device: graph replays stay in step)
    size_t tables_off;      // [3][world] device pointers: window bases, dense bound words, sparse bound words
    size_t total;
};
ShardWindowLayout shard_window_layout(int world, int max_batch, int max_k);
cudaError_t shard_next_epoch(unsigned* epoch, cudaStream_t st);
// keys [B][m] (local rows) -> best k per query as keys with global rows, into local_out [2][B][k] (list `list`) or,
// when local_out is NULL, into slot (epoch parity, rank) of every rank's window
cudaError_t shard_finalize_push(const uint64_t* keys, int B, int m, int k, int64_t id_offset, int list,
                                uint64_t* local_out, const ShardPush& dst, cudaStream_t st);
cudaError_t shard_signal(const ShardPush& dst, cudaStream_t st);
// gathered [G][2][B][k] keys, or NULL = read this rank's window after waiting for the world's flags;
// merged lists (optional stage_* outputs, [2][B][k] / [2][B]) and their RRF fusion
cudaError_t shard_fuse(const uint64_t* gathered, const ShardPush& win, int G, int B, int k, int limit, int tie_mode,
                       double* out_scores, int64_t* out_ids, int32_t* out_counts, float* stage_scores,
                       int64_t* stage_ids, int32_t* stage_counts, cudaStream_t st);

// ---- rerank.cu: mean-pooled hidden states -> scores [n] and the document order (score desc, higher index first)
cudaError_t rerank_meanpool(const float* qh, int Tq, const float* dh, int n, int T, int d, float* scores,
                            int32_t* order, cudaStream_t st);

// ---- compact.cu: filter-compacted copy of a dense space (rows that pass a filter, gathered contiguously)
size_t compact_tmp_bytes(int64_t nwords);
// prefix: [ceil(n/32) + 1] passing rows before each filter word; the last entry is the number of passing rows
cudaError_t compact_prefix(const uint32_t* mask, int64_t n, uint32_t* prefix, void* tmp, size_t tmp_bytes,
                           cudaStream_t st);
// rowmap: [nc] original row of every compact row (increasing); out: [nc, row_bytes]
cudaError_t compact_rows(const void* X, int row_bytes, const uint32_t* mask, int64_t n, const uint32_t* prefix,
                         int64_t nc, uint32_t* rowmap, void* out, cudaStream_t st);
// keys of compact rows -> keys of original rows (same order: rowmap is increasing)
cudaError_t remap_keys(uint64_t* keys, int64_t count, const uint32_t* rowmap, cudaStream_t st);
cudaError_t compact_rowmap(const uint32_t* mask, int64_t n, const uint32_t* prefix, uint32_t* rowmap, cudaStream_t st);
// brag_sparse_append: old posting i of term t -> out[dst_start[t] + i - term_ptr[t]]; delta posting j -> out[ddst[j]]
cudaError_t sparse_append_move(const int64_t* term_ptr, const uint2* post, int64_t V, int64_t nnz,
                               const int64_t* dst_start, const uint2* dpost, const int64_t* ddst, int64_t dnnz,
                               uint2* out, cudaStream_t st);
// filter-compacted copy of a sparse space: postings of the passing documents, renumbered (see compact.cu)
size_t sparse_compact_tmp_bytes(int64_t V, int64_t nnz);
int64_t sparse_compact_chunks(int64_t nnz);
cudaError_t sparse_compact_count(const int64_t* term_ptr, const uint2* post, int64_t V, int64_t nnz,
                                 const uint32_t* mask, int64_t* new_term_ptr, int64_t* chunk_off, void* tmp,
                                 size_t tmp_bytes, cudaStream_t st);
cudaError_t sparse_compact_select(const uint2* post, int64_t nnz, const uint32_t* mask, const uint32_t* prefix,
                                  const int64_t* chunk_off, uint2* out, int64_t nnz_c, cudaStream_t st);

// ---- bm25.cu
constexpr int BM25_SUB = 2048;  // documents per warp sub-tile
// postings are stored interleaved: post[i] = (doc id, float32 weight bits)
cudaError_t bm25_interleave(const uint32_t* doc, const float* w, int64_t nnz, uint2* post, cudaStream_t st);
// sub_off: [n_rows][n_docs / 2048 + 2] first posting of every 2048-document sub-tile, for the terms row_terms[0..n_rows)
// (the terms whose postings outweigh such a row; the others are searched, bm25_common.cuh)
cudaError_t bm25_build_suboffsets(const int64_t* term_ptr, const uint2* post, const int32_t* row_terms, int64_t n_rows,
                                  int64_t n_docs, uint32_t* sub_off, cudaStream_t st);
struct Bm25Plan {
    int splits; int lists_per_query; int grid; int workers;
    int visits_per_list;    // tiles a worker warp visits per work item (bm25x: the first one runs without a bound)
    int seed_tiles = 0;     // bm25x: tiles scanned per query by the seed launch (one warp per query, extra list)
    int seed_qpi = 0;       // bm25x: queries per seed work item
    size_t gthr_bytes, coarse_bytes, fine_bytes, scratch_bytes;
};
Bm25Plan bm25_plan(int64_t n_docs, int B, int k_sel, int num_sms);
// part: [lists_per_query][B][k_sel]; scratch: plan.scratch_bytes (work counter | bounds | histograms)
cudaError_t bm25_scan(const int64_t* term_ptr, const uint2* post, const uint32_t* sub_off, const int32_t* sub_row,
                      int64_t V, int64_t n_docs, const int32_t* q_ptr, const int32_t* q_terms,
                      const double* q_vals, int B, const uint32_t* mask, int k_sel, uint64_t* part,
                      const Bm25Plan& plan, void* scratch, cudaStream_t st, const ShardBounds& xb);


// ---- bm25x.cu: bitmap-filter variant of the sparse scan (same results, prunes on per-term weight bounds)
// wrange[t] = (min, max) weight of term t; tau[t] = low/high class threshold (>= ~70 % of the postings have
// w <= tau); *nonfinite_flag is set when any weight is NaN/Inf
cudaError_t bm25x_build_wrange(const int64_t* term_ptr, const uint2* post, int64_t V, int64_t nnz, float2* wrange,
                               float* tau, int* nonfinite_flag, cudaStream_t st);
// bits / bits_hi: [n_dense][bm25x_row_words(n_docs)] words; bit d of row r = term dense_terms[r] occurs in
// document d (bits) with a weight above tau (bits_hi)
cudaError_t bm25x_build_bitmaps(const int64_t* term_ptr, const uint2* post, const int32_t* dense_terms, int n_dense,
                                const float* tau, int64_t n_docs, uint32_t* bits, uint32_t* bits_hi, cudaStream_t st);
int64_t bm25x_row_words(int64_t n_docs);     // words per stored map row (sub-tile count rounded up to 4)
bool bm25x_supported(int max_terms, int k_sel);
Bm25Plan bm25x_plan(int64_t n_docs, int B, int k_sel, int max_terms, int num_sms);
// bm_row[t] = bitmap row of term t or -1; max_terms = longest query of the batch (<= 32)
cudaError_t bm25x_scan(const int64_t* term_ptr, const uint2* post, const uint32_t* sub_off, const int32_t* sub_row,
                       const int32_t* bm_row,
                       const uint32_t* bits, const uint32_t* bits_hi, const float2* wrange, const float* tau, int64_t V,
                       int64_t n_docs, const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals, int B,
                       int max_terms, const uint32_t* mask, int k_sel, uint64_t* part, const Bm25Plan& plan,
                       void* scratch, cudaStream_t st, const ShardBounds& xb);

}  // namespace brag
