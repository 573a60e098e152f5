/*
 * brag.h -- C ABI of the B200-native hybrid-retrieval engine (libbrag.so).
 *
 * This is the drop-in boundary for the hot path that the reference delegates to Qdrant:
 * every entry point below replaces one request the reference's vector-store adapter sends
 * through `qdrant_session.client` (reference: app/core/vector_store/qdrant/qdrant_handler.py).
 * The reference has no FFI of its own (it is Python talking HTTP to Qdrant), so the binding a
 * maintainer adds is the ctypes stub shown in INTEGRATION.md; the Python mirror of
 * AsyncQdrantClient that uses it is rag_application_b200/client.py.
 *
 * Conventions
 *  - plain pointers and sizes only; no exceptions cross the ABI; every call returns 0 on
 *    success, non-zero on failure, message via brag_last_error() (thread local).
 *  - `on_device` != 0 means every data pointer of that call is a device pointer on the
 *    engine's GPU (no copies, fully stream-ordered); 0 means host pointers (the call copies
 *    host->device, runs, copies results back and synchronises the stream before returning).
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *  - ids are int64 "global row ids" = local row + the engine's id offset (row sharding);
 *    global ids must stay below 2^32.  Result lists are padded with id -1 / score -inf and
 *    the true length is written to out_counts[b].
 *  - ranking order everywhere is (score desc, id asc).
 */
#ifndef BRAG_H
#define BRAG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct brag_engine brag_engine;

/* Storage of a dense space.  BRAG_I8 is the reference's "quantized" vector (qdrant_handler.py:144-146: int8 values held
 * in a COSINE space): rows are raw int8, d a multiple of 128, and the space ALWAYS scores by the cosine of the integer
 * vectors -- exact int32 dot, score = float32(dot / (sqrt(sum x^2) * sqrt(sum q^2))) evaluated in float64, 0 for a
 * zero row or query.  Its queries are float32 arrays holding integer values in [-128, 127] (what the reference
 * sends, :300-302); k <= 416. */
enum brag_dtype { BRAG_F32 = 0, BRAG_F16 = 1, BRAG_BF16 = 2, BRAG_I8 = 3 };
enum brag_rrf_tie { BRAG_TIE_ID = 0, BRAG_TIE_FIRST_SEEN = 1 };

/* Engine lifecycle.  Replaces AsyncQdrantClient(url=...) (qdrant_session.py:21). */
int brag_create(int device, brag_engine** out);
int brag_destroy(brag_engine* e);
const char* brag_last_error(void);
/* ABI version, for the loader's sanity check. */
int brag_version(void);

/* Row sharding: global id = local row + offset. */
int brag_set_id_offset(brag_engine* e, int64_t offset);

/* Corpus upload.  Replaces client.upsert(points) for one named dense vector
 * (qdrant_handler.py:155-190, schema :57-79).  rows = [n, d] row-major in `dtype`,
 * already L2-normalised by the caller when the space is COSINE.  The engine copies. */
int brag_dense_set(brag_engine* e, const char* name, const void* rows, int64_t n, int32_t d,
                   int32_t dtype, int32_t on_device);
/* SURVEY 8 row f2 (incremental upsert, qdrant_handler.py:190,259): append n_new rows to an existing dense space
 * without re-uploading it (capacity doubles when full; rows in the space's storage dtype, prepared like
 * brag_dense_set's), and overwrite existing rows in place (row_ids: host int64 [m] local rows; rows: host [m, d]).
 * Both synchronise the device and drop any filter-compacted copy of the space. */
int brag_dense_append(brag_engine* e, const char* name, const void* rows, int64_t n_new, int32_t d,
                      int32_t dtype, int32_t on_device);
int brag_dense_update(brag_engine* e, const char* name, const int64_t* row_ids, const void* rows, int64_t m);
int brag_dense_drop(brag_engine* e, const char* name);

/* Sparse corpus upload: term-major CSR of the sparse named vector (qdrant_handler.py:80-86,
 * :163).  term_ptr[V+1] int64, doc_ids[nnz] uint32 ascending inside each term, weights[nnz]
 * float32.  The engine copies and builds its per-sub-tile offset table. */
int brag_sparse_set_csr(brag_engine* e, const char* name, const int64_t* term_ptr,
                        const uint32_t* doc_ids, const float* weights, int64_t n_terms,
                        int64_t n_docs, int32_t on_device);
/* Append documents to a sparse space on the device (replaces the re-upload the reference's
 * client.upsert(points) of new points would otherwise cost: qdrant_handler.py:190,259).  The delta is a term-major CSR
 * over the GROWN vocabulary of n_terms_new rows: delta_term_ptr[n_terms_new + 1], doc_ids / weights per posting, every
 * document id in [current n_docs, n_docs_new) and strictly ascending within a term.  row_map[t] is the new row of the
 * space's current row t (strictly increasing), so a vocabulary kept in ascending external-index order stays in that
 * order and scores keep the summation order of a fresh upload.  Old posting lists are moved inside HBM, the delta
 * is spliced in after them and the scan indexes are rebuilt.  Host pointers only (the delta is small).  On failure
 * after the old index was released the space is dropped and must be uploaded again. */
int brag_sparse_append(brag_engine* e, const char* name, const int64_t* delta_term_ptr, const uint32_t* doc_ids,
                       const float* weights, const int32_t* row_map, int64_t n_terms_new, int64_t n_docs_new);
int brag_sparse_drop(brag_engine* e, const char* name);

/* Payload-filter bitmask (bit r set = row r passes).  Replaces query_filter=Filter(...)
 * (qdrant_handler.py:297,371).  mask_id >= 1; mask_id 0 in a query means "no filter". */
int brag_mask_set(brag_engine* e, int32_t mask_id, const uint32_t* bits, int64_t nbits,
                  int32_t on_device);

/* Exact dense top-k over the whole space: a leaf Prefetch(query=vec, using=name, limit=k)
 * or query_points(query=vec, using=name, limit=k, exact=True) (qdrant_handler.py:311-315,
 * :335-339, :363-369).  Q = [B, d] float32 (normalised by the caller for COSINE).
 * Scores are the fp32 re-scored values of the selected rows. */
int brag_dense_topk(brag_engine* e, const char* name, const float* Q, int32_t B, int32_t k,
                    int32_t mask_id, float* out_scores, int64_t* out_ids, int32_t* out_counts,
                    int32_t on_device, void* stream);

/* Candidate-restricted dense re-score: a Prefetch that has children
 * (qdrant_handler.py:317-329, :341-343, :363-372).  cand = [B, m] global ids, -1 = empty
 * slot, duplicates allowed (union semantics). */
int brag_dense_rescore(brag_engine* e, const char* name, const float* Q, int32_t B,
                       const int64_t* cand, int32_t m, int32_t k, int32_t mask_id,
                       float* out_scores, int64_t* out_ids, int32_t* out_counts,
                       int32_t on_device, void* stream);

/* Allocation-free variant of brag_dense_rescore.  It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same answers as brag_dense_rescore; device pointers only (Q, cand and the three outputs
 * live on the engine's GPU, Q 16-byte aligned), stream-ordered on `stream`, never synchronises, and calls no
 * cudaMalloc / cudaFree or any allocating routine: its [B][m] key scratch is a static device array of
 * BRAG_NOALLOC_MAX_KEYS keys (512 KB that every GPU loading the library holds until exit).  B * m beyond that limit is
 * refused with an error.  All calls on one GPU share the array: calls that overlap on the GPU (different streams,
 * different engines on the same device) race -- order them. */
#define BRAG_NOALLOC_MAX_KEYS 65536
int brag_dense_rescore_noalloc(brag_engine* e, const char* name, const float* Q, int32_t B,
                               const int64_t* cand, int32_t m, int32_t k, int32_t mask_id,
                               float* out_scores, int64_t* out_ids, int32_t* out_counts, void* stream);

/* Sparse (BM25) top-k: Prefetch(query=SparseVector, using=name, limit=k)
 * (qdrant_handler.py:347-354).  Query b owns q_terms/q_vals[q_ptr[b]:q_ptr[b+1]], term rows
 * strictly ascending; q_vals are float64 (Python floats; IDF already applied by the host).
 * Only documents sharing >=1 term are candidates.
 * max_terms: an upper bound of the longest query of the batch, in terms (it selects the kernel variant); -1 = let the
 * engine find it.  With host pointers that is a loop over q_ptr; with device pointers it costs a device->host copy of
 * q_ptr and a stream synchronisation in the middle of the call, so stream-ordered callers pass the bound (ABI 2).
 * With device pointers the bound is the caller's promise: terms of a query beyond it are ignored. */
int brag_sparse_topk(brag_engine* e, const char* name, const int32_t* q_ptr,
                     const int32_t* q_terms, const double* q_vals, int32_t B, int32_t k,
                     int32_t mask_id, int32_t max_terms, float* out_scores, int64_t* out_ids,
                     int32_t* out_counts, int32_t on_device, void* stream);
/* Document frequency of each term row (for Modifier.IDF on the host): df[V] int64, host. */
int brag_sparse_df(brag_engine* e, const char* name, int64_t* df_host, int64_t n_terms);

/* Reciprocal rank fusion: Prefetch(prefetch=[...], query=FusionQuery(RRF))
 * (qdrant_handler.py:357-360).  ids = [L, B, kmax], counts = [L, B]; score = sum 1/(2+rank)
 * in float64; out_scores float64 [B, limit].  L <= 16, L * kmax <= 8192, limit <= 8192. */
int brag_rrf(brag_engine* e, const int64_t* ids, const int32_t* counts, int32_t L, int32_t B,
             int32_t kmax, int32_t limit, int32_t tie_mode, double* out_scores,
             int64_t* out_ids, int32_t* out_counts, int32_t on_device, void* stream);

/* Rerank of the fused hits, everything after the encoder (replaces the tensor code of
 * rerank_documents, app/core/models/huggingface/huggingface.py:183-189, reached from qdrant_handler.py:401):
 * query_hidden [Tq, d] and doc_hidden [n, T, d] are the encoder's last hidden states (float32, padding positions
 * included, as the reference pools them); scores[i] = <mean_t query_hidden, mean_t doc_hidden[i]> and out_order is
 * np.argsort(scores)[::-1] with equal scores ordered higher index first.  1 <= n <= 4096. */
int brag_rerank_meanpool(brag_engine* e, const float* query_hidden, int32_t Tq, const float* doc_hidden, int32_t n,
                         int32_t T, int32_t d, int32_t* out_order, float* out_scores, int32_t on_device, void* stream);

/* k-way merge of per-shard top-k lists after the all-gather:
 * scores/ids = [G, B, k], counts = [G, B] -> [B, k]. */
int brag_merge_topk(brag_engine* e, const float* scores, const int64_t* ids,
                    const int32_t* counts, int32_t G, int32_t B, int32_t k, float* out_scores,
                    int64_t* out_ids, int32_t* out_counts, int32_t on_device, void* stream);

/* The north_star three-stage path in one call: dense top-k, sparse top-k, RRF(limit=k).
 * max_terms as in brag_sparse_topk. */
int brag_hybrid_topk(brag_engine* e, const char* dense_name, const char* sparse_name,
                     const float* Q, const int32_t* q_ptr, const int32_t* q_terms,
                     const double* q_vals, int32_t B, int32_t k, int32_t mask_id, int32_t max_terms,
                     double* out_scores, int64_t* out_ids, int32_t* out_counts,
                     int32_t on_device, void* stream);

/* ---- Row sharding across the GPUs of one box (the `brag_allgather_merge` of SURVEY 8b).
 * One engine per GPU, each holding the rows [id_offset, id_offset + n) of every space; a query batch is answered by
 * all of them.  A stage's per-shard list travels as sorted 64-bit keys, ordered(score) << 32 | (0xFFFFFFFF - global
 * row), 0 = empty slot: [2 lists (dense, sparse)][B][k].
 *
 * Collective variant (any transport): brag_shard_local_lists on every rank -> all-gather of the key arrays ->
 * brag_shard_fuse on every rank (merge of the G dense lists, merge of the G sparse lists, RRF of the two; one kernel;
 * it relies on what brag_shard_local_lists writes: every list sorted best first, empties last, rows disjoint across shards).
 * stage_* (all three or none): the two merged lists, [2][B][k] scores / ids and [2][B] counts. */
int brag_shard_local_lists(brag_engine* e, const char* dense_name, const char* sparse_name, const float* Q,
                           const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals, int32_t B, int32_t k,
                           int32_t mask_id, int32_t max_terms, uint64_t* out_keys, int32_t on_device, void* stream);
int brag_shard_fuse(brag_engine* e, const uint64_t* keys /* [G][2][B][k] */, int32_t G, int32_t B, int32_t k,
                    int32_t limit, int32_t tie_mode, double* out_scores, int64_t* out_ids, int32_t* out_counts,
                    float* stage_scores, int64_t* stage_ids, int32_t* stage_counts, int32_t on_device, void* stream);
/* Peer variant: the exchange is fused into the tail kernels over NVLink peer memory -- no collective call.
 * brag_shard_window_create allocates this engine's exchange window (for batches <= max_batch, lists <= max_k;
 * world * max_k <= 8192) and returns its 64-byte CUDA IPC handle; the caller carries the handles of all ranks to every
 * rank (any host transport) and calls brag_shard_connect(rank, world, handles [world][64]).  Engines living in ONE
 * process pass local_peers (the world's engine handles, in rank order) instead and handles = NULL.
 * brag_shard_hybrid_topk must then be called by every rank with the same batch, in the same order (it counts epochs):
 * each rank pushes its two lists straight into every peer's window while it scans, tells its score bounds to the peers
 * (any shard's k-th best is a lower bound of the global k-th best), and one kernel per rank waits for the world's
 * lists, merges and fuses them.  The result is identical on every rank.  The wait is bounded (a peer that never
 * arrives becomes a CUDA error after 8 s).  With on_device = 0 the call synchronises: engines of one process must
 * then be driven from one thread each, or with device pointers. */
#define BRAG_SHARD_HANDLE_BYTES 64
int brag_shard_window_create(brag_engine* e, int32_t world, int32_t max_batch, int32_t max_k, void* handle_out);
int brag_shard_connect(brag_engine* e, int32_t rank, int32_t world, const void* handles,
                       brag_engine* const* local_peers);
int brag_shard_hybrid_topk(brag_engine* e, const char* dense_name, const char* sparse_name, const float* Q,
                           const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals, int32_t B, int32_t k,
                           int32_t mask_id, int32_t max_terms, double* out_scores, int64_t* out_ids,
                           int32_t* out_counts, float* stage_scores, int64_t* stage_ids, int32_t* stage_counts,
                           int32_t on_device, void* stream);
int brag_shard_close(brag_engine* e);

/* Index persistence (the reference keeps Qdrant's index on a volume, docker-compose.yml:30-31, and
 * reopens a user's collection at start-up, qdrant_handler.py:94-100).  brag_save writes every dense and sparse space of
 * the engine -- rows, postings AND the scan indexes built from them -- to one flat file (header, section table,
 * 4096-byte-aligned raw arrays; written to path.tmp and renamed); brag_load puts them back into HBM with nothing to
 * rebuild and restores the id offset.  Spaces of the same name are replaced, others stay.  Filter masks and
 * filter-compacted copies are query state and are not saved.  One file per engine, i.e. per shard. */
int brag_save(brag_engine* e, const char* path);
int brag_load(brag_engine* e, const char* path);

/* Per-stage device timing (CUDA events on the call's stream).  Stages: 0 dense scan,
 * 1 dense select+rescore, 2 sparse scan, 3 sparse select, 4 rrf, 5 merge / sharded signal + fuse (includes the wait
 * for the slowest peer).
 * brag_profile_read returns accumulated ms and launch counts since the last reset and
 * synchronises the device. */
#define BRAG_N_STAGES 6
int brag_profile_enable(brag_engine* e, int32_t on);
int brag_profile_read(brag_engine* e, double* ms, int64_t* launches, int64_t* kernel_launches_total);
int brag_profile_reset(brag_engine* e);
/* Which dense scan kernel brag_dense_topk would pick for batch B: 0 = GEMV stream (float32 storage, or
 * d not a multiple of 64), 1 = TMA + tcgen05 (16-bit storage, any batch size). */
int brag_dense_path(brag_engine* e, const char* name, int32_t B);
/* Force a path (-1 auto, 0 GEMV, 1 tcgen05, 2 tcgen05 scanning the full space under a filter instead of its
 * filter-compacted copy) -- used by the parity tests. */
int brag_force_dense_path(brag_engine* e, int32_t path);
/* Sparse scan kernel (-1 auto, 0 scatter-accumulate, 1 bitmap filter).  Auto = bitmap filter unless a query of
 * the batch has more than 32 terms or the space holds non-finite weights; both give identical results. */
int brag_force_sparse_path(brag_engine* e, int32_t path);
/* Filtered sparse scans: once one filter (mask_id, contents) has served `queries` queries of a sparse space and keeps
 * at most half of its documents, an index of just the passing documents is built on the device and used for that
 * filter from then on (results identical).  Only calls with B >= 32 count towards `queries` (single queries are
 * latency-bound either way).  Default 256; 0 = build on first use; < 0 = never (an index already built is bypassed). */
int brag_sparse_compact_after(brag_engine* e, int64_t queries);
/* Per-(term, 2048-document sub-tile) posting offsets are kept only for terms with at least min_df postings; the short
 * lists of the other terms are searched by the scans (results identical).  0 = automatic: a row only where the term's
 * postings outweigh it (df >= (n_docs / 2048 + 2) / 2), so the table never exceeds the postings whatever the
 * vocabulary.  Applies to sparse spaces uploaded afterwards; the parity tests force the searched form with a large
 * value. */
int brag_sparse_offsets_min_df(brag_engine* e, int64_t min_df);
/* Which filter the space's compacted index was built for (0 = none built) and how many documents it holds. */
int brag_sparse_compact_info(brag_engine* e, const char* name, int32_t* mask_id, int64_t* n_docs);

/* ---- the sparse encoder the reference runs before it calls the vector store
 * (fastembed SparseTextEmbedding("Qdrant/bm25").embed, embedding_handler.py:41,123-138).  Host-only code.
 * tokens: the document after fastembed's normalisation (non-word characters -> space, lower-case, whitespace split),
 * concatenated; token i = tokens[tok_off[i] .. tok_off[i+1]).  Drops punctuation tokens, English stop words and
 * tokens longer than max_token_len characters, stems (Snowball English), id = |murmur3_x86_32(stem, seed 0)|,
 * value = tf*(k1+1)/(tf + k1*(1-b+b*len/avgdl)) (query_mode != 0: value = 1, fastembed's query_embed).
 * out_ids / out_vals need n_tokens slots; ids come in first-occurrence order. */
int brag_bm25_encode(const char* tokens, const int32_t* tok_off, int32_t n_tokens, double k1, double b,
                     double avgdl, int32_t max_token_len, int32_t query_mode, int64_t* out_ids,
                     double* out_vals, int32_t* out_n);
/* Snowball English (Porter2) stem of one lower-case word; MurmurHash3 x86_32 -- exposed for the known-answer tests. */
int brag_bm25_stem(const char* word, char* out, int32_t cap);
uint32_t brag_murmur3_32(const void* data, int32_t len, uint32_t seed);

#ifdef __cplusplus
}
#endif
#endif /* BRAG_H */
