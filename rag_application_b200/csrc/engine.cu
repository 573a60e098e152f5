// C ABI of libbrag.so (include/brag.h): engine handle, corpus upload, stage orchestration.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/brag.h"
#include "common.cuh"
#include "kernels.h"

namespace {

thread_local std::string g_err;

struct BragError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct CudaOom : BragError {
    using BragError::BragError;
};

#define CK(expr)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            char _b[512];                                                                          \
            snprintf(_b, sizeof _b, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            if (_e == cudaErrorMemoryAllocation) { cudaGetLastError(); throw CudaOom(_b); }         \
            throw BragError(_b);                                                                   \
        }                                                                                          \
    } while (0)

#define REQUIRE(cond, msg)                \
    do {                                  \
        if (!(cond)) throw BragError(msg); \
    } while (0)

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void* ensure(size_t bytes) {
        if (bytes > cap) {
            if (p) CK(cudaFree(p));
            p = nullptr;
            cap = 0;
            size_t want = std::max(bytes, (size_t)256);
            CK(cudaMalloc(&p, want));
            cap = want;
        }
        return p;
    }
    template <typename T> T* as(size_t n) { return reinterpret_cast<T*>(ensure(n * sizeof(T))); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct DenseSpace {
    void* x = nullptr;
    int64_t n = 0;
    int64_t cap = 0;               // rows allocated (brag_dense_append grows by doubling)
    int d = 0;
    int dtype = 0;
    bool umma_ok = false;
    alignas(64) unsigned char tmap_x[128];
    // int8 spaces (BRAG_I8): inverse row norms for the scan's epilogue, and the largest of them
    float* xinv = nullptr;
    int64_t xinv_cap = 0;
    float xinv_max = 0.f;
    float* cp_xinv = nullptr;
    int64_t cp_xinv_cap = 0;
    float cp_xinv_max = 0.f;
    // filter-compacted copy (compact.cu) of the rows passing the filter (mask_id, version) last scanned with
    int cp_mask_id = 0;
    uint64_t cp_version = 0;
    bool cp_use = false;           // false: the filter keeps too many / too few rows, scan the full space
    int64_t cp_n = 0;
    void* cp_x = nullptr;
    uint32_t* cp_rowmap = nullptr;
    size_t cp_x_cap = 0, cp_rowmap_cap = 0;
    alignas(64) unsigned char cp_tmap[128];
};

struct SparseSpace {
    int64_t* term_ptr = nullptr;
    uint2* post = nullptr;      // interleaved (doc id, weight bits)
    uint32_t* sub_off = nullptr;   // [n_sub][n_docs / 2048 + 2] offsets of the terms that have a row (bm25_common.cuh)
    int32_t* sub_row = nullptr;    // [V] row of a term in sub_off, or -1 (rare term: its short list is searched)
    int64_t n_sub = 0;
    // bitmap-filter scan (bm25x.cu): per-term weight range, document maps of the dense terms
    float2* wrange = nullptr;
    float* tau = nullptr;
    uint32_t* bits_hi = nullptr;
    int32_t* bm_row = nullptr;
    uint32_t* bits = nullptr;
    int n_dense = 0;
    bool bitmap_ok = false;
    int64_t V = 0, n_docs = 0, nnz = 0;
    std::vector<int64_t> term_ptr_host;
    // filter-compacted copy (compact.cu): an index of the documents passing filter (cp_mask_id, cp_version), built
    // once that filter has served cp_queries queries of this space
    SparseSpace* cp = nullptr;
    int cp_mask_id = 0;
    uint64_t cp_version = 0;
    int64_t cp_queries = 0;
    int cp_state = 0;              // 0 counting queries, 1 built and in use, 2 not worth it / did not fit
    uint32_t* cp_rowmap = nullptr;
};

struct MaskBuf {
    uint32_t* bits = nullptr;
    int64_t nbits = 0;
    size_t cap = 0;                // words allocated (the buffer is reused by later brag_mask_set calls that fit)
    uint64_t version = 0;          // bumped by every brag_mask_set: invalidates compacted copies
};

struct StageEvent {
    int stage;
    cudaEvent_t a, b;
};

// Row sharding (shard.cu): this rank's window, the peers' mappings of theirs, and the device tables the kernels read.
struct ShardState {
    bool connected = false;
    int rank = 0, world = 1;
    bool bounds = true;                    // exchange score bounds across the shards while scanning
    brag::ShardWindowLayout lay{};
    uint8_t* window = nullptr;             // own (cudaMalloc)
    uint8_t* peer[16] = {nullptr};         // every rank's window as seen from this device (peer[rank] = window)
    bool ipc_open[16] = {false};
};

// Where a stage's final list goes in a sharded call: packed keys with global rows, to a local buffer or into the
// world's windows -- and which cross-shard bound words its scan exchanges.
struct ListSink {
    int list = 0;                          // 0 dense, 1 sparse
    uint64_t* local_out = nullptr;         // [2][B][k] keys (collective variant); NULL = push into the windows
    brag::ShardPush push{};
    brag::ShardBounds xb{};
};

}  // namespace

struct brag_engine {
    int device = 0;
    int num_sms = 148;
    int64_t id_offset = 0;
    int force_path = -1;
    bool no_compact = false;       // parity tests: scan the full space under a filter
    int64_t suboff_min_df = 0;            // 0 = automatic; tests force rare-term searches with a large value
    int64_t sparse_compact_after = 256;   // queries under one filter before its sparse copy is built; < 0 never
    uint64_t mask_epoch = 0;
    int force_sparse_path = -1;    // -1 auto, 0 scatter-accumulate (bm25.cu), 1 bitmap filter (bm25x.cu)
    std::map<std::string, DenseSpace> dense;
    std::map<std::string, SparseSpace> sparse;
    std::map<int, MaskBuf> masks;
    // scratch (grow-only)
    DevBuf qs, q16, part, tmp_a, tmp_b, keys1, keys2, umma_scratch, work_counter, g_thr;
    DevBuf h_in[6], h_out[6];      // staging for host-pointer calls
    DevBuf hyb_ids, hyb_counts, hyb_scores, cp_prefix, cp_tmp, cp_chunks, shard_keys;
    ShardState shard;
    uint64_t* noalloc_keys = nullptr;      // device address of select.cu's static key scratch on this engine's GPU
    std::vector<int32_t> qptr_host;
    alignas(64) unsigned char tmap_q[128];
    // profiling
    bool prof = false;
    std::vector<StageEvent> events;
    std::vector<cudaEvent_t> pool;
    double ms[BRAG_N_STAGES] = {0};
    int64_t stage_launches[BRAG_N_STAGES] = {0};
    int64_t kernel_launches = 0;
};

namespace {

struct StageTimer {
    brag_engine* e;
    int stage;
    cudaStream_t st;
    cudaEvent_t a = nullptr, b = nullptr;
    StageTimer(brag_engine* e_, int stage_, cudaStream_t st_) : e(e_), stage(stage_), st(st_) {
        e->stage_launches[stage]++;
        if (!e->prof) return;
        a = get();
        b = get();
        cudaEventRecord(a, st);
    }
    cudaEvent_t get() {
        if (!e->pool.empty()) { cudaEvent_t ev = e->pool.back(); e->pool.pop_back(); return ev; }
        cudaEvent_t ev;
        CK(cudaEventCreate(&ev));
        return ev;
    }
    ~StageTimer() {
        if (!e->prof || !a) return;
        cudaEventRecord(b, st);
        e->events.push_back({stage, a, b});
    }
};

void use_device(brag_engine* e) { CK(cudaSetDevice(e->device)); }

int k_select(int k) { return std::min(512, ((k + 32) + 3) / 4 * 4); }
inline int dtype_bytes(int dtype) { return dtype == 0 ? 4 : (dtype == 3 ? 1 : 2); }

// inverse row norms of an int8 matrix (grow-only buffer) and their maximum
void i8_norms(const void* x, int64_t n, int d, float*& xinv, int64_t& cap, float& xinv_max) {
    if (n > cap) {
        if (xinv) CK(cudaFree(xinv));
        xinv = nullptr;
        cap = 0;
        CK(cudaMalloc(&xinv, (size_t)(n + n / 2 + 64) * sizeof(float)));
        cap = n + n / 2 + 64;
    }
    uint32_t* mx = nullptr;
    CK(cudaMalloc(&mx, sizeof(uint32_t)));
    cudaError_t ce = cudaMemset(mx, 0, sizeof(uint32_t));
    if (ce == cudaSuccess) ce = brag::i8_row_inv_norms(x, n, d, xinv, mx, 0);
    uint32_t bits = 0;
    if (ce == cudaSuccess) ce = cudaMemcpy(&bits, mx, sizeof(uint32_t), cudaMemcpyDeviceToHost);
    cudaFree(mx);
    CK(ce);
    memcpy(&xinv_max, &bits, sizeof(float));
}

// after the rows of a space changed: the scan's tensor map and, for int8 storage, the row norms
void refresh_dense(DenseSpace& s) {
    s.umma_ok = brag::dense_umma_supported(s.dtype, s.d, s.n) &&
                brag::make_tmap_2d(s.tmap_x, s.x, s.n, s.d, 256, s.dtype) == cudaSuccess;
    if (s.dtype == 3) {
        REQUIRE(s.umma_ok, "int8 spaces are scanned by the tensor-core kernel only: d must be a multiple of 128");
        i8_norms(s.x, s.n, s.d, s.xinv, s.xinv_cap, s.xinv_max);
    }
    s.cp_mask_id = 0;                                     // any filter-compacted copy is stale
    s.cp_use = false;
}

const uint32_t* mask_ptr(brag_engine* e, int mask_id, int64_t need_bits) {
    if (mask_id == 0) return nullptr;
    auto it = e->masks.find(mask_id);
    REQUIRE(it != e->masks.end(), "unknown mask_id");
    REQUIRE(it->second.nbits >= need_bits, "mask has fewer bits than the space has rows");
    return it->second.bits;
}

DenseSpace& dense_space(brag_engine* e, const char* name) {
    auto it = e->dense.find(name);
    REQUIRE(it != e->dense.end(), std::string("unknown dense space: ") + name);
    return it->second;
}
SparseSpace& sparse_space(brag_engine* e, const char* name) {
    auto it = e->sparse.find(name);
    REQUIRE(it != e->sparse.end(), std::string("unknown sparse space: ") + name);
    return it->second;
}

void free_dense(DenseSpace& s) {
    if (s.x) cudaFree(s.x);
    if (s.cp_x) cudaFree(s.cp_x);
    if (s.cp_rowmap) cudaFree(s.cp_rowmap);
    if (s.xinv) cudaFree(s.xinv);
    if (s.cp_xinv) cudaFree(s.cp_xinv);
    s.x = nullptr;
    s.cp_x = nullptr;
    s.cp_rowmap = nullptr;
    s.xinv = s.cp_xinv = nullptr;
    s.xinv_cap = s.cp_xinv_cap = 0;
}
void free_sparse(SparseSpace& s) {
    if (s.cp) { free_sparse(*s.cp); delete s.cp; }
    if (s.cp_rowmap) cudaFree(s.cp_rowmap);
    if (s.term_ptr) cudaFree(s.term_ptr);
    if (s.post) cudaFree(s.post);
    if (s.sub_off) cudaFree(s.sub_off);
    if (s.sub_row) cudaFree(s.sub_row);
    if (s.wrange) cudaFree(s.wrange);
    if (s.tau) cudaFree(s.tau);
    if (s.bits_hi) cudaFree(s.bits_hi);
    if (s.bm_row) cudaFree(s.bm_row);
    if (s.bits) cudaFree(s.bits);
    s = SparseSpace();
}

// Index of the bitmap-filter scan: weight range per term, and document maps for the frequent terms (df >= N/256),
// densest first, within as many bytes as the postings take.
void build_bitmap_index(SparseSpace& s) {
    CK(cudaMalloc(&s.wrange, (size_t)s.V * sizeof(float2)));
    CK(cudaMalloc(&s.tau, (size_t)s.V * sizeof(float)));
    int* flag = nullptr;
    CK(cudaMalloc(&flag, sizeof(int)));
    cudaError_t ce = brag::bm25x_build_wrange(s.term_ptr, s.post, s.V, s.nnz, s.wrange, s.tau, flag, 0);
    int nonfinite = 0;
    if (ce == cudaSuccess) ce = cudaMemcpy(&nonfinite, flag, sizeof(int), cudaMemcpyDeviceToHost);
    cudaFree(flag);
    CK(ce);
    const size_t row_bytes = (size_t)brag::bm25x_row_words(s.n_docs) * 4;
    // stored maps: terms with df >= n_docs / BRAG_BX_MAP_DIV, most frequent first, within BRAG_BX_MAP_BYTES bytes per
    // posting (tuning knobs, read once; see DESIGN 3.3 for the measured trade: map bytes against pool instructions)
    static const int map_div = [] { const char* e = getenv("BRAG_BX_MAP_DIV"); const int v = e ? atoi(e) : 256; return v < 8 ? 8 : (v > 4096 ? 4096 : v); }();
    static const int map_bytes = [] { const char* e = getenv("BRAG_BX_MAP_BYTES"); const int v = e ? atoi(e) : 8; return v < 1 ? 1 : (v > 16 ? 16 : v); }();
    const int64_t min_df = std::max<int64_t>(32, s.n_docs / map_div);
    std::vector<std::pair<int64_t, int32_t>> dense;     // (df, term)
    for (int64_t t = 0; t < s.V; ++t) {
        const int64_t df = s.term_ptr_host[t + 1] - s.term_ptr_host[t];
        if (df >= min_df) dense.push_back({df, (int32_t)t});
    }
    std::sort(dense.begin(), dense.end(), [](const std::pair<int64_t, int32_t>& a, const std::pair<int64_t, int32_t>& b) {
        return a.first != b.first ? a.first > b.first : a.second < b.second;
    });
    // two maps per dense term (present, high class), together at most as many bytes as the postings themselves
    const size_t budget = std::max<size_t>((size_t)64 << 20, (size_t)s.nnz * map_bytes);
    while (!dense.empty() && dense.size() * row_bytes * 2 > budget) dense.pop_back();
    std::vector<int32_t> rows((size_t)s.V, -1), terms(dense.size());
    for (size_t r = 0; r < dense.size(); ++r) { rows[(size_t)dense[r].second] = (int32_t)r; terms[r] = dense[r].second; }
    s.n_dense = (int)dense.size();
    CK(cudaMalloc(&s.bm_row, (size_t)s.V * sizeof(int32_t)));
    CK(cudaMemcpy(s.bm_row, rows.data(), (size_t)s.V * sizeof(int32_t), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&s.bits, std::max<size_t>(256, (size_t)s.n_dense * row_bytes)));
    CK(cudaMalloc(&s.bits_hi, std::max<size_t>(256, (size_t)s.n_dense * row_bytes)));
    if (s.n_dense) {
        int32_t* dterms = nullptr;
        CK(cudaMalloc(&dterms, terms.size() * sizeof(int32_t)));
        ce = cudaMemcpy(dterms, terms.data(), terms.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
        if (ce == cudaSuccess) ce = brag::bm25x_build_bitmaps(s.term_ptr, s.post, dterms, s.n_dense, s.tau, s.n_docs, s.bits, s.bits_hi, 0);
        if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
        cudaFree(dterms);
        CK(ce);
    }
    s.bitmap_ok = !nonfinite;
}

// everything the scans need beside term_ptr / post (both already on the device)
void build_sparse_index(SparseSpace& s, int64_t suboff_min_df) {
    const int64_t ns1 = (s.n_docs + brag::BM25_SUB - 1) / brag::BM25_SUB + 1;
    // an offset row (4 * ns1 bytes) only for terms whose postings (8 * df bytes) outweigh it: the table then never
    // exceeds the postings, whatever the vocabulary (a hashed-stem vocabulary is millions of rare terms)
    const int64_t min_df = suboff_min_df > 0 ? suboff_min_df : std::max<int64_t>(1, (ns1 + 1) / 2);
    std::vector<int32_t> rows((size_t)s.V, -1), terms;
    for (int64_t t = 0; t < s.V; ++t)
        if (s.term_ptr_host[t + 1] - s.term_ptr_host[t] >= min_df) { rows[(size_t)t] = (int32_t)terms.size(); terms.push_back((int32_t)t); }
    s.n_sub = (int64_t)terms.size();
    CK(cudaMalloc(&s.sub_row, (size_t)s.V * sizeof(int32_t)));
    CK(cudaMemcpy(s.sub_row, rows.data(), (size_t)s.V * sizeof(int32_t), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&s.sub_off, std::max<size_t>(256, (size_t)s.n_sub * ns1 * sizeof(uint32_t))));
    if (s.n_sub) {
        int32_t* dterms = nullptr;
        CK(cudaMalloc(&dterms, terms.size() * sizeof(int32_t)));
        cudaError_t ce = cudaMemcpy(dterms, terms.data(), terms.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
        if (ce == cudaSuccess) ce = brag::bm25_build_suboffsets(s.term_ptr, s.post, dterms, s.n_sub, s.n_docs, s.sub_off, 0);
        if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
        cudaFree(dterms);
        CK(ce);
    }
    build_bitmap_index(s);
    CK(cudaDeviceSynchronize());
}

int pick_path(brag_engine* e, const DenseSpace& s, int B) {
    if (!s.umma_ok) return 0;
    if (e->force_path == 0) return 0;
    if (e->force_path == 1) return 1;
    // measured on B200 (tools/_probe_paths.py): the TMA + tcgen05 scan streams 16-bit corpora faster than the
    // GEMV at EVERY batch size (B=1: 7.3 vs 7.1 TB/s at 10M x 1024, 5.5 vs 3.7 TB/s at 1M x 768), so the GEMV
    // is only the path of float32 storage and of row lengths the 64-element K slices cannot tile
    (void)B;
    return 1;
}

// ---- device-pointer implementations -------------------------------------------------------
// Compacted copy of the rows of `s` that pass filter `mask_id` (compact.cu), built on first use and kept until the
// filter changes.  Returns false when the filter is not worth compacting (keeps more than half of the rows, or so
// few that the masked scan is already cheap) or the copy does not fit; the caller then scans the full space.
bool ensure_compact(brag_engine* e, DenseSpace& s, int mask_id, const uint32_t* mask, cudaStream_t st) {
    const uint64_t version = e->masks[mask_id].version;
    if (s.cp_mask_id == mask_id && s.cp_version == version) return s.cp_use;
    s.cp_mask_id = mask_id;
    s.cp_version = version;
    s.cp_use = false;
    const int64_t nwords = (s.n + 31) / 32;
    uint32_t* prefix = e->cp_prefix.as<uint32_t>((size_t)nwords + 1);
    const size_t tmp_bytes = brag::compact_tmp_bytes(nwords);
    void* tmp = e->cp_tmp.ensure(tmp_bytes);
    CK(brag::compact_prefix(mask, s.n, prefix, tmp, tmp_bytes, st));
    uint32_t nc = 0;
    CK(cudaMemcpyAsync(&nc, prefix + nwords, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    e->kernel_launches += 2;
    if ((int64_t)nc * 2 > s.n || nc < 4096 || !brag::dense_umma_supported(s.dtype, s.d, nc)) return false;
    const size_t row_bytes = (size_t)s.d * dtype_bytes(s.dtype);
    const size_t need_x = (size_t)nc * row_bytes, need_map = (size_t)nc * sizeof(uint32_t);
    if (need_x > s.cp_x_cap) {
        if (s.cp_x) CK(cudaFree(s.cp_x));
        s.cp_x = nullptr;
        s.cp_x_cap = 0;
        if (cudaMalloc(&s.cp_x, need_x) != cudaSuccess) { cudaGetLastError(); s.cp_x = nullptr; return false; }
        s.cp_x_cap = need_x;
    }
    if (need_map > s.cp_rowmap_cap) {
        if (s.cp_rowmap) CK(cudaFree(s.cp_rowmap));
        s.cp_rowmap = nullptr;
        s.cp_rowmap_cap = 0;
        if (cudaMalloc(&s.cp_rowmap, need_map) != cudaSuccess) { cudaGetLastError(); s.cp_rowmap = nullptr; return false; }
        s.cp_rowmap_cap = need_map;
    }
    CK(brag::compact_rows(s.x, (int)row_bytes, mask, s.n, prefix, nc, s.cp_rowmap, s.cp_x, st));
    if (brag::make_tmap_2d(s.cp_tmap, s.cp_x, nc, s.d, 256, s.dtype) != cudaSuccess) return false;
    if (s.dtype == 3) {
        CK(cudaStreamSynchronize(st));
        i8_norms(s.cp_x, nc, s.d, s.cp_xinv, s.cp_xinv_cap, s.cp_xinv_max);
        e->kernel_launches += 1;
    }
    e->kernel_launches += 2;
    s.cp_n = nc;
    s.cp_use = true;
    return true;
}

// Sparse counterpart: an index over the documents passing filter `mask_id`, built once the filter has served
// `sparse_compact_after` queries of this space (the build reads every posting once, about what a few hundred queries
// cost) and kept until the filter or the space changes.
bool ensure_sparse_compact(brag_engine* e, SparseSpace& s, int mask_id, const uint32_t* mask, int B, cudaStream_t st) {
    const uint64_t version = e->masks[mask_id].version;
    // single queries under other filters (candidate and page masks of the plan executor) neither count nor evict
    if ((s.cp_mask_id != mask_id || s.cp_version != version) && B < 32 && e->sparse_compact_after != 0) return false;
    if (s.cp_mask_id != mask_id || s.cp_version != version) {
        if (s.cp) { CK(cudaDeviceSynchronize()); free_sparse(*s.cp); delete s.cp; s.cp = nullptr; }
        s.cp_mask_id = mask_id;
        s.cp_version = version;
        s.cp_queries = 0;
        s.cp_state = 0;
    }
    if (e->sparse_compact_after < 0) return false;
    // (a single query is as fast through the masked scan of the full index: C5 shard 0.105 vs 0.142 ms)
    if (s.cp_state == 1) return B >= 32 || e->sparse_compact_after == 0;
    if (s.cp_state == 2) return false;
    // single queries are latency-bound, not posting-bound (C5 shard: 0.13 ms either way): only batches count
    if (s.cp_queries < e->sparse_compact_after) { if (B >= 32) s.cp_queries += B; return false; }
    s.cp_state = 2;
    const int64_t nwords = (s.n_docs + 31) / 32;
    uint32_t* prefix = e->cp_prefix.as<uint32_t>((size_t)nwords + 1);
    const size_t tmp_bytes = std::max(brag::compact_tmp_bytes(nwords), brag::sparse_compact_tmp_bytes(s.V, s.nnz));
    void* tmp = e->cp_tmp.ensure(tmp_bytes + 64);
    CK(brag::compact_prefix(mask, s.n_docs, prefix, tmp, tmp_bytes, st));
    uint32_t nc = 0;
    CK(cudaMemcpyAsync(&nc, prefix + nwords, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    e->kernel_launches += 2;
    if ((int64_t)nc * 2 > s.n_docs || nc < 1024) return false;
    SparseSpace* c = new SparseSpace();
    try {
        c->V = s.V;
        c->n_docs = nc;
        CK(cudaMalloc(&c->term_ptr, (size_t)(s.V + 1) * sizeof(int64_t)));
        int64_t* chunk_off = e->cp_chunks.as<int64_t>((size_t)brag::sparse_compact_chunks(s.nnz) + 1);
        CK(brag::sparse_compact_count(s.term_ptr, s.post, s.V, s.nnz, mask, c->term_ptr, chunk_off, tmp, tmp_bytes, st));
        c->term_ptr_host.resize((size_t)s.V + 1);
        CK(cudaMemcpyAsync(c->term_ptr_host.data(), c->term_ptr, (size_t)(s.V + 1) * sizeof(int64_t),
                           cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        c->nnz = c->term_ptr_host[(size_t)s.V];
        CK(cudaMalloc(&c->post, std::max<size_t>((size_t)c->nnz, 1) * sizeof(uint2)));
        CK(brag::sparse_compact_select(s.post, s.nnz, mask, prefix, chunk_off, c->post, c->nnz, st));
        if (s.cp_rowmap) { CK(cudaFree(s.cp_rowmap)); s.cp_rowmap = nullptr; }
        CK(cudaMalloc(&s.cp_rowmap, (size_t)nc * sizeof(uint32_t)));
        CK(brag::compact_rowmap(mask, s.n_docs, prefix, s.cp_rowmap, st));
        CK(cudaStreamSynchronize(st));
        build_sparse_index(*c, e->suboff_min_df);
        e->kernel_launches += 8;
    } catch (const CudaOom&) {
        free_sparse(*c);
        delete c;
        return false;
    } catch (...) {
        free_sparse(*c);
        delete c;
        throw;
    }
    s.cp = c;
    s.cp_state = 1;
    return true;
}

void dense_topk_dev(brag_engine* e, DenseSpace& s, const float* Q, int B, int k, int mask_id, const uint32_t* mask,
                    float* out_scores, int64_t* out_ids, int32_t* out_counts, cudaStream_t st,
                    const ListSink* sink = nullptr) {
    const brag::ShardBounds xb = sink ? sink->xb : brag::ShardBounds();
    REQUIRE(B >= 1 && k >= 1 && k <= 512, "dense_topk: need B >= 1 and 1 <= k <= 512");
    REQUIRE((reinterpret_cast<uintptr_t>(Q) & 15) == 0, "dense_topk: the query array must be 16-byte aligned");
    const int k_sel = k_select(k);
    int path = pick_path(e, s, B);
    if (path == 1 && k_sel > brag::dense_umma_max_ksel()) path = 0;     // k > 416: the GEMV scan holds longer lists
    REQUIRE(s.dtype != 3 || path == 1, "dense_topk: int8 spaces take k <= 416 (tensor-core scan only)");
    uint64_t* cand = e->keys1.as<uint64_t>((size_t)B * k_sel);
    uint64_t* cand2 = e->keys2.as<uint64_t>((size_t)B * k_sel);
    // rows actually scanned: the whole space, or its filter-compacted copy
    const void* X = s.x;
    int64_t n = s.n;
    void* tmap_x = s.tmap_x;
    const uint32_t* scan_mask = mask;
    bool compact = false;
    bool finished = false;                 // the scan's tail already wrote the final list
    if (path == 1) {
        // tcgen05 scan; selection and exact re-score are its tail kernel (no partial lists to merge)
        brag::UmmaPlan plan;
        void* scratch = nullptr;
        brag::UmmaI8 i8;
        // (one stage-0 interval per group; the first one includes the preparation)
        std::optional<StageTimer> t0;
        t0.emplace(e, 0, st);
        {
            if (mask != nullptr && !e->no_compact && ensure_compact(e, s, mask_id, mask, st)) {
                compact = true;
                X = s.cp_x; n = s.cp_n; tmap_x = s.cp_tmap; scan_mask = nullptr;
            }
            if (s.dtype == 3) {
                i8.xinv = compact ? s.cp_xinv : s.xinv;
                i8.xinv_max = compact ? s.cp_xinv_max : s.xinv_max;
            }
            CK(brag::dense_umma_plan(n, s.d, B, k_sel, e->num_sms, &plan));
            const int Bpad = plan.n_qblocks * 128;
            void* q16 = e->q16.ensure((size_t)Bpad * s.d * dtype_bytes(s.dtype));
            float* qinv = s.dtype == 3 ? e->qs.as<float>((size_t)B) : nullptr;
            i8.qinv = qinv;
            CK(brag::scan_queries(Q, B, s.d, s.dtype, q16, Bpad, qinv, st));
            CK(brag::make_tmap_2d(e->tmap_q, q16, Bpad, s.d, 128, s.dtype));
            scratch = e->umma_scratch.ensure(plan.scratch_bytes);
            CK(brag::dense_umma_clear(scratch, plan, st));
            e->kernel_launches += 1;
        }
        brag::UmmaTail tail;
        tail.X = X; tail.dtype = s.dtype; tail.n = n; tail.d = s.d; tail.Q = Q;
        tail.k = k; tail.id_offset = e->id_offset;
        finished = !compact;               // a compacted copy's keys still need the map back to the space's rows
        if (finished && sink) {
            tail.sink.on = 1; tail.sink.list = sink->list; tail.sink.local_out = sink->local_out; tail.sink.push = sink->push;
        } else if (finished) {
            tail.out_scores = out_scores; tail.out_ids = out_ids; tail.out_counts = out_counts;
        } else {
            tail.keys_out = cand2;
        }
        for (int g = 0; g < brag::dense_umma_groups(plan); ++g) {
            {
                if (!t0) t0.emplace(e, 0, st);
                int launches = 0;
                CK(brag::dense_umma_scan(s.dtype, n, s.d, B, scan_mask, k_sel, scratch, plan, tmap_x, e->tmap_q, g, st,
                                         &launches, xb, i8));
                e->kernel_launches += launches;
                t0.reset();
            }
            {
                StageTimer t(e, 1, st);
                CK(brag::dense_umma_tail(B, k_sel, scratch, plan, g, tail, st, xb));
                e->kernel_launches += 1;
            }
        }
    } else {
        {
            StageTimer t(e, 0, st);
            // the GEMV scan multiplies in float32 anyway: it takes the caller's float32 query as it is, so its
            // candidates are ranked by (almost) the exact score whatever the storage dtype
            const int64_t ngroups = (s.n + 31) / 32;
            int grid = (int)std::min<int64_t>((int64_t)e->num_sms * 2, std::max<int64_t>(1, (ngroups + 15) / 16));
            const int P = grid;
            uint64_t* part = e->part.as<uint64_t>((size_t)P * B * k_sel);
            int launches = 0;
            CK(brag::dense_gemv_scan(s.x, s.dtype, s.n, s.d, Q, B, mask, k_sel, part, grid, st, &launches));
            size_t tmp = brag::merge_tmp_elems(P, B, k_sel);
            uint64_t* ta = e->tmp_a.as<uint64_t>(tmp);
            uint64_t* tb = e->tmp_b.as<uint64_t>(tmp);
            CK(brag::merge_key_lists(part, P, B, k_sel, ta, tb, cand, k_sel, st, &launches));
            e->kernel_launches += 1 + launches;
        }
        StageTimer t(e, 1, st);
        // exact scores: float64 dot of the stored row with the caller's float32 query
        CK(brag::rescore_keys(X, s.dtype, n, s.d, Q, B, cand, k_sel, cand2, st));
        e->kernel_launches += 1;
    }
    if (!finished) {
        StageTimer t(e, 1, st);
        if (compact) {
            CK(brag::remap_keys(cand2, (int64_t)B * k_sel, s.cp_rowmap, st));
            e->kernel_launches += 1;
        }
        if (sink)
            CK(brag::shard_finalize_push(cand2, B, k_sel, k, e->id_offset, sink->list, sink->local_out, sink->push, st));
        else
            CK(brag::finalize_keys(cand2, B, k_sel, k, false, e->id_offset, out_scores, out_ids, out_counts, st));
        e->kernel_launches += 1;
    }
}

void dense_rescore_dev(brag_engine* e, DenseSpace& s, const float* Q, int B, const int64_t* cand, int m, int k,
                       const uint32_t* mask, float* out_scores, int64_t* out_ids, int32_t* out_counts,
                       cudaStream_t st) {
    REQUIRE(B >= 1 && k >= 1 && m >= 1 && m <= 8192 && k <= 8192, "dense_rescore: need m,k in [1,8192]");
    REQUIRE((reinterpret_cast<uintptr_t>(Q) & 15) == 0, "dense_rescore: the query array must be 16-byte aligned");
    StageTimer t(e, 1, st);
    uint64_t* keys = e->keys1.as<uint64_t>((size_t)B * m);
    CK(brag::rescore_cands(s.x, s.dtype, s.n, s.d, Q, B, cand, m, e->id_offset, mask, keys, st));
    CK(brag::finalize_keys(keys, B, m, k, true, e->id_offset, out_scores, out_ids, out_counts, st));
    e->kernel_launches += 2;
}

// longest query of a batch whose q_ptr lives on the device (one small D2H copy + stream sync)
int max_terms_dev(brag_engine* e, const int32_t* q_ptr, int B, cudaStream_t st) {
    e->qptr_host.resize((size_t)B + 1);
    CK(cudaMemcpyAsync(e->qptr_host.data(), q_ptr, ((size_t)B + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    int mx = 0;
    for (int b = 0; b < B; ++b) mx = std::max(mx, e->qptr_host[b + 1] - e->qptr_host[b]);
    return mx;
}
int max_terms_host(const int32_t* q_ptr, int B) {
    int mx = 0;
    for (int b = 0; b < B; ++b) mx = std::max(mx, q_ptr[b + 1] - q_ptr[b]);
    return mx;
}
// host-pointer calls: the sparse queries are validated here (the device-pointer contract is the caller's)
void check_host_queries(const int32_t* q_ptr, const int32_t* q_terms, int B, int max_terms, const char* who) {
    for (int b = 0; b < B; ++b) {
        REQUIRE(q_ptr[b + 1] >= q_ptr[b], std::string(who) + ": q_ptr must be non-decreasing");
        REQUIRE(max_terms < 0 || q_ptr[b + 1] - q_ptr[b] <= max_terms, std::string(who) + ": a query is longer than max_terms");
        for (int i = q_ptr[b] + 1; i < q_ptr[b + 1]; ++i)
            REQUIRE(q_terms[i] > q_terms[i - 1], std::string(who) + ": query term rows must be strictly ascending");
    }
}

// max_terms: longest query of the batch, or -1 when only the device copy of q_ptr is at hand
void sparse_topk_dev(brag_engine* e, SparseSpace& full, const int32_t* q_ptr, const int32_t* q_terms,
                     const double* q_vals, int B, int k, int mask_id, const uint32_t* mask, float* out_scores,
                     int64_t* out_ids, int32_t* out_counts, int max_terms, cudaStream_t st,
                     const ListSink* sink = nullptr) {
    REQUIRE(B >= 1 && k >= 1 && k <= 512, "sparse_topk: need B >= 1 and 1 <= k <= 512");
    const brag::ShardBounds xb = sink ? sink->xb : brag::ShardBounds();
    // documents actually scanned: the whole space, or the index of the documents passing the filter
    const bool compact = mask != nullptr && !e->no_compact && ensure_sparse_compact(e, full, mask_id, mask, B, st);
    SparseSpace& s = compact ? *full.cp : full;
    if (compact) mask = nullptr;
    // bitmap-filter scan unless the space has non-finite weights or a query is longer than 32 terms
    bool bitmap = s.bitmap_ok && e->force_sparse_path != 0;
    // (checked first with the 32-term variant, which has the fewest lists: batch-1 calls on device pointers then
    // skip the copy of q_ptr that finding the longest query needs)
    if (bitmap && e->force_sparse_path < 0 && brag::bm25x_plan(s.n_docs, B, k, 32, e->num_sms).visits_per_list < 4)
        bitmap = false;
    if (bitmap) {
        if (max_terms < 0) max_terms = max_terms_dev(e, q_ptr, B, st);
        bitmap = brag::bm25x_supported(max_terms, k);
    }
    REQUIRE(bitmap || e->force_sparse_path != 1, "sparse_topk: the bitmap scan was forced but cannot run this batch");
    brag::Bm25Plan plan = bitmap ? brag::bm25x_plan(s.n_docs, B, k, max_terms, e->num_sms)
                                 : brag::bm25_plan(s.n_docs, B, k, e->num_sms);
    // A worker warp's first visit of a work item has no bound yet and scores every document that shares a term, so
    // the filter only pays when a list covers several tiles: small batches (one split per SM) and small corpora go
    // to the exhaustive kernel (measured crossover at C4: batch ~32, tools/bm25_probe.py)
    if (bitmap && e->force_sparse_path < 0 && plan.visits_per_list < 4) {
        bitmap = false;
        plan = brag::bm25_plan(s.n_docs, B, k, e->num_sms);
    }
    const int P = plan.lists_per_query;
    uint64_t* part = e->part.as<uint64_t>((size_t)P * B * k);
    uint64_t* merged = e->keys1.as<uint64_t>((size_t)B * k);
    void* scratch = e->g_thr.ensure(plan.scratch_bytes);
    {
        StageTimer t(e, 2, st);
        if (bitmap)
            CK(brag::bm25x_scan(s.term_ptr, s.post, s.sub_off, s.sub_row, s.bm_row, s.bits, s.bits_hi, s.wrange, s.tau, s.V, s.n_docs, q_ptr, q_terms,
                                q_vals, B, max_terms, mask, k, part, plan, scratch, st, xb));
        else
            CK(brag::bm25_scan(s.term_ptr, s.post, s.sub_off, s.sub_row, s.V, s.n_docs, q_ptr, q_terms, q_vals, B, mask, k,
                               part, plan, scratch, st, xb));
        e->kernel_launches += 1 + ((bitmap && plan.seed_tiles > 0) ? 1 : 0);
    }
    {
        StageTimer t(e, 3, st);
        int launches = 0;
        // both sparse kernels keep [work counters | bounds | coarse | fine histograms] in `scratch`: the bound of the
        // complete histograms lets the tail drop all but a few hundred of the P * k keys before it orders them
        uint8_t* sp = reinterpret_cast<uint8_t*>(scratch);
        uint32_t* gthr = reinterpret_cast<uint32_t*>(sp + 256);
        const uint32_t* coarse = reinterpret_cast<const uint32_t*>(sp + 256 + plan.gthr_bytes);
        const uint32_t* fine = reinterpret_cast<const uint32_t*>(sp + 256 + plan.gthr_bytes + plan.coarse_bytes);
        const bool finished = !compact;        // a compacted index's keys still need the map back to the space's rows
        brag::ShardSink sk;
        if (finished && sink) { sk.on = 1; sk.list = sink->list; sk.local_out = sink->local_out; sk.push = sink->push; }
        CK(brag::sparse_tail(part, P, B, k, coarse, fine, gthr, xb, k, finished ? nullptr : merged, e->id_offset, out_scores,
                             out_ids, out_counts, sk, st));
        if (!finished) {
            CK(brag::remap_keys(merged, (int64_t)B * k, full.cp_rowmap, st));
            if (sink)
                CK(brag::shard_finalize_push(merged, B, k, k, e->id_offset, sink->list, sink->local_out, sink->push, st));
            else
                CK(brag::finalize_keys(merged, B, k, k, false, e->id_offset, out_scores, out_ids, out_counts, st));
            launches += 2;
        }
        e->kernel_launches += launches + 1;
    }
}

// ---- host staging helpers -----------------------------------------------------------------
template <typename T>
const T* to_dev(DevBuf& buf, const T* host, size_t n, cudaStream_t st) {
    T* d = buf.as<T>(std::max<size_t>(n, 1));
    if (n) CK(cudaMemcpyAsync(d, host, n * sizeof(T), cudaMemcpyHostToDevice, st));
    return d;
}
template <typename T>
void to_host(T* host, const T* dev, size_t n, cudaStream_t st) {
    if (n && host) CK(cudaMemcpyAsync(host, dev, n * sizeof(T), cudaMemcpyDeviceToHost, st));
}

// ---- row sharding: window lifetime and the per-call kernel arguments ---------------------------
void shard_release(brag_engine* e) {
    ShardState& sh = e->shard;
    for (int g = 0; g < 16; ++g) {
        if (sh.ipc_open[g] && sh.peer[g]) cudaIpcCloseMemHandle(sh.peer[g]);
        sh.peer[g] = nullptr;
        sh.ipc_open[g] = false;
    }
    if (sh.window) cudaFree(sh.window);
    sh = ShardState();
}

brag::ShardPush shard_push_args(const ShardState& sh) {
    brag::ShardPush p;
    p.windows = reinterpret_cast<uint8_t* const*>(sh.window + sh.lay.tables_off);
    p.epoch = reinterpret_cast<const unsigned*>(sh.window + sh.lay.epoch_off);
    p.world = sh.world; p.rank = sh.rank;
    p.slots_off = sh.lay.slots_off; p.slot_bytes = sh.lay.slot_bytes; p.flags_off = sh.lay.flags_off;
    return p;
}

brag::ShardBounds shard_bound_args(const ShardState& sh, int list) {
    brag::ShardBounds xb;
    if (!sh.bounds || sh.world <= 1) return xb;
    xb.peers = reinterpret_cast<unsigned long long* const*>(sh.window + sh.lay.tables_off) + (size_t)(1 + list) * sh.world;
    xb.epoch = reinterpret_cast<const unsigned*>(sh.window + sh.lay.epoch_off);
    xb.world = sh.world; xb.rank = sh.rank;
    return xb;
}

template <typename F>
int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return 1;
    } catch (...) {
        g_err = "unknown error";
        return 2;
    }
}

}  // namespace

// =============================================================================================
extern "C" {

const char* brag_last_error(void) { return g_err.c_str(); }
int brag_version(void) { return 2; }

int brag_create(int device, brag_engine** out) {
    return guarded([&] {
        REQUIRE(out != nullptr, "brag_create: out is NULL");
        int count = 0;
        cudaError_t ce = cudaGetDeviceCount(&count);
        if (ce != cudaSuccess || count == 0)
            throw BragError("brag_create: no CUDA device available (this engine has no CPU fallback)");
        REQUIRE(device >= 0 && device < count, "brag_create: bad device index");
        CK(cudaSetDevice(device));
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, device));
        REQUIRE(prop.major >= 10, "brag_create: this library is built for sm_100a (B200) only");
        CK(brag::dense_umma_init_device());
        CK(brag::select_init_device());
        CK(brag::bm25_init_device());
        CK(brag::bm25x_init_device());
        CK(brag::shard_init_device());
        uint64_t* noalloc_keys = nullptr;
        CK(brag::rescore_noalloc_init_device(&noalloc_keys));
        auto* e = new brag_engine();
        e->device = device;
        e->num_sms = prop.multiProcessorCount;
        e->noalloc_keys = noalloc_keys;
        *out = e;
    });
}

int brag_destroy(brag_engine* e) {
    return guarded([&] {
        if (!e) return;
        cudaSetDevice(e->device);
        cudaDeviceSynchronize();
        for (auto& kv : e->dense) free_dense(kv.second);
        for (auto& kv : e->sparse) free_sparse(kv.second);
        for (auto& kv : e->masks) if (kv.second.bits) cudaFree(kv.second.bits);
        DevBuf* bufs[] = {&e->qs, &e->q16, &e->part, &e->tmp_a, &e->tmp_b, &e->keys1, &e->keys2, &e->umma_scratch,
                          &e->work_counter, &e->g_thr, &e->hyb_ids, &e->hyb_counts, &e->hyb_scores, &e->cp_prefix,
                          &e->cp_tmp, &e->cp_chunks};
        for (DevBuf* b : bufs) b->release();
        for (auto& b : e->h_in) b.release();
        for (auto& b : e->h_out) b.release();
        e->shard_keys.release();
        shard_release(e);
        for (auto& ev : e->events) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
        for (auto ev : e->pool) cudaEventDestroy(ev);
        delete e;
    });
}

int brag_set_id_offset(brag_engine* e, int64_t offset) {
    return guarded([&] {
        REQUIRE(e && offset >= 0, "brag_set_id_offset: bad arguments");
        e->id_offset = offset;
    });
}

int brag_dense_set(brag_engine* e, const char* name, const void* rows, int64_t n, int32_t d, int32_t dtype,
                   int32_t on_device) {
    return guarded([&] {
        REQUIRE(e && name && rows, "brag_dense_set: NULL argument");
        REQUIRE(dtype >= 0 && dtype <= 3, "brag_dense_set: dtype must be BRAG_F32/F16/BF16/I8");
        const int esz = dtype_bytes(dtype);
        REQUIRE(dtype != 3 || d % 128 == 0, "brag_dense_set: int8 rows must be a multiple of 128 elements");
        REQUIRE(n >= 1 && n < (int64_t)0xFFFFFFFFll, "brag_dense_set: need 1 <= n < 2^32-1");
        REQUIRE(d >= 1 && (d * esz) % 16 == 0, "brag_dense_set: row bytes must be a multiple of 16");
        REQUIRE(d * esz / 16 <= 256, "brag_dense_set: d too large (row > 4096 bytes)");
        use_device(e);
        auto it = e->dense.find(name);
        if (it != e->dense.end()) { CK(cudaDeviceSynchronize()); free_dense(it->second); e->dense.erase(it); }
        DenseSpace s;
        s.n = n; s.cap = n; s.d = d; s.dtype = dtype;
        const size_t bytes = (size_t)n * d * esz;
        CK(cudaMalloc(&s.x, bytes));
        cudaError_t ce = cudaMemcpy(s.x, rows, bytes, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice);
        if (ce != cudaSuccess) { cudaFree(s.x); CK(ce); }
        try { refresh_dense(s); } catch (...) { free_dense(s); throw; }
        e->dense[name] = s;
    });
}

int brag_dense_append(brag_engine* e, const char* name, const void* rows, int64_t n_new, int32_t d, int32_t dtype,
                      int32_t on_device) {
    return guarded([&] {
        REQUIRE(e && name && rows && n_new >= 1, "brag_dense_append: bad arguments");
        use_device(e);
        auto it = e->dense.find(name);
        REQUIRE(it != e->dense.end(), std::string("brag_dense_append: unknown dense space: ") + name);
        DenseSpace& s = it->second;
        REQUIRE(s.d == d && s.dtype == dtype, "brag_dense_append: dimension / storage type differ from the space");
        REQUIRE(s.n + n_new < (int64_t)0xFFFFFFFFll, "brag_dense_append: too many rows");
        const size_t row_bytes = (size_t)d * dtype_bytes(dtype);
        CK(cudaDeviceSynchronize());                      // no scan may still read the buffer that is about to move
        if (s.n + n_new > s.cap) {
            const int64_t cap = std::max<int64_t>(s.n + n_new, s.cap * 2);
            void* nx = nullptr;
            CK(cudaMalloc(&nx, (size_t)cap * row_bytes));
            cudaError_t ce = cudaMemcpy(nx, s.x, (size_t)s.n * row_bytes, cudaMemcpyDeviceToDevice);
            if (ce != cudaSuccess) { cudaFree(nx); CK(ce); }
            cudaFree(s.x);
            s.x = nx;
            s.cap = cap;
        }
        CK(cudaMemcpy(static_cast<char*>(s.x) + (size_t)s.n * row_bytes, rows, (size_t)n_new * row_bytes,
                      on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice));
        s.n += n_new;
        refresh_dense(s);
    });
}

int brag_dense_update(brag_engine* e, const char* name, const int64_t* row_ids, const void* rows, int64_t m) {
    return guarded([&] {
        REQUIRE(e && name && row_ids && rows && m >= 1, "brag_dense_update: bad arguments");
        use_device(e);
        DenseSpace& s = dense_space(e, name);
        const size_t row_bytes = (size_t)s.d * dtype_bytes(s.dtype);
        for (int64_t i = 0; i < m; ++i)
            REQUIRE(row_ids[i] >= 0 && row_ids[i] < s.n, "brag_dense_update: row id out of range");
        CK(cudaDeviceSynchronize());
        for (int64_t i = 0; i < m; ++i)
            CK(cudaMemcpyAsync(static_cast<char*>(s.x) + (size_t)row_ids[i] * row_bytes,
                               static_cast<const char*>(rows) + (size_t)i * row_bytes, row_bytes, cudaMemcpyHostToDevice, 0));
        CK(cudaDeviceSynchronize());
        if (s.dtype == 3) i8_norms(s.x, s.n, s.d, s.xinv, s.xinv_cap, s.xinv_max);
        s.cp_mask_id = 0;
        s.cp_use = false;
    });
}

int brag_dense_drop(brag_engine* e, const char* name) {
    return guarded([&] {
        REQUIRE(e && name, "brag_dense_drop: NULL argument");
        use_device(e);
        auto it = e->dense.find(name);
        if (it == e->dense.end()) return;
        CK(cudaDeviceSynchronize());
        free_dense(it->second);
        e->dense.erase(it);
    });
}

int brag_sparse_set_csr(brag_engine* e, const char* name, const int64_t* term_ptr, const uint32_t* doc_ids,
                        const float* weights, int64_t n_terms, int64_t n_docs, int32_t on_device) {
    return guarded([&] {
        REQUIRE(e && name && term_ptr, "brag_sparse_set_csr: NULL argument");
        REQUIRE(n_terms >= 1 && n_docs >= 1 && n_docs < (int64_t)0xFFFFFFFFll, "brag_sparse_set_csr: bad sizes");
        use_device(e);
        auto it = e->sparse.find(name);
        if (it != e->sparse.end()) { CK(cudaDeviceSynchronize()); free_sparse(it->second); e->sparse.erase(it); }
        SparseSpace s;
        s.V = n_terms; s.n_docs = n_docs;
        s.term_ptr_host.resize(n_terms + 1);
        const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
        CK(cudaMemcpy(s.term_ptr_host.data(), term_ptr, (n_terms + 1) * sizeof(int64_t),
                      on_device ? cudaMemcpyDeviceToHost : cudaMemcpyHostToHost));
        REQUIRE(s.term_ptr_host[0] == 0, "brag_sparse_set_csr: term_ptr[0] must be 0");
        for (int64_t t = 0; t < n_terms; ++t)
            REQUIRE(s.term_ptr_host[t + 1] >= s.term_ptr_host[t], "brag_sparse_set_csr: term_ptr must be non-decreasing");
        s.nnz = s.term_ptr_host[n_terms];
        REQUIRE(s.nnz == 0 || (doc_ids && weights), "brag_sparse_set_csr: NULL postings");
        try {
            CK(cudaMalloc(&s.term_ptr, (n_terms + 1) * sizeof(int64_t)));
            CK(cudaMemcpy(s.term_ptr, term_ptr, (n_terms + 1) * sizeof(int64_t), kind));
            CK(cudaMalloc(&s.post, std::max<size_t>(s.nnz, 1) * sizeof(uint2)));
            if (s.nnz) {
                if (on_device) {
                    CK(brag::bm25_interleave(doc_ids, weights, s.nnz, s.post, 0));
                } else {
                    uint32_t* td = nullptr;
                    float* tw = nullptr;
                    cudaError_t ce = cudaMalloc(&td, s.nnz * sizeof(uint32_t));
                    if (ce == cudaSuccess) ce = cudaMalloc(&tw, s.nnz * sizeof(float));
                    if (ce == cudaSuccess) ce = cudaMemcpy(td, doc_ids, s.nnz * sizeof(uint32_t), kind);
                    if (ce == cudaSuccess) ce = cudaMemcpy(tw, weights, s.nnz * sizeof(float), kind);
                    if (ce == cudaSuccess) ce = brag::bm25_interleave(td, tw, s.nnz, s.post, 0);
                    if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
                    if (td) cudaFree(td);
                    if (tw) cudaFree(tw);
                    CK(ce);
                }
            }
            build_sparse_index(s, e->suboff_min_df);
        } catch (...) {
            free_sparse(s);
            throw;
        }
        e->sparse[name] = std::move(s);
    });
}

int brag_sparse_append(brag_engine* e, const char* name, const int64_t* delta_term_ptr, const uint32_t* doc_ids,
                       const float* weights, const int32_t* row_map, int64_t n_terms_new, int64_t n_docs_new) {
    return guarded([&] {
        REQUIRE(e && name && delta_term_ptr && row_map, "brag_sparse_append: NULL argument");
        use_device(e);
        SparseSpace& s = sparse_space(e, name);
        REQUIRE(n_terms_new >= s.V && n_docs_new >= s.n_docs && n_docs_new < (int64_t)0xFFFFFFFFll,
                "brag_sparse_append: the space can only grow");
        REQUIRE(delta_term_ptr[0] == 0, "brag_sparse_append: delta_term_ptr[0] must be 0");
        const int64_t dnnz = delta_term_ptr[n_terms_new];
        REQUIRE(dnnz == 0 || (doc_ids && weights), "brag_sparse_append: NULL postings");
        for (int64_t t = 0; t < n_terms_new; ++t) {
            REQUIRE(delta_term_ptr[t + 1] >= delta_term_ptr[t], "brag_sparse_append: delta_term_ptr must be non-decreasing");
            for (int64_t j = delta_term_ptr[t]; j < delta_term_ptr[t + 1]; ++j) {
                REQUIRE((int64_t)doc_ids[j] >= s.n_docs && (int64_t)doc_ids[j] < n_docs_new,
                        "brag_sparse_append: delta documents must be new rows (>= the current document count)");
                REQUIRE(j == delta_term_ptr[t] || doc_ids[j] > doc_ids[j - 1],
                        "brag_sparse_append: documents must be strictly ascending within a term");
            }
        }
        for (int64_t t = 0; t < s.V; ++t)
            REQUIRE(row_map[t] >= 0 && row_map[t] < n_terms_new && (t == 0 || row_map[t] > row_map[t - 1]),
                    "brag_sparse_append: row_map must be strictly increasing inside the new vocabulary");
        // new term_ptr; where each old list and each delta posting lands
        std::vector<int64_t> new_ptr((size_t)n_terms_new + 1, 0), dst_start((size_t)std::max<int64_t>(s.V, 1), 0);
        for (int64_t t = 0; t < n_terms_new; ++t) new_ptr[(size_t)t + 1] = delta_term_ptr[t + 1] - delta_term_ptr[t];
        for (int64_t t = 0; t < s.V; ++t) new_ptr[(size_t)row_map[t] + 1] += s.term_ptr_host[t + 1] - s.term_ptr_host[t];
        for (int64_t t = 0; t < n_terms_new; ++t) new_ptr[(size_t)t + 1] += new_ptr[(size_t)t];
        for (int64_t t = 0; t < s.V; ++t) dst_start[(size_t)t] = new_ptr[(size_t)row_map[t]];
        std::vector<int64_t> ddst((size_t)std::max<int64_t>(dnnz, 1));
        std::vector<uint2> dpost((size_t)std::max<int64_t>(dnnz, 1));
        for (int64_t t = 0; t < n_terms_new; ++t) {
            const int64_t dcount = delta_term_ptr[t + 1] - delta_term_ptr[t];
            const int64_t at = new_ptr[(size_t)t + 1] - dcount;       // after the old postings of the term
            for (int64_t j = 0; j < dcount; ++j) {
                const int64_t src = delta_term_ptr[t] + j;
                ddst[(size_t)src] = at + j;
                uint32_t wb;
                std::memcpy(&wb, &weights[src], 4);
                dpost[(size_t)src] = make_uint2(doc_ids[src], wb);
            }
        }
        const int64_t nnz_new = new_ptr[(size_t)n_terms_new];
        CK(cudaDeviceSynchronize());
        SparseSpace g;
        g.V = n_terms_new; g.n_docs = n_docs_new; g.nnz = nnz_new;
        int64_t *d_dst_start = nullptr, *d_ddst = nullptr;
        uint2* d_dpost = nullptr;
        bool old_freed = false;
        try {
            CK(cudaMalloc(&g.term_ptr, (size_t)(n_terms_new + 1) * sizeof(int64_t)));
            CK(cudaMemcpy(g.term_ptr, new_ptr.data(), (size_t)(n_terms_new + 1) * sizeof(int64_t), cudaMemcpyHostToDevice));
            CK(cudaMalloc(&g.post, std::max<size_t>((size_t)nnz_new, 1) * sizeof(uint2)));
            CK(cudaMalloc(&d_dst_start, dst_start.size() * sizeof(int64_t)));
            CK(cudaMemcpy(d_dst_start, dst_start.data(), dst_start.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
            CK(cudaMalloc(&d_ddst, ddst.size() * sizeof(int64_t)));
            CK(cudaMemcpy(d_ddst, ddst.data(), ddst.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
            CK(cudaMalloc(&d_dpost, dpost.size() * sizeof(uint2)));
            CK(cudaMemcpy(d_dpost, dpost.data(), dpost.size() * sizeof(uint2), cudaMemcpyHostToDevice));
            CK(brag::sparse_append_move(s.term_ptr, s.post, s.V, s.nnz, d_dst_start, d_dpost, d_ddst, dnnz, g.post, 0));
            CK(cudaDeviceSynchronize());
            // the old postings are no longer needed: release them before the index of the grown space is built
            free_sparse(s);
            old_freed = true;
            g.term_ptr_host = std::move(new_ptr);
            build_sparse_index(g, e->suboff_min_df);
        } catch (...) {
            if (d_dst_start) cudaFree(d_dst_start);
            if (d_ddst) cudaFree(d_ddst);
            if (d_dpost) cudaFree(d_dpost);
            free_sparse(g);
            if (old_freed) e->sparse.erase(name);      // nothing left to serve: the caller must upload the space again
            throw;
        }
        cudaFree(d_dst_start);
        cudaFree(d_ddst);
        cudaFree(d_dpost);
        e->kernel_launches += 8;
        s = std::move(g);
    });
}

int brag_sparse_drop(brag_engine* e, const char* name) {
    return guarded([&] {
        REQUIRE(e && name, "brag_sparse_drop: NULL argument");
        use_device(e);
        auto it = e->sparse.find(name);
        if (it == e->sparse.end()) return;
        CK(cudaDeviceSynchronize());
        free_sparse(it->second);
        e->sparse.erase(it);
    });
}

int brag_sparse_df(brag_engine* e, const char* name, int64_t* df_host, int64_t n_terms) {
    return guarded([&] {
        REQUIRE(e && name && df_host, "brag_sparse_df: NULL argument");
        SparseSpace& s = sparse_space(e, name);
        REQUIRE(n_terms == s.V, "brag_sparse_df: n_terms mismatch");
        for (int64_t t = 0; t < n_terms; ++t) df_host[t] = s.term_ptr_host[t + 1] - s.term_ptr_host[t];
    });
}

int brag_mask_set(brag_engine* e, int32_t mask_id, const uint32_t* bits, int64_t nbits, int32_t on_device) {
    return guarded([&] {
        REQUIRE(e && bits && mask_id >= 1 && nbits >= 1, "brag_mask_set: bad arguments");
        use_device(e);
        MaskBuf& m = e->masks[mask_id];
        const size_t words = (size_t)((nbits + 31) / 32);
        // padded (zeros) so a warp's mask-word read past the last group stays in bounds; a buffer that fits is reused:
        // the plan executor of client.py sets several candidate / page masks per query
        if (m.bits) CK(cudaDeviceSynchronize());
        if (m.cap < words + 512) {
            if (m.bits) { CK(cudaFree(m.bits)); m.bits = nullptr; m.cap = 0; }
            CK(cudaMalloc(&m.bits, (words + 512) * sizeof(uint32_t)));
            m.cap = words + 512;
        }
        CK(cudaMemcpy(m.bits, bits, words * sizeof(uint32_t), on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice));
        CK(cudaMemset(m.bits + words, 0, (m.cap - words) * sizeof(uint32_t)));
        m.nbits = nbits;
        m.version = ++e->mask_epoch;
    });
}

int brag_dense_topk(brag_engine* e, const char* name, const float* Q, int32_t B, int32_t k, int32_t mask_id,
                    float* out_scores, int64_t* out_ids, int32_t* out_counts, int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && name && Q && out_scores && out_ids && out_counts, "brag_dense_topk: NULL argument");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        DenseSpace& s = dense_space(e, name);
        const uint32_t* mask = mask_ptr(e, mask_id, s.n);
        if (on_device) {
            dense_topk_dev(e, s, Q, B, k, mask_id, mask, out_scores, out_ids, out_counts, st);
        } else {
            const float* dq = to_dev(e->h_in[0], Q, (size_t)B * s.d, st);
            float* ds = e->h_out[0].as<float>((size_t)B * k);
            int64_t* di = e->h_out[1].as<int64_t>((size_t)B * k);
            int32_t* dc = e->h_out[2].as<int32_t>(B);
            dense_topk_dev(e, s, dq, B, k, mask_id, mask, ds, di, dc, st);
            to_host(out_scores, ds, (size_t)B * k, st);
            to_host(out_ids, di, (size_t)B * k, st);
            to_host(out_counts, dc, (size_t)B, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

static_assert(BRAG_NOALLOC_MAX_KEYS == brag::NOALLOC_MAX_KEYS, "include/brag.h and kernels.h disagree on the static scratch size");

int brag_dense_rescore_noalloc(brag_engine* e, const char* name, const float* Q, int32_t B, const int64_t* cand,
                               int32_t m, int32_t k, int32_t mask_id, float* out_scores, int64_t* out_ids,
                               int32_t* out_counts, void* stream) {
    return guarded([&] {
        REQUIRE(e && name && Q && cand && out_scores && out_ids && out_counts, "brag_dense_rescore_noalloc: NULL argument");
        REQUIRE(B >= 1 && k >= 1 && m >= 1 && m <= 8192 && k <= 8192, "brag_dense_rescore_noalloc: need m,k in [1,8192]");
        REQUIRE((int64_t)B * m <= BRAG_NOALLOC_MAX_KEYS,
                "brag_dense_rescore_noalloc: B * m exceeds BRAG_NOALLOC_MAX_KEYS (65536), the size of the static scratch; "
                "split the batch or use brag_dense_rescore");
        REQUIRE((reinterpret_cast<uintptr_t>(Q) & 15) == 0, "brag_dense_rescore_noalloc: the query array must be 16-byte aligned");
        use_device(e);
        DenseSpace& s = dense_space(e, name);
        const uint32_t* mask = mask_ptr(e, mask_id, s.n);
        // launches only: no cudaMalloc / cudaFree, no staging buffers, no events, no synchronisation
        CK(brag::dense_rescore_noalloc(e->noalloc_keys, s.x, s.dtype, s.n, s.d, Q, B, cand, m, k, e->id_offset, mask,
                                       out_scores, out_ids, out_counts, (cudaStream_t)stream));
        e->kernel_launches += 2;
    });
}

int brag_dense_rescore(brag_engine* e, const char* name, const float* Q, int32_t B, const int64_t* cand, int32_t m,
                       int32_t k, int32_t mask_id, float* out_scores, int64_t* out_ids, int32_t* out_counts,
                       int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && name && Q && cand && out_scores && out_ids && out_counts, "brag_dense_rescore: NULL argument");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        DenseSpace& s = dense_space(e, name);
        const uint32_t* mask = mask_ptr(e, mask_id, s.n);
        if (on_device) {
            dense_rescore_dev(e, s, Q, B, cand, m, k, mask, out_scores, out_ids, out_counts, st);
        } else {
            const float* dq = to_dev(e->h_in[0], Q, (size_t)B * s.d, st);
            const int64_t* dcand = to_dev(e->h_in[1], cand, (size_t)B * m, st);
            float* ds = e->h_out[0].as<float>((size_t)B * k);
            int64_t* di = e->h_out[1].as<int64_t>((size_t)B * k);
            int32_t* dc = e->h_out[2].as<int32_t>(B);
            dense_rescore_dev(e, s, dq, B, dcand, m, k, mask, ds, di, dc, st);
            to_host(out_scores, ds, (size_t)B * k, st);
            to_host(out_ids, di, (size_t)B * k, st);
            to_host(out_counts, dc, (size_t)B, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_sparse_topk(brag_engine* e, const char* name, const int32_t* q_ptr, const int32_t* q_terms,
                     const double* q_vals, int32_t B, int32_t k, int32_t mask_id, int32_t max_terms, float* out_scores,
                     int64_t* out_ids, int32_t* out_counts, int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && name && q_ptr && out_scores && out_ids && out_counts, "brag_sparse_topk: NULL argument");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        SparseSpace& s = sparse_space(e, name);
        const uint32_t* mask = mask_ptr(e, mask_id, s.n_docs);
        if (on_device) {
            sparse_topk_dev(e, s, q_ptr, q_terms, q_vals, B, k, mask_id, mask, out_scores, out_ids, out_counts,
                            max_terms, st);
        } else {
            const size_t nt = (size_t)q_ptr[B];
            check_host_queries(q_ptr, q_terms, B, max_terms, "brag_sparse_topk");
            const int32_t* dp = to_dev(e->h_in[0], q_ptr, (size_t)B + 1, st);
            const int32_t* dt = to_dev(e->h_in[1], q_terms, nt, st);
            const double* dv = to_dev(e->h_in[2], q_vals, nt, st);
            float* ds = e->h_out[0].as<float>((size_t)B * k);
            int64_t* di = e->h_out[1].as<int64_t>((size_t)B * k);
            int32_t* dc = e->h_out[2].as<int32_t>(B);
            sparse_topk_dev(e, s, dp, dt, dv, B, k, mask_id, mask, ds, di, dc,
                            max_terms < 0 ? max_terms_host(q_ptr, B) : max_terms, st);
            to_host(out_scores, ds, (size_t)B * k, st);
            to_host(out_ids, di, (size_t)B * k, st);
            to_host(out_counts, dc, (size_t)B, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_rrf(brag_engine* e, const int64_t* ids, const int32_t* counts, int32_t L, int32_t B, int32_t kmax,
             int32_t limit, int32_t tie_mode, double* out_scores, int64_t* out_ids, int32_t* out_counts,
             int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && ids && counts && out_scores && out_ids && out_counts, "brag_rrf: NULL argument");
        REQUIRE(L >= 1 && L <= 16 && B >= 1 && kmax >= 1 && limit >= 1 && (int64_t)L * kmax <= 8192 && limit <= 8192,
                "brag_rrf: need 1 <= L <= 16, L*kmax <= 8192, limit <= 8192");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        StageTimer t(e, 4, st);
        e->kernel_launches += 1;
        if (on_device) {
            CK(brag::rrf_fuse(ids, counts, L, B, kmax, limit, tie_mode, out_scores, out_ids, out_counts, st));
        } else {
            const int64_t* di = to_dev(e->h_in[0], ids, (size_t)L * B * kmax, st);
            const int32_t* dc = to_dev(e->h_in[1], counts, (size_t)L * B, st);
            double* os = e->h_out[0].as<double>((size_t)B * limit);
            int64_t* oi = e->h_out[1].as<int64_t>((size_t)B * limit);
            int32_t* oc = e->h_out[2].as<int32_t>(B);
            CK(brag::rrf_fuse(di, dc, L, B, kmax, limit, tie_mode, os, oi, oc, st));
            to_host(out_scores, os, (size_t)B * limit, st);
            to_host(out_ids, oi, (size_t)B * limit, st);
            to_host(out_counts, oc, (size_t)B, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_rerank_meanpool(brag_engine* e, const float* query_hidden, int32_t Tq, const float* doc_hidden, int32_t n,
                         int32_t T, int32_t d, int32_t* out_order, float* out_scores, int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && query_hidden && doc_hidden && out_order && out_scores, "brag_rerank_meanpool: NULL argument");
        REQUIRE(Tq >= 1 && T >= 1 && d >= 1 && n >= 1 && n <= 4096, "brag_rerank_meanpool: need Tq, T, d >= 1 and 1 <= n <= 4096");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        StageTimer t(e, 1, st);
        e->kernel_launches += 2;
        if (on_device) {
            CK(brag::rerank_meanpool(query_hidden, Tq, doc_hidden, n, T, d, out_scores, out_order, st));
        } else {
            const float* dq = to_dev(e->h_in[0], query_hidden, (size_t)Tq * d, st);
            const float* dd = to_dev(e->h_in[1], doc_hidden, (size_t)n * T * d, st);
            float* ds = e->h_out[0].as<float>((size_t)n);
            int32_t* dord = e->h_out[2].as<int32_t>((size_t)n);
            CK(brag::rerank_meanpool(dq, Tq, dd, n, T, d, ds, dord, st));
            to_host(out_scores, ds, (size_t)n, st);
            to_host(out_order, dord, (size_t)n, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_merge_topk(brag_engine* e, const float* scores, const int64_t* ids, const int32_t* counts, int32_t G,
                    int32_t B, int32_t k, float* out_scores, int64_t* out_ids, int32_t* out_counts,
                    int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && scores && ids && counts && out_scores && out_ids && out_counts, "brag_merge_topk: NULL argument");
        REQUIRE(G >= 1 && B >= 1 && k >= 1 && (int64_t)G * k <= 8192, "brag_merge_topk: need G*k <= 8192");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        StageTimer t(e, 5, st);
        e->kernel_launches += 1;
        if (on_device) {
            CK(brag::merge_scored_lists(scores, ids, counts, G, B, k, out_scores, out_ids, out_counts, st));
        } else {
            const float* ds = to_dev(e->h_in[0], scores, (size_t)G * B * k, st);
            const int64_t* di = to_dev(e->h_in[1], ids, (size_t)G * B * k, st);
            const int32_t* dc = to_dev(e->h_in[2], counts, (size_t)G * B, st);
            float* os = e->h_out[0].as<float>((size_t)B * k);
            int64_t* oi = e->h_out[1].as<int64_t>((size_t)B * k);
            int32_t* oc = e->h_out[2].as<int32_t>(B);
            CK(brag::merge_scored_lists(ds, di, dc, G, B, k, os, oi, oc, st));
            to_host(out_scores, os, (size_t)B * k, st);
            to_host(out_ids, oi, (size_t)B * k, st);
            to_host(out_counts, oc, (size_t)B, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_hybrid_topk(brag_engine* e, const char* dense_name, const char* sparse_name, const float* Q,
                     const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals, int32_t B, int32_t k,
                     int32_t mask_id, int32_t max_terms, double* out_scores, int64_t* out_ids, int32_t* out_counts,
                     int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && dense_name && sparse_name && Q && q_ptr && out_scores && out_ids && out_counts,
                "brag_hybrid_topk: NULL argument");
        REQUIRE(k >= 1 && k <= 512, "brag_hybrid_topk: need 1 <= k <= 512");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        DenseSpace& ds = dense_space(e, dense_name);
        SparseSpace& ss = sparse_space(e, sparse_name);
        const uint32_t* mask = mask_ptr(e, mask_id, std::max(ds.n, ss.n_docs));
        // stage lists live packed as [2][B][k] so RRF reads them in place
        int64_t* ids = e->hyb_ids.as<int64_t>((size_t)2 * B * k);
        int32_t* cnts = e->hyb_counts.as<int32_t>((size_t)2 * B);
        float* scs = e->hyb_scores.as<float>((size_t)2 * B * k);
        const float* dq = Q;
        const int32_t *dp = q_ptr, *dt = q_terms;
        const double* dv = q_vals;
        double* os = out_scores;
        int64_t* oi = out_ids;
        int32_t* oc = out_counts;
        if (!on_device) {
            const size_t nt = (size_t)q_ptr[B];
            check_host_queries(q_ptr, q_terms, B, max_terms, "brag_hybrid_topk");
            if (max_terms < 0) max_terms = max_terms_host(q_ptr, B);
            dq = to_dev(e->h_in[0], Q, (size_t)B * ds.d, st);
            dp = to_dev(e->h_in[1], q_ptr, (size_t)B + 1, st);
            dt = to_dev(e->h_in[2], q_terms, nt, st);
            dv = to_dev(e->h_in[3], q_vals, nt, st);
            os = e->h_out[0].as<double>((size_t)B * k);
            oi = e->h_out[1].as<int64_t>((size_t)B * k);
            oc = e->h_out[2].as<int32_t>(B);
        }
        dense_topk_dev(e, ds, dq, B, k, mask_id, mask, scs, ids, cnts, st);
        sparse_topk_dev(e, ss, dp, dt, dv, B, k, mask_id, mask, scs + (size_t)B * k, ids + (size_t)B * k, cnts + B,
                        max_terms, st);
        {
            StageTimer t(e, 4, st);
            CK(brag::rrf_fuse(ids, cnts, 2, B, k, k, BRAG_TIE_ID, os, oi, oc, st));
            e->kernel_launches += 1;
        }
        if (!on_device) {
            to_host(out_scores, os, (size_t)B * k, st);
            to_host(out_ids, oi, (size_t)B * k, st);
            to_host(out_counts, oc, (size_t)B, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

// ---- row sharding across the GPUs of one box (include/brag.h; kernels in shard.cu) ---------------
int brag_shard_window_create(brag_engine* e, int32_t world, int32_t max_batch, int32_t max_k, void* handle_out) {
    return guarded([&] {
        REQUIRE(e && handle_out, "brag_shard_window_create: NULL argument");
        REQUIRE(world >= 1 && world <= 16 && max_batch >= 1 && max_k >= 1 && max_k <= 512 && (int64_t)world * max_k <= 8192,
                "brag_shard_window_create: need 1 <= world <= 16, 1 <= max_k <= 512, world * max_k <= 8192");
        use_device(e);
        CK(cudaDeviceSynchronize());
        shard_release(e);
        ShardState& sh = e->shard;
        sh.lay = brag::shard_window_layout(world, max_batch, max_k);
        sh.world = world;
        CK(cudaMalloc(&sh.window, sh.lay.total));
        CK(cudaMemset(sh.window, 0, sh.lay.total));
        cudaIpcMemHandle_t h;
        static_assert(sizeof(h) == BRAG_SHARD_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
        CK(cudaIpcGetMemHandle(&h, sh.window));
        std::memcpy(handle_out, &h, sizeof h);
        const char* nb = getenv("BRAG_SHARD_BOUNDS");      // A/B knob: 0 = no cross-shard bound exchange
        sh.bounds = !(nb && atoi(nb) == 0);
    });
}

int brag_shard_connect(brag_engine* e, int32_t rank, int32_t world, const void* handles, brag_engine* const* local_peers) {
    return guarded([&] {
        REQUIRE(e && (handles || local_peers), "brag_shard_connect: need the world's window handles or its engines");
        ShardState& sh = e->shard;
        REQUIRE(sh.window && !sh.connected, "brag_shard_connect: call brag_shard_window_create first (once per connect)");
        REQUIRE(world == sh.world && rank >= 0 && rank < world, "brag_shard_connect: rank / world do not match the window");
        use_device(e);
        sh.rank = rank;
        for (int g = 0; g < world; ++g) {
            if (g == rank) { sh.peer[g] = sh.window; continue; }
            if (local_peers) {
                // engines of this process: enable peer access and use the peer's pointer as it is
                brag_engine* pe = local_peers[g];
                REQUIRE(pe && pe->shard.window && pe->shard.lay.total == sh.lay.total,
                        "brag_shard_connect: a local peer has no window of the same shape");
                if (pe->device != e->device) {
                    int can = 0;
                    CK(cudaDeviceCanAccessPeer(&can, e->device, pe->device));
                    REQUIRE(can, "brag_shard_connect: no peer access between the devices");
                    cudaError_t pe_rc = cudaDeviceEnablePeerAccess(pe->device, 0);
                    if (pe_rc == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                    else CK(pe_rc);
                }
                sh.peer[g] = pe->shard.window;
            } else {
                cudaIpcMemHandle_t h;
                std::memcpy(&h, static_cast<const uint8_t*>(handles) + (size_t)g * BRAG_SHARD_HANDLE_BYTES, sizeof h);
                void* p = nullptr;
                CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
                sh.peer[g] = static_cast<uint8_t*>(p);
                sh.ipc_open[g] = true;
            }
        }
        // device tables: [0] window bases, [1] dense bound words, [2] sparse bound words of every rank
        std::vector<void*> tab((size_t)3 * world);
        for (int g = 0; g < world; ++g) {
            tab[(size_t)g] = sh.peer[g];
            tab[(size_t)world + g] = sh.peer[g] + sh.lay.xb_off;
            tab[(size_t)2 * world + g] = sh.peer[g] + sh.lay.xb_off + (size_t)sh.lay.max_batch * sizeof(uint64_t);
        }
        CK(cudaMemcpy(sh.window + sh.lay.tables_off, tab.data(), tab.size() * sizeof(void*), cudaMemcpyHostToDevice));
        sh.connected = true;
    });
}

int brag_shard_close(brag_engine* e) {
    return guarded([&] {
        REQUIRE(e, "brag_shard_close: NULL engine");
        use_device(e);
        CK(cudaDeviceSynchronize());
        shard_release(e);
    });
}

// the two per-shard lists of a batch as packed keys: into `local_out` [2][B][k], or (NULL) into the world's windows
static void shard_lists_dev(brag_engine* e, DenseSpace& ds, SparseSpace& ss, const float* dq, const int32_t* dp,
                            const int32_t* dt, const double* dv, int B, int k, int mask_id, const uint32_t* mask,
                            int max_terms, uint64_t* local_out, cudaStream_t st) {
    ListSink sink;
    sink.local_out = local_out;
    if (local_out == nullptr) sink.push = shard_push_args(e->shard);
    sink.list = 0;
    if (local_out == nullptr) sink.xb = shard_bound_args(e->shard, 0);
    dense_topk_dev(e, ds, dq, B, k, mask_id, mask, nullptr, nullptr, nullptr, st, &sink);
    sink.list = 1;
    if (local_out == nullptr) sink.xb = shard_bound_args(e->shard, 1);
    sparse_topk_dev(e, ss, dp, dt, dv, B, k, mask_id, mask, nullptr, nullptr, nullptr, max_terms, st, &sink);
}

int brag_shard_local_lists(brag_engine* e, const char* dense_name, const char* sparse_name, const float* Q,
                           const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals, int32_t B, int32_t k,
                           int32_t mask_id, int32_t max_terms, uint64_t* out_keys, int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && dense_name && sparse_name && Q && q_ptr && out_keys, "brag_shard_local_lists: NULL argument");
        REQUIRE(k >= 1 && k <= 512, "brag_shard_local_lists: need 1 <= k <= 512");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        DenseSpace& ds = dense_space(e, dense_name);
        SparseSpace& ss = sparse_space(e, sparse_name);
        REQUIRE(e->id_offset + std::max(ds.n, ss.n_docs) <= (int64_t)0xFFFFFFFFll, "brag_shard_local_lists: global ids must stay below 2^32");
        const uint32_t* mask = mask_ptr(e, mask_id, std::max(ds.n, ss.n_docs));
        if (on_device) {
            shard_lists_dev(e, ds, ss, Q, q_ptr, q_terms, q_vals, B, k, mask_id, mask, max_terms, out_keys, st);
        } else {
            const size_t nt = (size_t)q_ptr[B];
            check_host_queries(q_ptr, q_terms, B, max_terms, "brag_shard_local_lists");
            if (max_terms < 0) max_terms = max_terms_host(q_ptr, B);
            const float* dq = to_dev(e->h_in[0], Q, (size_t)B * ds.d, st);
            const int32_t* dp = to_dev(e->h_in[1], q_ptr, (size_t)B + 1, st);
            const int32_t* dt = to_dev(e->h_in[2], q_terms, nt, st);
            const double* dv = to_dev(e->h_in[3], q_vals, nt, st);
            uint64_t* dk = e->h_out[0].as<uint64_t>((size_t)2 * B * k);
            shard_lists_dev(e, ds, ss, dq, dp, dt, dv, B, k, mask_id, mask, max_terms, dk, st);
            to_host(out_keys, dk, (size_t)2 * B * k, st);
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_shard_fuse(brag_engine* e, const uint64_t* keys, int32_t G, int32_t B, int32_t k, int32_t limit,
                    int32_t tie_mode, double* out_scores, int64_t* out_ids, int32_t* out_counts, float* stage_scores,
                    int64_t* stage_ids, int32_t* stage_counts, int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && keys && out_scores && out_ids && out_counts, "brag_shard_fuse: NULL argument");
        REQUIRE((stage_scores != nullptr) == (stage_ids != nullptr) && (stage_ids != nullptr) == (stage_counts != nullptr),
                "brag_shard_fuse: pass all three stage outputs or none");
        REQUIRE(G >= 1 && B >= 1 && k >= 1 && k <= 512 && (int64_t)G * k <= 8192 && limit >= 1 && limit <= 1024,
                "brag_shard_fuse: need G * k <= 8192, k <= 512, limit <= 1024");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        StageTimer t(e, 5, st);
        e->kernel_launches += 1;
        if (on_device) {
            CK(brag::shard_fuse(keys, brag::ShardPush(), G, B, k, limit, tie_mode, out_scores, out_ids, out_counts,
                                stage_scores, stage_ids, stage_counts, st));
        } else {
            const uint64_t* dk = to_dev(e->h_in[0], keys, (size_t)G * 2 * B * k, st);
            double* os = e->h_out[0].as<double>((size_t)B * limit);
            int64_t* oi = e->h_out[1].as<int64_t>((size_t)B * limit);
            int32_t* oc = e->h_out[2].as<int32_t>(B);
            float* ts = stage_scores ? e->h_out[3].as<float>((size_t)2 * B * k) : nullptr;
            int64_t* ti = stage_scores ? e->h_out[4].as<int64_t>((size_t)2 * B * k) : nullptr;
            int32_t* tc = stage_scores ? e->h_out[5].as<int32_t>((size_t)2 * B) : nullptr;
            CK(brag::shard_fuse(dk, brag::ShardPush(), G, B, k, limit, tie_mode, os, oi, oc, ts, ti, tc, st));
            to_host(out_scores, os, (size_t)B * limit, st);
            to_host(out_ids, oi, (size_t)B * limit, st);
            to_host(out_counts, oc, (size_t)B, st);
            if (stage_scores) {
                to_host(stage_scores, ts, (size_t)2 * B * k, st);
                to_host(stage_ids, ti, (size_t)2 * B * k, st);
                to_host(stage_counts, tc, (size_t)2 * B, st);
            }
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_shard_hybrid_topk(brag_engine* e, const char* dense_name, const char* sparse_name, const float* Q,
                           const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals, int32_t B, int32_t k,
                           int32_t mask_id, int32_t max_terms, double* out_scores, int64_t* out_ids,
                           int32_t* out_counts, float* stage_scores, int64_t* stage_ids, int32_t* stage_counts,
                           int32_t on_device, void* stream) {
    return guarded([&] {
        REQUIRE(e && dense_name && sparse_name && Q && q_ptr && out_scores && out_ids && out_counts,
                "brag_shard_hybrid_topk: NULL argument");
        REQUIRE((stage_scores != nullptr) == (stage_ids != nullptr) && (stage_ids != nullptr) == (stage_counts != nullptr),
                "brag_shard_hybrid_topk: pass all three stage outputs or none");
        ShardState& sh = e->shard;
        REQUIRE(sh.connected, "brag_shard_hybrid_topk: the engine is not connected to its peers (brag_shard_connect)");
        REQUIRE(B >= 1 && B <= sh.lay.max_batch && k >= 1 && k <= sh.lay.max_k,
                "brag_shard_hybrid_topk: batch or k exceed what the window was created for");
        use_device(e);
        cudaStream_t st = (cudaStream_t)stream;
        DenseSpace& ds = dense_space(e, dense_name);
        SparseSpace& ss = sparse_space(e, sparse_name);
        REQUIRE(e->id_offset + std::max(ds.n, ss.n_docs) <= (int64_t)0xFFFFFFFFll, "brag_shard_hybrid_topk: global ids must stay below 2^32");
        const uint32_t* mask = mask_ptr(e, mask_id, std::max(ds.n, ss.n_docs));
        const float* dq = Q;
        const int32_t *dp = q_ptr, *dt = q_terms;
        const double* dv = q_vals;
        double* os = out_scores;
        int64_t* oi = out_ids;
        int32_t* oc = out_counts;
        float* ts = stage_scores;
        int64_t* ti = stage_ids;
        int32_t* tc = stage_counts;
        if (!on_device) {
            const size_t nt = (size_t)q_ptr[B];
            check_host_queries(q_ptr, q_terms, B, max_terms, "brag_shard_hybrid_topk");
            if (max_terms < 0) max_terms = max_terms_host(q_ptr, B);
            dq = to_dev(e->h_in[0], Q, (size_t)B * ds.d, st);
            dp = to_dev(e->h_in[1], q_ptr, (size_t)B + 1, st);
            dt = to_dev(e->h_in[2], q_terms, nt, st);
            dv = to_dev(e->h_in[3], q_vals, nt, st);
            os = e->h_out[0].as<double>((size_t)B * k);
            oi = e->h_out[1].as<int64_t>((size_t)B * k);
            oc = e->h_out[2].as<int32_t>(B);
            if (stage_scores) {
                ts = e->h_out[3].as<float>((size_t)2 * B * k);
                ti = e->h_out[4].as<int64_t>((size_t)2 * B * k);
                tc = e->h_out[5].as<int32_t>((size_t)2 * B);
            }
        }
        // one epoch per call, bumped on the device so that a CUDA-graph replay of this call stays in step with its peers
        CK(brag::shard_next_epoch(reinterpret_cast<unsigned*>(sh.window + sh.lay.epoch_off), st));
        shard_lists_dev(e, ds, ss, dq, dp, dt, dv, B, k, mask_id, mask, max_terms, nullptr, st);
        const brag::ShardPush push = shard_push_args(sh);
        {
            StageTimer t(e, 5, st);
            CK(brag::shard_signal(push, st));
            CK(brag::shard_fuse(nullptr, push, sh.world, B, k, k, BRAG_TIE_ID, os, oi, oc, ts, ti, tc, st));
            e->kernel_launches += 3;
        }
        if (!on_device) {
            to_host(out_scores, os, (size_t)B * k, st);
            to_host(out_ids, oi, (size_t)B * k, st);
            to_host(out_counts, oc, (size_t)B, st);
            if (stage_scores) {
                to_host(stage_scores, ts, (size_t)2 * B * k, st);
                to_host(stage_ids, ti, (size_t)2 * B * k, st);
                to_host(stage_counts, tc, (size_t)2 * B, st);
            }
            CK(cudaStreamSynchronize(st));
        }
    });
}

int brag_profile_enable(brag_engine* e, int32_t on) {
    return guarded([&] {
        REQUIRE(e, "NULL engine");
        e->prof = on != 0;
    });
}

int brag_profile_reset(brag_engine* e) {
    return guarded([&] {
        REQUIRE(e, "NULL engine");
        use_device(e);
        CK(cudaDeviceSynchronize());
        for (auto& ev : e->events) { e->pool.push_back(ev.a); e->pool.push_back(ev.b); }
        e->events.clear();
        for (int i = 0; i < BRAG_N_STAGES; ++i) { e->ms[i] = 0; e->stage_launches[i] = 0; }
        e->kernel_launches = 0;
    });
}

int brag_profile_read(brag_engine* e, double* ms, int64_t* launches, int64_t* kernel_launches_total) {
    return guarded([&] {
        REQUIRE(e, "NULL engine");
        use_device(e);
        CK(cudaDeviceSynchronize());
        for (auto& ev : e->events) {
            float t = 0;
            CK(cudaEventElapsedTime(&t, ev.a, ev.b));
            e->ms[ev.stage] += t;
            e->pool.push_back(ev.a);
            e->pool.push_back(ev.b);
        }
        e->events.clear();
        for (int i = 0; i < BRAG_N_STAGES; ++i) {
            if (ms) ms[i] = e->ms[i];
            if (launches) launches[i] = e->stage_launches[i];
        }
        if (kernel_launches_total) *kernel_launches_total = e->kernel_launches;
    });
}

int brag_dense_path(brag_engine* e, const char* name, int32_t B) {
    int path = -1;
    int rc = guarded([&] {
        REQUIRE(e && name, "NULL argument");
        path = pick_path(e, dense_space(e, name), B);
    });
    return rc == 0 ? path : -1;
}

int brag_force_sparse_path(brag_engine* e, int32_t path) {
    return guarded([&] {
        REQUIRE(e && path >= -1 && path <= 1, "brag_force_sparse_path: path must be -1, 0 or 1");
        e->force_sparse_path = path;
    });
}

int brag_sparse_compact_after(brag_engine* e, int64_t queries) {
    return guarded([&] {
        REQUIRE(e, "brag_sparse_compact_after: NULL engine");
        e->sparse_compact_after = queries;
    });
}

int brag_sparse_offsets_min_df(brag_engine* e, int64_t min_df) {
    return guarded([&] {
        REQUIRE(e && min_df >= 0, "brag_sparse_offsets_min_df: bad arguments");
        e->suboff_min_df = min_df;
    });
}

int brag_sparse_compact_info(brag_engine* e, const char* name, int32_t* mask_id, int64_t* n_docs) {
    return guarded([&] {
        REQUIRE(e && name && mask_id && n_docs, "brag_sparse_compact_info: NULL argument");
        SparseSpace& s = sparse_space(e, name);
        const bool built = s.cp != nullptr && s.cp_state == 1;
        *mask_id = built ? s.cp_mask_id : 0;
        *n_docs = built ? s.cp->n_docs : 0;
    });
}

int brag_force_dense_path(brag_engine* e, int32_t path) {
    return guarded([&] {
        REQUIRE(e && path >= -1 && path <= 2, "brag_force_dense_path: path must be -1, 0, 1 or 2");
        e->force_path = path == 2 ? 1 : path;
        e->no_compact = path == 2;
    });
}

}  // extern "C"

// ---- index persistence (the reference keeps its index on a volume, docker-compose.yml:30-31, and
// reopens per-user collections, qdrant_handler.py:94-100) ------------------------------------------
// One flat file per engine (= per shard): a header, a section table, then 4096-byte-aligned raw arrays exactly as they
// live in HBM -- the dense rows, the postings AND the scan indexes built from them (sub-tile offsets, weight bounds,
// class thresholds, term maps), so loading is a straight copy with nothing to rebuild.
namespace {

constexpr char PERSIST_MAGIC[8] = {'B', 'R', 'A', 'G', 'I', 'D', 'X', '2'};

struct PersistSection {
    char name[96];
    int32_t kind;             // 0 dense rows, 1 sparse array
    int32_t which;            // sparse: 0 term_ptr 1 post 2 sub_off 3 wrange 4 tau 5 bm_row 6 bits 7 bits_hi 8 sub_row
    int64_t a, b, c, d;       // dense: n, d, dtype, 0      sparse: V, n_docs, nnz, n_dense | bitmap_ok << 31 | n_sub << 32
    uint64_t offset, bytes;
};

struct PersistHeader {
    char magic[8];
    int64_t id_offset;
    int64_t n_sections;
    uint64_t data_offset;
};

struct FileCloser {
    FILE* f;
    ~FileCloser() { if (f) fclose(f); }
};

void write_device_array(FILE* f, const void* dev, size_t bytes, void* stage, size_t stage_bytes) {
    for (size_t off = 0; off < bytes; off += stage_bytes) {
        const size_t n = std::min(stage_bytes, bytes - off);
        CK(cudaMemcpy(stage, static_cast<const char*>(dev) + off, n, cudaMemcpyDeviceToHost));
        REQUIRE(fwrite(stage, 1, n, f) == n, "brag_save: short write");
    }
}

void read_device_array(FILE* f, void* dev, size_t bytes, void* stage, size_t stage_bytes) {
    for (size_t off = 0; off < bytes; off += stage_bytes) {
        const size_t n = std::min(stage_bytes, bytes - off);
        REQUIRE(fread(stage, 1, n, f) == n, "brag_load: short read (truncated file)");
        CK(cudaMemcpy(static_cast<char*>(dev) + off, stage, n, cudaMemcpyHostToDevice));
    }
}

void pad_to(FILE* f, uint64_t target) {
    static const char zeros[4096] = {0};
    long at = ftell(f);
    REQUIRE(at >= 0 && (uint64_t)at <= target, "brag_save: layout error");
    uint64_t left = target - (uint64_t)at;
    while (left) {
        const size_t n = (size_t)std::min<uint64_t>(left, sizeof zeros);
        REQUIRE(fwrite(zeros, 1, n, f) == n, "brag_save: short write");
        left -= n;
    }
}

constexpr int N_SPARSE_ARRAYS = 9;
struct SparseArrays {
    const void* ptr[N_SPARSE_ARRAYS];
    size_t bytes[N_SPARSE_ARRAYS];
};

SparseArrays sparse_arrays(const SparseSpace& s) {
    const size_t ns1 = (size_t)((s.n_docs + brag::BM25_SUB - 1) / brag::BM25_SUB + 1);
    const size_t row_bytes = (size_t)brag::bm25x_row_words(s.n_docs) * 4;
    SparseArrays a;
    a.ptr[0] = s.term_ptr; a.bytes[0] = (size_t)(s.V + 1) * sizeof(int64_t);
    a.ptr[1] = s.post;     a.bytes[1] = (size_t)s.nnz * sizeof(uint2);
    a.ptr[2] = s.sub_off;  a.bytes[2] = (size_t)s.n_sub * ns1 * sizeof(uint32_t);
    a.ptr[3] = s.wrange;   a.bytes[3] = (size_t)s.V * sizeof(float2);
    a.ptr[4] = s.tau;      a.bytes[4] = (size_t)s.V * sizeof(float);
    a.ptr[5] = s.bm_row;   a.bytes[5] = (size_t)s.V * sizeof(int32_t);
    a.ptr[6] = s.bits;     a.bytes[6] = (size_t)s.n_dense * row_bytes;
    a.ptr[7] = s.bits_hi;  a.bytes[7] = (size_t)s.n_dense * row_bytes;
    a.ptr[8] = s.sub_row;  a.bytes[8] = (size_t)s.V * sizeof(int32_t);
    return a;
}

}  // namespace

extern "C" {

int brag_save(brag_engine* e, const char* path) {
    return guarded([&] {
        REQUIRE(e && path, "brag_save: NULL argument");
        use_device(e);
        CK(cudaDeviceSynchronize());
        std::vector<PersistSection> secs;
        std::vector<const void*> src;
        auto add = [&](const std::string& name, int kind, int which, int64_t a, int64_t b, int64_t c, int64_t d,
                       const void* p, size_t bytes) {
            PersistSection s;
            std::memset(&s, 0, sizeof s);
            REQUIRE(name.size() < sizeof s.name, "brag_save: space name too long");
            std::memcpy(s.name, name.data(), name.size());
            s.kind = kind; s.which = which; s.a = a; s.b = b; s.c = c; s.d = d; s.bytes = bytes;
            secs.push_back(s);
            src.push_back(p);
        };
        for (auto& kv : e->dense)
            add(kv.first, 0, 0, kv.second.n, kv.second.d, kv.second.dtype, 0, kv.second.x,
                (size_t)kv.second.n * kv.second.d * dtype_bytes(kv.second.dtype));
        for (auto& kv : e->sparse) {
            const SparseSpace& s = kv.second;
            const SparseArrays arr = sparse_arrays(s);
            for (int w = 0; w < N_SPARSE_ARRAYS; ++w)
                add(kv.first, 1, w, s.V, s.n_docs, s.nnz,
                    (int64_t)s.n_dense | ((int64_t)(s.bitmap_ok ? 1 : 0) << 31) | (s.n_sub << 32), arr.ptr[w], arr.bytes[w]);
        }
        PersistHeader h;
        std::memcpy(h.magic, PERSIST_MAGIC, 8);
        h.id_offset = e->id_offset;
        h.n_sections = (int64_t)secs.size();
        auto up = [](uint64_t x) { return (x + 4095) & ~(uint64_t)4095; };
        h.data_offset = up(sizeof h + secs.size() * sizeof(PersistSection));
        uint64_t at = h.data_offset;
        for (auto& s : secs) { s.offset = at; at = up(at + s.bytes); }
        const std::string tmp = std::string(path) + ".tmp";
        FILE* f = fopen(tmp.c_str(), "wb");
        REQUIRE(f != nullptr, std::string("brag_save: cannot open ") + tmp);
        FileCloser fc{f};
        const size_t stage_bytes = (size_t)64 << 20;
        void* stage = nullptr;
        CK(cudaMallocHost(&stage, stage_bytes));
        try {
            REQUIRE(fwrite(&h, sizeof h, 1, f) == 1, "brag_save: short write");
            if (!secs.empty()) REQUIRE(fwrite(secs.data(), sizeof(PersistSection), secs.size(), f) == secs.size(), "brag_save: short write");
            for (size_t i = 0; i < secs.size(); ++i) {
                pad_to(f, secs[i].offset);
                if (secs[i].bytes) write_device_array(f, src[i], secs[i].bytes, stage, stage_bytes);
            }
        } catch (...) {
            cudaFreeHost(stage);
            throw;
        }
        cudaFreeHost(stage);
        REQUIRE(fflush(f) == 0, "brag_save: flush failed");
        fclose(f);
        fc.f = nullptr;
        REQUIRE(rename(tmp.c_str(), path) == 0, "brag_save: cannot move the finished file into place");
    });
}

int brag_load(brag_engine* e, const char* path) {
    return guarded([&] {
        REQUIRE(e && path, "brag_load: NULL argument");
        use_device(e);
        FILE* f = fopen(path, "rb");
        REQUIRE(f != nullptr, std::string("brag_load: cannot open ") + path);
        FileCloser fc{f};
        PersistHeader h;
        REQUIRE(fread(&h, sizeof h, 1, f) == 1 && std::memcmp(h.magic, PERSIST_MAGIC, 8) == 0,
                "brag_load: not an index file of this library version");
        REQUIRE(h.n_sections >= 0 && h.n_sections < (1 << 20), "brag_load: corrupt header");
        std::vector<PersistSection> secs((size_t)h.n_sections);
        if (!secs.empty()) REQUIRE(fread(secs.data(), sizeof(PersistSection), secs.size(), f) == secs.size(), "brag_load: truncated section table");
        CK(cudaDeviceSynchronize());
        const size_t stage_bytes = (size_t)64 << 20;
        void* stage = nullptr;
        CK(cudaMallocHost(&stage, stage_bytes));
        std::map<std::string, SparseSpace> loading;
        try {
            for (auto& s : secs) {
                s.name[sizeof s.name - 1] = 0;
                const std::string name(s.name);
                REQUIRE(fseek(f, (long)s.offset, SEEK_SET) == 0, "brag_load: seek failed");
                if (s.kind == 0) {
                    auto it = e->dense.find(name);
                    if (it != e->dense.end()) { free_dense(it->second); e->dense.erase(it); }
                    DenseSpace ds;
                    ds.n = s.a; ds.cap = s.a; ds.d = (int)s.b; ds.dtype = (int)s.c;
                    REQUIRE(ds.n >= 1 && ds.d >= 1 && ds.dtype >= 0 && ds.dtype <= 3 &&
                            s.bytes == (uint64_t)ds.n * ds.d * dtype_bytes(ds.dtype), "brag_load: corrupt dense section");
                    CK(cudaMalloc(&ds.x, s.bytes));
                    try {
                        read_device_array(f, ds.x, s.bytes, stage, stage_bytes);
                    } catch (...) {
                        cudaFree(ds.x);
                        throw;
                    }
                    refresh_dense(ds);
                    e->dense[name] = ds;
                } else if (s.kind == 1) {
                    SparseSpace& sp = loading[name];
                    sp.V = s.a; sp.n_docs = s.b; sp.nnz = s.c;
                    sp.n_dense = (int)(s.d & 0x7FFFFFFFll);
                    sp.bitmap_ok = ((s.d >> 31) & 1) != 0;
                    sp.n_sub = s.d >> 32;
                    REQUIRE(s.which >= 0 && s.which < N_SPARSE_ARRAYS && sp.V >= 1 && sp.n_docs >= 1, "brag_load: corrupt sparse section");
                    REQUIRE(sparse_arrays(sp).bytes[s.which] == s.bytes, "brag_load: sparse section has the wrong size");
                    void* p = nullptr;
                    CK(cudaMalloc(&p, std::max<size_t>(s.bytes, 256)));
                    switch (s.which) {
                        case 0: sp.term_ptr = static_cast<int64_t*>(p); break;
                        case 1: sp.post = static_cast<uint2*>(p); break;
                        case 2: sp.sub_off = static_cast<uint32_t*>(p); break;
                        case 3: sp.wrange = static_cast<float2*>(p); break;
                        case 4: sp.tau = static_cast<float*>(p); break;
                        case 5: sp.bm_row = static_cast<int32_t*>(p); break;
                        case 6: sp.bits = static_cast<uint32_t*>(p); break;
                        case 7: sp.bits_hi = static_cast<uint32_t*>(p); break;
                        default: sp.sub_row = static_cast<int32_t*>(p); break;
                    }
                    if (s.bytes) read_device_array(f, p, s.bytes, stage, stage_bytes);
                    if (s.which == 0) {
                        sp.term_ptr_host.resize((size_t)sp.V + 1);
                        CK(cudaMemcpy(sp.term_ptr_host.data(), sp.term_ptr, (size_t)(sp.V + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
                        REQUIRE(sp.term_ptr_host[0] == 0 && sp.term_ptr_host[(size_t)sp.V] == sp.nnz, "brag_load: corrupt term_ptr");
                    }
                } else {
                    throw BragError("brag_load: unknown section kind");
                }
            }
            for (auto& kv : loading) {
                SparseSpace& sp = kv.second;
                REQUIRE(sp.term_ptr && sp.post && sp.sub_off && sp.sub_row && sp.wrange && sp.tau && sp.bm_row && sp.bits && sp.bits_hi,
                        "brag_load: a sparse space is missing sections");
            }
        } catch (...) {
            cudaFreeHost(stage);
            for (auto& kv : loading) free_sparse(kv.second);
            throw;
        }
        cudaFreeHost(stage);
        for (auto& kv : loading) {
            auto it = e->sparse.find(kv.first);
            if (it != e->sparse.end()) { free_sparse(it->second); e->sparse.erase(it); }
            e->sparse[kv.first] = std::move(kv.second);
        }
        e->id_offset = h.id_offset;
    });
}

}  // extern "C"
