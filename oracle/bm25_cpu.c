/* CPU restatement (C, pthreads) of the sparse stage for timing at sizes the numpy oracle cannot
 * reach.  TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.  PARITY UNPINNED (restates
 * qdrant-client local-mode sparse scoring; reference request qdrant_handler.py:347-354).
 *
 * Semantics identical to oracle.sparse_topk: per query, terms in ascending row order, float64
 * accumulation, float32 score, only touched documents are candidates, order (score desc, id asc).
 * Implementation: term-major inverted index + per-thread dense accumulator + size-k heap.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

typedef struct { float s; int64_t id; } ent_t;

/* a is "worse" than b under (score desc, id asc) */
static inline int worse(ent_t a, ent_t b) { return a.s < b.s || (a.s == b.s && a.id > b.id); }

static void sift_down(ent_t* h, int n, int i) {           /* min-heap on "worse" = root is the worst kept */
    for (;;) {
        int l = 2 * i + 1, r = l + 1, m = i;
        if (l < n && worse(h[l], h[m])) m = l;
        if (r < n && worse(h[r], h[m])) m = r;
        if (m == i) return;
        ent_t t = h[i]; h[i] = h[m]; h[m] = t; i = m;
    }
}

static int cmp_best_first(const void* pa, const void* pb) {
    ent_t a = *(const ent_t*)pa, b = *(const ent_t*)pb;
    return worse(a, b) ? 1 : (worse(b, a) ? -1 : 0);
}

typedef struct {
    const int64_t* term_ptr; const uint32_t* doc; const float* w; int64_t n_terms, n_docs;
    const int32_t *q_ptr, *q_terms; const double* q_vals; int32_t B, k; const uint8_t* mask;
    float* out_s; int64_t* out_i; int32_t* out_c; int32_t* next; int* failed;
} job_t;

static void* worker(void* arg) {
    job_t* J = (job_t*)arg;
    const int64_t n_docs = J->n_docs;
    const int k = J->k;
    double* acc = (double*)calloc((size_t)n_docs, sizeof(double));
    uint8_t* hit = (uint8_t*)calloc((size_t)n_docs, 1);
    uint32_t* touched = (uint32_t*)malloc((size_t)n_docs * sizeof(uint32_t));
    ent_t* heap = (ent_t*)malloc((size_t)k * sizeof(ent_t));
    if (!acc || !hit || !touched || !heap) {
        __atomic_store_n(J->failed, 1, __ATOMIC_RELAXED);
    } else {
        for (;;) {
            int32_t b = __atomic_fetch_add(J->next, 1, __ATOMIC_RELAXED);
            if (b >= J->B) break;
            int64_t nt = 0;
            for (int32_t i = J->q_ptr[b]; i < J->q_ptr[b + 1]; ++i) {
                int32_t t = J->q_terms[i];
                if (t < 0 || t >= J->n_terms) continue;
                const double qv = J->q_vals[i];
                for (int64_t p = J->term_ptr[t]; p < J->term_ptr[t + 1]; ++p) {
                    uint32_t d = J->doc[p];
                    if (!hit[d]) { hit[d] = 1; touched[nt++] = d; acc[d] = 0.0; }
                    acc[d] += qv * (double)J->w[p];
                }
            }
            int hn = 0;
            for (int64_t j = 0; j < nt; ++j) {
                uint32_t d = touched[j];
                hit[d] = 0;
                if (J->mask && !J->mask[d]) continue;
                ent_t e = {(float)acc[d], (int64_t)d};
                if (hn < k) {
                    heap[hn++] = e;
                    if (hn == k) for (int i = k / 2 - 1; i >= 0; --i) sift_down(heap, k, i);
                } else if (worse(heap[0], e)) {
                    heap[0] = e;
                    sift_down(heap, k, 0);
                }
            }
            qsort(heap, (size_t)hn, sizeof(ent_t), cmp_best_first);
            for (int j = 0; j < k; ++j) {
                J->out_s[(size_t)b * k + j] = j < hn ? heap[j].s : -INFINITY;
                J->out_i[(size_t)b * k + j] = j < hn ? heap[j].id : -1;
            }
            J->out_c[b] = hn;
        }
    }
    free(acc); free(hit); free(touched); free(heap);
    return NULL;
}

int bm25_topk_batch(const int64_t* term_ptr, const uint32_t* doc, const float* w, int64_t n_terms,
                    int64_t n_docs, const int32_t* q_ptr, const int32_t* q_terms, const double* q_vals,
                    int32_t B, int32_t k, const uint8_t* mask /* may be NULL, 1 byte per doc */,
                    float* out_s, int64_t* out_i, int32_t* out_c, int32_t nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    int32_t next = 0;
    int failed = 0;
    job_t J = {term_ptr, doc, w, n_terms, n_docs, q_ptr, q_terms, q_vals, B, k, mask, out_s, out_i, out_c, &next, &failed};
    pthread_t th[256];
    int started = 0;
    for (int i = 0; i < nthreads; ++i) {
        if (pthread_create(&th[i], NULL, worker, &J) != 0) break;
        ++started;
    }
    if (started == 0) worker(&J);
    for (int i = 0; i < started; ++i) pthread_join(th[i], NULL);
    return failed;
}
