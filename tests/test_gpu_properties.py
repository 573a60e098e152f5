"""GPU property tests: randomised shapes against the oracle (hypothesis), and size-independent
properties at BASELINE.json's full sizes, where the oracle is too slow to be the checker:
  * the two sparse kernels (exhaustive scatter / rank-safe bitmap filter) share no scoring code and must agree bit
    for bit on the C4 corpus (10 M documents, batch 1024, top-100);
  * dense top-k must not depend on the batch a query travels in, nor on whether a filter is scanned through the
    compacted copy or the full space (10 M x 1024 bf16)."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from oracle import oracle as O
from rag_application_b200 import synth

from test_gpu_parity import check_dense

pytestmark = pytest.mark.gpu

COMMON = dict(deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow],
              derandomize=True)


@settings(max_examples=25, **COMMON)
@given(n=st.integers(1, 4096), d=st.sampled_from([64, 384, 768, 1024]), k=st.integers(1, 128),
       B=st.sampled_from([1, 2, 9, 70]), dtype=st.sampled_from(["f32", "f16", "bf16"]),
       sel=st.sampled_from([None, 0.05, 0.5, 0.9]), seed=st.integers(0, 2 ** 16))
def test_dense_random_shapes_match_oracle(engine, n, d, k, B, dtype, sel, seed):
    x = synth.dense_corpus(n, d, seed=seed, clustered=bool(seed & 1))
    q = synth.dense_queries(x, B, seed=seed + 1)
    engine.dense_set("h", x, dtype)
    mask = None
    if sel is not None:
        mask = synth.random_mask(n, sel, seed=seed + 2)
        engine.mask_set(9, mask)
    s, i, c = engine.dense_topk("h", q, k, mask_id=9 if mask is not None else 0)
    xs = O.round_storage(x, dtype)
    for b in range(0, B, max(1, B // 3)):
        check_dense(xs, q[b], k, s[b], i[b], int(c[b]), mask=mask)


@settings(max_examples=20, **COMMON)
@given(n=st.integers(1, 4096), d=st.sampled_from([128, 384, 768, 1024]), k=st.integers(1, 128),
       B=st.sampled_from([1, 2, 9, 70]), sel=st.sampled_from([None, 0.05, 0.5, 0.9]), seed=st.integers(0, 2 ** 16))
def test_int8_random_shapes_match_oracle_bit_for_bit(engine, n, d, k, B, sel, seed):
    """SURVEY 8 a9 on drawn shapes: int8 rows, integer dots -- ids, counts and score bits equal the oracle's."""
    x = O.normalize_rows(synth.dense_corpus(n, d, seed=seed, clustered=bool(seed & 1)))
    q = O.normalize_rows(synth.dense_queries(x, B, seed=seed + 1))
    xi, qi = O.quantize_int8(x).astype(np.float32), O.quantize_int8(q).astype(np.float32)
    engine.dense_set("h8", xi, "i8")
    mask = None
    if sel is not None:
        mask = synth.random_mask(n, sel, seed=seed + 2)
        engine.mask_set(9, mask)
    s, i, c = engine.dense_topk("h8", qi, k, mask_id=9 if mask is not None else 0)
    for b in range(0, B, max(1, B // 3)):
        es, ei = O.dense_topk(xi, qi[b], k, mask=mask, int8=True)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei)
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))


@settings(max_examples=20, **COMMON)
@given(n_docs=st.integers(1, 30000), V=st.integers(1, 3000), L=st.integers(1, 80), k=st.integers(1, 128),
       B=st.sampled_from([1, 5, 40]), nt=st.integers(1, 20), path=st.sampled_from(["scatter", "bitmap", None]),
       sel=st.sampled_from([None, 0.1, 0.7]), neg=st.booleans(), seed=st.integers(0, 2 ** 16))
def test_sparse_random_shapes_bit_exact(engine, n_docs, V, L, k, B, nt, path, sel, neg, seed):
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, V, L, seed=seed)
    rng = np.random.default_rng(seed)
    if neg:
        w = (w * np.where(rng.random(len(w)) < 0.2, -1.0, 1.0)).astype(np.float32)
    csr = O.CSR(ptr, doc, w, n_docs)
    nt = min(nt, V)
    q_ptr = np.arange(0, (B + 1) * nt, nt, dtype=np.int32)
    q_terms = np.concatenate([np.sort(rng.choice(V, nt, replace=False)) for _ in range(B)]).astype(np.int32)
    q_vals = rng.random(B * nt) * (2.0 if neg else 1.0) - (0.5 if neg else 0.0)
    engine.sparse_set_csr("h", ptr, doc, w, n_docs)
    mask = None
    if sel is not None:
        mask = synth.random_mask(n_docs, sel, seed=seed + 2)
        engine.mask_set(9, mask)
    engine.force_sparse_path(path)
    try:
        s, i, c = engine.sparse_topk("h", q_ptr, q_terms, q_vals, k, mask_id=9 if mask is not None else 0)
    finally:
        engine.force_sparse_path(None)
    for b in range(B):
        es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k, mask=mask)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei), (b, path)
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))
        assert np.all(i[b][len(ei):] == -1)


@settings(max_examples=15, **COMMON)
@given(n_docs=st.integers(2100, 30000), V=st.integers(1, 3000), L=st.integers(1, 80), k=st.integers(1, 128),
       B=st.sampled_from([1, 5, 40]), nt=st.integers(1, 12), path=st.sampled_from(["scatter", "bitmap", None]),
       sel=st.sampled_from([0.5, 0.3, 0.06]), seed=st.integers(0, 2 ** 16))
def test_sparse_compacted_index_random_shapes_bit_exact(engine, n_docs, V, L, k, B, nt, path, sel, seed):
    """Filter-compacted sparse index (compact.cu) on random shapes: built on first use, oracle-exact.  (Filters keeping
    fewer than 1024 documents or more than half of them fall back to the masked scan: also exercised here.)"""
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, V, L, seed=seed)
    rng = np.random.default_rng(seed)
    csr = O.CSR(ptr, doc, w, n_docs)
    nt = min(nt, V)
    q_ptr = np.arange(0, (B + 1) * nt, nt, dtype=np.int32)
    q_terms = np.concatenate([np.sort(rng.choice(V, nt, replace=False)) for _ in range(B)]).astype(np.int32)
    q_vals = rng.random(B * nt)
    engine.sparse_set_csr("h", ptr, doc, w, n_docs)
    mask = synth.random_mask(n_docs, sel, seed=seed + 2)
    engine.mask_set(9, mask)
    engine.sparse_compact_after(0)
    engine.force_sparse_path(path)
    try:
        s, i, c = engine.sparse_topk("h", q_ptr, q_terms, q_vals, k, mask_id=9)
        s2, i2, c2 = engine.sparse_topk("h", q_ptr, q_terms, q_vals, k, mask_id=9)       # cached index
    finally:
        engine.force_sparse_path(None)
        engine.sparse_compact_after(256)
    assert np.array_equal(i, i2) and np.array_equal(c, c2) and np.array_equal(s.view(np.uint32), s2.view(np.uint32))
    for b in range(B):
        es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k, mask=mask)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei), (b, path)
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))


@settings(max_examples=15, **COMMON)
@given(n0=st.integers(1, 8000), extra=st.integers(0, 3000), V=st.integers(2, 2000), L=st.integers(1, 60),
       drop=st.sampled_from([0, 2, 5]), k=st.integers(1, 100), seed=st.integers(0, 2 ** 16))
def test_sparse_append_random_shapes_equal_fresh_upload(engine, n0, extra, V, L, drop, k, seed):
    """brag_sparse_append on random shapes: documents [n0, n0 + extra) spliced into an index of the first n0, with
    every `drop`-th vocabulary row unknown to the old index (inserted in the middle of the term order); answers and
    document frequencies must equal a fresh upload, and the oracle."""
    n1 = n0 + extra
    ptr, doc, w, _ = synth.zipf_corpus(n1, V, L, seed=seed)
    term = np.repeat(np.arange(V), np.diff(ptr))
    late = ((term % drop) == 1) if drop else np.zeros(len(term), bool)
    keep = ~(late & (doc < n0))
    term, doc, w = term[keep], doc[keep], w[keep]

    def csr(sel, rows, n_rows):
        p = np.zeros(n_rows + 1, np.int64)
        np.add.at(p, rows[sel] + 1, 1)
        return np.cumsum(p), doc[sel].astype(np.uint32), w[sel].astype(np.float32)

    full = csr(np.ones(len(term), bool), term, V)
    old_terms = np.flatnonzero((np.arange(V) % drop) != 1) if drop else np.arange(V)
    old_row = np.full(V, -1, np.int64)
    old_row[old_terms] = np.arange(len(old_terms))
    is_old = doc < n0
    o = csr(is_old, old_row[term], len(old_terms))
    d = csr(~is_old, term, V)
    engine.sparse_set_csr("f", full[0], full[1], full[2], n1)
    engine.sparse_set_csr("a", o[0], o[1], o[2], n0)
    engine.sparse_append("a", d[0], d[1], d[2], old_terms.astype(np.int32), n1)
    assert np.array_equal(engine.sparse_df("a", V), engine.sparse_df("f", V))
    rng = np.random.default_rng(seed)
    B, nt = 6, min(6, V)
    q_ptr = np.arange(0, (B + 1) * nt, nt, dtype=np.int32)
    q_terms = np.concatenate([np.sort(rng.choice(V, nt, replace=False)) for _ in range(B)]).astype(np.int32)
    q_vals = rng.random(B * nt)
    sa, ia, ca = engine.sparse_topk("a", q_ptr, q_terms, q_vals, k)
    sf, if_, cf = engine.sparse_topk("f", q_ptr, q_terms, q_vals, k)
    assert np.array_equal(ia, if_) and np.array_equal(ca, cf) and np.array_equal(sa.view(np.uint32), sf.view(np.uint32))
    oc = O.CSR(full[0], full[1], full[2], n1)
    for b in range(B):
        es, ei = O.sparse_topk(oc, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k)
        assert int(ca[b]) == len(ei) and np.array_equal(ia[b][:len(ei)], ei)
    engine.sparse_drop("a")
    engine.sparse_drop("f")


def test_full_size_sparse_kernels_agree():
    """C4 sparse stage at full size: 10 M documents, 50 k-term Zipf vocabulary, batch 1024, top-100."""
    import torch
    from rag_application_b200 import Engine
    dev = torch.device("cuda", 0)
    eng = Engine(0)
    try:
        n = 10_000_000
        tp, doc, w = synth.zipf_corpus_torch(n, 50_000, 285, dev)
        qp, qt, qv = synth.sparse_queries_torch(tp, 1024, 8)
        eng.sparse_set_csr("s", tp, doc, w, n)
        del doc, w
        torch.cuda.empty_cache()
        g = torch.Generator(device=dev)
        g.manual_seed(synth.SEED_MASK)
        eng.mask_set(1, (torch.rand(n, generator=g, device=dev) < 0.10).cpu().numpy())
        qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
        for mask_id in (0, 1):
            out = {}
            for path in ("scatter", "bitmap"):
                eng.force_sparse_path(path)
                out[path] = [t.cpu().numpy() for t in eng.sparse_topk("s", qp_d, qt_d, qv_d, 100, mask_id=mask_id)]
            eng.force_sparse_path(None)
            (s0, i0, c0), (s1, i1, c1) = out["scatter"], out["bitmap"]
            assert np.array_equal(c0, c1) and np.array_equal(i0, i1)
            assert np.array_equal(s0.view(np.uint32), s1.view(np.uint32))
            assert (c0 == 100).all() and (np.diff(s0, axis=1) <= 0).all()          # full, sorted lists
    finally:
        eng.close()


def test_full_size_dense_batch_and_filter_invariance():
    """C4 dense stage at full size (10 M x 1024 bf16): a query's list is the same alone, in a batch of 256, and --
    under a 10 % filter -- through the compacted copy or the full-space masked scan."""
    import torch
    from rag_application_b200 import Engine
    dev = torch.device("cuda", 0)
    eng = Engine(0)
    try:
        n, d, k = 10_000_000, 1024, 100
        x = synth.dense_corpus_torch(n, d, torch.bfloat16, dev)
        q = synth.dense_queries_torch(x, 256)
        eng.dense_set("x", x)
        del x
        torch.cuda.empty_cache()
        sb, ib, cb = [t.cpu().numpy() for t in eng.dense_topk("x", q, k)]
        assert (cb == k).all() and (np.diff(sb, axis=1) <= 0).all()
        for b in (0, 100, 255):
            s1, i1, c1 = [t.cpu().numpy() for t in eng.dense_topk("x", q[b:b + 1].contiguous(), k)]
            assert np.array_equal(i1[0], ib[b]) and np.array_equal(s1[0].view(np.uint32), sb[b].view(np.uint32))
        g = torch.Generator(device=dev)
        g.manual_seed(synth.SEED_MASK)
        m = torch.rand(n, generator=g, device=dev) < 0.10
        eng.mask_set(1, m.cpu().numpy())
        sc, ic, cc = [t.cpu().numpy() for t in eng.dense_topk("x", q, k, mask_id=1)]
        eng.force_dense_path("tcgen05_full")
        sf, if_, cf = [t.cpu().numpy() for t in eng.dense_topk("x", q, k, mask_id=1)]
        eng.force_dense_path(None)
        assert np.array_equal(ic, if_) and np.array_equal(cc, cf) and np.array_equal(sc.view(np.uint32), sf.view(np.uint32))
        assert m.cpu().numpy()[ic].all()
    finally:
        eng.close()
