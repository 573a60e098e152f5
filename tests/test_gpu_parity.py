"""GPU parity: every C-ABI stage against the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): bit-exact ids / ranks / counts for the integer work (BM25
candidate sets, RRF ranks, top-k id sets with ties broken by id); dense scores within 1e-3
relative (observed ~1e-7: candidates are re-scored with float64 accumulation)."""
import numpy as np
import pytest

from oracle import oracle as O
from rag_application_b200 import synth

pytestmark = pytest.mark.gpu

DENSE_RTOL = 1e-3      # tolerance stated by north_star
TIE_EPS = 4e-7         # |score gap| below which two rows count as a boundary near-tie


def check_dense(xs, qs, k, got_s, got_i, got_c, mask=None, id_offset=0):
    """Compare one query's GPU list with the oracle; near-ties at the k boundary may swap."""
    n = xs.shape[0]
    all_s = O.dense_scores(xs, qs)
    ids = np.arange(n)
    if mask is not None:
        ids = ids[mask]
    o = O.topk_order(all_s[ids], ids, k)
    exp_i, exp_s = ids[o], all_s[ids][o]
    assert got_c == len(exp_i)
    gi = got_i[:got_c] - id_offset
    gs = got_s[:got_c]
    assert np.all(got_i[got_c:] == -1)
    np.testing.assert_allclose(gs, all_s[gi], rtol=DENSE_RTOL, atol=1e-6)
    if np.array_equal(gi, exp_i):
        return
    # every disagreement must be explained by a near-tie in the oracle's own scores
    for pos in np.nonzero(gi != exp_i)[0]:
        assert abs(float(all_s[gi[pos]]) - float(exp_s[pos])) <= TIE_EPS * max(1.0, abs(float(exp_s[pos]))), \
            f"rank {pos}: got row {gi[pos]} ({all_s[gi[pos]]}) expected {exp_i[pos]} ({exp_s[pos]})"
    missing = set(exp_i) - set(gi)
    kth = float(exp_s[-1])
    for r in missing:
        assert abs(float(all_s[r]) - kth) <= TIE_EPS * max(1.0, abs(kth))


@pytest.mark.parametrize("dtype", ["f32", "f16", "bf16"])
@pytest.mark.parametrize("n,d", [(37, 64), (5000, 384), (4099, 768), (2500, 1024)])
@pytest.mark.parametrize("B", [1, 3, 8])
def test_dense_gemv(engine, dtype, n, d, B):
    x = synth.dense_corpus(n, d, seed=11 + n)
    q = synth.dense_queries(x, B, seed=7)
    xs = O.round_storage(x, dtype)
    engine.force_dense_path("gemv")
    engine.dense_set("t", x, dtype)
    for k in (10, 100):
        s, i, c = engine.dense_topk("t", q, k)
        for b in range(B):
            check_dense(xs, q[b], k, s[b], i[b], int(c[b]))
    engine.force_dense_path(None)


@pytest.mark.parametrize("dtype", ["f16", "bf16"])
@pytest.mark.parametrize("n,d", [(300, 64), (20000, 768), (9000, 1024), (70000, 128)])
@pytest.mark.parametrize("B", [16, 128, 200])
def test_dense_tcgen05(engine, dtype, n, d, B):
    x = synth.dense_corpus(n, d, seed=5 + n, clustered=True)
    q = synth.dense_queries(x, B, seed=9)
    xs = O.round_storage(x, dtype)
    engine.dense_set("t", x, dtype)
    engine.force_dense_path("tcgen05")
    assert engine.dense_path("t", B) == "tcgen05"
    k = 100
    s, i, c = engine.dense_topk("t", q, k)
    engine.force_dense_path(None)
    for b in range(0, B, max(1, B // 16)):
        check_dense(xs, q[b], k, s[b], i[b], int(c[b]))


@pytest.mark.parametrize("n,d", [(300, 64), (20000, 384), (9000, 1024), (70000, 96)])
@pytest.mark.parametrize("B", [1, 16, 200])
def test_dense_tcgen05_tf32_on_float32_storage(engine, n, d, B):
    """float32 spaces (the reference's own datatype, qdrant_handler.py:57-79) are scanned by the same kernel with
    kind::tf32 (rows a whole number of 32-element K slices); candidates are re-scored exactly as before, so the
    answer equals the GEMV path's and the oracle's."""
    x = synth.dense_corpus(n, d, seed=5 + n, clustered=True)
    q = synth.dense_queries(x, B, seed=9)
    engine.dense_set("t", x, "f32")
    assert engine.dense_path("t", B) == "tcgen05"
    k = 100
    s, i, c = engine.dense_topk("t", q, k)
    engine.force_dense_path("gemv")
    s0, i0, c0 = engine.dense_topk("t", q, k)
    engine.force_dense_path(None)
    assert np.array_equal(i, i0) and np.array_equal(c, c0) and np.array_equal(s.view(np.uint32), s0.view(np.uint32))
    for b in range(0, B, max(1, B // 16)):
        check_dense(x, q[b], k, s[b], i[b], int(c[b]))


def _int8_rows(n, d, seed, B=0):
    """Unit-norm float rows quantised the way the reference does (qdrant_handler.py:144-146), as float32 arrays."""
    x = O.normalize_rows(synth.dense_corpus(n, d, seed=seed, clustered=True))
    xi = O.quantize_int8(x).astype(np.float32)
    if not B:
        return xi
    q = O.normalize_rows(synth.dense_queries(x, B, seed=seed + 1))
    return xi, O.quantize_int8(q).astype(np.float32)


def check_int8(xi, qi, k, s, i, c, mask=None, id_offset=0):
    """int8 spaces are integer work: ids, counts AND score bits equal the oracle's (exact int dot, one float64
    normalisation)."""
    es, ei = O.dense_topk(xi, qi, k, mask=mask, int8=True)
    assert int(c) == len(ei)
    assert np.array_equal(i[:len(ei)] - id_offset, ei), (i[:len(ei)][:8], ei[:8])
    assert np.array_equal(s[:len(ei)].view(np.uint32), es.view(np.uint32))
    assert np.all(i[len(ei):] == -1)


@pytest.mark.parametrize("n,d", [(300, 128), (20000, 768), (9000, 1024), (160_000, 128)])
@pytest.mark.parametrize("B", [1, 16, 200])
def test_dense_int8_space_bit_exact(engine, n, d, B):
    """the `quantized` space held as int8 rows and scanned with tcgen05 kind::i8 (exact int32 dots)."""
    xi, qi = _int8_rows(n, d, 100 + n, B)
    engine.dense_set("q8", xi, "i8")
    assert engine.dense_path("q8", B) == "tcgen05"
    for k in (10, 100):
        s, i, c = engine.dense_topk("q8", qi, k)
        for b in range(0, B, max(1, B // 16)):
            check_int8(xi, qi[b], k, s[b], i[b], c[b])


def test_dense_int8_mask_offset_append_update_rescore(engine):
    n, d, B, k = 30000, 256, 40, 60
    xi, qi = _int8_rows(n, d, 300, B)
    engine.dense_set("q8", xi[:20000], "i8")
    engine.dense_append("q8", xi[20000:])
    xi[[3, 777, 25000]] = xi[[9, 9, 9]]                     # overwritten rows tie exactly: lower id first
    engine.dense_update("q8", [3, 777, 25000], xi[[3, 777, 25000]])
    engine.set_id_offset(5_000_000)
    try:
        for sel in (1.0, 0.3, 0.01):                        # unfiltered, compacted copy, masked scan
            mask = None if sel == 1.0 else synth.random_mask(n, sel, seed=7)
            if mask is not None:
                engine.mask_set(6, mask)
            s, i, c = engine.dense_topk("q8", qi, k, mask_id=0 if mask is None else 6)
            for b in range(0, B, 5):
                check_int8(xi, qi[b], k, s[b], i[b], c[b], mask=mask, id_offset=5_000_000)
        s, i, c = engine.dense_topk("q8", xi[9:10], 4)
        assert (i[0] - 5_000_000).tolist() == [3, 9, 777, 25000] and np.all(s[0] == np.float32(1.0))
    finally:
        engine.set_id_offset(0)
    rng = np.random.default_rng(1)
    cand = rng.integers(0, n, size=(B, 70)).astype(np.int64)
    cand[:, 60:] = -1
    s, i, c = engine.dense_rescore("q8", qi, cand, 30)
    for b in range(0, B, 7):
        es, ei = O.dense_topk(xi, qi[b], 30, cand=cand[b][cand[b] >= 0], int8=True)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei)
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))
    # a zero row and a zero query score 0, never NaN
    z = xi[:256].copy()
    z[5] = 0
    engine.dense_set("z8", z, "i8")
    s, i, c = engine.dense_topk("z8", np.zeros((1, d), np.float32), 3)
    assert int(c[0]) == 3 and i[0].tolist() == [0, 1, 2] and np.all(s[0] == 0)
    s, i, c = engine.dense_topk("z8", qi[:1], 256)
    assert int(c[0]) == 256 and np.isfinite(s[0]).all() and s[0][list(i[0]).index(5)] == 0


def test_dense_int8_rejects_what_it_cannot_hold(engine):
    with pytest.raises(ValueError):
        engine.dense_set("bad", np.full((4, 128), 0.5, np.float32), "i8")       # not integer-valued
    with pytest.raises(Exception):
        engine.dense_set("bad", np.zeros((4, 96), np.float32), "i8")            # rows not a multiple of 128
    engine.dense_set("ok", _int8_rows(600, 128, 1), "i8")
    with pytest.raises(Exception):
        engine.dense_topk("ok", np.ones((1, 128), np.float32), 500)               # k > 416 on an int8 space


@pytest.mark.parametrize("masked", [False, True])
def test_dense_tcgen05_seeded_two_phase(engine, masked):
    """Corpora of >= 4*148 tiles take the two-phase scan (seeding scan -> per-query bound -> main scan)."""
    n, d, B, k = 160_000, 64, 130, 100
    x = synth.dense_corpus(n, d, seed=77, clustered=True)
    q = synth.dense_queries(x, B, seed=78)
    xs = O.round_storage(x, "bf16")
    engine.dense_set("t", x, "bf16")
    mask = synth.random_mask(n, 0.10) if masked else None
    if masked:
        engine.mask_set(7, mask)
    engine.force_dense_path("tcgen05")
    s, i, c = engine.dense_topk("t", q, k, mask_id=7 if masked else 0)
    engine.force_dense_path("gemv")
    s0, i0, c0 = engine.dense_topk("t", q[:8], k, mask_id=7 if masked else 0)
    engine.force_dense_path(None)
    assert np.array_equal(i[:8], i0) and np.array_equal(s[:8], s0)
    for b in range(0, B, 13):
        check_dense(xs, q[b], k, s[b], i[b], int(c[b]), mask=mask)


@pytest.mark.parametrize("n,B", [(50_001, 512), (38_000, 256), (160_000, 1024), (50_001, 384), (38_500, 600), (160_000, 896)])
def test_dense_tcgen05_shared_row_ranges(engine, n, B):
    """Corpora of >= 148 tiles: pairs of CTAs scan the same rows, each against half of the query blocks
    (dense_umma.cu: share); with an odd number of blocks (B = 384, 600, 896 -> 3, 5, 7 blocks)
    the last block is split between the pair by rows.  Ragged last chunk, payload filter, every query block checked."""
    d, k = 64, 100
    x = synth.dense_corpus(n, d, seed=81, clustered=True)
    q = synth.dense_queries(x, B, seed=82)
    xs = O.round_storage(x, "f16")
    engine.dense_set("t", x, "f16")
    mask = synth.random_mask(n, 0.5)
    engine.mask_set(9, mask)
    engine.force_dense_path("tcgen05")
    s, i, c = engine.dense_topk("t", q, k, mask_id=9)
    engine.force_dense_path("gemv")
    rows = np.arange(0, B, 37)
    s0, i0, c0 = engine.dense_topk("t", q[rows], k, mask_id=9)
    engine.force_dense_path(None)
    assert np.array_equal(i[rows], i0) and np.array_equal(s[rows], s0)
    for b in range(5, B, 61):
        check_dense(xs, q[b], k, s[b], i[b], int(c[b]), mask=mask)


def test_dense_paths_agree(engine):
    """Both scan kernels feed the same re-score: final lists must be identical."""
    x = synth.dense_corpus(30000, 256, seed=3)
    q = synth.dense_queries(x, 24, seed=4)
    engine.dense_set("t", x, "bf16")
    engine.force_dense_path("gemv")
    s0, i0, c0 = engine.dense_topk("t", q, 50)
    engine.force_dense_path("tcgen05")
    s1, i1, c1 = engine.dense_topk("t", q, 50)
    engine.force_dense_path(None)
    assert np.array_equal(i0, i1) and np.array_equal(c0, c1) and np.array_equal(s0, s1)


@pytest.mark.parametrize("path", ["gemv", "tcgen05"])
def test_dense_mask_and_offset(engine, path):
    n, d, k = 6000, 128, 100
    x = synth.dense_corpus(n, d, seed=21)
    B = 2 if path == "gemv" else 130
    q = synth.dense_queries(x, B, seed=22)
    mask = synth.random_mask(n, 0.10)
    xs = O.round_storage(x, "f16")
    engine.dense_set("t", x, "f16")
    engine.mask_set(1, mask)
    engine.set_id_offset(1_000_000)
    engine.force_dense_path(path)
    s, i, c = engine.dense_topk("t", q, k, mask_id=1)
    engine.force_dense_path(None)
    engine.set_id_offset(0)
    for b in range(0, B, max(1, B // 8)):
        check_dense(xs, q[b], k, s[b], i[b], int(c[b]), mask=mask, id_offset=1_000_000)
    # a mask that keeps fewer rows than k: short list, padded with -1
    tiny = np.zeros(n, dtype=bool)
    tiny[[5, 77, 4000]] = True
    engine.mask_set(2, tiny)
    s, i, c = engine.dense_topk("t", q[:1], k, mask_id=2)
    assert int(c[0]) == 3 and set(i[0][:3]) == {5, 77, 4000} and np.all(i[0][3:] == -1)


@pytest.mark.parametrize("B", [1, 40])
def test_dense_filter_compacted_copy(engine, B):
    """A filter that keeps <= half of the rows (and >= 4096 of them) is scanned through a gathered copy of the
    passing rows (compact.cu): same ids and scores as the oracle and as the full-space masked scan; a new filter
    under the same mask_id rebuilds the copy."""
    n, d, k = 60000, 128, 50
    x = synth.dense_corpus(n, d, seed=31, clustered=True)
    q = synth.dense_queries(x, B, seed=32)
    engine.dense_set("x", x, "bf16")
    xs = O.round_storage(x, "bf16")
    for seed, sel in ((41, 0.10), (42, 0.30)):
        mask = synth.random_mask(n, sel, seed=seed)
        engine.mask_set(5, mask)
        engine.force_dense_path("tcgen05")
        try:
            s, i, c = engine.dense_topk("x", q, k, mask_id=5)
            s2, i2, c2 = engine.dense_topk("x", q, k, mask_id=5)       # second call: cached copy
            engine.force_dense_path("tcgen05_full")
            s3, i3, c3 = engine.dense_topk("x", q, k, mask_id=5)
        finally:
            engine.force_dense_path(None)
        assert np.array_equal(i, i2) and np.array_equal(s.view(np.uint32), s2.view(np.uint32))
        assert np.array_equal(i, i3) and np.array_equal(s.view(np.uint32), s3.view(np.uint32)) and np.array_equal(c, c3)
        assert mask[i[np.arange(k)[None, :] < c[:, None]]].all()
        for b in range(0, B, max(1, B // 8)):
            check_dense(xs, q[b], k, s[b], i[b], int(c[b]), mask=mask)


@pytest.mark.parametrize("k", [416, 417, 512])
@pytest.mark.parametrize("dtype,d", [("bf16", 64), ("f16", 768)])
def test_dense_longest_lists_on_16bit_storage(engine, k, dtype, d):
    """k up to 512 on 16-bit spaces: lists longer than the tcgen05 scan's per-CTA candidate slots (k > 416) are
    scanned by the GEMV kernel; same oracle either way (the paged long limits of client.py ask for k = 512)."""
    n = 5000
    x = synth.dense_corpus(n, d, seed=81, clustered=True)
    q = synth.dense_queries(x, 3, seed=82)
    engine.dense_set("x", x, dtype)
    xs = O.round_storage(x, dtype)
    mask = synth.random_mask(n, 0.5, seed=83)
    engine.mask_set(4, mask)
    for mid, m in ((0, None), (4, mask)):
        s, i, c = engine.dense_topk("x", q, k, mask_id=mid)
        for b in range(3):
            check_dense(xs, q[b], k, s[b], i[b], int(c[b]), mask=m)


@pytest.mark.parametrize("k", [300, 512])
def test_sparse_and_hybrid_longest_lists(engine, k):
    """k up to 512 through the sparse scan (auto path) and the fused hybrid call, against the oracle."""
    n, d, B = 30000, 128, 5
    ptr, doc, w, _ = synth.zipf_corpus(n, 2000, 60, seed=91)
    csr = _oracle_csr(ptr, doc, w, n)
    engine.sparse_set_csr("s", ptr, doc, w, n)
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=92, skip_top=8)
    s, i, c = engine.sparse_topk("s", q_ptr, q_terms, q_vals, k)
    for b in range(B):
        es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei)
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))
    x = synth.dense_corpus(n, d, seed=93, clustered=True)
    q = synth.dense_queries(x, B, seed=94)
    engine.dense_set("x", x, "bf16")
    col = O.OracleCollection()
    col.add_dense("dense", x, "dot", "bf16")
    col.add_sparse("sparse", csr)
    hs, hi, hc = engine.hybrid_topk("x", "s", q, q_ptr, q_terms, q_vals, k)
    for b in range(B):
        es, ei = O.hybrid_simple(col, q[b], q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k)
        assert int(hc[b]) == len(ei) and set(hi[b][:len(ei)].tolist()) == set(ei.tolist())


def test_dense_exact_ties_broken_by_id(engine):
    """Duplicate rows score identically: the lower id must win (north_star: ties broken by id)."""
    x = synth.dense_corpus(512, 64, seed=1)
    x[100:110] = x[7]
    engine.dense_set("t", x, "f32")
    s, i, c = engine.dense_topk("t", x[7:8], 5)
    assert list(i[0]) == [7, 100, 101, 102, 103]


@pytest.mark.parametrize("B", [1, 3, 130])
def test_dense_every_row_ties(engine, B):
    """A corpus of identical rows: every candidate reaches every bound, so the per-CTA lists overflow and the tail
    kernel's selection takes its multi-round path; the answer is still the lowest ids, in order."""
    n, d, k = 40_000, 64, 50
    x = np.tile(synth.dense_corpus(1, d, seed=5), (n, 1))
    q = np.tile(x[:1], (B, 1))
    for dtype in ("f16", "f32"):
        engine.dense_set("ties", x, dtype)
        s, i, c = engine.dense_topk("ties", q, k)
        assert np.all(c == k) and np.array_equal(i, np.tile(np.arange(k), (B, 1)))
        assert np.all(s == s[0, 0])


def test_dense_rescore_union_dedup_mask(engine):
    n, d = 3000, 768
    x = synth.dense_corpus(n, d, seed=31)
    q = synth.dense_queries(x, 4, seed=32)
    engine.dense_set("t", x, "f32")
    rng = np.random.default_rng(0)
    cand = rng.integers(0, n, size=(4, 90)).astype(np.int64)
    cand[:, 10:20] = cand[:, 0:10]      # duplicates (union semantics)
    cand[:, 80:] = -1                   # empty slots
    mask = synth.random_mask(n, 0.5)
    engine.mask_set(3, mask)
    for mid, m in ((0, None), (3, mask)):
        s, i, c = engine.dense_rescore("t", q, cand, 40, mask_id=mid)
        for b in range(4):
            es, ei = O.dense_topk(x, q[b], 40, mask=m, cand=cand[b][cand[b] >= 0])
            assert int(c[b]) == len(ei)
            assert np.array_equal(i[b][:len(ei)], ei)
            np.testing.assert_allclose(s[b][:len(ei)], es, rtol=1e-6)


# ----------------------------------------------------------------------------- sparse
def _oracle_csr(ptr, doc, w, n_docs):
    return O.CSR(ptr, doc, w, n_docs)


@pytest.fixture(params=["scatter", "bitmap"])
def sparse_path(request, engine):
    """Both sparse scan kernels (bm25.cu scatter-accumulate, bm25x.cu bitmap filter) must give the oracle's bits."""
    engine.force_sparse_path(request.param)
    yield request.param
    engine.force_sparse_path(None)


@pytest.mark.parametrize("n_docs,V,L", [(300, 500, 20), (5000, 2000, 60), (20000, 5000, 120)])
@pytest.mark.parametrize("regime", ["zipf", "doc"])
def test_sparse_topk_bit_exact(engine, sparse_path, n_docs, V, L, regime):
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, V, L, seed=n_docs)
    csr = _oracle_csr(ptr, doc, w, n_docs)
    B = 33
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n_docs, B, 8, regime, seed=2, skip_top=8)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    for k in (10, 100):
        s, i, c = engine.sparse_topk("s", q_ptr, q_terms, q_vals, k)
        for b in range(B):
            t, v = q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]]
            es, ei = O.sparse_topk(csr, t, v, k)
            assert int(c[b]) == len(ei)
            assert np.array_equal(i[b][:len(ei)], ei), f"query {b}"
            assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32)), "scores must be bit-exact"
            assert np.all(i[b][len(ei):] == -1)


def test_sparse_overlap_only_and_batch1(engine, sparse_path):
    """Docs sharing no term are not candidates: rare terms give short lists."""
    ptr, doc, w, _ = synth.zipf_corpus(4000, 3000, 40, seed=8)
    csr = _oracle_csr(ptr, doc, w, 4000)
    df = np.diff(ptr)
    rare = np.nonzero((df > 0) & (df < 4))[0][:3].astype(np.int32)
    engine.sparse_set_csr("s", ptr, doc, w, 4000)
    q_ptr = np.array([0, len(rare)], np.int32)
    vals = np.array([0.5, 1.25, 2.0][:len(rare)], np.float64)
    s, i, c = engine.sparse_topk("s", q_ptr, rare, vals, 100)
    es, ei = O.sparse_topk(csr, rare, vals, 100)
    assert 0 < len(ei) < 100 and int(c[0]) == len(ei)
    assert np.array_equal(i[0][:len(ei)], ei)
    # empty query and out-of-vocabulary term: no candidates at all
    s, i, c = engine.sparse_topk("s", np.array([0, 0, 1], np.int32), np.array([3000 + 5], np.int32), np.array([1.0]), 10)
    assert list(c) == [0, 0] and np.all(i == -1)


def test_sparse_idf_and_mask(engine, sparse_path):
    n_docs = 6000
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, 1500, 50, seed=5)
    csr = _oracle_csr(ptr, doc, w, n_docs)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    df = engine.sparse_df("s", 1500)
    assert np.array_equal(df, np.diff(ptr))
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n_docs, 9, 6, "doc", seed=3)
    mask = synth.random_mask(n_docs, 0.10)
    engine.mask_set(4, mask)
    vals_idf = np.concatenate([O.apply_idf(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]])
                               for b in range(9)])
    s, i, c = engine.sparse_topk("s", q_ptr, q_terms, vals_idf, 50, mask_id=4)
    for b in range(9):
        t = q_terms[q_ptr[b]:q_ptr[b + 1]]
        es, ei = O.sparse_topk(csr, t, vals_idf[q_ptr[b]:q_ptr[b + 1]], 50, mask=mask)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei)
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))


def test_sparse_large_k(engine, sparse_path):
    """k up to 512 (fewer worker warps per CTA: the per-warp candidate list grows)."""
    n_docs = 30000
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, 3000, 60, seed=17)
    csr = _oracle_csr(ptr, doc, w, n_docs)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n_docs, 5, 8, "zipf", seed=18, skip_top=8)
    for k in (200, 512):
        s, i, c = engine.sparse_topk("s", q_ptr, q_terms, q_vals, k)
        for b in range(5):
            es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k)
            assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei)
            assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))


def test_sparse_many_terms_per_query(engine):
    """More than one 16-term group per query (document-sized queries)."""
    n_docs = 3000
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, 800, 80, seed=6)
    csr = _oracle_csr(ptr, doc, w, n_docs)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    rng = np.random.default_rng(1)
    terms = np.sort(rng.choice(800, 45, replace=False)).astype(np.int32)
    vals = rng.random(45) * 2 - 0.5     # includes negative weights
    s, i, c = engine.sparse_topk("s", np.array([0, 45], np.int32), terms, vals, 64)
    es, ei = O.sparse_topk(csr, terms, vals, 64)
    assert np.array_equal(i[0][:len(ei)], ei)
    assert np.array_equal(s[0][:len(ei)].view(np.uint32), es.view(np.uint32))


@pytest.mark.parametrize("n_terms", [12, 16, 17, 32])
def test_sparse_bitmap_term_groups_and_signs(engine, n_terms):
    """The bitmap scan's 16- and 32-term variants, with negative and zero query values (a term's contribution
    is bounded by max(0, qv*maxw, qv*minw), so pruning stays rank-safe for any sign)."""
    n_docs = 9000
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, 900, 70, seed=16)
    w = (w * np.where(np.arange(len(w)) % 7 == 0, -1.0, 1.0)).astype(np.float32)    # some negative weights too
    csr = _oracle_csr(ptr, doc, w, n_docs)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    rng = np.random.default_rng(n_terms)
    B = 7
    q_ptr = np.arange(0, (B + 1) * n_terms, n_terms, dtype=np.int32)
    q_terms = np.concatenate([np.sort(rng.choice(900, n_terms, replace=False)) for _ in range(B)]).astype(np.int32)
    q_vals = rng.random(B * n_terms) * 2 - 0.6
    q_vals[::5] = 0.0
    engine.force_sparse_path("bitmap")
    try:
        s, i, c = engine.sparse_topk("s", q_ptr, q_terms, q_vals, 40)
    finally:
        engine.force_sparse_path(None)
    for b in range(B):
        es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], 40)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei), f"query {b}"
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))


@pytest.mark.parametrize("masked", [False, True])
def test_sparse_bitmap_warm_bound_large(engine, masked):
    """Enough documents for the global bound to warm up and the count filter to prune (dense-term maps, many
    sub-tiles, several splits); both kernels must agree with the oracle and with each other."""
    n_docs = 150000
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, 3000, 100, seed=21)
    csr = _oracle_csr(ptr, doc, w, n_docs)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    B = 48
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n_docs, B, 8, "zipf", seed=22, skip_top=16)
    mask = synth.random_mask(n_docs, 0.10) if masked else None
    if masked:
        engine.mask_set(6, mask)
    out = {}
    for path in ("scatter", "bitmap"):
        engine.force_sparse_path(path)
        try:
            out[path] = [engine.sparse_topk("s", q_ptr, q_terms, q_vals, k, mask_id=6 if masked else 0) for k in (10, 100)]
        finally:
            engine.force_sparse_path(None)
    for kk, k in enumerate((10, 100)):
        s, i, c = out["bitmap"][kk]
        s0, i0, c0 = out["scatter"][kk]
        assert np.array_equal(i, i0) and np.array_equal(c, c0) and np.array_equal(s.view(np.uint32), s0.view(np.uint32))
        for b in range(B):
            es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k, mask=mask)
            assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei), f"query {b}"
            assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))


@pytest.mark.parametrize("B", [1, 40])
def test_sparse_filter_compacted_index(engine, B):
    """A filter that keeps <= half of the documents gets, once it has served enough queries, an index of just the
    passing documents (compact.cu): ids, scores and counts equal the masked scan of the full space and the oracle,
    through both scan kernels and through the fused hybrid call; a new filter under the same mask_id rebuilds it."""
    n_docs, k, d = 120000, 40, 64
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, 3000, 60, seed=51)
    csr = _oracle_csr(ptr, doc, w, n_docs)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    x = synth.dense_corpus(n_docs, d, seed=53)
    q = synth.dense_queries(x, B, seed=54)
    engine.dense_set("x", x, "bf16")
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n_docs, B, 8, "zipf", seed=52, skip_top=8)
    for seed, sel in ((61, 0.10), (62, 0.35)):
        mask = synth.random_mask(n_docs, sel, seed=seed)
        engine.mask_set(7, mask)
        got = {}
        try:
            engine.sparse_compact_after(-1)
            full = engine.sparse_topk("s", q_ptr, q_terms, q_vals, k, mask_id=7)
            hyb_full = engine.hybrid_topk("x", "s", q, q_ptr, q_terms, q_vals, k, mask_id=7)
            engine.sparse_compact_after(0)
            for path in (None, "scatter", "bitmap"):
                engine.force_sparse_path(path)
                got[path] = engine.sparse_topk("s", q_ptr, q_terms, q_vals, k, mask_id=7)
            engine.force_sparse_path(None)
            hyb = engine.hybrid_topk("x", "s", q, q_ptr, q_terms, q_vals, k, mask_id=7)
        finally:
            engine.force_sparse_path(None)
            engine.sparse_compact_after(256)
        s0, i0, c0 = full
        for path, (s, i, c) in got.items():
            assert np.array_equal(i, i0) and np.array_equal(c, c0), path
            assert np.array_equal(s.view(np.uint32), s0.view(np.uint32)), path
        for a, b in zip(hyb, hyb_full):
            assert np.array_equal(np.asarray(a), np.asarray(b), equal_nan=np.asarray(a).dtype.kind == "f")
        assert mask[i0[np.arange(k)[None, :] < c0[:, None]]].all()
        for b in range(0, B, max(1, B // 8)):
            es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k, mask=mask)
            assert int(c0[b]) == len(ei) and np.array_equal(i0[b][:len(ei)], ei), f"query {b}"
            assert np.array_equal(s0[b][:len(ei)].view(np.uint32), es.view(np.uint32))


def test_sparse_append_equals_a_fresh_upload(engine):
    """brag_sparse_append splices the postings of new documents into the index in HBM (old lists moved on
    the device, vocabulary rows inserted in the middle of the term order, scan indexes rebuilt); every answer must be
    bit-identical to a fresh upload of the grown corpus, through both scan kernels."""
    n0, n1, V, k = 90000, 100000, 3000, 100
    ptr, doc, w, _ = synth.zipf_corpus(n1, V, 60, seed=71)
    term = np.repeat(np.arange(V), np.diff(ptr))
    late = (term % 7) == 3                         # terms that only the appended documents use
    keep = ~(late & (doc < n0))
    term, doc, w = term[keep], doc[keep], w[keep]

    def csr(sel, rows, n_rows):
        p = np.zeros(n_rows + 1, np.int64)
        np.add.at(p, rows[sel] + 1, 1)
        return np.cumsum(p), doc[sel].astype(np.uint32), w[sel].astype(np.float32)

    full_ptr, full_doc, full_w = csr(np.ones(len(term), bool), term, V)
    old_terms = np.flatnonzero((np.arange(V) % 7) != 3)
    old_row = np.full(V, -1, np.int64)
    old_row[old_terms] = np.arange(len(old_terms))
    is_old = doc < n0
    o_ptr, o_doc, o_w = csr(is_old, old_row[term], len(old_terms))
    d_ptr, d_doc, d_w = csr(~is_old, term, V)
    engine.sparse_set_csr("f", full_ptr, full_doc, full_w, n1)
    engine.sparse_set_csr("a", o_ptr, o_doc, o_w, n0)
    engine.sparse_append("a", d_ptr, d_doc, d_w, old_terms.astype(np.int32), n1)
    assert np.array_equal(engine.sparse_df("a", V), engine.sparse_df("f", V))
    B = 48
    q_ptr, q_terms, q_vals = synth.sparse_queries(full_ptr, full_doc, n1, B, 8, "zipf", seed=72, skip_top=16)
    csr_o = _oracle_csr(full_ptr, full_doc, full_w, n1)
    for path in ("scatter", "bitmap"):
        engine.force_sparse_path(path)
        try:
            sa, ia, ca = engine.sparse_topk("a", q_ptr, q_terms, q_vals, k)
            sf, if_, cf = engine.sparse_topk("f", q_ptr, q_terms, q_vals, k)
        finally:
            engine.force_sparse_path(None)
        assert np.array_equal(ia, if_) and np.array_equal(ca, cf) and np.array_equal(sa.view(np.uint32), sf.view(np.uint32))
        assert (ia[:, 0] >= 0).all() and (ia >= n0).any()
    for b in range(0, B, 6):
        es, ei = O.sparse_topk(csr_o, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k)
        assert int(ca[b]) == len(ei) and np.array_equal(ia[b][:len(ei)], ei)
        assert np.array_equal(sa[b][:len(ei)].view(np.uint32), es.view(np.uint32))
    with pytest.raises(Exception):                 # documents below the current count cannot be appended
        engine.sparse_append("a", d_ptr, d_doc, d_w, np.arange(V, dtype=np.int32), n1 + 10)
    engine.sparse_drop("a")
    engine.sparse_drop("f")


def test_sparse_compacted_index_lifecycle(engine):
    """Default policy: the index of a filter's passing documents appears once that filter has served 256 batched
    queries; single queries under other filters (the plan executor's candidate / page masks) neither count nor evict
    it; a batch under another filter, or new contents under the same mask_id, start over.  Answers stay exact."""
    n_docs, k, B = 60000, 20, 40
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, 2000, 50, seed=101)
    csr = _oracle_csr(ptr, doc, w, n_docs)
    engine.sparse_set_csr("s", ptr, doc, w, n_docs)
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n_docs, B, 8, "zipf", seed=102, skip_top=8)
    m7, m8 = synth.random_mask(n_docs, 0.2, seed=103), synth.random_mask(n_docs, 0.3, seed=104)
    engine.mask_set(7, m7)
    engine.mask_set(8, m8)

    def exact(out, mask, nq):
        s, i, c = out
        for b in range(nq):
            es, ei = O.sparse_topk(csr, q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k, mask=mask)
            assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei)
            assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32))

    for call in range(8):
        assert engine.sparse_compact_info("s")[0] == 0, call           # 7 x 40 queries counted, none built yet
        out = engine.sparse_topk("s", q_ptr, q_terms, q_vals, k, mask_id=7)
    assert engine.sparse_compact_info("s") == (7, int(m7.sum()))
    exact(out, m7, 6)
    one = engine.sparse_topk("s", q_ptr[:2], q_terms[:q_ptr[1]], q_vals[:q_ptr[1]], k, mask_id=8)
    exact(one, m8, 1)
    assert engine.sparse_compact_info("s")[0] == 7                      # a single query under another filter: kept
    exact(engine.sparse_topk("s", q_ptr, q_terms, q_vals, k, mask_id=7), m7, 6)
    exact(engine.sparse_topk("s", q_ptr, q_terms, q_vals, k, mask_id=8), m8, 6)
    assert engine.sparse_compact_info("s")[0] == 0                      # a batch under another filter starts over
    engine.mask_set(7, m8)                                              # same id, new contents
    exact(engine.sparse_topk("s", q_ptr, q_terms, q_vals, k, mask_id=7), m8, 6)


def test_sparse_nonfinite_weights_use_the_scatter_scan(engine):
    """A NaN/Inf weight makes per-term bounds meaningless: the space falls back to the exhaustive kernel."""
    ptr, doc, w, _ = synth.zipf_corpus(2000, 300, 30, seed=3)
    w = w.copy()
    w[5] = np.inf
    engine.sparse_set_csr("s", ptr, doc, w, 2000)
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, 2000, 3, 4, "zipf", seed=4, skip_top=2)
    s, i, c = engine.sparse_topk("s", q_ptr, q_terms, q_vals, 10)      # auto: must not raise
    engine.force_sparse_path("bitmap")
    try:
        with pytest.raises(Exception):
            engine.sparse_topk("s", q_ptr, q_terms, q_vals, 10)
    finally:
        engine.force_sparse_path(None)


# ----------------------------------------------------------------------------- fusion / merge
@pytest.mark.parametrize("tie", ["id", "first_seen"])
def test_rrf_ranks_exact(engine, tie):
    rng = np.random.default_rng(5)
    L, B, kmax = 3, 17, 60
    ids = np.full((L, B, kmax), -1, np.int64)
    counts = np.zeros((L, B), np.int32)
    for l in range(L):
        for b in range(B):
            c = int(rng.integers(0, kmax + 1))
            counts[l, b] = c
            ids[l, b, :c] = rng.choice(150, size=c, replace=False)
    for limit in (10, 100):
        s, i, c = engine.rrf(ids, counts, limit, tie=0 if tie == "id" else 1)
        for b in range(B):
            es, ei = O.rrf([ids[l, b, :counts[l, b]] for l in range(L)], limit=limit, tie=tie)
            assert int(c[b]) == len(ei)
            assert np.array_equal(i[b][:len(ei)], ei)
            assert np.array_equal(s[b][:len(ei)], es), "float64 RRF scores must be bit-exact"


def test_merge_topk_matches_global_sort(engine):
    rng = np.random.default_rng(9)
    G, B, k = 8, 5, 100
    scores = np.full((G, B, k), -np.inf, np.float32)
    ids = np.full((G, B, k), -1, np.int64)
    counts = np.zeros((G, B), np.int32)
    for g in range(G):
        for b in range(B):
            c = int(rng.integers(0, k + 1))
            sc = np.sort(rng.integers(0, 50, c).astype(np.float32))[::-1]   # many ties
            counts[g, b] = c
            scores[g, b, :c] = sc
            ids[g, b, :c] = g * 1000 + np.arange(c)
    s, i, c = engine.merge_topk(scores, ids, counts)
    for b in range(B):
        fs = np.concatenate([scores[g, b, :counts[g, b]] for g in range(G)])
        fi = np.concatenate([ids[g, b, :counts[g, b]] for g in range(G)])
        o = O.topk_order(fs, fi, k)
        assert int(c[b]) == len(o)
        assert np.array_equal(i[b][:len(o)], fi[o]) and np.array_equal(s[b][:len(o)], fs[o])


def test_hybrid_three_stage(engine):
    """north_star path in one call == oracle dense top-k + sparse top-k + RRF."""
    n, d, k, B = 8000, 384, 100, 12
    x = synth.dense_corpus(n, d, seed=41, clustered=True)
    q = synth.dense_queries(x, B, seed=42)
    ptr, doc, w, _ = synth.zipf_corpus(n, 3000, 80, seed=43)
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=44, skip_top=8)
    col = O.OracleCollection()
    col.add_dense("dense", x, "dot", "bf16")
    col.add_sparse("sparse", O.CSR(ptr, doc, w, n))
    engine.dense_set("dense", x, "bf16")
    engine.sparse_set_csr("sparse", ptr, doc, w, n)
    mask = synth.random_mask(n, 0.3)
    engine.mask_set(5, mask)
    for mid, m in ((0, None), (5, mask)):
        s, i, c = engine.hybrid_topk("dense", "sparse", q, q_ptr, q_terms, q_vals, k, mask_id=mid)
        for b in range(B):
            es, ei = O.hybrid_simple(col, q[b], q_terms[q_ptr[b]:q_ptr[b + 1]], q_vals[q_ptr[b]:q_ptr[b + 1]], k, mask=m)
            assert int(c[b]) == len(ei)
            assert np.array_equal(i[b][:len(ei)], ei), f"query {b} mask {mid}"
            assert np.array_equal(s[b][:len(ei)], es)


def test_hybrid_call_replays_from_a_cuda_graph(engine):
    """The device-pointer hybrid call is stream-ordered and, once warmed up, allocation- and sync-free, so a serving loop
    can capture it once and replay it per query (bench.py batch1.graph_replay_p50_ms): the replay must answer like the
    eager call, also after new query contents were copied into the captured buffers."""
    import torch
    n, d, k = 20000, 256, 50
    x = synth.dense_corpus(n, d, seed=61, clustered=True)
    q = synth.dense_queries(x, 2, seed=62)
    ptr, doc, w, _ = synth.zipf_corpus(n, 3000, 80, seed=63)
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n, 2, 8, "zipf", seed=64, skip_top=8)
    engine.dense_set("dense", x, "bf16")
    engine.sparse_set_csr("sparse", ptr, doc, w, n)
    dev = torch.device("cuda", 0)
    tq = torch.from_numpy(q).to(dev)
    tt, tv = torch.from_numpy(q_terms).to(dev), torch.from_numpy(q_vals).to(dev)
    q1, qp1 = tq[:1].clone(), torch.tensor([0, 8], dtype=torch.int32, device=dev)
    qt1, qv1 = tt[:8].clone(), tv[:8].clone()

    def call():
        return engine.hybrid_topk("dense", "sparse", q1, qp1, qt1, qv1, k)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            call()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        captured = call()
    for b in range(2):
        q1.copy_(tq[b:b + 1]); qt1.copy_(tt[8 * b:8 * b + 8]); qv1.copy_(tv[8 * b:8 * b + 8])
        graph.replay()
        torch.cuda.synchronize()
        got = [t.clone() for t in captured]
        want = call()
        torch.cuda.synchronize()
        assert all(torch.equal(a, b_) for a, b_ in zip(got, want)), f"query {b}"
        assert int(got[2][0]) > 0


@pytest.mark.parametrize("G", [2, 4, 8])
def test_logical_shards_on_one_gpu_match_unsharded(engine, G):
    """Row sharding emulated on one device: G engines own disjoint row ranges and answer
    with global ids; gathered lists -> brag_merge_topk x2 -> brag_rrf must equal the unsharded hybrid call."""
    from rag_application_b200 import Engine
    from rag_application_b200.sharded import shard_bounds
    n, d, k, B, V = 9001, 128, 50, 20, 2000
    x = synth.dense_corpus(n, d, seed=51, clustered=True)
    q = synth.dense_queries(x, B, seed=52)
    ptr, doc, w, _ = synth.zipf_corpus(n, V, 50, seed=53)
    q_ptr, q_terms, q_vals = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=54, skip_top=8)
    engine.dense_set("dense", x, "bf16")
    engine.sparse_set_csr("sparse", ptr, doc, w, n)
    es, ei, ec = engine.hybrid_topk("dense", "sparse", q, q_ptr, q_terms, q_vals, k)
    term_of = np.repeat(np.arange(V), np.diff(ptr))
    lists = []
    shards = []
    for r in range(G):
        lo, hi = shard_bounds(n, G, r)
        e = Engine(0)
        shards.append(e)
        e.set_id_offset(lo)
        e.dense_set("dense", x[lo:hi], "bf16")
        keep = (doc >= lo) & (doc < hi)
        p = np.zeros(V + 1, np.int64)
        np.add.at(p, term_of[keep] + 1, 1)
        e.sparse_set_csr("sparse", np.cumsum(p), (doc[keep] - lo).astype(np.uint32), w[keep], hi - lo)
        lists.append((e.dense_topk("dense", q, k), e.sparse_topk("sparse", q_ptr, q_terms, q_vals, k)))
    merged = []
    for which in range(2):
        s = np.stack([lists[r][which][0] for r in range(G)])
        i = np.stack([lists[r][which][1] for r in range(G)])
        c = np.stack([lists[r][which][2] for r in range(G)])
        merged.append(engine.merge_topk(s, i, c))
    ids = np.stack([merged[0][1], merged[1][1]])
    cnt = np.stack([merged[0][2], merged[1][2]])
    s, i, c = engine.rrf(ids, cnt, k)
    # the same through the sharded tail kernels (brag_shard_local_lists on every shard -> brag_shard_fuse)
    keys = np.stack([e.shard_local_lists("dense", "sparse", q, q_ptr, q_terms, q_vals, k) for e in shards])
    fs, fi, fc, (ts, ti, tc) = engine.shard_fuse(keys, k, stages=True)
    for e in shards:
        e.close()
    assert np.array_equal(c, ec) and np.array_equal(i, ei) and np.array_equal(s, es)
    assert np.array_equal(fc, ec) and np.array_equal(fi, ei) and np.array_equal(fs, es)
    for l in range(2):
        assert np.array_equal(ti[l], merged[l][1]) and np.array_equal(tc[l], merged[l][2])
        assert np.array_equal(ts[l].view(np.uint32), merged[l][0].view(np.uint32))


def test_save_load_bit_equal(engine, tmp_path):
    """brag_save -> a new engine -> brag_load answers bit for bit like the engine that saved (dense through
    both scans, sparse through both kernels -- i.e. the saved scan indexes are the ones in use -- and the fused call)."""
    from rag_application_b200 import Engine
    n, d, k, B, V = 30_000, 128, 60, 40, 3000
    x = synth.dense_corpus(n, d, seed=71, clustered=True)
    q = synth.dense_queries(x, B, seed=72)
    ptr, doc, w, _ = synth.zipf_corpus(n, V, 70, seed=73)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=74, skip_top=8)
    a = Engine(0)
    b = Engine(0)
    try:
        a.set_id_offset(1000)
        a.dense_set("d16", x, "bf16")
        a.dense_set("d32", x[:5000], "f32")
        a.sparse_set_csr("s", ptr, doc, w, n)
        path = str(tmp_path / "index.brag")
        a.save(path)
        b.load(path, a.host_state())

        def answers(e):
            out = [e.dense_topk("d16", q, k), e.dense_topk("d32", q[:4], k), e.hybrid_topk("d16", "s", q, qp, qt, qv, k)]
            for p in ("scatter", "bitmap"):
                e.force_sparse_path(p)
                out.append(e.sparse_topk("s", qp, qt, qv, k))
            e.force_sparse_path(None)
            return out
        for ra, rb in zip(answers(a), answers(b)):
            for ta, tb in zip(ra, rb):
                assert np.array_equal(np.asarray(ta).view(np.uint8), np.asarray(tb).view(np.uint8))
        assert int(b.dense_topk("d16", q[:1], 1)[1][0, 0]) >= 1000            # the id offset came back too
        # the loaded spaces take appends like uploaded ones
        b.dense_append("d16", x[:16])
        assert b.dense_topk("d16", q[:2], k)[2].tolist() == [k, k]
        with pytest.raises(Exception):
            b.load(str(tmp_path / "missing.brag"))
    finally:
        a.close()
        b.close()


@pytest.mark.parametrize("masked", [False, True])
def test_sparse_rare_terms_without_offset_rows(engine, masked):
    """Terms whose postings do not outweigh a per-sub-tile offset row get none (the table was V x N/2048
    words whatever the vocabulary); the scans search their short lists instead.  Forced for every term here: both
    kernels, long and short queries, must answer bit for bit like the index with a full table, and like the oracle."""
    n_docs, V, L, k, B = 60_000, 4000, 60, 50, 48
    ptr, doc, w, _ = synth.zipf_corpus(n_docs, V, L, seed=201)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n_docs, B, 8, "zipf", seed=202, skip_top=8)
    qp2, qt2, qv2 = synth.sparse_queries(ptr, doc, n_docs, 6, 24, "doc", seed=203)       # > 16 terms: second term group
    mask = synth.random_mask(n_docs, 0.4) if masked else None
    if masked:
        engine.mask_set(6, mask)
    mid = 6 if masked else 0
    engine.sparse_set_csr("full", ptr, doc, w, n_docs)
    engine.sparse_offsets_min_df(1 << 40)
    engine.sparse_set_csr("rare", ptr, doc, w, n_docs)
    engine.sparse_offsets_min_df(300)                      # mixed: head terms keep rows, the tail is searched
    engine.sparse_set_csr("mixed", ptr, doc, w, n_docs)
    engine.sparse_offsets_min_df(0)
    csr = O.CSR(ptr, doc, w, n_docs)
    try:
        for path in ("scatter", "bitmap"):
            engine.force_sparse_path(path)
            for (a, b_, c_) in ((qp, qt, qv), (qp2, qt2, qv2)):
                want = engine.sparse_topk("full", a, b_, c_, k, mask_id=mid)
                for name in ("rare", "mixed"):
                    got = engine.sparse_topk(name, a, b_, c_, k, mask_id=mid)
                    assert np.array_equal(got[2], want[2]) and np.array_equal(got[1], want[1]), (path, name)
                    assert np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)), (path, name)
                for b in range(0, len(a) - 1, 5):
                    es, ei = O.sparse_topk(csr, b_[a[b]:a[b + 1]], c_[a[b]:a[b + 1]], k, mask=mask)
                    assert int(want[2][b]) == len(ei) and np.array_equal(want[1][b][:len(ei)], ei)
    finally:
        engine.force_sparse_path(None)
        for name in ("full", "rare", "mixed"):
            engine.sparse_drop(name)


def test_sparse_index_footprint_with_a_hashed_vocabulary(engine):
    """2 M distinct term rows (what murmur3-hashed stems give) over 300 k documents: the index must stay within a small
    multiple of the postings (round 1: 2 M x 148 offset words = 1.2 GB beside 100 MB of postings)."""
    import torch
    rng = np.random.default_rng(5)
    n_docs, V, per_doc = 300_000, 2_000_000, 40
    term = np.concatenate([rng.integers(0, 2000, (n_docs, per_doc // 2)),              # a frequent head ...
                           rng.integers(0, V, (n_docs, per_doc // 2))], axis=1)       # ... and a long hashed tail
    key = np.unique(term.astype(np.int64) * n_docs + np.arange(n_docs)[:, None])
    t, d = key // n_docs, (key % n_docs).astype(np.uint32)
    ptr = np.zeros(V + 1, np.int64)
    np.add.at(ptr, t + 1, 1)
    ptr = np.cumsum(ptr)
    w = (rng.random(len(d)).astype(np.float32) + 0.1)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    engine.sparse_set_csr("hv", ptr, d, w, n_docs)
    free1, _ = torch.cuda.mem_get_info()
    used, postings = free0 - free1, len(d) * 8
    assert used < 4 * postings + (64 << 20), (used, postings)
    q_terms = np.sort(np.concatenate([rng.choice(2000, 3, replace=False), 2000 + rng.choice(V - 2000, 5, replace=False)])).astype(np.int32)
    q_ptr = np.array([0, 8], np.int32)
    q_vals = np.ones(8)
    csr = O.CSR(ptr, d, w, n_docs)
    es, ei = O.sparse_topk(csr, q_terms, q_vals, 20)
    for path in ("scatter", "bitmap"):
        engine.force_sparse_path(path)
        s, i, c = engine.sparse_topk("hv", q_ptr, q_terms, q_vals, 20)
        assert int(c[0]) == len(ei) and np.array_equal(i[0][:len(ei)], ei) and np.array_equal(s[0][:len(ei)].view(np.uint32), es.view(np.uint32))
    engine.force_sparse_path(None)
    engine.sparse_drop("hv")
