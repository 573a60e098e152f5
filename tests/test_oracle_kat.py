"""CPU: the oracle against the committed known-answer vectors (tests/golden/kat.json, produced by
the independent brute force in tests/golden/make_golden.py) and against hand-computed values
(SURVEY.md 8c known-answer tests 1-5 and 7).  No GPU, no native library."""
import json
import math
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kat.json")


@pytest.fixture(scope="module")
def kat():
    with open(GOLD) as f:
        return json.load(f)


def pairs(scores, ids):
    return [[float(s), int(i)] for s, i in zip(scores, ids)]


def assert_ranked(got, exp, rtol=1e-6):
    assert [i for _, i in got] == [i for _, i in exp]
    np.testing.assert_allclose([s for s, _ in got], [s for s, _ in exp], rtol=rtol, atol=1e-7)


# ------------------------------------------------------------------ (1) dense, zero vector, exact tie
def test_dense_small(kat):
    c = kat["dense_small"]
    rows = np.array(c["rows"], np.float32)
    col = O.OracleCollection()
    col.add_dense("cos", rows, "cosine")
    col.add_dense("dot", rows, "dot")
    s, i = O.dense_topk(col.dense["cos"], col.prep_query("cos", c["q"]), 3)
    assert_ranked(pairs(s, i), c["cosine_top3"])
    assert list(i) == [0, 1, 3]                       # tie between rows 1 and 3 -> lower id first
    s, i = O.dense_topk(col.dense["dot"], col.prep_query("dot", c["q"]), 3)
    assert_ranked(pairs(s, i), c["dot_top3"])
    mask = np.array([0, 1, 1, 1, 1], bool)
    s, i = O.dense_topk(col.dense["cos"], col.prep_query("cos", c["q"]), 3, mask=mask)
    assert_ranked(pairs(s, i), c["cosine_top3_mask"])
    assert np.all(np.isfinite(col.dense["cos"][2]))   # zero row: eps guard, no NaN


# ------------------------------------------------------------------ (2) sparse overlap-only, IDF
def _csr(docs, n_terms):
    return O.build_csr([(d[0], d[1]) for d in docs], n_terms)


def test_sparse_small(kat):
    c = kat["sparse_small"]
    csr = _csr(c["docs"], 16)
    o = np.argsort(c["q_idx"])
    qi, qv = np.array(c["q_idx"])[o], np.array(c["q_val"])[o]
    s, i = O.sparse_topk(csr, qi, qv, 10)
    assert_ranked(pairs(s, i), c["top10"], rtol=0)
    assert 2 not in i and 5 not in i                  # no shared index -> not a candidate
    # hand-computed: doc0 = 1*2 + 2*1 + .5*.5 = 4.25 ; doc3 = .75 + .5 = 1.25 ; doc4 = 1 ; doc1 = .5
    assert pairs(s, i) == [[4.25, 0], [1.25, 3], [1.0, 4], [0.5, 1]]
    s, i = O.sparse_topk(csr, qi, qv, 2)
    assert_ranked(pairs(s, i), c["top2"], rtol=0)
    # Modifier.IDF
    vals = O.apply_idf(csr, qi, qv)
    exp = np.array(c["q_val_idf"])[o]
    np.testing.assert_allclose(vals, exp, rtol=1e-15)
    assert vals[0] == 2.0 * math.log(1.0 + (6 - 2 + 0.5) / (2 + 0.5))   # term 1: df=2, N=6
    s, i = O.sparse_topk(csr, qi, vals, 10)
    assert_ranked(pairs(s, i), c["top10_idf"], rtol=0)
    # two-pointer form agrees with the inverted-index form
    for d, (di, dv) in enumerate(c["docs"]):
        r = O.sparse_dot(qi, vals, di, np.array(dv, np.float32))
        if d in (2, 5):
            assert r is None
        else:
            assert r == s[list(i).index(d)]


# ------------------------------------------------------------------ (3) RRF
def test_rrf(kat):
    c = kat["rrf"]
    s, i = O.rrf(c["lists"])
    assert list(i) == [10, 12, 11, 13]                # a, c, b, d
    assert list(s) == [1 / 2 + 1 / 3, 1 / 4 + 1 / 2, 1 / 3, 1 / 4]
    assert_ranked(pairs(s, i), c["fused"], rtol=0)
    s, i = O.rrf(c["long_lists"])                     # default fusion limit
    assert len(i) == O.DEFAULT_FUSION_LIMIT == 10
    assert_ranked(pairs(s, i), c["long_default_limit"], rtol=0)
    s, i = O.rrf(c["long_lists"], limit=100)
    assert_ranked(pairs(s, i), c["long_limit_100"], rtol=0)
    # first-seen tie order (upstream's stable sort) differs from id order on exact ties
    s2, i2 = O.rrf([[5, 3], [9, 1]], tie="first_seen")
    assert list(i2) == [5, 9, 3, 1]
    s3, i3 = O.rrf([[5, 3], [9, 1]], tie="id")
    assert list(i3) == [5, 9, 1, 3]


# ------------------------------------------------------------------ (5) int8 "quantisation"
def test_quantize_trunc(kat):
    c = kat["quant"]
    assert O.quantize_int8(c["x"]).astype(np.float32).tolist() == c["q"]
    assert O.quantize_int8([0.999, -0.5, 1.0]).tolist() == [126, -63, 127]


def test_int8_cosine_by_hand_and_against_the_float_space(kat):
    """int8 STORAGE of the `quantized` space (BRAG_I8): hand-computed cosines, zero vectors, and how far the scores are
    from the same values held in a float32 COSINE space (float32 rounding of the norms only); Q1 of the reference tree
    is the same ranked list either way."""
    x = np.zeros((4, 8), np.float32)
    x[0, :2] = [3, 4]
    x[1, :2] = [4, 3]
    x[2, :3] = [-3, -4, 12]
    q = np.zeros(8, np.float32)
    q[:2] = [4, 3]
    s = O.int8_cosine_scores(x, q)
    assert s.tolist() == [np.float32(24 / 25), np.float32(1.0), np.float32(-24 / 65), 0.0]
    assert O.int8_cosine_scores(x, np.zeros(8, np.float32)).tolist() == [0.0] * 4
    es, ei = O.dense_topk(x, q, 3, int8=True)
    assert ei.tolist() == [1, 0, 3] and es.tolist() == [np.float32(1.0), np.float32(0.96), 0.0]
    c = kat["tree"]
    xt, qt = _tree_inputs(c)
    xi = np.stack([O.quantize_int8(r) for r in xt]).astype(np.float32)
    qi = O.quantize_int8(qt).astype(np.float32)
    f32_space = O.dense_scores(O.normalize_rows(xi), O.normalize_rows(qi[None])[0])
    i8_space = O.int8_cosine_scores(xi, qi)
    assert np.abs(f32_space - i8_space).max() <= 3e-7
    lim = c["search_params"]["quantized_limit"]
    _, ids = O.dense_topk(xi, qi, lim, int8=True)
    assert ids.tolist() == [r for _, r in c["stages"]["Q1"]]


# ------------------------------------------------------------------ (4) the reference's prefetch tree
def _tree_inputs(c):
    rng = np.random.Generator(np.random.PCG64(c["seed"]))
    n, d, V = c["n"], c["d"], c["V"]
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = (x[17] + 0.2 * rng.standard_normal(d)).astype(np.float32)
    q /= np.linalg.norm(q)
    return x, q


def build_reference_collection(x, docs, V):
    """The six named vectors of qdrant_handler.py:57-86 / :144-163."""
    col = O.OracleCollection()
    col.add_dense("dense", x, "cosine")
    for k in (64, 128, 256):
        col.add_dense(f"matryoshka_{k}", x[:, :k], "cosine")
    col.add_dense("quantized", np.stack([O.quantize_int8(r) for r in x]).astype(np.float32), "cosine")
    col.add_sparse("sparse", _csr(docs, V))
    return col


def test_reference_tree(kat):
    c = kat["tree"]
    x, q = _tree_inputs(c)
    col = build_reference_collection(x, c["docs"], c["V"])
    sp = c["search_params"]
    st = c["stages"]
    # stage by stage
    s, i = O.run_node(col, O.Node(list(q[:64]), "matryoshka_64", sp["matryoshka_64_limit"]))
    assert_ranked(pairs(s, i), st["P1"])
    qq = O.quantize_int8(q).astype(np.float32).tolist()
    s, i = O.run_node(col, O.Node(qq, "quantized", sp["quantized_limit"]))
    assert_ranked(pairs(s, i), st["Q1"])
    s, i = O.run_node(col, O.Node(("sparse", c["q_idx"], c["q_val"]), "sparse", sp["sparse_limit"]))
    assert_ranked(pairs(s, i), st["S1"], rtol=0)
    # whole tree
    s, i = O.run_node(col, O.reference_tree(q, c["q_idx"], c["q_val"], sp))
    assert_ranked(pairs(s, i), c["final"])
    mask = np.array(c["mask"], bool)
    s, i = O.run_node(col, O.reference_tree(q, c["q_idx"], c["q_val"], sp, mask=mask))
    assert_ranked(pairs(s, i), c["final_masked"])
    assert all(mask[j] for j in i)
    # server semantics: the root filter reaches every prefetch
    s, i = O.run_node(col, O.reference_tree(q, c["q_idx"], c["q_val"], sp, mask=mask), propagate=True)
    assert_ranked(pairs(s, i), c["final_masked_server"])
    assert all(mask[j] for j in i) and len(i) == sp["final_limit"]


# ------------------------------------------------------------------ (7) shard invariance (oracle level)
@pytest.mark.parametrize("G", [2, 4, 8])
def test_shard_invariance(G):
    from rag_application_b200 import synth
    from rag_application_b200.sharded import shard_bounds
    n, d, k = 3000, 64, 20
    x = synth.dense_corpus(n, d, seed=3)
    q = synth.dense_queries(x, 1, seed=4)[0]
    ptr, doc, w, _ = synth.zipf_corpus(n, 400, 30, seed=5)
    csr = O.CSR(ptr, doc, w, n)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n, 1, 6, "zipf", seed=6, skip_top=4)
    es, ei = O.dense_topk(x, q, k)
    ss, si = O.sparse_topk(csr, qt, qv, k)
    ds_all, di_all, ss_all, si_all = [], [], [], []
    term_of = np.repeat(np.arange(400), np.diff(ptr))
    for r in range(G):
        lo, hi = shard_bounds(n, G, r)
        s, i = O.dense_topk(x[lo:hi], q, k)
        ds_all.append(s)
        di_all.append(i + lo)
        keep = (doc >= lo) & (doc < hi)
        sub = O.build_csr([], 400)
        p = np.zeros(401, np.int64)
        np.add.at(p, term_of[keep] + 1, 1)
        sub = O.CSR(np.cumsum(p), (doc[keep] - lo).astype(np.uint32), w[keep], hi - lo)
        s, i = O.sparse_topk(sub, qt, qv, k)
        ss_all.append(s)
        si_all.append(i + lo)
    for parts_s, parts_i, exp_s, exp_i in ((ds_all, di_all, es, ei), (ss_all, si_all, ss, si)):
        fs, fi = np.concatenate(parts_s), np.concatenate(parts_i)
        o = O.topk_order(fs, fi, k)
        assert np.array_equal(fi[o], exp_i) and np.array_equal(fs[o], exp_s)


# ------------------------------------------------------------------ storage rounding / BM25 weights
def test_round_storage_bf16_rne():
    x = np.array([1.0, 1.00390625, 1.01171875, -2.5, 3.1415927], np.float32)
    r = O.round_storage(x, "bf16")
    assert r[0] == 1.0 and r[1] == 1.0 and r[2] == np.float32(1.015625) and r[3] == -2.5
    import torch
    t = torch.from_numpy(x).to(torch.bfloat16).float().numpy()
    assert np.array_equal(r, t)
    assert np.array_equal(O.round_storage(x, "f16"), x.astype(np.float16).astype(np.float32))


def test_bm25_tf_weight_formula():
    # fastembed Qdrant/bm25 defaults k1=1.2 b=0.75 avgdl=256
    w = O.bm25_tf_weight([1, 2, 3], [256, 256, 128])
    exp = [1 * 2.2 / (1 + 1.2), 2 * 2.2 / (2 + 1.2), 3 * 2.2 / (3 + 1.2 * (0.25 + 0.75 * 0.5))]
    np.testing.assert_allclose(w, np.array(exp, np.float32), rtol=1e-7)
    assert w.dtype == np.float32


def test_fast_cpu_matches_oracle():
    """The timed CPU baseline (oracle/fast_cpu.py + bm25_cpu.c) returns what the oracle returns."""
    from oracle import fast_cpu
    from rag_application_b200 import synth
    n, d, k, B = 5000, 128, 50, 6
    x = O.round_storage(synth.dense_corpus(n, d, seed=1, clustered=True), "bf16")
    q = O.round_storage(synth.dense_queries(x, B, seed=2), "bf16")
    ptr, doc, w, _ = synth.zipf_corpus(n, 800, 40, seed=3)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=4, skip_top=4)
    mask = synth.random_mask(n, 0.3)
    ds, di = fast_cpu.dense_topk_batch(x, q, k)
    ss, si, sc = fast_cpu.sparse_topk_batch(ptr, doc, w, n, qp, qt, qv, k, mask=mask, threads=2)
    csr = O.CSR(ptr, doc, w, n)
    for b in range(B):
        es, ei = O.dense_topk(x, q[b], k)
        assert np.array_equal(di[b], ei) and np.array_equal(ds[b], es)
        es, ei = O.sparse_topk(csr, qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]], k, mask=mask)
        assert sc[b] == len(ei) and np.array_equal(si[b][:len(ei)], ei)
        assert np.array_equal(ss[b][:len(ei)].view(np.uint32), es.view(np.uint32))
