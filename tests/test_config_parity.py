"""GPU parity at the sizes the numbers are quoted on (BASELINE.json configs C1-C5), through the C ABI.

C1 is small enough for the oracle itself (and for the FAITHFUL qdrant-local restatement: float32 `X @ q` + argsort, one
two-pointer sparse loop per document).  From C2 up the checker is tools/torch_checker.py -- plain torch on the device,
float64, written from the definitions, sharing no code with libbrag.so (tests/test_checker_cpu.py pins it to the oracle);
C3 is additionally compared bit for bit with the oracle's C inverted index on all 1024 queries.
Bars: sparse / RRF ids, counts and float32 score bits exact; dense ids equal except near-ties of the checker's own
scores (|gap| <= 4e-7 relative), dense scores within 1e-3 relative (north_star) of the float64 dot with the FLOAT32
query."""
import numpy as np
import pytest

from oracle import oracle as O
from rag_application_b200 import synth
from tools import torch_checker as T

pytestmark = pytest.mark.gpu

SAMPLE = 16


def _sample(B, n=SAMPLE):
    return sorted(set(np.linspace(0, B - 1, n).astype(int).tolist()))


def _check_dense(X, q_dev, rows, got, k, mask=None, id_offset=0):
    """got = (scores, ids, counts) numpy for the whole batch; rows = sampled query indices."""
    s, i, c = got
    n_equal = 0
    for b in rows:
        es, ei, all_s = T.dense_topk(X, q_dev[b], k, mask=mask, id_offset=id_offset)
        assert int(c[b]) == len(ei), (b, int(c[b]), len(ei))
        r = T.dense_lists_agree(s[b], i[b], es, ei, all_scores=all_s, id_offset=id_offset)
        assert r["explained"], (b, r, i[b][:8], ei[:8])
        assert r["max_rel_err"] <= 1e-3
        n_equal += r["ids_equal"]
    return n_equal


def _check_sparse(tp, doc, w, n_docs, qp, qt, qv, rows, got, k, mask=None, id_offset=0):
    s, i, c = got
    for b in rows:
        es, ei = T.sparse_topk(tp, doc, w, n_docs, qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]], k, mask=mask, id_offset=id_offset)
        assert int(c[b]) == len(ei), (b, int(c[b]), len(ei))
        assert np.array_equal(i[b][:len(ei)], ei), b
        assert np.array_equal(s[b][:len(ei)].view(np.uint32), es.view(np.uint32)), b


def _np(ts):
    return [t.cpu().numpy() for t in ts]


# ------------------------------------------------------------------------------------------ C1
def test_c1_reference_cpu_config(engine):
    """configs[0]: 10 000 chunks, 384-d float32 COSINE + BM25 sparse, RRF top-10, 100 queries -- the reference's own
    CPU-runnable case (request shape: qdrant_handler.py:296-372, float32 vectors :57-79)."""
    n, d, k, B = 10_000, 384, 10, 100
    x = O.normalize_rows(synth.dense_corpus(n, d, seed=synth.SEED_CORPUS))
    q = O.normalize_rows(synth.dense_queries(x, B, seed=synth.SEED_QUERY))
    ptr, doc, w, _ = synth.zipf_corpus(n, 50_000, 285, seed=synth.SEED_CORPUS)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=synth.SEED_QUERY)
    engine.dense_set("c1d", x, "f32")
    engine.sparse_set_csr("c1s", ptr, doc, w, n)
    assert engine.dense_path("c1d", B) == "tcgen05"               # float32 storage: the kind::tf32 scan
    s, i, c = engine.hybrid_topk("c1d", "c1s", q, qp, qt, qv, k)
    ds, di, dc = engine.dense_topk("c1d", q, k)
    ss, si, sc = engine.sparse_topk("c1s", qp, qt, qv, k)
    col = O.OracleCollection()
    col.add_dense("dense", x, "cosine", "f32")
    col.add_sparse("sparse", O.CSR(ptr, doc, w, n))
    for b in range(B):
        t, v = qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]]
        es, ei = O.hybrid_simple(col, q[b], t, v, k)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei), b
        assert np.array_equal(s[b][:len(ei)], es)
    # the faithful path (what qdrant-local executes): float32 matmul + full argsort; per-document two-pointer loop
    order = np.argsort(doc, kind="stable")
    term_of = np.repeat(np.arange(len(ptr) - 1), np.diff(ptr))
    dptr = np.searchsorted(doc[order], np.arange(n + 1))
    for b in range(0, B, 20):
        f32 = x @ q[b]
        top = np.argsort(f32)[::-1][:k]
        assert set(top.tolist()) == set(di[b].tolist()) or np.allclose(np.sort(f32[top]), np.sort(f32[di[b]]), rtol=0, atol=4e-7)
        np.testing.assert_allclose(ds[b], f32[di[b]], rtol=1e-3)
        t, v = qt[qp[b]:qp[b + 1]].tolist(), qv[qp[b]:qp[b + 1]].tolist()
        scored = []
        for dd in range(n):
            seg = order[dptr[dd]:dptr[dd + 1]]
            r = O.sparse_dot(t, v, term_of[seg].tolist(), w[seg].tolist())
            if r is not None:
                scored.append((-float(r), dd, r))
        scored.sort()
        exp_i = np.array([e[1] for e in scored[:k]], np.int64)
        exp_s = np.array([e[2] for e in scored[:k]], np.float32)
        assert int(sc[b]) == len(exp_i) and np.array_equal(si[b][:len(exp_i)], exp_i)
        assert np.array_equal(ss[b][:len(exp_i)].view(np.uint32), exp_s.view(np.uint32))
    engine.dense_drop("c1d")
    engine.sparse_drop("c1s")


# ------------------------------------------------------------------------------------------ C2
def test_c2_dense_1m_768_f16():
    """configs[1]: dense exact top-100, 1 M x 768 float16, batch 1 and 1024."""
    import torch
    from rag_application_b200 import Engine
    dev = torch.device("cuda", 0)
    eng = Engine(0)
    try:
        n, d, k = 1 << 20, 768, 100
        x = synth.dense_corpus_torch(n, d, torch.float16, dev)
        q = synth.dense_queries_torch(x, 1024)
        eng.dense_set("x", x)
        got = _np(eng.dense_topk("x", q, k))
        assert (got[2] == k).all()
        neq = _check_dense(x, q, _sample(1024, 32), got, k)
        assert neq >= 30                                    # near-tie swaps are rare
        for b in (0, 511):
            one = _np(eng.dense_topk("x", q[b:b + 1].contiguous(), k))
            assert np.array_equal(one[1][0], got[1][b]) and np.array_equal(one[0][0].view(np.uint32), got[0][b].view(np.uint32))
        # odd numbers of query blocks (no shared row ranges) and a ragged batch answer like the full batch
        for B in (128, 384, 1000):
            part = _np(eng.dense_topk("x", q[:B].contiguous(), k))
            assert np.array_equal(part[1], got[1][:B]) and np.array_equal(part[0].view(np.uint32), got[0][:B].view(np.uint32))
    finally:
        eng.close()


# ------------------------------------------------------------------------------------------ C3
def test_c3_bm25_1m_docs():
    """configs[2]: BM25 top-100 over 1 M docs, 50 k-term Zipf vocabulary, 8-term queries, batch 1024 -- every query
    against the oracle's C inverted index (bit-exact ids and float32 scores), a sample against the torch checker."""
    import torch
    from oracle import fast_cpu
    from rag_application_b200 import Engine
    dev = torch.device("cuda", 0)
    eng = Engine(0)
    try:
        n, k, B = 1_000_000, 100, 1024
        tp, doc, w = synth.zipf_corpus_torch(n, 50_000, 285, dev)
        qp, qt, qv = synth.sparse_queries_torch(tp, B, 8)
        eng.sparse_set_csr("s", tp, doc, w, n)
        qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
        got = _np(eng.sparse_topk("s", qp_d, qt_d, qv_d, k, max_terms=8))
        qpn, qtn, qvn = qp.numpy(), qt.numpy(), qv.numpy()
        es, ei, ec = fast_cpu.sparse_topk_batch(tp.cpu().numpy(), doc.cpu().numpy().astype(np.uint32), w.cpu().numpy(), n,
                                                qpn, qtn, qvn, k)
        assert np.array_equal(got[2], ec)
        for b in range(B):
            c = int(ec[b])
            assert np.array_equal(got[1][b][:c], ei[b][:c]), b
            assert np.array_equal(got[0][b][:c].view(np.uint32), es[b][:c].view(np.uint32)), b
        _check_sparse(tp, doc, w, n, qpn, qtn, qvn, _sample(B), got, k)
        # without the max_terms promise (engine reads q_ptr back) and through host buffers: same answer
        again = _np(eng.sparse_topk("s", qp_d, qt_d, qv_d, k))
        host = eng.sparse_topk("s", qpn, qtn, qvn, k)
        for a in (again, host):
            assert np.array_equal(a[1], got[1]) and np.array_equal(np.asarray(a[0]).view(np.uint32), got[0].view(np.uint32))
    finally:
        eng.close()


# ------------------------------------------------------------------------------------------ C4
@pytest.fixture(scope="module")
def c4():
    """configs[3] on one GPU: 10 M x 1024 bfloat16 + 10 M documents (2.1 G postings), batch 1024, top-100."""
    import torch
    from rag_application_b200 import Engine
    dev = torch.device("cuda", 0)
    eng = Engine(0)
    n, d, B = 10_000_000, 1024, 1024
    x = synth.dense_corpus_torch(n, d, torch.bfloat16, dev)
    q = synth.dense_queries_torch(x, B)
    eng.dense_set("dense", x)
    tp, doc, w = synth.zipf_corpus_torch(n, 50_000, 285, dev)
    qp, qt, qv = synth.sparse_queries_torch(tp, B, 8)
    eng.sparse_set_csr("sparse", tp, doc, w, n)
    g = torch.Generator(device=dev)
    g.manual_seed(synth.SEED_MASK)
    mask = torch.rand(n, generator=g, device=dev) < 0.10
    eng.mask_set(1, mask.cpu().numpy())
    ctx = dict(eng=eng, n=n, d=d, B=B, k=100, x=x, q=q, tp=tp, doc=doc, w=w, qp=qp, qt=qt, qv=qv, mask=mask, dev=dev)
    yield ctx
    eng.close()
    ctx.clear()
    torch.cuda.empty_cache()


@pytest.mark.parametrize("masked", [False, True])
def test_c4_dense_stage(c4, masked):
    got = _np(c4["eng"].dense_topk("dense", c4["q"], c4["k"], mask_id=1 if masked else 0))
    assert (got[2] == c4["k"]).all() and (np.diff(got[0], axis=1) <= 0).all()
    _check_dense(c4["x"], c4["q"], _sample(c4["B"]), got, c4["k"], mask=c4["mask"] if masked else None)


@pytest.mark.parametrize("masked", [False, True])
def test_c4_sparse_stage(c4, masked):
    dev = c4["dev"]
    got = _np(c4["eng"].sparse_topk("sparse", c4["qp"].to(dev), c4["qt"].to(dev), c4["qv"].to(dev), c4["k"],
                                    mask_id=1 if masked else 0, max_terms=8))
    _check_sparse(c4["tp"], c4["doc"], c4["w"], c4["n"], c4["qp"].numpy(), c4["qt"].numpy(), c4["qv"].numpy(),
                  _sample(c4["B"]), got, c4["k"], mask=c4["mask"] if masked else None)


def test_c4_hybrid_step(c4):
    """The benchmarked call (brag_hybrid_topk, device pointers, batch 1024): fused list == RRF of the checker's lists."""
    dev, k = c4["dev"], c4["k"]
    qp_d, qt_d, qv_d = c4["qp"].to(dev), c4["qt"].to(dev), c4["qv"].to(dev)
    s, i, c = _np(c4["eng"].hybrid_topk("dense", "sparse", c4["q"], qp_d, qt_d, qv_d, k, max_terms=8))
    qp, qt, qv = c4["qp"].numpy(), c4["qt"].numpy(), c4["qv"].numpy()
    dense = _np(c4["eng"].dense_topk("dense", c4["q"], k))
    for b in _sample(c4["B"]):
        ds, di, all_s = T.dense_topk(c4["x"], c4["q"][b], k)
        r = T.dense_lists_agree(dense[0][b], dense[1][b], ds, di, all_scores=all_s)
        assert r["explained"]
        ss, si = T.sparse_topk(c4["tp"], c4["doc"], c4["w"], c4["n"], qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]], k)
        es, ei = T.rrf([dense[1][b] if not r["ids_equal"] else di, si], k)
        assert int(c[b]) == len(ei) and np.array_equal(i[b][:len(ei)], ei), b
        assert np.array_equal(s[b][:len(ei)], es)


def test_c4_sharded_kernels_on_one_gpu(c4):
    """The sharded tail (brag_shard_local_lists -> brag_shard_fuse) on the full-size corpus, one shard: same fused answer
    as brag_hybrid_topk, and the merged stage lists are the per-stage calls' lists."""
    import torch
    dev, k, eng = c4["dev"], c4["k"], c4["eng"]
    qp_d, qt_d, qv_d = c4["qp"].to(dev), c4["qt"].to(dev), c4["qv"].to(dev)
    want = _np(eng.hybrid_topk("dense", "sparse", c4["q"], qp_d, qt_d, qv_d, k, max_terms=8))
    keys = eng.shard_local_lists("dense", "sparse", c4["q"], qp_d, qt_d, qv_d, k, max_terms=8)
    s, i, c, (ts, ti, tc) = eng.shard_fuse(keys[None].contiguous(), k, stages=True)
    got = _np((s, i, c))
    assert np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2]) and np.array_equal(got[0], want[0])
    dense = _np(eng.dense_topk("dense", c4["q"], k))
    sparse = _np(eng.sparse_topk("sparse", qp_d, qt_d, qv_d, k, max_terms=8))
    ts, ti, tc = _np((ts, ti, tc))
    for l, ref in enumerate((dense, sparse)):
        assert np.array_equal(ti[l], ref[1]) and np.array_equal(tc[l], ref[2])
        assert np.array_equal(ts[l].view(np.uint32), ref[0].view(np.uint32))
    torch.cuda.synchronize()


# ------------------------------------------------------------------------------------------ C5 (one of 8 shards)
def test_c5_shard_with_10pct_filter():
    """configs[4], the shard one GPU of eight holds: 12.5 M x 384 float16 + 12.5 M documents, Bernoulli(0.10) payload
    filter (seed 4242), ids offset to the shard's place in the 100 M-row corpus; batch 1 and a 256-query batch, before and
    after the filter-compacted copies exist."""
    import torch
    from rag_application_b200 import Engine
    dev = torch.device("cuda", 0)
    eng = Engine(0)
    try:
        n, d, k, B = 12_500_000, 384, 100, 256
        off = 3 * n                                       # shard 3 of 8
        eng.set_id_offset(off)
        x = synth.dense_corpus_torch(n, d, torch.float16, dev, seed=synth.SEED_CORPUS + 3)
        q = synth.dense_queries_torch(x, B)
        eng.dense_set("dense", x)
        g = torch.Generator(device=dev)
        g.manual_seed(synth.SEED_MASK)
        mask = torch.rand(n, generator=g, device=dev) < 0.10
        eng.mask_set(1, mask.cpu().numpy())
        rows = _sample(B)
        for mid, m in ((1, mask), (0, None)):
            got = _np(eng.dense_topk("dense", q, k, mask_id=mid))
            _check_dense(x, q, rows, got, k, mask=m, id_offset=off)
            one = _np(eng.dense_topk("dense", q[:1].contiguous(), k, mask_id=mid))
            assert np.array_equal(one[1][0], got[1][0])
        if mask is not None:
            eng.force_dense_path("tcgen05_full")          # masked scan of the full space == scan of the compacted copy
            full = _np(eng.dense_topk("dense", q, k, mask_id=1))
            eng.force_dense_path(None)
            comp = _np(eng.dense_topk("dense", q, k, mask_id=1))
            assert np.array_equal(full[1], comp[1]) and np.array_equal(full[0].view(np.uint32), comp[0].view(np.uint32))
        del x
        eng.dense_drop("dense")
        torch.cuda.empty_cache()
        tp, doc, w = synth.zipf_corpus_torch(n, 50_000, 285, dev, seed=synth.SEED_CORPUS + 3)
        qp, qt, qv = synth.sparse_queries_torch(tp, B, 8)
        eng.sparse_set_csr("sparse", tp, doc, w, n)
        qp_d, qt_d, qv_d = qp.to(dev), qt.to(dev), qv.to(dev)
        qpn, qtn, qvn = qp.numpy(), qt.numpy(), qv.numpy()
        eng.sparse_compact_after(-1)                      # masked scan of the full index
        got = _np(eng.sparse_topk("sparse", qp_d, qt_d, qv_d, k, mask_id=1, max_terms=8))
        _check_sparse(tp, doc, w, n, qpn, qtn, qvn, rows, got, k, mask=mask, id_offset=off)
        eng.sparse_compact_after(0)                       # index of the passing documents, built on first use
        comp = _np(eng.sparse_topk("sparse", qp_d, qt_d, qv_d, k, mask_id=1, max_terms=8))
        assert eng.sparse_compact_info("sparse")[0] == 1
        assert np.array_equal(comp[1], got[1]) and np.array_equal(comp[0].view(np.uint32), got[0].view(np.uint32))
        nomask = _np(eng.sparse_topk("sparse", qp_d, qt_d, qv_d, k, max_terms=8))
        _check_sparse(tp, doc, w, n, qpn, qtn, qvn, rows[:4], nomask, k, id_offset=off)
    finally:
        eng.close()
