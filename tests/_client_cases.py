"""Shared bodies of the client tests: the same requests the reference's QdrantHandler issues
(qdrant_handler.py:57-100 create_collection, :155-193 upsert, :305-372 hybrid_search), run against a
B200QdrantClient built either on the real engine (GPU tests) or on the oracle stand-in (CPU tests),
and compared with the committed known-answer vectors."""
import asyncio
import json
import os

import numpy as np

from rag_application_b200 import models as M
from rag_application_b200.client import B200QdrantClient

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kat.json")


def load_tree_case():
    with open(GOLD) as f:
        c = json.load(f)["tree"]
    rng = np.random.Generator(np.random.PCG64(c["seed"]))
    n, d = c["n"], c["d"]
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = (x[17] + 0.2 * rng.standard_normal(d)).astype(np.float32)
    q /= np.linalg.norm(q)
    return c, x, q


async def reference_create_collection(client, user_id, dense=768, sizes=(64, 128, 256)):
    """qdrant_handler.py:57-100."""
    vectors_config = {
        "dense": M.VectorParams(size=dense, distance=M.Distance.COSINE, on_disk=True),
        "quantized": M.VectorParams(size=dense, distance=M.Distance.COSINE, on_disk=True),
    }
    for dim in sizes:
        vectors_config[f"matryoshka_{dim}"] = M.VectorParams(size=dim, distance=M.Distance.COSINE, on_disk=False)
    await client.create_collection(
        collection_name=user_id, vectors_config=vectors_config,
        sparse_vectors_config={"sparse": M.SparseVectorParams(index=M.SparseIndexParams(on_disk=False))},
        optimizers_config=M.OptimizersConfigDiff(memmap_threshold=20000), on_disk_payload=True)


def reference_points(x, docs, id_of=lambda i: f"chunk-{i:04d}"):
    """qdrant_handler.py:134-187: derived vectors exactly as the reference computes them."""
    pts = []
    for i, row in enumerate(x):
        emb = row.tolist()
        quant = np.clip((np.array(emb) * 127).astype(np.int8), -128, 127).tolist()
        e = np.array(emb)
        pts.append(M.PointStruct(
            id=id_of(i),
            vector={"dense": emb, "matryoshka_64": e[:64].tolist(), "matryoshka_128": e[:128].tolist(),
                    "matryoshka_256": e[:256].tolist(), "quantized": quant,
                    "sparse": {"indices": docs[i][0], "values": docs[i][1]}},
            payload={"document_id": f"doc-{i % 7}", "content": f"text {i}", "page_number": i % 5, "row": i}))
    return pts


async def reference_hybrid_search(client, user_id, dense_vector, sparse_vector, sp, query_filter=None):
    """qdrant_handler.py:296-372, request construction only (no reranker)."""
    quantized_query = np.clip((np.array(dense_vector) * 127).astype(np.int8), -128, 127).tolist()
    matryoshka_prefetch = M.Prefetch(
        prefetch=[M.Prefetch(
            prefetch=[M.Prefetch(
                prefetch=[M.Prefetch(query=dense_vector[:64], using="matryoshka_64", limit=sp["matryoshka_64_limit"])],
                query=dense_vector[:128], using="matryoshka_128", limit=sp["matryoshka_128_limit"])],
            query=dense_vector[:256], using="matryoshka_256", limit=sp["matryoshka_256_limit"])],
        query=dense_vector, using="dense", limit=sp["dense_limit"])
    quantized_dense_prefetch = M.Prefetch(
        prefetch=[M.Prefetch(query=quantized_query, using="quantized", limit=sp["quantized_limit"])],
        query=dense_vector, using="dense", limit=sp["dense_limit"])
    sparse_search_prefetch = M.Prefetch(
        query=M.SparseVector(indices=sparse_vector["indices"], values=sparse_vector["values"]),
        using="sparse", limit=sp["sparse_limit"])
    sparse_dense_rrf = M.Prefetch(prefetch=[quantized_dense_prefetch, sparse_search_prefetch],
                                  query=M.models.FusionQuery(fusion=M.models.Fusion.RRF))
    res = await client.query_points(
        collection_name=str(user_id), prefetch=[matryoshka_prefetch, sparse_dense_rrf], query=dense_vector,
        using="dense", limit=sp["final_limit"], search_params={"hnsw_ef": 128, "exact": True},
        with_payload=True, query_filter=query_filter)
    return res.points


def run_reference_flow(engine_factory=None, rtol=1e-5, **client_kw):
    c, x, q = load_tree_case()

    async def flow():
        client = B200QdrantClient(engine_factory=engine_factory, **client_kw)
        user = "user-42"
        assert [col.name for col in (await client.get_collections()).collections] == []
        await reference_create_collection(client, user)
        assert user in [col.name for col in (await client.get_collections()).collections]
        pts = reference_points(x, c["docs"])
        await client.upsert(collection_name=user, points=pts[:120])
        await client.upsert(collection_name=user, points=pts[120:])
        assert (await client.get_collection(collection_name=user)).points_count == c["n"]
        sp = dict(c["search_params"], hnsw_ef=128)
        sv = {"indices": c["q_idx"], "values": c["q_val"]}
        got = await reference_hybrid_search(client, user, q.tolist(), sv, sp)
        exp = c["final"]
        assert [p.payload["row"] for p in got] == [r for _, r in exp]
        assert [p.id for p in got] == [f"chunk-{r:04d}" for _, r in exp]
        np.testing.assert_allclose([p.score for p in got], [s for s, _ in exp], rtol=rtol, atol=1e-6)
        assert got[0].payload["content"] == f"text {exp[0][1]}"
        # root payload filter: a Qdrant server (what the reference talks to) merges it into every prefetch -- the
        # client's default; filter_propagation="local" keeps qdrant-client local mode (top level only)
        mask = np.array(c["mask"], bool)
        keep_rows = [int(i) for i in np.nonzero(mask)[0]]
        flt = M.Filter(must=[M.FieldCondition(key="row", match=M.MatchAny(any=keep_rows))])
        got = await reference_hybrid_search(client, user, q.tolist(), sv, sp, query_filter=flt)
        assert [p.payload["row"] for p in got] == [r for _, r in c["final_masked_server"]]
        client.core.filter_propagation = "local"
        got = await reference_hybrid_search(client, user, q.tolist(), sv, sp, query_filter=flt)
        assert [p.payload["row"] for p in got] == [r for _, r in c["final_masked"]]
        client.core.filter_propagation = "server"
        # the batched shape and the tree of the same shape agree under a filter too
        fb = await client.hybrid_search_batch(user, q[None, :], [sv], limit=10, query_filter=flt)
        fo = await client.query_points(
            collection_name=user, query=M.FusionQuery(fusion=M.Fusion.RRF), limit=10, query_filter=flt,
            prefetch=[M.Prefetch(query=q.tolist(), using="dense", limit=10),
                      M.Prefetch(query=M.SparseVector(indices=c["q_idx"], values=c["q_val"]), using="sparse", limit=10)])
        assert [p.id for p in fb[0]] == [p.id for p in fo.points] and all(mask[int(p.id[-4:])] for p in fb[0])
        assert (await client.count(collection_name=user, count_filter=flt)).count == int(mask.sum())
        # Filter(**dict) spelling the reference uses (qdrant_handler.py:297)
        flt2 = M.Filter(**{"must": [{"key": "page_number", "match": {"value": 3}}]})
        cnt = (await client.count(collection_name=user, count_filter=flt2)).count
        assert cnt == sum(1 for i in range(c["n"]) if i % 5 == 3)
        res = await client.query_points(collection_name=user, query=q.tolist(), using="dense", limit=5, query_filter=flt2)
        assert len(res.points) == 5 and all(p.payload["page_number"] == 3 for p in res.points)
        # sparse-only leaf == golden S1, default fusion limit is 10
        res = await client.query_points(collection_name=user, query=M.SparseVector(indices=c["q_idx"], values=c["q_val"]),
                                        using="sparse", limit=sp["sparse_limit"])
        s1 = c["stages"]["S1"]
        assert [p.payload["row"] for p in res.points] == [r for _, r in s1]
        assert [np.float32(p.score) for p in res.points] == [np.float32(s) for s, _ in s1]
        # batched three-stage shape agrees with the per-query tree of the same shape
        batch = await client.hybrid_search_batch(user, np.stack([q, x[5]]), [sv, {"indices": c["docs"][5][0], "values": c["docs"][5][1]}],
                                                 limit=10, with_payload=True)
        one = await client.query_points(
            collection_name=user, query=M.FusionQuery(fusion=M.Fusion.RRF), limit=10,
            prefetch=[M.Prefetch(query=q.tolist(), using="dense", limit=10),
                      M.Prefetch(query=M.SparseVector(indices=c["q_idx"], values=c["q_val"]), using="sparse", limit=10)])
        assert [p.id for p in batch[0]] == [p.id for p in one.points]
        assert [p.score for p in batch[0]] == [p.score for p in one.points]
        # upsert of an existing id overwrites in place; delete drops the collection
        await client.upsert(collection_name=user, points=[pts[17]])
        assert (await client.get_collection(collection_name=user)).points_count == c["n"]
        await client.delete_collection(collection_name=user)
        assert [col.name for col in (await client.get_collections()).collections] == []
        await client.close()

    asyncio.run(flow())


def run_incremental_upsert(engine_factory=None, storage=None, **client_kw):
    """points upserted after the first query are appended / overwritten in HBM
    (brag_dense_append / brag_dense_update / brag_sparse_append) and must answer exactly like a collection uploaded
    in one go; overwriting a stored sparse vector falls back to a rebuild of that space."""
    from rag_application_b200.client import B200QdrantCore
    rng = np.random.Generator(np.random.PCG64(99))
    n, d = 900, 64
    x = rng.standard_normal((n + 50, d)).astype(np.float32)
    sp = [(np.sort(rng.choice(500, 6, replace=False)), rng.random(6).astype(np.float32) + 0.1) for _ in range(n + 50)]

    def point(i, vec=None):
        return M.PointStruct(id=i, vector={"dense": (x[i] if vec is None else vec).tolist(),
                                           "sparse": M.SparseVector(indices=sp[i][0].tolist(), values=sp[i][1].tolist())},
                             payload={"i": int(i)})

    def make():
        kw = dict({"engine_factory": engine_factory} if engine_factory else {}, **client_kw)
        core = B200QdrantCore(storage=storage, **kw)
        core.create_collection("c", vectors_config={"dense": M.VectorParams(size=d, distance=M.Distance.COSINE)},
                               sparse_vectors_config={"sparse": M.SparseVectorParams()})
        return core

    def ask(core, q):
        r = core.query_points("c", prefetch=[M.Prefetch(query=q.tolist(), using="dense", limit=40),
                                             M.Prefetch(query=M.SparseVector(indices=sp[3][0].tolist(), values=sp[3][1].tolist()),
                                                        using="sparse", limit=40)],
                              query=M.FusionQuery(fusion=M.Fusion.RRF), limit=20).points
        return [(p.id, p.score) for p in r]

    q = x[5] + 0.05 * rng.standard_normal(d).astype(np.float32)
    inc = make()
    inc.upsert("c", [point(i) for i in range(n)])
    first = ask(inc, q)                                           # commits n rows
    calls = []
    eng = inc.engine
    for name in ("dense_set", "dense_append", "dense_update", "sparse_set_csr", "sparse_append"):
        orig = getattr(eng, name)
        setattr(eng, name, (lambda o, nm: (lambda *a, **k: (calls.append(nm), o(*a, **k))[1]))(orig, name))
    changed = -x[7]
    inc.upsert("c", [point(i) for i in range(n, n + 50)] + [point(7, changed)])
    second = ask(inc, q)
    assert "dense_append" in calls and "dense_update" in calls and "dense_set" not in calls
    assert "sparse_append" in calls and "sparse_set_csr" not in calls
    x2 = x.copy()
    x2[7] = changed
    full = make()
    full.upsert("c", [point(i, x2[i]) for i in range(n + 50)])
    assert ask(full, q) == second
    # new vocabulary rows in the middle of the term order, then an overwritten sparse vector (rebuild)
    del calls[:]
    sp.append((np.array([3, 250, 777, 900]), np.array([0.9, 0.8, 0.7, 0.6], np.float32)))
    x = np.concatenate([x, x[5:6] * 0.9 + 0.1 * x[6:7]])
    x2 = np.concatenate([x2, x[-1:]])
    inc.upsert("c", [point(n + 50)])
    third = ask(inc, q)
    assert "sparse_append" in calls and "sparse_set_csr" not in calls
    full.upsert("c", [point(n + 50, x2[n + 50])])
    assert ask(full, q) == third
    del calls[:]
    sp[11] = (sp[3][0].copy(), (sp[3][1] * 2).astype(np.float32))
    inc.upsert("c", [point(11, x2[11])])
    fourth = ask(inc, q)
    assert "sparse_set_csr" in calls and fourth != third
    full.upsert("c", [point(11, x2[11])])
    assert ask(full, q) == fourth
    inc.close()
    full.close()


def run_long_limits(engine_factory=None, storage=None, **client_kw):
    """Prefetch limits above the engine's list length (the reference lets an LLM choose them,
    hybrid_search_workflow.py:28-66) are served in pages: ids must come back in exactly the brute-force order."""
    from rag_application_b200.client import B200QdrantCore
    rng = np.random.Generator(np.random.PCG64(7))
    n, d, V = 1500, 32, 300
    x = rng.standard_normal((n, d)).astype(np.float32)
    sp = [(np.sort(rng.choice(V, 6, replace=False)), (rng.random(6) + 0.1).astype(np.float32)) for _ in range(n)]
    kw = dict({"engine_factory": engine_factory} if engine_factory else {}, **client_kw)
    core = B200QdrantCore(storage=storage, **kw)
    core.create_collection("c", vectors_config={"dense": M.VectorParams(size=d, distance=M.Distance.COSINE)},
                           sparse_vectors_config={"sparse": M.SparseVectorParams()})
    core.upsert("c", [M.PointStruct(id=i, vector={"dense": x[i].tolist(),
                                                  "sparse": M.SparseVector(indices=sp[i][0].tolist(), values=sp[i][1].tolist())},
                                    payload={"even": i % 2 == 0}) for i in range(n)])
    q = rng.standard_normal(d).astype(np.float32)
    xn = x / np.linalg.norm(x, axis=1, keepdims=True)
    cos = xn.astype(np.float64) @ (q / np.linalg.norm(q)).astype(np.float64)
    for flt, ok in ((None, np.ones(n, bool)),
                    (M.Filter(must=[M.FieldCondition(key="even", match=M.MatchValue(value=True))]), np.arange(n) % 2 == 0)):
        want = [int(i) for i in np.argsort(-cos, kind="stable") if ok[i]][:1300]
        got = core.query_points("c", query=q.tolist(), using="dense", limit=1300, query_filter=flt, with_payload=False).points
        assert len(got) == min(1300, int(ok.sum()))
        ids = [p.id for p in got]
        # float32 scores of near-equal neighbours may swap places: compare as sets per page and check monotone scores
        assert set(ids) == set(want) or len(set(ids) ^ set(want)) <= 4
        assert all(a.score >= b.score - 1e-6 for a, b in zip(got, got[1:]))
        assert len(set(ids)) == len(ids)
    # nested prefetch whose inner list is longer than one page, fused with a long sparse list (RRF over 2 x 700 entries)
    inner = M.Prefetch(query=q.tolist(), using="dense", limit=700)
    nested = core.query_points("c", prefetch=[inner], query=(-q).tolist(), using="dense", limit=650, with_payload=False).points
    top700 = [int(i) for i in np.argsort(-cos, kind="stable")][:700]
    assert len(nested) == 650 and set(p.id for p in nested) <= set(top700)
    assert all(a.score >= b.score - 1e-6 for a, b in zip(nested, nested[1:]))
    worst = sorted(top700, key=lambda i: cos[i])[:600]          # re-ranked by -q: the worst of the 700 come first
    assert len(set(p.id for p in nested[:600]) ^ set(worst)) <= 4
    core.ENGINE_MAX_CAND = 256                                   # force the sliced re-score of the 700 candidates
    sliced = core.query_points("c", prefetch=[inner], query=(-q).tolist(), using="dense", limit=650, with_payload=False).points
    del core.ENGINE_MAX_CAND
    assert [p.id for p in sliced] == [p.id for p in nested] and [p.score for p in sliced] == [p.score for p in nested]
    qi = np.sort(rng.choice(V, 80, replace=False))
    qv = (rng.random(80) + 0.1).astype(np.float32)
    fused = core.query_points("c", prefetch=[inner, M.Prefetch(query=M.SparseVector(indices=qi.tolist(), values=qv.tolist()),
                                                               using="sparse", limit=700)],
                              query=M.FusionQuery(fusion=M.Fusion.RRF), limit=900, with_payload=False).points
    assert len(fused) == 900 and len(set(p.id for p in fused)) == 900
    assert all(a.score >= b.score for a, b in zip(fused, fused[1:]))
    sc = np.zeros(n)
    hit = np.zeros(n, bool)
    lut = dict(zip(qi.tolist(), qv.tolist()))
    for i, (ti, tv) in enumerate(sp):
        for t, v in zip(ti.tolist(), tv.tolist()):
            if t in lut:
                sc[i] += float(np.float32(lut[t])) * float(v)
                hit[i] = True
    got = core.query_points("c", query=M.SparseVector(indices=qi.tolist(), values=qv.tolist()), using="sparse",
                            limit=1300, with_payload=False).points
    assert len(got) == min(1300, int(hit.sum())) and len(got) > 600
    ids = [p.id for p in got]
    assert len(set(ids)) == len(ids) and all(hit[i] for i in ids)
    assert all(a.score >= b.score for a, b in zip(got, got[1:]))
    assert np.allclose([p.score for p in got], sc[ids], rtol=1e-5)
    worst_kept = min(p.score for p in got)
    assert all(sc[i] <= worst_kept + 1e-6 for i in range(n) if hit[i] and i not in set(ids))
    core.close()


def run_save_load(tmpdir, engine_factory=None, storage=None, **client_kw):
    """SURVEY 8 f2 persistence: save -> a NEW client -> load answers exactly like the one that saved, and goes on taking
    incremental upserts (the loaded spaces are appended to in HBM, not re-uploaded)."""
    from rag_application_b200.client import B200QdrantCore
    rng = np.random.Generator(np.random.PCG64(123))
    n, d = 700, 64
    x = rng.standard_normal((n + 40, d)).astype(np.float32)
    sp = [(np.sort(rng.choice(400, 6, replace=False)), rng.random(6).astype(np.float32) + 0.1) for _ in range(n + 40)]
    kw = dict({"engine_factory": engine_factory} if engine_factory else {}, **client_kw)

    def point(i):
        return M.PointStruct(id=f"p{i}", vector={"dense": x[i].tolist(),
                                                 "sparse": M.SparseVector(indices=sp[i][0].tolist(), values=sp[i][1].tolist())},
                             payload={"i": int(i), "page_number": int(i % 4)})

    def ask(core, flt=None):
        r = core.query_points("c", prefetch=[M.Prefetch(query=x[9].tolist(), using="dense", limit=40),
                                             M.Prefetch(query=M.SparseVector(indices=sp[3][0].tolist(), values=sp[3][1].tolist()),
                                                        using="sparse", limit=40)],
                              query=M.FusionQuery(fusion=M.Fusion.RRF), limit=20, query_filter=flt).points
        return [(p.id, p.score, p.payload["i"]) for p in r]

    a = B200QdrantCore(storage=storage, **kw)
    a.create_collection("c", vectors_config={"dense": M.VectorParams(size=d, distance=M.Distance.COSINE)},
                        sparse_vectors_config={"sparse": M.SparseVectorParams(modifier=M.Modifier.IDF)})
    a.upsert("c", [point(i) for i in range(n)])
    flt = M.Filter(must=[M.FieldCondition(key="page_number", match=M.MatchValue(value=2))])
    want, want_f = ask(a), ask(a, flt)
    a.save(str(tmpdir))
    b = B200QdrantCore(storage=storage, **kw)
    calls = []
    for name in ("dense_set", "sparse_set_csr"):
        orig = getattr(b.engine, name)
        setattr(b.engine, name, (lambda o, nm: (lambda *aa, **k: (calls.append(nm), o(*aa, **k))[1]))(orig, name))
    b.load(str(tmpdir))
    assert [c.name for c in b.get_collections().collections] == ["c"] and b.get_collection("c").points_count == n
    assert ask(b) == want and ask(b, flt) == want_f
    assert calls == [], "load must not re-upload or rebuild a space"
    # both go on: the same upserts give the same answers (append path on the loaded spaces)
    for core in (a, b):
        core.upsert("c", [point(i) for i in range(n, n + 40)])
    assert ask(b) == ask(a) and ask(b, flt) == ask(a, flt)
    assert calls == []
    a.close()
    b.close()
