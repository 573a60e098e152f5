"""CPU: the Qdrant-shaped client (plan executor, filter compiler, id/payload mapping, lazy upload)
on the oracle-backed stand-in engine, against the committed known-answer vectors."""
import numpy as np
import pytest

from _oracle_engine import OracleEngine
from _client_cases import run_incremental_upsert, run_long_limits, run_reference_flow
from rag_application_b200 import models as M
from rag_application_b200.client import B200QdrantCore, compile_filter


def test_reference_flow_on_oracle_engine():
    run_reference_flow(engine_factory=OracleEngine)


def test_reference_flow_with_the_quantized_space_stored_as_int8():
    """storage_overrides={"quantized": "i8"}: the reference's int8 vector kept as int8 rows; same golden
    answers, and what cannot be held as int8 is refused."""
    run_reference_flow(engine_factory=OracleEngine, storage_overrides={"quantized": "i8"})
    core = B200QdrantCore(engine_factory=OracleEngine, storage_overrides={"q8": "i8", "bad": "i8"})
    with pytest.raises(ValueError):
        core.create_collection("c", vectors_config={"bad": M.VectorParams(size=96, distance=M.Distance.COSINE)})
    with pytest.raises(ValueError):
        core.create_collection("c", vectors_config={"q8": M.VectorParams(size=128, distance=M.Distance.DOT)})
    core.create_collection("c", vectors_config={"q8": M.VectorParams(size=128, distance=M.Distance.COSINE)})
    core.upsert("c", points=[M.PointStruct(id=1, vector={"q8": [0.5] * 128}, payload={})])
    with pytest.raises(ValueError):
        core.query_points("c", query=[1.0] * 128, using="q8", limit=1)          # rows are not integer-valued
    core.upsert("c", points=[M.PointStruct(id=1, vector={"q8": [3.0] * 128}, payload={}),
                             M.PointStruct(id=2, vector={"q8": [-3.0] * 128}, payload={})])
    with pytest.raises(ValueError):
        core.query_points("c", query=[0.25] * 128, using="q8", limit=1)         # nor may the query be fractional
    res = core.query_points("c", query=[7.0] * 128, using="q8", limit=2)
    assert [(p.id, p.score) for p in res.points] == [(1, 1.0), (2, -1.0)]


def test_incremental_upsert_on_oracle_engine():
    run_incremental_upsert(engine_factory=OracleEngine)


def test_filter_compiler():
    payloads = [{"a": 1, "t": ["x", "y"], "n": {"k": 5}}, {"a": 2, "t": "x"}, None, {"a": 1, "n": {"k": 9}}]
    ids = ["p0", "p1", "p2", "p3"]
    f = M.Filter(must=[M.FieldCondition(key="a", match=M.MatchValue(value=1))])
    assert compile_filter(f, payloads, ids).tolist() == [True, False, False, True]
    f = M.Filter(must_not=[{"key": "t", "match": {"value": "x"}}])
    assert compile_filter(f, payloads, ids).tolist() == [False, False, True, True]
    f = M.Filter(should=[M.FieldCondition(key="n.k", range=M.Range(gte=6)), M.HasIdCondition(has_id=["p1"])])
    assert compile_filter(f, payloads, ids).tolist() == [False, True, False, True]
    f = M.Filter(must=[M.Filter(should=[{"key": "a", "match": {"any": [2, 3]}}]), {"key": "t", "match": {"value": "x"}}])
    assert compile_filter(f, payloads, ids).tolist() == [False, True, False, False]
    assert compile_filter(None, payloads, ids).all()
    with pytest.raises(ValueError):
        compile_filter(M.Filter(must=[{"nonsense": 1}]), payloads, ids)


def test_errors_and_edge_cases():
    core = B200QdrantCore(engine_factory=OracleEngine)
    core.create_collection("c", vectors_config={"v": M.VectorParams(size=4, distance=M.Distance.DOT)},
                           sparse_vectors_config={"s": M.SparseVectorParams(modifier=M.Modifier.IDF)})
    with pytest.raises(ValueError):
        core.create_collection("c", vectors_config={})
    with pytest.raises(ValueError):
        core.query_points("nope", query=[0, 0, 0, 0], using="v")
    # empty collection answers with no points
    assert core.query_points("c", query=[1, 0, 0, 0], using="v", limit=3).points == []
    with pytest.raises(ValueError):
        core.upsert("c", [M.PointStruct(id=1, vector={"v": [1, 2, 3]})])          # wrong dim
    with pytest.raises(ValueError):
        core.upsert("c", [M.PointStruct(id=1, vector={"zzz": [1, 2, 3, 4]})])     # unknown vector name
    with pytest.raises(ValueError):
        core.upsert("c", [M.PointStruct(id=1, vector={"s": {"indices": [3, 3], "values": [1.0, 2.0]}})])
    core.upsert("c", [M.PointStruct(id=7, vector={"v": [1, 0, 0, 0], "s": M.SparseVector([10, 2**31 - 5], [1.0, 2.0])}, payload={"k": 1}),
                      M.PointStruct(id=8, vector={"v": [0, 2, 0, 0]}, payload={"k": 2}),
                      M.PointStruct(id="u-9", vector={"s": {"indices": [10], "values": [3.0]}})])
    r = core.query_points("c", query=[1, 1, 0, 0], using="v", limit=10).points
    assert [(p.id, p.score) for p in r] == [(8, 2.0), (7, 1.0)]                    # point u-9 has no dense vector
    r = core.query_points("c", query=M.SparseVector([10, 999], [1.0, 1.0]), using="s", limit=10).points
    # IDF: N = 2 points carry the sparse vector, df(10) = 2 -> ln(1 + 0.5/2.5)
    import math
    w = math.log(1 + 0.5 / 2.5)
    assert [p.id for p in r] == ["u-9", 7]
    np.testing.assert_allclose([p.score for p in r], [3.0 * w, 1.0 * w], rtol=1e-6)
    assert core.query_points("c", query=M.SparseVector([999], [1.0]), using="s").points == []   # unknown term only
    # offset / score_threshold / with_payload
    r = core.query_points("c", query=[1, 1, 0, 0], using="v", limit=1, offset=1, with_payload=False).points
    assert [(p.id, p.payload) for p in r] == [(7, None)]
    assert core.query_points("c", query=[1, 1, 0, 0], using="v", score_threshold=1.5).points[0].id == 8
    assert core.count("c").count == 3 and core.get_collection("c").points_count == 3
    with pytest.raises(NotImplementedError):
        core.query_points("c", query=M.FusionQuery(fusion=M.Fusion.DBSF), prefetch=[M.Prefetch(query=[1, 0, 0, 0], using="v")])
    assert core.delete_collection("c") and not core.delete_collection("c")


def test_limits_above_the_engine_list_length_on_oracle_engine():
    run_long_limits(engine_factory=OracleEngine)


def test_unsupported_filter_conditions_are_refused():
    """A condition the compiler does not evaluate raises instead of matching everything."""
    import pytest
    from rag_application_b200.client import compile_filter
    payloads = [{"a": 1, "t": "x y"}, {"a": 2}, None]
    ids = [0, 1, 2]
    for bad in ({"must": [{"key": "t", "match": {"text": "x"}}]},
                {"must": [{"key": "a", "match": {"except": [1]}}]},
                {"must": [{"key": "a", "values_count": {"gt": 1}}]},
                {"must_not": [{"key": "a", "is_empty": True}]},
                {"should": [{"key": "a"}]},
                {"must": [{"key": "a", "match": {}}]}):
        with pytest.raises(NotImplementedError):
            compile_filter(bad, payloads, ids)
    assert compile_filter({"must": [{"key": "a", "match": {"value": 1}}]}, payloads, ids).tolist() == [True, False, False]


def test_upsert_replaces_the_whole_point_and_batch_skips_points_without_the_dense_vector():
    import asyncio
    import numpy as np
    from _oracle_engine import OracleEngine
    from rag_application_b200 import models as M
    from rag_application_b200.client import B200QdrantClient

    async def flow():
        cl = B200QdrantClient(engine_factory=OracleEngine)
        await cl.create_collection(collection_name="c", vectors_config={"dense": M.VectorParams(size=4, distance=M.Distance.DOT)},
                                   sparse_vectors_config={"sparse": M.SparseVectorParams()})
        rng = np.random.default_rng(3)
        pts = [M.PointStruct(id=i, vector={"dense": rng.standard_normal(4).tolist(), "sparse": {"indices": [1, 2 + i % 3], "values": [1.0, 0.5]}},
                             payload={"i": i}) for i in range(12)]
        await cl.upsert(collection_name="c", points=pts)
        q = [1.0, 0.0, 0.0, 0.0]
        before = await cl.query_points(collection_name="c", query=q, using="dense", limit=12)
        assert len(before.points) == 12
        # overwrite point 5 with a sparse-only point: its dense vector is gone, not kept
        await cl.upsert(collection_name="c", points=[M.PointStruct(id=5, vector={"sparse": {"indices": [1], "values": [9.0]}}, payload={"i": 5})])
        after = await cl.query_points(collection_name="c", query=q, using="dense", limit=12)
        assert 5 not in [p.id for p in after.points] and len(after.points) == 11
        sv = {"indices": [1], "values": [1.0]}
        batch = await cl.hybrid_search_batch("c", np.array([q], np.float32), [sv], limit=12)
        one = await cl.query_points(collection_name="c", query=M.FusionQuery(fusion=M.Fusion.RRF), limit=12,
                                    prefetch=[M.Prefetch(query=q, using="dense", limit=12),
                                              M.Prefetch(query=M.SparseVector(**sv), using="sparse", limit=12)])
        assert [p.id for p in batch[0]] == [p.id for p in one.points]
        assert 5 in [p.id for p in batch[0]]            # still a sparse hit
    asyncio.run(flow())


def test_client_flows_on_a_row_sharded_collection():
    """B200QdrantClient(devices=[...]): the same reference flows, every collection row-sharded over three engines
    (multi.py, blocks of 16 rows so the 200-point corpus really spreads) -- identical answers."""
    from _client_cases import run_incremental_upsert, run_long_limits, run_reference_flow
    from _oracle_engine import OracleEngine
    kw = dict(engine_factory=OracleEngine, devices=[0, 1, 2], shard_block_rows=16)
    run_reference_flow(**kw)
    run_incremental_upsert(**kw)
    run_long_limits(**kw)


def test_multi_engine_row_mapping():
    from _oracle_engine import OracleEngine
    from rag_application_b200.multi import MultiEngine
    import numpy as np
    m = MultiEngine([0, 1, 2], OracleEngine, block_rows=8)
    rows = np.arange(1000)
    sh, loc = m._shard_of(rows), m._local_of(rows)
    for g in range(3):
        mine = rows[sh == g]
        assert np.array_equal(m._global_of(g, loc[sh == g]), mine)
        assert np.array_equal(loc[sh == g], np.arange(len(mine)))          # local rows are dense and keep the order
        for n in (0, 5, 8, 23, 24, 25, 1000):
            assert m._local_n(g, n) == int((mine < n).sum())
    m.close()


def test_save_load_on_oracle_engine(tmp_path):
    from _client_cases import run_save_load
    from _oracle_engine import OracleEngine
    run_save_load(tmp_path / "one", engine_factory=OracleEngine)
    run_save_load(tmp_path / "three", engine_factory=OracleEngine, devices=[0, 1, 2], shard_block_rows=16)
