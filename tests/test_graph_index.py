"""db.index.vector.queryNodes on the dense scan (graph_index.py) -- CPU mirror engine here, the CUDA
engine under -m gpu, both against a brute-force numpy ranking."""
import numpy as np
import pytest

from _oracle_engine import OracleEngine
from rag_application_b200.graph_index import B200GraphVectorIndexes


def _nodes(n, d, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    x = rng.standard_normal((n, d)).astype(np.float32)
    props = [{"id": f"n{i}", "user_id": "u1" if i % 3 else "u2", "type": "person" if i % 2 else "place"} for i in range(n)]
    return x, props


def _check(engine, d, storage):
    g = B200GraphVectorIndexes(engine, storage=storage)
    g.create_vector_index("entity_embedding_index", d)
    g.create_vector_index("relation_embedding_index", d)
    x, props = _nodes(700, d, 5)
    y, rprops = _nodes(300, d, 6)
    rprops = [{**p, "id": "r" + p["id"]} for p in rprops]
    g.set_nodes("entity_embedding_index", x, props)
    g.set_nodes("relation_embedding_index", y, rprops)
    q = x[17] + 0.2 * y[3]
    cos = (x @ q) / (np.linalg.norm(x, axis=1) * np.linalg.norm(q))
    top = np.argsort(-cos, kind="stable")[:25]
    got = g.query_nodes("entity_embedding_index", 25, q)
    tol = 1e-2 if storage != "f32" else 1e-5
    ids = [int(p["id"][1:]) for p, _ in got]
    if storage == "f32":
        assert ids == top.tolist()
    else:       # rows rounded to 16 bits: near-equal neighbours may swap, and the last places may differ
        assert len(ids) == 25 and len(set(ids) & set(top.tolist())) >= 23
        assert all(cos[a] >= cos[b] - tol for a, b in zip(ids, ids[1:]))
    assert np.allclose([s for _, s in got], (1 + cos[ids]) / 2, atol=tol)
    # the filter runs on the k rows the index returned (fewer than k may survive)
    rows = g.search("entity_embedding_index", 25, q, {"user_id": "u1", "type": None})
    assert [r["id"] for r in rows] == [f"n{i}" for i in ids if i % 3]
    assert all(a["score"] >= b["score"] for a, b in zip(rows, rows[1:]))
    merged = g.search_across_spaces(["entity_embedding_index", "relation_embedding_index"], q, 10, {"user_id": "u1"})
    assert len(merged) == 10 and all(a["score"] >= b["score"] for a, b in zip(merged, merged[1:]))
    assert {r["id"][0] for r in merged} <= {"n", "r"}
    with pytest.raises(ValueError):
        g.query_nodes("entity_embedding_index", 5, q[:-1])
    g.set_nodes("relation_embedding_index", np.zeros((0, d), np.float32), [])
    assert g.query_nodes("relation_embedding_index", 5, q) == []


def test_query_nodes_on_the_cpu_mirror():
    _check(OracleEngine(0), 256, "f32")


@pytest.mark.gpu
@pytest.mark.parametrize("d,storage", [(256, "f32"), (256, "bf16"), (2048, "f16")])
def test_query_nodes_on_the_gpu(engine, d, storage):
    _check(engine, d, storage)
