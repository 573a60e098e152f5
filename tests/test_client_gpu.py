"""GPU: the reference's own request sequence (create_collection -> upsert -> hybrid_search's nested
prefetch tree -> count) through B200QdrantClient on the real engine, against the known-answer vectors."""
import pytest

from _client_cases import run_reference_flow

pytestmark = pytest.mark.gpu


def test_reference_flow_on_b200():
    run_reference_flow(engine_factory=None)


def test_reference_flow_with_the_quantized_space_stored_as_int8_on_b200():
    """Q1 of the reference's tree runs on int8 rows through the kind::i8 scan."""
    run_reference_flow(engine_factory=None, storage_overrides={"quantized": "i8"})


def test_limits_above_the_engine_list_length_on_b200():
    from _client_cases import run_long_limits
    run_long_limits()


@pytest.mark.parametrize("storage", [None, "bf16"])
def test_incremental_upsert_on_b200(storage):
    from _client_cases import run_incremental_upsert
    run_incremental_upsert(engine_factory=None, storage=storage)


def test_dense_append_update_equal_a_fresh_upload(engine):
    """C ABI level (f2): set(a) + append(b) + update(rows) answers exactly like set(final)."""
    import numpy as np
    from rag_application_b200 import synth
    x = synth.dense_corpus(7000, 128, seed=3)
    q = synth.dense_queries(x, 70, seed=4)
    for dtype in ("f32", "bf16"):
        engine.dense_set("inc", x[:3000], dtype)
        engine.dense_topk("inc", q, 10)
        engine.dense_append("inc", x[3000:3001])                 # one row: forces a capacity growth
        engine.dense_append("inc", x[3001:])
        rows = np.array([0, 17, 6999])
        x2 = x.copy()
        x2[rows] = -x[rows]
        engine.dense_update("inc", rows, x2[rows])
        got = engine.dense_topk("inc", q, 50)
        engine.dense_set("ref", x2, dtype)
        want = engine.dense_topk("ref", q, 50)
        for a, b in zip(got, want):
            assert np.array_equal(a, b)


def test_save_load_on_b200(tmp_path):
    from _client_cases import run_save_load
    run_save_load(tmp_path / "bf16", storage="bf16")
    run_save_load(tmp_path / "f32")
