"""the tensor code of the reference's rerank (huggingface.py:183-189) -- oracle pinned against a
float64 restatement, the host mirror of `rerank_with_colbert` (qdrant_handler.py:388-412) on the CPU mirror engine, and
the CUDA kernel against the oracle."""
import numpy as np
import pytest

from oracle import oracle as O
from rag_application_b200.rerank import B200Reranker

from _oracle_engine import OracleEngine


def _hidden(n, T, d, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.standard_normal((n, T, d)).astype(np.float32)


def test_oracle_rerank_matches_float64_and_pins_ties():
    q, d = _hidden(1, 9, 48, 1)[0], _hidden(20, 13, 48, 2)
    d[7] = d[3]
    d[15] = d[3]
    order, scores = O.rerank_meanpool(q, d)
    ref = d.astype(np.float64).mean(1) @ q.astype(np.float64).mean(0)
    assert np.allclose(scores, ref, rtol=1e-5, atol=1e-6)
    assert sorted(order.tolist()) == list(range(20))
    assert all(scores[order[i]] >= scores[order[i + 1]] for i in range(19))
    pos = {int(r): p for p, r in enumerate(order)}
    assert pos[15] < pos[7] < pos[3] and pos[3] - pos[15] == 2      # equal scores: higher index first


def test_rerank_with_colbert_host_mirror():
    texts = {f"doc {i}": i for i in range(6)}
    table = _hidden(7, 5, 32, 3)

    def encode(batch):
        return np.stack([table[texts.get(t, 6)] for t in batch])

    rr = B200Reranker(OracleEngine(0), encode)
    docs = list(texts)
    results = [{"id": i} for i in range(6)]
    order = rr.rerank_documents("what", docs, 100)
    want, _ = O.rerank_meanpool(table[6], table[:6])
    assert order == want.tolist()
    assert rr.rerank_with_colbert("what", docs, results, 100) == [results[i] for i in order]
    assert rr.rerank_with_colbert("what", [], results, 100) is results          # empty order: unchanged
    bad = B200Reranker(OracleEngine(0), lambda batch: (_ for _ in ()).throw(RuntimeError("encoder down")))
    assert bad.rerank_with_colbert("what", docs, results, 100) is results       # failure: unchanged


@pytest.mark.gpu
@pytest.mark.parametrize("n,T,Tq,d", [(1, 4, 3, 64), (10, 37, 9, 768), (257, 16, 5, 1024), (4096, 3, 2, 96)])
def test_rerank_kernel_matches_the_oracle(engine, n, T, Tq, d):
    q, h = _hidden(1, Tq, d, 11)[0], _hidden(n, T, d, 12)
    if n >= 10:
        h[n // 2] = h[1]                                  # an exact tie
    order, scores = engine.rerank_meanpool(q, h)
    eo, es = O.rerank_meanpool(q, h)
    tol = 1e-5 * np.abs(es).max() + 1e-6                  # float32 pooling, different summation order
    assert np.abs(scores - es).max() <= tol
    assert sorted(order.tolist()) == list(range(n))
    assert all(scores[order[i]] >= scores[order[i + 1]] for i in range(n - 1))
    if n >= 10:
        pos = {int(r): p for p, r in enumerate(order)}
        assert scores[n // 2] == scores[1] and pos[n // 2] + 1 == pos[1]
    # where the oracle's neighbours are further apart than the tolerance the two orders agree
    gap = np.abs(np.diff(es[eo]))
    firm = np.concatenate([[True], gap > 2 * tol]) & np.concatenate([gap > 2 * tol, [True]])
    assert np.array_equal(order[firm], eo[firm])


@pytest.mark.gpu
def test_rerank_kernel_device_pointers(engine):
    import torch
    q, h = _hidden(1, 7, 256, 21)[0], _hidden(33, 12, 256, 22)
    o_h, s_h = engine.rerank_meanpool(q, h)
    o_d, s_d = engine.rerank_meanpool(torch.from_numpy(q).cuda(), torch.from_numpy(h).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(o_h, o_d.cpu().numpy()) and np.array_equal(s_h, s_d.cpu().numpy())
    with pytest.raises(Exception):
        engine.rerank_meanpool(q, np.zeros((5000, 2, 256), np.float32))
