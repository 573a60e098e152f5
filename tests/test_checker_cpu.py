"""CPU: the torch checker used at the BASELINE config sizes (tools/torch_checker.py, shares no code with libbrag.so or
oracle/) agrees with the oracle on small seeded inputs -- so "GPU == checker" at 10 M rows means "GPU == oracle"."""
import numpy as np
import torch

from oracle import oracle as O
from rag_application_b200 import synth
from tools import torch_checker as T


def test_checker_dense_matches_oracle():
    n, d, k = 3000, 96, 40
    x = synth.dense_corpus(n, d, seed=5, clustered=True)
    q = synth.dense_queries(x, 4, seed=6)
    mask = synth.random_mask(n, 0.3)
    for dtype, tdt in (("f32", torch.float32), ("f16", torch.float16), ("bf16", torch.bfloat16)):
        xs = O.round_storage(x, dtype)
        X = torch.from_numpy(xs).to(tdt)
        assert np.array_equal(X.float().numpy(), xs)
        for b in range(4):
            for m in (None, mask):
                es, ei = O.dense_topk(xs, q[b], k, mask=m)
                s, i, _ = T.dense_topk(X, torch.from_numpy(q[b]), k, mask=None if m is None else torch.from_numpy(m), id_offset=7)
                assert np.array_equal(i - 7, ei) and np.array_equal(s.view(np.uint32), es.view(np.uint32))


def test_checker_sparse_and_rrf_match_oracle():
    n = 4000
    ptr, doc, w, _ = synth.zipf_corpus(n, 700, 40, seed=7)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n, 6, 8, "zipf", seed=8, skip_top=4)
    mask = synth.random_mask(n, 0.4)
    csr = O.CSR(ptr, doc, w, n)
    tdoc, tw = torch.from_numpy(doc.astype(np.int32)), torch.from_numpy(w)
    lists = []
    for b in range(6):
        t, v = qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]]
        for m in (None, mask):
            es, ei = O.sparse_topk(csr, t, v, 30, mask=m)
            s, i = T.sparse_topk(ptr, tdoc, tw, n, t, v, 30, mask=None if m is None else torch.from_numpy(m))
            assert np.array_equal(i, ei) and np.array_equal(s.view(np.uint32), es.view(np.uint32))
        lists.append(ei)
    for a, b in ((0, 1), (2, 3), (4, 5)):
        es, ei = O.rrf([lists[a], lists[b]], limit=25)
        s, i = T.rrf([lists[a], lists[b]], 25)
        assert np.array_equal(i, ei) and np.array_equal(s, es)
