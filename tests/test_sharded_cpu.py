"""CPU, world_size 2 over gloo: the row-sharded hybrid path (rag_application_b200/sharded.py) with
the oracle-backed stand-in engine.  Every rank must return the unsharded oracle's fused list."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _inputs():
    from rag_application_b200 import synth
    n, d, B, k = 4001, 64, 5, 20
    x = synth.dense_corpus(n, d, seed=11, clustered=True)
    q = synth.dense_queries(x, B, seed=12)
    ptr, doc, w, _ = synth.zipf_corpus(n, 600, 30, seed=13)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n, B, 6, "zipf", seed=14, skip_top=4)
    return n, d, B, k, x, q, ptr, doc, w, qp, qt, qv


def _shard_csr(ptr, doc, w, lo, hi, V):
    term_of = np.repeat(np.arange(V), np.diff(ptr))
    keep = (doc >= lo) & (doc < hi)
    p = np.zeros(V + 1, np.int64)
    np.add.at(p, term_of[keep] + 1, 1)
    return np.cumsum(p), (doc[keep] - lo).astype(np.uint32), w[keep]


def _worker(rank, world, port, q_out):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "tests"))
    import torch
    import torch.distributed as dist
    from _oracle_engine import OracleEngine
    from rag_application_b200.sharded import ShardedHybrid, shard_bounds
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n, d, B, k, x, q, ptr, doc, w, qp, qt, qv = _inputs()
        lo, hi = shard_bounds(n, world, rank)
        eng = OracleEngine()
        eng.dense_set("dense", x[lo:hi], "bf16")
        sp, sd, sw = _shard_csr(ptr, doc, w, lo, hi, 600)
        eng.sparse_set_csr("sparse", sp, sd, sw, hi - lo)
        sh = ShardedHybrid(eng, rank, world, lo)
        s, i, c = sh.hybrid_topk("dense", "sparse", torch.from_numpy(q), torch.from_numpy(qp), torch.from_numpy(qt),
                                 torch.from_numpy(qv), k)
        q_out.put((rank, s.numpy().copy(), i.numpy().copy(), c.numpy().copy()))
    finally:
        dist.destroy_process_group()


def test_shard_bounds_cover():
    from rag_application_b200.sharded import shard_bounds
    for n in (1, 7, 4001, 10_000_000):
        for G in (1, 2, 3, 8):
            b = [shard_bounds(n, G, r) for r in range(G)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[r][1] == b[r + 1][0] for r in range(G - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_key_format_roundtrip():
    """keys_to_lists / lists_to_keys mirror csrc/common.cuh make_key: order-preserving, ids in the low word."""
    from rag_application_b200.sharded import keys_to_lists, lists_to_keys
    rng = np.random.default_rng(0)
    s = rng.standard_normal((3, 7)).astype(np.float32)
    s[0, 0], s[0, 1], s[1, 2] = 0.0, -0.0, np.float32(-1e-30)
    i = rng.integers(0, 1 << 32, (3, 7)).astype(np.int64)
    i[2, 3:] = -1
    k = lists_to_keys(s, i)
    s2, i2 = keys_to_lists(k)
    assert np.array_equal(i2, i)
    assert np.array_equal(s2[i >= 0], (s + np.float32(0))[i >= 0]) and np.all(np.isneginf(s2[i < 0]))
    # a bigger key is a better entry under (score desc, id asc)
    flat_s, flat_i, flat_k = s[i >= 0], i[i >= 0], k.view(np.uint64)[i >= 0]
    order = np.lexsort((flat_i, -flat_s.astype(np.float64)))
    assert np.array_equal(np.sort(flat_k)[::-1], flat_k[order])


@pytest.mark.timeout(180)
def test_sharded_hybrid_world2_gloo():
    import torch.multiprocessing as mp
    from oracle import oracle as O
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q_out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q_out.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    n, d, B, k, x, q, ptr, doc, w, qp, qt, qv = _inputs()
    col = O.OracleCollection()
    col.add_dense("dense", x, "dot", "bf16")
    col.add_sparse("sparse", O.CSR(ptr, doc, w, n))
    res.sort(key=lambda t: t[0])
    for rank, s, i, c in res:
        for b in range(B):
            es, ei = O.hybrid_simple(col, q[b], qt[qp[b]:qp[b + 1]], qv[qp[b]:qp[b + 1]], k)
            assert int(c[b]) == len(ei)
            assert np.array_equal(i[b][:len(ei)], ei), (rank, b)
            assert np.array_equal(s[b][:len(ei)], es)
    assert np.array_equal(res[0][2], res[1][2])
