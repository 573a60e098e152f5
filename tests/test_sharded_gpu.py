"""GPU, two or more devices on one box (skipped on a single-GPU box): the row-sharded paths on real engines.
  * the peer exchange (brag_shard_hybrid_topk: lists pushed over NVLink peer memory by the tail kernels, cross-shard
    bounds) between engines of one process, one host thread per engine -- equal to the unsharded call, bit for bit;
  * the drop-in client on a row-sharded collection, B200QdrantClient(devices=[0, 1]) -- the golden reference flow."""
import threading

import numpy as np
import pytest

from rag_application_b200 import synth

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def _shard_csr(ptr, doc, w, lo, hi, V):
    term_of = np.repeat(np.arange(V), np.diff(ptr))
    keep = (doc >= lo) & (doc < hi)
    p = np.zeros(V + 1, np.int64)
    np.add.at(p, term_of[keep] + 1, 1)
    return np.cumsum(p), (doc[keep] - lo).astype(np.uint32), w[keep]


@pytest.mark.parametrize("masked", [False, True])
def test_peer_exchange_two_engines_one_process(engine, masked):
    if _n_gpus() < 2:
        pytest.skip("needs two GPUs")
    from rag_application_b200 import Engine
    from rag_application_b200.sharded import shard_bounds
    G = min(_n_gpus(), 4)
    n, d, k, B, V = 60_000, 128, 50, 96, 3000
    x = synth.dense_corpus(n, d, seed=91, clustered=True)
    q = synth.dense_queries(x, B, seed=92)
    ptr, doc, w, _ = synth.zipf_corpus(n, V, 60, seed=93)
    qp, qt, qv = synth.sparse_queries(ptr, doc, n, B, 8, "zipf", seed=94, skip_top=8)
    mask = synth.random_mask(n, 0.3) if masked else None
    engine.dense_set("dense", x, "bf16")
    engine.sparse_set_csr("sparse", ptr, doc, w, n)
    if masked:
        engine.mask_set(3, mask)
    want = engine.hybrid_topk("dense", "sparse", q, qp, qt, qv, k, mask_id=3 if masked else 0)
    shards = [Engine(g) for g in range(G)]
    try:
        for g, e in enumerate(shards):
            lo, hi = shard_bounds(n, G, g)
            e.set_id_offset(lo)
            e.dense_set("dense", x[lo:hi], "bf16")
            sp, sd, sw = _shard_csr(ptr, doc, w, lo, hi, V)
            e.sparse_set_csr("sparse", sp, sd, sw, hi - lo)
            if masked:
                e.mask_set(3, mask[lo:hi])
            e.shard_window_create(G, 128, 64)
        for g, e in enumerate(shards):
            e.shard_connect(g, G, local_peers=shards)
        for rep in range(3):                      # several epochs: both slot parities, flags from earlier calls around
            out = [None] * G

            def run(g):
                out[g] = shards[g].shard_hybrid_topk("dense", "sparse", q, qp, qt, qv, k, mask_id=3 if masked else 0,
                                                     stages=True)
            ts = [threading.Thread(target=run, args=(g,)) for g in range(G)]
            for t in ts:
                t.start()
            for t in ts:
                t.join(timeout=120)
            for g in range(G):
                s, i, c, _ = out[g]
                assert np.array_equal(c, want[2]) and np.array_equal(i, want[1]) and np.array_equal(s, want[0]), (rep, g)
        for e in shards:
            e.shard_close()
    finally:
        for e in shards:
            e.close()


def test_client_on_a_row_sharded_collection_two_gpus():
    if _n_gpus() < 2:
        pytest.skip("needs two GPUs")
    from _client_cases import run_incremental_upsert, run_reference_flow
    run_reference_flow(devices=[0, 1], shard_block_rows=32)
    run_incremental_upsert(devices=[0, 1], shard_block_rows=32)
