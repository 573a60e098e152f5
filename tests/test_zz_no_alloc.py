"""GPU: brag_dense_rescore_noalloc (include/brag.h) -- the allocation-free variant of the candidate re-score
(no counterpart in the reference) -- against the existing brag_dense_rescore, through
the C ABI, on two calls in a row.  Both run the same two kernels, so ids, counts and score BITS must be equal (the tests'
tolerance for re-scored candidates is exact).  The file sorts last on purpose."""
import numpy as np
import pytest

from rag_application_b200 import synth

pytestmark = pytest.mark.gpu

MAX_KEYS = 65536            # BRAG_NOALLOC_MAX_KEYS


def _case(engine, torch, n, d, B, m, seed, dtype):
    dev = torch.device("cuda", 0)
    x = synth.dense_corpus(n, d, seed=seed)
    q = synth.dense_queries(x, B, seed=seed + 1)
    engine.dense_set("na", x, dtype)
    rng = np.random.default_rng(seed)
    cand = rng.integers(0, n, size=(B, m)).astype(np.int64)
    cand[:, : m // 8] = cand[:, m // 8: 2 * (m // 8)]          # duplicates (union semantics)
    cand[:, -3:] = -1                                           # empty slots
    return torch.from_numpy(q).to(dev), torch.from_numpy(cand).to(dev)


def _outs(torch, B, k):
    dev = torch.device("cuda", 0)
    return (torch.full((B, k), 7.0, dtype=torch.float32, device=dev), torch.full((B, k), 7, dtype=torch.int64, device=dev),
            torch.full((B,), 7, dtype=torch.int32, device=dev))


def _same(torch, got, want):
    torch.cuda.synchronize()
    return (torch.equal(got[1], want[1]) and torch.equal(got[2], want[2]) and
            torch.equal(got[0].view(torch.int32), want[0].view(torch.int32)))


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_noalloc_rescore_equals_the_allocating_entry_point_twice_in_a_row(engine, dtype):
    import torch
    n, d, k = 5000, 256, 40
    mask = synth.random_mask(n, 0.5, seed=3)
    engine.mask_set(7, mask)
    # two different batches back to back: the second call finds the first one's keys in the shared static scratch
    qa, ca = _case(engine, torch, n, d, 16, 96, 11, dtype)
    qb, cb = qa[:5].contiguous(), torch.flip(ca[:5], dims=[1]).contiguous()[:, :50].contiguous()
    for mid in (0, 7):
        want_a = engine.dense_rescore("na", qa, ca, k, mask_id=mid)
        want_b = engine.dense_rescore("na", qb, cb, k, mask_id=mid)
        out_a, out_b = _outs(torch, 16, k), _outs(torch, 5, k)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        engine.dense_rescore_noalloc("na", qa, ca, k, out_a, mask_id=mid)
        engine.dense_rescore_noalloc("na", qb, cb, k, out_b, mask_id=mid)
        assert _same(torch, out_a, want_a) and _same(torch, out_b, want_b)
        # and again, the other way round, into fresh outputs
        out_a2, out_b2 = _outs(torch, 16, k), _outs(torch, 5, k)
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
        engine.dense_rescore_noalloc("na", qb, cb, k, out_b2, mask_id=mid)
        engine.dense_rescore_noalloc("na", qa, ca, k, out_a2, mask_id=mid)
        assert _same(torch, out_a2, want_a) and _same(torch, out_b2, want_b)
        # no device memory was taken by the calls themselves (the first pair may have loaded the kernels' module)
        assert torch.cuda.mem_get_info()[0] >= free1 and free1 >= free0 - (64 << 20)
        if mid == 7:
            rows = out_a[1][out_a[1] >= 0].cpu().numpy()
            assert mask[rows].all()


def test_noalloc_rescore_at_the_limit_and_beyond(engine):
    import torch
    n, d, k = 3000, 64, 10
    q, c = _case(engine, torch, n, d, 8, 8192, 5, "f16")          # B * m = 65536 = the whole static array
    want = engine.dense_rescore("na", q, c, k)
    out = _outs(torch, 8, k)
    engine.dense_rescore_noalloc("na", q, c, k, out)
    assert _same(torch, out, want)
    q9 = torch.cat([q, q[:1]]).contiguous()
    c9 = torch.cat([c, c[:1]]).contiguous()                       # 9 * 8192 keys: one row too many
    with pytest.raises(Exception, match="BRAG_NOALLOC_MAX_KEYS"):
        engine.dense_rescore_noalloc("na", q9, c9, k, _outs(torch, 9, k))
    with pytest.raises(ValueError):
        engine.dense_rescore_noalloc("na", q.cpu().numpy(), c.cpu().numpy(), k, out)      # host buffers are refused
    # the limit is what the header says
    import re, os
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "brag.h")).read()
    assert int(re.search(r"#define BRAG_NOALLOC_MAX_KEYS (\d+)", hdr).group(1)) == MAX_KEYS
