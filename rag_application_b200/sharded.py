"""Row-sharded hybrid search across the GPUs of one box.

One process per GPU.  Rank r owns corpus rows / documents [offset_r, offset_r + n_r) and answers
with GLOBAL ids.  Per query batch there is exactly one exchange step: an all-gather of the two
per-shard ranked lists (dense and sparse are merged separately, because RRF needs global ranks),
followed on every rank by the k-way merge kernel and the RRF kernel.  Queries are replicated.

`merge_lists` / `fuse` are pure functions of gathered tensors so the same code path is exercised
by the world_size-2 gloo tests on CPU (tests/test_sharded_cpu.py) with a numpy stand-in for the
two kernels' contracts."""
from __future__ import annotations

from typing import Callable, Optional, Tuple


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous split; the first n % world ranks get one extra row."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_lists(dense, sparse):
    """(scores [B,k] f32, ids [B,k] i64, counts [B] i32) x 2 -> one contiguous int64 payload
    [2, B, 2k+1] so the exchange is a single collective.  Scores travel as raw bits."""
    import torch
    out = []
    for s, i, c in (dense, sparse):
        sb = s.contiguous().view(torch.int32).to(torch.int64)
        out.append(torch.cat([sb, i, c.to(torch.int64)[:, None]], dim=1))
    return torch.stack(out).contiguous()


def unpack_lists(payload, k: int):
    """[G, 2, B, 2k+1] -> per list: (scores [G,B,k] f32, ids [G,B,k] i64, counts [G,B] i32)."""
    import torch
    res = []
    for l in range(2):
        p = payload[:, l]
        s = p[:, :, :k].to(torch.int32).view(torch.float32).contiguous()
        i = p[:, :, k:2 * k].contiguous()
        c = p[:, :, 2 * k].to(torch.int32).contiguous()
        res.append((s, i, c))
    return res


class ShardedHybrid:
    """Hybrid top-k over a row-sharded corpus.  `engine` is this rank's rag_application_b200.Engine
    (or any object with dense_topk / sparse_topk / merge_topk / rrf of the same contract)."""

    def __init__(self, engine, rank: int, world: int, id_offset: int, group=None):
        self.engine, self.rank, self.world, self.group = engine, rank, world, group
        engine.set_id_offset(id_offset)

    def local_lists(self, dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id=0):
        d = self.engine.dense_topk(dense_name, Q, k, mask_id)
        s = self.engine.sparse_topk(sparse_name, q_ptr, q_terms, q_vals, k, mask_id)
        return d, s

    def exchange(self, payload):
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return payload[None]
        # flat buffers: NCCL accepts any [world, ...] view, gloo insists on a concatenated dim 0
        flat = payload.reshape(-1)
        out = torch.empty((self.world * flat.numel(),), dtype=payload.dtype, device=payload.device)
        dist.all_gather_into_tensor(out, flat, group=self.group)
        return out.view((self.world,) + tuple(payload.shape))

    def hybrid_topk(self, dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id=0):
        """Device tensors in, device tensors out: (rrf scores f64 [B,k], ids [B,k], counts [B]),
        identical on every rank."""
        import torch
        d, s = self.local_lists(dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id)
        gathered = self.exchange(pack_lists(d, s))
        (ds, di, dc), (ss, si, sc) = unpack_lists(gathered, k)
        md = self.engine.merge_topk(ds, di, dc)
        ms = self.engine.merge_topk(ss, si, sc)
        ids = torch.stack([md[1], ms[1]]).contiguous()
        cnt = torch.stack([md[2], ms[2]]).contiguous()
        return self.engine.rrf(ids, cnt, k)
