"""Row-sharded hybrid search across the GPUs of one box.

One engine per GPU.  Rank r owns corpus rows / documents [offset_r, offset_r + n_r) and answers with
GLOBAL ids.  A stage's per-shard ranked list travels as sorted 64-bit keys (ordered score << 32 |
0xFFFFFFFF - global row): dense and sparse lists are merged separately (RRF needs global ranks) and
then fused, all in ONE kernel per rank (brag_shard_fuse / brag_shard_hybrid_topk).  Queries are replicated.

Two exchanges, same kernels, same results:
  "peer"        the product path on a box: every rank pushes its lists straight into every peer's window over
                NVLink while it scans (brag_shard_hybrid_topk; no collective call, CUDA-graph capturable), and the
                shards tell each other their running score bounds;
  "collective"  brag_shard_local_lists -> torch.distributed all-gather of the keys -> brag_shard_fuse; any
                backend (the world_size-2 gloo tests on CPU run it with the oracle-backed stand-in engine), and
                the checker of the peer path in bench.py."""
from __future__ import annotations

from typing import Tuple

import numpy as np


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous split; the first n % world ranks get one extra row."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def keys_to_lists(keys) -> Tuple[np.ndarray, np.ndarray]:
    """Packed keys (any shape, int64 / uint64 bits) -> (float32 scores, int64 global ids); empty slots -> (-inf, -1).
    The inverse of the engine's key format (csrc/common.cuh: make_key), for checkers."""
    u = np.ascontiguousarray(keys).view(np.uint64)
    o = (u >> np.uint64(32)).astype(np.uint32)
    bits = np.where(o & np.uint32(0x80000000), o & np.uint32(0x7FFFFFFF), ~o).astype(np.uint32)
    s = bits.view(np.float32).copy()
    i = (np.uint64(0xFFFFFFFF) - (u & np.uint64(0xFFFFFFFF))).astype(np.int64)
    empty = u == 0
    s[empty] = -np.inf
    i[empty] = -1
    return s, i


def lists_to_keys(scores, ids) -> np.ndarray:
    """(float32 scores, int64 global ids, -1 = empty) -> packed keys as int64 bits (csrc/common.cuh: make_key)."""
    s = np.ascontiguousarray(scores, dtype=np.float32) + np.float32(0.0)        # -0 -> +0
    u = s.view(np.uint32)
    o = np.where(u & np.uint32(0x80000000), ~u, u | np.uint32(0x80000000)).astype(np.uint64)
    i = np.asarray(ids, dtype=np.int64)
    k = (o << np.uint64(32)) | (np.uint64(0xFFFFFFFF) - np.where(i >= 0, i, 0).astype(np.uint64))
    k[i < 0] = 0
    return k.view(np.int64)


class ShardedHybrid:
    """Hybrid top-k over a row-sharded corpus, one process per GPU.  `engine` is this rank's
    rag_application_b200.Engine (or any object with shard_local_lists / shard_fuse of the same contract)."""

    def __init__(self, engine, rank: int, world: int, id_offset: int, group=None, exchange: str = "auto",
                 max_batch: int = 4096, max_k: int = 128):
        self.engine, self.rank, self.world, self.group = engine, rank, world, group
        engine.set_id_offset(id_offset)
        if exchange == "auto":
            exchange = "peer" if world > 1 and hasattr(engine, "shard_window_create") else "collective"
        self.exchange = exchange
        if exchange == "peer":
            import torch.distributed as dist
            handle = engine.shard_window_create(world, max_batch, max_k)
            handles = [None] * world
            if world > 1:
                dist.all_gather_object(handles, handle, group=group)
            else:
                handles = [handle]
            engine.shard_connect(rank, world, handles=handles)
            if world > 1:
                dist.barrier(group=group)          # nobody pushes before everybody has mapped everybody

    def gather_keys(self, keys):
        """[2, B, k] keys of this rank -> [world, 2, B, k] (one all-gather)."""
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return keys[None]
        flat = keys.reshape(-1)          # NCCL accepts any [world, ...] view, gloo insists on a concatenated dim 0
        out = torch.empty((self.world * flat.numel(),), dtype=keys.dtype, device=keys.device)
        dist.all_gather_into_tensor(out, flat, group=self.group)
        return out.view((self.world,) + tuple(keys.shape))

    def hybrid_topk(self, dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id=0, max_terms=-1,
                    stages=False, exchange=None, out=None):
        """Device tensors in, device tensors out: (rrf scores f64 [B,k], ids [B,k], counts [B]) -- identical on every
        rank -- plus the merged dense / sparse lists when `stages`."""
        if (exchange or self.exchange) == "peer":
            return self.engine.shard_hybrid_topk(dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id,
                                                 max_terms=max_terms, stages=stages, out=out)
        keys = self.engine.shard_local_lists(dense_name, sparse_name, Q, q_ptr, q_terms, q_vals, k, mask_id,
                                             max_terms=max_terms)
        return self.engine.shard_fuse(self.gather_keys(keys), k, stages=stages)
