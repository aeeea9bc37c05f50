"""Multi-GPU plumbing: one process per GPU, torch.distributed (NCCL on GPUs, gloo in the CPU tests).

Embeddings shard naturally (the reference embeds texts independently,
AbstractHuggingFaceEmbeddingService.java:191-206): every rank embeds a contiguous slice of the
sequence list, balanced by token count, with a full weight replica and NO data-path collective.
Top-k has one exchange step: the corpus is row-sharded, queries are replicated, each rank produces
its local [B,k] (score,id) lists, one all-gather moves them (80 B – 80 KiB per rank) and every rank
merges the `world` sorted lists per query (lsc_topk_merge_dev).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np


def split_by_tokens(cu_seqlens: np.ndarray, world: int) -> List[Tuple[int, int]]:
    """Contiguous [start, end) sequence ranges, one per rank, balanced by TOKEN count so that packed
    variable-length batches stay even.  Every sequence belongs to exactly one rank; ranks may be empty."""
    cu = np.asarray(cu_seqlens, np.int64)
    n = len(cu) - 1
    total = int(cu[-1])
    bounds = [0]
    for r in range(1, world):
        target = total * r / world
        # first sequence boundary whose prefix token count reaches the target
        i = int(np.searchsorted(cu, target, side="left"))
        i = min(max(i, bounds[-1]), n)
        bounds.append(i)
    bounds.append(n)
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def shard_rows(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Row range [lo, hi) of the corpus shard owned by `rank`."""
    return n_rows * rank // world, n_rows * (rank + 1) // world


class DataParallelEmbedder:
    """Embeds this rank's slice of a global batch.  `encoder` is a langstream_b200._native.Encoder
    (or anything with .embed(token_ids, cu_seqlens) -> [n, hidden] float32)."""

    def __init__(self, encoder, rank: int, world: int):
        self.encoder, self.rank, self.world = encoder, rank, world

    def embed_local(self, token_ids: np.ndarray, cu_seqlens: np.ndarray) -> Tuple[Tuple[int, int], np.ndarray]:
        s0, s1 = split_by_tokens(cu_seqlens, self.world)[self.rank]
        cu = np.asarray(cu_seqlens, np.int64)
        ids = np.asarray(token_ids, np.int32)[cu[s0]:cu[s1]]
        local_cu = (cu[s0:s1 + 1] - cu[s0]).astype(np.int32)
        return (s0, s1), self.encoder.embed(ids, local_cu)

    def embed_gathered(self, token_ids, cu_seqlens):
        """Convenience for callers that want the whole result on every rank (not the hot path)."""
        import torch
        import torch.distributed as dist

        (s0, s1), local = self.embed_local(token_ids, cu_seqlens)
        parts: List[Optional[tuple]] = [None] * self.world
        dist.all_gather_object(parts, (s0, s1, torch.from_numpy(np.ascontiguousarray(local))))
        n = len(cu_seqlens) - 1
        out = np.zeros((n, local.shape[1] if local.size else parts[0][2].shape[1]), np.float32)
        for a, b, t in parts:
            if b > a:
                out[a:b] = t.numpy()
        return out


class ShardedIndex:
    """Row-sharded top-k.  `local_search(queries[B,D] tensor, k) -> (ids[B,k] int64, scores[B,k] fp32)`
    and `merge(all_ids[W,B,k], all_scores[W,B,k]) -> (ids[B,k], scores[B,k])` run on this rank's
    device; on GPUs they are lsc_index_search_dev and lsc_topk_merge_dev (see `for_gpu`)."""

    def __init__(self, local_search, merge, world: int):
        self.local_search, self.merge, self.world = local_search, merge, world

    @classmethod
    def for_gpu(cls, index, world: int, device):
        import torch
        from . import _native

        def local_search(q, k):
            B = q.shape[0]
            ids = torch.empty(B, k, dtype=torch.int64, device=device)
            sc = torch.empty(B, k, dtype=torch.float32, device=device)
            index.search_dev(q, B, k, ids, sc, torch.cuda.current_stream().cuda_stream)
            return ids, sc

        def merge(all_ids, all_sc):
            W, B, k = all_ids.shape
            ids = torch.empty(B, k, dtype=torch.int64, device=device)
            sc = torch.empty(B, k, dtype=torch.float32, device=device)
            _native.topk_merge_dev(all_ids, all_sc, W, B, k, ids, sc, torch.cuda.current_stream().cuda_stream)
            return ids, sc

        return cls(local_search, merge, world)

    def search(self, queries, k: int):
        import torch
        import torch.distributed as dist

        ids, sc = self.local_search(queries, k)
        if self.world == 1:
            return ids, sc
        all_ids = torch.empty((self.world,) + tuple(ids.shape), dtype=ids.dtype, device=ids.device)
        all_sc = torch.empty((self.world,) + tuple(sc.shape), dtype=sc.dtype, device=sc.device)
        # all_gather into the rows of the stacked buffers (works with NCCL and gloo alike)
        dist.all_gather(list(all_ids.unbind(0)), ids.contiguous())
        dist.all_gather(list(all_sc.unbind(0)), sc.contiguous())
        return self.merge(all_ids, all_sc)
