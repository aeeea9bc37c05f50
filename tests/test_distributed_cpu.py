"""The N>1 paths with world_size 2 on CPU (gloo): data-parallel embeddings without a collective on
the data path, and the row-sharded top-k with its all-gather + merge.  The per-rank compute is a
fake built on the oracle; the -m gpu suite covers the same merge on the CUDA kernels
(test_merge_shards_equals_single_index)."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


def _worker(rank, world, port, out):
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from fakes import FakeEncoder, FakeIndex
        from langstream_b200 import weights as W
        from langstream_b200.distributed import DataParallelEmbedder, ShardedIndex, shard_rows
        from oracle import topk_oracle

        # ---- embeddings: each rank embeds its own slice ----
        spec = W.TINY
        blob = W.random_init(spec, 1, True)
        ids, cu = W.synthetic_token_ids(11, 30, seed=5, min_len=1, vocab=spec.vocab)
        emb = DataParallelEmbedder(FakeEncoder(spec, blob), rank, world)
        (s0, s1), local = emb.embed_local(ids, cu)
        full = emb.embed_gathered(ids, cu)
        assert np.array_equal(full[s0:s1], local)

        # ---- top-k: row-sharded corpus, all-gather of [B,k] candidates, merge ----
        n, dim, B, k = 999, 64, 7, 10
        rows = topk_oracle.to_bf16_rounded(topk_oracle.normalize_rows(
            np.random.default_rng(42).standard_normal((n, dim), dtype=np.float32)))
        q = topk_oracle.to_bf16_rounded(topk_oracle.normalize_rows(
            np.random.default_rng(43).standard_normal((B, dim), dtype=np.float32)))
        lo, hi = shard_rows(n, world, rank)
        shard = FakeIndex(dim)
        shard.upsert(np.arange(lo, hi), rows[lo:hi])

        def local_search(queries, kk):
            i, s = shard.search(queries.numpy(), kk)
            return torch.from_numpy(i), torch.from_numpy(s)

        def merge(all_ids, all_sc):  # same rule as lsc_topk_merge_dev: score desc, ties to the lower shard
            W_, B_, k_ = all_ids.shape
            ids_ = all_ids.permute(1, 0, 2).reshape(B_, W_ * k_).numpy()
            sc_ = all_sc.permute(1, 0, 2).reshape(B_, W_ * k_).numpy()
            sc_ = np.where(ids_ >= 0, sc_, -np.inf)
            order = np.argsort(-sc_, axis=1, kind="stable")[:, :k_]
            return (torch.from_numpy(np.take_along_axis(ids_, order, 1)),
                    torch.from_numpy(np.take_along_axis(sc_, order, 1)))

        got_ids, got_sc = ShardedIndex(local_search, merge, world).search(torch.from_numpy(q), k)
        want_ids, want_sc = topk_oracle.search_numpy(rows, q, k)
        assert np.array_equal(got_ids.numpy(), want_ids) and np.allclose(got_sc.numpy(), want_sc, atol=1e-6)
        out[rank] = (s0, s1, full.tobytes())
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    world = 2
    port = 29600 + os.getpid() % 200
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert sorted(out.keys()) == [0, 1]
    assert out[0][2] == out[1][2]                    # both ranks hold the same gathered result
    assert out[0][0] == 0 and out[0][1] == out[1][0]  # contiguous, disjoint slices


def test_single_process_matches_oracle_for_each_slice():
    sys.path.insert(0, str(ROOT / "tests"))
    from fakes import FakeEncoder
    from langstream_b200 import weights as W
    from langstream_b200.distributed import DataParallelEmbedder
    from oracle.encoder_oracle import bert_numpy

    spec = W.TINY
    blob = W.random_init(spec, 1, True)
    ids, cu = W.synthetic_token_ids(9, 20, seed=2, min_len=0, vocab=spec.vocab)
    want = bert_numpy(spec, blob, ids, cu, dtype=np.float32)
    got = np.zeros_like(want)
    for r in range(3):
        (a, b), part = DataParallelEmbedder(FakeEncoder(spec, blob), r, 3).embed_local(ids, cu)
        got[a:b] = part
    assert np.allclose(got, want, atol=1e-6)
