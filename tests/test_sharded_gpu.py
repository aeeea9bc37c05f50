"""The row-sharded table (lsc_sharded_index, lsc_exchange, lsc_index_search_sharded_dev) through the C ABI against
the CPU oracle and against one unsharded index.  On a single GPU the shards share the device (and split its SMs);
with more GPUs visible the same tests also run with one shard per GPU, and the IPC form of the exchange - one
process per GPU, the shape torchrun / one agent pod per GPU gives - runs in two spawned processes."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _quantised(n, dim, seed):
    from oracle import topk_oracle

    x = np.random.default_rng(seed).standard_normal((n, dim), dtype=np.float32)
    return topk_oracle.to_bf16_rounded(topk_oracle.normalize_rows(x))


def _device_sets():
    import torch

    n = torch.cuda.device_count()
    sets = [[0, 0], [0, 0, 0, 0]]
    if n >= 2:
        sets.append(list(range(min(n, 8))))
    return sets


@pytest.mark.parametrize("B,k", [(1, 10), (64, 10), (129, 10), (300, 16), (5, 20), (40, 32)])
def test_sharded_search_matches_oracle_and_single_index(native_lib, cuda_device, B, k):
    from langstream_b200._native import Index, ShardedIndex
    from oracle import topk_oracle

    dim, n = 384, 30_011
    rows = _quantised(n, dim, 1)
    q = _quantised(B, dim, 2 + B)
    ids = np.arange(n, dtype=np.int64) * 3 + 7
    full = topk_oracle.scores_numpy(rows, q)
    whole = Index(dim, capacity=n)
    whole.upsert(ids, rows)
    want_ids, want_sc = whole.search(q, k)
    whole.close()
    for devices in _device_sets():
        sh = ShardedIndex(dim, devices)
        for o in range(0, n, 7000):  # several batches: new ids keep going to the emptiest shard
            sh.upsert(ids[o:o + 7000], rows[o:o + 7000])
        assert len(sh) == n
        sizes = [len(sh.shard(i)) for i in range(len(devices))]
        assert max(sizes) - min(sizes) <= 1, sizes
        got_ids, got_sc = sh.search(q, k)
        topk_oracle.assert_topk_matches((got_ids - 7) // 3, got_sc, full, k, tol=TOL)
        # same rows, same bf16 operands, same tensor-core accumulation: the scores are bit-identical to the
        # unsharded index's, and so are the ids wherever scores do not tie
        assert np.array_equal(got_sc, want_sc), devices
        differ = got_ids != want_ids
        assert not differ.any() or np.array_equal(got_sc[differ], want_sc[differ]), devices
        again_ids, again_sc = sh.search(q, k)   # the receive buffers alternate between searches
        assert np.array_equal(again_ids, got_ids) and np.array_equal(again_sc, got_sc)
        sh.close()


def test_sharded_upsert_replace_delete(native_lib, cuda_device):
    """JdbcWriter semantics across shards: an existing id is replaced where it lives, a null value deletes,
    unknown ids are ignored (JdbcWriter.java:139-179)."""
    from langstream_b200._native import ShardedIndex

    dim = 64
    e = np.eye(dim, dtype=np.float32)
    for devices in _device_sets():
        sh = ShardedIndex(dim, devices)
        sh.upsert([10, 20, 30, 40, 50], e[:5])
        assert len(sh) == 5
        ids, sc = sh.search(e[1], 3)
        assert ids[0, 0] == 20 and abs(sc[0, 0] - 1.0) < 1e-3
        sh.upsert([20], e[7:8])                       # replace: still 5 rows, id 20 now points along e7
        assert len(sh) == 5
        assert sh.search(e[7], 1)[0][0, 0] == 20
        assert abs(sh.search(e[1], 1)[1][0, 0]) < 1e-6
        sh.delete([10, 999, 40])
        assert len(sh) == 3
        ids, _ = sh.search(e[0], 5)
        assert sorted(ids[0][ids[0] >= 0].tolist()) == [20, 30, 50] and (ids[0, 3:] == -1).all()
        sh.upsert([7, 7], np.stack([e[8], e[9]]))     # duplicate id in one batch: last wins
        assert len(sh) == 4 and sh.search(e[9], 1)[0][0, 0] == 7
        sh.delete([20, 30, 50, 7])
        assert len(sh) == 0
        assert (sh.search(e[0], 2)[0] == -1).all()
        sh.close()


def test_sharded_duplicates_keep_storage_order(native_lib, cuda_device):
    """Duplicate rows (equal scores) across contiguous shards come back in the order of one unsharded index:
    rank first, then position - checked through the device-pointer call with shards filled range by range."""
    import torch
    from langstream_b200._native import Exchange, Index

    dim, B, k, n, world = 64, 6, 10, 4000, 4
    base = _quantised(30, dim, 5)
    rows = base[np.random.default_rng(6).integers(0, 30, n)]
    q = base[:B]
    whole = Index(dim)
    whole.upsert(np.arange(n), rows)
    want_ids, want_sc = whole.search(q, k)
    whole.close()
    os.environ["LSC_TOPK_MAX_CTAS"] = "32"  # four shards side by side on one GPU
    try:
        shards = [Index(dim) for _ in range(world)]
    finally:
        del os.environ["LSC_TOPK_MAX_CTAS"]
    xs = [Exchange(0, r, world, 64) for r in range(world)]
    Exchange.connect_local(xs)
    cuts = [0, 900, 2100, 3000, n]
    for r in range(world):
        shards[r].upsert(np.arange(cuts[r], cuts[r + 1]), rows[cuts[r]:cuts[r + 1]])
        shards[r].reserve(B, k)  # no allocation (= device-wide synchronise) once a peer's scan is waiting
    dq = torch.from_numpy(q).to(cuda_device)
    streams = [torch.cuda.Stream() for _ in range(world)]
    outs = [(torch.empty(B, k, dtype=torch.int64, device=cuda_device), torch.empty(B, k, dtype=torch.float32, device=cuda_device))
            for _ in range(world)]
    torch.cuda.synchronize()
    for rep in range(3):
        for r in range(world):
            shards[r].search_sharded_dev(xs[r], dq, B, k, outs[r][0], outs[r][1], streams[r].cuda_stream)
        for r in range(world):
            shards[r].check(streams[r].cuda_stream)
        for o_ids, o_sc in outs:  # every rank holds the merged result
            assert (o_ids.cpu().numpy() == want_ids).all()
            assert np.array_equal(o_sc.cpu().numpy(), want_sc)
    for s in shards:
        s.close()
    for x in xs:
        x.close()


def _ipc_rank(rank, world, port, n, dim, B, k, devices, result_dir):
    import torch
    import torch.distributed as dist

    os.environ["LSC_TOPK_MAX_CTAS"] = "64"  # the ranks may share one GPU
    from langstream_b200._native import Exchange, Index

    dev = devices[rank]
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    rows = _quantised(n, dim, 1)
    q = _quantised(B, dim, 2)
    lo, hi = n * rank // world, n * (rank + 1) // world
    ix = Index(dim, capacity=hi - lo, device=dev)
    ix.upsert(np.arange(lo, hi), rows[lo:hi])
    x = Exchange(dev, rank, world, 256)
    handles = [None] * world
    dist.all_gather_object(handles, x.ipc_handle())
    x.connect_ipc(handles)
    dist.barrier()
    d = torch.device("cuda", dev)
    dq = torch.from_numpy(q).to(d)
    o_ids = torch.empty(B, k, dtype=torch.int64, device=d)
    o_sc = torch.empty(B, k, dtype=torch.float32, device=d)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        ix.search_sharded_dev(x, dq, B, k, o_ids, o_sc, st)
        ix.check(st)
    np.save(os.path.join(result_dir, f"ids{rank}.npy"), o_ids.cpu().numpy())
    np.save(os.path.join(result_dir, f"sc{rank}.npy"), o_sc.cpu().numpy())
    dist.barrier()
    ix.close()
    x.close()
    dist.destroy_process_group()


def test_exchange_between_processes_over_ipc(native_lib, cuda_device, tmp_path):
    """One process per shard, receive buffers mapped through CUDA IPC handles that travel over the processes' own
    channel (gloo here, as in bench.py under torchrun): every rank ends with the oracle's answer.  Needs two GPUs:
    two processes on ONE GPU are time-sliced, never resident together, so their scans cannot wait for each other."""
    import torch
    import torch.multiprocessing as mp
    from oracle import topk_oracle

    world = 2
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    devices = [0, 1]
    n, dim, B, k = 20_000, 128, 70, 10
    port = 29600 + os.getpid() % 300
    mp.spawn(_ipc_rank, args=(world, port, n, dim, B, k, devices, str(tmp_path)), nprocs=world, join=True)
    full = topk_oracle.scores_numpy(_quantised(n, dim, 1), _quantised(B, dim, 2))
    for r in range(world):
        ids = np.load(tmp_path / f"ids{r}.npy")
        sc = np.load(tmp_path / f"sc{r}.npy")
        topk_oracle.assert_topk_matches(ids, sc, full, k, tol=TOL)
