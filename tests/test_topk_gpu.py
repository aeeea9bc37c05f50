"""Brute-force cosine top-k through the C ABI against the CPU oracle (bit-exact ids except ties
within 1e-5, the tolerance BASELINE.json states)."""
import json
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).parent / "golden"
TOL = 1e-5


def _rand(n, dim, seed):
    return np.random.default_rng(seed).standard_normal((n, dim), dtype=np.float32)


def _quantised(n, dim, seed):
    """Random unit rows already rounded to bf16: the index's normalise+round is idempotent on them
    (re-normalising moves a value by ~1e-4 relative, far inside a bf16 half-ulp), so the oracle and
    the index hold EXACTLY the same operands and only the fp32 accumulation order differs.  With raw
    fp32 inputs a 1-ulp difference in the norm can flip one bf16 rounding (~1e-5 on a score)."""
    from oracle import topk_oracle

    return topk_oracle.to_bf16_rounded(topk_oracle.normalize_rows(_rand(n, dim, seed)))


def test_golden_small_corpus(native_lib, cuda_device):
    """tests/golden/topk_small.npz: scalar ReRankAgent-form oracle, 2000 x 384, 5 queries, k=10."""
    from langstream_b200._native import Index
    from oracle import topk_oracle

    fx = np.load(GOLDEN / "topk_small.npz")
    rows = _rand(int(fx["n"]), int(fx["dim"]), int(fx["corpus_seed"]))
    q = _rand(int(fx["B"]), int(fx["dim"]), int(fx["query_seed"]))
    ix = Index(int(fx["dim"]))
    ix.upsert(np.arange(rows.shape[0]), rows)
    ids, scores = ix.search(q, int(fx["k"]))
    # raw fp32 inputs: against the fp32 scalar oracle the bf16 storage moves scores by < 4e-3 and
    # ids may swap inside that margin (exact-operand parity is test_search_matches_oracle)
    full = topk_oracle.scores_numpy(rows, q)
    topk_oracle.assert_topk_matches(ids, scores, full, int(fx["k"]), tol=1e-4)
    exp_ids, exp_scores = fx["expected_ids"], fx["expected_scores"]
    assert np.abs(scores - exp_scores).max() < 4e-3
    for b in range(ids.shape[0]):
        for i in ids[b]:
            assert topk_oracle.cosine_scalar(rows[i], q[b]) >= exp_scores[b, -1] - 4e-3
    assert exp_ids.shape == ids.shape
    ix.close()


@pytest.mark.parametrize("n,B,k", [(1, 1, 1), (5, 3, 10), (255, 1, 10), (256, 2, 10), (257, 64, 10),
                                   (40000, 1, 10), (40000, 64, 20), (100003, 130, 10), (30000, 300, 32),
                                   (5000, 7, 128),
                                   # more than one query block, k <= 16: the CTA-pair scan (256 queries per pass)
                                   (300, 129, 10), (129, 256, 10), (100003, 257, 10), (5000, 512, 16), (70000, 1000, 10)])
def test_search_matches_oracle(native_lib, cuda_device, n, B, k):
    from langstream_b200._native import Index
    from oracle import topk_oracle

    dim = 384
    rows = _quantised(n, dim, 42 + n)
    q = _quantised(B, dim, 43 + B)
    ix = Index(dim, capacity=n)
    ix.upsert(np.arange(n), rows)
    assert len(ix) == n
    ids, scores = ix.search(q, k)
    full = topk_oracle.scores_numpy(rows, q)
    topk_oracle.assert_topk_matches(ids, scores, full, k, tol=TOL)
    ix.close()


@pytest.mark.parametrize("dim", [64, 128, 256, 512])
def test_other_dimensions(native_lib, cuda_device, dim):
    from langstream_b200._native import Index
    from oracle import topk_oracle

    rows, q = _quantised(3000, dim, 1), _quantised(9, dim, 2)
    ix = Index(dim)
    ix.upsert(np.arange(3000), rows)
    ids, scores = ix.search(q, 10)
    topk_oracle.assert_topk_matches(ids, scores, topk_oracle.scores_numpy(rows, q), 10, tol=TOL)
    ix.close()


def test_rerank_it_rows(native_lib, cuda_device):
    """The reference's own pipeline fixture (RerankAgentRunnerIT): 10 rows [i,2,3,4,5] padded to 64
    dims, query [1,2,3,4,5], LIMIT 20 returns all 10 rows, best first = text1."""
    from langstream_b200._native import Index
    from oracle import topk_oracle

    fx = json.loads((GOLDEN / "rerank_it.json").read_text())
    rows = np.zeros((10, 64), np.float32)
    for i, r in enumerate(fx["rows"]):
        rows[i, :5] = r["embeddings_vector"]
    q = np.zeros((1, 64), np.float32)
    q[0, :5] = fx["query_embeddings"]
    ix = Index(64)
    ix.upsert(np.arange(10), rows)
    ids, scores = ix.search(q, fx["limit"])
    assert ids[0, 0] == 1
    assert sorted(ids[0, :10].tolist()) == list(range(10))
    assert (ids[0, 10:] == -1).all()
    topk_oracle.assert_topk_matches(ids, scores, topk_oracle.scores_numpy(rows, q), fx["limit"], tol=TOL)
    ix.close()


def test_upsert_replace_delete_semantics(native_lib, cuda_device):
    """JdbcWriter semantics: UPDATE existing id, INSERT new id, DELETE on null value
    (JdbcWriterTest.java:134-170 does the same round trip on one row)."""
    from langstream_b200._native import Index

    dim = 64
    ix = Index(dim)
    e = np.eye(dim, dtype=np.float32)
    ix.upsert([10, 20, 30], e[:3])
    ids, scores = ix.search(e[1], 3)
    assert ids[0, 0] == 20 and abs(scores[0, 0] - 1.0) < 1e-3
    ix.upsert([20], e[5:6])            # replace id 20 with another direction
    assert len(ix) == 3
    ids, _ = ix.search(e[5], 1)
    assert ids[0, 0] == 20
    ids, scores = ix.search(e[1], 3)
    assert abs(scores[0, 0]) < 1e-6     # nobody points along e1 any more
    ix.delete([10, 999])               # unknown id ignored
    assert len(ix) == 2
    ids, _ = ix.search(e[0], 3)
    assert 10 not in ids[0].tolist() and ids[0, 2] == -1
    ix.upsert([7, 7], np.stack([e[8], e[9]]))  # duplicate id in one batch: last wins
    assert len(ix) == 3
    ids, _ = ix.search(e[9], 1)
    assert ids[0, 0] == 7
    ix.delete([20, 30, 7])
    assert len(ix) == 0
    ids, scores = ix.search(e[0], 2)
    assert (ids == -1).all()
    ix.close()


def test_zero_vector_scores_zero(native_lib, cuda_device):
    """ReRankAgent.cosineSimilarity returns 0 when a norm is 0 (ReRankAgent.java:262-273)."""
    from langstream_b200._native import Index

    dim = 64
    rows = np.zeros((3, dim), np.float32)
    rows[1, 0] = -1.0
    rows[2, 0] = 2.0
    ix = Index(dim)
    ix.upsert([0, 1, 2], rows)
    q = np.zeros((1, dim), np.float32)
    q[0, 0] = 3.0
    ids, scores = ix.search(q, 3)
    assert ids[0].tolist() == [2, 0, 1]
    assert np.allclose(scores[0], [1.0, 0.0, -1.0], atol=1e-6)
    ix.close()


def test_save_load_roundtrip(native_lib, cuda_device, tmp_path):
    from langstream_b200._native import Index

    rows, q = _rand(1000, 128, 3), _rand(4, 128, 4)
    ix = Index(128)
    ix.upsert(np.arange(1000) * 3, rows)
    a = ix.search(q, 5)
    ix.save(tmp_path / "ix.bin")
    ix.close()
    ix2 = Index.load(tmp_path / "ix.bin", dim=128)
    assert len(ix2) == 1000
    b = ix2.search(q, 5)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
    ix2.close()


def test_full_size_properties(native_lib, cuda_device):
    """2M x 384 on device (size-independent checks + the on-device fp32 checker kernel)."""
    import torch
    from langstream_b200._native import Index
    from oracle import topk_oracle

    n, dim, B, k = 2_000_000, 384, 16, 10
    g = torch.Generator(device=cuda_device).manual_seed(42)
    rows = torch.randn(n, dim, generator=g, device=cuda_device)
    q = torch.randn(B, dim, generator=torch.Generator(device=cuda_device).manual_seed(43), device=cuda_device)
    q[3] = rows[123456]  # a planted exact match must come first with score ~1
    ix = Index(dim, capacity=n)
    st = torch.cuda.current_stream().cuda_stream
    ix.upsert_dev(np.arange(n, dtype=np.int64), rows, n, st)
    out_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
    out_scores = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
    ix.search_dev(q, B, k, out_ids, out_scores, st)
    full = torch.empty(B, n, dtype=torch.float32, device=cuda_device)
    ix.scores_naive_dev(q, B, full, st)
    torch.cuda.synchronize()
    ids, scores = out_ids.cpu().numpy(), out_scores.cpu().numpy()
    assert ids[3, 0] == 123456 and abs(scores[3, 0] - 1.0) < 5e-3
    assert (np.diff(scores, axis=1) <= 0).all()  # sorted best first
    ref_scores, ref_ids = torch.topk(full, k, dim=1)
    ref_scores = ref_scores.cpu().numpy()
    # k-th best score agrees, and every returned id's true score is within tol of the oracle's
    assert np.abs(scores - ref_scores).max() <= TOL
    got_true = torch.gather(full, 1, out_ids).cpu().numpy()
    assert np.abs(got_true - scores).max() <= TOL
    # idempotence: searching twice gives the same answer
    ix.search_dev(q, B, k, out_ids, out_scores, st)
    torch.cuda.synchronize()
    assert (out_ids.cpu().numpy() == ids).all()
    ix.close()


def test_merge_shards_equals_single_index(native_lib, cuda_device):
    """Row-sharded search + lsc_topk_merge_dev == one index over all rows (the N>1 path, emulated on
    one GPU with two shard handles)."""
    import torch
    from langstream_b200._native import Index, topk_merge_dev

    dim, B, k = 384, 33, 10
    rows, q = _rand(20000, dim, 8), _rand(B, dim, 9)
    whole = Index(dim)
    whole.upsert(np.arange(20000), rows)
    want_ids, want_scores = whole.search(q, k)
    parts = [Index(dim), Index(dim)]
    parts[0].upsert(np.arange(0, 12000), rows[:12000])
    parts[1].upsert(np.arange(12000, 20000), rows[12000:])
    dq = torch.from_numpy(q).to(cuda_device)
    ids = torch.empty(2, B, k, dtype=torch.int64, device=cuda_device)
    sc = torch.empty(2, B, k, dtype=torch.float32, device=cuda_device)
    st = torch.cuda.current_stream().cuda_stream
    for p in range(2):
        parts[p].search_dev(dq, B, k, ids[p], sc[p], st)
    out_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
    out_sc = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
    topk_merge_dev(ids, sc, 2, B, k, out_ids, out_sc, st)
    torch.cuda.synchronize()
    assert (out_ids.cpu().numpy() == want_ids).all()
    assert np.array_equal(out_sc.cpu().numpy(), want_scores)
    for p in parts + [whole]:
        p.close()


def test_bad_arguments_raise(native_lib, cuda_device):
    from langstream_b200._native import Index, NativeError

    with pytest.raises(NativeError):
        Index(100)  # not a multiple of 64
    ix = Index(64)
    with pytest.raises(NativeError):
        ix.search(np.zeros((1, 64), np.float32), 0)
    with pytest.raises(NativeError):
        ix.search(np.zeros((1, 64), np.float32), 129)
    ix.close()
