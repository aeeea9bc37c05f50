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


def test_batched_delete_matches_a_host_model(native_lib, cuda_device):
    """lsc_index_delete applies a whole batch with one gather kernel; the surviving table must be what deleting
    one id at a time gives (JdbcWriter.java:139-179 deletes per record): same size, and a search over it
    returns exactly the oracle's answer over the surviving rows - including chains where the row moved into a
    freed slot is itself deleted, or moved again, later in the same call."""
    from langstream_b200._native import Index
    from oracle import topk_oracle

    dim, n = 128, 6000
    rng = np.random.default_rng(5)
    rows = _quantised(n, dim, 11)
    ids = np.arange(n, dtype=np.int64) * 7 + 1
    ix = Index(dim, capacity=n)
    ix.upsert(ids, rows)
    alive = dict(zip(ids.tolist(), range(n)))
    for batch in (np.concatenate([ids[-40:], ids[:25], [999_999_999]]),      # the tail itself + the head + unknown id
                  rng.choice(ids[25:-40], 1500, replace=False),             # a large random batch
                  np.concatenate([ids[3000:3010], ids[3000:3010]])):        # ids repeated inside one call
        ix.delete(batch)
        for i in batch.tolist():
            alive.pop(int(i), None)
        assert len(ix) == len(alive)
        keep = np.fromiter(alive.values(), dtype=np.int64)
        q = _quantised(9, dim, 100 + len(alive))
        got_ids, got_sc = ix.search(q, 10)
        full = topk_oracle.scores_numpy(rows, q)
        masked = np.full_like(full, -np.inf)
        masked[:, keep] = full[:, keep]
        # map returned ids back to original row numbers; the order among exact ties follows storage slots, which
        # deletes permute, so ties are exempt exactly as the parity rule says
        got_rows = np.where(got_ids >= 0, (got_ids - 1) // 7, -1)
        assert all(int(i) in alive for i in got_ids.reshape(-1) if i >= 0)
        topk_oracle.assert_topk_matches(got_rows, got_sc, masked, 10, tol=TOL)
    # re-insert some deleted ids: they append at the end and are found again
    back = ids[:25]
    ix.upsert(back, rows[:25])
    got_ids, got_sc = ix.search(rows[3], 1)
    assert got_ids[0, 0] == ids[3] and abs(got_sc[0, 0] - 1.0) < 1e-2
    ix.close()


def test_merge_orders_duplicate_rows_like_one_index(native_lib, cuda_device):
    """Crawled corpora hold duplicate chunks: equal scores across shards must come back in global storage order
    (shard first, then position), exactly as the unsharded index orders them - also at the k boundary."""
    import torch
    from langstream_b200._native import Index, topk_merge_dev

    dim, B, k, n = 64, 5, 10, 3000
    base = _quantised(40, dim, 21)
    rng = np.random.default_rng(22)
    rows = base[rng.integers(0, 40, n)]          # every row is one of 40 vectors: ~75 copies of each
    q = base[:B]
    whole = Index(dim)
    whole.upsert(np.arange(n), rows)
    want_ids, want_sc = whole.search(q, k)
    cuts = [0, 700, 1500, 2300, n]
    parts = []
    for a, b in zip(cuts[:-1], cuts[1:]):
        p = Index(dim)
        p.upsert(np.arange(a, b), rows[a:b])
        parts.append(p)
    dq = torch.from_numpy(q).to(cuda_device)
    st = torch.cuda.current_stream().cuda_stream
    ids = torch.empty(len(parts), B, k, dtype=torch.int64, device=cuda_device)
    sc = torch.empty(len(parts), B, k, dtype=torch.float32, device=cuda_device)
    for p in range(len(parts)):
        parts[p].search_dev(dq, B, k, ids[p], sc[p], st)
    out_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
    out_sc = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
    topk_merge_dev(ids, sc, len(parts), B, k, out_ids, out_sc, st)
    torch.cuda.synchronize()
    assert (out_ids.cpu().numpy() == want_ids).all()
    assert np.array_equal(out_sc.cpu().numpy(), want_sc)
    for p in parts + [whole]:
        p.close()


def test_searches_on_different_streams_do_not_overlap(native_lib, cuda_device):
    """The index handle's query / candidate scratch is shared by all calls: lsc_index_search_dev on two caller
    streams and the host call on the handle's own stream still each get their own answer."""
    import torch
    from langstream_b200._native import Index

    dim, n, k = 384, 300_000, 10
    g = torch.Generator(device=cuda_device).manual_seed(4)
    rows = torch.randn(n, dim, generator=g, device=cuda_device)
    ix = Index(dim, capacity=n)
    ix.upsert_dev(np.arange(n, dtype=np.int64), rows, n, torch.cuda.current_stream().cuda_stream)
    qs = [torch.randn(B, dim, generator=g, device=cuda_device) for B in (64, 3, 128, 17)]
    want = [ix.search(q.cpu().numpy(), k) for q in qs]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    for rep in range(3):
        outs = []
        for j, q in enumerate(qs):
            B = q.shape[0]
            o_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
            o_sc = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
            ix.search_dev(q, B, k, o_ids, o_sc, streams[j & 1].cuda_stream)
            outs.append((o_ids, o_sc))
        h_ids, h_sc = ix.search(qs[0].cpu().numpy(), k)
        torch.cuda.synchronize()
        assert (h_ids == want[0][0]).all()
        for (o_ids, o_sc), (w_ids, w_sc) in zip(outs, want):
            assert (o_ids.cpu().numpy() == w_ids).all() and np.array_equal(o_sc.cpu().numpy(), w_sc)
    ix.close()


@pytest.fixture(scope="module")
def corpus_10m(native_lib, cuda_device):
    """BASELINE configs[3]'s corpus on one GPU: 10M x 384, randn seed 42 per 1M-row chunk (bench.py's generator)."""
    import torch
    from langstream_b200._native import Index

    n, dim, chunk = 10_000_000, 384, 1 << 20
    ix = Index(dim, capacity=n)
    st = torch.cuda.current_stream().cuda_stream
    for o in range(0, n, chunk):
        m = min(chunk, n - o)
        g = torch.Generator(device=cuda_device).manual_seed(42 + o // chunk)
        vec = torch.randn(m, dim, generator=g, device=cuda_device)
        ix.upsert_dev(np.arange(o, o + m, dtype=np.int64), vec, m, st)
    del vec
    torch.cuda.empty_cache()
    yield ix
    ix.close()
    torch.cuda.empty_cache()


@pytest.mark.parametrize("B", [1, 64, 1024])
def test_config4_full_size_10m(corpus_10m, cuda_device, B):
    """BASELINE configs[3] / SURVEY.md 8(d) config 4 at its own size: 10M x 384, k = 10, B = 1 / 64 / 1024 (the
    last one is the only check that runs topk_scan_rows_kernel<2>, the CTA-pair scan, over a full-length corpus).
    Checker: lsc_index_scores_naive_dev (exact fp32 dot products on CUDA cores over the same bf16 operands) in
    chunks of 16 queries + a plain top-k of those scores.  Rule: ids identical except ties within 1e-5."""
    import json
    import torch

    ix, n, dim, k = corpus_10m, 10_000_000, 384, 10
    st = torch.cuda.current_stream().cuda_stream
    q = torch.randn(B, dim, generator=torch.Generator(device=cuda_device).manual_seed(43), device=cuda_device)
    out_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
    out_sc = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
    ix.search_dev(q, B, k, out_ids, out_sc, st)
    torch.cuda.synchronize()
    full = torch.empty(16, n, dtype=torch.float32, device=cuda_device)
    mismatches, worst_score, worst_rank = 0, 0.0, 0.0
    for b0 in range(0, B, 16):
        nb = min(16, B - b0)
        ix.scores_naive_dev(q[b0:b0 + nb].contiguous(), nb, full, st)
        view = full.view(-1)[: nb * n].view(nb, n)
        ref_sc, ref_ids = torch.topk(view, k, dim=1)
        got_ids, got_sc = out_ids[b0:b0 + nb], out_sc[b0:b0 + nb]
        true_sc = torch.gather(view, 1, got_ids)               # ids are row numbers here
        worst_score = max(worst_score, (true_sc - got_sc).abs().max().item())
        worst_rank = max(worst_rank, (true_sc - ref_sc).abs().max().item())
        mismatches += int((got_ids != ref_ids).sum().item())
        srt = torch.sort(got_ids, dim=1).values
        assert (srt[:, 1:] != srt[:, :-1]).all(), "duplicate ids in a result list"
        assert (got_sc[:, 1:] <= got_sc[:, :-1]).all(), "result list not sorted best first"
    rec = {"case": f"topk_10Mx384_B{B}_k10", "queries": B, "id_mismatches_all_within_tie_tol": mismatches,
           "max_abs_score_err": worst_score, "max_abs_rank_score_gap": worst_rank, "tol": TOL}
    print("PARITY", json.dumps(rec))
    out = Path(__file__).resolve().parents[1] / "gpurun_out"
    try:
        out.mkdir(exist_ok=True)
        with open(out / "parity_margins.jsonl", "a") as f:
            f.write(json.dumps(rec) + "\n")
    except OSError:
        pass
    assert worst_score <= TOL, f"returned scores off by {worst_score}"
    assert worst_rank <= TOL, f"a returned row is not within the tie tolerance of the oracle's row at its rank ({worst_rank})"
