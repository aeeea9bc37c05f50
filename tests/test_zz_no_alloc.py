"""lsc_topk_search_static_dev: the top-k entry point for callers that cannot allocate device memory
(their GPU is fully reserved by a co-tenant framework, or they run on a latency budget that forbids
cudaMalloc's device-wide synchronisation).  Its scratch is a fixed array inside the library that all
calls share and that keeps stale contents, so the check is: two different searches in a row through
the C ABI, each compared with the regular entry point (lsc_index_search_dev) on the same rows.
Collected last on purpose (file name), and self-contained so it also passes alone in a fresh process.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5  # the score tolerance the other top-k tests use


def _unit_bf16_rows(n, dim, seed):
    from oracle import topk_oracle

    x = np.random.default_rng(seed).standard_normal((n, dim), dtype=np.float32)
    return topk_oracle.to_bf16_rounded(topk_oracle.normalize_rows(x))


def test_static_scratch_search_matches_regular_search_twice(native_lib, cuda_device):
    import torch
    from langstream_b200._native import Index, NativeError, topk_search_static_dev, LSC_E_UNSUPPORTED

    dim, n = 384, 70_001
    rows = _unit_bf16_rows(n, dim, 2024)       # exactly representable: both paths hold identical bf16 rows
    ids = np.arange(n, dtype=np.int64) * 5 + 3  # ids differ from row numbers
    index = Index(dim, capacity=n)
    index.upsert(ids, rows)
    d_rows = torch.from_numpy(rows).to(cuda_device).to(torch.bfloat16).contiguous()
    d_ids = torch.from_numpy(ids).to(cuda_device)
    st = torch.cuda.current_stream().cuda_stream

    # first call: a full query block with the largest k; second call: fewer queries and a smaller k,
    # so anything left in the shared scratch by the first call would show up as a wrong answer
    for call, (B, k, seed) in enumerate([(128, 32, 7), (5, 10, 8)]):
        q = torch.from_numpy(_unit_bf16_rows(B, dim, seed)).to(cuda_device)
        want_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
        want_sc = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
        index.search_dev(q, B, k, want_ids, want_sc, st)
        got_ids = torch.full((B, k), -7, dtype=torch.int64, device=cuda_device)
        got_sc = torch.full((B, k), float("nan"), dtype=torch.float32, device=cuda_device)
        topk_search_static_dev(d_rows, d_ids, n, dim, q, B, k, got_ids, got_sc, st)
        torch.cuda.synchronize()
        assert torch.equal(got_ids, want_ids), f"call {call}: ids differ"
        assert (got_sc - want_sc).abs().max().item() <= TOL, f"call {call}: scores differ"

    # NULL id table: row numbers come back
    B, k = 3, 4
    q = torch.from_numpy(_unit_bf16_rows(B, dim, 9)).to(cuda_device)
    want_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
    want_sc = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
    index.search_dev(q, B, k, want_ids, want_sc, st)
    got_ids = torch.empty(B, k, dtype=torch.int64, device=cuda_device)
    got_sc = torch.empty(B, k, dtype=torch.float32, device=cuda_device)
    topk_search_static_dev(d_rows, None, n, dim, q, B, k, got_ids, got_sc, st)
    torch.cuda.synchronize()
    assert torch.equal(got_ids * 5 + 3, want_ids)

    # sizes beyond the fixed scratch are refused with a clear error, not truncated
    for bad_B, bad_k in [(129, 10), (1, 33)]:
        with pytest.raises(NativeError) as ei:
            topk_search_static_dev(d_rows, d_ids, n, dim, q, bad_B, bad_k, got_ids, got_sc, st)
        assert ei.value.code == LSC_E_UNSUPPORTED and "fixed scratch" in ei.value.message
    index.close()
