"""lsc_topk_search_static_dev: the top-k entry point for callers that cannot allocate device memory
(their GPU is fully reserved by a co-tenant framework, or they run on a latency budget that forbids
cudaMalloc's device-wide synchronisation).  Its scratch is a fixed array inside the library that all
calls share and that keeps stale contents, so the check is: two different searches in a row through
the C ABI, each compared with the regular entry point (lsc_index_search_dev) on the same rows and with
the CPU oracle, with the driver's free-memory figure unchanged across the calls (the one promise this
entry point makes on top of the regular search).
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

    from oracle import topk_oracle

    # first call: a full query block with the largest k; second call: fewer queries and a smaller k,
    # so anything left in the shared scratch by the first call would show up as a wrong answer.
    # Every tensor the loop needs exists before the first call (torch's own allocator must not move the
    # free-memory figure), and call 0 doubles as the documented set-up call of the device.
    cases = [(128, 32, 7), (5, 10, 8)]
    bufs = []
    for B, k, seed in cases:
        q_host = _unit_bf16_rows(B, dim, seed)
        bufs.append((q_host, torch.from_numpy(q_host).to(cuda_device),
                     torch.empty(B, k, dtype=torch.int64, device=cuda_device),
                     torch.empty(B, k, dtype=torch.float32, device=cuda_device),
                     torch.full((B, k), -7, dtype=torch.int64, device=cuda_device),
                     torch.full((B, k), float("nan"), dtype=torch.float32, device=cuda_device)))
    for (B, k, _), (_, q, want_ids, want_sc, _, _) in zip(cases, bufs):
        index.search_dev(q, B, k, want_ids, want_sc, st)  # the regular entry point may allocate: do it up front
    torch.cuda.synchronize()
    free_before = None
    for call, ((B, k, _), (q_host, q, want_ids, want_sc, got_ids, got_sc)) in enumerate(zip(cases, bufs)):
        topk_search_static_dev(d_rows, d_ids, n, dim, q, B, k, got_ids, got_sc, st)
        torch.cuda.synchronize()
        free_now = torch.cuda.mem_get_info()[0]
        if free_before is None:
            free_before = free_now  # after the set-up call
        assert free_now == free_before, f"call {call}: free device memory moved by {free_before - free_now} bytes"
        assert torch.equal(got_ids, want_ids), f"call {call}: ids differ"
        assert (got_sc - want_sc).abs().max().item() <= TOL, f"call {call}: scores differ"
        # and directly against the CPU oracle (ids are 5 * row + 3)
        slot_of = (got_ids.cpu().numpy() - 3) // 5
        topk_oracle.assert_topk_matches(slot_of, got_sc.cpu().numpy(), topk_oracle.scores_numpy(rows, q_host), k, tol=TOL)
    # a third and fourth call change nothing either
    for (B, k, _), (_, q, _, _, got_ids, got_sc) in zip(cases, bufs):
        topk_search_static_dev(d_rows, d_ids, n, dim, q, B, k, got_ids, got_sc, st)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before

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


@pytest.mark.parametrize("n", [1, 100, 300])
def test_static_scratch_search_on_tiny_tables(native_lib, cuda_device, n):
    """Fewer rows than one scan tile (the regular index never sees that: its slab starts at 1024 rows) and
    fewer rows than k: the tail of every result list is the empty marker (-1, -inf), as the oracle says."""
    import torch
    from langstream_b200._native import topk_search_static_dev
    from oracle import topk_oracle

    dim = 128
    rows = _unit_bf16_rows(n, dim, 31 + n)
    d_rows = torch.from_numpy(rows).to(cuda_device).to(torch.bfloat16).contiguous()
    st = torch.cuda.current_stream().cuda_stream
    for B, k in [(3, 16), (128, 32)]:
        q_host = _unit_bf16_rows(B, dim, 77 + B)
        q = torch.from_numpy(q_host).to(cuda_device)
        got_ids = torch.full((B, k), -7, dtype=torch.int64, device=cuda_device)
        got_sc = torch.full((B, k), float("nan"), dtype=torch.float32, device=cuda_device)
        topk_search_static_dev(d_rows, None, n, dim, q, B, k, got_ids, got_sc, st)
        torch.cuda.synchronize()
        topk_oracle.assert_topk_matches(got_ids.cpu().numpy(), got_sc.cpu().numpy(),
                                        topk_oracle.scores_numpy(rows, q_host), k, tol=TOL)
        if n < k:
            assert (got_ids[:, n:] == -1).all()
