"""Packed variable-length attention kernel alone against an fp32 torch reference of the same op."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _reference(qkv, cu, H, heads):
    import torch

    d = H // heads
    out = torch.zeros(qkv.shape[0], H, dtype=torch.float32, device=qkv.device)
    x = qkv.float()
    for i in range(len(cu) - 1):
        a, b = int(cu[i]), int(cu[i + 1])
        if b == a:
            continue
        q = x[a:b, :H].view(b - a, heads, d).transpose(0, 1)
        k = x[a:b, H:2 * H].view(b - a, heads, d).transpose(0, 1)
        v = x[a:b, 2 * H:].view(b - a, heads, d).transpose(0, 1)
        p = torch.softmax(q @ k.transpose(1, 2) / d ** 0.5, dim=-1)
        out[a:b] = (p @ v).transpose(0, 1).reshape(b - a, H)
    return out


@pytest.mark.parametrize("H,heads,lens,scale", [
    (384, 12, [128] * 6, 1.0),
    (384, 12, [1, 7, 16, 63, 64, 65, 100, 127, 128, 0, 33], 1.0),
    (384, 12, [128, 96, 50], 6.0),          # peaked softmax (large logits)
    (768, 12, [512, 300, 64, 1, 257], 1.0),  # d_head 64, several key blocks with online rescaling
    (768, 12, [512, 511], 4.0),
    (128, 4, [40, 5], 1.0),
    # d_head 64 on the tcgen05 kernel: more items than SMs, every residue of the 128-row / 64-key tiling, rounds
    # with 1-4 active query tiles, sequences longer than 512 (two rounds of query tiles), empty sequences
    (768, 12, [1, 63, 64, 65, 127, 128, 129, 191, 192, 193, 255, 256, 257, 383, 384, 385, 0, 511, 513, 640, 700], 1.0),
    (256, 4, [33, 600, 2, 130], 3.0),
])
def test_attention_matches_fp32_reference(native_lib, cuda_device, H, heads, lens, scale):
    import torch
    from langstream_b200 import _native

    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T = int(cu[-1])
    g = torch.Generator(device="cpu").manual_seed(len(lens) * 1000 + H)
    qkv = (torch.randn(T, 3 * H, generator=g) * scale).to(torch.bfloat16).to(cuda_device)
    # the kernel's input layout: Q and K thirds bf16, V third fp16 (what the QKV GEMM epilogue writes)
    packed = qkv.clone()
    packed[:, 2 * H:] = qkv[:, 2 * H:].to(torch.float16).view(torch.bfloat16)
    ctx = torch.full((T, H), float("nan"), dtype=torch.bfloat16, device=cuda_device)
    d_cu = torch.from_numpy(cu).to(cuda_device)
    _native.debug_attention_dev(packed, d_cu, len(lens), T, H, heads, max(lens), ctx,
                                torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ref = _reference(qkv, cu, H, heads)
    got = ctx.float()
    assert torch.isfinite(got).all()
    err = (got - ref).abs()
    # bf16 output rounding (2^-8 relative) + fp16 probabilities (2^-11) on values of magnitude <= |V|max
    tol = 0.01 * scale + 0.01 * ref.abs()
    assert bool((err <= tol).all()), f"max err {err.max().item():.4f} (ref magnitude {ref.abs().max().item():.2f})"
    assert err.mean().item() < 2e-3 * scale


@pytest.mark.parametrize("lens", [
    [128] * 3 + [1] * 200 + [64, 64, 65, 63, 127, 1, 128, 2, 126],       # full tiles, 128 one-token sequences per tile, exact fits
    [17, 40, 9, 0, 0, 61, 1, 128, 3, 3, 3, 100, 28, 29, 70, 58, 0],      # empty sequences inside and at the end of tiles
    list(np.random.default_rng(5).integers(1, 129, 400)),                # UniformInt[1,128], more tiles than SMs x heads / 12
])
def test_packed_tiles_match_fp32_reference(native_lib, cuda_device, lens, monkeypatch):
    """Short sequences sharing a 128-row attention tile (block-diagonal mask): every sequence still attends only to
    itself - same fp32 reference, same tolerance as the one-tile-per-sequence path."""
    import torch
    from langstream_b200 import _native

    monkeypatch.setenv("LSC_DEBUG_ATTN_PACK", "1")
    H, heads = 384, 12
    lens = [int(x) for x in lens]
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T = int(cu[-1])
    g = torch.Generator(device="cpu").manual_seed(len(lens))
    qkv = torch.randn(T, 3 * H, generator=g).to(torch.bfloat16).to(cuda_device)
    packed = qkv.clone()
    packed[:, 2 * H:] = qkv[:, 2 * H:].to(torch.float16).view(torch.bfloat16)
    ctx = torch.full((T, H), float("nan"), dtype=torch.bfloat16, device=cuda_device)
    d_cu = torch.from_numpy(cu).to(cuda_device)
    _native.debug_attention_dev(packed, d_cu, len(lens), T, H, heads, max(lens), ctx, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ref = _reference(qkv, cu, H, heads)
    got = ctx.float()
    assert torch.isfinite(got).all()
    err = (got - ref).abs()
    assert bool((err <= 0.01 + 0.01 * ref.abs()).all()), f"max err {err.max().item():.4f}"
    assert err.mean().item() < 2e-3


@pytest.mark.parametrize("form", ["blocks", "one_tile_per_sequence"])
def test_ragged_kernel_every_length(native_lib, cuda_device, form, monkeypatch):
    """The length-adaptive kernels - tiles of 32-row blocks of sequences from anywhere in the batch
    (attention_tc_blocks.cu, the product path) and one adaptive tile per sequence (attention_tc_ragged.cu): every
    length 1..128 several times, in random order, empty sequences between them, many more items than SMs x groups -
    so that short items follow long ones in the same shared-memory stage and TMEM slot (stale rows behind the loaded
    part must never reach a stored row) and every combination 4 | 3+1 | 2+2 | 2+1+1 | 1+1+1+1 of blocks occurs."""
    import torch
    from langstream_b200 import _native

    if form != "blocks":
        monkeypatch.setenv("LSC_ATTN_NO_BLOCKS", "1")

    H, heads = 384, 12
    rng = np.random.default_rng(11)
    lens = list(range(1, 129)) * 3 + [0] * 7
    rng.shuffle(lens)
    lens = [int(x) for x in lens]
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T = int(cu[-1])
    g = torch.Generator(device="cpu").manual_seed(77)
    qkv = (torch.randn(T, 3 * H, generator=g) * 2.0).to(torch.bfloat16).to(cuda_device)
    packed = qkv.clone()
    packed[:, 2 * H:] = qkv[:, 2 * H:].to(torch.float16).view(torch.bfloat16)
    d_cu = torch.from_numpy(cu).to(cuda_device)
    ref = _reference(qkv, cu, H, heads)
    for _ in range(2):  # the second run starts from the first run's leftovers in shared memory / TMEM
        ctx = torch.full((T, H), float("nan"), dtype=torch.bfloat16, device=cuda_device)
        _native.debug_attention_dev(packed, d_cu, len(lens), T, H, heads, max(lens), ctx, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        got = ctx.float()
        assert torch.isfinite(got).all()
        err = (got - ref).abs()
        assert bool((err <= 0.02 + 0.01 * ref.abs()).all()), f"max err {err.max().item():.4f}"
        assert err.mean().item() < 4e-3


@pytest.mark.parametrize("case", ["uniform", "ones", "mixed_with_empty", "many"])
def test_block_tile_list(native_lib, cuda_device, case):
    """The tile list the block-tile attention kernel walks: every non-empty sequence exactly once, its blocks consecutive
    inside one tile, used blocks a prefix, the per-block fields consistent, and as few tiles as the 4 | 3+1 | 2+2 | 2+1+1 |
    1+1+1+1 scheme promises."""
    import torch
    from langstream_b200 import _native

    rng = np.random.default_rng(3)
    lens = {"uniform": rng.integers(16, 129, 1024), "ones": rng.integers(1, 33, 777),
            "mixed_with_empty": np.array([0, 128, 1, 0, 33, 64, 65, 96, 97, 32, 0, 5, 100, 2, 70, 0]),
            "many": rng.integers(0, 129, 20000)}[case].astype(np.int64)
    n = len(lens)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    d_cu = torch.from_numpy(cu).to(cuda_device)
    d_out = torch.full((8 + 9 * n,), -99, dtype=torch.int32, device=cuda_device)
    _native.check(_native.lib().lsc_debug_attention_blocks_dev(d_cu.data_ptr(), n, d_out.data_ptr(),
                                                               torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    out = d_out.cpu().numpy()
    n_tiles = int(out[0])
    rec = out[8:8 + 8 * n_tiles].reshape(n_tiles, 8)
    np_of = (lens + 31) // 32
    c = [int((np_of == k).sum()) for k in range(5)]
    m31 = min(c[3], c[1])
    odd2 = c[2] & 1
    with_odd = min(2, c[1] - m31) if odd2 else 0
    assert n_tiles == c[4] + c[3] + c[2] // 2 + odd2 + (c[1] - m31 - with_odd + 3) // 4
    start_of = {int(cu[i]): i for i in range(n) if lens[i] > 0}
    seen = set()
    for t in range(n_tiles):
        row0, info = rec[t, :4], rec[t, 4:].astype(np.uint32)
        nb = int((row0 >= 0).sum())
        assert nb >= 1 and (row0[:nb] >= 0).all() and (row0[nb:] < 0).all()      # used blocks are a prefix
        assert all(int(x >> 20) & 15 == nb for x in info)
        b = 0
        while b < nb:
            seq = start_of[int(row0[b])]                                         # a sequence starts here
            s, npq = int(lens[seq]), int(np_of[seq])
            assert seq not in seen and b + npq <= nb
            seen.add(seq)
            for blk in range(npq):
                x = int(info[b + blk])
                assert int(row0[b + blk]) == int(cu[seq]) + 32 * blk
                assert (x & 15, (x >> 4) & 15, (x >> 8) & 255, (x >> 16) & 15) == (b, npq, s, blk)
            b += npq
    assert seen == set(start_of.values())
