"""tcgen05 GEMM alone against torch (fp32 reference of the same op) through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ref(A, W, bias, res, epi):
    import torch

    y = A.float() @ W.float().t() + bias
    if epi == 1:
        y = torch.nn.functional.gelu(y)  # exact erf
    if epi == 2:
        y = y + res.float()
    return y


@pytest.mark.parametrize("impl", [1, 0])
@pytest.mark.parametrize("epi", [0, 1, 2])
@pytest.mark.parametrize("M,N,K", [(128, 384, 384), (300, 1152, 384), (1000, 384, 1536), (257, 1536, 384),
                                   (64, 256, 64), (4096, 768, 768), (129, 128, 3072)])
def test_gemm_matches_torch(native_lib, cuda_device, M, N, K, epi, impl):
    import torch
    from langstream_b200 import _native

    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K + epi)
    A = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).to(cuda_device)
    W = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).to(cuda_device)
    bias = torch.randn(N, generator=g).to(cuda_device)
    res = torch.randn(M, N, generator=g).to(torch.bfloat16).to(cuda_device)
    Cout = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=cuda_device)
    stream = torch.cuda.current_stream().cuda_stream
    _native.debug_gemm_dev(A, W, bias, res if epi == 2 else None, Cout, M, N, K, epi, impl, stream)
    torch.cuda.synchronize()
    ref = _ref(A, W, bias, res, epi)
    got = Cout.float()
    assert torch.isfinite(got).all()
    err = (got - ref).abs()
    tol = 0.02 + 0.01 * ref.abs()  # bf16 output rounding (2^-8 relative) + accumulation order
    assert bool((err <= tol).all()), f"max err {err.max().item()} at {np.unravel_index(err.argmax().item(), err.shape)}"


def test_gemm_tc_matches_validation_kernel_bitwise_close(native_lib, cuda_device):
    import torch
    from langstream_b200 import _native

    M, N, K = 777, 1152, 384
    g = torch.Generator(device="cpu").manual_seed(5)
    A = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda_device)
    W = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).to(cuda_device)
    bias = torch.randn(N, generator=g).to(cuda_device)
    c0 = torch.empty((M, N), dtype=torch.bfloat16, device=cuda_device)
    c1 = torch.empty_like(c0)
    st = torch.cuda.current_stream().cuda_stream
    _native.debug_gemm_dev(A, W, bias, None, c0, M, N, K, 0, 0, st)
    _native.debug_gemm_dev(A, W, bias, None, c1, M, N, K, 0, 1, st)
    torch.cuda.synchronize()
    d = (c0.float() - c1.float()).abs()
    # both accumulate in fp32; they may differ by one bf16 ulp where the sum order differs
    assert d.max().item() <= 0.02 + 0.01 * c1.float().abs().max().item()
    assert (d > 0).float().mean().item() < 0.2
