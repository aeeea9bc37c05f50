"""Micro-benchmark of the GEMM kernel alone through lsc_debug_gemm_dev (CUDA events, 20 launches).
    python profiles/gemm_bench.py            # encoder shapes, all epilogues
    LSC_GEMM_BN=256 python profiles/gemm_bench.py
epilogue 3 = debug 'discard' (TMA + MMA main loop only)."""
import os, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200 import _native

dev = torch.device("cuda:0")
T = 131072
shapes = [("qkv", T, 1152, 384, 0), ("out", T, 384, 384, 2), ("ffn1", T, 1536, 384, 1), ("ffn2", T, 384, 1536, 2),
          ("out_disc", T, 384, 384, 3), ("out_ldonly", T, 384, 384, 4), ("out_bias", T, 384, 384, 0), ("qkv_disc", T, 1152, 384, 3), ("qkv_ldonly", T, 1152, 384, 4), ("ffn2_ldonly", T, 384, 1536, 4), ("ffn1_disc", T, 1536, 384, 3), ("ffn2_disc", T, 384, 1536, 3),
          ("base_ffn1", 65536, 3072, 768, 1), ("base_ffn2", 65536, 768, 3072, 2), ("base_disc", 65536, 3072, 768, 3)]
st = torch.cuda.current_stream().cuda_stream
print("LSC_GEMM_BN =", os.environ.get("LSC_GEMM_BN"), "DBG =", os.environ.get("LSC_GEMM_DBG"))
only = os.environ.get("ONLY")
if only: shapes = [s for s in shapes if s[0] in only.split(",")]
for name, M, N, K, epi in shapes:
    A = torch.randn(M, K, device=dev).to(torch.bfloat16)
    W = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
    b = torch.randn(N, device=dev)
    R = torch.randn(M, N, device=dev).to(torch.bfloat16)
    C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        _native.debug_gemm_dev(A, W, b, R, C, M, N, K, epi, 0, st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        _native.debug_gemm_dev(A, W, b, R, C, M, N, K, epi, 0, st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"{name:10s} M={M} N={N} K={K} epi={epi}: {ms*1000:8.1f} us  {2*M*N*K/ms/1e9:8.1f} TFLOP/s")
