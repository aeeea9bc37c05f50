"""Micro-benchmark of the attention kernel alone (lsc_debug_attention_dev, CUDA events)."""
import os, sys
os.environ.setdefault("LSC_DEBUG_ATTN_NO_ORDER", "1")  # the debug entry would allocate + sort per call; the encoder sorts once per pass
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200 import _native
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
for name, n, L, H, heads, ragged in [("minilm 1024x128", 1024, 128, 384, 12, False), ("minilm ragged", 1024, 128, 384, 12, True),
                                     ("base 256x512", 256, 512, 768, 12, False)]:
    lens = np.random.default_rng(0).integers(16, L + 1, size=n) if ragged else np.full(n, L)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T = int(cu[-1])
    qkv = torch.randn(T, 3 * H, device=dev).to(torch.bfloat16)
    qkv[:, 2 * H:] = qkv[:, 2 * H:].to(torch.float16).view(torch.bfloat16)
    ctx = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
    d_cu = torch.from_numpy(cu).to(dev)
    for _ in range(3): _native.debug_attention_dev(qkv, d_cu, n, T, H, heads, int(lens.max()), ctx, st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): _native.debug_attention_dev(qkv, d_cu, n, T, H, heads, int(lens.max()), ctx, st)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1000
    flops = 4.0 * H * float((lens.astype(np.float64) ** 2).sum())
    print(f"{name:18s}: {us:8.1f} us   {flops/us/1e6:7.1f} TFLOP/s   {(T*3*H*2+T*H*2)/us/1e3:7.0f} GB/s")
