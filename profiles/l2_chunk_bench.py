"""Does chunking the token dimension keep GEMM outputs L2-resident for their consumer?
QKV GEMM -> attention, and FFN-up -> FFN-down, full batch vs chunks that reuse one small buffer."""
import sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200 import _native
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
H, heads, L, F = 384, 12, 128, 1536
n_seq = 1024; T = n_seq * L
x = torch.randn(T, H, device=dev).to(torch.bfloat16)
Wqkv = (torch.randn(3 * H, H, device=dev) * 0.05).to(torch.bfloat16); bqkv = torch.randn(3 * H, device=dev)
W1 = (torch.randn(F, H, device=dev) * 0.05).to(torch.bfloat16); b1 = torch.randn(F, device=dev)
W2 = (torch.randn(H, F, device=dev) * 0.05).to(torch.bfloat16); b2 = torch.randn(H, device=dev)
ctx = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
out = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
def timeit(fn, iters=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1000
for nchunk in (1, 2, 4, 8, 16):
    cs = n_seq // nchunk; ct = cs * L
    qkv = torch.empty(ct, 3 * H, device=dev, dtype=torch.bfloat16)   # ONE chunk-sized buffer, reused
    ffn = torch.empty(ct, F, device=dev, dtype=torch.bfloat16)
    cu = torch.arange(0, ct + 1, L, device=dev, dtype=torch.int32)
    def qkv_attn():
        for c in range(nchunk):
            _native.debug_gemm_dev(x[c * ct:(c + 1) * ct], Wqkv, bqkv, None, qkv, ct, 3 * H, H, 0, 0, st)
            _native.debug_attention_dev(qkv, cu, cs, ct, H, heads, L, ctx[c * ct:(c + 1) * ct], st)
    def ffn_pair():
        for c in range(nchunk):
            _native.debug_gemm_dev(x[c * ct:(c + 1) * ct], W1, b1, None, ffn, ct, F, H, 1, 0, st)
            _native.debug_gemm_dev(ffn, W2, b2, x[c * ct:(c + 1) * ct], out[c * ct:(c + 1) * ct], ct, H, F, 2, 0, st)
    print(f"chunks={nchunk:2d} ({ct} tokens, qkv {ct*3*H*2/1e6:.0f} MB, ffn {ct*F*2/1e6:.0f} MB): "
          f"qkv+attention {timeit(qkv_attn):7.1f} us   ffn_up+ffn_down {timeit(ffn_pair):7.1f} us")
