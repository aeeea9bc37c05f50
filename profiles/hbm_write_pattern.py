"""How fast can HBM absorb writes made of short row segments (the shape of a GEMM epilogue's output)?"""
import torch
dev = torch.device("cuda:0")
M, N = 131072, 1152
out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
src = torch.randn(M, N, device=dev).to(torch.bfloat16)
def t(fn, iters=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters / 1000
nbytes = M * N * 2
s = t(lambda: out.fill_(1.0)); print("fill whole tensor: %.0f GB/s" % (nbytes / s / 1e9))
s = t(lambda: out.copy_(src)); print("copy whole tensor: %.0f GB/s written (+ same read)" % (nbytes / s / 1e9))
for w in (32, 64, 96, 192, 384):
    def f():
        for j in range(0, N, w):
            out[:, j:j + w].fill_(1.0)
    s = t(f, 5); print("fill in column blocks of %4d (%4d B segments, sequential launches): %.0f GB/s" % (w, w * 2, nbytes / s / 1e9))
