"""Pure-read / pure-write / copy HBM bandwidth with torch ops (CUDA events), to calibrate the
write-heavy GEMM epilogues."""
import torch
dev = torch.device("cuda:0")
n = 1 << 29  # 512 Mi bf16 = 1 GiB
a = torch.empty(n, dtype=torch.bfloat16, device=dev).normal_()
b = torch.empty_like(a)
def t(fn, iters=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters / 1000
s = t(lambda: b.fill_(1.0)); print("pure write (fill_): %.0f GB/s" % (n * 2 / s / 1e9))
s = t(lambda: b.zero_()); print("pure write (zero_ / memset): %.0f GB/s" % (n * 2 / s / 1e9))
s = t(lambda: b.copy_(a)); print("copy: %.0f GB/s (read+write)" % (2 * n * 2 / s / 1e9))
s = t(lambda: a.sum()); print("pure read (sum): %.0f GB/s" % (n * 2 / s / 1e9))
af = a.view(torch.int16)
s = t(lambda: torch.add(a, a, out=b)); print("read 1x + write 1x (add a,a): %.0f GB/s" % (2 * n * 2 / s / 1e9))
