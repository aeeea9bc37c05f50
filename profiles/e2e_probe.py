"""Where does the end-to-end loop lose time against the device-resident loop?  Same process, 20 / 200 steps."""
import sys, time
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200 import weights as W
from langstream_b200._native import Encoder
spec = W.MINILM_L6
enc = Encoder(spec, W.random_init(spec, 0))
ids, cu = W.synthetic_token_ids(1024, 128, seed=1234)
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
out = torch.empty(1024, spec.hidden, device=dev)
def dev_loop(n):
    for _ in range(5): enc.embed_dev(d_ids, d_cu, 1024, int(cu[-1]), 128, out, st)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): enc.embed_dev(d_ids, d_cu, 1024, int(cu[-1]), 128, out, st)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
bufs = []
for _ in range(2):
    bufs.append((torch.from_numpy(ids).pin_memory().numpy(), torch.from_numpy(cu).pin_memory().numpy(), torch.empty(1024, spec.hidden).pin_memory().numpy()))
def e2e_loop(n):
    for w in range(4): enc.wait(enc.submit(*bufs[w & 1]))
    t0 = time.perf_counter(); sub = 0.0
    pending = enc.submit(*bufs[0])
    for i in range(1, n):
        a = time.perf_counter(); nxt = enc.submit(*bufs[i & 1]); sub += time.perf_counter() - a
        enc.wait(pending); pending = nxt
    enc.wait(pending)
    return (time.perf_counter() - t0) / n * 1e3, sub / max(n - 1, 1) * 1e3
for n in (20, 200):
    print(n, "steps: device loop %.3f ms/step" % dev_loop(n), "e2e loop %.3f ms/step (host time inside submit %.3f ms)" % e2e_loop(n), flush=True)
