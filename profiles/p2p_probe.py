"""Does enabling peer access change the scan time?  One process, GPU 0: time search_dev, connect an exchange to GPU 1, time again."""
import os, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200._native import Index, Exchange
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
ix = Index(384, capacity=rows)
for o in range(0, rows, 1 << 20):
    m = min(1 << 20, rows - o)
    v = torch.randn(m, 384, generator=torch.Generator(device=dev).manual_seed(42 + o), device=dev)
    ix.upsert_dev(np.arange(o, o + m, dtype=np.int64), v, m, st)
ix.reserve(1024, 10)
def t(B, reps=20):
    q = torch.randn(B, 384, generator=torch.Generator(device=dev).manual_seed(43), device=dev)
    ids = torch.empty(B, 10, dtype=torch.int64, device=dev); sc = torch.empty(B, 10, device=dev)
    for _ in range(3): ix.search_dev(q, B, 10, ids, sc, st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): ix.search_dev(q, B, 10, ids, sc, st)
    e1.record(); torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / reps, 4)
print("before peer access:", {B: t(B) for B in (1, 64)}, flush=True)
if torch.cuda.device_count() > 1:
    a, b = Exchange(0, 0, 2, 1024), Exchange(1, 1, 2, 1024)
    Exchange.connect_local([a, b])
    torch.cuda.set_device(0)
    print("after peer access: ", {B: t(B) for B in (1, 64)}, flush=True)
