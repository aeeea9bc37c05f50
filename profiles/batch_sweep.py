"""Throughput of the encoder as a function of the per-pass batch (are intermediates L2-resident?)."""
import sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200 import weights as W
from langstream_b200._native import Encoder
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
spec = W.MINILM_L6
enc = Encoder(spec, W.random_init(spec, 0))
for n in (32, 64, 128, 256, 512, 1024, 2048):
    ids, cu = W.synthetic_token_ids(n, 128, seed=1)
    d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
    out = torch.empty(n, spec.hidden, device=dev)
    reps = max(4, 4096 // n)
    for _ in range(3): enc.embed_dev(d_ids, d_cu, n, int(cu[-1]), 128, out, st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): enc.embed_dev(d_ids, d_cu, n, int(cu[-1]), 128, out, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"n_seq={n:5d}: {ms:8.3f} ms/pass  {n/ms*1000:10.0f} emb/s")
