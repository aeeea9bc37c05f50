"""Per-kernel-class time of one MiniLM pass over a ragged batch (UniformInt[16,128], 1024 sequences) and over the
fixed-length batch, with the encoder's event profiling.  LSC_ATTN_NO_PACK=1 switches the attention tile packing off."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from langstream_b200 import weights as W
from langstream_b200._native import Encoder

dev = torch.device("cuda:0")
spec = W.MINILM_L6
enc = Encoder(spec, W.random_init(spec, 0))
st = torch.cuda.current_stream().cuda_stream
for name, min_len in (("fixed128", None), ("ragged16-128", 16), ("ragged1-64", 1)):
    ids, cu = W.synthetic_token_ids(1024, 128 if name != "ragged1-64" else 64, seed=4321, min_len=min_len)
    d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
    out = torch.empty(1024, spec.hidden, device=dev)
    T, mx = int(cu[-1]), int(np.diff(cu).max())
    for _ in range(3):
        enc.embed_dev(d_ids, d_cu, 1024, T, mx, out, st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        enc.embed_dev(d_ids, d_cu, 1024, T, mx, out, st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    enc.set_profiling(True)
    for _ in range(5):
        enc.embed_dev(d_ids, d_cu, 1024, T, mx, out, st)
    pr = enc.read_profile()
    enc.set_profiling(False)
    print(name, "tokens", T, "ms/pass %.3f" % ms, "Mtok/s %.2f" % (T / ms / 1e3),
          {k: round(v[0] / 5, 3) for k, v in pr.items() if v[1]})
