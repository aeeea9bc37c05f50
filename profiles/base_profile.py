"""Per-kernel-class time of the base shape (12L/768/12h/3072), 512 sequences x 512 tokens (SURVEY config 3 pass)."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200 import weights as W
from langstream_b200._native import Encoder
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
spec = W.BERT_BASE
enc = Encoder(spec, W.random_init(spec, 0))
n, s = 512, 512
ids, cu = W.synthetic_token_ids(n, s, seed=1)
d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
out = torch.empty(n, spec.hidden, device=dev)
for _ in range(2): enc.embed_dev(d_ids, d_cu, n, int(cu[-1]), s, out, st)
torch.cuda.synchronize()
enc.set_profiling(True)
for _ in range(3): enc.embed_dev(d_ids, d_cu, n, int(cu[-1]), s, out, st)
prof = enc.read_profile()
tot = sum(v[0] for v in prof.values())
for k, (ms, cnt) in prof.items():
    if cnt: print(f"{k:10s} {ms/3:8.2f} ms/pass  {100*ms/tot:5.1f}%  {cnt//3} launches")
print("total", tot / 3)
