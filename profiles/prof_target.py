"""Short workloads for ncu (kept tiny: ncu replays every kernel ~40x under --set full).

    python profiles/prof_target.py encoder   # 2 passes of MiniLM-L6 over 1024 x 128 tokens
    python profiles/prof_target.py topk      # 10M x 384 corpus, B=1, 64 and 256 top-10, 2 searches each
"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from langstream_b200 import weights as W  # noqa: E402
from langstream_b200._native import Encoder, Index  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "encoder"
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
if what == "encoder":
    spec = W.MINILM_L6
    enc = Encoder(spec, W.random_init(spec, 0))
    ids, cu = W.synthetic_token_ids(1024, 128, seed=1234)
    d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
    out = torch.empty(1024, spec.hidden, device=dev)
    for _ in range(2):
        enc.embed_dev(d_ids, d_cu, 1024, int(cu[-1]), 128, out, st)
    torch.cuda.synchronize()
    print("encoder ok", float(out.norm(dim=1).mean()))
    # one ragged pass (UniformInt[16,128]): attention tiles of 32-row blocks
    ids, cu = W.synthetic_token_ids(1024, 128, seed=1234, min_len=16)
    d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
    enc.embed_dev(d_ids, d_cu, 1024, int(cu[-1]), int(np.diff(cu).max()), out, st)
    torch.cuda.synchronize()
    print("ragged ok", float(out.norm(dim=1).mean()))
else:
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
    ix = Index(384, capacity=n)
    chunk = 1 << 20
    for o in range(0, n, chunk):
        m = min(chunk, n - o)
        v = torch.randn(m, 384, device=dev, generator=torch.Generator(device=dev).manual_seed(42 + o // chunk))
        ix.upsert_dev(np.arange(o, o + m, dtype=np.int64), v, m, st)
    for B in (1, 64, 256):  # 256: the CTA-pair scan
        q = torch.randn(B, 384, device=dev, generator=torch.Generator(device=dev).manual_seed(43))
        oi = torch.empty(B, 10, dtype=torch.int64, device=dev)
        os_ = torch.empty(B, 10, dtype=torch.float32, device=dev)
        for _ in range(2):
            ix.search_dev(q, B, 10, oi, os_, st)
    torch.cuda.synchronize()
    print("topk ok", oi[0, :3].tolist())
