"""Micro-benchmark of the top-k search alone (one GPU): ms per batch for several batch sizes, run once per
combination of the debug switches LSC_TOPK_UNFUSED / LSC_TOPK_NO_SHARED_THRESHOLDS (each in its own process).
    python profiles/topk_bench.py [rows] [reps]"""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def child(rows, reps):
    sys.path.insert(0, str(ROOT))
    import numpy as np
    import torch
    from langstream_b200._native import Index

    dev = torch.device("cuda:0")
    ix = Index(384, capacity=rows)
    st = torch.cuda.current_stream().cuda_stream
    id0 = int(os.environ.get("TOPK_ID_OFFSET", "0"))  # a shard's ids do not start at 0
    for o in range(0, rows, 1 << 20):
        m = min(1 << 20, rows - o)
        v = torch.randn(m, 384, generator=torch.Generator(device=dev).manual_seed(42 + o + id0), device=dev)
        ix.upsert_dev(np.arange(id0 + o, id0 + o + m, dtype=np.int64), v, m, st)
    if os.environ.get("TOPK_RESERVE"):
        ix.reserve(int(os.environ["TOPK_RESERVE"]), 10)
    out = {}
    for B in [int(x) for x in os.environ.get("TOPK_BATCHES", "1,16,64,128,256,1024").split(",")]:
        q = torch.randn(B, 384, generator=torch.Generator(device=dev).manual_seed(43), device=dev)
        ids = torch.empty(B, 10, dtype=torch.int64, device=dev)
        sc = torch.empty(B, 10, dtype=torch.float32, device=dev)
        for _ in range(3):
            ix.search_dev(q, B, 10, ids, sc, st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            ix.search_dev(q, B, 10, ids, sc, st)
        e1.record()
        torch.cuda.synchronize()
        out[B] = round(e0.elapsed_time(e1) / reps, 4)
    print(json.dumps(out))


if __name__ == "__main__":
    if os.environ.get("TOPK_BENCH_CHILD"):
        child(int(sys.argv[1]), int(sys.argv[2]))
        sys.exit(0)
    rows = sys.argv[1] if len(sys.argv) > 1 else "10000000"
    reps = sys.argv[2] if len(sys.argv) > 2 else "10"
    for unfused in ((False,) if os.environ.get("TOPK_ONE") else (False, True)):
        for no_thr in ((False,) if os.environ.get("TOPK_ONE") else (False, True)):
            env = dict(os.environ, TOPK_BENCH_CHILD="1")
            if unfused:
                env["LSC_TOPK_UNFUSED"] = "1"
            if no_thr:
                env["LSC_TOPK_NO_SHARED_THRESHOLDS"] = "1"
            r = subprocess.run([sys.executable, __file__, rows, reps], env=env, capture_output=True, text=True)
            print(f"unfused={int(unfused)} no_shared_thresholds={int(no_thr)} rows={rows}: ms/batch",
                  r.stdout.strip() or r.stderr[-500:], flush=True)
