"""Where does the native ingest loop (tokenise -> embed -> upsert) spend its time?  One GPU, 262,144 documents."""
import sys, time
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench
from langstream_b200 import weights as W
from langstream_b200._native import Encoder, Index, WordPieceTokenizer
spec = W.MINILM_L6
enc = Encoder(spec, W.random_init(spec, 0))
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
words = [f"w{i}" for i in range(bench.PIPE_VOCAB_WORDS)]
W_obj = np.array(words, dtype=object)
tok = WordPieceTokenizer("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words) + "\n", max_length=128, n_threads=16)
n_docs, batch = 262144, 8192
texts = [[" ".join(r) for r in W_obj[bench.pipe_doc_words(np.arange(o, o + batch, dtype=np.int64))]] for o in range(0, n_docs, batch)]
t0 = time.perf_counter(); toks = [tok.encode_packed(t) for t in texts]; t_tok = time.perf_counter() - t0
H = spec.hidden
E = torch.empty(n_docs, H, device=dev)
ix = Index(H, capacity=n_docs)
def run(upsert, sync_each):
    o = 0
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for ids, cu in toks:
        m = len(cu) - 1
        d_ids = torch.from_numpy(ids).pin_memory().to(dev, non_blocking=True)
        d_cu = torch.from_numpy(cu).pin_memory().to(dev, non_blocking=True)
        enc.embed_dev(d_ids, d_cu, m, int(cu[-1]), int(np.diff(cu).max()), E[o:o + m], st)
        if upsert: ix.upsert_dev(np.arange(o, o + m, dtype=np.int64), E[o:o + m], m, st)
        elif sync_each: torch.cuda.synchronize()
        o += m
    torch.cuda.synchronize(); return n_docs / (time.perf_counter() - t0)
run(False, False)
print("tokenise alone: %.0f docs/s" % (n_docs / t_tok))
print("embed only, no sync: %.0f docs/s" % run(False, False))
print("embed only, sync per batch: %.0f docs/s" % run(False, True))
print("embed + upsert: %.0f docs/s" % run(True, False))
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ids, cu = toks[0]; d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
ev0.record()
for _ in range(5): enc.embed_dev(d_ids, d_cu, batch, int(cu[-1]), int(np.diff(cu).max()), E[:batch], st)
ev1.record(); torch.cuda.synchronize()
print("GPU time per batch of 8192 docs: %.2f ms -> %.0f docs/s" % (ev0.elapsed_time(ev1) / 5, batch / (ev0.elapsed_time(ev1) / 5e3)))

# the same with the tokeniser in a producer thread, as bench.py runs it: where does the consumer wait?
import queue, threading
for nthreads in (16, 12, 8):
    tok2 = WordPieceTokenizer("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words) + "\n", max_length=128, n_threads=nthreads)
    raw = [[t.encode() for t in tt] for tt in texts]
    qd = queue.Queue(maxsize=2)
    t_tok = [0.0]
    def producer():
        for tt in raw:
            a = time.perf_counter(); r = tok2.encode_packed(tt); t_tok[0] += time.perf_counter() - a
            qd.put(r)
        qd.put(None)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    th = threading.Thread(target=producer, daemon=True); th.start()
    o = 0; w_get = w_enq = w_up = 0.0
    while True:
        a = time.perf_counter(); item = qd.get(); b = time.perf_counter(); w_get += b - a
        if item is None: break
        ids, cu = item; m = len(cu) - 1
        d_ids = torch.from_numpy(ids).pin_memory().to(dev, non_blocking=True)
        d_cu = torch.from_numpy(cu).pin_memory().to(dev, non_blocking=True)
        enc.embed_dev(d_ids, d_cu, m, int(cu[-1]), int(np.diff(cu).max()), E[o:o + m], st)
        c = time.perf_counter(); w_enq += c - b
        ix.upsert_dev(np.arange(o, o + m, dtype=np.int64), E[o:o + m], m, st)
        w_up += time.perf_counter() - c
        o += m
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"{nthreads} tokenizer threads: {n_docs / dt:.0f} docs/s; consumer waited for tokens {w_get:.2f} s, enqueue {w_enq:.2f} s, upsert (sync) {w_up:.2f} s of {dt:.2f} s; producer inside encode_packed {t_tok[0]:.2f} s")
