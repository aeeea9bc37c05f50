#!/usr/bin/env python
"""Benchmark of the hot path (BASELINE.json): MiniLM-L6 embeddings/s at 1024 x 128 tokens per GPU,
plus brute-force top-10 queries/s over a 10M x 384 bf16 corpus.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle port)

One "step" = one pass of the encoder over one batch of 1024 synthetic 128-token sequences per GPU
(weak scaling: every rank embeds its own batch, no collective on the data path).  The timed region
is bracketed by barrier + synchronize, timed with CUDA events on the launching stream, MAX over
ranks.  `value` has the token ids resident in HBM; `e2e` goes through the C-ABI host-buffer call
(lsc_encoder_embed: H2D of ids, D2H of embeddings inside the timed region).
Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_SEQ = 1024          # sequences per GPU per step (BASELINE configs[1])
SEQ_LEN = 128
TOPK_ROWS = 10_000_000
TOPK_DIM = 384
TOPK_K = 10
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


_REAL_STDOUT = None


def capture_stdout():
    """Rank 0 must print exactly ONE line on stdout, but libraries write there too (NCCL prints its
    version banner from C).  Point fd 1 at stderr for the run and keep the real stdout for emit()."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line: dict):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            d = json.loads(p.read_text())
            return {k: float(d[k]) for k in FALLBACK_PEAKS}, "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return dict(FALLBACK_PEAKS), "fallback (B200_PROFILING.md)"


def flops_per_sequence(spec, s):
    """L * (24 H^2 s + 4 H s^2); dense part only = L * 24 H^2 s."""
    H, L = spec.hidden, spec.layers
    return L * (24 * H * H * s + 4 * H * s * s), L * 24 * H * H * s


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.nv = None

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------
# reference arm: the reference's CPU path (oracle port), all host threads
# ---------------------------------------------------------------------------------------------
def cpu_reference_rate(n_seq_sample: int, min_seconds: float, max_seconds: float, warmup: int = 1):
    """embeddings/s of the reference-shaped CPU path: HuggingFace BertModel fp32 on torch CPU, ONE
    text per forward (AbstractHuggingFaceEmbeddingService.java:191-206), mean-pool + L2."""
    import torch
    from langstream_b200 import weights as W
    from oracle.encoder_oracle import HFOracle

    spec = W.MINILM_L6
    oracle = HFOracle(spec, W.random_init(spec, seed=0))
    ids, cu = W.synthetic_token_ids(n_seq_sample, SEQ_LEN, seed=1234)
    for _ in range(warmup):
        oracle.embed_one_by_one(ids[: cu[4]], cu[:5])
    done, t0 = 0, time.perf_counter()
    while True:
        oracle.embed_one_by_one(ids, cu)
        done += n_seq_sample
        el = time.perf_counter() - t0
        if el >= min_seconds or el >= max_seconds:
            break
    return done / el, torch.get_num_threads(), done, el


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    from langstream_b200 import weights as W
    from oracle.encoder_oracle import HFOracle

    # torchrun exports OMP_NUM_THREADS=1 to its workers; the reference arm is one process and gets the whole host
    torch.set_num_threads(os.cpu_count() or 1)
    spec = W.MINILM_L6
    sample = 32  # sequences per step: BASELINE configs[0] (batch 32 x 128 on CPU)
    oracle = HFOracle(spec, W.random_init(spec, seed=0))
    ids, cu = W.synthetic_token_ids(sample, SEQ_LEN, seed=1234)
    for _ in range(max(args.warmup, 1)):
        oracle.embed_one_by_one(ids, cu)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.embed_one_by_one(ids, cu)
    el = time.perf_counter() - t0
    value = sample * args.steps / el
    cores = torch.get_num_threads()
    line = {
        "impl": "reference", "metric": "embeddings/s", "value": value, "unit": "embeddings/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * el / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "MiniLM-L6 (6L,384d,12h,ffn1536) random-init, 128-token sequences, "
                               f"reference shape: one text per forward; each step = {sample} sequences "
                               "(bounded sample of the 1024x128 batch)"},
        "cpu_baseline": {"value": value, "unit": "embeddings/s", "cores": cores, "kind": "port",
                         "sample": f"{sample} sequences x {SEQ_LEN} tokens per step, {args.steps} steps, "
                                   "HuggingFace BertModel fp32 on torch CPU, one text per forward"},
        "e2e": {"value": value, "unit": "embeddings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------
# top-k half of the metric
# ---------------------------------------------------------------------------------------------
def topk_cpu_baseline(max_seconds: float = 12.0):
    """BASELINE.md 3, rows B4 and B5, on the box's host cores.
    B4: fp32 `queries @ corpus.T` + top-k over a 1M x 384 slice (L2-normalised, bf16-rounded, widened) for
        B = 1 / 64 / 1024, QPS at 1M rows and EXTRAPOLATED x10 (a scan is linear in the rows) to 10M.
    B5: the scalar per-row cosine of ReRankAgent.cosineSimilarity (ReRankAgent.java:254-273,361-372) + stable
        top-10 on 10k rows, one thread: a per-row cost figure, not extrapolated."""
    import torch
    from oracle import topk_oracle

    torch.set_num_threads(os.cpu_count() or 1)
    n, dim = 1_000_000, TOPK_DIM
    g = torch.Generator().manual_seed(42)
    corpus = torch.nn.functional.normalize(torch.randn(n, dim, generator=g), dim=1).bfloat16().float()
    out = {"kind": "port", "cores": torch.get_num_threads(), "rows": n,
           "what": "B4: torch CPU fp32 queries @ corpus.T + topk on a 1M-row slice, x10-extrapolated to 10M rows; "
                   "B5: scalar ReRankAgent-form cosine + stable top-10 on 10k rows (1 thread)", "batches": {}}
    budget = max_seconds / 3.0
    for B in (1, 64, 1024):
        q = torch.nn.functional.normalize(torch.randn(B, dim, generator=torch.Generator().manual_seed(43)), dim=1).bfloat16().float()
        torch.topk(q @ corpus.T, TOPK_K, dim=1)
        done, t0 = 0, time.perf_counter()
        while True:
            torch.topk(q @ corpus.T, TOPK_K, dim=1)
            done += 1
            el = time.perf_counter() - t0
            if el >= budget or done >= 50:
                break
        qps_1m = B * done / el
        out["batches"][str(B)] = {"qps_at_1M_rows": qps_1m, "qps_at_10M_rows_extrapolated": qps_1m / 10.0,
                                  "sample": f"{done} batches in {el:.2f} s"}
    rows10k = corpus[:10_000].numpy()
    q1 = corpus[123:124].numpy()
    topk_oracle.search_scalar(rows10k[:100], q1, TOPK_K)
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < 1.0:
        topk_oracle.search_scalar(rows10k, q1, TOPK_K)
        reps += 1
    el = time.perf_counter() - t0
    out["scalar_form"] = {"rows": 10_000, "ns_per_row": 1e9 * el / (reps * 10_000), "queries_per_s_at_10k_rows": reps / el,
                          "cores": 1}
    return out


class _LazyRows:
    """Host row store stand-in for the batching figure: row i is {"id": i} (a real table keeps every column)."""

    def __getitem__(self, i):
        return {"id": int(i)}

    def __len__(self):
        return 0


def query_batching_figure(dev, stream, n_rows: int = 1_000_000):
    """SURVEY.md 8(f) row 4 through the datasource mirror (`CudaVectorDataSource.fetch_data`, the call QueryStep
    makes, QueryStep.java:174-198) on a 1M x 384 table: one query per call the way the reference's runtime issues
    them (SingleRecordAgentProcessor.java:31-47), against 64 concurrent callers and 1024 requests in flight that the
    cross-caller micro-batcher answers with one search per <=128 queries.  Host-side Python is part of the figure."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from langstream_b200 import vector

    name = "bench_docs"
    table = vector.get_table(name, TOPK_DIM, "vector", dev.index or 0)
    for o in range(0, n_rows, 1 << 20):
        m = min(1 << 20, n_rows - o)
        v = torch.randn(m, TOPK_DIM, generator=torch.Generator(device=dev).manual_seed(7 + o), device=dev)
        table.index.upsert_dev(np.arange(o, o + m, dtype=np.int64), v, m, stream)
    table.rows = _LazyRows()
    rng = np.random.default_rng(11)
    queries = rng.standard_normal((1024, TOPK_DIM)).astype(np.float32)
    sql = f"SELECT id FROM {name} ORDER BY cosine_similarity(vector,cast(? as float array)) DESC LIMIT {TOPK_K}"
    out = {"table": f"{n_rows} x {TOPK_DIM}", "k": TOPK_K}
    plain = vector.CudaVectorDataSource()
    plain.initialize({"service": "cuda"})
    for i in range(5):
        plain.fetch_data(sql, [queries[i]])
    t0 = time.perf_counter()
    first = [plain.fetch_data(sql, [queries[i]]) for i in range(200)]
    out["one_query_per_call_qps"] = 200 / (time.perf_counter() - t0)
    ds = vector.CudaVectorDataSource()
    ds.initialize({"service": "cuda", "query-batch-size": 128, "query-flush-interval-ms": 1.0})
    for f in [ds.fetch_data_async(sql, [queries[i]]) for i in range(128)]:
        f.result()
    t0 = time.perf_counter()
    futs = [ds.fetch_data_async(sql, [queries[i]]) for i in range(1024)]
    res = [f.result() for f in futs]
    out["1024_requests_in_flight_qps"] = 1024 / (time.perf_counter() - t0)
    out["same_rows_as_one_at_a_time"] = bool(res[:200] == first)
    with ThreadPoolExecutor(64) as ex:
        t0 = time.perf_counter()
        list(ex.map(lambda i: ds.fetch_data(sql, [queries[i]]), range(1024)))
        out["64_blocking_callers_qps"] = 1024 / (time.perf_counter() - t0)
    out["searches_issued"] = ds._batcher.batches
    out["queries_answered"] = ds._batcher.queries
    t0 = time.perf_counter()
    many = ds.fetch_data_many(sql, [[queries[i]] for i in range(1024)])
    out["fetch_data_many_1024_qps"] = 1024 / (time.perf_counter() - t0)
    out["same_rows_many"] = bool(many[:200] == first)
    ds.close()
    plain.close()
    vector.drop_table(name)
    return out


def run_topk(args, world, rank, local_rank, dev, stream, barrier, max_over_ranks, pk, peak_tf):
    """Top-10 queries/s over a 10M x 384 bf16 corpus row-sharded over the ranks, B = 1 / 64 / 1024.

    One step = one query batch answered on every rank: N = 1: lsc_index_search_dev (normalise + ONE fused scan launch per
    pass); N > 1: lsc_index_search_sharded_dev - the same launch also stores the shard's candidates into every
    peer's receive buffer over NVLink and merges the peers' (lsc_exchange over CUDA IPC handles, exchanged once
    through torch.distributed).  `nccl_allgather_qps` times the unfused form (search, one NCCL all-gather per
    array, merge launch) beside it.  `e2e` starts from pinned HOST queries and ends with the merged ids / scores
    in host memory every step.  `parity` compares the timed result with exact fp32 scores of every shard row
    (lsc_index_scores_naive_dev, CUDA cores) - ids identical except ties within 1e-5."""
    import torch
    import torch.distributed as dist
    from langstream_b200._native import Exchange, Index, topk_merge_dev

    rows_total = args.topk_rows
    lo = rows_total * rank // world
    hi = rows_total * (rank + 1) // world
    n_local = hi - lo
    ix = Index(TOPK_DIM, capacity=n_local, device=local_rank)
    chunk = 1 << 20
    for o in range(0, n_local, chunk):
        m = min(chunk, n_local - o)
        g = torch.Generator(device=dev).manual_seed(42 + (lo + o) // chunk)
        vec = torch.randn(m, TOPK_DIM, generator=g, device=dev)
        ix.upsert_dev(np.arange(lo + o, lo + o + m, dtype=np.int64), vec, m, stream)
    del vec
    xg, exchange_kind = None, "none (one GPU)"
    if world > 1:
        try:
            xg = Exchange(local_rank, rank, world, 1024)
            handles = [None] * world
            dist.all_gather_object(handles, xg.ipc_handle())
            xg.connect_ipc(handles)
            exchange_kind = "fused into the scan launch: peer-to-peer stores + per-query flags (lsc_exchange, CUDA IPC)"
        except Exception as e:  # noqa: BLE001 - no P2P between these GPUs: the NCCL form is the product path then
            xg, exchange_kind = None, f"NCCL all-gather + merge launch (peer mapping failed: {e})"
        ok = torch.tensor([1 if xg is not None else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            xg = None
    ix.reserve(1024, TOPK_K)
    topk = {"corpus": f"{rows_total} x {TOPK_DIM} bf16, seed 42, row-sharded over {world} GPU(s)", "k": TOPK_K,
            "exchange": exchange_kind, "batches": {}}
    steps_topk = max(3, min(args.steps, 20))
    topk_traffic = None
    try:
        topk_traffic = json.loads((ROOT / "profiles" / "r02_traffic.json").read_text())["topk_scan_kernel"]["dram_bytes_per_launch"]
    except Exception:
        pass
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for B in (1, 64, 1024):
        g = torch.Generator(device=dev).manual_seed(43)
        q = torch.randn(B, TOPK_DIM, generator=g, device=dev)
        fin_ids = torch.empty(B, TOPK_K, dtype=torch.int64, device=dev)
        fin_sc = torch.empty(B, TOPK_K, dtype=torch.float32, device=dev)
        loc_ids = torch.empty(B, TOPK_K, dtype=torch.int64, device=dev)
        loc_sc = torch.empty(B, TOPK_K, dtype=torch.float32, device=dev)
        all_ids = torch.empty(world, B, TOPK_K, dtype=torch.int64, device=dev)
        all_sc = torch.empty(world, B, TOPK_K, dtype=torch.float32, device=dev)

        def q_step():
            if world == 1:
                ix.search_dev(q, B, TOPK_K, fin_ids, fin_sc, stream)
            elif xg is not None:
                ix.search_sharded_dev(xg, q, B, TOPK_K, fin_ids, fin_sc, stream)
            else:
                q_step_nccl()

        def q_step_nccl():
            ix.search_dev(q, B, TOPK_K, loc_ids, loc_sc, stream)
            dist.all_gather_into_tensor(all_ids, loc_ids)
            dist.all_gather_into_tensor(all_sc, loc_sc)
            topk_merge_dev(all_ids, all_sc, world, B, TOPK_K, fin_ids, fin_sc, stream)

        def timed(fn, n):
            barrier()
            ev0.record()
            for _ in range(n):
                fn()
            ev1.record()
            barrier()
            return max_over_ranks(ev0.elapsed_time(ev1))

        for _ in range(3):
            q_step()
        ms_q = timed(q_step, steps_topk)            # the QPS: nothing but the search in the timed region
        ix.set_profiling(True)                      # same steps again with every launch bracketed by events
        timed(q_step, steps_topk)
        pr = ix.read_profile()
        ix.set_profiling(False)
        ix.check(stream)
        scan_ms, scan_launches = pr["scan"]
        per_launch_ms = scan_ms / max(scan_launches, 1)
        bytes_per_launch = n_local * TOPK_DIM * 2
        gbs = bytes_per_launch / (per_launch_ms / 1000.0) / 1e9 if per_launch_ms > 0 else 0.0
        entry = {"qps": B * steps_topk / (ms_q / 1000.0), "ms_per_batch": ms_q / steps_topk,
                 "scan_launches_per_batch": scan_launches / steps_topk, "scan_ms_per_launch": per_launch_ms,
                 "scan_launch_covers": "scan + merge of the per-CTA lists" + (" + peer exchange + final merge" if xg is not None else "")}
        if B <= 128:
            entry["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s",
                                 "frac": gbs / pk["hbm_gbs"], "bytes_per_launch": bytes_per_launch,
                                 "traffic": topk_traffic if (world == 1 and rows_total == TOPK_ROWS) else None}
        else:
            tf = 2.0 * B * n_local * TOPK_DIM * steps_topk / (scan_ms / 1000.0) / 1e12 if scan_ms > 0 else 0.0
            entry["roofline"] = {"bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s",
                                 "frac": tf / peak_tf,
                                 "hbm_frac": (bytes_per_launch / (per_launch_ms / 1000.0) / 1e9 / pk["hbm_gbs"]) if per_launch_ms > 0 else None}
        if world > 1 and xg is not None:
            for _ in range(2):
                q_step_nccl()
            ms_n = timed(q_step_nccl, steps_topk)
            entry["nccl_allgather_qps"] = B * steps_topk / (ms_n / 1000.0)

        # ---- parity of the timed result (every rank checks its shard exactly; merged over ranks with torch) ----
        q_step()
        torch.cuda.synchronize()
        got_ids, got_sc = fin_ids.clone(), fin_sc.clone()
        ref_sc = torch.empty(B, TOPK_K, dtype=torch.float32, device=dev)
        ref_id = torch.empty(B, TOPK_K, dtype=torch.int64, device=dev)
        full = torch.empty(16 * n_local, dtype=torch.float32, device=dev)
        for b0 in range(0, B, 16):
            nb = min(16, B - b0)
            ix.scores_naive_dev(q[b0:b0 + nb].contiguous(), nb, full, stream)
            s, i = torch.topk(full[: nb * n_local].view(nb, n_local), TOPK_K, dim=1)
            ref_sc[b0:b0 + nb], ref_id[b0:b0 + nb] = s, i + lo
        del full
        if world > 1:
            g_sc = torch.empty(world, B, TOPK_K, dtype=torch.float32, device=dev)
            g_id = torch.empty(world, B, TOPK_K, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(g_sc, ref_sc)
            dist.all_gather_into_tensor(g_id, ref_id)
            cat_sc = g_sc.permute(1, 0, 2).reshape(B, world * TOPK_K)
            cat_id = g_id.permute(1, 0, 2).reshape(B, world * TOPK_K)
            ref_sc, pos = torch.topk(cat_sc, TOPK_K, dim=1)
            ref_id = torch.gather(cat_id, 1, pos)
        # rank by rank: the returned row's score must equal the exact k-th score within the tie tolerance
        rank_gap = float((got_sc - ref_sc).abs().max().item())
        differ = got_ids != ref_id
        beyond_tol = int((differ & ((got_sc - ref_sc).abs() > 1e-5)).sum().item())
        entry["parity"] = {"topk_mismatch": beyond_tol, "ids_differing_within_tie_tol": int(differ.sum().item()) - beyond_tol,
                           "max_abs_score_gap": rank_gap, "tol": 1e-5, "queries_checked": B,
                           "checker": "exact fp32 dot products of every shard row (CUDA cores) + top-k, merged over ranks"}

        # ---- e2e: pinned host queries in, merged ids + scores in host memory out, every step ----
        h_q = q.cpu().pin_memory()
        h_ids = torch.empty(B, TOPK_K, dtype=torch.int64).pin_memory()
        h_sc = torch.empty(B, TOPK_K, dtype=torch.float32).pin_memory()
        if world == 1:
            hq_np, hi_np, hs_np = h_q.numpy(), h_ids.numpy(), h_sc.numpy()
            from langstream_b200._native import lib as _lib, check as _check, _ptr

            def e2e_step():  # the host-buffer C-ABI call a datasource makes
                _check(_lib().lsc_index_search(ix._h, _ptr(hq_np), B, TOPK_K, _ptr(hi_np), _ptr(hs_np)))
            api = "lsc_index_search (host queries -> host ids + scores)"
        else:
            def e2e_step():
                q.copy_(h_q, non_blocking=True)
                q_step()
                h_ids.copy_(fin_ids, non_blocking=True)
                h_sc.copy_(fin_sc, non_blocking=True)
                torch.cuda.current_stream().synchronize()
            api = ("pinned host queries -> lsc_index_search_sharded_dev on every rank -> merged ids + scores copied to "
                   "host memory, synchronised every step")
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps_topk):
            e2e_step()
        torch.cuda.synchronize()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1000.0)
        entry["e2e"] = {"value": B * steps_topk / (e2e_ms / 1000.0), "unit": "queries/s", "api": api,
                        "h2d_bytes_per_step": int(B * TOPK_DIM * 4), "d2h_bytes_per_step": int(B * TOPK_K * 12)}
        topk["batches"][str(B)] = entry
    topk["gpu_launches_per_batch"] = {"1": 2, "64": 2, "1024": 6}  # normalise + fused scan; B > 256: normalise + 4 scans + merge
    if rank == 0 and world == 1:
        try:
            topk["query_batching"] = query_batching_figure(dev, stream)
        except Exception as e:  # noqa: BLE001 - an extra figure must not take the benchmark down
            topk["query_batching"] = {"error": str(e)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            topk["cpu_baseline"] = topk_cpu_baseline()
        except Exception as e:  # noqa: BLE001
            topk["cpu_baseline"] = {"error": str(e)}
    ix.close()
    if xg is not None:
        xg.close()
    return topk


# ---------------------------------------------------------------------------------------------
# config 5: the whole pipeline, native-only
# ---------------------------------------------------------------------------------------------
PIPE_WORDS_PER_DOC = 57      # x ("w12345" + blank) ~ 400 characters, the chunk size of the reference's crawler example
PIPE_VOCAB_WORDS = 30000


def _splitmix(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    return x ^ (x >> np.uint64(31))


def pipe_doc_words(doc_ids: np.ndarray) -> np.ndarray:
    """[m, 57] word indices of documents `doc_ids`: a pure function of (seed 7, document id, position), so that any
    rank can rebuild the text of any document it gets back from a query."""
    with np.errstate(over="ignore"):
        base = doc_ids.astype(np.uint64)[:, None] * np.uint64(PIPE_WORDS_PER_DOC) + np.arange(PIPE_WORDS_PER_DOC, dtype=np.uint64)[None, :]
        return (_splitmix(base + np.uint64(7) * np.uint64(0x1000003)) % np.uint64(PIPE_VOCAB_WORDS)).astype(np.int64)


def run_pipeline(args, world, rank, local_rank, dev, stream, barrier, max_over_ranks, enc, spec):
    """BASELINE configs[4] / SURVEY.md 8(d) config 5 as a NATIVE-ONLY end-to-end figure (the JVM stages around it -
    runtime-tester, Kafka, text-splitter - cannot run here; the shape is examples/applications/docker-chatbot/
    crawler.yaml:62-125 + chatbot.yaml:40-61): n synthetic ~400-character documents (seed 7), split over the ranks ->
    lsc_tokenizer_encode -> lsc_encoder_embed_dev -> lsc_index_upsert_dev (GPU-resident sink, no List<Double> in
    between) ; then 1024 query texts -> tokenise -> embed -> top-20 over all shards (lsc_index_search_sharded_dev)
    -> MMR re-rank on the device (lsc_rerank_term_statistics on the host, one lsc_rerank_mmr_batch_dev launch per batch).  Text
    generation is outside the timed regions (it stands in for the topic the documents arrive on)."""
    import queue
    import torch
    import torch.distributed as dist
    from langstream_b200._native import Exchange, Index, WordPieceTokenizer, rerank_mmr_batch_dev, rerank_term_statistics

    n_docs = args.pipeline_docs
    lo, hi = n_docs * rank // world, n_docs * (rank + 1) // world
    n_local = hi - lo
    words = [f"w{i}" for i in range(PIPE_VOCAB_WORDS)]
    W_obj = np.array(words, dtype=object)
    # the rank's share of the host cores, less two: the thread that feeds the GPU (and the driver's) must not wait for a core
    threads = max(1, (os.cpu_count() or 1) // world - (2 if (os.cpu_count() or 1) // world > 4 else 0))
    tok = WordPieceTokenizer("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words) + "\n", max_length=128, n_threads=threads)

    def texts_of(doc_ids):
        return [" ".join(r) for r in W_obj[pipe_doc_words(np.asarray(doc_ids, np.int64))]]

    batch = 8192
    # the "topic": documents arrive as UTF-8 bytes, as they do from a broker
    doc_texts = [[t.encode("utf-8") for t in texts_of(np.arange(o, min(o + batch, hi)))] for o in range(lo, hi, batch)]
    H = spec.hidden
    E = torch.empty(n_local, H, device=dev)                    # the table's vector column, kept for the re-ranker
    ix = Index(H, capacity=n_local, device=local_rank)
    xg = None
    if world > 1:
        xg = Exchange(local_rank, rank, world, 256)
        handles = [None] * world
        dist.all_gather_object(handles, xg.ipc_handle())
        xg.connect_ipc(handles)
    ix.reserve(64, 20)

    # ---- ingest: tokeniser thread -> (H2D, encoder, sink) on the stream --------------------------------
    def ingest_once(texts):
        ids, cu = tok.encode_packed(texts)
        return ids, cu

    ids, cu = ingest_once(doc_texts[0][:256])                  # warm-up of every stage
    d_ids, d_cu = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
    warm = torch.empty(256, H, device=dev)
    enc.embed_dev(d_ids, d_cu, 256, int(cu[-1]), int(np.diff(cu).max()), warm, stream)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    qd: "queue.Queue" = queue.Queue(maxsize=2)

    def producer():
        for texts in doc_texts:
            qd.put(ingest_once(texts))
        qd.put(None)

    # CPython hands the GIL over every 5 ms by default: the thread that feeds the GPU would wait that long for the
    # tokeniser thread's Python glue (pointer arrays, copies) every batch, with the GPU idle
    old_switch = sys.getswitchinterval()
    sys.setswitchinterval(1e-4)
    th = threading.Thread(target=producer, daemon=True)
    th.start()
    o, tokens = 0, 0
    while True:
        item = qd.get()
        if item is None:
            break
        ids, cu = item
        m = len(cu) - 1
        d_ids = torch.from_numpy(ids).pin_memory().to(dev, non_blocking=True)
        d_cu = torch.from_numpy(cu).pin_memory().to(dev, non_blocking=True)
        enc.embed_dev(d_ids, d_cu, m, int(cu[-1]), int(np.diff(cu).max()), E[o:o + m], stream)
        ix.upsert_dev(np.arange(lo + o, lo + o + m, dtype=np.int64), E[o:o + m], m, stream)  # returns when the rows are queryable
        o += m
        tokens += int(cu[-1])
    torch.cuda.synchronize()
    ingest_ms = max_over_ranks((time.perf_counter() - t0) * 1000.0)
    th.join()
    sys.setswitchinterval(old_switch)
    del doc_texts

    # ---- queries: 1024 texts (the first 8 words of a document each), 64 per batch ----------------------
    n_q, qb, k = 1024, 64, 20
    src_docs = (np.arange(n_q, dtype=np.int64) * 997 + 13) % n_docs
    q_texts = [" ".join(r[:8]) for r in W_obj[pipe_doc_words(src_docs)]]
    d_hit_ids = torch.empty(qb, k, dtype=torch.int64, device=dev)
    d_hit_sc = torch.empty(qb, k, dtype=torch.float32, device=dev)
    d_qe = torch.empty(qb, H, device=dev)
    self_hits, reranked = 0, 0

    def answer(b0):
        nonlocal self_hits, reranked
        texts = q_texts[b0:b0 + qb]
        ids, cu = tok.encode_packed(texts)
        d_i, d_c = torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev)
        enc.embed_dev(d_i, d_c, qb, int(cu[-1]), int(np.diff(cu).max()), d_qe, stream)
        if xg is not None:
            ix.search_sharded_dev(xg, d_qe, qb, k, d_hit_ids, d_hit_sc, stream)
        else:
            ix.search_dev(d_qe, qb, k, d_hit_ids, d_hit_sc, stream)
        # the hits' vectors: every rank contributes the rows it stores
        local = (d_hit_ids >= lo) & (d_hit_ids < hi)
        vec = torch.zeros(qb, k, H, device=dev)
        vec[local] = E[(d_hit_ids[local] - lo)]
        if world > 1:
            dist.all_reduce(vec)
        hit_ids = d_hit_ids.cpu().numpy()
        self_hits += int((hit_ids == src_docs[b0:b0 + qb, None]).any(axis=1).sum())
        # this rank re-ranks its share of the batch (MMR, lambda 0.5, BM25 k1 1.5 b 0.75 - the chatbot example's values)
        mine = [j for j in range(qb) if (b0 + j) % world == rank]
        hit_texts = texts_of(hit_ids[mine].reshape(-1))
        n_terms = 8
        tf_all = np.zeros((len(mine), k, n_terms), np.int32)
        len_all = np.zeros((len(mine), k), np.int32)
        for jj, j in enumerate(mine):  # ReRankAgent.tokenise + term counts, natively (lsc_rerank_term_statistics)
            tf, dl = rerank_term_statistics(hit_texts[jj * k:(jj + 1) * k], texts[j], n_terms)
            tf_all[jj, :, :tf.shape[1]] = tf
            len_all[jj] = dl
        d_tf = torch.from_numpy(tf_all).to(dev)
        d_len = torch.from_numpy(len_all).to(dev)
        d_vm = vec[mine].contiguous()
        d_om = torch.empty(len(mine), k + 1, dtype=torch.int32, device=dev)   # per query: k picked positions + the count
        rerank_mmr_batch_dev(d_vm, len(mine), k, H, d_tf, d_len, n_terms, 5, 0.5, 1.5, 0.75, d_om, k + 1, d_om[:, k:], k + 1,
                             stream)                                  # ONE launch for the rank's share of the batch
        order_m = d_om.cpu().numpy()   # synchronises: the batch's answers are in host memory
        reranked += sum(int(order_m[jj, k]) == 5 for jj in range(len(mine)))

    answer(0)                                                  # warm-up
    self_hits, reranked = 0, 0
    barrier()
    t0 = time.perf_counter()
    for b0 in range(0, n_q, qb):
        answer(b0)
    torch.cuda.synchronize()
    query_ms = max_over_ranks((time.perf_counter() - t0) * 1000.0)
    if world > 1:
        t = torch.tensor([reranked], device=dev)
        dist.all_reduce(t)
        reranked = int(t.item())
    out = {"documents": n_docs, "chars_per_document": "~400 (57 pseudo-words, seed 7)", "tokens_per_document": tokens / max(n_local, 1),
           "ingest": {"value": n_docs / (ingest_ms / 1000.0), "unit": "documents/s", "seconds": ingest_ms / 1000.0,
                      "stages": "lsc_tokenizer_encode (%d host threads per rank) -> H2D -> lsc_encoder_embed_dev -> lsc_index_upsert_dev" % threads},
           "query": {"value": n_q / (query_ms / 1000.0), "unit": "queries/s", "queries": n_q, "batch": qb, "k": k,
                     "stages": "tokenise -> embed -> top-20 over all shards -> gather hit vectors -> lsc_rerank_term_statistics (host) -> "
                               "lsc_rerank_mmr_batch_dev (5 of 20, one launch per batch) -> D2H",
                     "source_document_in_top20": self_hits / n_q, "queries_reranked_to_5": reranked},
           "scope": "native-only e2e: the JVM stages of the reference pipeline (runtime-tester, broker, text-splitter) cannot run "
                    "in this image"}
    ix.close()
    if xg is not None:
        xg.close()
    tok.close()
    return out

# ---------------------------------------------------------------------------------------------
# CUDA arm
# ---------------------------------------------------------------------------------------------
def run_cuda(args):
    import torch
    import torch.distributed as dist
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder, Index, topk_merge_dev

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pk, pk_src = peaks()
    spec = W.MINILM_L6
    blob = W.random_init(spec, seed=0)
    enc = Encoder(spec, blob, device=local_rank)
    stream = torch.cuda.current_stream().cuda_stream

    # every rank embeds its own batch (pure data parallelism, no collective)
    ids, cu = W.synthetic_token_ids(N_SEQ, SEQ_LEN, seed=1234 + rank)
    d_ids = torch.from_numpy(ids).to(dev)
    d_cu = torch.from_numpy(cu).to(dev)
    d_out = torch.empty(N_SEQ, spec.hidden, device=dev)
    T = int(cu[-1])

    def step_dev():
        enc.embed_dev(d_ids, d_cu, N_SEQ, T, SEQ_LEN, d_out, stream)

    for _ in range(max(args.warmup, 3)):
        step_dev()
    barrier()
    # ---- headline: K steps, nothing but the kernels in the timed region ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = enc.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step_dev()
    ev1.record()
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.stop()
    launches = enc.launch_count - launches0
    value = world * N_SEQ * args.steps / (ms_total / 1000.0)

    # parity of what the timed steps produced: the first sequences of this rank's batch against the fp32 oracle
    # (one text per forward, as the reference runs them); worst margin over the ranks
    parity = {"tolerance": {"cos_min": 0.999, "max_abs": 2e-3}}
    try:
        from oracle.encoder_oracle import HFOracle
        torch.set_num_threads(max(1, (os.cpu_count() or 1) // world))
        n_chk = 16
        want = HFOracle(spec, blob).embed_one_by_one(ids[: cu[n_chk]], cu[: n_chk + 1])
        got = d_out[:n_chk].cpu().numpy()
        cosv = (got * want).sum(1) / (np.linalg.norm(got, axis=1) * np.linalg.norm(want, axis=1))
        cos_min, max_abs = float(cosv.min()), float(np.abs(got - want).max())
        if world > 1:
            t = torch.tensor([-cos_min, max_abs], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            cos_min, max_abs = -float(t[0].item()), float(t[1].item())
        parity.update({"cos_min": cos_min, "max_abs": max_abs, "rows_checked_per_gpu": n_chk,
                       "oracle": "HuggingFace BertModel fp32 on torch CPU (parity unpinned against the reference's own "
                                 "tests: its one golden vector needs e5-small weights)"})
    except Exception as e:  # noqa: BLE001
        parity["error"] = str(e)

    # ---- same K steps again with every launch bracketed by CUDA events on the launching stream:
    #      per-kernel durations for the roofline (the events add a few microseconds per launch, which is
    #      why this pass is not the headline) ----
    enc.set_profiling(True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step_dev()
    ev1.record()
    barrier()
    ms_profiled = max_over_ranks(ev0.elapsed_time(ev1))
    prof = enc.read_profile()
    enc.set_profiling(False)

    # roofline of the dominant kernel: the fused feed-forward kernel (two of the four GEMMs of a layer, half of
    # the step) when the model shape supports it, else the tcgen05 GEMM class.  Algorithmic dense FLOPs of the
    # launches of that class / their event-timed duration.
    total_flop, dense_flop = flops_per_sequence(spec, SEQ_LEN)
    H, F, L = spec.hidden, spec.ffn, spec.layers
    tokens = N_SEQ * SEQ_LEN
    ffn_flop_step = 4.0 * tokens * H * F * L            # FFN-up + FFN-down
    gemm_flop_step = dense_flop * N_SEQ                 # all four GEMM shapes
    gemm_ms, gemm_launches = prof["gemm"]
    ffn_ms, ffn_launches = prof["ffn_fused"]
    peak_tf = pk["bf16_tflops_sustained"]
    if ffn_launches > 0:
        dom_name = "ffn_fused_kernel (tcgen05 cta_group::2: FFN-up + GELU + FFN-down + residual, LayerNorm folded in)"
        dom_ms, dom_launches, dom_flop = ffn_ms, ffn_launches, ffn_flop_step * args.steps
        traffic_key = "ffn_fused_kernel"
        other_flop = (gemm_flop_step - ffn_flop_step) * args.steps
    else:
        dom_name = "gemm_tc_kernel (tcgen05/TMEM/TMA, all 4 encoder GEMM shapes, LayerNorm folded in)"
        dom_ms, dom_launches, dom_flop = gemm_ms, gemm_launches, gemm_flop_step * args.steps
        traffic_key = "gemm_tc_kernel"
        other_flop = 0.0
    dom_tflops = dom_flop / (dom_ms / 1000.0) / 1e12 if dom_ms > 0 else 0.0
    traffic = None
    tpath = ROOT / "profiles" / "r02_traffic.json"
    if tpath.exists():
        try:
            traffic = json.loads(tpath.read_text())[traffic_key]["dram_bytes_per_launch_avg"]
        except Exception:
            traffic = None
    roofline = {
        "bound": "tensor", "kernel": dom_name,
        "achieved": dom_tflops, "peak": peak_tf, "unit": "TFLOP/s",
        "frac": dom_tflops / peak_tf if peak_tf else None, "traffic": traffic,
        "traffic_source": "ncu --set full capture under profiles/ (dram read+write per launch)",
        "peak_source": pk_src + ", sustained bf16 (kernel timed inside a long step)",
        "flops_per_launch": dom_flop / max(dom_launches, 1),
        "avg_launch_ms": dom_ms / max(dom_launches, 1),
        "other_gemm_tflops": (other_flop / (gemm_ms / 1000.0) / 1e12) if (other_flop > 0 and gemm_ms > 0) else None,
        "kernel_share_of_step": {k: v[0] / max(sum(x[0] for x in prof.values()), 1e-9) for k, v in prof.items()},
        "ms_per_step_with_event_profiling": ms_profiled / args.steps,
        "whole_step_tflops": total_flop * N_SEQ * args.steps / (ms_total / 1000.0) / 1e12,
    }

    # ragged variant of the same batch size (packed variable length, UniformInt[16,128])
    ids_v, cu_v = W.synthetic_token_ids(N_SEQ, SEQ_LEN, seed=4321 + rank, min_len=16)
    dv_ids, dv_cu = torch.from_numpy(ids_v).to(dev), torch.from_numpy(cu_v).to(dev)
    Tv, maxv = int(cu_v[-1]), int(np.diff(cu_v).max())
    for _ in range(3):
        enc.embed_dev(dv_ids, dv_cu, N_SEQ, Tv, maxv, d_out, stream)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        enc.embed_dev(dv_ids, dv_cu, N_SEQ, Tv, maxv, d_out, stream)
    ev1.record()
    barrier()
    ms_var = max_over_ranks(ev0.elapsed_time(ev1))
    varlen = {"value": world * N_SEQ * args.steps / (ms_var / 1000.0), "unit": "embeddings/s",
              "lengths": "UniformInt[16,128] packed", "tokens_per_step": Tv}

    # config 3 of SURVEY.md 8(d), as an extra figure (not the headline): the base shape (12 layers, H = 768,
    # 12 heads of 64, FFN 3072), 4096 sequences x 512 tokens SPLIT over the N GPUs (strong scaling: contiguous
    # slices of the sequence list, no collective), fixed length and UniformInt[64,512] packed; each GPU runs its
    # slice in passes of at most 262,144 tokens (the pass size the host API cuts batches into).
    base = None
    if not args.no_base:
        bspec = W.BERT_BASE
        benc = Encoder(bspec, W.random_init(bspec, 0), device=local_rank)
        b_total_seqs, b_s = 4096, 512
        s_lo, s_hi = b_total_seqs * rank // world, b_total_seqs * (rank + 1) // world
        base = {"workload": f"base shape 12L/768/12h/3072, {b_total_seqs} sequences x {b_s} tokens split over {world} GPU(s), "
                            "passes of <= 262,144 tokens", "scaling": "strong"}
        for label, min_len in (("fixed_512", None), ("ragged_64_512", 64)):
            a_ids, a_cu = W.synthetic_token_ids(b_total_seqs, b_s, seed=99, min_len=min_len)
            a_cu = a_cu.astype(np.int64)
            passes, p0 = [], s_lo
            while p0 < s_hi:  # whole sequences, <= 262,144 tokens per pass
                p1 = p0 + 1
                while p1 < s_hi and a_cu[p1 + 1] - a_cu[p0] <= (1 << 18):
                    p1 += 1
                ids_p = torch.from_numpy(a_ids[a_cu[p0]:a_cu[p1]].copy()).to(dev)
                cu_p = torch.from_numpy((a_cu[p0:p1 + 1] - a_cu[p0]).astype(np.int32)).to(dev)
                out_p = torch.empty(p1 - p0, bspec.hidden, device=dev)
                passes.append((ids_p, cu_p, p1 - p0, int(a_cu[p1] - a_cu[p0]), int(np.diff(a_cu[p0:p1 + 1]).max()), out_p))
                p0 = p1

            def base_step():
                for ids_p, cu_p, n_p, t_p, m_p, out_p in passes:
                    benc.embed_dev(ids_p, cu_p, n_p, t_p, m_p, out_p, stream)

            base_step()
            barrier()
            reps = 2
            ev0.record()
            for _ in range(reps):
                base_step()
            ev1.record()
            barrier()
            ms_base = max_over_ranks(ev0.elapsed_time(ev1))
            lens = np.diff(a_cu)
            flop = float(sum(bspec.layers * (24 * bspec.hidden ** 2 * int(l) + 4 * bspec.hidden * int(l) ** 2) for l in lens))
            base[label] = {"value": b_total_seqs * reps / (ms_base / 1000.0), "unit": "embeddings/s",
                           "ms_per_4096_sequences": ms_base / reps, "passes_per_gpu": len(passes),
                           "tokens_total": int(a_cu[-1]),
                           "whole_job_tflops": flop * reps / (ms_base / 1000.0) / 1e12,
                           "frac_of_sustained_peak_per_gpu": flop * reps / (ms_base / 1000.0) / 1e12 / world / peak_tf}
            del passes
        benc.close()
        torch.cuda.empty_cache()

    # e2e: the C-ABI host-buffer call, pinned host memory, copies inside the timed region
    h_ids = torch.from_numpy(ids).pin_memory()
    h_cu = torch.from_numpy(cu).pin_memory()
    h_out = torch.empty(N_SEQ, spec.hidden).pin_memory()
    h_ids_np, h_cu_np, h_out_np = h_ids.numpy(), h_cu.numpy(), h_out.numpy()
    # two batches in flight through lsc_encoder_submit / lsc_encoder_wait (the asynchronous form of the service
    # contract): the upload of step i+1 and the download of step i-1 overlap the kernels of step i.  Every step
    # uploads its ids from pinned host memory and its embeddings are in host memory when its wait returns.
    h_out2 = torch.empty(N_SEQ, spec.hidden).pin_memory()
    h_ids2, h_cu2 = h_ids.clone().pin_memory(), h_cu.clone().pin_memory()
    bufs = [(h_ids_np, h_cu_np, h_out_np), (h_ids2.numpy(), h_cu2.numpy(), h_out2.numpy())]
    for _ in range(3):
        enc.embed(h_ids_np, h_cu_np, out=h_out_np)
    for w in range(max(args.warmup, 3)):  # warm-up of the asynchronous path too: its two slots allocate on first use
        enc.wait(enc.submit(*bufs[w & 1]))
    barrier()
    t0 = time.perf_counter()
    pending = enc.submit(*bufs[0])
    for i in range(1, args.steps):
        nxt = enc.submit(*bufs[i & 1])
        enc.wait(pending)  # step i-1 is complete in host memory
        pending = nxt
    enc.wait(pending)
    e2e_s = time.perf_counter() - t0
    e2e_ms = max_over_ranks(e2e_s * 1000.0)
    # and the blocking call, one batch at a time
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        enc.embed(h_ids_np, h_cu_np, out=h_out_np)  # blocks until the embeddings are in host memory
    torch.cuda.synchronize()
    blocking_ms = max_over_ranks((time.perf_counter() - t0) * 1000.0)
    e2e = {"value": world * N_SEQ * args.steps / (e2e_ms / 1000.0), "unit": "embeddings/s",
           "h2d_bytes_per_step": int(h_ids_np.nbytes + h_cu_np.nbytes),
           "d2h_bytes_per_step": int(h_out_np.nbytes),
           "api": "lsc_encoder_submit / lsc_encoder_wait, two batches in flight (host token ids -> host fp32 embeddings)",
           "blocking_call_value": world * N_SEQ * args.steps / (blocking_ms / 1000.0),
           "blocking_call_api": "lsc_encoder_embed, one batch at a time"}

    # from texts: lsc_tokenizer_encode (native WordPiece, all host threads) + lsc_encoder_embed, per step 1024
    # synthetic texts of 126 pseudo-words (= 128 tokens with [CLS] / [SEP]); extra figure, not the headline
    text_e2e = None
    try:
        from langstream_b200._native import WordPieceTokenizer
        words = [f"w{i}" for i in range(spec.vocab - 5)]
        tok_threads = max(1, (os.cpu_count() or 1) // world)  # the ranks of one box share its host cores
        tok = WordPieceTokenizer("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words) + "\n",
                                 max_length=SEQ_LEN, n_threads=tok_threads)
        trng = np.random.default_rng(77 + rank)
        texts = [" ".join(words[j] for j in trng.integers(0, len(words), SEQ_LEN - 2)) for _ in range(N_SEQ)]
        t_ids, t_cu = tok.encode_packed(texts)
        assert int(t_cu[-1]) == N_SEQ * SEQ_LEN
        reps = max(3, min(args.steps, 10))
        enc.embed(t_ids, t_cu, out=h_out_np)
        barrier()
        t0 = time.perf_counter()
        tok_s = 0.0
        pending, keep = None, []
        for r in range(reps):  # tokenise batch r while the GPU works on batch r-1 (submit / wait)
            a0 = time.perf_counter()
            t_ids, t_cu = tok.encode_packed(texts)
            tok_s += time.perf_counter() - a0
            keep = keep[-1:] + [(t_ids, t_cu)]  # the arrays of the batches in flight stay referenced
            t = enc.submit(t_ids, t_cu, bufs[r & 1][2])
            if pending is not None:
                enc.wait(pending)
            pending = t
        enc.wait(pending)
        text_ms = max_over_ranks((time.perf_counter() - t0) * 1000.0)
        text_e2e = {"value": world * N_SEQ * reps / (text_ms / 1000.0), "unit": "embeddings/s",
                    "api": "lsc_tokenizer_encode + lsc_encoder_submit / lsc_encoder_wait (host texts -> host fp32 embeddings), "
                           "tokenisation of batch i+1 overlapping the kernels of batch i",
                    "tokenizer_texts_per_s": N_SEQ * reps / tok_s, "tokenizer_threads": tok_threads}
        tok.close()
    except Exception as e:  # noqa: BLE001 - an extra figure must not take the benchmark down
        text_e2e = {"error": str(e)}

    # ---- top-k over a 10M x 384 corpus, row-sharded across ranks -------------------------------
    topk = None
    if not args.no_topk:
        topk = run_topk(args, world, rank, local_rank, dev, stream, barrier, max_over_ranks, pk, peak_tf)

    pipeline = None
    if not args.no_pipeline:
        try:
            pipeline = run_pipeline(args, world, rank, local_rank, dev, stream, barrier, max_over_ranks, enc, spec)
        except Exception as e:  # noqa: BLE001 - an extra figure must not take the benchmark down
            import traceback
            pipeline = {"error": str(e), "trace": traceback.format_exc()[-600:]}

    # ---- CPU baseline (rank 0, N=1 only): reference-shaped oracle on a bounded sample ------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rate, cores, done, el = cpu_reference_rate(32, min_seconds=10.0, max_seconds=30.0)
        cpu_baseline = {"value": rate, "unit": "embeddings/s", "cores": cores, "kind": "port",
                        "sample": f"{done} sequences x {SEQ_LEN} tokens in {el:.1f} s, one text per forward "
                                  "(HuggingFace BertModel fp32 on torch CPU = the graph DJL runs)"}

    if rank == 0:
        line = {
            "metric": "embeddings/s", "value": value, "unit": "embeddings/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": f"all-MiniLM-L6-v2 architecture (6L,384d,12h,ffn1536) random-init seed 0, "
                                   f"{N_SEQ} x {SEQ_LEN} synthetic tokens per GPU per step, packed",
                       "parallelism": f"dp{world} (one replica per GPU, no collective)",
                       "l2": "per-step working set ~1.1 GB of activations >> 126 MB L2; no explicit flush"},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "parity": parity,
            "gpu_launches": int(launches), "clocks": clocks, "varlen": varlen, "base_4096x512": base, "text_e2e": text_e2e, "topk": topk, "pipeline_e2e": pipeline,
        }
        emit(line)
    enc.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-topk", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-base", action="store_true", help="skip the base-shape (config 3) extra figure")
    ap.add_argument("--topk-rows", type=int, default=TOPK_ROWS)
    ap.add_argument("--no-pipeline", action="store_true", help="skip the config-5 pipeline figure")
    ap.add_argument("--pipeline-docs", type=int, default=1_000_000)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and args.impl == "cuda" and world == 1 and args.gpus > 1:
        # launched without torchrun: re-launch one rank per GPU
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29531", __file__] + sys.argv[1:]
        return subprocess.call(cmd)
    capture_stdout()
    if args.impl == "reference":
        return run_reference(args)
    return run_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
