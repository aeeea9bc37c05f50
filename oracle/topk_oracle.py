"""Python side of the top-k oracle.  TEST INFRASTRUCTURE ONLY (see topk_oracle.c for citations).

`search_scalar`  — the literal ReRankAgent-form cosine + stable top-k, in C (oracle/_build/liboracle.so),
                   for small N.
`search_numpy`   — the fast form used at larger N: rows and queries L2-normalised in fp32, rounded to
                   bf16 (what the index stores), widened back, fp32 matmul, stable descending sort
                   with ties broken by lower row index.
`assert_topk_matches` — the parity rule of BASELINE.json: ids identical except where scores tie
                   within `tol` (1e-5).
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = _HERE / "_build" / "liboracle.so"
_lib = None


def build() -> Path:
    src = _HERE / "topk_oracle.c"
    if not _LIB.exists() or _LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE)], check=True, capture_output=True)
    return _LIB


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(_LIB))
        _lib.oracle_cosine.restype = C.c_float
        _lib.oracle_cosine.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        _lib.oracle_search.restype = None
        _lib.oracle_search.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                       C.c_void_p, C.c_void_p]
    return _lib


def cosine_scalar(a: np.ndarray, b: np.ndarray) -> float:
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return float(_load().oracle_cosine(a.ctypes.data, b.ctypes.data, int(a.size)))


def search_scalar(rows: np.ndarray, queries: np.ndarray, k: int):
    rows = np.ascontiguousarray(rows, np.float32)
    queries = np.ascontiguousarray(queries, np.float32).reshape(-1, rows.shape[1])
    B = queries.shape[0]
    ids = np.empty((B, k), np.int64)
    scores = np.empty((B, k), np.float32)
    _load().oracle_search(rows.ctypes.data, rows.shape[0], rows.shape[1], queries.ctypes.data, B, k,
                          ids.ctypes.data, scores.ctypes.data)
    return ids, scores


def to_bf16_rounded(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 (round to nearest even) -> fp32, in NumPy."""
    u = np.ascontiguousarray(x, np.float32).view(np.uint32).astype(np.uint64)
    rounded = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return rounded.astype(np.uint32).view(np.float32).reshape(np.shape(x))


def normalize_rows(x: np.ndarray) -> np.ndarray:
    x = np.asarray(x, np.float32)
    n = np.sqrt((x * x).sum(-1, keepdims=True, dtype=np.float32))
    return np.where(n > 0, x / np.where(n > 0, n, 1), 0).astype(np.float32)


def scores_numpy(rows: np.ndarray, queries: np.ndarray) -> np.ndarray:
    r = to_bf16_rounded(normalize_rows(rows))
    q = to_bf16_rounded(normalize_rows(queries))
    return (q.astype(np.float32) @ r.astype(np.float32).T).astype(np.float32)


def topk_stable(scores: np.ndarray, k: int):
    """Top-k per row of `scores`: score descending, ties by lower column index."""
    B, n = scores.shape
    kk = min(k, n)
    order = np.argsort(-scores, axis=1, kind="stable")[:, :kk]
    ids = np.full((B, k), -1, np.int64)
    out = np.full((B, k), -np.inf, np.float32)
    ids[:, :kk] = order
    out[:, :kk] = np.take_along_axis(scores, order, axis=1)
    return ids, out


def search_numpy(rows: np.ndarray, queries: np.ndarray, k: int):
    return topk_stable(scores_numpy(rows, np.asarray(queries, np.float32).reshape(-1, rows.shape[1])), k)


def assert_topk_matches(got_ids, got_scores, full_scores, k, tol=1e-5, id_of_slot=None):
    """Parity rule: position by position the returned id must be the oracle's id, unless the two
    candidates' oracle scores differ by <= tol (a tie).  Also checks the returned scores."""
    got_ids = np.asarray(got_ids)
    got_scores = np.asarray(got_scores)
    ref_ids, ref_scores = topk_stable(np.asarray(full_scores, np.float32), k)
    B = ref_ids.shape[0]
    for b in range(B):
        for j in range(k):
            rid = ref_ids[b, j]
            gid = got_ids[b, j]
            if rid < 0:
                assert gid == -1, f"query {b} pos {j}: expected empty slot, got id {gid}"
                continue
            gslot = gid if id_of_slot is None else id_of_slot[gid]
            assert gslot >= 0, f"query {b} pos {j}: got empty slot, expected row {rid}"
            assert abs(float(got_scores[b, j]) - float(full_scores[b, gslot])) <= tol, (
                f"query {b} pos {j}: returned score {got_scores[b, j]} but row {gslot} scores {full_scores[b, gslot]}")
            if gslot != rid:
                assert abs(float(full_scores[b, gslot]) - float(ref_scores[b, j])) <= tol, (
                    f"query {b} pos {j}: got row {gslot} (score {full_scores[b, gslot]}), "
                    f"oracle row {rid} (score {ref_scores[b, j]})")
        valid = got_ids[b][got_ids[b] >= 0]
        assert len(set(valid.tolist())) == len(valid), f"query {b}: duplicate ids {got_ids[b]}"
