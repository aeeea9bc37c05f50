"""ctypes binding of include/langstream_cuda.h.

The library is loaded from langstream_b200/_lib/ (built in-tree by `langstream_b200.build`).
Nothing here computes anything: a missing library or missing GPU raises, it never falls back.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "_lib" / "liblangstream_cuda.so"

LSC_OK = 0
LSC_E_INVALID = -1
LSC_E_CUDA = -2
LSC_E_NOMEM = -3
LSC_E_UNSUPPORTED = -4
LSC_E_IO = -5
LSC_E_TOO_LONG = -6


class NativeError(RuntimeError):
    """A C-ABI call returned a negative code; `.code` is the LSC_E_* value."""

    def __init__(self, code: int, message: str):
        super().__init__(f"langstream-cuda error {code}: {message}")
        self.code = code
        self.message = message


class EncoderConfig(C.Structure):
    _fields_ = [
        ("layers", C.c_int32),
        ("hidden", C.c_int32),
        ("heads", C.c_int32),
        ("ffn", C.c_int32),
        ("vocab", C.c_int32),
        ("max_position", C.c_int32),
        ("type_vocab", C.c_int32),
        ("ln_eps", C.c_float),
        ("gemm_impl", C.c_int32),
        ("arch", C.c_int32),
        ("rel_buckets", C.c_int32),
        ("rel_max_distance", C.c_int32),
        ("reserved", C.c_int32 * 4),
    ]


_vp = C.c_void_p
_i32 = C.c_int32
_i64 = C.c_int64

# name -> (restype, argtypes); mirrors include/langstream_cuda.h one to one.
SIGNATURES = {
    "lsc_version": (C.c_char_p, []),
    "lsc_last_error": (C.c_char_p, []),
    "lsc_device_count": (C.c_int, []),
    "lsc_encoder_weights_nbytes": (C.c_size_t, [C.POINTER(EncoderConfig)]),
    "lsc_encoder_create": (C.c_int, [C.POINTER(EncoderConfig), _vp, C.c_size_t, C.c_int, C.POINTER(_vp)]),
    "lsc_encoder_embed": (C.c_int, [_vp, _vp, _vp, _i32, _vp]),
    "lsc_encoder_embed_dev": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "lsc_encoder_submit": (C.c_int, [_vp, _vp, _vp, _i32, _vp, C.POINTER(_i64)]),
    "lsc_encoder_wait": (C.c_int, [_vp, _i64]),
    "lsc_encoder_set_profiling": (C.c_int, [_vp, C.c_int]),
    "lsc_encoder_read_profile": (C.c_int, [_vp, _vp, _vp, _i32]),
    "lsc_encoder_dim": (C.c_int, [_vp]),
    "lsc_encoder_launch_count": (_i64, [_vp]),
    "lsc_encoder_destroy": (None, [_vp]),
    "lsc_index_create": (C.c_int, [_i32, _i64, C.c_int, C.POINTER(_vp)]),
    "lsc_index_upsert": (C.c_int, [_vp, _vp, _vp, _i32]),
    "lsc_index_upsert_dev": (C.c_int, [_vp, _vp, _vp, _i32, _vp]),
    "lsc_index_delete": (C.c_int, [_vp, _vp, _i32]),
    "lsc_index_size": (_i64, [_vp]),
    "lsc_index_search": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _vp]),
    "lsc_index_search_dev": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "lsc_index_scores_naive_dev": (C.c_int, [_vp, _vp, _i32, _vp, _vp]),
    "lsc_topk_merge_dev": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
    "lsc_exchange_create": (C.c_int, [C.c_int, _i32, _i32, _i32, C.POINTER(_vp)]),
    "lsc_exchange_ipc_handle": (C.c_int, [_vp, _vp]),
    "lsc_exchange_connect_ipc": (C.c_int, [_vp, _vp]),
    "lsc_exchange_connect_local": (C.c_int, [C.POINTER(_vp), _i32]),
    "lsc_exchange_destroy": (None, [_vp]),
    "lsc_index_search_sharded_dev": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "lsc_index_check": (C.c_int, [_vp, _vp]),
    "lsc_index_reserve": (C.c_int, [_vp, _i32, _i32]),
    "lsc_sharded_index_create": (C.c_int, [_i32, _i64, C.POINTER(C.c_int), _i32, C.POINTER(_vp)]),
    "lsc_sharded_index_upsert": (C.c_int, [_vp, _vp, _vp, _i32]),
    "lsc_sharded_index_delete": (C.c_int, [_vp, _vp, _i32]),
    "lsc_sharded_index_size": (_i64, [_vp]),
    "lsc_sharded_index_search": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _vp]),
    "lsc_sharded_index_shards": (_i32, [_vp]),
    "lsc_sharded_index_shard": (_vp, [_vp, _i32]),
    "lsc_sharded_index_destroy": (None, [_vp]),
    "lsc_topk_search_static_dev": (C.c_int, [_vp, _vp, _i64, _i32, _vp, _i32, _i32, _vp, _vp, _vp]),
    "lsc_index_set_profiling": (C.c_int, [_vp, C.c_int]),
    "lsc_index_read_profile": (C.c_int, [_vp, _vp, _vp, _i32]),
    "lsc_index_launch_count": (_i64, [_vp]),
    "lsc_index_save": (C.c_int, [_vp, C.c_char_p]),
    "lsc_index_load": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(_vp)]),
    "lsc_index_destroy": (None, [_vp]),
    "lsc_tokenizer_create": (C.c_int, [C.c_char_p, C.c_size_t, C.c_int, _i32, _i32, C.POINTER(_vp)]),
    "lsc_tokenizer_encode": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _i64, _vp]),
    "lsc_tokenizer_set_special": (C.c_int, [_vp, C.c_char_p, C.c_char_p, C.c_char_p]),
    "lsc_tokenizer_vocab_size": (_i32, [_vp]),
    "lsc_tokenizer_destroy": (None, [_vp]),
    "lsc_rerank_mmr": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _i32, _i32, C.c_double, C.c_double, C.c_double, _vp, _vp]),
    "lsc_rerank_mmr_dev": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _i32, _i32, C.c_double, C.c_double, C.c_double, _vp, _vp, _vp]),
    "lsc_rerank_mmr_batch_dev": (C.c_int, [_vp, _i32, _i32, _i32, _vp, _vp, _i32, _i32, C.c_double, C.c_double, C.c_double,
                                          _vp, _i32, _vp, _i32, _vp]),
    "lsc_rerank_term_statistics": (C.c_int, [_vp, _vp, _i32, C.c_char_p, _i32, _i32, _vp, _vp, _vp]),
    "lsc_debug_attention_dev": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "lsc_debug_attention_blocks_dev": (C.c_int, [_vp, _i32, _vp, _vp]),
    "lsc_debug_mpnet_bias_table": (C.c_int, [_vp, _i32, _i32, _i32, _vp]),
    "lsc_debug_gemm_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the native library; raises if it has not been built."""
    global _lib
    if _lib is None:
        import os
        path = Path(os.environ.get("LSC_NATIVE_LIB") or LIB_PATH)  # LSC_NATIVE_LIB: A/B builds of the same library (profiles/)
        if not path.exists():
            raise RuntimeError(
                f"{path} is missing: build it with `python -m langstream_b200.build` "
                "(there is no CPU fallback for this path)"
            )
        handle = C.CDLL(str(path))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def last_error() -> str:
    return lib().lsc_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != LSC_OK:
        raise NativeError(rc, last_error())


# ---------------------------------------------------------------------------------------------
# Handle wrappers (numpy for host buffers; raw device pointers + stream for the *_dev calls)
# ---------------------------------------------------------------------------------------------
import numpy as _np


def _ptr(a) -> int:
    """Address of a C-contiguous numpy array, or the data_ptr of a torch tensor, or an int."""
    if a is None:
        return 0
    if isinstance(a, int):
        return a
    if isinstance(a, _np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


class Encoder:
    """One encoder replica on one GPU (lsc_encoder)."""

    def __init__(self, spec, blob, device: int = 0, gemm_impl: int = 0):
        L = lib()
        cfg = EncoderConfig(spec.layers, spec.hidden, spec.heads, spec.ffn, spec.vocab, spec.max_position,
                            spec.type_vocab, spec.ln_eps, gemm_impl, getattr(spec, "arch_id", 0),
                            getattr(spec, "rel_buckets", 0), getattr(spec, "rel_max_distance", 0))
        blob = _np.ascontiguousarray(blob, _np.float32)
        need = L.lsc_encoder_weights_nbytes(C.byref(cfg))
        if need == 0:
            raise NativeError(LSC_E_INVALID, last_error())
        h = _vp()
        check(L.lsc_encoder_create(C.byref(cfg), blob.ctypes.data, blob.nbytes, device, C.byref(h)))
        self._h = h
        self.spec = spec
        self.device = device
        self.dim = spec.hidden

    def embed(self, token_ids, cu_seqlens, out=None):
        """Host API: packed int32 token ids + cu_seqlens -> [n_seq, hidden] fp32 (numpy)."""
        ids = _np.ascontiguousarray(token_ids, _np.int32)
        cu = _np.ascontiguousarray(cu_seqlens, _np.int32)
        n = int(cu.size) - 1
        if out is None:
            out = _np.empty((max(n, 0), self.dim), _np.float32)
        check(lib().lsc_encoder_embed(self._h, _ptr(ids), _ptr(cu), n, _ptr(out)))
        return out

    def embed_dev(self, d_token_ids, d_cu_seqlens, n_seq, total_tokens, max_seqlen, d_out, stream=0):
        check(lib().lsc_encoder_embed_dev(self._h, _ptr(d_token_ids), _ptr(d_cu_seqlens), n_seq, total_tokens,
                                          max_seqlen, _ptr(d_out), stream))

    @property
    def launch_count(self) -> int:
        return int(lib().lsc_encoder_launch_count(self._h))

    PROFILE_CLASSES = ("embed_ln", "gemm", "attention", "layernorm", "pool_l2", "ffn_fused")

    def submit(self, token_ids, cu_seqlens, out) -> int:
        """lsc_encoder_submit: enqueue a batch and return its ticket.  The three arrays (int32 ids, int32
        cu_seqlens, float32 [n, hidden] out) must stay alive and untouched until wait(ticket)."""
        n = int(len(cu_seqlens) - 1)
        t = _i64(0)
        check(lib().lsc_encoder_submit(self._h, _ptr(token_ids), _ptr(cu_seqlens), n, _ptr(out), C.byref(t)))
        return int(t.value)

    def wait(self, ticket: int) -> None:
        check(lib().lsc_encoder_wait(self._h, int(ticket)))

    def set_profiling(self, enabled: bool):
        check(lib().lsc_encoder_set_profiling(self._h, int(enabled)))

    def read_profile(self):
        """{class: (milliseconds, launches)} accumulated since the previous read."""
        n = len(self.PROFILE_CLASSES)
        ms = _np.zeros(n, _np.float32)
        cnt = _np.zeros(n, _np.int64)
        check(lib().lsc_encoder_read_profile(self._h, _ptr(ms), _ptr(cnt), n))
        return {c: (float(ms[i]), int(cnt[i])) for i, c in enumerate(self.PROFILE_CLASSES)}

    def close(self):
        if self._h:
            lib().lsc_encoder_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Index:
    """One vector-index shard on one GPU (lsc_index)."""

    def __init__(self, dim: int, capacity: int = 0, device: int = 0, _handle=None):
        if _handle is None:
            h = _vp()
            check(lib().lsc_index_create(dim, capacity, device, C.byref(h)))
            _handle = h
        self._h = _handle
        self.dim = dim
        self.device = device

    @classmethod
    def load(cls, path: str, device: int = 0, dim: int | None = None):
        h = _vp()
        check(lib().lsc_index_load(str(path).encode(), device, C.byref(h)))
        ix = cls(dim or 0, device=device, _handle=h)
        return ix

    def save(self, path: str):
        check(lib().lsc_index_save(self._h, str(path).encode()))

    def __len__(self) -> int:
        return int(lib().lsc_index_size(self._h))

    def upsert(self, ids, vecs):
        ids = _np.ascontiguousarray(ids, _np.int64)
        vecs = _np.ascontiguousarray(vecs, _np.float32).reshape(ids.size, -1)
        if vecs.shape[1] != self.dim:
            raise ValueError(f"vectors have dimension {vecs.shape[1]}, index has {self.dim}")
        check(lib().lsc_index_upsert(self._h, _ptr(ids), _ptr(vecs), int(ids.size)))

    def upsert_dev(self, ids, d_vecs, n, stream=0):
        ids = _np.ascontiguousarray(ids, _np.int64)
        check(lib().lsc_index_upsert_dev(self._h, _ptr(ids), _ptr(d_vecs), int(n), stream))

    def delete(self, ids):
        ids = _np.ascontiguousarray(ids, _np.int64)
        check(lib().lsc_index_delete(self._h, _ptr(ids), int(ids.size)))

    def search(self, queries, k: int):
        q = _np.ascontiguousarray(queries, _np.float32).reshape(-1, self.dim)
        B = q.shape[0]
        ids = _np.empty((B, k), _np.int64)
        scores = _np.empty((B, k), _np.float32)
        check(lib().lsc_index_search(self._h, _ptr(q), B, k, _ptr(ids), _ptr(scores)))
        return ids, scores

    def search_dev(self, d_queries, B, k, d_out_ids, d_out_scores, stream=0):
        check(lib().lsc_index_search_dev(self._h, _ptr(d_queries), B, k, _ptr(d_out_ids), _ptr(d_out_scores), stream))

    def search_sharded_dev(self, exchange, d_queries, B, k, d_out_ids, d_out_scores, stream=0):
        """This shard's part of a search over a row-sharded table; every rank of `exchange`'s group makes the same
        call and every rank receives the merged [B, k] result (lsc_index_search_sharded_dev)."""
        check(lib().lsc_index_search_sharded_dev(self._h, exchange._h, _ptr(d_queries), B, k, _ptr(d_out_ids),
                                                 _ptr(d_out_scores), stream))

    def reserve(self, B: int, k: int):
        """Allocate the scratch of a later search of up to B queries / k results now (lsc_index_reserve)."""
        check(lib().lsc_index_reserve(self._h, int(B), int(k)))

    def check(self, stream=0):
        """Synchronise `stream` and raise if a wait inside an earlier search gave up (lsc_index_check)."""
        check(lib().lsc_index_check(self._h, stream))

    def scores_naive_dev(self, d_queries, B, d_scores, stream=0):
        check(lib().lsc_index_scores_naive_dev(self._h, _ptr(d_queries), B, _ptr(d_scores), stream))

    @property
    def launch_count(self) -> int:
        return int(lib().lsc_index_launch_count(self._h))

    PROFILE_CLASSES = ("normalize", "scan", "merge")

    def set_profiling(self, enabled: bool):
        check(lib().lsc_index_set_profiling(self._h, int(enabled)))

    def read_profile(self):
        n = len(self.PROFILE_CLASSES)
        ms = _np.zeros(n, _np.float32)
        cnt = _np.zeros(n, _np.int64)
        check(lib().lsc_index_read_profile(self._h, _ptr(ms), _ptr(cnt), n))
        return {c: (float(ms[i]), int(cnt[i])) for i, c in enumerate(self.PROFILE_CLASSES)}

    def close(self):
        if self._h and getattr(self, "_owned", True):
            lib().lsc_index_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Exchange:
    """One GPU's end of the sharded search's candidate exchange (lsc_exchange)."""

    HANDLE_BYTES = 64

    def __init__(self, device: int, rank: int, world: int, max_queries: int = 1024):
        h = _vp()
        check(lib().lsc_exchange_create(device, rank, world, max_queries, C.byref(h)))
        self._h = h
        self.rank, self.world, self.device = rank, world, device

    def ipc_handle(self) -> bytes:
        buf = C.create_string_buffer(self.HANDLE_BYTES)
        check(lib().lsc_exchange_ipc_handle(self._h, buf))
        return buf.raw

    def connect_ipc(self, handles) -> None:
        """handles: every rank's ipc_handle() bytes, in rank order."""
        blob = b"".join(handles)
        if len(blob) != self.world * self.HANDLE_BYTES:
            raise ValueError("one 64-byte handle per rank is needed")
        check(lib().lsc_exchange_connect_ipc(self._h, blob))

    @staticmethod
    def connect_local(endpoints) -> None:
        arr = (_vp * len(endpoints))(*[e._h for e in endpoints])
        check(lib().lsc_exchange_connect_local(arr, len(endpoints)))

    def close(self):
        if self._h:
            lib().lsc_exchange_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


class ShardedIndex:
    """A table spread over several GPUs behind one handle and one host thread (lsc_sharded_index)."""

    def __init__(self, dim: int, devices, capacity_per_shard: int = 0):
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        h = _vp()
        check(lib().lsc_sharded_index_create(dim, capacity_per_shard, devs, len(devices), C.byref(h)))
        self._h = h
        self.dim = dim
        self.devices = list(devices)

    def __len__(self) -> int:
        return int(lib().lsc_sharded_index_size(self._h))

    def upsert(self, ids, vecs):
        ids = _np.ascontiguousarray(ids, _np.int64)
        vecs = _np.ascontiguousarray(vecs, _np.float32).reshape(ids.size, -1)
        if vecs.shape[1] != self.dim:
            raise ValueError(f"vectors have dimension {vecs.shape[1]}, index has {self.dim}")
        check(lib().lsc_sharded_index_upsert(self._h, _ptr(ids), _ptr(vecs), int(ids.size)))

    def delete(self, ids):
        ids = _np.ascontiguousarray(ids, _np.int64)
        check(lib().lsc_sharded_index_delete(self._h, _ptr(ids), int(ids.size)))

    def search(self, queries, k: int):
        q = _np.ascontiguousarray(queries, _np.float32).reshape(-1, self.dim)
        B = q.shape[0]
        ids = _np.empty((B, k), _np.int64)
        scores = _np.empty((B, k), _np.float32)
        check(lib().lsc_sharded_index_search(self._h, _ptr(q), B, k, _ptr(ids), _ptr(scores)))
        return ids, scores

    def shard(self, i: int) -> "Index":
        """The i-th shard as a borrowed Index (closing it does not destroy the shard)."""
        h = lib().lsc_sharded_index_shard(self._h, i)
        if not h:
            raise IndexError(i)
        ix = Index.__new__(Index)
        ix._h, ix.dim, ix.device = _vp(h), self.dim, self.devices[i]
        ix._owned = False
        return ix

    def close(self):
        if self._h:
            lib().lsc_sharded_index_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def topk_merge_dev(d_ids, d_scores, parts, B, k, d_out_ids, d_out_scores, stream=0):
    check(lib().lsc_topk_merge_dev(_ptr(d_ids), _ptr(d_scores), parts, B, k, _ptr(d_out_ids), _ptr(d_out_scores), stream))


def topk_search_static_dev(d_rows_bf16, d_row_ids, n_rows, dim, d_queries, B, k, d_out_ids, d_out_scores, stream=0):
    """Top-k for callers that must not trigger a device allocation (GPU memory fully reserved by a
    co-tenant framework, or a latency budget that rules out cudaMalloc's device-wide sync): scratch
    comes from a fixed array inside the library.  B <= 128, k <= 32, dim <= 512; one call at a time
    per GPU (calls share the scratch)."""
    check(lib().lsc_topk_search_static_dev(_ptr(d_rows_bf16), _ptr(d_row_ids), int(n_rows), int(dim),
                                           _ptr(d_queries), int(B), int(k), _ptr(d_out_ids),
                                           _ptr(d_out_scores), stream))


def debug_attention_dev(d_qkv, d_cu_seqlens, n_seq, total_tokens, hidden, heads, max_seqlen, d_ctx, stream=0):
    check(lib().lsc_debug_attention_dev(_ptr(d_qkv), _ptr(d_cu_seqlens), n_seq, total_tokens, hidden, heads,
                                        max_seqlen, _ptr(d_ctx), stream))


def debug_gemm_dev(d_A, d_W, d_bias, d_residual, d_C, M, N, K, epilogue, impl, stream=0):
    check(lib().lsc_debug_gemm_dev(_ptr(d_A), _ptr(d_W), _ptr(d_bias), _ptr(d_residual), _ptr(d_C), M, N, K,
                                   epilogue, impl, stream))


def rerank_mmr(doc_vecs, term_freq, doc_len, max_docs: int, lam: float, k1: float, b: float):
    """lsc_rerank_mmr on host arrays: doc_vecs [n, dim] fp32, term_freq [n, t] int32, doc_len [n] int32.
    Returns the picked indices in MMR order."""
    doc_vecs = _np.ascontiguousarray(doc_vecs, _np.float32)
    n = int(doc_vecs.shape[0])
    dim = int(doc_vecs.shape[1]) if doc_vecs.ndim == 2 and n > 0 else 1
    term_freq = _np.ascontiguousarray(term_freq, _np.int32).reshape(n, -1) if n > 0 else _np.zeros((0, 0), _np.int32)
    t = int(term_freq.shape[1]) if n > 0 else 0
    doc_len = _np.ascontiguousarray(doc_len, _np.int32)
    order = _np.empty(max(n, 1), _np.int32)
    count = _np.zeros(1, _np.int32)
    check(lib().lsc_rerank_mmr(_ptr(doc_vecs), n, dim, _ptr(term_freq), _ptr(doc_len), t, int(max_docs),
                               float(lam), float(k1), float(b), _ptr(order), _ptr(count)))
    return [int(i) for i in order[: int(count[0])]]


def rerank_term_statistics(texts, query_text: str, max_terms: int = 64):
    """lsc_rerank_term_statistics: (term_freq [n, n_terms] int32, doc_len [n] int32) of ReRankAgent.tokenise, natively
    (no GPU).  The same result as langstream_b200.rerank.term_statistics (the Python restatement)."""
    raw = [t.encode("utf-8") if isinstance(t, str) else bytes(t) for t in texts]
    n = len(raw)
    ptrs = (C.c_char_p * max(n, 1))(*raw)
    lens = _np.asarray([len(r) for r in raw], _np.int32)
    tf = _np.zeros((n, max_terms), _np.int32)
    doc_len = _np.zeros(n, _np.int32)
    n_terms = _np.zeros(1, _np.int32)
    q = query_text.encode("utf-8") if isinstance(query_text, str) else bytes(query_text)
    check(lib().lsc_rerank_term_statistics(C.cast(ptrs, _vp), _ptr(lens) if n else None, n, q, len(q), int(max_terms),
                                           _ptr(tf) if n else None, _ptr(doc_len) if n else None, _ptr(n_terms)))
    return tf[:, : int(n_terms[0])].copy(), doc_len


def rerank_mmr_batch_dev(d_doc_vecs, n_queries: int, n_docs: int, dim: int, d_term_freq, d_doc_len, n_terms: int,
                         max_docs: int, lam: float, k1: float, b: float, d_out_order, order_stride: int, d_out_count,
                         count_stride: int, stream=None) -> None:
    """lsc_rerank_mmr_batch_dev: one launch for a batch of queries (device pointers / torch tensors)."""
    check(lib().lsc_rerank_mmr_batch_dev(_ptr(d_doc_vecs), int(n_queries), int(n_docs), int(dim), _ptr(d_term_freq),
                                         _ptr(d_doc_len), int(n_terms), int(max_docs), float(lam), float(k1), float(b),
                                         _ptr(d_out_order), int(order_stride), _ptr(d_out_count), int(count_stride), stream))


class WordPieceTokenizer:
    """lsc_tokenizer_*: the BERT WordPiece pipeline of the HuggingFace `tokenizers` library (what DJL's
    HuggingFaceTokenizer runs in the reference), multi-threaded, output packed for Encoder.embed."""

    def __init__(self, vocab_txt, do_lower_case: bool = True, max_length: int = 512, n_threads: int = 0,
                 cls_token: str | None = None, sep_token: str | None = None, unk_token: str | None = None):
        """cls_token / sep_token / unk_token: template tokens other than [CLS] / [SEP] / [UNK] (MPNet: "<s>", "</s>")."""
        if isinstance(vocab_txt, str):
            vocab_txt = vocab_txt.encode("utf-8")
        h = _vp()
        check(lib().lsc_tokenizer_create(vocab_txt, len(vocab_txt), int(do_lower_case), int(max_length), int(n_threads),
                                         C.byref(h)))
        self._h = h
        self.max_length = int(max_length)
        if cls_token or sep_token or unk_token:
            enc = lambda t: t.encode("utf-8") if t else None  # noqa: E731
            check(lib().lsc_tokenizer_set_special(h, enc(cls_token), enc(sep_token), enc(unk_token)))

    @classmethod
    def from_file(cls, path, **kw):
        with open(path, "rb") as f:
            return cls(f.read(), **kw)

    @property
    def vocab_size(self) -> int:
        return int(lib().lsc_tokenizer_vocab_size(self._h))

    def encode_packed(self, texts):
        """-> (token_ids int32 [total], cu_seqlens int32 [n + 1])"""
        raw = [t.encode("utf-8") if isinstance(t, str) else bytes(t) for t in texts]
        n = len(raw)
        ptrs = (C.c_char_p * max(n, 1))(*raw)
        lens = _np.asarray([len(r) for r in raw], _np.int32)
        ids = _np.empty(max(n * self.max_length, 1), _np.int32)
        cu = _np.zeros(n + 1, _np.int32)
        check(lib().lsc_tokenizer_encode(self._h, C.cast(ptrs, _vp), _ptr(lens) if n else None, n, _ptr(ids),
                                         int(ids.size), _ptr(cu)))
        return ids[: int(cu[-1])].copy(), cu

    def __call__(self, texts):
        """The Tokenizer protocol of embeddings.py: a list of id lists."""
        ids, cu = self.encode_packed(texts)
        return [ids[cu[i]: cu[i + 1]].tolist() for i in range(len(cu) - 1)]

    def close(self):
        if self._h:
            lib().lsc_tokenizer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
