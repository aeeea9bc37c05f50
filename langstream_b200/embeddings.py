"""Host-side mirror of the reference's embeddings SPI for the `cuda` provider.

Java is absent from this image, so the classes a maintainer would add to langstream-ai-agents
(sources under java/) are mirrored here in Python with the same names, argument
meaning and error behaviour, on top of the C ABI:

    EmbeddingsService          com/datastax/oss/streaming/ai/embeddings/EmbeddingsService.java:24-37
    ServiceProvider            com/datastax/oss/streaming/ai/services/ServiceProvider.java:24-31
    ServiceProviderProvider    ai/langstream/ai/agents/services/ServiceProviderProvider.java:22-28
    (reference implementation being replaced: HuggingFaceProvider.java:47-139,
     AbstractHuggingFaceEmbeddingService.java:110-223, HuggingFaceEmbeddingService.java:38-63;
     all under langstream-agents/langstream-ai-agents/src/main/java)

Tokenisation stays on the host: the service takes any callable texts -> token-id lists.
"""
from __future__ import annotations

import threading
import time
from abc import ABC, abstractmethod
from concurrent.futures import Future
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

from . import _native
from .weights import EncoderSpec, MINILM_L6, BERT_BASE, MPNET_BASE, random_init

Tokenizer = Callable[[Sequence[str]], List[List[int]]]


class EmbeddingsService(ABC):
    """EmbeddingsService.java:24-37 — `computeEmbeddings(List<String>)` returns a future of one
    vector (list of double) per text, index-aligned with the input."""

    @abstractmethod
    def compute_embeddings(self, texts: Sequence[str]) -> "Future[List[List[float]]]":
        ...

    def close(self) -> None:  # default no-op, as in the interface
        pass


class HashingTokenizer:
    """Deterministic stand-in tokenizer for synthetic data (no vocabulary file is available
    offline): whitespace split, each word hashed into [1000, vocab), [CLS]=101 / [SEP]=102 added,
    truncated to max_length like a HuggingFace tokenizer with truncation=True.  Production uses the
    model's real tokenizer (DJL HuggingFaceTokenizer on the Java side)."""

    def __init__(self, vocab: int = 30522, max_length: int = 512):
        self.vocab, self.max_length = vocab, max_length

    def __call__(self, texts: Sequence[str]) -> List[List[int]]:
        out = []
        span = max(self.vocab - 1000, 1)
        for t in texts:
            ids = [101]
            for w in t.split():
                h = 2166136261
                for ch in w.encode("utf-8"):
                    h = ((h ^ ch) * 16777619) & 0xFFFFFFFF
                ids.append(1000 + h % span if self.vocab > 1000 else h % self.vocab)
            ids = ids[: self.max_length - 1] + [102]
            out.append(ids)
        return out


class CudaEmbeddingsService(EmbeddingsService):
    """`EmbeddingsService` backed by lsc_encoder_embed on one GPU.

    Callers hand over one record at a time in the reference runtime
    (CompositeAgentProcessor.java:225-227; OrderedAsyncBatchExecutor.java:113-118 flushes every add
    when flush-interval is 0), so requests from all callers are merged by a micro-batcher: a batch is
    sent to the GPU when `max_batch` sequences are waiting or the oldest request is `flush_interval_ms`
    old.  Errors never raise from compute_embeddings: they fail the returned future with
    RuntimeError("failed to compute embeddings"), as AbstractHuggingFaceEmbeddingService.java:213-223.
    """

    def __init__(self, spec: EncoderSpec, weights: np.ndarray, tokenizer: Optional[Tokenizer] = None,
                 device: int = 0, max_batch: int = 1024, flush_interval_ms: float = 2.0,
                 encoder_factory=None):
        self.spec = spec
        self.tokenizer = tokenizer or HashingTokenizer(spec.vocab, spec.max_len)
        factory = encoder_factory or (lambda: _native.Encoder(spec, weights, device=device))
        self._encoder = factory()
        self._max_batch = int(max_batch)
        self._flush_s = float(flush_interval_ms) / 1000.0
        self._lock = threading.Condition()
        self._pending: List[tuple] = []  # (token id lists, future, enqueue time)
        self._closed = False
        self._worker = threading.Thread(target=self._run, name="lsc-embeddings-batcher", daemon=True)
        self._worker.start()
        self.embeddings_computed = 0
        self.batches = 0

    # -- EmbeddingsService -----------------------------------------------------------------------
    def compute_embeddings(self, texts: Sequence[str]) -> "Future[List[List[float]]]":
        fut: Future = Future()
        try:
            if self._closed:
                raise RuntimeError("service is closed")
            ids = self.tokenizer(list(texts))
            if len(ids) != len(texts):
                raise ValueError("tokenizer returned a different number of sequences")
        except Exception as e:  # noqa: BLE001 - mirrors the catch-all of the reference
            err = RuntimeError("failed to compute embeddings")
            err.__cause__ = e
            fut.set_exception(err)
            return fut
        if not ids:
            fut.set_result([])
            return fut
        with self._lock:
            self._pending.append((ids, fut, time.monotonic()))
            self._lock.notify()
        return fut

    def close(self) -> None:
        with self._lock:
            self._closed = True
            self._lock.notify()
        self._worker.join(timeout=10)
        self._encoder.close()

    # -- batching --------------------------------------------------------------------------------
    def _run(self) -> None:
        while True:
            with self._lock:
                while not self._pending and not self._closed:
                    self._lock.wait()
                if not self._pending and self._closed:
                    return
                # wait for more work until the window closes or the batch is full
                while not self._closed:
                    waiting = sum(len(p[0]) for p in self._pending)
                    age = time.monotonic() - self._pending[0][2]
                    if waiting >= self._max_batch or age >= self._flush_s:
                        break
                    self._lock.wait(timeout=max(self._flush_s - age, 0.0002))
                batch, n = [], 0
                while self._pending and (not batch or n + len(self._pending[0][0]) <= self._max_batch):
                    item = self._pending.pop(0)
                    batch.append(item)
                    n += len(item[0])
            self._execute(batch)

    def _execute(self, batch) -> None:
        try:
            seqs = [s for ids, _, _ in batch for s in ids]
            lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=len(seqs))
            cu = np.zeros(len(seqs) + 1, np.int32)
            cu[1:] = np.cumsum(lens)
            flat = np.fromiter((t for s in seqs for t in s), dtype=np.int32, count=int(cu[-1]))
            emb = self._encoder.embed(flat, cu)
            self.embeddings_computed += len(seqs)
            self.batches += 1
            o = 0
            for ids, fut, _ in batch:
                # float -> double widening, one fresh list per text (HuggingFaceEmbeddingService.java:52-63);
                # nothing returned aliases native memory
                fut.set_result(emb[o:o + len(ids)].astype(np.float64).tolist())
                o += len(ids)
        except Exception as e:  # noqa: BLE001
            for _, fut, _ in batch:  # the whole batch fails together (ComputeAIEmbeddingsStep.java:215-220)
                err = RuntimeError("failed to compute embeddings")
                err.__cause__ = e
                if not fut.done():
                    fut.set_exception(err)


URL_PREFIXES_SYSTEM_PROP = "ALLOWED_HF_URLS"
DJL_BASE_URL = "djl://ai.djl.huggingface.pytorch"


def allowed_url_prefixes() -> List[str]:
    """AbstractHuggingFaceEmbeddingService.getHuggingFaceAllowedUrlPrefixes (:55-70): the comma-separated
    ALLOWED_HF_URLS from the environment, default "file://,djl://ai.djl.huggingface.pytorch"."""
    import os

    prop = os.environ.get(URL_PREFIXES_SYSTEM_PROP) or ""
    if not prop:
        prop = "file://," + DJL_BASE_URL
    return prop.split(",")


def check_if_url_is_allowed(model_url: str) -> None:
    """checkIfUrlIsAllowed (:174-181): IllegalAccessException("modelUrl is not allowed: ...") in the reference."""
    if not any(model_url.startswith(p) for p in allowed_url_prefixes()):
        raise PermissionError(f"modelUrl is not allowed: {model_url}")


KNOWN_ARCHITECTURES: Dict[str, EncoderSpec] = {
    "all-MiniLM-L6-v2": MINILM_L6,
    "sentence-transformers/all-MiniLM-L6-v2": MINILM_L6,
    "BAAI/bge-base-en-v1.5": BERT_BASE,
    "bge-base": BERT_BASE,
    "all-mpnet-base-v2": MPNET_BASE,
    "sentence-transformers/all-mpnet-base-v2": MPNET_BASE,
}


class CudaServiceProvider:
    """`ServiceProvider` (ServiceProvider.java:24-31) for the `cuda` resource."""

    def __init__(self, provider_configuration: dict):
        self.provider_configuration = dict(provider_configuration or {})
        self._services: List[CudaEmbeddingsService] = []

    def get_embeddings_service(self, additional_configuration: dict) -> EmbeddingsService:
        """Step keys as in HuggingFaceProvider.getEmbeddingsService (HuggingFaceProvider.java:82-139):
        `model` names the architecture, `model-url` (file://…npy) points at a weight blob;
        `options.random-init-seed` builds HuggingFace-default random weights instead (benchmarks).
        `options.vocab-file` (the model's vocab.txt) selects the native WordPiece tokenizer
        (lsc_tokenizer_*: the BERT pipeline of the HuggingFace tokenizers library, multi-threaded), with
        `options.do-lower-case` (default true) and `options.tokenizer-threads` (default: all cores); without it the
        hashing stand-in is used (synthetic data)."""
        cfg = dict(additional_configuration or {})
        model = cfg.get("model")
        if not model or model not in KNOWN_ARCHITECTURES:
            raise ValueError(f"unknown model {model!r}; known: {sorted(KNOWN_ARCHITECTURES)}")
        spec = KNOWN_ARCHITECTURES[model]
        options = cfg.get("options") or {}
        url = cfg.get("model-url")
        if url:
            check_if_url_is_allowed(str(url))
            if not str(url).startswith("file://"):  # allowed by the operator, but this provider only reads local files
                raise ValueError(f"the cuda provider loads weights from file:// urls only, got {url}")
            blob = np.load(str(url)[len("file://"):]).astype(np.float32)
        else:
            blob = random_init(spec, seed=int(options.get("random-init-seed", 0)))
        tokenizer = None
        if options.get("vocab-file"):
            tokenizer = _native.WordPieceTokenizer.from_file(
                str(options["vocab-file"]), do_lower_case=bool(options.get("do-lower-case", True)),
                max_length=spec.max_len, n_threads=int(options.get("tokenizer-threads", 0)),
                # MPNet's tokenizer is the BERT pipeline with the template <s> A </s>
                cls_token="<s>" if spec.arch == "mpnet" else None, sep_token="</s>" if spec.arch == "mpnet" else None)
            if tokenizer.vocab_size > spec.vocab:
                raise ValueError(f"vocab-file has {tokenizer.vocab_size} tokens, the model {spec.vocab}")
        svc = CudaEmbeddingsService(
            spec, blob, tokenizer=tokenizer, device=int(self.provider_configuration.get("device", 0)),
            max_batch=int(cfg.get("batch-size", 1024)),
            flush_interval_ms=float(cfg.get("flush-interval", 2)),
        )
        self._services.append(svc)
        return svc

    def get_completions_service(self, additional_configuration: dict):
        raise ValueError("the cuda provider does not offer a completions service")

    def close(self) -> None:
        for s in self._services:
            s.close()


class CudaServiceProviderProvider:
    """`ServiceProviderProvider` (ServiceProviderProvider.java:22-28): claims agent configurations
    that carry a `cuda` entry (HuggingFace claims `huggingface`, HuggingFaceProvider.java:52-55)."""

    def supports(self, agent_configuration: dict) -> bool:
        return "cuda" in (agent_configuration or {})

    def create_implementation(self, agent_configuration: dict, metrics_reporter=None) -> CudaServiceProvider:
        return CudaServiceProvider(agent_configuration.get("cuda") or {})
