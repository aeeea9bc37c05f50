"""Host-side mirror of the reference's `re-rank` agent on the CUDA re-rank entry point.

Mirrors ReRankAgent (langstream-agents/langstream-ai-agents/src/main/java/ai/langstream/ai/agents/rerank/
ReRankAgent.java): configuration keys and defaults of `init` (:68-95), `processRecord` (:99-136), the
pass-through rules of `rerank` (:138-153: empty query text or algorithm "none" return the list unchanged),
and `tokenise` (:314-316).  The numeric part - BM25 over the remaining pool, diversity, the MMR selection -
runs in `lsc_rerank_mmr` (csrc/rerank.cu); this module only prepares what the reference derives from the texts:
document lengths in tokens and the frequency of each query token per document.

There is no CPU fallback: without the native library the agent raises.
"""
from __future__ import annotations

import json
import re
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

from . import _native
from .vector import Record, evaluate

ALGORITHM_NONE = "none"
ALGORITHM_MMR = "MMR"

MAX_DOCS_PER_CALL = 128  # lsc_rerank_mmr limit (the reference's default `max` is 100)
MAX_QUERY_TERMS = 64


def java_split_whitespace(text: str) -> List[str]:
    """Java's `text.split("\\\\s+")` (ReRankAgent.tokenise): trailing empty strings are removed, a leading
    empty string (text starting with whitespace) is kept, "" gives [""], all-whitespace gives []."""
    parts = re.split(r"[ \t\n\x0b\f\r]+", text)  # Java's \s without UNICODE_CHARACTER_CLASS: these six characters only
    while len(parts) > 1 and parts[-1] == "":
        parts.pop()
    if parts == [""] and text != "":
        return []
    return parts


def term_statistics(texts: Sequence[str], query_text: str):
    """What calculateBM25Scores (:275-302) reads from the texts: per document its token count and, for
    every query token (in query order, duplicates included), how often the document contains it."""
    q_terms = java_split_whitespace(query_text)
    n, t = len(texts), len(q_terms)
    tf = np.zeros((n, t), np.int32)
    doc_len = np.zeros(n, np.int32)
    for i, text in enumerate(texts):
        terms = java_split_whitespace(text)
        doc_len[i] = len(terms)
        counts: Dict[str, int] = {}
        for term in terms:
            counts[term] = counts.get(term, 0) + 1
        for j, term in enumerate(q_terms):
            tf[i, j] = counts.get(term, 0)
    return tf, doc_len


def _to_array_of_float(value) -> Optional[np.ndarray]:
    """ReRankAgent.toArrayOfFloat (:345-359): None stays None, a collection becomes float[]."""
    if value is None:
        return None
    if isinstance(value, (str, bytes)) or not hasattr(value, "__iter__"):
        raise ValueError(f"Cannot convert {value!r} to list of float")
    return np.asarray([np.float32(v) for v in value], np.float32)


class CudaReRankAgent:
    """`type: "re-rank"` with the reference's configuration keys."""

    def __init__(self):
        self.field = self.output_field = None
        self.algorithm = ALGORITHM_NONE
        self.max = 100
        self.mmr_lambda, self.bm25_k1, self.bm25_b = 0.5, 1.5, 0.75
        self.query_text = self.query_embeddings = self.text_field = self.embeddings_field = ""

    def init(self, configuration: Dict[str, Any]) -> None:
        def required(key):
            v = configuration.get(key)
            if v is None or str(v) == "":
                raise ValueError(f"Missing required field '{key}' in re-rank agent")
            return str(v)

        self.field = required("field")
        self.output_field = required("output-field")
        self.algorithm = str(configuration.get("algorithm", ALGORITHM_NONE))
        self.query_embeddings = str(configuration.get("query-embeddings", ""))
        self.query_text = str(configuration.get("query-text", ""))
        self.embeddings_field = str(configuration.get("embeddings-field", ""))
        self.text_field = str(configuration.get("text-field", ""))
        self.max = int(configuration.get("max", 100))
        self.mmr_lambda = float(configuration.get("lambda", 0.5))
        self.bm25_k1 = float(configuration.get("k1", 1.5))
        self.bm25_b = float(configuration.get("b", 0.75))

    # -- the reference's processRecord: read the list, re-rank, write it to output-field ---------------
    def process_record(self, record: Optional[Record]) -> List[Record]:
        if record is None:
            return []
        value = record.value
        as_text = isinstance(value, (str, bytes))
        obj = json.loads(value) if as_text else dict(value)
        work = Record(key=record.key, value=obj, headers=dict(record.headers))
        current = evaluate(self.field, work)
        current = list(current) if current is not None else []
        query_text = evaluate(self.query_text, work) if self.query_text else None
        query_emb = _to_array_of_float(evaluate(self.query_embeddings, work)) if self.query_embeddings else None
        result = self.rerank(current, query_text, query_emb)
        self._set_result(obj, self.output_field, result)
        out_value = json.dumps(obj) if as_text else obj
        return [Record(key=record.key, value=out_value, headers=dict(record.headers))]

    @staticmethod
    def _set_result(obj: Dict[str, Any], field: str, result) -> None:
        parts = field.split(".")
        if parts[0] != "value" or len(parts) < 2:
            raise ValueError(f"output-field {field!r}: only value.<field> is supported")
        cur = obj
        for p in parts[1:-1]:
            cur = cur.setdefault(p, {})
        cur[parts[-1]] = result

    def rerank(self, current: List[Any], query_text: Optional[str], query_embeddings=None) -> List[Any]:
        if query_text is None or query_text == "":
            return list(current)
        if self.algorithm == ALGORITHM_NONE:
            return list(current)
        if self.algorithm != ALGORITHM_MMR:
            raise RuntimeError("IllegalStateException")  # the reference's default branch
        if not current or self.max <= 0:
            return []
        texts, vecs = [], []
        for o in current:
            ctx = Record(value=o)
            text = evaluate(self.text_field, ctx)
            emb = _to_array_of_float(evaluate(self.embeddings_field, ctx))
            if emb is None:
                raise ValueError(f"Embeddings are null in record: {o}")
            if text is None:
                raise ValueError(f"Text is null in record: {o}")
            if text == "" or emb.size == 0:
                raise ValueError("re-rank: documents with empty text or embeddings are not supported")
            texts.append(str(text))
            vecs.append(emb)
        if len({v.size for v in vecs}) != 1:
            raise ValueError("Arrays must have the same length")  # dotProduct's check (:361-364)
        if len(current) > MAX_DOCS_PER_CALL:
            raise ValueError(f"re-rank: at most {MAX_DOCS_PER_CALL} documents per record")
        try:  # ReRankAgent.tokenise + the term counts, natively (term_statistics above is the Python restatement of it)
            tf, doc_len = _native.rerank_term_statistics(texts, query_text, MAX_QUERY_TERMS)
        except _native.NativeError as e:
            if e.code == _native.LSC_E_UNSUPPORTED:
                raise ValueError(f"re-rank: at most {MAX_QUERY_TERMS} query terms") from e
            raise
        order = _native.rerank_mmr(np.stack(vecs), tf, doc_len, self.max, self.mmr_lambda, self.bm25_k1, self.bm25_b)
        return [current[i] for i in order]

    def close(self) -> None:
        pass
