"""MMR re-rank restatement — TEST INFRASTRUCTURE ONLY.

Restates ReRankAgent (langstream-agents/langstream-ai-agents/src/main/java/ai/langstream/ai/agents/
rerank/ReRankAgent.java): rerankMMR :155-180, calculateTopDocumentMMR_BM25_CosineSimilarity :182-222,
calculateDiversity :224-251, calculateBM25Scores :275-302, calculateIDF :335-343, tokenise :314-316.
Used to pin the top-k oracle against the reference's known-answer pipeline test
(RerankAgentRunnerIT.java:62-92): sink 10 rows -> cosine query -> MMR picks 8 in an exact order that
depends on the cosine ordering of the query result.
"""
from __future__ import annotations

import math
import re
from typing import List, Sequence, Tuple

from .topk_oracle import cosine_scalar


def _tokenise(text: str) -> List[str]:
    """ReRankAgent.tokenise (:314-316) = Java's text.split("\\s+"): trailing empty strings are dropped, a
    leading one (text starting with whitespace) is kept, and the empty text gives one empty token."""
    parts = re.split(r"[ \t\n\x0b\f\r]+", text)  # Java's \s (no UNICODE_CHARACTER_CLASS): exactly these six characters
    while len(parts) > 1 and parts[-1] == "":
        parts.pop()
    if parts == [""] and text != "":  # text of whitespace only: Java returns an empty array
        return []
    return parts


def _bm25(docs: Sequence[Tuple[str, Sequence[float]]], query_text: str, k1: float, b: float) -> List[float]:
    n = len(docs)
    avgdl = sum(len(_tokenise(t)) for t, _ in docs) / n
    tfs = []
    for t, _ in docs:
        tf = {}
        for term in _tokenise(t):
            tf[term] = tf.get(term, 0) + 1
        tfs.append(tf)
    scores = [0.0] * n
    for i, (t, _) in enumerate(docs):
        dl = float(len(_tokenise(t)))
        for term in _tokenise(query_text):
            tf = tfs[i].get(term, 0)
            df = sum(1 for m in tfs if term in m)
            idf = math.log((n - df + 0.5) / (df + 0.5) + 1.0)
            scores[i] += idf * ((tf * (k1 + 1)) / (tf + k1 * (1 - b + b * (dl / avgdl))))
    return scores


def rerank_mmr(docs: Sequence[Tuple[str, Sequence[float]]], max_docs: int, query_text: str,
               lam: float, k1: float, b: float) -> List[int]:
    """Returns indices into `docs` in MMR order."""
    remaining = list(range(len(docs)))
    ranked: List[int] = []
    while remaining and len(ranked) < max_docs:
        cur = [docs[i] for i in remaining]
        bm = _bm25(cur, query_text, k1, b)
        top, top_score = None, -math.inf
        for j, i in enumerate(remaining):
            if not ranked:
                diversity = 0.0
            else:
                diversity = sum(cosine_scalar(docs[i][1], docs[r][1]) for r in ranked) / len(ranked)
            score = lam * bm[j] - (1 - lam) * diversity
            if score > top_score:
                top, top_score = i, score
        ranked.append(top)
        remaining.remove(top)
    return ranked
