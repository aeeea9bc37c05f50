"""MMR re-rank on device against the oracle's restatement of ReRankAgent and the
reference's known-answer pipeline test (RerankAgentRunnerIT.java:62-92, 204-245)."""
import json
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).parent / "golden"

WORDS = ["alpha", "beta", "gamma", "delta", "what", "is", "a", "question", "this", "text", "vector", "db"]


def _random_case(rng, n, dim, n_query_terms):
    texts = []
    for _ in range(n):
        length = int(rng.integers(1, 12))
        texts.append(" ".join(rng.choice(WORDS, size=length)))
    vecs = rng.standard_normal((n, dim)).astype(np.float32)
    if n >= 4:
        vecs[1] = vecs[0]          # duplicate embedding: cosine exactly 1
        vecs[2] = 0.0              # zero vector: cosine defined as 0 by the reference
        texts[3] = texts[0]        # identical text: identical BM25 score, the first one must win
    query = " ".join(rng.choice(WORDS, size=n_query_terms))
    return texts, vecs, query


@pytest.mark.parametrize("n,dim,terms,max_docs,lam", [
    (1, 5, 3, 8, 0.5), (5, 5, 4, 8, 0.5), (20, 384, 5, 8, 0.5), (20, 384, 5, 20, 0.3), (64, 64, 9, 100, 0.7),
    (100, 384, 6, 100, 0.5), (128, 384, 12, 40, 0.9), (10, 384, 1, 3, 0.0), (10, 384, 2, 10, 1.0)])
def test_mmr_order_matches_oracle(native_lib, cuda_device, n, dim, terms, max_docs, lam):
    from langstream_b200 import _native
    from langstream_b200.rerank import term_statistics
    from oracle import rerank_oracle

    rng = np.random.default_rng(1000 + n * 7 + terms)
    texts, vecs, query = _random_case(rng, n, dim, terms)
    k1, b = 1.5, 0.75
    expected = rerank_oracle.rerank_mmr(list(zip(texts, vecs)), max_docs, query, lam, k1, b)
    tf, doc_len = term_statistics(texts, query)
    got = _native.rerank_mmr(vecs, tf, doc_len, max_docs, lam, k1, b)
    assert got == expected
    # twice in a row: no state survives a call
    assert _native.rerank_mmr(vecs, tf, doc_len, max_docs, lam, k1, b) == expected


def test_rerank_it_known_answer_on_device(native_lib, cuda_device):
    """The reference's own expected list: text1, 9, 0, 8, 7, 2, 6, 3 (RerankAgentRunnerIT.java:230-245)."""
    from langstream_b200.rerank import CudaReRankAgent
    from langstream_b200.vector import Record

    fx = json.loads((GOLDEN / "rerank_it.json").read_text())
    q = np.asarray(fx["query_embeddings"], np.float32)
    rows = fx["rows"]
    # what the cosine query returns (LIMIT 20 >= 10 rows): all rows, by descending cosine, ties by insertion
    cos = [float(np.dot(np.float32(r["embeddings_vector"]), q) /
                 (np.linalg.norm(np.float32(r["embeddings_vector"])) * np.linalg.norm(q))) for r in rows]
    related = [rows[i] for i in sorted(range(len(rows)), key=lambda i: (-cos[i], i))]
    related = [{"text": r["text"], "embeddings_vector": r["embeddings_vector"]} for r in related]
    agent = CudaReRankAgent()
    m = fx["mmr"]
    agent.init({"max": m["max"], "field": "value.related_documents", "query-text": "value.question",
                "query-embeddings": "value.question_embeddings", "output-field": "value.related_documents",
                "text-field": "record.text", "embeddings-field": "record.embeddings_vector", "algorithm": "MMR",
                "lambda": m["lambda"], "k1": m["k1"], "b": m["b"]})
    rec = Record(value=json.dumps({"question": fx["query_text"], "question_embeddings": fx["query_embeddings"],
                                   "related_documents": related}))
    out = agent.process_record(rec)
    assert len(out) == 1
    assert json.loads(out[0].value)["related_documents"] == fx["expected_related_documents"]


def test_pass_through_and_limits(native_lib, cuda_device):
    from langstream_b200 import _native
    from langstream_b200.rerank import CudaReRankAgent

    docs = [{"text": f"t{i}", "embeddings_vector": [1.0, float(i)]} for i in range(4)]
    base = {"field": "value.docs", "output-field": "value.docs", "text-field": "record.text",
            "embeddings-field": "record.embeddings_vector"}
    a = CudaReRankAgent()
    a.init(dict(base, algorithm="none"))
    assert a.rerank(docs, "some query") == docs              # ReRankAgent.java:150-151
    a.init(dict(base, algorithm="MMR"))
    assert a.rerank(docs, "") == docs and a.rerank(docs, None) == docs   # :143-145
    assert len(a.rerank(docs, "t1 t2")) == 4
    a.init(dict(base, algorithm="MMR", max=2))
    assert len(a.rerank(docs, "t1 t2")) == 2
    with pytest.raises(ValueError):
        a.rerank([{"text": "x", "embeddings_vector": None}], "q")    # :120-122
    with pytest.raises(RuntimeError, match="128"):
        _native.rerank_mmr(np.zeros((129, 4), np.float32), np.zeros((129, 1), np.int32), np.ones(129, np.int32), 5, 0.5, 1.5, 0.75)


def test_device_pointer_variant(native_lib, cuda_device):
    import ctypes
    import torch
    from langstream_b200 import _native
    from langstream_b200.rerank import term_statistics

    rng = np.random.default_rng(5)
    texts, vecs, query = _random_case(rng, 30, 384, 4)
    tf, doc_len = term_statistics(texts, query)
    expected = _native.rerank_mmr(vecs, tf, doc_len, 10, 0.5, 1.5, 0.75)
    dev = torch.device("cuda:0")
    d_v, d_tf, d_len = torch.from_numpy(vecs).to(dev), torch.from_numpy(tf).to(dev), torch.from_numpy(doc_len).to(dev)
    d_order = torch.full((30,), -1, dtype=torch.int32, device=dev)
    d_count = torch.zeros(1, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    rc = _native.lib().lsc_rerank_mmr_dev(d_v.data_ptr(), 30, 384, d_tf.data_ptr(), d_len.data_ptr(), tf.shape[1], 10,
                                          ctypes.c_double(0.5), ctypes.c_double(1.5), ctypes.c_double(0.75),
                                          d_order.data_ptr(), d_count.data_ptr(), st)
    assert rc == 0
    torch.cuda.synchronize()
    assert int(d_count.item()) == 10 and d_order[:10].tolist() == expected


def test_batch_of_queries_in_one_launch(native_lib, cuda_device):
    """lsc_rerank_mmr_batch_dev: 37 queries with 20 documents each in one launch pick exactly what 37 single calls pick
    (term statistics from the native lsc_rerank_term_statistics, padded to a common number of query terms)."""
    import torch
    from langstream_b200 import _native

    rng = np.random.default_rng(9)
    B, n, dim, T = 37, 20, 384, 8
    V = np.zeros((B, n, dim), np.float32)
    TF = np.zeros((B, n, T), np.int32)
    DL = np.zeros((B, n), np.int32)
    expected = []
    for q in range(B):
        texts, vecs, query = _random_case(rng, n, dim, int(rng.integers(1, T + 1)))
        tf, dl = _native.rerank_term_statistics(texts, query, max_terms=T)
        V[q], DL[q] = vecs, dl
        TF[q, :, :tf.shape[1]] = tf   # zero columns = query terms no document contains: idf x 0, no effect on the order
        expected.append(_native.rerank_mmr(vecs, TF[q], dl, 5, 0.5, 1.5, 0.75))
    dev = torch.device("cuda:0")
    d_v, d_tf, d_dl = torch.from_numpy(V).to(dev), torch.from_numpy(TF).to(dev), torch.from_numpy(DL).to(dev)
    d_out = torch.full((B, n + 1), -7, dtype=torch.int32, device=dev)   # per query: order, then the count
    _native.rerank_mmr_batch_dev(d_v, B, n, dim, d_tf, d_dl, T, 5, 0.5, 1.5, 0.75, d_out, n + 1, d_out[:, n:], n + 1,
                                 torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out = d_out.cpu().numpy()
    for q in range(B):
        assert int(out[q, n]) == 5 and out[q, :5].tolist() == expected[q]
