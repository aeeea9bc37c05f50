"""The host mirrors on the real CUDA handles: the reference's RerankAgentRunnerIT flow
(vector-db-sink -> query-vector-db with the reference's SQL -> MMR) and the embeddings service."""
import json
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).parent / "golden"


def test_rerank_it_pipeline_on_gpu(native_lib, cuda_device):
    from langstream_b200 import vector
    from langstream_b200.vector import CudaVectorDataSource, CudaVectorDatabaseWriter, Record
    from oracle import rerank_oracle

    fx = json.loads((GOLDEN / "rerank_it.json").read_text())
    vector.drop_table("documents")
    w = CudaVectorDatabaseWriter({"service": "cuda", "device": 0})
    w.initialise({"table-name": "documents", "fields": [
        {"name": "filename", "expression": "value.filename", "primary-key": True},
        {"name": "chunk_id", "expression": "value.chunk_id", "primary-key": True},
        {"name": "embeddings_vector", "expression": "fn:toListOfFloat(value.embeddings_vector)"},
        {"name": "text", "expression": "value.text"}]})
    for i, r in enumerate(fx["rows"]):
        msg = {"filename": f"doc{i}.pdf", "chunk_id": 1, "embeddings_vector": r["embeddings_vector"],
               "language": "en", "text": r["text"], "chunk_num_tokens": 10}
        assert w.upsert(Record(value=json.dumps(msg)), {}).result(30) is None
    ds = CudaVectorDataSource()
    ds.initialize({"service": "cuda"})
    related = ds.fetch_data(fx["sql"], [fx["query_embeddings"]])
    assert len(related) == 10
    assert related[0] == {"text": "text1", "embeddings_vector": [1.0, 2.0, 3.0, 4.0, 5.0]}
    docs = [(d["text"], np.array(d["embeddings_vector"], np.float32)) for d in related]
    m = fx["mmr"]
    # ... -> re-rank on the device (CudaReRankAgent), the oracle's order and the reference's expected list
    from langstream_b200.rerank import CudaReRankAgent
    agent = CudaReRankAgent()
    agent.init({"max": m["max"], "field": "value.related_documents", "query-text": "value.question",
                "query-embeddings": "value.question_embeddings", "output-field": "value.related_documents",
                "text-field": "record.text", "embeddings-field": "record.embeddings_vector", "algorithm": "MMR",
                "lambda": m["lambda"], "k1": m["k1"], "b": m["b"]})
    out = agent.process_record(Record(value={"question": fx["query_text"], "question_embeddings": fx["query_embeddings"],
                                             "related_documents": related}))
    order = rerank_oracle.rerank_mmr(docs, int(m["max"]), fx["query_text"], m["lambda"], m["k1"], m["b"])
    assert out[0].value["related_documents"] == [related[i] for i in order] == fx["expected_related_documents"]
    # delete on null value, then the row is gone
    w.upsert(Record(value=None, key=None), {})  # pk evaluates to (None, None): unknown -> ignored
    assert len(ds.fetch_data(fx["sql"], [fx["query_embeddings"]])) == 10
    vector.drop_table("documents")


def test_embeddings_service_on_gpu(native_lib, cuda_device):
    """HuggingFaceEmbeddingServiceTest.java:27-44: two texts -> two vectors of the model dimension;
    here also checked against the oracle on the same tokens."""
    from langstream_b200 import weights as W
    from langstream_b200.embeddings import CudaServiceProviderProvider, HashingTokenizer
    from oracle.encoder_oracle import bert_numpy

    provider = CudaServiceProviderProvider().create_implementation({"cuda": {"device": 0}})
    svc = provider.get_embeddings_service({"model": "all-MiniLM-L6-v2", "options": {"random-init-seed": 0},
                                           "batch-size": 64, "flush-interval": 5})
    texts = ["something to embed", "A second, somewhat longer text with a few more words in it."]
    futures = [svc.compute_embeddings([t]) for t in texts] + [svc.compute_embeddings(texts)]
    res = [f.result(timeout=60) for f in futures]
    assert [len(r) for r in res] == [1, 1, 2] and all(len(v) == 384 for r in res for v in r)
    assert np.allclose(res[0][0], res[2][0], atol=1e-6) and np.allclose(res[1][0], res[2][1], atol=1e-6)
    spec = W.MINILM_L6
    ids = HashingTokenizer(spec.vocab, spec.max_position)(texts)
    cu = np.concatenate([[0], np.cumsum([len(s) for s in ids])]).astype(np.int32)
    want = bert_numpy(spec, W.random_init(spec, 0), np.concatenate(ids).astype(np.int32), cu, dtype=np.float32)
    got = np.array(res[2], np.float32)
    cos = (got * want).sum(1) / (np.linalg.norm(got, axis=1) * np.linalg.norm(want, axis=1))
    assert cos.min() >= 0.999 and np.abs(got - want).max() <= 2e-3
    provider.close()


def test_gpu_resident_sink_and_persistence(native_lib, cuda_device, tmp_path):
    """SURVEY.md §8(f) row 2 on the CUDA handles: embeddings stay on the device between the encoder and the
    index (`upsert_from_device`), a batched host upsert, save to / restore from the persistent state
    directory, and the `cuda-table` asset."""
    import torch
    from langstream_b200 import vector, weights as W
    from langstream_b200._native import Encoder
    from langstream_b200.vector import CudaAssetManagerProvider, CudaVectorDataSource, CudaVectorTable

    vector.drop_table("chunks")
    spec = W.MINILM_L6
    enc = Encoder(spec, W.random_init(spec, 0))
    n = 96
    ids, cu = W.synthetic_token_ids(n, 32, seed=11, min_len=8)
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    d_out = torch.empty(n, spec.hidden, device=dev)
    enc.embed_dev(torch.from_numpy(ids).to(dev), torch.from_numpy(cu).to(dev), n, int(cu[-1]), 32, d_out, st)
    torch.cuda.synchronize()
    a = CudaAssetManagerProvider().create_instance("cuda-table")
    a.initialize({"config": {"table-name": "chunks", "dimensions": spec.hidden, "vector-column": "embeddings_vector",
                             "persistent-state-directory": str(tmp_path)}})
    assert not a.asset_exists()
    a.deploy_asset()
    t = vector.get_table("chunks")
    half = n // 2
    t.upsert_from_device([(i,) for i in range(half)], [{"chunk": i, "text": f"text {i}"} for i in range(half)],
                         d_out[:half], st)
    host = d_out[half:].cpu().numpy()
    t.upsert_many([(i,) for i in range(half, n)],
                  [{"chunk": i, "text": f"text {i}", "embeddings_vector": host[i - half].tolist()} for i in range(half, n)])
    ds = CudaVectorDataSource()
    ds.initialize({"service": "cuda"})
    sql = "SELECT chunk,text FROM chunks ORDER BY cosine_similarity(embeddings_vector, CAST(? AS FLOAT ARRAY)) DESC LIMIT 3"
    emb = d_out.cpu().numpy()
    for probe in (0, half - 1, half, n - 1):       # every row finds itself first, device-fed or host-fed
        got = ds.fetch_data(sql, [emb[probe].tolist()])
        assert got[0] == {"chunk": probe, "text": f"text {probe}"}
    before = [ds.fetch_data(sql, [emb[p].tolist()]) for p in (3, 70)]
    t.save(tmp_path)
    vector.drop_table("chunks")
    assert a.asset_exists()                        # on disk
    a.deploy_asset()                               # restores it
    assert len(vector.get_table("chunks").index) == n
    assert [ds.fetch_data(sql, [emb[p].tolist()]) for p in (3, 70)] == before
    assert a.delete_asset_if_exists() and not CudaVectorTable.exists_on_disk(tmp_path, "chunks")
    enc.close()


def test_embeddings_service_with_native_tokenizer(native_lib, cuda_device, tmp_path):
    """texts -> lsc_tokenizer_encode -> lsc_encoder_embed: the service with `options.vocab-file`; the vectors
    equal the oracle's on the ids the HuggingFace tokenizer produces for the same texts."""
    import sys
    sys.path.insert(0, str(Path(__file__).parent))
    import tokenizer_corpus as tc
    transformers = pytest.importorskip("transformers")
    from langstream_b200 import weights as W
    from langstream_b200.embeddings import CudaServiceProviderProvider
    from oracle.encoder_oracle import bert_numpy

    vp = tmp_path / "vocab.txt"
    vp.write_text(tc.vocab_text(), encoding="utf-8")
    provider = CudaServiceProviderProvider().create_implementation({"cuda": {"device": 0}})
    svc = provider.get_embeddings_service({"model": "all-MiniLM-L6-v2", "batch-size": 64, "flush-interval": 5,
                                           "options": {"random-init-seed": 0, "vocab-file": str(vp)}})
    texts = ["Hello, World!", "unable unablely the vector data base", tc.FIXED_TEXTS[4], "something to embed"]
    got = np.asarray(svc.compute_embeddings(texts).result(timeout=60))
    hf = transformers.BertTokenizer(str(vp), do_lower_case=True)
    ids = hf(texts)["input_ids"]
    spec = W.MINILM_L6
    blob = W.random_init(spec, 0)
    for i, seq in enumerate(ids):
        ref = bert_numpy(spec, blob, np.asarray(seq, np.int32), np.asarray([0, len(seq)], np.int32))[0]
        cos = float(np.dot(got[i], ref) / (np.linalg.norm(got[i]) * np.linalg.norm(ref)))
        assert cos >= 0.999 and np.abs(got[i] - ref).max() <= 2e-3
    provider.close()


def test_chatbot_flow_through_the_agent_mirrors(native_lib, cuda_device):
    """The reference's docker-chatbot pipeline shape on the CUDA path end to end: compute-ai-embeddings (step with
    batching) -> vector-db-sink -> query-vector-db -> re-rank, every stage through the agent mirrors."""
    from langstream_b200 import vector
    from langstream_b200.agents import ComputeAIEmbeddingsStep, QueryVectorDBAgent
    from langstream_b200.embeddings import CudaServiceProviderProvider
    from langstream_b200.rerank import CudaReRankAgent
    from langstream_b200.vector import CudaVectorDatabaseWriter, Record

    vector.drop_table("chunks2")
    provider = CudaServiceProviderProvider().create_implementation({"cuda": {"device": 0}})
    svc = provider.get_embeddings_service({"model": "all-MiniLM-L6-v2", "options": {"random-init-seed": 0},
                                           "batch-size": 64, "flush-interval": 5})
    step = ComputeAIEmbeddingsStep.from_config({"text": "{{ value.text }}", "embeddings-field": "value.embeddings_vector",
                                                "batch-size": 16, "flush-interval": 10, "concurrency": 2}, svc)
    step.start()
    topics = ["stream", "vector", "kafka", "pulsar", "gpu", "tensor", "index", "query", "agent", "record"]
    docs = [Record(key=f"doc{i}", value={"filename": f"doc{i}.pdf", "chunk_id": i,
                                         "text": f"{topics[i % 10]} chunk number {i} about {topics[(i * 3) % 10]} things"})
            for i in range(40)]
    futs = [step.process_async(r) for r in docs]
    for f in futs:
        f.result(60)
    assert all(len(r.value["embeddings_vector"]) == 384 for r in docs)
    w = CudaVectorDatabaseWriter({"service": "cuda", "device": 0})
    w.initialise({"table-name": "chunks2", "fields": [
        {"name": "filename", "expression": "value.filename", "primary-key": True},
        {"name": "chunk_id", "expression": "value.chunk_id", "primary-key": True},
        {"name": "embeddings_vector", "expression": "fn:toListOfFloat(value.embeddings_vector)"},
        {"name": "text", "expression": "value.text"}]})
    assert w.upsert_batch(docs).result(30) is None
    # a question identical to one chunk's text: that chunk must come back first, and survive the re-rank
    question = Record(key="q", value={"question": docs[17].value["text"]})
    qstep = ComputeAIEmbeddingsStep("{{ value.question }}", "value.question_embeddings", None, 1, 0, 1, svc)
    qstep.process(question)
    agent = QueryVectorDBAgent()
    agent.init({"datasource": {"service": "cuda"},
                "query": "SELECT text,embeddings_vector FROM chunks2 ORDER BY cosine_similarity(embeddings_vector, CAST(? AS FLOAT ARRAY)) DESC LIMIT 20",
                "fields": ["value.question_embeddings"], "output-field": "value.related_documents"})
    out = agent.process_record(question)[0]
    related = out.value["related_documents"]
    assert len(related) == 20 and related[0]["text"] == docs[17].value["text"]
    rr = CudaReRankAgent()
    rr.init({"max": 5, "field": "value.related_documents", "query-text": "value.question",
             "query-embeddings": "value.question_embeddings", "output-field": "value.related_documents",
             "text-field": "record.text", "embeddings-field": "record.embeddings_vector", "algorithm": "MMR"})
    final = rr.process_record(out)[0].value["related_documents"]
    assert len(final) == 5 and final[0]["text"] == docs[17].value["text"]
    step.batch_executor.stop()
    qstep.batch_executor.stop()
    provider.close()
    vector.drop_table("chunks2")


def test_query_micro_batcher_on_gpu(native_lib, cuda_device):
    """64 concurrent one-query callers through CudaVectorDataSource.fetch_data with the micro-batcher on: the same
    rows as the one-at-a-time path, from a handful of searches; and a list of records through the batched
    query-vector-db agent mirror."""
    import threading

    from langstream_b200 import vector
    from langstream_b200.agents import QueryVectorDBAgent
    from langstream_b200.vector import CudaVectorDataSource, Record

    n, dim = 50_000, 384
    rng = np.random.default_rng(21)
    vecs = rng.standard_normal((n, dim)).astype(np.float32)
    t = vector.get_table("qb_docs", dim, "vector", 0)
    try:
        t.upsert_many([(i,) for i in range(n)], [{"id": i, "vector": vecs[i], "text": f"t{i}"} for i in range(n)])
        queries = vecs[:64] + 0.05 * rng.standard_normal((64, dim)).astype(np.float32)
        sql = "SELECT id,text FROM qb_docs ORDER BY cosine_similarity(vector,cast(? as float array)) DESC LIMIT 10"
        plain = CudaVectorDataSource()
        plain.initialize({"service": "cuda"})
        want = [plain.fetch_data(sql, [q]) for q in queries]
        assert all(w[0]["id"] == i for i, w in enumerate(want))
        ds = CudaVectorDataSource()
        ds.initialize({"service": "cuda", "query-batch-size": 64, "query-flush-interval-ms": 20})
        got = [None] * 64
        start = threading.Barrier(64)

        def caller(i):
            start.wait()
            got[i] = ds.fetch_data(sql, [queries[i]])

        threads = [threading.Thread(target=caller, args=(i,)) for i in range(64)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        assert got == want
        assert ds._batcher.batches <= 8
        assert ds.fetch_data_many(sql, [[q] for q in queries]) == want
        agent = QueryVectorDBAgent()
        agent.init({"datasource": {"service": "cuda"}, "output-field": "value.hits", "query": sql, "fields": ["value.e"]})
        emitted = []
        agent.process([Record(key=i, value={"e": queries[i].tolist()}) for i in range(64)],
                      lambda src, res, err: emitted.append((res, err)))
        assert [e[0][0].value["hits"] for e in emitted] == want and all(e[1] is None for e in emitted)
        agent.close()
        ds.close()
        plain.close()
    finally:
        vector.drop_table("qb_docs")
