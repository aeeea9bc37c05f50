"""Host-side mirrors of the reference SPI (CPU: fakes built on the oracle stand in for the GPU handles).
The flows follow the reference's own tests: ComputeAIEmbeddingsTest / HuggingFaceEmbeddingServiceTest
(index alignment, dimension), JdbcWriterTest (upsert / update / delete round trip) and
RerankAgentRunnerIT (sink -> cosine SQL -> MMR)."""
import json
from pathlib import Path

import numpy as np
import pytest

from fakes import FakeEncoder, FakeIndex

GOLDEN = Path(__file__).parent / "golden"


@pytest.fixture()
def fake_tables():
    from langstream_b200 import vector

    yield lambda d, dev: FakeIndex(d, dev)
    for name in list(vector._TABLES):
        vector.drop_table(name)


def test_provider_lookup_and_config_errors():
    from langstream_b200.embeddings import CudaServiceProviderProvider

    p = CudaServiceProviderProvider()
    assert p.supports({"cuda": {"device": 0}}) and not p.supports({"huggingface": {}})
    sp = p.create_implementation({"cuda": {"device": 0}})
    with pytest.raises(ValueError):
        sp.get_completions_service({})          # like the no-op provider: no completions here
    with pytest.raises(ValueError):
        sp.get_embeddings_service({"model": "no-such-model"})
    with pytest.raises(PermissionError):       # ALLOWED_HF_URLS default: file:// and the DJL zoo prefix
        sp.get_embeddings_service({"model": "all-MiniLM-L6-v2", "model-url": "http://example.com/w.npy"})


def test_allowed_hf_urls_prefix_list(monkeypatch):
    """AbstractHuggingFaceEmbeddingService.java:55-70,174-181: comma-separated prefixes from ALLOWED_HF_URLS."""
    from langstream_b200 import embeddings as E

    monkeypatch.delenv("ALLOWED_HF_URLS", raising=False)
    assert E.allowed_url_prefixes() == ["file://", "djl://ai.djl.huggingface.pytorch"]
    E.check_if_url_is_allowed("file:///models/w.npy")
    E.check_if_url_is_allowed("djl://ai.djl.huggingface.pytorch/sentence-transformers/all-MiniLM-L6-v2")
    with pytest.raises(PermissionError, match="modelUrl is not allowed"):
        E.check_if_url_is_allowed("s3://bucket/w.npy")
    monkeypatch.setenv("ALLOWED_HF_URLS", "file://,s3://,https://models.example.com/")
    E.check_if_url_is_allowed("s3://bucket/w.npy")
    E.check_if_url_is_allowed("https://models.example.com/minilm")
    with pytest.raises(PermissionError):
        E.check_if_url_is_allowed("https://elsewhere.example.com/minilm")
    sp = E.CudaServiceProviderProvider().create_implementation({"cuda": {}})
    with pytest.raises(ValueError, match="file:// urls only"):   # allowed, but this provider reads local files
        sp.get_embeddings_service({"model": "all-MiniLM-L6-v2", "model-url": "s3://bucket/w.npy"})


def test_embeddings_service_alignment_batching_and_errors():
    from langstream_b200 import weights as W
    from langstream_b200.embeddings import CudaEmbeddingsService, HashingTokenizer
    from oracle.encoder_oracle import bert_numpy

    spec = W.TINY
    blob = W.random_init(spec, 3, True)
    enc = FakeEncoder(spec, blob)
    svc = CudaEmbeddingsService(spec, blob, encoder_factory=lambda: enc, max_batch=64, flush_interval_ms=30)
    texts_a = ["something to embed", "another text"]            # HuggingFaceEmbeddingServiceTest: 2 texts
    texts_b = ["a b c d e", "", "x"]
    fa, fb = svc.compute_embeddings(texts_a), svc.compute_embeddings(texts_b)
    ra, rb = fa.result(timeout=30), fb.result(timeout=30)
    assert len(ra) == 2 and len(rb) == 3 and all(len(v) == spec.hidden for v in ra + rb)
    assert isinstance(ra[0][0], float)                           # List<List<Double>>
    assert enc.calls == [5]                                      # both callers merged into ONE pass
    tok = HashingTokenizer(spec.vocab, spec.max_position)
    ids = tok(texts_a + texts_b)
    cu = np.concatenate([[0], np.cumsum([len(s) for s in ids])]).astype(np.int32)
    want = bert_numpy(spec, blob, np.concatenate(ids).astype(np.int32), cu, dtype=np.float32)
    assert np.allclose(np.array(ra + rb), want, atol=1e-6)       # index-aligned with the inputs
    assert svc.compute_embeddings([]).result(timeout=5) == []
    # a failing batch fails the future, it does not raise (AbstractHuggingFaceEmbeddingService.java:213-223)
    svc.tokenizer = lambda texts: [[1] * (spec.max_position + 5) for _ in texts]
    f = svc.compute_embeddings(["too long"])
    with pytest.raises(RuntimeError, match="failed to compute embeddings"):
        f.result(timeout=30)
    assert "number of tokens is too high" in str(f.exception().__cause__)
    svc.close()


def test_expression_evaluation_subset():
    from langstream_b200.vector import Record, evaluate

    r = Record(key={"name": "doc1"}, value={"text": "t", "chunk": {"id": 7}, "embeddings": [1.5, 2, 3]},
               headers={"h": "v"})
    assert evaluate("value.text", r) == "t" and evaluate("value.chunk.id", r) == 7
    assert evaluate("key.name", r) == "doc1" and evaluate("properties.h", r) == "v"
    assert evaluate("fn:toListOfFloat(value.embeddings)", r) == [1.5, 2.0, 3.0]
    assert evaluate("fn:toListOfFloat(value.missing)", r) is None
    assert evaluate("value.missing.deeper", r) is None
    assert evaluate("value.text", Record(value=json.dumps({"text": "from json"}))) == "from json"
    with pytest.raises(ValueError):
        evaluate("unknown.root", r)


def test_writer_round_trip_like_jdbc_writer_test(fake_tables):
    """JdbcWriterTest.java:49-176: insert a row, query it with the cosine SQL, update it, delete it."""
    from langstream_b200.vector import (CudaVectorDataSourceProvider, CudaVectorDatabaseWriterProvider, Record)

    ds_cfg = {"service": "cuda"}
    wp = CudaVectorDatabaseWriterProvider(fake_tables)
    assert wp.supports(ds_cfg) and not wp.supports({"service": "jdbc"})
    w = wp.create_implementation(ds_cfg)
    w.initialise({"table-name": "documents", "fields": [
        {"name": "name", "expression": "key.name", "primary-key": True},
        {"name": "vector", "expression": "fn:toListOfFloat(value.vector)"},
        {"name": "text", "expression": "value.text"},
        {"name": "chunk_id", "expression": "value.chunk_id"}]})
    vec = [float(i) for i in range(32)]
    rec = Record(key={"name": "do'c1"}, value={"vector": vec, "text": "Lorem ipsum...", "chunk_id": 1})
    assert w.upsert(rec, {}).result(timeout=5) is None
    ds = CudaVectorDataSourceProvider(fake_tables).create_data_source_implementation(ds_cfg)
    ds.initialize(ds_cfg)
    sql = "SELECT chunk_id,name,text FROM documents ORDER BY cosine_similarity(vector,cast(? as float array)) DESC LIMIT 10"
    rows = ds.fetch_data(sql, [vec])
    assert rows == [{"chunk_id": 1, "name": "do'c1", "text": "Lorem ipsum..."}]
    w.upsert(Record(key={"name": "do'c1"}, value={"vector": vec, "text": "Lorem ipsum changed...", "chunk_id": 1}), {}).result(5)
    assert ds.fetch_data(sql, [vec])[0]["text"] == "Lorem ipsum changed..."
    w.upsert(Record(key={"name": "do'c1"}, value=None), {}).result(5)   # null value => delete
    assert ds.fetch_data(sql, [vec]) == []
    # errors complete the future exceptionally instead of raising (JdbcWriter.java:175-177)
    bad = w.upsert(Record(key={"name": "x"}, value={"vector": [1.0, 2.0], "text": "t", "chunk_id": 2}), {})
    assert isinstance(bad.exception(timeout=5), ValueError)
    with pytest.raises(ValueError):
        ds.fetch_data("DROP TABLE documents", [])
    w.upsert(rec, {}).result(5)
    with pytest.raises(ValueError):
        ds.fetch_data("SELECT nope FROM documents ORDER BY cosine_similarity(vector,cast(? as float array)) DESC LIMIT 1", [vec])


def test_rerank_it_pipeline_on_host_mirrors(fake_tables):
    """RerankAgentRunnerIT.java:95-264 end to end: vector-db-sink x10 -> query-vector-db with the
    reference's exact SQL -> MMR re-rank -> the exact expected documents, vectors returned unchanged."""
    from langstream_b200.vector import CudaVectorDataSource, CudaVectorDatabaseWriter, Record
    from oracle import rerank_oracle

    fx = json.loads((GOLDEN / "rerank_it.json").read_text())
    w = CudaVectorDatabaseWriter({"service": "cuda"}, fake_tables)
    w.initialise({"table-name": "documents", "fields": [
        {"name": "filename", "expression": "value.filename", "primary-key": True},
        {"name": "chunk_id", "expression": "value.chunk_id", "primary-key": True},
        {"name": "embeddings_vector", "expression": "fn:toListOfFloat(value.embeddings_vector)"},
        {"name": "text", "expression": "value.text"}]})
    for i, r in enumerate(fx["rows"]):
        msg = {"filename": f"doc{i}.pdf", "chunk_id": 1, "embeddings_vector": r["embeddings_vector"],
               "language": "en", "text": r["text"], "chunk_num_tokens": 10}
        w.upsert(Record(value=json.dumps(msg)), {}).result(5)
    ds = CudaVectorDataSource(fake_tables)
    related = ds.fetch_data(fx["sql"], [fx["query_embeddings"]])
    assert len(related) == 10 and related[0] == {"text": "text1", "embeddings_vector": [1.0, 2.0, 3.0, 4.0, 5.0]}
    docs = [(d["text"], np.array(d["embeddings_vector"], np.float32)) for d in related]
    m = fx["mmr"]
    order = rerank_oracle.rerank_mmr(docs, int(m["max"]), fx["query_text"], m["lambda"], m["k1"], m["b"])
    assert [related[i] for i in order] == fx["expected_related_documents"]
    # the JSON query form returns the same rows plus the similarity column
    q = json.dumps({"table": "documents", "vector": "?", "topK": 3, "select": ["text"]}).replace('"?"', "?")
    top3 = ds.fetch_data(q, [fx["query_embeddings"]])
    assert [r["text"] for r in top3] == [r["text"] for r in related[:3]] and "similarity" in top3[0]


def test_split_by_tokens_covers_everything_once():
    from langstream_b200.distributed import shard_rows, split_by_tokens

    rng = np.random.default_rng(0)
    for world in (1, 2, 3, 8):
        lens = rng.integers(0, 129, size=200)
        cu = np.concatenate([[0], np.cumsum(lens)])
        parts = split_by_tokens(cu, world)
        assert parts[0][0] == 0 and parts[-1][1] == 200
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        tokens = [cu[e] - cu[s] for s, e in parts]
        assert max(tokens) - min(tokens) <= 2 * 128          # balanced to within one sequence or two
        rows = [shard_rows(1000003, world, r) for r in range(world)]
        assert rows[0][0] == 0 and rows[-1][1] == 1000003 and all(a[1] == b[0] for a, b in zip(rows, rows[1:]))
    assert split_by_tokens(np.array([0]), 4) == [(0, 0)] * 4   # empty batch


def test_batched_sink_persistence_and_asset_manager(fake_tables, tmp_path):
    """batched upserts, the table restored from the agent's persistent state
    directory, and a `cuda-table` asset mirroring `jdbc-table` (JdbcAssetsManagerProvider.java:33-144)."""
    from langstream_b200 import vector
    from langstream_b200.vector import (CudaAssetManagerProvider, CudaVectorDataSource,
                                        CudaVectorDatabaseWriterProvider, Record)

    loader = lambda path, dev, dim: FakeIndex.load(path, dev, dim)  # noqa: E731
    cfg = {"table-name": "docs", "persistent-state-directory": str(tmp_path), "fields": [
        {"name": "id", "expression": "key.id", "primary-key": True},
        {"name": "vector", "expression": "fn:toListOfFloat(value.vector)"},
        {"name": "text", "expression": "value.text"}]}
    w = CudaVectorDatabaseWriterProvider(fake_tables, loader).create_implementation({"service": "cuda"})
    w.initialise(cfg)
    rng = np.random.default_rng(3)
    vecs = rng.standard_normal((12, 8)).astype(np.float32)
    recs = [Record(key={"id": i}, value={"vector": vecs[i].tolist(), "text": f"t{i}"}) for i in range(12)]
    recs.insert(6, Record(key={"id": 2}, value=None))                      # delete in the middle of the batch
    recs.append(Record(key={"id": 3}, value={"vector": vecs[0].tolist(), "text": "t3 again"}))  # later row wins
    assert w.upsert_batch(recs).result(5) is None
    ds = CudaVectorDataSource(fake_tables)
    ds.initialize({"service": "cuda"})
    sql = "SELECT id,text FROM docs ORDER BY cosine_similarity(vector, CAST(? AS FLOAT ARRAY)) DESC LIMIT 20"
    got = ds.fetch_data(sql, [vecs[0].tolist()])
    assert len(got) == 11 and {"id": 2, "text": "t2"} not in got
    assert got[0] == {"id": 0, "text": "t0"} and got[1] == {"id": 3, "text": "t3 again"}   # tie: earlier row first
    w.close()                                                                 # saves the table
    assert (tmp_path / "docs.lscidx").exists() and (tmp_path / "docs.rows.json").exists()
    vector.drop_table("docs")
    # a new agent instance restores it
    w2 = CudaVectorDatabaseWriterProvider(fake_tables, loader).create_implementation({"service": "cuda"})
    w2.initialise(cfg)
    assert ds.fetch_data(sql, [vecs[0].tolist()]) == got
    # assets: exists (loaded / on disk), delete, re-create empty
    am = CudaAssetManagerProvider(fake_tables, loader)
    assert am.supports("cuda-table") and not am.supports("jdbc-table")
    a = am.create_instance("cuda-table")
    a.initialize({"config": {"table-name": "docs", "dimensions": 8, "vector-column": "vector",
                             "persistent-state-directory": str(tmp_path)}})
    assert a.asset_exists()
    assert a.delete_asset_if_exists() is True and not a.asset_exists()
    assert a.delete_asset_if_exists() is False
    assert not (tmp_path / "docs.lscidx").exists()
    a.deploy_asset()
    assert a.asset_exists() and ds.fetch_data(sql, [vecs[0].tolist()]) == []


def _fill_docs(index_factory, n=300, dim=64, seed=3):
    from langstream_b200 import vector

    rng = np.random.default_rng(seed)
    vecs = rng.standard_normal((n, dim)).astype(np.float32)
    t = vector.get_table("documents", dim, "vector", 0, index_factory)
    t.upsert_many([(f"d{i}",) for i in range(n)], [{"name": f"d{i}", "vector": vecs[i].tolist(), "text": f"text {i}"} for i in range(n)])
    return vecs


def test_query_micro_batcher_answers_concurrent_callers_with_one_search(fake_tables):
    """many callers, each with ONE query (the shape SingleRecordAgentProcessor.java:31-47
    gives), are answered by a handful of batched searches - and every caller gets exactly the rows its own
    synchronous fetch_data would have returned."""
    import threading
    from langstream_b200.vector import CudaVectorDataSourceProvider

    vecs = _fill_docs(fake_tables)
    sql = "SELECT name,text FROM documents ORDER BY cosine_similarity(vector,cast(? as float array)) DESC LIMIT 5"
    plain = CudaVectorDataSourceProvider(fake_tables).create_data_source_implementation({"service": "cuda"})
    plain.initialize({"service": "cuda"})
    want = [plain.fetch_data(sql, [vecs[i].tolist()]) for i in range(64)]
    ds = CudaVectorDataSourceProvider(fake_tables).create_data_source_implementation({"service": "cuda"})
    ds.initialize({"service": "cuda", "query-batch-size": 32, "query-flush-interval-ms": 50})
    got = [None] * 64
    start = threading.Barrier(64)

    def caller(i):
        start.wait()
        got[i] = ds.fetch_data(sql, [vecs[i].tolist()])

    threads = [threading.Thread(target=caller, args=(i,)) for i in range(64)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert got == want
    assert ds._batcher.queries == 64 and ds._batcher.batches <= 8   # 64 one-query calls, a handful of searches
    # asynchronous form + per-request errors
    futs = [ds.fetch_data_async(sql, [vecs[i].tolist()]) for i in range(10)] + [ds.fetch_data_async(sql, [[1.0, 2.0]])]
    assert [f.result(5) for f in futs[:10]] == want[:10]
    assert isinstance(futs[10].exception(5), ValueError)
    ds.close()
    plain.close()


def test_query_agent_processes_a_list_of_records_with_one_search(fake_tables):
    """AgentProcessor.process(List<Record>, sink): one emit per record, in order, errors per record; the results are
    those of the per-record path (QueryVectorDBAgent.java:66-79)."""
    from langstream_b200.agents import QueryVectorDBAgent
    from langstream_b200.vector import CudaVectorDataSourceProvider, Record

    vecs = _fill_docs(fake_tables)
    cfg = {"datasource": {"service": "cuda"}, "output-field": "value.related",
           "query": "SELECT name FROM documents ORDER BY cosine_similarity(vector,cast(? as float array)) DESC LIMIT 3",
           "fields": ["value.embeddings"]}
    agent = QueryVectorDBAgent([CudaVectorDataSourceProvider(fake_tables)])
    agent.init(cfg)
    records = [Record(key=f"k{i}", value={"q": i, "embeddings": vecs[i].tolist()}) for i in range(20)]
    records[7] = Record(key="bad", value={"q": 7, "embeddings": [1.0, 2.0, 3.0]})   # wrong dimension: this record fails
    emitted = []
    agent.process(records, lambda src, res, err: emitted.append((src, res, err)))
    assert [e[0] for e in emitted] == records
    for i, (src, res, err) in enumerate(emitted):
        if i == 7:
            assert res == [] and isinstance(err, ValueError)
            continue
        assert err is None and len(res) == 1
        assert res[0].value["related"] == agent.process_record(records[i])[0].value["related"]
        assert res[0].value["related"][0] == {"name": f"d{i}"}
    calls = []
    orig = agent.data_source.fetch_data_many
    agent.data_source.fetch_data_many = lambda q, p: (calls.append(len(p)), orig(q, p))[1]
    agent.process(records[:5], lambda *a: None)
    assert calls == [5]
    agent.close()
