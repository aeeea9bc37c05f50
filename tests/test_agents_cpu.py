"""Host mirrors of the agents around the hot path (langstream_b200/agents.py), following the reference's own tests:
ComputeAIEmbeddingsTest.java:44-187 (field placement with a mock service, loop-over) and the ordering contract of
OrderedAsyncBatchExecutor."""
import json
import threading
import time
from concurrent.futures import Future

import pytest

from langstream_b200.agents import ComputeAIEmbeddingsStep, OrderedAsyncBatchExecutor, QueryStep, render_template
from langstream_b200.vector import Record


class MockEmbeddingsService:
    """MockEmbeddingsService of the reference tests: a fixed vector per text."""

    def __init__(self, table=None, fail=False, delay=0.0):
        self.table, self.fail, self.delay, self.calls = dict(table or {}), fail, delay, []
        self.closed = False

    def compute_embeddings(self, texts):
        self.calls.append(list(texts))
        fut: Future = Future()

        def run():
            if self.delay:
                time.sleep(self.delay)
            if self.fail:
                fut.set_exception(RuntimeError("failed to compute embeddings"))
            else:
                fut.set_result([self.table.get(t, [float(len(t))]) for t in texts])

        threading.Thread(target=run, daemon=True).start()
        return fut

    def close(self):
        self.closed = True


def test_value_and_key_fields_like_compute_ai_embeddings_test():
    svc = MockEmbeddingsService({"Jane The Princess ": [1.0, 2.0, 3.0], "key1": [4.0, 5.0]})
    step = ComputeAIEmbeddingsStep("{{ value.firstName }} {{ value.lastName }}", "value.newField", None, 1, 500, 1, svc)
    step.start()
    rec = Record(key="test-key", value={"firstName": "Jane", "lastName": "The Princess "})
    step.process(rec)
    assert rec.value == {"firstName": "Jane", "lastName": "The Princess ", "newField": [1.0, 2.0, 3.0]}
    # JSON text value stays JSON text (testJson); key fields are addressable (testKeyValueAvro)
    step2 = ComputeAIEmbeddingsStep("{{ key.keyField1 }}", "value.newField", None, 1, 500, 1, svc)
    rec2 = Record(key={"keyField1": "key1"}, value=json.dumps({"valueField1": "value1"}))
    step2.process(rec2)
    assert json.loads(rec2.value) == {"valueField1": "value1", "newField": [4.0, 5.0]}
    step.close()
    assert svc.closed


def test_loop_over_like_the_reference():
    svc = MockEmbeddingsService({"Jane The Princess": [1.0, 2.0, 3.0], "George The Prince": [1.0, 5.0, 3.0]})
    step = ComputeAIEmbeddingsStep("{{ record.firstName }} {{ record.lastName }}", "record.newField",
                                   "value.documents_to_retrieve", 1, 0, 1, svc)
    rec = Record(value=json.dumps({"documents_to_retrieve": [{"firstName": "Jane", "lastName": "The Princess"},
                                                             {"firstName": "George", "lastName": "The Prince"}]}))
    step.process(rec)
    assert json.loads(rec.value) == {"documents_to_retrieve": [
        {"firstName": "Jane", "lastName": "The Princess", "newField": [1.0, 2.0, 3.0]},
        {"firstName": "George", "lastName": "The Prince", "newField": [1.0, 5.0, 3.0]}]}
    with pytest.raises(ValueError, match="record.xxx"):
        ComputeAIEmbeddingsStep("{{ record.a }}", "value.newField", "value.docs", 1, 0, 1, svc)
    with pytest.raises(ValueError, match="record.xxx.yyy"):
        ComputeAIEmbeddingsStep("{{ record.a }}", "record.a.b", "value.docs", 1, 0, 1, svc)


def test_batching_order_and_whole_batch_failure():
    svc = MockEmbeddingsService(delay=0.01)
    step = ComputeAIEmbeddingsStep.from_config({"text": "{{ value.t }}", "embeddings-field": "value.e", "batch-size": 4,
                                                "flush-interval": 20, "concurrency": 1}, svc)
    step.start()
    recs = [Record(key="k", value={"t": "x" * (i + 1)}) for i in range(10)]
    futs = [step.process_async(r) for r in recs]
    for f in futs:
        f.result(5)
    assert [r.value["e"] for r in recs] == [[float(i + 1)] for i in range(10)]
    assert [len(c) for c in svc.calls] == [4, 4, 2]                       # two full batches, the rest on the timer
    assert [t for c in svc.calls for t in c] == ["x" * (i + 1) for i in range(10)]   # order kept
    step.close()
    bad = ComputeAIEmbeddingsStep("{{ value.t }}", "value.e", None, 3, 0, 1, MockEmbeddingsService(fail=True))
    r = Record(value={"t": "a"})
    f = bad.process_async(r)
    with pytest.raises(RuntimeError, match="failed to compute embeddings"):
        f.result(5)
    assert "e" not in r.value


def test_ordered_async_batch_executor_buckets():
    seen, lock = [], threading.Lock()

    def processor(batch, handle):
        def run():
            time.sleep(0.005)
            with lock:
                seen.append(list(batch))
            handle.set_result(None)
        threading.Thread(target=run, daemon=True).start()

    ex = OrderedAsyncBatchExecutor(2, processor, 0, 3, lambda item: item[0])
    ex.start()
    for i in range(12):
        ex.add((i % 3 - 1, i))            # hashes -1, 0, 1 -> Math.abs(hash % 3) buckets 1, 0, 1
    ex.stop()
    time.sleep(0.3)
    flat = [x for b in seen for x in b]
    assert sorted(x[1] for x in flat) == list(range(12))
    for h in (-1, 0, 1):                    # per-hash arrival order is preserved
        assert [x[1] for x in flat if x[0] == h] == [i for i in range(12) if i % 3 - 1 == h]
    assert all(len(b) == 1 for b in seen)   # flush-interval <= 0: every add cuts a batch


def test_render_template():
    ctx = {"value": {"a": {"b": 3}, "s": "x"}, "key": json.dumps({"k": "v"}), "record": {"t": "doc"}}
    assert render_template("{{ value.a.b }}-{{value.s}}-{{ key.k }}-{{{ record.t }}}-{{ value.missing }}", ctx) == "3-x-v-doc-"


class FakeDataSource:
    def __init__(self):
        self.queries = []

    def fetch_data(self, query, params):
        self.queries.append((query, params))
        return [{"text": f"r{len(self.queries)}", "p": params[0]}]

    def execute_statement(self, query, generated_keys, params):
        return {"count": 1, "generatedKeys": {k: 7 for k in generated_keys}}


def test_query_step_modes():
    ds = FakeDataSource()
    step = QueryStep.from_config({"query": "SELECT ?", "fields": ["value.q"], "output-field": "value.out"}, ds)
    rec = Record(value={"q": [1.0, 2.0]})
    step.process(rec)
    assert rec.value["out"] == [{"text": "r1", "p": [1.0, 2.0]}]
    first = QueryStep(ds, "SELECT ?", ["value.q"], "value.out", only_first=True)
    first.process(rec)
    assert rec.value["out"] == {"text": "r2", "p": [1.0, 2.0]}
    loop = QueryStep(ds, "SELECT ?", ["record.v"], "value.out", loop_over="value.items")
    rec2 = Record(value=json.dumps({"items": [{"v": 1}, {"v": 2}]}))
    loop.process(rec2)
    assert [r["p"] for r in json.loads(rec2.value)["out"]] == [1, 2]
    ex = QueryStep(ds, "INSERT ?", ["value.q"], "value.res", generated_keys=["id"], mode="execute")
    ex.process(rec)
    assert rec.value["res"] == {"count": 1, "generatedKeys": {"id": 7}}
    with pytest.raises(RuntimeError, match="Unknown mode"):
        QueryStep(ds, "x", [], "value.o", mode="nope").process(rec)


def test_agent_emits_exactly_once_per_record_asynchronously():
    """GenAIToolKitAgent.process (GenAIToolKitAgent.java:70-91): one emit per input record, from whatever thread the
    service completes on, errors as the emit's error; an empty list is an IllegalStateException."""
    from langstream_b200.agents import GenAIToolKitAgent

    svc = MockEmbeddingsService(delay=0.02)
    agent = GenAIToolKitAgent(ComputeAIEmbeddingsStep.from_config(
        {"text": "{{ value.t }}", "embeddings-field": "value.e", "batch-size": 4, "flush-interval": 10, "concurrency": 2}, svc))
    agent.start()
    recs = [Record(key=f"k{i}", value={"t": "y" * (i + 1)}) for i in range(9)]
    emitted, threads, done = [], set(), threading.Event()

    def sink(source, results, error):
        emitted.append((source, results, error))
        threads.add(threading.get_ident())
        if len(emitted) == len(recs):
            done.set()

    agent.process(recs, sink)                     # returns at once: nothing has been computed yet
    assert len(emitted) < len(recs)
    assert done.wait(5)
    assert sorted(id(e[0]) for e in emitted) == sorted(id(r) for r in recs)       # exactly one emit per record
    assert threading.get_ident() not in threads                                   # completed on other threads
    for source, results, error in emitted:
        assert error is None and len(results) == 1
        assert results[0].value["e"] == [float(len(source.value["t"]))]
        assert "e" not in source.value                                             # the source record is not mutated
    assert agent.processed_in == 9 and agent.processed_out == 9
    with pytest.raises(RuntimeError, match="cannot be null or empty"):
        agent.process([], sink)
    agent.close()
    assert svc.closed
    failing = GenAIToolKitAgent(ComputeAIEmbeddingsStep("{{ value.t }}", "value.e", None, 2, 0, 1, MockEmbeddingsService(fail=True)))
    got = []
    ev = threading.Event()
    failing.process([Record(value={"t": "a"}), Record(value={"t": "b"})], lambda s, r, e: (got.append((r, e)), len(got) == 2 and ev.set()))
    assert ev.wait(5)
    assert all(r is None and isinstance(e, RuntimeError) for r, e in got) and failing.processed_out == 0
