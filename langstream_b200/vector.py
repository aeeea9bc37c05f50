"""Host-side mirror of the reference's vector datasource / sink SPI for the local GPU store.

Mirrors (same names, argument meaning, error behaviour; paths under the reference tree):
    QueryStepDataSource        langstream-agents/langstream-ai-agents/src/main/java/com/datastax/oss/streaming/ai/datasource/QueryStepDataSource.java:22-37
    DataSourceProvider         …/ai/langstream/ai/agents/datasource/DataSourceProvider.java:21-26
    VectorDatabaseWriter       langstream-api/src/main/java/ai/langstream/api/database/VectorDatabaseWriter.java:26-39
    VectorDatabaseWriterProvider  …/database/VectorDatabaseWriterProvider.java:20-25
replacing JdbcDataSourceProvider.JdbcDataSourceImpl.fetchData (…/datasource/impl/JdbcDataSourceProvider.java:103-124)
and JdbcWriter.JdbcVectorDatabaseWriter (langstream-agents/langstream-vector-agents/src/main/java/ai/langstream/agents/vector/jdbc/JdbcWriter.java:60-179).

The GPU holds the L2-normalised bf16 copy of the vector column (what the scan needs); every other
column — and the original vector, which `SELECT … embeddings_vector` must return unchanged
(RerankAgentRunnerIT.java:62-92) — lives in a host row store keyed by the primary key.
"""
from __future__ import annotations

import json
import re
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple

import numpy as np

from . import _native


@dataclass
class Record:
    """ai.langstream.api.runner.code.Record: key, value (None = tombstone), headers."""
    key: Any = None
    value: Any = None
    headers: Dict[str, Any] = field(default_factory=dict)


# ---------------------------------------------------------------------------------------------
# expression evaluation: the subset of the JSTL language the vector agents' examples use
# (value.x.y, key.x, properties.x, fn:toListOfFloat(expr), literals)
# ---------------------------------------------------------------------------------------------
def _as_mapping(obj):
    if isinstance(obj, (bytes, str)):
        try:
            return json.loads(obj)
        except Exception:  # noqa: BLE001
            return obj
    return obj


def evaluate(expression: str, record: Record):
    expr = expression.strip()
    m = re.fullmatch(r"fn:toListOfFloat\((.*)\)", expr)
    if m:
        # JstlFunctions.toListOfFloat (…/commons/jstl/JstlFunctions.java:93-112): null stays null
        inner = evaluate(m.group(1), record)
        if inner is None:
            return None
        return [float(np.float32(v)) for v in inner]
    if expr.startswith("[") and expr.endswith("]"):
        return json.loads(expr)
    if re.fullmatch(r"-?\d+", expr):
        return int(expr)
    if re.fullmatch(r"-?\d*\.\d+", expr):
        return float(expr)
    if len(expr) >= 2 and expr[0] == expr[-1] and expr[0] in "'\"":
        return expr[1:-1]
    parts = expr.split(".")
    root = {"value": record.value, "key": record.key, "properties": record.headers,
            "record": record.value}.get(parts[0], None)
    if parts[0] not in ("value", "key", "properties", "record"):
        raise ValueError(f"cannot evaluate expression {expression!r}")
    cur = _as_mapping(root)
    for p in parts[1:]:
        if cur is None:
            return None
        cur = _as_mapping(cur)
        cur = cur.get(p) if isinstance(cur, dict) else None
    return cur


# ---------------------------------------------------------------------------------------------
# the store: one GPU index + host rows per table
# ---------------------------------------------------------------------------------------------
class CudaVectorTable:
    def __init__(self, name: str, dim: int, vector_column: str, device: int = 0,
                 index_factory: Optional[Callable[[int, int], Any]] = None):
        self.name, self.dim, self.vector_column = name, dim, vector_column
        # the scan kernel wants a multiple of 64 dimensions: zero padding leaves every cosine unchanged
        self.store_dim = (dim + 63) // 64 * 64
        make = index_factory or (lambda d, dev: _native.Index(d, device=dev))
        self.index = make(self.store_dim, device)
        self.rows: Dict[int, dict] = {}          # internal id -> full row (all columns, original vector)
        self.id_of_pk: Dict[Tuple, int] = {}
        self._next_id = 0
        self.lock = threading.Lock()

    def upsert(self, pk: Tuple, row: dict) -> None:
        vec = row.get(self.vector_column)
        if vec is None:
            raise ValueError(f"column {self.vector_column!r} is null")
        v = np.asarray(vec, np.float32).reshape(-1)
        if v.size != self.dim:
            raise ValueError(f"vector has {v.size} dimensions, table {self.name!r} has {self.dim}")
        with self.lock:
            rid = self.id_of_pk.get(pk)
            if rid is None:
                rid = self._next_id
                self._next_id += 1
                self.id_of_pk[pk] = rid
            self.rows[rid] = dict(row)
            self.index.upsert(np.array([rid], np.int64), self._pad(v)[None, :])

    def _assign_ids(self, pks: List[Tuple]) -> np.ndarray:
        ids = np.empty(len(pks), np.int64)
        for i, pk in enumerate(pks):
            rid = self.id_of_pk.get(pk)
            if rid is None:
                rid = self._next_id
                self._next_id += 1
                self.id_of_pk[pk] = rid
            ids[i] = rid
        return ids

    def upsert_many(self, pks: List[Tuple], rows: List[dict]) -> None:
        """One native call for a whole batch (the reference does one SQL round trip
        per record, JdbcWriter.java:139-179).  A primary key repeated inside the batch keeps its last row."""
        if len(pks) != len(rows):
            raise ValueError("pks and rows differ in length")
        if not pks:
            return
        last = {pk: i for i, pk in enumerate(pks)}
        keep = sorted(last.values())
        vecs = np.zeros((len(keep), self.store_dim), np.float32)
        for j, i in enumerate(keep):
            vec = rows[i].get(self.vector_column)
            if vec is None:
                raise ValueError(f"column {self.vector_column!r} is null")
            v = np.asarray(vec, np.float32).reshape(-1)
            if v.size != self.dim:
                raise ValueError(f"vector has {v.size} dimensions, table {self.name!r} has {self.dim}")
            vecs[j, : self.dim] = v
        with self.lock:
            ids = self._assign_ids([pks[i] for i in keep])
            for rid, i in zip(ids, keep):
                self.rows[int(rid)] = dict(rows[i])
            self.index.upsert(ids, vecs)

    def upsert_from_device(self, pks: List[Tuple], rows: List[dict], d_vectors, stream: int = 0) -> None:
        """Feed the index straight from device memory - e.g. the [n, dim] fp32 output of the embeddings
        service - instead of List<Double> -> List<Float> -> SQL (JstlFunctions.java:93-112).  `rows` hold the
        other columns; the vector column of the host rows is filled from one batched read-back so that
        `SELECT embeddings_vector` keeps working.  Primary keys must be distinct within the batch."""
        n = len(pks)
        if n != len(rows) or len(set(pks)) != n:
            raise ValueError("pks must be distinct and match rows")
        if self.store_dim != self.dim:
            raise ValueError("device-fed upserts need a dimension that is a multiple of 64")
        if tuple(d_vectors.shape) != (n, self.dim):
            raise ValueError(f"expected a [{n}, {self.dim}] device tensor")
        host = d_vectors.detach().to("cpu").numpy()  # one copy for the whole batch
        with self.lock:
            ids = self._assign_ids(list(pks))
            for j, rid in enumerate(ids):
                row = dict(rows[j])
                row[self.vector_column] = host[j]
                self.rows[int(rid)] = row
            self.index.upsert_dev(ids, d_vectors, n, stream)

    # -- persistence: the index file of lsc_index_save + the host rows next to it ---------------------
    def save(self, directory) -> None:
        """Write <directory>/<name>.lscidx (ids + bf16 rows, lsc_index_save) and <name>.rows.json; the
        directory is the agent's persistent state directory (AgentContext.java:64-66)."""
        from pathlib import Path
        d = Path(directory)
        d.mkdir(parents=True, exist_ok=True)
        with self.lock:
            self.index.save(str(d / f"{self.name}.lscidx"))
            meta = {"name": self.name, "dim": self.dim, "vector_column": self.vector_column, "next_id": self._next_id,
                    "pks": [[list(pk), rid] for pk, rid in self.id_of_pk.items()],
                    "rows": {str(rid): {k: (np.asarray(v).tolist() if isinstance(v, np.ndarray) else v)
                                        for k, v in row.items()} for rid, row in self.rows.items()}}
            tmp = d / f"{self.name}.rows.json.tmp"
            tmp.write_text(json.dumps(meta))
            tmp.replace(d / f"{self.name}.rows.json")

    @classmethod
    def load(cls, directory, name: str, device: int = 0, index_loader=None) -> "CudaVectorTable":
        from pathlib import Path
        d = Path(directory)
        meta = json.loads((d / f"{name}.rows.json").read_text())
        t = cls.__new__(cls)
        t.name, t.dim, t.vector_column = meta["name"], int(meta["dim"]), meta["vector_column"]
        t.store_dim = (t.dim + 63) // 64 * 64
        load = index_loader or (lambda path, dev, dim: _native.Index.load(path, dev, dim))
        t.index = load(str(d / f"{name}.lscidx"), device, t.store_dim)
        t.rows = {int(k): v for k, v in meta["rows"].items()}
        t.id_of_pk = {tuple(pk): int(rid) for pk, rid in meta["pks"]}
        t._next_id = int(meta["next_id"])
        t.lock = threading.Lock()
        if len(t.index) != len(t.rows):
            raise ValueError(f"table {name!r}: index file holds {len(t.index)} rows, row file {len(t.rows)}")
        return t

    @staticmethod
    def exists_on_disk(directory, name: str) -> bool:
        from pathlib import Path
        d = Path(directory)
        return (d / f"{name}.rows.json").exists() and (d / f"{name}.lscidx").exists()

    def _pad(self, v: np.ndarray) -> np.ndarray:
        if v.size == self.store_dim:
            return np.ascontiguousarray(v, np.float32)
        out = np.zeros(self.store_dim, np.float32)
        out[: v.size] = v
        return out

    def delete(self, pk: Tuple) -> None:
        with self.lock:
            rid = self.id_of_pk.pop(pk, None)
            if rid is not None:
                self.rows.pop(rid, None)
                self.index.delete(np.array([rid], np.int64))

    def search(self, vector, k: int) -> List[Tuple[dict, float]]:
        q = np.asarray(vector, np.float32).reshape(1, -1)
        if q.shape[1] != self.dim:
            raise ValueError(f"query vector has {q.shape[1]} dimensions, table {self.name!r} has {self.dim}")
        with self.lock:
            ids, scores = self.index.search(self._pad(q[0])[None, :], k)
            return [(self.rows[int(i)], float(s)) for i, s in zip(ids[0], scores[0]) if i >= 0]

    def search_many(self, vectors, k: int) -> List[List[Tuple[dict, float]]]:
        """One native call (one pass over the corpus per 128 / 256 queries) for a whole batch of query vectors."""
        q = np.asarray(vectors, np.float32).reshape(-1, self.dim)
        if q.shape[0] == 0:
            return []
        if self.store_dim != self.dim:
            padded = np.zeros((q.shape[0], self.store_dim), np.float32)
            padded[:, : self.dim] = q
            q = padded
        with self.lock:
            ids, scores = self.index.search(np.ascontiguousarray(q), k)
            return [[(self.rows[int(i)], float(s)) for i, s in zip(ids[b], scores[b]) if i >= 0] for b in range(q.shape[0])]

    def close(self) -> None:
        self.index.close()


class QueryBatcher:
    """Cross-caller micro-batcher for vector queries.

    The reference answers one query per record, synchronously (SingleRecordAgentProcessor.java:31-47; fused
    pipelines feed List.of(record), CompositeAgentProcessor.java:225-227), which on a GPU means one pass over the
    whole corpus per record.  A pass serves up to 128 queries for the same HBM traffic, so queries from all
    callers (agent threads, gateway sessions) are collected here and sent as one batch when `max_batch` are
    waiting or the oldest is `flush_interval_ms` old - the same window logic as CudaEmbeddingsService."""

    def __init__(self, max_batch: int = 64, flush_interval_ms: float = 1.0):
        self.max_batch, self._flush_s = int(max_batch), float(flush_interval_ms) / 1000.0
        self._cv = threading.Condition()
        self._pending: List[tuple] = []  # (table, k, vector, future, enqueue time)
        self._closed = False
        self.batches = 0
        self.queries = 0
        self._worker = threading.Thread(target=self._run, name="lsc-query-batcher", daemon=True)
        self._worker.start()

    def submit(self, table: CudaVectorTable, vector, k: int) -> "Future[List[Tuple[dict, float]]]":
        fut: Future = Future()
        v = np.asarray(vector, np.float32).reshape(-1)
        if v.size != table.dim:
            fut.set_exception(ValueError(f"query vector has {v.size} dimensions, table {table.name!r} has {table.dim}"))
            return fut
        with self._cv:
            if self._closed:
                fut.set_exception(RuntimeError("datasource is closed"))
                return fut
            self._pending.append((table, int(k), v, fut, time.monotonic()))
            self._cv.notify()
        return fut

    def close(self) -> None:
        with self._cv:
            self._closed = True
            self._cv.notify()
        self._worker.join(timeout=10)

    def _run(self) -> None:
        while True:
            with self._cv:
                while not self._pending and not self._closed:
                    self._cv.wait()
                if not self._pending and self._closed:
                    return
                while not self._closed:
                    age = time.monotonic() - self._pending[0][4]
                    if len(self._pending) >= self.max_batch or age >= self._flush_s:
                        break
                    self._cv.wait(timeout=max(self._flush_s - age, 0.0002))
                # one native call per (table, k): take the oldest request's group, leave the rest for the next turn
                table, k = self._pending[0][0], self._pending[0][1]
                batch = [p for p in self._pending if p[0] is table and p[1] == k][: self.max_batch]
                taken = set(id(p) for p in batch)
                self._pending = [p for p in self._pending if id(p) not in taken]
            try:
                hits = table.search_many(np.stack([p[2] for p in batch]), k)
                self.batches += 1
                self.queries += len(batch)
                for p, h in zip(batch, hits):
                    p[3].set_result(h)
            except Exception as e:  # noqa: BLE001 - every caller of the batch sees the failure, as its own call would
                for p in batch:
                    if not p[3].done():
                        p[3].set_exception(e)


_TABLES: Dict[str, CudaVectorTable] = {}
_TABLES_LOCK = threading.Lock()


def get_table(name: str, dim: Optional[int] = None, vector_column: Optional[str] = None, device: int = 0,
              index_factory=None) -> CudaVectorTable:
    """Tables are process-wide, like a database shared by the sink and the query agent of one pod."""
    with _TABLES_LOCK:
        t = _TABLES.get(name)
        if t is None:
            if dim is None or vector_column is None:
                raise ValueError(f"table {name!r} does not exist")
            t = CudaVectorTable(name, dim, vector_column, device, index_factory)
            _TABLES[name] = t
        return t


def table_exists(name: str) -> bool:
    with _TABLES_LOCK:
        return name in _TABLES


def register_table(table: CudaVectorTable) -> None:
    with _TABLES_LOCK:
        old = _TABLES.get(table.name)
        _TABLES[table.name] = table
    if old is not None and old is not table:
        old.close()


def drop_table(name: str) -> None:
    with _TABLES_LOCK:
        t = _TABLES.pop(name, None)
    if t is not None:
        t.close()


# ---------------------------------------------------------------------------------------------
# query-vector-db side
# ---------------------------------------------------------------------------------------------
_SQL = re.compile(
    r"^\s*SELECT\s+(?P<cols>.+?)\s+FROM\s+(?P<table>[\w.]+)\s+ORDER\s+BY\s+cosine_similarity\s*\(\s*(?P<vec>\w+)\s*,"
    r"\s*CAST\s*\(\s*\?\s+AS\s+FLOAT\s+ARRAY\s*\)\s*\)\s+DESC\s+LIMIT\s+(?P<k>\d+)\s*;?\s*$",
    re.IGNORECASE | re.DOTALL)


def interpolate(query: str, params: List[Any]) -> str:
    """InterpolationUtils.interpolate (langstream-agents/langstream-vector-agents/…/InterpolationUtils.java:30-52):
    each '?' is replaced, in order, by the JSON form of the next parameter."""
    if query is None or "?" not in query:
        return query
    for value in params:
        q = query.find("?")
        if q < 0:
            return query
        query = query[:q] + json.dumps(value) + query[q + 1:]
    return query


class CudaVectorDataSource:
    """QueryStepDataSource for `service: "cuda"`.

    fetch_data accepts
      * the SQL shape of the reference's examples (examples/applications/docker-chatbot/chatbot.yaml:44,
        RerankAgentRunnerIT.java:204):
        SELECT cols FROM table ORDER BY cosine_similarity(col, CAST(? AS FLOAT ARRAY)) DESC LIMIT k
      * a JSON query in the style of the other vector datasources (PineconeDataSource.java:291-313):
        {"table": "documents", "vector": ?, "topK": 10, "select": ["text"], "include-similarity": true}
    and returns one map per row with the selected columns (plus `similarity` when asked for, the
    column name AstraVectorDBDataSource.java:142-143 uses).  Errors raise unchecked exceptions, which
    SingleRecordAgentProcessor turns into a per-record error (SingleRecordAgentProcessor.java:41-45).
    """

    def __init__(self, index_factory=None):
        self._config: dict = {}
        self._index_factory = index_factory
        self._batcher: Optional[QueryBatcher] = None

    def initialize(self, config: dict) -> None:
        """Datasource configuration; besides `service: cuda` / `device`:
        `query-batch-size` (default 1 = every fetchData is its own search) and `query-flush-interval-ms`
        (default 1.0) switch on the cross-caller micro-batcher: concurrent fetch_data calls are answered by one
        search of up to that many queries."""
        self._config = dict(config or {})
        size = int(self._config.get("query-batch-size", 1) or 1)
        if size > 1:
            self._batcher = QueryBatcher(size, float(self._config.get("query-flush-interval-ms", 1.0)))

    def _resolve(self, query: str, params: List[Any]):
        """(table, query vector, k, projection of one hit) of a query string + positional parameters."""
        m = _SQL.match(query or "")
        if m:
            if len(params) < 1:
                raise ValueError("the query has one '?' parameter but none was given")
            cols = [c.strip() for c in m.group("cols").split(",")]
            table = get_table(m.group("table"))
            if m.group("vec") != table.vector_column:
                raise ValueError(f"{m.group('vec')!r} is not the vector column of table {table.name!r}")
            return table, params[0], int(m.group("k")), (lambda row, score: self._project(row, score, cols, False))
        try:
            spec = json.loads(interpolate(query, params))
        except Exception as e:  # noqa: BLE001
            raise ValueError(f"cannot parse query: {e}") from e  # IllegalArgumentException in buildObjectFromJson
        if not isinstance(spec, dict) or "vector" not in spec:
            raise ValueError("query must be the cosine SQL form or a JSON object with a 'vector'")
        table = get_table(spec.get("table") or self._config.get("table-name") or "documents")
        cols = spec.get("select") or ["*"]
        sim = bool(spec.get("include-similarity", True))
        return table, spec["vector"], int(spec.get("topK", 10)), (lambda row, score: self._project(row, score, cols, sim))

    def fetch_data(self, query: str, params: List[Any]) -> List[Dict[str, Any]]:
        table, vector, k, project = self._resolve(query, params)
        if self._batcher is not None:  # joins whatever other callers are asking at this moment
            hits = self._batcher.submit(table, vector, k).result()
        else:
            hits = table.search(vector, k)
        return [project(row, score) for row, score in hits]

    def fetch_data_async(self, query: str, params: List[Any]) -> "Future[List[Dict[str, Any]]]":
        """fetch_data that returns at once (needs the micro-batcher): the shape an asynchronous agent wants - it
        keeps many records in flight and lets them meet in one search."""
        out: Future = Future()
        try:
            table, vector, k, project = self._resolve(query, params)
            if self._batcher is None:
                out.set_result([project(row, score) for row, score in table.search(vector, k)])
                return out
            inner = self._batcher.submit(table, vector, k)
        except Exception as e:  # noqa: BLE001
            out.set_exception(e)
            return out

        def done(f):
            try:
                out.set_result([project(row, score) for row, score in f.result()])
            except Exception as e:  # noqa: BLE001
                out.set_exception(e)

        inner.add_done_callback(done)
        return out

    def fetch_data_many(self, query: str, params_list: List[List[Any]]) -> List[Any]:
        """One search for a whole list of parameter sets of the same query (a batch of records handed to
        AgentProcessor.process at once).  Element i is the row list of params_list[i], or the exception that
        params_list[i] alone would have raised."""
        out: List[Any] = [None] * len(params_list)
        groups: Dict[tuple, list] = {}
        for i, params in enumerate(params_list):
            try:
                table, vector, k, project = self._resolve(query, params)
                v = np.asarray(vector, np.float32).reshape(-1)
                if v.size != table.dim:
                    raise ValueError(f"query vector has {v.size} dimensions, table {table.name!r} has {table.dim}")
                groups.setdefault((id(table), k), []).append((i, table, v, project))
            except Exception as e:  # noqa: BLE001
                out[i] = e
        for (_, k), items in groups.items():
            try:
                hits = items[0][1].search_many(np.stack([it[2] for it in items]), k)
                for (i, _, _, project), h in zip(items, hits):
                    out[i] = [project(row, score) for row, score in h]
            except Exception as e:  # noqa: BLE001
                for i, _, _, _ in items:
                    out[i] = e
        return out

    @staticmethod
    def _project(row: dict, score: float, cols: List[str], with_similarity: bool) -> Dict[str, Any]:
        if cols == ["*"]:
            out = dict(row)
        else:
            missing = [c for c in cols if c not in row]
            if missing:
                raise ValueError(f"unknown column(s) {missing}")
            out = {c: row[c] for c in cols}
        if with_similarity:
            out["similarity"] = score
        return out

    def execute_statement(self, query: str, generated_keys: List[str], params: List[Any]) -> Dict[str, Any]:
        self.fetch_data(query, params)  # the interface's default behaviour (QueryStepDataSource.java:30-34)
        return {}

    def close(self) -> None:
        if self._batcher is not None:
            self._batcher.close()
            self._batcher = None


class CudaVectorDataSourceProvider:
    """DataSourceProvider: keyed on `service == "cuda"` like the other providers
    (e.g. PineconeDataSource.java:58-61)."""

    def __init__(self, index_factory=None):
        self._index_factory = index_factory

    def supports(self, data_source_config: dict) -> bool:
        return (data_source_config or {}).get("service") == "cuda"

    def create_data_source_implementation(self, data_source_config: dict) -> CudaVectorDataSource:
        return CudaVectorDataSource(self._index_factory)


# ---------------------------------------------------------------------------------------------
# vector-db-sink side
# ---------------------------------------------------------------------------------------------
class CudaVectorDatabaseWriter:
    """VectorDatabaseWriter for `service: "cuda"`; agent configuration as JdbcWriter's
    (`table-name`, `fields: [{name, expression, primary-key}]`, JdbcWriter.java:74-95), plus
    `vector-column` (default: the first field whose value is a list) and `dimensions`."""

    def __init__(self, datasource_config: dict, index_factory=None, index_loader=None):
        self.datasource_config = dict(datasource_config or {})
        self._index_factory = index_factory
        self._index_loader = index_loader
        self.table_name: Optional[str] = None
        self.primary_key: Dict[str, str] = {}
        self.columns: Dict[str, str] = {}
        self.vector_column: Optional[str] = None
        self.dim: Optional[int] = None
        self._table: Optional[CudaVectorTable] = None
        self._lock = threading.Lock()
        self.state_dir = None

    def upsert_batch(self, records: List[Record]) -> "Future[None]":
        """All records of a batch in one native call; null values (deletes) are applied in order around it.
        The future completes once every row is queryable, like `upsert`."""
        handle: Future = Future()
        try:
            with self._lock:
                pks, rows = [], []

                def flush():
                    if pks:
                        self._ensure_table(rows[0]).upsert_many(list(pks), list(rows))
                        pks.clear()
                        rows.clear()

                for record in records:
                    pk = tuple(evaluate(e, record) for e in self.primary_key.values())
                    if record.value is not None:
                        row = {n: evaluate(e, record) for n, e in self.primary_key.items()}
                        row.update({n: evaluate(e, record) for n, e in self.columns.items()})
                        pks.append(pk)
                        rows.append(row)
                    else:
                        flush()
                        if self._table is None and table_exists(self.table_name):
                            self._table = get_table(self.table_name)
                        if self._table is not None:
                            self._table.delete(pk)
                flush()
            handle.set_result(None)
        except Exception as e:  # noqa: BLE001
            handle.set_exception(e)
        return handle

    def initialise(self, agent_configuration: dict) -> None:
        cfg = dict(agent_configuration or {})
        self.table_name = cfg.get("table-name")
        if not self.table_name:
            raise ValueError("table-name is required")
        for f in cfg.get("fields", []):
            target = self.primary_key if bool(f.get("primary-key", False)) else self.columns
            target[str(f["name"])] = str(f["expression"])
        if not self.primary_key:
            raise ValueError("at least one field must be marked primary-key")
        self.vector_column = cfg.get("vector-column")
        self.dim = cfg.get("dimensions")
        # restore the table from the agent's persistent state directory, if any
        self.state_dir = cfg.get("persistent-state-directory")
        if self.state_dir and not table_exists(self.table_name) and \
                CudaVectorTable.exists_on_disk(self.state_dir, self.table_name):
            register_table(CudaVectorTable.load(self.state_dir, self.table_name,
                                                int(self.datasource_config.get("device", 0)), self._index_loader))

    def upsert(self, record: Record, context: Dict[str, Any]) -> "Future[None]":
        handle: Future = Future()
        try:
            with self._lock:  # the reference's upsert is synchronized (JdbcWriter.java:139)
                pk = tuple(evaluate(e, record) for e in self.primary_key.values())
                if record.value is not None:
                    row = {n: evaluate(e, record) for n, e in self.primary_key.items()}
                    row.update({n: evaluate(e, record) for n, e in self.columns.items()})
                    table = self._ensure_table(row)
                    table.upsert(pk, row)
                else:  # null value => delete (VectorDatabaseWriter.java:30-33)
                    if self._table is None:
                        try:
                            self._table = get_table(self.table_name)
                        except ValueError:
                            self._table = None
                    if self._table is not None:
                        self._table.delete(pk)
            handle.set_result(None)  # completes only once the row is queryable
        except Exception as e:  # noqa: BLE001 - completeExceptionally (JdbcWriter.java:175-177)
            handle.set_exception(e)
        return handle

    def _ensure_table(self, row: dict) -> CudaVectorTable:
        if self._table is None:
            col = self.vector_column
            if col is None:
                col = next((n for n, v in row.items() if isinstance(v, (list, tuple, np.ndarray))), None)
                if col is None:
                    raise ValueError("no vector column found among the fields")
            dim = int(self.dim) if self.dim else len(row[col])
            self._table = get_table(self.table_name, dim, col, int(self.datasource_config.get("device", 0)),
                                    self._index_factory)
        return self._table

    def close(self) -> None:
        # durable before the agent goes away: the source offsets were committed when the futures completed
        if self.state_dir and self.table_name and table_exists(self.table_name):
            get_table(self.table_name).save(self.state_dir)


class CudaTableAssetManager:
    """AssetManager for `asset-type: "cuda-table"`, the counterpart of `jdbc-table`
    (JdbcAssetsManagerProvider.java:33-144): `assets:` in an application can create / drop a GPU table.
    Asset config: `table-name`, `dimensions`, `vector-column`, optional `datasource.configuration.device`
    and `persistent-state-directory` (an existing saved table counts as existing and is loaded on deploy)."""

    def __init__(self, index_factory=None, index_loader=None):
        self._index_factory, self._index_loader = index_factory, index_loader
        self.config: Dict[str, Any] = {}

    def initialize(self, asset_definition: Dict[str, Any]) -> None:
        self.config = dict(asset_definition.get("config", asset_definition) or {})
        if not self.config.get("table-name"):
            raise ValueError("table-name is required")

    def _device(self) -> int:
        ds = (self.config.get("datasource") or {}).get("configuration") or {}
        return int(ds.get("device", 0))

    def asset_exists(self) -> bool:
        name, d = self.config["table-name"], self.config.get("persistent-state-directory")
        return table_exists(name) or bool(d and CudaVectorTable.exists_on_disk(d, name))

    def deploy_asset(self) -> None:
        name, d = self.config["table-name"], self.config.get("persistent-state-directory")
        if table_exists(name):
            return
        if d and CudaVectorTable.exists_on_disk(d, name):
            register_table(CudaVectorTable.load(d, name, self._device(), self._index_loader))
            return
        dim, col = self.config.get("dimensions"), self.config.get("vector-column")
        if not dim or not col:
            raise ValueError("dimensions and vector-column are required to create a cuda-table")
        get_table(name, int(dim), str(col), self._device(), self._index_factory)

    def delete_asset_if_exists(self) -> bool:
        if not self.asset_exists():
            return False
        name, d = self.config["table-name"], self.config.get("persistent-state-directory")
        drop_table(name)
        if d:
            from pathlib import Path
            for suffix in (".lscidx", ".rows.json"):
                p = Path(d) / f"{name}{suffix}"
                if p.exists():
                    p.unlink()
        return True

    def close(self) -> None:
        pass


class CudaAssetManagerProvider:
    def __init__(self, index_factory=None, index_loader=None):
        self._index_factory, self._index_loader = index_factory, index_loader

    def supports(self, asset_type: str) -> bool:
        return asset_type == "cuda-table"

    def create_instance(self, asset_type: str) -> CudaTableAssetManager:
        if asset_type != "cuda-table":
            raise ValueError(asset_type)  # IllegalArgumentException in the reference
        return CudaTableAssetManager(self._index_factory, self._index_loader)


class CudaVectorDatabaseWriterProvider:
    def __init__(self, index_factory=None, index_loader=None):
        self._index_factory, self._index_loader = index_factory, index_loader

    def supports(self, data_source_config: dict) -> bool:
        return (data_source_config or {}).get("service") == "cuda"

    def create_implementation(self, datasource_config: dict) -> CudaVectorDatabaseWriter:
        return CudaVectorDatabaseWriter(datasource_config, self._index_factory, self._index_loader)
