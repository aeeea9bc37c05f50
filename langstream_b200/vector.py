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

    def close(self) -> None:
        self.index.close()


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

    def initialize(self, config: dict) -> None:
        self._config = dict(config or {})

    def fetch_data(self, query: str, params: List[Any]) -> List[Dict[str, Any]]:
        m = _SQL.match(query or "")
        if m:
            if len(params) < 1:
                raise ValueError("the query has one '?' parameter but none was given")
            cols = [c.strip() for c in m.group("cols").split(",")]
            table = get_table(m.group("table"))
            if m.group("vec") != table.vector_column:
                raise ValueError(f"{m.group('vec')!r} is not the vector column of table {table.name!r}")
            hits = table.search(params[0], int(m.group("k")))
            return [self._project(row, score, cols, False) for row, score in hits]
        try:
            spec = json.loads(interpolate(query, params))
        except Exception as e:  # noqa: BLE001
            raise ValueError(f"cannot parse query: {e}") from e  # IllegalArgumentException in buildObjectFromJson
        if not isinstance(spec, dict) or "vector" not in spec:
            raise ValueError("query must be the cosine SQL form or a JSON object with a 'vector'")
        table = get_table(spec.get("table") or self._config.get("table-name") or "documents")
        hits = table.search(spec["vector"], int(spec.get("topK", 10)))
        cols = spec.get("select") or ["*"]
        return [self._project(row, score, cols, bool(spec.get("include-similarity", True))) for row, score in hits]

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
        pass


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

    def __init__(self, datasource_config: dict, index_factory=None):
        self.datasource_config = dict(datasource_config or {})
        self._index_factory = index_factory
        self.table_name: Optional[str] = None
        self.primary_key: Dict[str, str] = {}
        self.columns: Dict[str, str] = {}
        self.vector_column: Optional[str] = None
        self.dim: Optional[int] = None
        self._table: Optional[CudaVectorTable] = None
        self._lock = threading.Lock()

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
        pass


class CudaVectorDatabaseWriterProvider:
    def __init__(self, index_factory=None):
        self._index_factory = index_factory

    def supports(self, data_source_config: dict) -> bool:
        return (data_source_config or {}).get("service") == "cuda"

    def create_implementation(self, datasource_config: dict) -> CudaVectorDatabaseWriter:
        return CudaVectorDatabaseWriter(datasource_config, self._index_factory)
