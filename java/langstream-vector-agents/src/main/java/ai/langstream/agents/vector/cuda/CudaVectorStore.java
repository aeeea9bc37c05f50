package ai.langstream.agents.vector.cuda;

import ai.langstream.ai.agents.commons.MutableRecord;
import ai.langstream.ai.agents.commons.jstl.JstlEvaluator;
import ai.langstream.ai.agents.datasource.DataSourceProvider;
import ai.langstream.api.database.VectorDatabaseWriter;
import ai.langstream.api.database.VectorDatabaseWriterProvider;
import ai.langstream.api.runner.code.Record;
import ai.langstream.cuda.LangstreamCuda;
import com.datastax.oss.streaming.ai.datasource.QueryStepDataSource;
import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.nio.FloatBuffer;
import java.nio.LongBuffer;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.LinkedHashMap;
import java.util.List;
import java.util.Map;
import java.util.concurrent.CompletableFuture;
import java.util.concurrent.ConcurrentHashMap;
import java.util.regex.Matcher;
import java.util.regex.Pattern;

/**
 * Local GPU vector store for `service: "cuda"`: the datasource (query-vector-db) and the writer
 * (vector-db-sink) of one pod share a table registry. The GPU keeps the L2-normalised bf16 scan copy of
 * the vector column (lsc_index_*); all columns, including the original vector, stay in a host row map.
 * Register the two providers in META-INF/services (files next to this source tree).
 */
public final class CudaVectorStore {

    static final class Table {
        final long index;
        final int dim;          // user dimension
        final int storeDim;     // padded to a multiple of 64 for the scan kernel (zero padding keeps cosine)
        final String vectorColumn;
        final Map<List<Object>, Long> idOfPk = new HashMap<>();
        final Map<Long, Map<String, Object>> rows = new HashMap<>();
        long nextId;

        Table(int dim, String vectorColumn, int device) {
            this.dim = dim;
            this.storeDim = (dim + 63) / 64 * 64;
            this.vectorColumn = vectorColumn;
            this.index = LangstreamCuda.indexCreate(storeDim, 1024, device);
        }
    }

    static final Map<String, Table> TABLES = new ConcurrentHashMap<>();
    static final Pattern COSINE_SQL = Pattern.compile(
            "^\\s*SELECT\\s+(.+?)\\s+FROM\\s+([\\w.]+)\\s+ORDER\\s+BY\\s+cosine_similarity\\s*\\(\\s*(\\w+)\\s*,"
                    + "\\s*CAST\\s*\\(\\s*\\?\\s+AS\\s+FLOAT\\s+ARRAY\\s*\\)\\s*\\)\\s+DESC\\s+LIMIT\\s+(\\d+)\\s*;?\\s*$",
            Pattern.CASE_INSENSITIVE | Pattern.DOTALL);

    private CudaVectorStore() {}

    static FloatBuffer padded(List<? extends Number> v, int storeDim) {
        FloatBuffer b = ByteBuffer.allocateDirect(storeDim * 4).order(ByteOrder.nativeOrder()).asFloatBuffer();
        for (int i = 0; i < v.size(); i++) {
            b.put(i, v.get(i).floatValue());
        }
        return b;
    }

    // ------------------------------------------------------------------ query-vector-db
    public static class Provider implements DataSourceProvider {
        @Override
        public boolean supports(Map<String, Object> dataSourceConfig) {
            return "cuda".equals(dataSourceConfig.get("service"));
        }

        @Override
        public QueryStepDataSource createDataSourceImplementation(Map<String, Object> dataSourceConfig) {
            return new DataSource();
        }
    }

    static class DataSource implements QueryStepDataSource {
        @Override
        @SuppressWarnings("unchecked")
        public List<Map<String, Object>> fetchData(String query, List<Object> params) {
            // the SQL shape of the examples (chatbot.yaml:44): only the datasource changes in the app
            Matcher m = COSINE_SQL.matcher(query);
            if (!m.matches()) {
                throw new IllegalArgumentException("unsupported query for the cuda datasource: " + query);
            }
            String[] cols = m.group(1).split("\\s*,\\s*");
            Table t = TABLES.get(m.group(2));
            int k = Integer.parseInt(m.group(4));
            if (t == null) {
                throw new IllegalArgumentException("table " + m.group(2) + " does not exist");
            }
            List<? extends Number> q = (List<? extends Number>) params.get(0);
            if (q.size() != t.dim) {
                throw new IllegalArgumentException("query vector has " + q.size() + " dimensions, table has " + t.dim);
            }
            LongBuffer ids = ByteBuffer.allocateDirect(k * 8).order(ByteOrder.nativeOrder()).asLongBuffer();
            FloatBuffer scores = ByteBuffer.allocateDirect(k * 4).order(ByteOrder.nativeOrder()).asFloatBuffer();
            List<Map<String, Object>> out = new ArrayList<>();
            synchronized (t) {
                LangstreamCuda.indexSearch(t.index, padded(q, t.storeDim), 1, k, ids, scores);
                for (int i = 0; i < k && ids.get(i) >= 0; i++) {
                    Map<String, Object> row = t.rows.get(ids.get(i));
                    Map<String, Object> projected = new LinkedHashMap<>();
                    for (String c : cols) {
                        if (!row.containsKey(c)) {
                            throw new IllegalArgumentException("unknown column " + c);
                        }
                        projected.put(c, row.get(c));
                    }
                    out.add(projected);
                }
            }
            return out;
        }
    }

    // ------------------------------------------------------------------ vector-db-sink
    public static class WriterProvider implements VectorDatabaseWriterProvider {
        @Override
        public boolean supports(Map<String, Object> dataSourceConfig) {
            return "cuda".equals(dataSourceConfig.get("service"));
        }

        @Override
        public VectorDatabaseWriter createImplementation(Map<String, Object> datasourceConfig) {
            return new Writer(datasourceConfig);
        }
    }

    static class Writer implements VectorDatabaseWriter {
        private final Map<String, Object> datasourceConfig;
        private String tableName;
        private final LinkedHashMap<String, JstlEvaluator<Object>> primaryKey = new LinkedHashMap<>();
        private final LinkedHashMap<String, JstlEvaluator<Object>> columns = new LinkedHashMap<>();

        Writer(Map<String, Object> datasourceConfig) {
            this.datasourceConfig = datasourceConfig;
        }

        @Override
        @SuppressWarnings("unchecked")
        public void initialise(Map<String, Object> agentConfiguration) {
            // same agent configuration as JdbcWriter (JdbcWriter.java:74-95)
            tableName = (String) agentConfiguration.get("table-name");
            List<Map<String, Object>> fields =
                    (List<Map<String, Object>>) agentConfiguration.getOrDefault("fields", List.of());
            for (Map<String, Object> f : fields) {
                boolean pk = Boolean.parseBoolean(f.getOrDefault("primary-key", "false").toString());
                JstlEvaluator<Object> ev = new JstlEvaluator<>("${" + f.get("expression") + "}", Object.class);
                (pk ? primaryKey : columns).put(f.get("name").toString(), ev);
            }
        }

        @Override
        @SuppressWarnings("unchecked")
        public synchronized CompletableFuture<?> upsert(Record record, Map<String, Object> context) {
            CompletableFuture<Void> handle = new CompletableFuture<>();
            try {
                MutableRecord ctx = MutableRecord.recordToMutableRecord(record, true);
                List<Object> pk = new ArrayList<>();
                Map<String, Object> row = new LinkedHashMap<>();
                primaryKey.forEach((n, ev) -> { Object v = ev.evaluate(ctx); pk.add(v); row.put(n, v); });
                if (record.value() != null) {
                    columns.forEach((n, ev) -> row.put(n, ev.evaluate(ctx)));
                    String vecCol = row.entrySet().stream().filter(e -> e.getValue() instanceof List)
                            .map(Map.Entry::getKey).findFirst()
                            .orElseThrow(() -> new IllegalArgumentException("no vector column among the fields"));
                    List<? extends Number> vec = (List<? extends Number>) row.get(vecCol);
                    int device = Integer.parseInt(datasourceConfig.getOrDefault("device", "0").toString());
                    Table t = TABLES.computeIfAbsent(tableName, n -> new Table(vec.size(), vecCol, device));
                    synchronized (t) {
                        long id = t.idOfPk.computeIfAbsent(pk, p -> t.nextId++);
                        t.rows.put(id, row);
                        LongBuffer ids = ByteBuffer.allocateDirect(8).order(ByteOrder.nativeOrder()).asLongBuffer();
                        ids.put(0, id);
                        LangstreamCuda.indexUpsert(t.index, ids, padded(vec, t.storeDim), 1);
                    }
                } else {                                   // null value => delete (VectorDatabaseWriter.java:30-33)
                    Table t = TABLES.get(tableName);
                    if (t != null) {
                        synchronized (t) {
                            Long id = t.idOfPk.remove(pk);
                            if (id != null) {
                                t.rows.remove(id);
                                LongBuffer ids = ByteBuffer.allocateDirect(8).order(ByteOrder.nativeOrder()).asLongBuffer();
                                ids.put(0, id);
                                LangstreamCuda.indexDelete(t.index, ids, 1);
                            }
                        }
                    }
                }
                handle.complete(null);                     // the row is queryable: the source offset may be committed
            } catch (Exception e) {
                handle.completeExceptionally(e);
            }
            return handle;
        }
    }
}
