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
        final long index;       // lsc_index (one GPU) ...
        final long sharded;     // ... or lsc_sharded_index (rows spread over `devices`), exactly one of the two is set
        final int dim;          // user dimension
        final int storeDim;     // padded to a multiple of 64 for the scan kernel (zero padding keeps cosine)
        final String vectorColumn;
        final Map<List<Object>, Long> idOfPk = new HashMap<>();
        final Map<Long, Map<String, Object>> rows = new HashMap<>();
        long nextId;

        Table(int dim, String vectorColumn, int[] devices) {
            this.dim = dim;
            this.storeDim = (dim + 63) / 64 * 64;
            this.vectorColumn = vectorColumn;
            if (devices.length > 1) {   // one host thread drives all GPUs through the sharded handle
                this.sharded = LangstreamCuda.shardedIndexCreate(storeDim, 1024, devices);
                this.index = 0;
            } else {
                this.index = LangstreamCuda.indexCreate(storeDim, 1024, devices[0]);
                this.sharded = 0;
                LangstreamCuda.indexReserve(index, 128, 32);   // no allocation on the query path later
            }
        }

        private Table(long index, int dim, String vectorColumn) {
            this.index = index;
            this.sharded = 0;
            this.dim = dim;
            this.storeDim = (dim + 63) / 64 * 64;
            this.vectorColumn = vectorColumn;
        }

        void upsert(LongBuffer ids, FloatBuffer vecs, int n) {
            if (sharded != 0) LangstreamCuda.shardedIndexUpsert(sharded, ids, vecs, n);
            else LangstreamCuda.indexUpsert(index, ids, vecs, n);
        }

        void delete(LongBuffer ids, int n) {
            if (sharded != 0) LangstreamCuda.shardedIndexDelete(sharded, ids, n);
            else LangstreamCuda.indexDelete(index, ids, n);
        }

        void search(FloatBuffer queries, int b, int k, LongBuffer outIds, FloatBuffer outScores) {
            if (sharded != 0) LangstreamCuda.shardedIndexSearch(sharded, queries, b, k, outIds, outScores);
            else LangstreamCuda.indexSearch(index, queries, b, k, outIds, outScores);
        }

        /**
         * Durable copy in the agent's persistent state directory (AgentContext.java:64-66): name.lscidx (ids + bf16
         * rows, lsc_index_save) and name.rows.json (every column, the primary keys, the id counter). Single-GPU
         * tables only: a sharded table is rebuilt from its source topic.
         */
        synchronized void save(java.nio.file.Path dir, String name) throws java.io.IOException {
            if (sharded != 0) throw new UnsupportedOperationException("persistence of sharded tables is not implemented");
            java.nio.file.Files.createDirectories(dir);
            LangstreamCuda.indexSave(index, dir.resolve(name + ".lscidx").toString());
            Map<String, Object> meta = new LinkedHashMap<>();
            meta.put("dim", dim);
            meta.put("vector_column", vectorColumn);
            meta.put("next_id", nextId);
            List<Object> pks = new ArrayList<>();
            idOfPk.forEach((pk, id) -> pks.add(List.of(pk, id)));
            meta.put("pks", pks);
            Map<String, Object> r = new LinkedHashMap<>();
            rows.forEach((id, row) -> r.put(Long.toString(id), row));
            meta.put("rows", r);
            java.nio.file.Path tmp = dir.resolve(name + ".rows.json.tmp");
            new com.fasterxml.jackson.databind.ObjectMapper().writeValue(tmp.toFile(), meta);
            java.nio.file.Files.move(tmp, dir.resolve(name + ".rows.json"), java.nio.file.StandardCopyOption.REPLACE_EXISTING);
        }

        @SuppressWarnings("unchecked")
        static Table load(java.nio.file.Path dir, String name, int device) throws java.io.IOException {
            Map<String, Object> meta = new com.fasterxml.jackson.databind.ObjectMapper()
                    .readValue(dir.resolve(name + ".rows.json").toFile(), Map.class);
            Table t = new Table(LangstreamCuda.indexLoad(dir.resolve(name + ".lscidx").toString(), device),
                    ((Number) meta.get("dim")).intValue(), (String) meta.get("vector_column"));
            t.nextId = ((Number) meta.get("next_id")).longValue();
            for (List<Object> e : (List<List<Object>>) meta.get("pks")) {
                t.idOfPk.put((List<Object>) e.get(0), ((Number) e.get(1)).longValue());
            }
            ((Map<String, Map<String, Object>>) meta.get("rows")).forEach((id, row) -> t.rows.put(Long.parseLong(id), row));
            if (LangstreamCuda.indexSize(t.index) != t.rows.size()) {
                throw new java.io.IOException("table " + name + ": index file and row file disagree");
            }
            return t;
        }

        static boolean existsOnDisk(java.nio.file.Path dir, String name) {
            return java.nio.file.Files.exists(dir.resolve(name + ".rows.json")) && java.nio.file.Files.exists(dir.resolve(name + ".lscidx"));
        }

        void close() {
            if (sharded != 0) LangstreamCuda.shardedIndexDestroy(sharded);
            else LangstreamCuda.indexDestroy(index);
        }
    }

    /** `device: 0` or `devices: [0, 1, ...]` of the datasource configuration. */
    @SuppressWarnings("unchecked")
    static int[] devicesOf(Map<String, Object> datasourceConfig) {
        Object list = datasourceConfig.get("devices");
        if (list instanceof List<?> l && !l.isEmpty()) {
            return l.stream().mapToInt(d -> Integer.parseInt(d.toString())).toArray();
        }
        return new int[] {Integer.parseInt(datasourceConfig.getOrDefault("device", "0").toString())};
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

    /**
     * query-vector-db side. Two query forms: the cosine SQL of the reference's examples (chatbot.yaml:44), and a JSON
     * query in the style of the other vector datasources, {"table": "documents", "vector": ?, "topK": 10,
     * "select": ["text"], "include-similarity": true} - the latter adds a `similarity` column (the name
     * AstraVectorDBDataSource.java:142-143 uses).
     *
     * <p>`query-batch-size` > 1 in the datasource configuration turns on the cross-caller micro-batcher: fetchData
     * is called once per record from the runner thread of every agent replica / gateway session
     * (SingleRecordAgentProcessor.java:31-47), and one pass over the corpus serves up to 128 queries, so concurrent
     * calls wait up to `query-flush-interval-ms` for each other and share one search.
     */
    static class DataSource implements QueryStepDataSource {
        private record Pending(Table table, int k, FloatBuffer query, CompletableFuture<List<Map.Entry<Long, Float>>> hits, long enqueuedNanos) {}

        private final java.util.ArrayDeque<Pending> pending = new java.util.ArrayDeque<>();
        private Thread batcher;
        private int maxBatch = 1;
        private long flushNanos = 1_000_000;
        private volatile boolean closed;

        @Override
        public void initialize(Map<String, Object> config) {
            maxBatch = Integer.parseInt(config.getOrDefault("query-batch-size", "1").toString());
            flushNanos = (long) (1e6 * Double.parseDouble(config.getOrDefault("query-flush-interval-ms", "1").toString()));
            if (maxBatch > 1) {
                batcher = new Thread(this::run, "lsc-query-batcher");
                batcher.setDaemon(true);
                batcher.start();
            }
        }

        @Override
        @SuppressWarnings("unchecked")
        public List<Map<String, Object>> fetchData(String query, List<Object> params) {
            Matcher m = COSINE_SQL.matcher(query);
            String[] cols;
            Table t;
            int k;
            boolean similarity = false;
            List<? extends Number> q;
            if (m.matches()) {
                cols = m.group(1).split("\\s*,\\s*");
                t = TABLES.get(m.group(2));
                k = Integer.parseInt(m.group(4));
                if (t == null) throw new IllegalArgumentException("table " + m.group(2) + " does not exist");
                if (!t.vectorColumn.equals(m.group(3))) throw new IllegalArgumentException(m.group(3) + " is not the vector column");
                q = (List<? extends Number>) params.get(0);
            } else {
                Map<String, Object> spec = ai.langstream.agents.vector.InterpolationUtils.buildObjectFromJson(query, Map.class, params);
                if (!spec.containsKey("vector")) throw new IllegalArgumentException("unsupported query for the cuda datasource: " + query);
                t = TABLES.get(spec.getOrDefault("table", "documents").toString());
                if (t == null) throw new IllegalArgumentException("table " + spec.get("table") + " does not exist");
                k = Integer.parseInt(spec.getOrDefault("topK", "10").toString());
                List<String> sel = (List<String>) spec.getOrDefault("select", List.of("*"));
                cols = sel.toArray(new String[0]);
                similarity = Boolean.parseBoolean(spec.getOrDefault("include-similarity", "true").toString());
                q = (List<? extends Number>) spec.get("vector");
            }
            if (q.size() != t.dim) {
                throw new IllegalArgumentException("query vector has " + q.size() + " dimensions, table has " + t.dim);
            }
            List<Map.Entry<Long, Float>> hits;
            if (batcher != null) {
                Pending p = new Pending(t, k, padded(q, t.storeDim), new CompletableFuture<>(), System.nanoTime());
                synchronized (pending) {
                    pending.add(p);
                    pending.notifyAll();
                }
                hits = p.hits().join();          // unchecked CompletionException on failure, a per-record error upstream
            } else {
                hits = search(t, List.of(padded(q, t.storeDim)), k).get(0);
            }
            List<Map<String, Object>> out = new ArrayList<>();
            synchronized (t) {
                for (Map.Entry<Long, Float> h : hits) {
                    Map<String, Object> row = t.rows.get(h.getKey());
                    Map<String, Object> projected = new LinkedHashMap<>();
                    if (cols.length == 1 && cols[0].equals("*")) {
                        projected.putAll(row);
                    } else {
                        for (String c : cols) {
                            if (!row.containsKey(c)) throw new IllegalArgumentException("unknown column " + c);
                            projected.put(c, row.get(c));
                        }
                    }
                    if (similarity) projected.put("similarity", h.getValue());
                    out.add(projected);
                }
            }
            return out;
        }

        /** One native call for a batch of (padded) query vectors of one table. */
        private static List<List<Map.Entry<Long, Float>>> search(Table t, List<FloatBuffer> queries, int k) {
            int b = queries.size();
            FloatBuffer q = ByteBuffer.allocateDirect(b * t.storeDim * 4).order(ByteOrder.nativeOrder()).asFloatBuffer();
            for (int i = 0; i < b; i++) {
                for (int d = 0; d < t.storeDim; d++) q.put(i * t.storeDim + d, queries.get(i).get(d));
            }
            LongBuffer ids = ByteBuffer.allocateDirect(b * k * 8).order(ByteOrder.nativeOrder()).asLongBuffer();
            FloatBuffer scores = ByteBuffer.allocateDirect(b * k * 4).order(ByteOrder.nativeOrder()).asFloatBuffer();
            synchronized (t) {
                t.search(q, b, k, ids, scores);
            }
            List<List<Map.Entry<Long, Float>>> out = new ArrayList<>(b);
            for (int i = 0; i < b; i++) {
                List<Map.Entry<Long, Float>> hits = new ArrayList<>(k);
                for (int j = 0; j < k && ids.get(i * k + j) >= 0; j++) hits.add(Map.entry(ids.get(i * k + j), scores.get(i * k + j)));
                out.add(hits);
            }
            return out;
        }

        private void run() {
            while (true) {
                List<Pending> batch = new ArrayList<>();
                synchronized (pending) {
                    try {
                        while (pending.isEmpty() && !closed) pending.wait();
                        if (pending.isEmpty()) return;
                        while (!closed) {
                            long age = System.nanoTime() - pending.peek().enqueuedNanos();
                            if (pending.size() >= maxBatch || age >= flushNanos) break;
                            pending.wait(Math.max((flushNanos - age) / 1_000_000, 1));
                        }
                        Pending first = pending.peek();   // one native call per (table, k)
                        for (java.util.Iterator<Pending> it = pending.iterator(); it.hasNext() && batch.size() < maxBatch; ) {
                            Pending p = it.next();
                            if (p.table() == first.table() && p.k() == first.k()) {
                                batch.add(p);
                                it.remove();
                            }
                        }
                    } catch (InterruptedException e) {
                        Thread.currentThread().interrupt();
                        return;
                    }
                }
                try {
                    List<List<Map.Entry<Long, Float>>> hits =
                            search(batch.get(0).table(), batch.stream().map(Pending::query).toList(), batch.get(0).k());
                    for (int i = 0; i < batch.size(); i++) batch.get(i).hits().complete(hits.get(i));
                } catch (Throwable t) {
                    for (Pending p : batch) p.hits().completeExceptionally(t);
                }
            }
        }

        @Override
        public void close() {
            synchronized (pending) {
                closed = true;
                pending.notifyAll();
            }
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
        private java.nio.file.Path stateDir;
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
            // restore the table from the agent's persistent state directory, if it was saved there (close() below)
            Object dir = agentConfiguration.get("persistent-state-directory");
            stateDir = dir == null ? null : java.nio.file.Path.of(dir.toString());
            if (stateDir != null && !TABLES.containsKey(tableName) && Table.existsOnDisk(stateDir, tableName)) {
                try {
                    TABLES.put(tableName, Table.load(stateDir, tableName, devicesOf(datasourceConfig)[0]));
                } catch (java.io.IOException e) {
                    throw new IllegalStateException("cannot restore table " + tableName, e);
                }
            }
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
                    Table t = TABLES.computeIfAbsent(tableName, n -> new Table(vec.size(), vecCol, devicesOf(datasourceConfig)));
                    synchronized (t) {
                        long id = t.idOfPk.computeIfAbsent(pk, p -> t.nextId++);
                        t.rows.put(id, row);
                        LongBuffer ids = ByteBuffer.allocateDirect(8).order(ByteOrder.nativeOrder()).asLongBuffer();
                        ids.put(0, id);
                        t.upsert(ids, padded(vec, t.storeDim), 1);
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
                                t.delete(ids, 1);
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

        @Override
        public void close() throws Exception {
            // durable before the agent goes away: the source offsets were committed when the futures completed
            Table t = TABLES.get(tableName);
            if (stateDir != null && t != null && t.sharded == 0) {
                t.save(stateDir, tableName);
            }
        }
    }

    // ------------------------------------------------------------------ assets: `asset-type: "cuda-table"`
    /**
     * Counterpart of the jdbc-table asset (JdbcAssetsManagerProvider.java:33-144): `assets:` in an application can
     * create / drop a GPU table. Asset config: table-name, dimensions, vector-column, datasource (the resolved
     * resource: device / devices), optional persistent-state-directory (a saved table counts as existing and is loaded
     * on deploy). Register in META-INF/services/ai.langstream.api.runner.assets.AssetManagerProvider.
     */
    public static class AssetManagerProvider implements ai.langstream.api.runner.assets.AssetManagerProvider {
        @Override
        public boolean supports(String assetType) {
            return "cuda-table".equals(assetType);
        }

        @Override
        public ai.langstream.api.runner.assets.AssetManager createInstance(String assetType) {
            if (!supports(assetType)) throw new IllegalArgumentException(assetType);
            return new TableAssetManager();
        }
    }

    static class TableAssetManager implements ai.langstream.api.runner.assets.AssetManager {
        private Map<String, Object> config;

        @Override
        public void initialize(ai.langstream.api.model.AssetDefinition assetDefinition) {
            config = assetDefinition.getConfig();
            if (config.get("table-name") == null) throw new IllegalArgumentException("table-name is required");
        }

        private String name() {
            return config.get("table-name").toString();
        }

        private java.nio.file.Path dir() {
            Object d = config.get("persistent-state-directory");
            return d == null ? null : java.nio.file.Path.of(d.toString());
        }

        @SuppressWarnings("unchecked")
        private int[] devices() {
            Map<String, Object> ds = (Map<String, Object>) config.getOrDefault("datasource", Map.of());
            return devicesOf((Map<String, Object>) ds.getOrDefault("configuration", ds));
        }

        @Override
        public boolean assetExists() {
            return TABLES.containsKey(name()) || (dir() != null && Table.existsOnDisk(dir(), name()));
        }

        @Override
        public void deployAsset() throws Exception {
            if (TABLES.containsKey(name())) return;
            if (dir() != null && Table.existsOnDisk(dir(), name())) {
                TABLES.put(name(), Table.load(dir(), name(), devices()[0]));
                return;
            }
            Object dim = config.get("dimensions"), col = config.get("vector-column");
            if (dim == null || col == null) throw new IllegalArgumentException("dimensions and vector-column are required to create a cuda-table");
            TABLES.put(name(), new Table(Integer.parseInt(dim.toString()), col.toString(), devices()));
        }

        @Override
        public boolean deleteAssetIfExists() throws Exception {
            if (!assetExists()) return false;
            Table t = TABLES.remove(name());
            if (t != null) t.close();
            if (dir() != null) {
                java.nio.file.Files.deleteIfExists(dir().resolve(name() + ".lscidx"));
                java.nio.file.Files.deleteIfExists(dir().resolve(name() + ".rows.json"));
            }
            return true;
        }

        @Override
        public void close() {}
    }
}
