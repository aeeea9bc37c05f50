package ai.langstream.ai.agents.services.impl;

import ai.langstream.ai.agents.services.ServiceProviderProvider;
import ai.langstream.api.runner.code.MetricsReporter;
import ai.langstream.cuda.LangstreamCuda;
import com.datastax.oss.streaming.ai.completions.CompletionsService;
import com.datastax.oss.streaming.ai.embeddings.EmbeddingsService;
import com.datastax.oss.streaming.ai.services.ServiceProvider;
import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.nio.FloatBuffer;
import java.nio.IntBuffer;
import java.nio.channels.FileChannel;
import java.nio.file.Path;
import java.nio.file.StandardOpenOption;
import java.util.AbstractList;
import java.util.ArrayList;
import java.util.List;
import java.util.Map;
import java.util.concurrent.CompletableFuture;

/**
 * `cuda` AI service: a ServiceProviderProvider found by ServiceLoader (register this class in
 * META-INF/services/ai.langstream.ai.agents.services.ServiceProviderProvider). It claims agent
 * configurations that carry a "cuda" entry; HuggingFaceProvider claims "huggingface"
 * (HuggingFaceProvider.java:52-55), so the two never compete.
 */
public class CudaProvider implements ServiceProviderProvider {

    @Override
    public boolean supports(Map<String, Object> agentConfiguration) {
        return agentConfiguration.containsKey("cuda");
    }

    @Override
    @SuppressWarnings("unchecked")
    public ServiceProvider createImplementation(
            Map<String, Object> agentConfiguration, MetricsReporter metricsReporter) {
        return new CudaServiceProvider(
                (Map<String, Object>) agentConfiguration.get("cuda"), metricsReporter);
    }

    static class CudaServiceProvider implements ServiceProvider {
        private final Map<String, Object> providerConfiguration;
        private final MetricsReporter metrics;
        private final List<CudaEmbeddingsService> services = new ArrayList<>();

        CudaServiceProvider(Map<String, Object> providerConfiguration, MetricsReporter metrics) {
            this.providerConfiguration = providerConfiguration == null ? Map.of() : providerConfiguration;
            this.metrics = metrics;
        }

        @Override
        public CompletionsService getCompletionsService(Map<String, Object> additionalConfiguration) {
            throw new IllegalArgumentException("the cuda provider does not offer a completions service");
        }

        @Override
        public EmbeddingsService getEmbeddingsService(Map<String, Object> stepConfig) throws Exception {
            // same step keys as HuggingFaceProvider.getEmbeddingsService (:82-139)
            String modelUrl = (String) stepConfig.get("model-url"); // directory: config.json, vocab.txt, weights.f32
            checkIfUrlIsAllowed(modelUrl);
            if (!modelUrl.startsWith("file://")) {
                throw new IllegalArgumentException("the cuda provider loads models from file:// urls only, got " + modelUrl);
            }
            int device = Integer.parseInt(providerConfiguration.getOrDefault("device", "0").toString());
            int maxBatch = Integer.parseInt(stepConfig.getOrDefault("batch-size", "1024").toString());
            long flushMicros = (long) (1000 * Double.parseDouble(stepConfig.getOrDefault("flush-interval", "2").toString()));
            CudaEmbeddingsService svc = new CudaEmbeddingsService(
                    Path.of(modelUrl.substring("file://".length())), device, maxBatch, Math.max(flushMicros, 100), metrics);
            services.add(svc);
            return svc;
        }

        /** AbstractHuggingFaceEmbeddingService.checkIfUrlIsAllowed (:174-181) with the same ALLOWED_HF_URLS list (:55-70). */
        static void checkIfUrlIsAllowed(String modelUrl) throws IllegalAccessException {
            String prop = System.getenv("ALLOWED_HF_URLS");
            if (prop == null || prop.isEmpty()) {
                prop = System.getProperty("ALLOWED_HF_URLS");
            }
            if (prop == null || prop.isEmpty()) {
                prop = "file://,djl://ai.djl.huggingface.pytorch";
            }
            if (modelUrl != null) {
                for (String prefix : prop.split(",")) {
                    if (modelUrl.startsWith(prefix)) {
                        return;
                    }
                }
            }
            throw new IllegalAccessException("modelUrl is not allowed: " + modelUrl);
        }

        @Override
        public void close() {
            services.forEach(CudaEmbeddingsService::close);
        }
    }

    /**
     * EmbeddingsService (EmbeddingsService.java:24-37) over lsc_tokenizer_encode + lsc_encoder_submit / lsc_encoder_wait.
     *
     * <p>The runtime hands texts over one record at a time (CompositeAgentProcessor.java:225-227; with
     * flush-interval 0 OrderedAsyncBatchExecutor flushes every add, :113-118), so requests from all callers meet
     * in a micro-batcher: a batch goes to the GPU when maxBatch texts are waiting or the oldest request is
     * flushMicros old. One batcher thread tokenises batch i+1 (native, multi-threaded) while the GPU works on
     * batch i; two page-locked staging slots are reused for the whole life of the service. Never throws from
     * computeEmbeddings: failures complete the future exceptionally
     * (AbstractHuggingFaceEmbeddingService.computeEmbeddings :213-223).
     */
    static class CudaEmbeddingsService implements EmbeddingsService {
        private record Request(List<String> texts, CompletableFuture<List<List<Double>>> future, long enqueuedNanos) {}

        /** One staging slot: direct buffers handed to the native module, valid until its ticket has been waited for. */
        private final class Slot {
            ByteBuffer utf8 = direct(1 << 20);
            IntBuffer textNbytes, tokens, cu;
            FloatBuffer out;
            List<Request> requests;
            long ticket = -1;
            int nTexts;

            void ensure(int nTexts, int utf8Bytes) {
                if (utf8.capacity() < utf8Bytes) utf8 = direct(utf8Bytes + utf8Bytes / 4);
                if (textNbytes == null || textNbytes.capacity() < nTexts) {
                    int cap = Math.max(nTexts, maxBatch);
                    textNbytes = direct(cap * 4).asIntBuffer();
                    cu = direct((cap + 1) * 4).asIntBuffer();
                    tokens = direct(cap * maxPosition * 4).asIntBuffer();
                    out = direct(cap * dim * 4).asFloatBuffer();
                }
            }
        }

        private final long encoder;
        private final long tokenizer;
        private final int dim;
        private final int maxPosition;
        private final int maxBatch;
        private final long flushNanos;
        private final java.util.ArrayDeque<Request> pending = new java.util.ArrayDeque<>();
        private final Slot[] slots = {null, null};
        private final Thread batcher;
        private volatile boolean closed;
        private final MetricsReporter.Counter embeddingsComputed;

        CudaEmbeddingsService(Path modelDir, int device, int maxBatch, long flushMicros, MetricsReporter metrics) throws Exception {
            Map<String, Object> cfg = EmbeddingsService.createObjectMapper()
                    .readValue(modelDir.resolve("config.json").toFile(), Map.class);
            // config.json's model_type picks the graph: "bert" (all-MiniLM, BGE, e5) or "mpnet" (all-mpnet-base-v2:
            // no token types, positions from 2, bucketed relative-position bias in the attention kernel)
            boolean mpnet = "mpnet".equals(String.valueOf(cfg.get("model_type")));
            int positions = ((Number) cfg.get("max_position_embeddings")).intValue();
            this.maxPosition = mpnet ? positions - 2 : positions;
            this.maxBatch = maxBatch;
            this.flushNanos = flushMicros * 1000;
            boolean lower = Boolean.parseBoolean(String.valueOf(cfg.getOrDefault("do_lower_case", "true")));
            this.tokenizer = LangstreamCuda.tokenizerCreate(
                    java.nio.file.Files.readAllBytes(modelDir.resolve("vocab.txt")), lower, maxPosition, 0);
            if (mpnet) {
                LangstreamCuda.tokenizerSetSpecial(tokenizer, "<s>", "</s>", null);  // template <s> A </s>
            }
            try (FileChannel ch = FileChannel.open(modelDir.resolve("weights.f32"), StandardOpenOption.READ)) {
                ByteBuffer weights = ByteBuffer.allocateDirect((int) ch.size()).order(ByteOrder.nativeOrder());
                while (weights.hasRemaining() && ch.read(weights) >= 0) {}
                weights.flip();
                this.encoder = LangstreamCuda.encoderCreate(
                        ((Number) cfg.get("num_hidden_layers")).intValue(),
                        ((Number) cfg.get("hidden_size")).intValue(),
                        ((Number) cfg.get("num_attention_heads")).intValue(),
                        ((Number) cfg.get("intermediate_size")).intValue(),
                        ((Number) cfg.get("vocab_size")).intValue(),
                        positions,
                        mpnet ? 0 : ((Number) cfg.getOrDefault("type_vocab_size", 2)).intValue(),
                        ((Number) cfg.getOrDefault("layer_norm_eps", 1e-12)).floatValue(),
                        mpnet ? 1 : 0,
                        mpnet ? ((Number) cfg.getOrDefault("relative_attention_num_buckets", 32)).intValue() : 0,
                        mpnet ? 128 : 0,
                        weights, device);
            }
            this.dim = LangstreamCuda.encoderDim(encoder);
            this.embeddingsComputed = metrics == null ? null
                    : metrics.counter("embeddings_computed", "embeddings computed on the GPU");
            slots[0] = new Slot();
            slots[1] = new Slot();
            this.batcher = new Thread(this::run, "lsc-embeddings-batcher");
            this.batcher.setDaemon(true);
            this.batcher.start();
        }

        @Override
        public CompletableFuture<List<List<Double>>> computeEmbeddings(List<String> texts) {
            CompletableFuture<List<List<Double>>> f = new CompletableFuture<>();
            if (texts.isEmpty()) {
                f.complete(List.of());
                return f;
            }
            synchronized (pending) {
                if (closed) {
                    f.completeExceptionally(new RuntimeException("failed to compute embeddings", new IllegalStateException("closed")));
                    return f;
                }
                pending.add(new Request(List.copyOf(texts), f, System.nanoTime()));
                pending.notifyAll();
            }
            return f;
        }

        /** Batcher thread: collect -> tokenise -> submit; then complete the batch submitted one turn earlier. */
        private void run() {
            int turn = 0;
            Slot inFlight = null;
            while (true) {
                List<Request> batch = new ArrayList<>();
                synchronized (pending) {
                    try {
                        while (pending.isEmpty() && !closed && inFlight == null) pending.wait();
                        if (!pending.isEmpty()) {
                            // the window: full batch, or the oldest request is old enough (do not hold back a batch
                            // while an earlier one is still to be completed)
                            while (!closed && inFlight == null) {
                                int waiting = pending.stream().mapToInt(r -> r.texts().size()).sum();
                                long age = System.nanoTime() - pending.peek().enqueuedNanos();
                                if (waiting >= maxBatch || age >= flushNanos) break;
                                pending.wait(Math.max((flushNanos - age) / 1_000_000, 1));
                            }
                            int n = 0;
                            while (!pending.isEmpty() && (batch.isEmpty() || n + pending.peek().texts().size() <= maxBatch)) {
                                Request r = pending.poll();
                                batch.add(r);
                                n += r.texts().size();
                            }
                        }
                    } catch (InterruptedException e) {
                        Thread.currentThread().interrupt();
                        return;
                    }
                }
                Slot next = null;
                if (!batch.isEmpty()) {
                    next = slots[turn++ & 1];
                    try {
                        submit(next, batch);
                    } catch (Throwable t) {   // the whole batch fails together (ComputeAIEmbeddingsStep.java:215-220)
                        for (Request r : batch) r.future().completeExceptionally(new RuntimeException("failed to compute embeddings", t));
                        next = null;
                    }
                }
                if (inFlight != null) complete(inFlight);   // batch i-1: its download overlapped the work above
                inFlight = next;
                if (closed && inFlight == null) {
                    synchronized (pending) {
                        if (pending.isEmpty()) return;
                    }
                }
            }
        }

        private void submit(Slot s, List<Request> batch) {
            int n = 0, bytes = 0;
            List<byte[]> raw = new ArrayList<>();
            for (Request r : batch) {
                for (String t : r.texts()) {
                    byte[] b = t.getBytes(java.nio.charset.StandardCharsets.UTF_8);
                    raw.add(b);
                    bytes += b.length;
                    n++;
                }
            }
            s.ensure(n, bytes);
            s.utf8.clear();
            for (int i = 0; i < n; i++) {
                s.utf8.put(raw.get(i));
                s.textNbytes.put(i, raw.get(i).length);
            }
            // host tokenisation stays on the host, but natively and on all cores (lsc_tokenizer_encode); an over-long
            // text is truncated to max_position by the tokenizer, like DJL's truncation=true default
            LangstreamCuda.tokenizerEncode(tokenizer, s.utf8, s.textNbytes, n, s.tokens, s.cu);
            s.requests = batch;
            s.nTexts = n;
            s.ticket = LangstreamCuda.encoderSubmit(encoder, s.tokens, s.cu, n, s.out);
        }

        private void complete(Slot s) {
            try {
                LangstreamCuda.encoderWait(encoder, s.ticket);
                if (embeddingsComputed != null) embeddingsComputed.count(s.nTexts);
                int row = 0;
                for (Request r : s.requests) {
                    // float -> double widening (HuggingFaceEmbeddingService.convertOutput :52-63) without boxing dim
                    // Doubles per text: one float[] per text behind a List<Double> view; nothing aliases the slot
                    List<List<Double>> result = new ArrayList<>(r.texts().size());
                    for (int i = 0; i < r.texts().size(); i++, row++) {
                        final float[] v = new float[dim];
                        s.out.position(row * dim);
                        s.out.get(v);
                        result.add(new AbstractList<>() {
                            @Override public Double get(int j) { return (double) v[j]; }
                            @Override public int size() { return v.length; }
                        });
                    }
                    r.future().complete(result);
                }
            } catch (Throwable t) {
                for (Request r : s.requests) r.future().completeExceptionally(new RuntimeException("failed to compute embeddings", t));
            }
        }

        private static ByteBuffer direct(int bytes) {
            return ByteBuffer.allocateDirect(Math.max(bytes, 4)).order(ByteOrder.nativeOrder());
        }

        @Override
        public void close() {
            synchronized (pending) {
                closed = true;
                pending.notifyAll();
            }
            try {
                batcher.join(10_000);
            } catch (InterruptedException e) {
                Thread.currentThread().interrupt();
            }
            LangstreamCuda.tokenizerDestroy(tokenizer);
            LangstreamCuda.encoderDestroy(encoder);
        }
    }
}
