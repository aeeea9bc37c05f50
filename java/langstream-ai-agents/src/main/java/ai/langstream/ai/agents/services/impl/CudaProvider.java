package ai.langstream.ai.agents.services.impl;

import ai.langstream.ai.agents.services.ServiceProviderProvider;
import ai.langstream.api.runner.code.MetricsReporter;
import ai.langstream.cuda.LangstreamCuda;
import ai.djl.huggingface.tokenizers.Encoding;
import ai.djl.huggingface.tokenizers.HuggingFaceTokenizer;
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
import java.util.concurrent.ExecutorService;
import java.util.concurrent.Executors;

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
            String modelUrl = (String) stepConfig.get("model-url"); // directory: config.json, tokenizer.json, weights.f32
            if (modelUrl == null || !modelUrl.startsWith("file://")) {
                throw new IllegalAccessException("modelUrl is not allowed: " + modelUrl);
            }
            int device = Integer.parseInt(providerConfiguration.getOrDefault("device", "0").toString());
            CudaEmbeddingsService svc =
                    new CudaEmbeddingsService(Path.of(modelUrl.substring("file://".length())), device, metrics);
            services.add(svc);
            return svc;
        }

        @Override
        public void close() {
            services.forEach(CudaEmbeddingsService::close);
        }
    }

    /** EmbeddingsService (EmbeddingsService.java:24-37) over lsc_encoder_embed. Thread-safe. */
    static class CudaEmbeddingsService implements EmbeddingsService {
        private final long encoder;
        private final int dim;
        private final int maxPosition;
        private final HuggingFaceTokenizer tokenizer;
        private final ExecutorService gpuThread = Executors.newSingleThreadExecutor();
        private final MetricsReporter.Counter embeddingsComputed;

        CudaEmbeddingsService(Path modelDir, int device, MetricsReporter metrics) throws Exception {
            Map<String, Object> cfg = EmbeddingsService.createObjectMapper()
                    .readValue(modelDir.resolve("config.json").toFile(), Map.class);
            this.maxPosition = ((Number) cfg.get("max_position_embeddings")).intValue();
            this.tokenizer = HuggingFaceTokenizer.newInstance(modelDir.resolve("tokenizer.json"));
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
                        maxPosition,
                        ((Number) cfg.getOrDefault("type_vocab_size", 2)).intValue(),
                        ((Number) cfg.getOrDefault("layer_norm_eps", 1e-12)).floatValue(),
                        weights, device);
            }
            this.dim = LangstreamCuda.encoderDim(encoder);
            this.embeddingsComputed = metrics == null ? null
                    : metrics.counter("embeddings_computed", "embeddings computed on the GPU");
        }

        @Override
        public CompletableFuture<List<List<Double>>> computeEmbeddings(List<String> texts) {
            // Never throws: failures complete the future exceptionally, as
            // AbstractHuggingFaceEmbeddingService.computeEmbeddings does (:213-223).
            return CompletableFuture.supplyAsync(() -> {
                try {
                    return embed(texts);
                } catch (Exception e) {
                    throw new RuntimeException("failed to compute embeddings", e);
                }
            }, gpuThread);
        }

        private List<List<Double>> embed(List<String> texts) {
            int n = texts.size();
            long[][] ids = new long[n][];
            int total = 0;
            for (int i = 0; i < n; i++) {
                Encoding enc = tokenizer.encode(texts.get(i));   // host tokenisation (adds [CLS]/[SEP], truncates)
                ids[i] = enc.getIds();
                total += ids[i].length;
            }
            IntBuffer tokens = direct(total * 4).asIntBuffer();
            IntBuffer cu = direct((n + 1) * 4).asIntBuffer();
            int offset = 0;
            cu.put(0, 0);
            for (int i = 0; i < n; i++) {
                for (long t : ids[i]) {
                    tokens.put(offset++, (int) t);
                }
                cu.put(i + 1, offset);
            }
            FloatBuffer out = direct(n * dim * 4).asFloatBuffer();
            LangstreamCuda.encoderEmbed(encoder, tokens, cu, n, out);   // packed variable-length batch
            if (embeddingsComputed != null) {
                embeddingsComputed.count(n);
            }
            // float -> double widening (HuggingFaceEmbeddingService.convertOutput :52-63) without boxing
            // n*dim Doubles up front: one float[] per text, exposed as a List<Double> view.
            List<List<Double>> result = new ArrayList<>(n);
            for (int i = 0; i < n; i++) {
                final float[] row = new float[dim];
                out.position(i * dim);
                out.get(row);
                result.add(new AbstractList<>() {
                    @Override public Double get(int j) { return (double) row[j]; }
                    @Override public int size() { return row.length; }
                });
            }
            return result;
        }

        private static ByteBuffer direct(int bytes) {
            return ByteBuffer.allocateDirect(Math.max(bytes, 4)).order(ByteOrder.nativeOrder());
        }

        @Override
        public void close() {
            gpuThread.shutdown();
            LangstreamCuda.encoderDestroy(encoder);
        }
    }
}
