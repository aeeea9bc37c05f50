/*
 * CPU oracle for the brute-force cosine top-k behind `query-vector-db`.  TEST INFRASTRUCTURE ONLY —
 * never linked into or called from the product path.
 *
 * The reference delegates the scan to HerdDB 0.28.0 (third-party, not in the tree):
 *   SELECT … ORDER BY cosine_similarity(col, CAST(? AS FLOAT ARRAY)) DESC LIMIT k
 *   (examples/applications/docker-chatbot/chatbot.yaml:44; executed by JdbcDataSourceImpl.fetchData,
 *    langstream-agents/langstream-ai-agents/src/main/java/ai/langstream/ai/agents/datasource/impl/JdbcDataSourceProvider.java:103-124).
 * The only in-tree statement of the cosine is ReRankAgent
 * (langstream-agents/langstream-ai-agents/src/main/java/ai/langstream/ai/agents/rerank/ReRankAgent.java):
 *   euclideanNorm :254-260   fp32 running sum of squares, (float) Math.sqrt
 *   cosineSimilarity :262-273 0 when either norm is 0, else dot / (norm1 * norm2)
 *   dotProduct :361-372      fp32 running sum
 * which this file restates literally (fp32, sequential).  Ordering: score descending, ties by lower
 * row index (the tie rule is ours: HerdDB's is unknown and no reference test pins it).
 *
 * PARITY PIN STATUS: pinned against the known-answer pipeline test
 * langstream-runtime/langstream-runtime-impl/src/test/java/ai/langstream/kafka/RerankAgentRunnerIT.java:62-92,204
 * (tests/golden/rerank_it.json).  HerdDB's own top-k selection is NOT pinned by any reference test
 * (LIMIT 20 over 10 rows there; one row in JdbcWriterTest) — "parity unpinned" at that boundary.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static float euclidean_norm(const float* a, int n) {
    float sum = 0.0f;
    for (int i = 0; i < n; ++i) sum += a[i] * a[i];
    return (float)sqrt((double)sum);
}

static float dot_product(const float* a, const float* b, int n) {
    float r = 0.0f;
    for (int i = 0; i < n; ++i) r += a[i] * b[i];
    return r;
}

float oracle_cosine(const float* a, const float* b, int n) {
    float n1 = euclidean_norm(a, n);
    if (n1 == 0) return 0;
    float n2 = euclidean_norm(b, n);
    if (n2 == 0) return 0.0f;
    return dot_product(a, b, n) / (n1 * n2);
}

/* scores[r] = cosine(rows[r], q) for every row */
void oracle_cosine_scores(const float* rows, int64_t n, int dim, const float* q, float* scores) {
    for (int64_t r = 0; r < n; ++r) scores[r] = oracle_cosine(rows + r * dim, q, dim);
}

/* Stable top-k by insertion: score descending, ties by lower index.  ids/out_scores hold k entries,
 * unfilled entries are id -1 / score -inf. */
void oracle_topk(const float* scores, int64_t n, int k, int64_t* ids, float* out_scores) {
    int cnt = 0;
    for (int i = 0; i < k; ++i) {
        ids[i] = -1;
        out_scores[i] = -INFINITY;
    }
    for (int64_t r = 0; r < n; ++r) {
        float s = scores[r];
        if (cnt == k && !(s > out_scores[k - 1])) continue;
        int pos = cnt < k ? cnt : k - 1;
        while (pos > 0 && out_scores[pos - 1] < s) {
            out_scores[pos] = out_scores[pos - 1];
            ids[pos] = ids[pos - 1];
            --pos;
        }
        out_scores[pos] = s;
        ids[pos] = r;
        if (cnt < k) ++cnt;
    }
}

/* The whole query: B queries x n rows -> [B][k]. */
void oracle_search(const float* rows, int64_t n, int dim, const float* queries, int B, int k,
                   int64_t* out_ids, float* out_scores) {
    float* scores = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
    for (int b = 0; b < B; ++b) {
        oracle_cosine_scores(rows, n, dim, queries + (size_t)b * dim, scores);
        oracle_topk(scores, n, k, out_ids + (size_t)b * k, out_scores + (size_t)b * k);
    }
    free(scores);
}
