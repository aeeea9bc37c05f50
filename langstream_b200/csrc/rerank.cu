// MMR re-rank on device: the step that follows query-vector-db in the reference's
// chatbot pipeline.  Restates ReRankAgent (langstream-agents/langstream-ai-agents/src/main/java/ai/langstream/
// ai/agents/rerank/ReRankAgent.java): rerankMMR :155-180, calculateTopDocumentMMR_BM25_CosineSimilarity
// :182-222, calculateDiversity :224-251, cosineSimilarity / euclideanNorm / dotProduct :254-273, :361-372,
// calculateBM25Scores :275-302, calculateIDF :335-343.
//
// The text side (tokenise = split on \s+, term frequencies) stays on the host: the kernel receives, per
// document, its length in tokens and the frequency of every QUERY term (one column per query token, duplicates
// included, in query order - which is what the reference's loops visit), and the document embeddings.
// Arithmetic is kept in the reference's types and order so that the selection is identical, not just close:
// fp32 sequential sums without FMA contraction for norms and dot products, (float) sqrt of the widened sum,
// fp64 for BM25, diversity and the MMR score, first maximum wins (strict >).
//
// One CTA does the whole selection: the pairwise cosine matrix (n <= 128 documents: 64 KB of shared memory)
// is computed once, then one iteration per picked document with BM25 re-evaluated over the documents still
// remaining (N, avgdl and the document frequencies change as documents leave the pool, as in the reference).
#include <algorithm>
#include <cstring>
#include <mutex>
#include <utility>
#include <vector>

#include "common.h"

namespace lsc {
namespace {

constexpr int RR_MAX_DOCS = 128;
constexpr int RR_MAX_TERMS = 64;
constexpr int RR_THREADS = 256;

__global__ void __launch_bounds__(RR_THREADS, 1)
rerank_mmr_kernel(const float* __restrict__ vecs_all, int n, int dim, const int32_t* __restrict__ tf_all,
                  const int32_t* __restrict__ doc_len_all, int n_terms, int max_docs, double lambda, double k1, double b,
                  int32_t* __restrict__ out_order_all, int order_stride, int32_t* __restrict__ out_count_all, int count_stride) {
    // one CTA per query of a batch (blockIdx.x): every query brings its own n documents
    const float* __restrict__ vecs = vecs_all + static_cast<size_t>(blockIdx.x) * n * dim;
    const int32_t* __restrict__ tf = tf_all + static_cast<size_t>(blockIdx.x) * n * n_terms;
    const int32_t* __restrict__ doc_len = doc_len_all + static_cast<size_t>(blockIdx.x) * n;
    int32_t* __restrict__ out_order = out_order_all + static_cast<size_t>(blockIdx.x) * order_stride;
    int32_t* __restrict__ out_count = out_count_all + static_cast<size_t>(blockIdx.x) * count_stride;
    extern __shared__ float cosm[];  // [n][n]
    __shared__ float norm[RR_MAX_DOCS];
    __shared__ double score[RR_MAX_DOCS];
    __shared__ double idf[RR_MAX_TERMS];
    __shared__ int remaining[RR_MAX_DOCS];
    __shared__ int ranked[RR_MAX_DOCS];
    __shared__ int n_rem, n_rank;
    __shared__ double avgdl;

    const int tid = threadIdx.x;
    // euclideanNorm: float accumulation in index order, (float) Math.sqrt(double)
    for (int i = tid; i < n; i += RR_THREADS) {
        const float* v = vecs + static_cast<size_t>(i) * dim;
        float s = 0.0f;
        for (int k = 0; k < dim; ++k) s = __fadd_rn(s, __fmul_rn(v[k], v[k]));
        norm[i] = static_cast<float>(sqrt(static_cast<double>(s)));
        remaining[i] = i;
    }
    if (tid == 0) {
        n_rem = n;
        n_rank = 0;
    }
    __syncthreads();
    // cosineSimilarity for every pair (symmetric in fp32: products and the norm product commute)
    for (int p = tid; p < n * n; p += RR_THREADS) {
        const int i = p / n, j = p % n;
        if (j < i) continue;
        float c = 0.0f;
        if (norm[i] != 0.0f && norm[j] != 0.0f) {
            const float* a = vecs + static_cast<size_t>(i) * dim;
            const float* bb = vecs + static_cast<size_t>(j) * dim;
            float dot = 0.0f;
            for (int k = 0; k < dim; ++k) dot = __fadd_rn(dot, __fmul_rn(a[k], bb[k]));
            c = __fdiv_rn(dot, __fmul_rn(norm[i], norm[j]));
        }
        cosm[i * n + j] = c;
        cosm[j * n + i] = c;
    }
    __syncthreads();

    while (true) {
        const int N = n_rem, R = n_rank;
        if (N == 0 || R >= max_docs) break;
        // calculateAverageDocumentLength and calculateIDF over the documents still in the pool
        if (tid == 0) {
            int total = 0;
            for (int j = 0; j < N; ++j) total += doc_len[remaining[j]];
            avgdl = static_cast<double>(total) / static_cast<double>(N);
        }
        for (int t = tid; t < n_terms; t += RR_THREADS) {
            int df = 0;
            for (int j = 0; j < N; ++j) df += tf[remaining[j] * n_terms + t] > 0 ? 1 : 0;
            idf[t] = log(__dadd_rn(__ddiv_rn(__dadd_rn(static_cast<double>(N - df), 0.5), __dadd_rn(static_cast<double>(df), 0.5)), 1.0));
        }
        __syncthreads();
        for (int j = tid; j < N; j += RR_THREADS) {
            const int d = remaining[j];
            const double dl = static_cast<double>(doc_len[d]);
            // k1 * (1 - b + b * (documentLength / avgdl)), evaluated left to right as written in the reference
            const double norm_len = __dmul_rn(k1, __dadd_rn(__dsub_rn(1.0, b), __dmul_rn(b, __ddiv_rn(dl, avgdl))));
            double rel = 0.0;
            for (int t = 0; t < n_terms; ++t) {
                const double f = static_cast<double>(tf[d * n_terms + t]);
                const double numerator = __dmul_rn(f, __dadd_rn(k1, 1.0));
                const double denominator = __dadd_rn(f, norm_len);
                rel = __dadd_rn(rel, __dmul_rn(idf[t], __ddiv_rn(numerator, denominator)));
            }
            double diversity = 0.0;
            if (R > 0) {
                double sum = 0.0;
                for (int r = 0; r < R; ++r) sum = __dadd_rn(sum, static_cast<double>(cosm[d * n + ranked[r]]));
                diversity = __ddiv_rn(sum, static_cast<double>(R));
            }
            score[j] = __dsub_rn(__dmul_rn(lambda, rel), __dmul_rn(__dsub_rn(1.0, lambda), diversity));
        }
        __syncthreads();
        if (tid == 0) {
            int top = -1;
            double top_score = -INFINITY;
            for (int j = 0; j < N; ++j) {
                if (score[j] > top_score) {  // strict: the first maximum wins; NaN never wins
                    top_score = score[j];
                    top = j;
                }
            }
            if (top < 0) {
                // the reference throws IllegalStateException("topDocument is null"): every score was NaN
                n_rem = 0;
                n_rank = -1 - R;
            } else {
                ranked[R] = remaining[top];
                out_order[R] = remaining[top];
                for (int j = top; j + 1 < N; ++j) remaining[j] = remaining[j + 1];
                n_rem = N - 1;
                n_rank = R + 1;
            }
        }
        __syncthreads();
    }
    if (tid == 0) *out_count = n_rank;
}

int check_args(int n, int dim, int n_terms, int max_docs) {
    if (n < 0 || n > RR_MAX_DOCS) return fail(LSC_E_UNSUPPORTED, "rerank: at most 128 documents per call");
    if (dim < 1) return fail(LSC_E_INVALID, "rerank: dim must be positive");
    if (n_terms < 0 || n_terms > RR_MAX_TERMS) return fail(LSC_E_UNSUPPORTED, "rerank: at most 64 query terms");
    if (max_docs < 0) return fail(LSC_E_INVALID, "rerank: max must not be negative");
    return LSC_OK;
}

int launch(const float* d_vecs, int n, int dim, const int32_t* d_tf, const int32_t* d_len, int n_terms, int max_docs,
           double lambda, double k1, double b, int32_t* d_order, int32_t* d_count, cudaStream_t st, int batch = 1,
           int order_stride = 0, int count_stride = 0) {
    const size_t smem = static_cast<size_t>(n) * n * sizeof(float);
    static bool configured[64] = {};
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(rerank_mmr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      RR_MAX_DOCS * RR_MAX_DOCS * static_cast<int>(sizeof(float))));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    rerank_mmr_kernel<<<batch, RR_THREADS, smem, st>>>(d_vecs, n, dim, d_tf, d_len, n_terms, max_docs, lambda, k1, b, d_order,
                                                       order_stride, d_count, count_stride);
    LSC_CUDA(cudaGetLastError());
    return LSC_OK;
}

constexpr int kMaxDevices = 64;
struct HostStaging {
    std::mutex mu;
    unsigned char* d = nullptr;  // device block
    unsigned char* h = nullptr;  // pinned host twin
    size_t cap = 0;
    cudaStream_t stream = nullptr;
};
HostStaging g_staging[kMaxDevices];

}  // namespace
}  // namespace lsc

using namespace lsc;

extern "C" int lsc_rerank_mmr_dev(const float* d_doc_vecs, int32_t n_docs, int32_t dim, const int32_t* d_term_freq,
                                  const int32_t* d_doc_len, int32_t n_terms, int32_t max_docs, double lambda, double k1,
                                  double b, int32_t* d_out_order, int32_t* d_out_count, void* stream) {
    LSC_TRY(check_args(n_docs, dim, n_terms, max_docs));
    if (d_out_order == nullptr || d_out_count == nullptr) return fail(LSC_E_INVALID, "lsc_rerank_mmr_dev: null output");
    if (n_docs > 0 && (d_doc_vecs == nullptr || d_doc_len == nullptr || (n_terms > 0 && d_term_freq == nullptr)))
        return fail(LSC_E_INVALID, "lsc_rerank_mmr_dev: null input");
    return launch(d_doc_vecs, n_docs, dim, d_term_freq, d_doc_len, n_terms, max_docs, lambda, k1, b, d_out_order,
                  d_out_count, static_cast<cudaStream_t>(stream));
}

extern "C" int lsc_rerank_mmr_batch_dev(const float* d_doc_vecs, int32_t n_queries, int32_t n_docs, int32_t dim,
                                        const int32_t* d_term_freq, const int32_t* d_doc_len, int32_t n_terms,
                                        int32_t max_docs, double lambda, double k1, double b, int32_t* d_out_order,
                                        int32_t order_stride, int32_t* d_out_count, int32_t count_stride, void* stream) {
    LSC_TRY(check_args(n_docs, dim, n_terms, max_docs));
    if (n_queries < 0 || n_queries > 65535) return fail(LSC_E_INVALID, "lsc_rerank_mmr_batch_dev: n_queries must be in [0, 65535]");
    if (n_queries == 0) return LSC_OK;
    if (order_stride < n_docs || count_stride < 1) return fail(LSC_E_INVALID, "lsc_rerank_mmr_batch_dev: strides too small");
    if (d_out_order == nullptr || d_out_count == nullptr) return fail(LSC_E_INVALID, "lsc_rerank_mmr_batch_dev: null output");
    if (n_docs > 0 && (d_doc_vecs == nullptr || d_doc_len == nullptr || (n_terms > 0 && d_term_freq == nullptr)))
        return fail(LSC_E_INVALID, "lsc_rerank_mmr_batch_dev: null input");
    return launch(d_doc_vecs, n_docs, dim, d_term_freq, d_doc_len, n_terms, max_docs, lambda, k1, b, d_out_order, d_out_count,
                  static_cast<cudaStream_t>(stream), n_queries, order_stride, count_stride);
}

// ReRankAgent.tokenise (ReRankAgent.java:314-316): Java's text.split("\\s+") with \s = [ \t\n\x0B\f\r] - trailing empty
// strings are dropped, a leading empty string (text starting with white space) is kept, "" gives one empty token,
// an all-white-space text gives none.
static void java_split(const char* s, size_t n, std::vector<std::pair<const char*, size_t>>* out) {
    out->clear();
    auto ws = [](char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\x0B' || c == '\f' || c == '\r'; };
    if (n == 0) {
        out->emplace_back(s, 0);
        return;
    }
    size_t i = 0;
    if (ws(s[0])) {
        while (i < n && ws(s[i])) ++i;
        if (i == n) return;           // nothing but white space: every token is empty and trailing
        out->emplace_back(s, 0);      // the leading empty token
    }
    while (i < n) {
        const size_t b0 = i;
        while (i < n && !ws(s[i])) ++i;
        out->emplace_back(s + b0, i - b0);
        while (i < n && ws(s[i])) ++i;
    }
}

extern "C" int lsc_rerank_term_statistics(const char* const* doc_texts, const int32_t* doc_nbytes, int32_t n_docs,
                                          const char* query_text, int32_t query_nbytes, int32_t max_terms,
                                          int32_t* term_freq, int32_t* doc_len, int32_t* n_terms) {
    if (n_docs < 0 || max_terms < 0 || n_terms == nullptr || (n_docs > 0 && (doc_texts == nullptr || doc_len == nullptr)))
        return fail(LSC_E_INVALID, "lsc_rerank_term_statistics: bad argument");
    std::vector<std::pair<const char*, size_t>> q, d;
    const char* qt = query_text == nullptr ? "" : query_text;
    java_split(qt, query_text == nullptr ? 0 : (query_nbytes >= 0 ? static_cast<size_t>(query_nbytes) : strlen(qt)), &q);
    *n_terms = static_cast<int32_t>(q.size());
    if (static_cast<int64_t>(q.size()) > max_terms)
        return fail(LSC_E_UNSUPPORTED, "lsc_rerank_term_statistics: the query has more terms than max_terms");
    if (n_docs > 0 && max_terms > 0 && term_freq == nullptr) return fail(LSC_E_INVALID, "lsc_rerank_term_statistics: term_freq is null");
    for (int32_t i = 0; i < n_docs; ++i) {
        const char* t = doc_texts[i] == nullptr ? "" : doc_texts[i];
        const size_t nb = doc_texts[i] == nullptr ? 0 : (doc_nbytes != nullptr ? static_cast<size_t>(doc_nbytes[i]) : strlen(t));
        java_split(t, nb, &d);
        doc_len[i] = static_cast<int32_t>(d.size());
        int32_t* row = term_freq + static_cast<size_t>(i) * max_terms;
        for (int32_t j = 0; j < max_terms; ++j) row[j] = 0;
        // at most 64 query terms against a few hundred document tokens: a plain comparison loop beats building a map
        for (size_t j = 0; j < q.size(); ++j) {
            int32_t c = 0;
            for (const auto& tok : d) c += (tok.second == q[j].second && memcmp(tok.first, q[j].first, tok.second) == 0) ? 1 : 0;
            row[j] = c;
        }
    }
    return LSC_OK;
}

extern "C" int lsc_rerank_mmr(const float* doc_vecs, int32_t n_docs, int32_t dim, const int32_t* term_freq,
                              const int32_t* doc_len, int32_t n_terms, int32_t max_docs, double lambda, double k1,
                              double b, int32_t* out_order, int32_t* out_count) {
    LSC_TRY(check_args(n_docs, dim, n_terms, max_docs));
    if (out_order == nullptr || out_count == nullptr) return fail(LSC_E_INVALID, "lsc_rerank_mmr: null output");
    if (n_docs > 0 && (doc_vecs == nullptr || doc_len == nullptr || (n_terms > 0 && term_freq == nullptr)))
        return fail(LSC_E_INVALID, "lsc_rerank_mmr: null input");
    if (n_docs == 0 || max_docs == 0) {
        *out_count = 0;
        return LSC_OK;
    }
    const size_t nv = static_cast<size_t>(n_docs) * dim * sizeof(float);
    const size_t nt = static_cast<size_t>(n_docs) * (n_terms > 0 ? n_terms : 1) * sizeof(int32_t);
    const size_t nl = static_cast<size_t>(n_docs) * sizeof(int32_t);
    const size_t no = static_cast<size_t>(n_docs + 1) * sizeof(int32_t);
    // one block: vectors | term frequencies | lengths | order + count
    const size_t o_t = (nv + 255) & ~static_cast<size_t>(255), o_l = (o_t + nt + 255) & ~static_cast<size_t>(255),
                 o_o = (o_l + nl + 255) & ~static_cast<size_t>(255);
    const size_t total = o_o + no;
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(LSC_E_UNSUPPORTED, "lsc_rerank_mmr: device ordinal too large");
    // Per-device staging kept between calls (a device block, its pinned host twin and a stream): at <= 128
    // documents the kernel takes microseconds, so a cudaMalloc / cudaFree pair and blocking copies on the
    // default stream per call were most of the call.  Calls on one device are serialised by the mutex.
    HostStaging& hs = g_staging[dev];
    std::lock_guard<std::mutex> lock(hs.mu);
    if (hs.stream == nullptr) LSC_CUDA(cudaStreamCreateWithFlags(&hs.stream, cudaStreamNonBlocking));
    if (total > hs.cap) {
        if (hs.d) cudaFree(hs.d);
        if (hs.h) cudaFreeHost(hs.h);
        hs.d = hs.h = nullptr;
        hs.cap = 0;
        const size_t cap = std::max<size_t>(total, 1 << 20);
        LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&hs.d), cap));
        LSC_CUDA(cudaMallocHost(reinterpret_cast<void**>(&hs.h), cap));
        hs.cap = cap;
    }
    memcpy(hs.h, doc_vecs, nv);
    if (n_terms > 0) memcpy(hs.h + o_t, term_freq, nt);
    memcpy(hs.h + o_l, doc_len, nl);
    LSC_CUDA(cudaMemcpyAsync(hs.d, hs.h, o_o, cudaMemcpyHostToDevice, hs.stream));  // one upload
    int32_t* d_order = reinterpret_cast<int32_t*>(hs.d + o_o);
    LSC_TRY(launch(reinterpret_cast<const float*>(hs.d), n_docs, dim, reinterpret_cast<const int32_t*>(hs.d + o_t),
                   reinterpret_cast<const int32_t*>(hs.d + o_l), n_terms, max_docs, lambda, k1, b, d_order,
                   d_order + n_docs, hs.stream));
    LSC_CUDA(cudaMemcpyAsync(hs.h + o_o, d_order, no, cudaMemcpyDeviceToHost, hs.stream));
    LSC_CUDA(cudaStreamSynchronize(hs.stream));
    const int32_t* host = reinterpret_cast<const int32_t*>(hs.h + o_o);
    const int32_t count = host[n_docs];
    if (count < 0) return fail(LSC_E_INVALID, "rerank: no document could be selected (all scores are NaN)");
    for (int32_t i = 0; i < count; ++i) out_order[i] = host[i];
    *out_count = count;
    return LSC_OK;
}
