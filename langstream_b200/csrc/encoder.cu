// Encoder handle: owns one model replica on one GPU and runs the BERT forward pass
// (embeddings+LN -> L x [QKV GEMM, attention, out GEMM+residual, LN, FFN-up GEMM+GELU,
// FFN-down GEMM+residual, LN] -> mean-pool + L2) over a packed, variable-length batch.
// It is the native replacement of what the reference reaches through DJL:
//   AbstractHuggingFaceEmbeddingService ctor (model load)  …/embeddings/AbstractHuggingFaceEmbeddingService.java:110-172
//   compute(): predictor.predict(text) per text            …/embeddings/AbstractHuggingFaceEmbeddingService.java:183-207
// (paths under langstream-agents/langstream-ai-agents/src/main/java/com/datastax/oss/streaming/ai).
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.h"

namespace lsc {
namespace {

struct LayerW {
    __nv_bfloat16* w_qkv = nullptr;  // [3H,H]
    float* b_qkv = nullptr;          // [3H]
    __nv_bfloat16* w_o = nullptr;    // [H,H]
    float* b_o = nullptr;
    float* ln1_g = nullptr;
    float* ln1_b = nullptr;
    __nv_bfloat16* w_1 = nullptr;  // [F,H]
    float* b_1 = nullptr;
    __nv_bfloat16* w_2 = nullptr;  // [H,F]
    float* b_2 = nullptr;
    float* ln2_g = nullptr;
    float* ln2_b = nullptr;
    // LayerNorm folding (tcgen05 path): w_qkv / w_1 hold W diag(gamma_in); c = column sums of those
    // folded bf16 weights; d = bias + W beta_in; in_g / in_b = the LayerNorm applied to this layer's input
    // (ones / zeros for layer 0, whose input the embedding kernel already normalised).
    float* c_qkv = nullptr;
    float* d_qkv = nullptr;
    float* c_1 = nullptr;
    float* d_1 = nullptr;
    float* in_g = nullptr;
    float* in_b = nullptr;
    CUtensorMap tm_qkv, tm_o, tm_1, tm_2;
    CUtensorMap tm_1f, tm_2f;  // 128-row boxes of W1' / W2 for the fused feed-forward kernel
};

}  // namespace
}  // namespace lsc

using namespace lsc;

struct lsc_encoder {
    lsc_encoder_config cfg;
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    std::mutex mu;
    std::vector<void*> owned;  // every device allocation of the weights
    float* d_word = nullptr;
    float* d_pos = nullptr;
    float* d_type = nullptr;
    float* d_rel = nullptr;      // arch 1: relative-position bias table for the attention kernel (attention_rel_table)
    const float* d_pos0 = nullptr;  // position row of a sequence's first token (arch 1: row 2 of d_pos)
    int max_len = 0;             // longest sequence the position table serves
    float* d_emb_g = nullptr;
    float* d_emb_b = nullptr;
    std::vector<LayerW> layers;
    // activation workspace, sized for cap_tokens
    int64_t cap_tokens = 0;
    __nv_bfloat16 *x = nullptr, *x1 = nullptr, *tmp = nullptr, *qkv = nullptr, *ctx = nullptr, *ffn = nullptr;
    // staging for the host-pointer API
    int64_t cap_stage_tokens = 0, cap_stage_seqs = 0;
    int32_t* d_ids = nullptr;
    int32_t* d_cu = nullptr;
    float* d_out = nullptr;
    // asynchronous host-buffer path (lsc_encoder_submit / lsc_encoder_wait): two staging slots, copies on their
    // own streams so that batch i+1's upload and batch i-1's download overlap batch i's kernels
    struct AsyncSlot {
        int32_t* d_ids = nullptr;
        int32_t* d_cu = nullptr;
        float* d_out = nullptr;
        int64_t cap_tokens = 0, cap_seqs = 0;
        std::vector<int32_t> cu_host;
        cudaEvent_t ev_in = nullptr, ev_fwd = nullptr, ev_done = nullptr;
        int64_t ticket = -1;   // ticket in flight in this slot (-1: free)
        bool sync_done = false;  // the submission was served synchronously (multi-pass batch)
    };
    AsyncSlot slots[2];
    cudaStream_t stream_in = nullptr, stream_out = nullptr;
    int64_t next_ticket = 0;
    int32_t* d_tiles = nullptr;  // attention tile list of the current pass (attention_tc.cu): 2 * n_seq + 2 ints, then T row words
    int64_t cap_tiles = 0;
    float* stats_x = nullptr;   // [cap_tokens][stats_p][2] partial row statistics of x / tmp (LN folding)
    float* stats_1 = nullptr;
    int stats_p = 0;    // partials per row of stats_1 (written by the attention-output GEMM)
    int stats_px = 0;   // partials per row of stats_x (4 when the fused feed-forward kernel writes it)
    std::atomic<int64_t> launches{0};
    Profiler prof;  // classes: 0 embed+LN, 1 GEMM, 2 attention, 3 LayerNorm, 4 pool
    // All passes of a handle share the activation workspace, and the *_dev / submit calls return without
    // synchronising: every pass records this event at its end and the next pass - on whatever stream - waits
    // for it first, so passes enqueued on different streams are serialised on the GPU as well as on the host.
    cudaEvent_t ev_last = nullptr;
    bool ev_recorded = false;
    // tensor maps of the activation buffers for a given token count (they change only with T or a re-allocation)
    struct ActMaps {
        int64_t T = -1;
        CUtensorMap tm_x, tm_tmp, tm_ctx, tm_ffn, tc_qkv, tc_tmp, tc_x, tc_ffn;
    };
    ActMaps act_maps[4];
    int act_next = 0;
};

namespace lsc {
namespace {

constexpr int64_t kChunkTokens = 1 << 18;  // tokens per forward pass of the host-pointer API

template <typename T>
int dev_alloc(lsc_encoder* e, T** p, size_t count) {
    void* q = nullptr;
    LSC_CUDA(cudaMalloc(&q, count * sizeof(T)));
    e->owned.push_back(q);
    *p = static_cast<T*>(q);
    return LSC_OK;
}

int upload_f32(lsc_encoder* e, float** dst, const float* src, size_t count) {
    LSC_TRY(dev_alloc(e, dst, count));
    LSC_CUDA(cudaMemcpy(*dst, src, count * sizeof(float), cudaMemcpyHostToDevice));
    return LSC_OK;
}

int upload_bf16(lsc_encoder* e, __nv_bfloat16** dst, const float* const* parts, const size_t* counts,
                int nparts) {
    size_t total = 0;
    for (int i = 0; i < nparts; ++i) total += counts[i];
    std::vector<__nv_bfloat16> h(total);
    size_t o = 0;
    for (int i = 0; i < nparts; ++i)
        for (size_t j = 0; j < counts[i]; ++j) h[o++] = __float2bfloat16_rn(parts[i][j]);
    LSC_TRY(dev_alloc(e, dst, total));
    LSC_CUDA(cudaMemcpy(*dst, h.data(), total * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
    return LSC_OK;
}

// W' = W diag(gamma) as bf16 [N,K]; c_n = sum_k bf16(W'_nk); d_n = bias_n + sum_k beta_k W_nk.
int upload_folded(lsc_encoder* e, const float* W, const float* bias, const float* gamma, const float* beta,
                  size_t N, size_t K, __nv_bfloat16** d_w, float** d_c, float** d_d) {
    std::vector<__nv_bfloat16> h(N * K);
    std::vector<float> c(N), d(N);
    for (size_t n = 0; n < N; ++n) {
        double cs = 0.0, ds = bias[n];
        for (size_t k = 0; k < K; ++k) {
            const float w = W[n * K + k];
            const __nv_bfloat16 q = __float2bfloat16_rn(w * gamma[k]);
            h[n * K + k] = q;
            cs += static_cast<double>(__bfloat162float(q));
            ds += static_cast<double>(beta[k]) * static_cast<double>(w);
        }
        c[n] = static_cast<float>(cs);
        d[n] = static_cast<float>(ds);
    }
    LSC_TRY(dev_alloc(e, d_w, N * K));
    LSC_CUDA(cudaMemcpy(*d_w, h.data(), N * K * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
    LSC_TRY(upload_f32(e, d_c, c.data(), N));
    LSC_TRY(upload_f32(e, d_d, d.data(), N));
    return LSC_OK;
}

int check_config(const lsc_encoder_config* c) {
    if (c == nullptr) return fail(LSC_E_INVALID, "encoder config is null");
    if (c->layers <= 0 || c->hidden <= 0 || c->heads <= 0 || c->ffn <= 0 || c->vocab <= 0 ||
        c->max_position <= 0 || c->type_vocab < 0)
        return fail(LSC_E_INVALID, "encoder config: all sizes must be positive");
    if (c->arch != 0 && c->arch != 1) return fail(LSC_E_UNSUPPORTED, "encoder config: arch must be 0 (BERT) or 1 (MPNet)");
    if (c->arch == 0 && c->type_vocab <= 0) return fail(LSC_E_INVALID, "encoder config: a BERT model has token-type embeddings");
    if (c->arch == 1) {
        if (c->type_vocab != 0) return fail(LSC_E_INVALID, "encoder config: an MPNet model has no token-type embeddings (type_vocab 0)");
        if (c->max_position <= 2) return fail(LSC_E_INVALID, "encoder config: MPNet position ids start at 2");
        if (c->hidden / c->heads != 64) return fail(LSC_E_UNSUPPORTED, "encoder config: MPNet needs a head dimension of 64");
        if (c->rel_buckets < 4 || c->rel_buckets > 256 || c->rel_buckets % 4 != 0 || c->rel_max_distance > 128 ||
            c->rel_max_distance <= c->rel_buckets / 4)
            return fail(LSC_E_UNSUPPORTED, "encoder config: relative-position buckets must be a multiple of 4 and saturate within 128 positions");
        if (c->gemm_impl != 0) return fail(LSC_E_UNSUPPORTED, "encoder config: MPNet runs on the product path only");
    }
    if (c->hidden % c->heads != 0) return fail(LSC_E_INVALID, "encoder config: hidden % heads != 0");
    const int d = c->hidden / c->heads;
    if (d != 32 && d != 64) return fail(LSC_E_UNSUPPORTED, "encoder config: head dimension must be 32 or 64");
    const int H = c->hidden;
    if (!(H == 128 || H == 256 || H == 384 || H == 512 || H == 768 || H == 1024))
        return fail(LSC_E_UNSUPPORTED, "encoder config: hidden must be one of 128,256,384,512,768,1024");
    if (c->ffn % 128 != 0) return fail(LSC_E_UNSUPPORTED, "encoder config: ffn must be a multiple of 128");
    if (c->gemm_impl != 0 && c->gemm_impl != 1) return fail(LSC_E_INVALID, "encoder config: gemm_impl must be 0 or 1");
    return LSC_OK;
}

size_t weights_count(const lsc_encoder_config* c) {
    const size_t H = c->hidden, F = c->ffn;
    size_t n = (static_cast<size_t>(c->vocab) + c->max_position + c->type_vocab) * H + 2 * H;
    if (c->arch == 1) n += static_cast<size_t>(c->rel_buckets) * c->heads;
    const size_t per_layer = 3 * (H * H + H) + (H * H + H) + 2 * H + (F * H + F) + (H * F + H) + 2 * H;
    return n + per_layer * c->layers;
}

int ensure_workspace(lsc_encoder* e, int64_t T) {
    if (T <= e->cap_tokens) return LSC_OK;
    const int64_t cap = std::max<int64_t>(T, 4096);
    const size_t H = e->cfg.hidden, F = e->cfg.ffn;
    LSC_CUDA(cudaDeviceSynchronize());  // earlier passes (on any stream) may still use the old buffers
    __nv_bfloat16** bufs[6] = {&e->x, &e->x1, &e->tmp, &e->qkv, &e->ctx, &e->ffn};
    for (auto b : bufs) {
        if (*b) cudaFree(*b);
        *b = nullptr;
    }
    if (e->stats_x) cudaFree(e->stats_x);
    if (e->stats_1) cudaFree(e->stats_1);
    e->stats_x = e->stats_1 = nullptr;
    e->cap_tokens = 0;
    for (auto& m : e->act_maps) m.T = -1;  // they describe the buffers just freed
    const size_t sizes[6] = {H, H, H, 3 * H, H, F};
    for (int i = 0; i < 6; ++i) {
        void* p = nullptr;
        LSC_CUDA(cudaMalloc(&p, static_cast<size_t>(cap) * sizes[i] * sizeof(__nv_bfloat16)));
        *bufs[i] = static_cast<__nv_bfloat16*>(p);
    }
    if (e->stats_p > 0) {
        const size_t n = static_cast<size_t>(cap) * std::max(e->stats_p, e->stats_px) * 2 * sizeof(float);
        LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->stats_x), n));
        LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->stats_1), n));
    }
    e->cap_tokens = cap;
    return LSC_OK;
}

int ensure_staging(lsc_encoder* e, int64_t T, int64_t n_seq) {
    if (T > e->cap_stage_tokens) {
        if (e->d_ids) cudaFree(e->d_ids);
        e->d_ids = nullptr;
        e->cap_stage_tokens = 0;
        const int64_t cap = std::max<int64_t>(T, 4096);
        LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->d_ids), cap * sizeof(int32_t)));
        e->cap_stage_tokens = cap;
    }
    if (n_seq > e->cap_stage_seqs) {
        if (e->d_cu) cudaFree(e->d_cu);
        if (e->d_out) cudaFree(e->d_out);
        e->d_cu = nullptr;
        e->d_out = nullptr;
        e->cap_stage_seqs = 0;
        const int64_t cap = std::max<int64_t>(n_seq, 64);
        LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->d_cu), (cap + 1) * sizeof(int32_t)));
        LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->d_out), cap * e->cfg.hidden * sizeof(float)));
        e->cap_stage_seqs = cap;
    }
    return LSC_OK;
}

int gemm(lsc_encoder* e, const CUtensorMap& tmA, const __nv_bfloat16* A, const CUtensorMap& tmW,
         const __nv_bfloat16* W, const float* bias, const __nv_bfloat16* residual, const CUtensorMap& tmC,
         __nv_bfloat16* C, int M, int N, int K, int epi, int f16_from_col, const LnArgs& ln, cudaStream_t st) {
    if (e->cfg.gemm_impl == 1) return gemm_ref_launch(A, W, bias, residual, C, M, N, K, epi, f16_from_col, st);
    return gemm_tc_launch(tmA, tmW, tmC, bias, residual, M, N, K, epi, f16_from_col, ln, e->num_sms, st);
}

#define LSC_LAUNCH(cls, expr)         \
    do {                              \
        e->prof.begin(cls, st);       \
        int _n = (expr);              \
        e->prof.end(st);              \
        if (_n < 0) return _n;        \
        nlaunch += _n;                \
    } while (0)

// The forward pass proper; everything is enqueued on `st`, nothing synchronises.
int forward(lsc_encoder* e, const int32_t* d_token_ids, const int32_t* d_cu, int n_seq, int T,
            int max_seqlen, float* d_out, cudaStream_t st) {
    const lsc_encoder_config& c = e->cfg;
    const int H = c.hidden, F = c.ffn;
    int64_t nlaunch = 0;
    if (T > 0) LSC_TRY(ensure_workspace(e, T));
    // Ragged batches of the MiniLM kernel class (d_head 32, s <= 128, some sequence shorter than its tile) run on
    // attention tiles composed of 32-row blocks of sequences from anywhere in the batch (attention_tc_blocks.cu); the
    // tile list is built once per pass and serves every layer.  Debug switches restore the older forms: sequences
    // shorter than 48 tokens on average packed into tiles of neighbours (attention_tc_packed.cu), everything else one
    // length-adaptive tile per sequence in descending length (attention_tc_ragged.cu).
    static const bool no_blocks = getenv("LSC_ATTN_NO_BLOCKS") != nullptr;
    static const bool no_pack = getenv("LSC_ATTN_NO_PACK") != nullptr;
    const bool ragged = T > 0 && n_seq > 1 && attention_wants_order(n_seq, T, H, c.heads, max_seqlen);
    const bool block_tiles = ragged && !no_blocks;
    const bool pack_tiles = !block_tiles && !no_pack && T > 0 && H / c.heads == 32 && max_seqlen <= 128 && n_seq > 1 &&
                            n_seq <= attention_tiles_max_seqs() &&
                            static_cast<int64_t>(n_seq) * 48 > static_cast<int64_t>(T);
    const bool order_seqs = !block_tiles && !pack_tiles && ragged;
    const int64_t tiles_need = block_tiles ? attention_blocks_scratch_ints(n_seq)
                               : pack_tiles ? 2 * static_cast<int64_t>(n_seq) + 2 + T : (order_seqs ? n_seq : 0);
    if (tiles_need > e->cap_tiles) {
        LSC_CUDA(cudaDeviceSynchronize());
        if (e->d_tiles) cudaFree(e->d_tiles);
        e->d_tiles = nullptr;
        e->cap_tiles = 0;
        const int64_t cap = std::max<int64_t>(tiles_need + tiles_need / 4, 1 << 16);
        LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->d_tiles), cap * sizeof(int32_t)));
        e->cap_tiles = cap;
    }
    if (e->ev_recorded) LSC_CUDA(cudaStreamWaitEvent(st, e->ev_last, 0));
    const int32_t* tiles = nullptr;
    const uint32_t* rowinfo = nullptr;
    if (pack_tiles) {
        uint32_t* ri = reinterpret_cast<uint32_t*>(e->d_tiles + 2 * n_seq + 2);
        LSC_LAUNCH(2, attention_tiles_launch(d_cu, n_seq, T, e->d_tiles, ri, st));
        tiles = e->d_tiles;
        rowinfo = ri;
    }
    const int32_t* blocks = nullptr;
    if (block_tiles) {
        LSC_LAUNCH(2, attention_blocks_launch(d_cu, n_seq, e->d_tiles, st));
        blocks = e->d_tiles;
    }
    const int32_t* order = nullptr;
    if (order_seqs) {
        LSC_LAUNCH(2, attention_order_launch(d_cu, n_seq, e->d_tiles, st));
        order = e->d_tiles;
    }
    if (T > 0) {
        if (c.gemm_impl == 0) {
            // ---- product path: tcgen05 GEMMs, every LayerNorm folded into the neighbouring epilogues ----
            // x and tmp hold PRE-LayerNorm rows plus per-row partial statistics; a normalised activation is
            // never written to memory.
            lsc_encoder::ActMaps* am = nullptr;
            for (auto& m : e->act_maps)
                if (m.T == T) am = &m;
            if (am == nullptr) {
                am = &e->act_maps[e->act_next];
                e->act_next = (e->act_next + 1) % 4;
                am->T = -1;
                LSC_TRY(make_tmap_bf16_store32(&am->tc_qkv, e->qkv, T, 3 * H));
                LSC_TRY(make_tmap_bf16_store32(&am->tc_tmp, e->tmp, T, H));
                LSC_TRY(make_tmap_bf16_store32(&am->tc_x, e->x, T, H));
                LSC_TRY(make_tmap_bf16_store32(&am->tc_ffn, e->ffn, T, F));
                LSC_TRY(make_tmap_bf16_sw128(&am->tm_x, e->x, T, H, 128));
                LSC_TRY(make_tmap_bf16_sw128(&am->tm_tmp, e->tmp, T, H, 128));
                LSC_TRY(make_tmap_bf16_sw128(&am->tm_ctx, e->ctx, T, H, 128));
                LSC_TRY(make_tmap_bf16_sw128(&am->tm_ffn, e->ffn, T, F, 128));
                am->T = T;
            }
            const CUtensorMap &tm_x = am->tm_x, &tm_tmp = am->tm_tmp, &tm_ctx = am->tm_ctx, &tm_ffn = am->tm_ffn,
                              &tc_qkv = am->tc_qkv, &tc_tmp = am->tc_tmp, &tc_x = am->tc_x, &tc_ffn = am->tc_ffn;
            const int P = e->stats_p, PX = e->stats_px;
            const float inv_h = 1.0f / static_cast<float>(H);
            LSC_LAUNCH(0, embed_ln_launch(d_token_ids, d_cu, n_seq, T, H, e->max_len, c.vocab, e->d_word, e->d_pos0,
                                          e->d_type, e->d_emb_g, e->d_emb_b, c.ln_eps, e->x, nullptr, e->stats_x, PX, st));
            for (int l = 0; l < c.layers; ++l) {
                const LayerW& w = e->layers[l];
                LnArgs ln_in;  // consumer of x: LN_in folded
                ln_in.c = w.c_qkv; ln_in.stats_in = e->stats_x; ln_in.p_in = PX; ln_in.inv_h = inv_h; ln_in.eps = c.ln_eps;
                LSC_LAUNCH(1, gemm(e, tm_x, e->x, w.tm_qkv, w.w_qkv, w.d_qkv, nullptr, tc_qkv, e->qkv, T, 3 * H, H,
                                   EPI_LNIN_BIAS, 2 * H, ln_in, st));
                LSC_LAUNCH(2, attention_launch(e->qkv, d_cu, n_seq, T, H, c.heads, max_seqlen, e->ctx, st, tiles, rowinfo, order, e->d_rel, blocks));
                LnArgs ln_o;  // tmp = ctx.Wo^T + b + LN_in(x); statistics of tmp -> stats_1
                ln_o.stats_in = e->stats_x; ln_o.p_in = PX; ln_o.stats_out = e->stats_1; ln_o.gamma = w.in_g; ln_o.beta = w.in_b;
                ln_o.inv_h = inv_h; ln_o.eps = c.ln_eps;
                ln_o.has_res_map = 1; ln_o.res_map = tc_x;  // the residual x through TMA
                LSC_LAUNCH(1, gemm(e, tm_ctx, e->ctx, w.tm_o, w.w_o, w.b_o, e->x, tc_tmp, e->tmp, T, H, H, EPI_RES_LN, H, ln_o, st));
                if (ffn_fused_supported(H, F)) {
                    // x = GELU(LN1(tmp).W1^T + b1).W2^T + b2 + LN1(tmp) in one kernel; statistics of x -> stats_x
                    LSC_LAUNCH(5, ffn_fused_launch(tm_tmp, w.tm_1f, w.tm_2f, tc_x, T, H, F, w.c_1, w.d_1, w.b_2, w.ln1_g,
                                                   w.ln1_b, e->stats_1, P, e->stats_x, PX, e->tmp, c.ln_eps, e->num_sms, st));
                    continue;
                }
                LnArgs ln_1;  // consumer of tmp: LN1 folded, then GELU
                ln_1.c = w.c_1; ln_1.stats_in = e->stats_1; ln_1.p_in = P; ln_1.inv_h = inv_h; ln_1.eps = c.ln_eps;
                LSC_LAUNCH(1, gemm(e, tm_tmp, e->tmp, w.tm_1, w.w_1, w.d_1, nullptr, tc_ffn, e->ffn, T, F, H,
                                   EPI_LNIN_GELU, F, ln_1, st));
                LnArgs ln_2;  // x = ffn.W2^T + b + LN1(tmp); statistics of x -> stats_x
                ln_2.stats_in = e->stats_1; ln_2.p_in = P; ln_2.stats_out = e->stats_x; ln_2.gamma = w.ln1_g; ln_2.beta = w.ln1_b;
                ln_2.inv_h = inv_h; ln_2.eps = c.ln_eps;
                ln_2.has_res_map = 1; ln_2.res_map = tc_tmp;
                LSC_LAUNCH(1, gemm(e, tm_ffn, e->ffn, w.tm_2, w.w_2, w.b_2, e->tmp, tc_x, e->x, T, H, F, EPI_RES_LN, H, ln_2, st));
            }
        } else {
            // ---- validation path: CUDA-core GEMMs and explicit LayerNorm kernels (unfused) ----
            CUtensorMap none;
            const LnArgs no_ln{};
            LSC_LAUNCH(0, embed_ln_launch(d_token_ids, d_cu, n_seq, T, H, e->max_len, c.vocab, e->d_word, e->d_pos0,
                                          e->d_type, e->d_emb_g, e->d_emb_b, c.ln_eps, e->x, nullptr, nullptr, 0, st));
            for (int l = 0; l < c.layers; ++l) {
                const LayerW& w = e->layers[l];
                LSC_LAUNCH(1, gemm(e, none, e->x, none, w.w_qkv, w.b_qkv, nullptr, none, e->qkv, T, 3 * H, H, EPI_BIAS, 2 * H, no_ln, st));
                LSC_LAUNCH(2, attention_launch(e->qkv, d_cu, n_seq, T, H, c.heads, max_seqlen, e->ctx, st, tiles, rowinfo, order, e->d_rel, blocks));
                LSC_LAUNCH(1, gemm(e, none, e->ctx, none, w.w_o, w.b_o, e->x, none, e->tmp, T, H, H, EPI_BIAS_RESIDUAL, H, no_ln, st));
                LSC_LAUNCH(3, layernorm_launch(e->tmp, w.ln1_g, w.ln1_b, c.ln_eps, e->x1, T, H, st));
                LSC_LAUNCH(1, gemm(e, none, e->x1, none, w.w_1, w.b_1, nullptr, none, e->ffn, T, F, H, EPI_BIAS_GELU, F, no_ln, st));
                LSC_LAUNCH(1, gemm(e, none, e->ffn, none, w.w_2, w.b_2, e->x1, none, e->tmp, T, H, F, EPI_BIAS_RESIDUAL, H, no_ln, st));
                LSC_LAUNCH(3, layernorm_launch(e->tmp, w.ln2_g, w.ln2_b, c.ln_eps, e->x, T, H, st));
            }
        }
    }
    if (c.gemm_impl == 0 && T > 0) {
        const LayerW& last = e->layers[c.layers - 1];
        LSC_LAUNCH(4, pool_l2_launch(e->x, d_cu, n_seq, H, d_out, e->stats_x, e->stats_px, last.ln2_g, last.ln2_b, c.ln_eps, st));
    } else {
        LSC_LAUNCH(4, pool_l2_launch(e->x, d_cu, n_seq, H, d_out, nullptr, 0, nullptr, nullptr, c.ln_eps, st));
    }
    LSC_CUDA(cudaEventRecord(e->ev_last, st));
    e->ev_recorded = true;
    e->launches += nlaunch;
    return LSC_OK;
}

}  // namespace
}  // namespace lsc

extern "C" {

size_t lsc_encoder_weights_nbytes(const lsc_encoder_config* cfg) {
    if (check_config(cfg) != LSC_OK) return 0;
    return weights_count(cfg) * sizeof(float);
}

void lsc_encoder_destroy(lsc_encoder* e) {
    if (e == nullptr) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->stream_out) cudaStreamSynchronize(e->stream_out);
    for (auto& sl : e->slots) {
        if (sl.d_ids) cudaFree(sl.d_ids);
        if (sl.d_cu) cudaFree(sl.d_cu);
        if (sl.d_out) cudaFree(sl.d_out);
        if (sl.ev_in) cudaEventDestroy(sl.ev_in);
        if (sl.ev_fwd) cudaEventDestroy(sl.ev_fwd);
        if (sl.ev_done) cudaEventDestroy(sl.ev_done);
    }
    if (e->stream_in) cudaStreamDestroy(e->stream_in);
    if (e->stream_out) cudaStreamDestroy(e->stream_out);
    for (void* p : e->owned) cudaFree(p);
    void* ws[] = {e->x, e->x1, e->tmp, e->qkv, e->ctx, e->ffn, e->d_ids, e->d_cu, e->d_out, e->stats_x, e->stats_1, e->d_tiles};
    for (void* p : ws)
        if (p) cudaFree(p);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->ev_last) cudaEventDestroy(e->ev_last);
    delete e;
}

int lsc_encoder_create(const lsc_encoder_config* cfg, const void* weights, size_t nbytes, int device,
                       lsc_encoder** out) {
    if (out == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_create: out is null");
    *out = nullptr;
    LSC_TRY(check_config(cfg));
    if (weights == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_create: weights is null");
    if (nbytes != weights_count(cfg) * sizeof(float))
        return fail(LSC_E_INVALID, "lsc_encoder_create: weight blob size does not match the config");
    int ndev = 0;
    LSC_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(LSC_E_INVALID, "lsc_encoder_create: no such device");
    LSC_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    LSC_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(LSC_E_UNSUPPORTED, "langstream-cuda needs a Blackwell (sm_100) GPU; there is no fallback path");

    lsc_encoder* e = new lsc_encoder();
    e->cfg = *cfg;
    e->device = device;
    e->num_sms = prop.multiProcessorCount;
    int rc = LSC_OK;
    auto body = [&]() -> int {
        LSC_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
        LSC_CUDA(cudaEventCreateWithFlags(&e->ev_last, cudaEventDisableTiming));
        const size_t H = cfg->hidden, F = cfg->ffn;
        const float* p = static_cast<const float*>(weights);
        LSC_TRY(upload_f32(e, &e->d_word, p, static_cast<size_t>(cfg->vocab) * H));
        p += static_cast<size_t>(cfg->vocab) * H;
        LSC_TRY(upload_f32(e, &e->d_pos, p, static_cast<size_t>(cfg->max_position) * H));
        p += static_cast<size_t>(cfg->max_position) * H;
        if (cfg->type_vocab > 0) {
            LSC_TRY(upload_f32(e, &e->d_type, p, static_cast<size_t>(cfg->type_vocab) * H));
            p += static_cast<size_t>(cfg->type_vocab) * H;
        } else {  // no token-type embeddings: the embedding kernel adds a zero row
            const std::vector<float> zeros(H, 0.f);
            LSC_TRY(upload_f32(e, &e->d_type, zeros.data(), H));
        }
        // MPNet numbers positions from padding_idx + 1 = 2 (MPNetEmbeddings: create_position_ids_from_input_ids)
        const int pos0 = cfg->arch == 1 ? 2 : 0;
        e->d_pos0 = e->d_pos + static_cast<size_t>(pos0) * H;
        e->max_len = cfg->max_position - pos0;
        LSC_TRY(upload_f32(e, &e->d_emb_g, p, H));
        p += H;
        LSC_TRY(upload_f32(e, &e->d_emb_b, p, H));
        p += H;
        if (cfg->arch == 1) {
            std::vector<float> tbl;
            attention_rel_table(p, cfg->rel_buckets, cfg->rel_max_distance, cfg->heads, &tbl);
            LSC_TRY(upload_f32(e, &e->d_rel, tbl.data(), tbl.size()));
            p += static_cast<size_t>(cfg->rel_buckets) * cfg->heads;
        }
        e->layers.resize(cfg->layers);
        const float* prev_ln2_g = nullptr;
        const float* prev_ln2_b = nullptr;
        if (cfg->gemm_impl == 0) {
            e->stats_p = gemm_tc_stats_parts(static_cast<int>(H));
            e->stats_px = ffn_fused_supported(static_cast<int>(H), static_cast<int>(F)) ? 4 : e->stats_p;
        }
        for (int l = 0; l < cfg->layers; ++l) {
            LayerW& w = e->layers[l];
            const float* qw = p;            const float* qb = qw + H * H;
            const float* kw = qb + H;       const float* kb = kw + H * H;
            const float* vw = kb + H;       const float* vb = vw + H * H;
            p = vb + H;
            const size_t hh = H * H, fh = F * H;
            // the LayerNorm applied to this layer's input: previous layer's output LayerNorm
            // (layer 0: identity, the embedding kernel normalises)
            std::vector<float> in_g(H, 1.0f), in_b(H, 0.0f);
            if (l > 0) {
                memcpy(in_g.data(), prev_ln2_g, H * sizeof(float));
                memcpy(in_b.data(), prev_ln2_b, H * sizeof(float));
            }
            std::vector<float> wq(3 * hh), bq(3 * H);
            memcpy(wq.data(), qw, hh * sizeof(float));
            memcpy(wq.data() + hh, kw, hh * sizeof(float));
            memcpy(wq.data() + 2 * hh, vw, hh * sizeof(float));
            memcpy(bq.data(), qb, H * sizeof(float));
            memcpy(bq.data() + H, kb, H * sizeof(float));
            memcpy(bq.data() + 2 * H, vb, H * sizeof(float));
            const float* ow = p;             const float* ob = ow + hh;
            const float* l1g = ob + H;       const float* l1b = l1g + H;
            const float* w1 = l1b + H;       const float* b1 = w1 + fh;
            const float* w2 = b1 + F;        const float* b2 = w2 + fh;
            const float* l2g = b2 + H;       const float* l2b = l2g + H;
            p = l2b + H;
            if (cfg->gemm_impl == 0) {
                LSC_TRY(upload_folded(e, wq.data(), bq.data(), in_g.data(), in_b.data(), 3 * H, H, &w.w_qkv, &w.c_qkv, &w.d_qkv));
                LSC_TRY(upload_folded(e, w1, b1, l1g, l1b, F, H, &w.w_1, &w.c_1, &w.d_1));
                LSC_TRY(upload_f32(e, &w.in_g, in_g.data(), H));
                LSC_TRY(upload_f32(e, &w.in_b, in_b.data(), H));
            } else {
                const float* wp[1] = {wq.data()};
                const size_t wc[1] = {3 * hh};
                LSC_TRY(upload_bf16(e, &w.w_qkv, wp, wc, 1));
                LSC_TRY(upload_bf16(e, &w.w_1, &w1, &fh, 1));
            }
            LSC_TRY(upload_f32(e, &w.b_qkv, bq.data(), 3 * H));
            LSC_TRY(upload_bf16(e, &w.w_o, &ow, &hh, 1));
            LSC_TRY(upload_f32(e, &w.b_o, ob, H));
            LSC_TRY(upload_f32(e, &w.ln1_g, l1g, H));
            LSC_TRY(upload_f32(e, &w.ln1_b, l1b, H));
            LSC_TRY(upload_f32(e, &w.b_1, b1, F));
            LSC_TRY(upload_bf16(e, &w.w_2, &w2, &fh, 1));
            LSC_TRY(upload_f32(e, &w.b_2, b2, H));
            LSC_TRY(upload_f32(e, &w.ln2_g, l2g, H));
            LSC_TRY(upload_f32(e, &w.ln2_b, l2b, H));
            prev_ln2_g = l2g;
            prev_ln2_b = l2b;
            if (cfg->gemm_impl == 0) {
                const int bn3 = gemm_tc_w_box_rows(3 * static_cast<int>(H));
                const int bnh = gemm_tc_w_box_rows(static_cast<int>(H));
                const int bnf = gemm_tc_w_box_rows(static_cast<int>(F));
                if (bn3 == 0 || bnh == 0 || bnf == 0)
                    return fail(LSC_E_UNSUPPORTED, "encoder: hidden and ffn must be multiples of 128");
                LSC_TRY(make_tmap_bf16_sw128(&w.tm_qkv, w.w_qkv, 3 * H, H, bn3));
                LSC_TRY(make_tmap_bf16_sw128(&w.tm_o, w.w_o, H, H, bnh));
                LSC_TRY(make_tmap_bf16_sw128(&w.tm_1, w.w_1, F, H, bnf));
                LSC_TRY(make_tmap_bf16_sw128(&w.tm_2, w.w_2, H, F, bnh));
                if (ffn_fused_supported(static_cast<int>(H), static_cast<int>(F))) {
                    LSC_TRY(make_tmap_bf16_sw128(&w.tm_1f, w.w_1, F, H, ffn_fused_w_box_rows()));
                    LSC_TRY(make_tmap_bf16_sw128(&w.tm_2f, w.w_2, H, F, ffn_fused_w_box_rows()));
                }
            }
        }
        return LSC_OK;
    };
    rc = body();
    if (rc != LSC_OK) {
        std::string msg = lsc_last_error();
        lsc_encoder_destroy(e);
        set_error(msg);
        return rc;
    }
    *out = e;
    return LSC_OK;
}

int lsc_encoder_set_profiling(lsc_encoder* e, int enabled) {
    if (e == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_set_profiling: encoder is null");
    std::lock_guard<std::mutex> lock(e->mu);
    e->prof.enabled = enabled != 0;
    return LSC_OK;
}

int lsc_encoder_read_profile(lsc_encoder* e, float* ms_by_class, int64_t* launches_by_class, int32_t nclass) {
    if (e == nullptr || ms_by_class == nullptr || launches_by_class == nullptr || nclass < 1 || nclass > Profiler::kMaxClasses)
        return fail(LSC_E_INVALID, "lsc_encoder_read_profile: bad argument");
    std::lock_guard<std::mutex> lock(e->mu);
    LSC_CUDA(cudaSetDevice(e->device));
    return e->prof.read(ms_by_class, launches_by_class, nclass);
}

int lsc_encoder_dim(const lsc_encoder* e) { return e ? e->cfg.hidden : LSC_E_INVALID; }

int64_t lsc_encoder_launch_count(const lsc_encoder* e) { return e ? e->launches.load() : 0; }

int lsc_encoder_embed_dev(lsc_encoder* e, const int32_t* d_token_ids, const int32_t* d_cu_seqlens,
                          int32_t n_seq, int32_t total_tokens, int32_t max_seqlen, float* d_out,
                          void* stream) {
    if (e == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_embed_dev: encoder is null");
    if (n_seq < 0 || total_tokens < 0 || max_seqlen < 0) return fail(LSC_E_INVALID, "lsc_encoder_embed_dev: negative size");
    if (n_seq == 0) return LSC_OK;
    if (d_cu_seqlens == nullptr || d_out == nullptr || (total_tokens > 0 && d_token_ids == nullptr))
        return fail(LSC_E_INVALID, "lsc_encoder_embed_dev: null pointer");
    if (max_seqlen > e->max_len)
        return fail(LSC_E_TOO_LONG, "Illegal input, maybe the number of tokens is too high");
    std::lock_guard<std::mutex> lock(e->mu);
    LSC_CUDA(cudaSetDevice(e->device));
    return forward(e, d_token_ids, d_cu_seqlens, n_seq, total_tokens, max_seqlen, d_out,
                   static_cast<cudaStream_t>(stream));
}

// argument checks shared by the host-buffer entry points; *total receives the token count
static int validate_batch(const char* who, lsc_encoder* e, const int32_t* token_ids, const int32_t* cu_seqlens,
                          int32_t n_seq, const float* out, int64_t* total_out) {
    if (cu_seqlens == nullptr || out == nullptr) return fail(LSC_E_INVALID, std::string(who) + ": null pointer");
    if (cu_seqlens[0] != 0) return fail(LSC_E_INVALID, std::string(who) + ": cu_seqlens[0] must be 0");
    for (int i = 0; i < n_seq; ++i) {
        const int len = cu_seqlens[i + 1] - cu_seqlens[i];
        if (len < 0) return fail(LSC_E_INVALID, std::string(who) + ": cu_seqlens must be non-decreasing");
        if (len > e->max_len)
            return fail(LSC_E_TOO_LONG, "Illegal input, maybe the number of tokens is too high");
    }
    const int64_t total = cu_seqlens[n_seq];
    if (total > 0 && token_ids == nullptr) return fail(LSC_E_INVALID, std::string(who) + ": token_ids is null");
    for (int64_t t = 0; t < total; ++t)
        if (token_ids[t] < 0 || token_ids[t] >= e->cfg.vocab)
            return fail(LSC_E_INVALID, std::string(who) + ": token id outside the vocabulary");
    *total_out = total;
    return LSC_OK;
}

// The blocking host-buffer pass; the caller holds e->mu and has made the device current.
static int embed_locked(lsc_encoder* e, const int32_t* token_ids, const int32_t* cu_seqlens, int32_t n_seq, float* out) {
    const int H = e->cfg.hidden;
    // Chunk the batch so one pass holds at most kChunkTokens tokens.
    std::vector<int32_t> cu_local;
    int s0 = 0;
    while (s0 < n_seq) {
        int s1 = s0;
        int maxlen = 0;
        while (s1 < n_seq && (s1 == s0 || cu_seqlens[s1 + 1] - cu_seqlens[s0] <= kChunkTokens)) {
            maxlen = std::max(maxlen, cu_seqlens[s1 + 1] - cu_seqlens[s1]);
            ++s1;
        }
        const int ns = s1 - s0;
        const int T = cu_seqlens[s1] - cu_seqlens[s0];
        cu_local.resize(ns + 1);
        for (int i = 0; i <= ns; ++i) cu_local[i] = cu_seqlens[s0 + i] - cu_seqlens[s0];
        LSC_TRY(ensure_staging(e, T, ns));
        if (T > 0)
            LSC_CUDA(cudaMemcpyAsync(e->d_ids, token_ids + cu_seqlens[s0], static_cast<size_t>(T) * sizeof(int32_t),
                                     cudaMemcpyHostToDevice, e->stream));
        LSC_CUDA(cudaMemcpyAsync(e->d_cu, cu_local.data(), static_cast<size_t>(ns + 1) * sizeof(int32_t),
                                 cudaMemcpyHostToDevice, e->stream));
        LSC_TRY(forward(e, e->d_ids, e->d_cu, ns, T, maxlen, e->d_out, e->stream));
        LSC_CUDA(cudaMemcpyAsync(out + static_cast<size_t>(s0) * H, e->d_out, static_cast<size_t>(ns) * H * sizeof(float),
                                 cudaMemcpyDeviceToHost, e->stream));
        LSC_CUDA(cudaStreamSynchronize(e->stream));  // cu_local and staging are reused by the next chunk
        s0 = s1;
    }
    return LSC_OK;
}

int lsc_encoder_embed(lsc_encoder* e, const int32_t* token_ids, const int32_t* cu_seqlens, int32_t n_seq,
                      float* out) {
    if (e == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_embed: encoder is null");
    if (n_seq < 0) return fail(LSC_E_INVALID, "lsc_encoder_embed: negative n_seq");
    if (n_seq == 0) return LSC_OK;
    int64_t total = 0;
    LSC_TRY(validate_batch("lsc_encoder_embed", e, token_ids, cu_seqlens, n_seq, out, &total));
    std::lock_guard<std::mutex> lock(e->mu);
    LSC_CUDA(cudaSetDevice(e->device));
    return embed_locked(e, token_ids, cu_seqlens, n_seq, out);
}

int lsc_encoder_submit(lsc_encoder* e, const int32_t* token_ids, const int32_t* cu_seqlens, int32_t n_seq, float* out,
                       int64_t* ticket) {
    if (e == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_submit: encoder is null");
    if (ticket == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_submit: ticket is null");
    if (n_seq < 0) return fail(LSC_E_INVALID, "lsc_encoder_submit: negative n_seq");
    int64_t total = 0;
    if (n_seq > 0) LSC_TRY(validate_batch("lsc_encoder_submit", e, token_ids, cu_seqlens, n_seq, out, &total));
    std::lock_guard<std::mutex> lock(e->mu);
    LSC_CUDA(cudaSetDevice(e->device));
    const int64_t tk = e->next_ticket;
    lsc_encoder::AsyncSlot& sl = e->slots[tk & 1];
    if (sl.ticket >= 0) {  // two submissions in flight: the older one has to finish before its slot is reused
        if (!sl.sync_done) LSC_CUDA(cudaEventSynchronize(sl.ev_done));
        sl.ticket = -1;
    }
    if (e->stream_in == nullptr) {
        LSC_CUDA(cudaStreamCreateWithFlags(&e->stream_in, cudaStreamNonBlocking));
        LSC_CUDA(cudaStreamCreateWithFlags(&e->stream_out, cudaStreamNonBlocking));
    }
    if (sl.ev_in == nullptr) {
        LSC_CUDA(cudaEventCreateWithFlags(&sl.ev_in, cudaEventDisableTiming));
        LSC_CUDA(cudaEventCreateWithFlags(&sl.ev_fwd, cudaEventDisableTiming));
        LSC_CUDA(cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming));
    }
    sl.sync_done = false;
    if (n_seq == 0 || total > kChunkTokens) {
        // empty, or more than one pass: served synchronously on the compute stream (ordered behind earlier tickets)
        // (the handle stays locked throughout: the ticket and its slot cannot be handed out twice)
        if (n_seq > 0) LSC_TRY(embed_locked(e, token_ids, cu_seqlens, n_seq, out));
        sl.sync_done = true;
    } else {
        const int H = e->cfg.hidden;
        if (total > sl.cap_tokens) {
            if (sl.d_ids) cudaFree(sl.d_ids);
            sl.d_ids = nullptr;
            sl.cap_tokens = 0;
            const int64_t cap = std::max<int64_t>(total, 4096);
            LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&sl.d_ids), cap * sizeof(int32_t)));
            sl.cap_tokens = cap;
        }
        if (n_seq > sl.cap_seqs) {
            if (sl.d_cu) cudaFree(sl.d_cu);
            if (sl.d_out) cudaFree(sl.d_out);
            sl.d_cu = nullptr;
            sl.d_out = nullptr;
            sl.cap_seqs = 0;
            const int64_t cap = std::max<int64_t>(n_seq, 64);
            LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&sl.d_cu), (cap + 1) * sizeof(int32_t)));
            LSC_CUDA(cudaMalloc(reinterpret_cast<void**>(&sl.d_out), cap * H * sizeof(float)));
            sl.cap_seqs = cap;
        }
        int maxlen = 0;
        for (int i = 0; i < n_seq; ++i) maxlen = std::max(maxlen, cu_seqlens[i + 1] - cu_seqlens[i]);
        sl.cu_host.assign(cu_seqlens, cu_seqlens + n_seq + 1);  // stays alive until the copy has run
        // upload on the input stream, kernels on the compute stream (shared workspace: batches are serialised
        // there), download on the output stream
        if (total > 0)
            LSC_CUDA(cudaMemcpyAsync(sl.d_ids, token_ids, static_cast<size_t>(total) * sizeof(int32_t), cudaMemcpyHostToDevice,
                                     e->stream_in));
        LSC_CUDA(cudaMemcpyAsync(sl.d_cu, sl.cu_host.data(), static_cast<size_t>(n_seq + 1) * sizeof(int32_t),
                                 cudaMemcpyHostToDevice, e->stream_in));
        LSC_CUDA(cudaEventRecord(sl.ev_in, e->stream_in));
        LSC_CUDA(cudaStreamWaitEvent(e->stream, sl.ev_in, 0));
        LSC_TRY(forward(e, sl.d_ids, sl.d_cu, n_seq, static_cast<int>(total), maxlen, sl.d_out, e->stream));
        LSC_CUDA(cudaEventRecord(sl.ev_fwd, e->stream));
        LSC_CUDA(cudaStreamWaitEvent(e->stream_out, sl.ev_fwd, 0));
        LSC_CUDA(cudaMemcpyAsync(out, sl.d_out, static_cast<size_t>(n_seq) * H * sizeof(float), cudaMemcpyDeviceToHost,
                                 e->stream_out));
        LSC_CUDA(cudaEventRecord(sl.ev_done, e->stream_out));
    }
    sl.ticket = tk;
    e->next_ticket = tk + 1;
    *ticket = tk;
    return LSC_OK;
}

int lsc_encoder_wait(lsc_encoder* e, int64_t ticket) {
    if (e == nullptr) return fail(LSC_E_INVALID, "lsc_encoder_wait: encoder is null");
    cudaEvent_t ev = nullptr;
    {
        std::lock_guard<std::mutex> lock(e->mu);
        if (ticket < 0 || ticket >= e->next_ticket) return fail(LSC_E_INVALID, "lsc_encoder_wait: unknown ticket");
        lsc_encoder::AsyncSlot& sl = e->slots[ticket & 1];
        if (sl.ticket != ticket) return LSC_OK;  // already completed (its slot has been reused or released)
        if (sl.sync_done) {
            sl.ticket = -1;
            return LSC_OK;
        }
        ev = sl.ev_done;
    }
    LSC_CUDA(cudaSetDevice(e->device));
    LSC_CUDA(cudaEventSynchronize(ev));
    std::lock_guard<std::mutex> lock(e->mu);
    if (e->slots[ticket & 1].ticket == ticket) e->slots[ticket & 1].ticket = -1;
    return LSC_OK;
}

}  // extern "C"
