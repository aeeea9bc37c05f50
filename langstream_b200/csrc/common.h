// Host-side plumbing shared by the translation units: error reporting across the C ABI,
// CUDA call checking, TMA tensor-map encoding (driver entry point fetched at run time, so the
// library has no link-time dependency on libcuda), and the kernel launchers' prototypes.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stddef.h>

#include <string>
#include <vector>

#include "../../include/langstream_cuda.h"

namespace lsc {

// ---- errors ------------------------------------------------------------------------------
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);
int fail_cuda(cudaError_t e, const char* what, const char* file, int line);

#define LSC_CUDA(expr)                                                        \
    do {                                                                      \
        cudaError_t _e = (expr);                                              \
        if (_e != cudaSuccess) return ::lsc::fail_cuda(_e, #expr, __FILE__, __LINE__); \
    } while (0)

#define LSC_TRY(expr)             \
    do {                          \
        int _rc = (expr);         \
        if (_rc != LSC_OK) return _rc; \
    } while (0)

// ---- optional per-kernel-class timing (CUDA events on the launching stream) -------------------
// Used by bench.py to attribute the step time to kernels; off by default (zero cost).
struct Profiler {
    static constexpr int kMaxClasses = 8;
    bool enabled = false;
    struct Rec { cudaEvent_t a, b; int cls; };
    std::vector<Rec> recs;
    std::vector<cudaEvent_t> pool;
    cudaEvent_t get() {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        return e;
    }
    void begin(int cls, cudaStream_t st) {
        if (!enabled) return;
        Rec r{get(), get(), cls};
        cudaEventRecord(r.a, st);
        recs.push_back(r);
    }
    void end(cudaStream_t st) {
        if (!enabled || recs.empty()) return;
        cudaEventRecord(recs.back().b, st);
    }
    // Accumulated milliseconds and launch counts per class since the last read; waits for the events.
    int read(float* ms, int64_t* counts, int nclass) {
        for (int i = 0; i < nclass; ++i) { ms[i] = 0.f; counts[i] = 0; }
        for (auto& r : recs) {
            cudaError_t e = cudaEventSynchronize(r.b);
            if (e != cudaSuccess) return fail_cuda(e, "cudaEventSynchronize", __FILE__, __LINE__);
            float t = 0.f;
            cudaEventElapsedTime(&t, r.a, r.b);
            if (r.cls >= 0 && r.cls < nclass) { ms[r.cls] += t; counts[r.cls] += 1; }
            pool.push_back(r.a);
            pool.push_back(r.b);
        }
        recs.clear();
        return LSC_OK;
    }
    ~Profiler() {
        for (auto& r : recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
        for (auto e : pool) cudaEventDestroy(e);
    }
};

// ---- TMA tensor maps -----------------------------------------------------------------------
// 2-D bf16 row-major tensor [rows, cols] (cols contiguous), box = [box_rows, 64 cols] with
// SWIZZLE_128B: the layout tcgen05 K-major SW128 descriptors expect.  Out-of-bounds elements
// are filled with zeros.
int make_tmap_bf16_sw128(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                         uint32_t box_rows);
// Store map of a bf16 row-major [rows, cols] output: box = [32 rows, 32 cols], SWIZZLE_64B (the GEMM
// epilogue stages 32x32 blocks in shared memory and writes them with TMA; rows past `rows` are clipped).
int make_tmap_bf16_store32(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols);
// Any 16-bit row-major tensor: box [box_rows, box_cols], swizzle 0 / 32 / 64 / 128 bytes.
int make_tmap_bf16_box(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                       uint32_t box_cols, int swizzle_bytes);

// ---- kernel launchers (defined in the .cu files) ---------------------------------------------
enum GemmEpilogue {
    EPI_BIAS = 0,           // C = A.W^T + bias
    EPI_BIAS_GELU = 1,      // C = gelu(A.W^T + bias)
    EPI_BIAS_RESIDUAL = 2,  // C = A.W^T + bias + residual
    EPI_DISCARD = 3,        // debug: main loop only
    EPI_LOADONLY = 4,       // debug: main loop + TMEM reads
    EPI_LNIN_BIAS = 5,      // C = LayerNorm(A).W^T + bias, with the LayerNorm folded into the epilogue
    EPI_LNIN_GELU = 6,      // C = gelu(LayerNorm(A).W^T + bias), same folding
    EPI_RES_LN = 7          // C = A.W^T + bias + LayerNorm(residual); also emits row statistics of C
};

// Operands of the LayerNorm-folding epilogues.  A LayerNorm between two GEMMs is never materialised:
// the producer GEMM (EPI_RES_LN) writes the pre-LayerNorm rows and, per row, partial (sum, sum of
// squares) pairs — one per (column tile, warp half) = 2*N/BN per row; the consumer GEMM
// (EPI_LNIN_*) multiplies the RAW rows by W' = W diag(gamma) and finishes the algebra per element:
//   LN(a).W^T + b = rstd * (a.W'^T) - rstd * mean * c + d,  c_n = sum_k W'_nk,  d_n = b_n + sum_k beta_k W_nk.
struct LnArgs {
    const float* c = nullptr;          // [N]  column sums of the folded weights        (EPI_LNIN_*)
    const float* stats_in = nullptr;   // [M][p_in][2] partials of the tensor being normalised
    int p_in = 0;
    float* stats_out = nullptr;        // [M][2*N/BN][2] partials of the rows written  (EPI_RES_LN)
    const float* gamma = nullptr;      // [N]  LayerNorm weight / bias applied to the residual (EPI_RES_LN)
    const float* beta = nullptr;
    float inv_h = 0.f;                 // 1 / row length of the normalised tensor
    float eps = 0.f;
    // Residual epilogues: a store32-type tensor map of the residual tensor (same geometry as the output).  When set,
    // the epilogue warps fetch their residual blocks with TMA into their staging blocks while the accumulator is still
    // being computed, instead of per-thread global loads on the epilogue's critical path.
    int has_res_map = 0;
    CUtensorMap res_map;
};

// Output columns >= f16_from_col are written as fp16 instead of bf16 (pass N for "none"): the QKV GEMM
// emits its V third in fp16, the operand type of the attention kernel's P·V product.
// tcgen05 GEMM: C[M,N] = epi(A[M,K]·W[N,K]^T + bias).  tmapA: box 128 rows; tmapW: box BN rows,
// where BN = gemm_tc_block_n(N).  Returns number of kernel launches (1) or <0 on error.
int gemm_tc_block_n(int N);
int gemm_tc_stats_parts(int N);  // per-row statistic partials written by an EPI_RES_LN GEMM of N columns
int gemm_tc_w_box_rows(int N);  // rows of the weight map's TMA box (BN, or BN / 2 with CTA pairs)
int gemm_tc_launch(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const CUtensorMap& tmapC,
                   const float* bias, const __nv_bfloat16* residual, int M, int N, int K, int epilogue,
                   int f16_from_col, const LnArgs& ln, int num_sms, cudaStream_t stream);
// Fused feed-forward block (ffn_fused.cu): x_out = GELU(LN(tmp).W1^T + b1).W2^T + b2 + LN(tmp), the [T, F]
// intermediate never leaves the SM.  tmapA: tmp, box 128 rows; tmapW1 / tmapW2: W1' [F, H] and W2 [H, F],
// box 128 rows; tmapC: store map of x_out.  c1 / d1: the LN fold of W1 (see LnArgs).  Returns launches (1).
bool ffn_fused_supported(int H, int F);
int ffn_fused_w_box_rows();  // rows of the W1' / W2 maps' TMA boxes (128, or 64 with CTA pairs)
int ffn_fused_launch(const CUtensorMap& tmapA, const CUtensorMap& tmapW1, const CUtensorMap& tmapW2,
                     const CUtensorMap& tmapC, int M, int H, int F, const float* c1, const float* d1, const float* b2,
                     const float* gamma, const float* beta, const float* stats_in, int p_in, float* stats_out,
                     int p_out, const __nv_bfloat16* residual, float eps, int num_sms, cudaStream_t stream);
// CUDA-core validation GEMM (debug).
int gemm_ref_launch(const __nv_bfloat16* A, const __nv_bfloat16* W, const float* bias,
                    const __nv_bfloat16* residual, __nv_bfloat16* C, int M, int N, int K, int epilogue,
                    int f16_from_col, cudaStream_t stream);

// word + position + token-type embedding gather, then LayerNorm -> bf16 [T,H].
int embed_ln_launch(const int32_t* token_ids, const int32_t* cu_seqlens, int n_seq, int T, int H,
                    int max_position, int vocab, const float* word, const float* pos,
                    const float* type0, const float* gamma, const float* beta, float eps,
                    __nv_bfloat16* out, int32_t* seq_of_token, float* stats_out, int stats_p,
                    cudaStream_t stream);
// LayerNorm over rows of a bf16 [T,H] tensor (input already holds residual + sublayer output).
int layernorm_launch(const __nv_bfloat16* in, const float* gamma, const float* beta, float eps,
                     __nv_bfloat16* out, int T, int H, cudaStream_t stream);
// Packed variable-length multi-head self-attention over qkv [T,3H] -> ctx [T,H].
// qkv: Q and K thirds bf16, V third fp16.  Dispatches to the tcgen05 kernel (attention_tc.cu) for
// d_head 32 and sequences of at most 128 tokens, else to the mma.sync kernel (attention.cu).
// tiles (optional, only used by the d_head 32 / s <= 128 kernel): the packing of short sequences into 128-row
// attention tiles built by attention_tiles_launch for the same cu_seqlens (tiles: 2 * n_seq + 2 ints; rowinfo: one
// word per token row = the key range of the row's own sequence inside its tile).
// order (optional, ragged batches of the same kernel class): the sequences by descending length.
// blocks (optional, ragged batches of that kernel class; takes precedence over tiles / order): the tile list of
// attention_blocks_launch - tiles of 32-row blocks of sequences from anywhere in the batch.
// rel_tbl (optional, d_head 64 only): relative-position bias added to the scores (MPNet), see attention_rel_table.
int attention_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int T, int H, int heads,
                     int max_seqlen, __nv_bfloat16* ctx, cudaStream_t stream, const int32_t* tiles = nullptr,
                     const uint32_t* rowinfo = nullptr, const int32_t* order = nullptr, const float* rel_tbl = nullptr,
                     const int32_t* blocks = nullptr);
// true when attention_launch would take a length-adaptive kernel, which wants `blocks` (attention_blocks_launch) or an
// `order` (attention_order_launch)
bool attention_wants_order(int n_seq, int T, int H, int heads, int max_seqlen);
int attention_tc_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int T, int H, int heads,
                        int max_seqlen, __nv_bfloat16* ctx, int num_sms, cudaStream_t stream);
// the same for batches in which not every sequence fills its tile: every stage sized by the item's length (attention_tc_ragged.cu)
// order (optional): the sequences by descending length, from attention_order_launch (n_seq ints, once per pass)
int attention_tc_ragged_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, const int32_t* order, int n_seq, int T, int H,
                               int heads, int max_seqlen, __nv_bfloat16* ctx, int num_sms, cudaStream_t stream);
int attention_order_launch(const int32_t* cu_seqlens, int n_seq, int32_t* order, cudaStream_t stream);
// the same, third form (attention_tc_blocks.cu): tiles composed of 32-row blocks of sequences from anywhere in the batch;
// attention_blocks_launch builds the tile list once per pass into `tiles` (attention_blocks_scratch_ints(n_seq) ints)
int64_t attention_blocks_scratch_ints(int n_seq);
int attention_blocks_launch(const int32_t* cu_seqlens, int n_seq, int32_t* tiles, cudaStream_t stream);
int attention_tc_blocks_launch(const __nv_bfloat16* qkv, const int32_t* tiles, int n_seq, int T, int H, int heads,
                               int max_seqlen, __nv_bfloat16* ctx, int num_sms, cudaStream_t stream);
int attention_tc_packed_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int T, int H, int heads,
                               int max_seqlen, const int32_t* tiles, const uint32_t* rowinfo, __nv_bfloat16* ctx,
                               int num_sms, cudaStream_t stream);
int attention_tiles_max_seqs();
int attention_tiles_launch(const int32_t* cu_seqlens, int n_seq, int T, int32_t* tiles, uint32_t* rowinfo, cudaStream_t stream);
// tcgen05 attention for d_head 64 and any sequence length (attention_tc64.cu).
bool attention_tc64_supported(int H, int heads);
// rel_tbl (optional): [heads][ATT_REL_W] fp32 relative-position bias from attention_rel_table (MPNet)
constexpr int ATT_REL_HALF = 191;
constexpr int ATT_REL_W = 2 * ATT_REL_HALF + 2;  // one pad entry: rows stay 16-byte multiples
void attention_rel_table(const float* w, int buckets, int max_distance, int heads, std::vector<float>* out);
int attention_tc64_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int T, int H, int heads,
                          __nv_bfloat16* ctx, int num_sms, cudaStream_t stream, const float* rel_tbl = nullptr);
// Masked mean-pool over each sequence's tokens + L2 normalisation -> fp32 [n_seq,H].  With `stats`
// (per-row partial sums, see LnArgs) x holds pre-LayerNorm rows and the final LayerNorm (gamma, beta)
// is applied on the fly.
int pool_l2_launch(const __nv_bfloat16* x, const int32_t* cu_seqlens, int n_seq, int H, float* out,
                   const float* stats, int stats_p, const float* gamma, const float* beta, float eps,
                   cudaStream_t stream);

}  // namespace lsc
