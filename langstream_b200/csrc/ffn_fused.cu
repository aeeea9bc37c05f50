// The feed-forward block of one encoder layer as ONE kernel:
//   x_out = GELU(LN1(tmp) . W1^T + b1) . W2^T + b2 + LN1(tmp)          (+ row statistics of x_out)
// replacing libtorch's intermediate.dense -> gelu -> output.dense -> residual add -> LayerNorm chain in the
// DJL forward pass (reference call site: AI/com/datastax/oss/streaming/ai/embeddings/
// AbstractHuggingFaceEmbeddingService.java:194-196, predictor.predict).  As two GEMM launches the
// [T, F] intermediate (403 MB per layer at T = 131,072, F = 1536) is written to HBM and read back, and both
// launches run at the HBM rate rather than the tensor rate; here it never leaves the SM.
//
// Per 128-row tile, with the FFN dimension walked in chunks of 128 columns:
//   G1(j):  Hacc[128 x 128]  = A[128 x H] . W1'[chunk j]^T           (tcgen05.mma, A resident in smem)
//   epi(j): h_j = GELU(rstd * Hacc - rstd * mean * c1 + d1) -> bf16 -> smem, in the UMMA K-major SW128 layout
//   G2(j):  OUT[128 x H]    += h_j . W2[:, chunk j]^T                (A operand = h_j from smem)
// and after the last chunk OUT + b2 + LN1(tmp) goes out through TMA stores with its row statistics.
// LayerNorm folding is the same as in gemm_tcgen05.cu (W1' = W1 diag(gamma1); c1, d1 precomputed).
//
//   warp 0      TMA producer: the A tile once per tile; W1' / W2 boxes of 128 rows x 64 K through a ring
//   warp 1      TMEM allocation + the single-thread MMA issuer; issue order G1(0), G1(1), G2(0), G1(2), G2(1) ...
//               so the tensor pipe works on G2(j-1) / G1(j+1) while the epilogue warps convert chunk j
//   warps 2-9   epilogue: TMEM -> registers -> GELU -> smem (per chunk) and the final residual/LN epilogue
// TMEM: OUT in columns [0, H), Hacc in [384, 512).  Shared memory: A 16 KB x H/64, h two buffers of 32 KB,
// weight ring 4 x 16 KB.  Supported: H <= 384, H % 64 == 0, F % 128 == 0 (MiniLM: H = 384, F = 1536);
// other shapes keep the two-GEMM path.
#include <cstdlib>

#include "common.h"
#include "epilogue_math.cuh"
#include "ptx.cuh"

namespace lsc {

namespace {

constexpr int FF_BM = 128;
constexpr int FF_BK = 64;
constexpr int FF_CH = 128;   // FFN columns per chunk
constexpr int FF_EPI_WARPS = 8;
constexpr int FF_THREADS = 64 + FF_EPI_WARPS * 32;
constexpr int FF_STAGES = 4;
constexpr int FF_BOX_BYTES = 128 * FF_BK * 2;      // 16 KB: every TMA box is 128 rows x 64 bf16
constexpr int FF_HBUF_BYTES = FF_BM * FF_CH * 2;   // 32 KB: one h chunk = two K slices of 16 KB
constexpr int FF_HACC_COL = 384;                   // TMEM column of the chunk accumulator
constexpr int FF_MAX_H = 384;

// barrier slots (8 bytes each) after the ring's full/empty pairs
enum {
    FB_A_FULL = 2 * FF_STAGES,
    FB_A_EMPTY,
    FB_HACC_FULL,
    FB_HACC_EMPTY,
    FB_SH_FULL0,
    FB_SH_FULL1,
    FB_SH_EMPTY0,
    FB_SH_EMPTY1,
    FB_OUT_FULL,
    FB_OUT_EMPTY,
    FB_COUNT
};

struct FfnArgs {
    const float* c1;         // [F] LN fold of FFN-up: sum_k gamma_k W1[n,k]
    const float* d1;         // [F] folded bias of FFN-up
    const float* b2;         // [H]
    const float* gamma;      // [H] LN1 weight (for the residual)
    const float* beta;       // [H]
    const float* stats_in;   // [T][p_in][2] partial (sum, sum sq) of the tmp rows
    float* stats_out;        // [T][p_out][2] partial statistics of the rows written
    const __nv_bfloat16* residual;  // tmp (pre-LN rows)
    int p_in, p_out;
    float inv_h, eps;
};

__device__ __forceinline__ void sts_u32x4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__global__ void __launch_bounds__(FF_THREADS, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW1,
                 const __grid_constant__ CUtensorMap tmapW2, const __grid_constant__ CUtensorMap tmapC, int M, int H,
                 int F, const FfnArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t sA = smem_u32(smem);                                  // A tile: H/64 slices of 16 KB
    const uint32_t sH = sA + (FF_MAX_H / FF_BK) * FF_BOX_BYTES;          // two h buffers (also the store staging)
    const uint32_t sW = sH + 2 * FF_HBUF_BYTES;                          // weight ring
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (FF_MAX_H / FF_BK) * FF_BOX_BYTES + 2 * FF_HBUF_BYTES +
                                                 FF_STAGES * FF_BOX_BYTES);
    const uint32_t bar0 = smem_u32(bars);
    auto bar = [&](int slot) { return bar0 + 8u * static_cast<uint32_t>(slot); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + FB_COUNT);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapA);
        tma_prefetch_desc(&tmapW1);
        tma_prefetch_desc(&tmapW2);
        tma_prefetch_desc(&tmapC);
        for (int s = 0; s < FF_STAGES; ++s) {
            mbar_init(bar(s), 1);
            mbar_init(bar(FF_STAGES + s), 1);
        }
        mbar_init(bar(FB_A_FULL), 1);
        mbar_init(bar(FB_A_EMPTY), 1);
        mbar_init(bar(FB_HACC_FULL), 1);
        mbar_init(bar(FB_HACC_EMPTY), FF_EPI_WARPS);
        mbar_init(bar(FB_SH_FULL0), FF_EPI_WARPS);
        mbar_init(bar(FB_SH_FULL1), FF_EPI_WARPS);
        mbar_init(bar(FB_SH_EMPTY0), 1);
        mbar_init(bar(FB_SH_EMPTY1), 1);
        mbar_init(bar(FB_OUT_FULL), 1);
        mbar_init(bar(FB_OUT_EMPTY), FF_EPI_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int num_tiles = (M + FF_BM - 1) / FF_BM;
    const int num_kb = H / FF_BK;   // K slices of G1
    const int nb_out = H / 128;     // 128-column blocks of OUT
    const int FC = F / FF_CH;       // chunks

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            uint32_t t = 0;
            auto push = [&](const CUtensorMap* m, int c0, int c1) {
                mbar_wait(bar(FF_STAGES + stage), phase ^ 1);
                mbar_arrive_expect_tx(bar(stage), FF_BOX_BYTES);
                tma_load_2d(sW + stage * FF_BOX_BYTES, m, bar(stage), c0, c1);
                if (++stage == FF_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            };
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
                mbar_wait(bar(FB_A_EMPTY), (t & 1) ^ 1);
                mbar_arrive_expect_tx(bar(FB_A_FULL), num_kb * FF_BOX_BYTES);
                for (int kb = 0; kb < num_kb; ++kb)
                    tma_load_2d(sA + kb * FF_BOX_BYTES, &tmapA, bar(FB_A_FULL), kb * FF_BK, tile * FF_BM);
                for (int j = 0; j <= FC; ++j) {  // same order as the MMA issuer: G1(j), then G2(j-1)
                    if (j < FC)
                        for (int kb = 0; kb < num_kb; ++kb) push(&tmapW1, kb * FF_BK, j * FF_CH);
                    if (j >= 1)
                        for (int ks = 0; ks < FF_CH / FF_BK; ++ks)
                            for (int nb = 0; nb < nb_out; ++nb) push(&tmapW2, (j - 1) * FF_CH + ks * FF_BK, nb * 128);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(FF_BM, 128);
            int stage = 0;
            uint32_t phase = 0;
            uint32_t t = 0, g = 0;  // tiles / chunks done by this CTA (barrier parities)
            auto g2 = [&](int jj, uint32_t gg) {
                const uint32_t b = gg & 1, u = gg >> 1;
                mbar_wait(bar(FB_SH_FULL0 + b), u & 1);
                if (jj == 0) mbar_wait(bar(FB_OUT_EMPTY), (t & 1) ^ 1);  // previous tile's OUT has been read
                tc_fence_after();
                for (int ks = 0; ks < FF_CH / FF_BK; ++ks) {
                    const uint64_t da = umma_desc_sw128(sH + b * FF_HBUF_BYTES + ks * FF_BOX_BYTES);
                    for (int nb = 0; nb < nb_out; ++nb) {
                        mbar_wait(bar(stage), phase);
                        tc_fence_after();
                        const uint64_t db = umma_desc_sw128(sW + stage * FF_BOX_BYTES);
#pragma unroll
                        for (int k = 0; k < FF_BK / 16; ++k)
                            umma_bf16(tmem_base + nb * 128, da + 2 * k, db + 2 * k, idesc, (jj | ks | k) != 0 ? 1u : 0u);
                        umma_commit(bar(FF_STAGES + stage));
                        if (++stage == FF_STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                }
                umma_commit(bar(FB_SH_EMPTY0 + b));  // the h buffer may be rewritten when these MMAs retire
            };
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
                mbar_wait(bar(FB_A_FULL), t & 1);
                tc_fence_after();
                for (int j = 0; j < FC; ++j, ++g) {
                    mbar_wait(bar(FB_HACC_EMPTY), (g & 1) ^ 1);  // chunk accumulator has been read
                    tc_fence_after();
                    for (int kb = 0; kb < num_kb; ++kb) {
                        mbar_wait(bar(stage), phase);
                        tc_fence_after();
                        const uint64_t da = umma_desc_sw128(sA + kb * FF_BOX_BYTES);
                        const uint64_t db = umma_desc_sw128(sW + stage * FF_BOX_BYTES);
#pragma unroll
                        for (int k = 0; k < FF_BK / 16; ++k)
                            umma_bf16(tmem_base + FF_HACC_COL, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                        umma_commit(bar(FF_STAGES + stage));
                        if (++stage == FF_STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                    umma_commit(bar(FB_HACC_FULL));
                    if (j == FC - 1) umma_commit(bar(FB_A_EMPTY));  // last read of the A tile
                    if (j >= 1) g2(j - 1, g - 1);
                }
                g2(FC - 1, g - 1);
                umma_commit(bar(FB_OUT_FULL));
            }
        }
    } else {
        // ===== epilogue warps =====
        // Warp e owns TMEM lane quadrant (warp & 3) = 32 tile rows; thread = row.  Of a 128-column chunk it
        // converts columns [64 * half, 64 * half + 64) = K slice `half` of the h buffer; of OUT it owns the
        // column half [H/2 * half, +H/2) in blocks of 32 columns (TMA stores through 32x32 staging blocks,
        // as in gemm_tcgen05.cu; the staging lives in h buffer 0, which is idle during the final epilogue).
        const int ew = warp - 2;
        const int quad = warp & 3;
        const int half = ew >> 2;
        const int r = quad * 32 + lane;  // row inside the tile
        const uint32_t swz = static_cast<uint32_t>(r & 7);
        const uint32_t stg_addr = sH + ew * 4096;
        const uint32_t sw64 = (static_cast<uint32_t>(lane) >> 1) & 3;  // SWIZZLE_64B of the store staging
        const uint32_t tlane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        const int nch_out = H / 2 / 32;
        uint32_t t = 0, g = 0;
        int buf = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
            const int row0 = tile * FF_BM + quad * 32;
            const int grow = row0 + lane;
            const bool row_ok = grow < M;
            // LayerNorm statistics of this thread's row of tmp (LN1 feeds both the FFN and the residual)
            float rstd = 1.f, nm = 0.f;  // nm = -mean * rstd
            {
                float sum = 0.f, sq = 0.f;
                if (row_ok) {
                    const float2* sp = reinterpret_cast<const float2*>(a.stats_in) + static_cast<size_t>(grow) * a.p_in;
                    for (int p = 0; p < a.p_in; ++p) {
                        const float2 s2 = __ldg(sp + p);
                        sum += s2.x;
                        sq += s2.y;
                    }
                }
                const float mean = sum * a.inv_h;
                const float var = fmaxf(sq * a.inv_h - mean * mean, 0.f);
                rstd = row_ok ? rsqrtf(var + a.eps) : 1.f;
                nm = -mean * rstd;
            }
            uint32_t va[32], vb[32];
            // ---- chunks: Hacc -> GELU -> h buffer ----
            for (int j = 0; j < FC; ++j, ++g) {
                const uint32_t b = g & 1, u = g >> 1;
                mbar_wait(bar(FB_HACC_FULL), g & 1);
                tc_fence_after();
                const uint32_t tcol = tlane + FF_HACC_COL + half * 64;
                tmem_ld_32x32(tcol, va);
                tmem_ld_32x32(tcol + 32, vb);
                mbar_wait(bar(FB_SH_EMPTY0 + b), (u & 1) ^ 1);  // G2 of two chunks ago has read this buffer
                const uint32_t hrow = sH + b * FF_HBUF_BYTES + half * FF_BOX_BYTES + static_cast<uint32_t>(r) * 128;
                const int col0 = j * FF_CH + half * 64;
                tmem_ld_wait();
                // both loads have landed: the accumulator can be overwritten by G1 of the next chunk
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(FB_HACC_EMPTY));
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t(&v)[32] = c ? vb : va;
                    const float4* cp = reinterpret_cast<const float4*>(a.c1 + col0 + c * 32);
                    const float4* dp = reinterpret_cast<const float4*>(a.d1 + col0 + c * 32);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 c0 = __ldg(cp + 2 * i), c1 = __ldg(cp + 2 * i + 1);
                        const float4 d0 = __ldg(dp + 2 * i), d1 = __ldg(dp + 2 * i + 1);
                        const float cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
                        const float dd[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
                        float x[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e)
                            x[e] = fmaf(rstd, __uint_as_float(v[8 * i + e]), fmaf(nm, cc[e], dd[e]));
#pragma unroll
                        for (int e = 0; e < 8; e += 2) gelu2(x[e], x[e + 1]);
                        // K-major SWIZZLE_128B: 16-byte chunk index XOR (row & 7)
                        const uint32_t chunk = static_cast<uint32_t>(c * 4 + i) ^ swz;
                        sts_u32x4(hrow + chunk * 16, pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]),
                                  pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
                    }
                }
                fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's smem reads
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(FB_SH_FULL0 + b));
            }
            // ---- final: OUT + b2 + LN1(tmp) -> x, with row statistics ----
            const int col_base = half * (H / 2);
            const __nv_bfloat16* rrow = a.residual + static_cast<size_t>(row_ok ? grow : 0) * H + col_base;
            float osum = 0.f, osq = 0.f;
            mbar_wait(bar(FB_OUT_FULL), t & 1);
            tc_fence_after();
            const uint32_t tout = tlane + col_base;
            tmem_ld_32x32(tout, va);
            for (int c = 0; c < nch_out; ++c) {
                uint32_t res[16];
                ldg256(rrow + c * 32, res);
                ldg256(rrow + c * 32 + 16, res + 8);
                tmem_ld_wait();
                // ping-pong without run-time register indexing: copy, then prefetch the next block
                uint32_t v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = va[i];
                if (c + 1 < nch_out) {
                    tmem_ld_32x32(tout + (c + 1) * 32, va);
                } else {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(FB_OUT_EMPTY));  // OUT may be overwritten by the next tile
                }
                const int col = col_base + c * 32;
                const float4* bp = reinterpret_cast<const float4*>(a.b2 + col);
                const float4* gp = reinterpret_cast<const float4*>(a.gamma + col);
                const float4* ep = reinterpret_cast<const float4*>(a.beta + col);
                if (lane == 0) tma_store_wait_read<1>();  // the block filled two rounds ago has been read
                __syncwarp();
                const uint32_t blk = stg_addr + buf * 2048 + lane * 64;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 b0 = __ldg(bp + 2 * i), b1 = __ldg(bp + 2 * i + 1);
                    const float4 g0 = __ldg(gp + 2 * i), g1 = __ldg(gp + 2 * i + 1);
                    const float4 e0 = __ldg(ep + 2 * i), e1 = __ldg(ep + 2 * i + 1);
                    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                    const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
                    const float ee[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
                    float x[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float raw = (e & 1) ? bf16_hi(res[4 * i + (e >> 1)]) : bf16_lo(res[4 * i + (e >> 1)]);
                        x[e] = __uint_as_float(v[8 * i + e]) + bb[e] + fmaf(raw, rstd * gg[e], fmaf(nm, gg[e], ee[e]));
                        osum += x[e];
                        osq = fmaf(x[e], x[e], osq);
                    }
                    sts_u32x4(blk + ((static_cast<uint32_t>(i) ^ sw64) * 16), pack_bf16x2(x[0], x[1]),
                              pack_bf16x2(x[2], x[3]), pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&tmapC, stg_addr + buf * 2048, col, row0);
                    tma_store_commit();
                }
                buf ^= 1;
            }
            if (row_ok) {
                float2* so = reinterpret_cast<float2*>(a.stats_out) + static_cast<size_t>(grow) * a.p_out;
                so[half] = make_float2(osum, osq);
                for (int p = half + 2; p < a.p_out; p += 2) so[p] = make_float2(0.f, 0.f);
            }
            // the staging blocks sit in h buffer 0: every warp's stores must have been read before any warp
            // writes the next tile's first chunk there
            if (lane == 0) tma_store_wait_read<0>();
            asm volatile("bar.sync 1, %0;" ::"n"(FF_EPI_WARPS * 32) : "memory");
        }
        if (lane == 0) tma_store_wait_all<0>();  // global writes complete before the kernel ends
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

bool ffn_fused_supported(int H, int F) {
    static const bool off = getenv("LSC_FFN_UNFUSED") != nullptr;  // debug: keep the two-GEMM path
    return !off && H > 0 && H <= FF_MAX_H && H % 128 == 0 && F > 0 && F % FF_CH == 0;
}

int ffn_fused_launch(const CUtensorMap& tmapA, const CUtensorMap& tmapW1, const CUtensorMap& tmapW2,
                     const CUtensorMap& tmapC, int M, int H, int F, const float* c1, const float* d1, const float* b2,
                     const float* gamma, const float* beta, const float* stats_in, int p_in, float* stats_out,
                     int p_out, const __nv_bfloat16* residual, float eps, int num_sms, cudaStream_t stream) {
    if (M <= 0) return 0;
    if (!ffn_fused_supported(H, F)) return fail(LSC_E_UNSUPPORTED, "fused FFN: unsupported hidden / ffn size");
    if (!c1 || !d1 || !b2 || !gamma || !beta || !stats_in || !stats_out || !residual || p_in < 1 || p_out < 2)
        return fail(LSC_E_INVALID, "fused FFN: missing argument");
    constexpr int SMEM = (FF_MAX_H / FF_BK) * FF_BOX_BYTES + 2 * FF_HBUF_BYTES + FF_STAGES * FF_BOX_BYTES + 256 + 1024;
    static bool configured[64] = {};
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(ffn_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    FfnArgs a;
    a.c1 = c1; a.d1 = d1; a.b2 = b2; a.gamma = gamma; a.beta = beta;
    a.stats_in = stats_in; a.stats_out = stats_out; a.residual = residual;
    a.p_in = p_in; a.p_out = p_out; a.inv_h = 1.0f / static_cast<float>(H); a.eps = eps;
    const int tiles = (M + FF_BM - 1) / FF_BM;
    const int grid = tiles < num_sms ? tiles : num_sms;
    ffn_fused_kernel<<<grid, FF_THREADS, SMEM, stream>>>(tmapA, tmapW1, tmapW2, tmapC, M, H, F, a);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
