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
//   warps 2-17  epilogue: TMEM -> registers -> GELU -> smem (per chunk) and the final residual/LN epilogue
// TMEM: OUT in columns [0, H), Hacc in [384, 512).  Shared memory: A 16 KB x H/64, h two buffers of 32 KB,
// weight ring 48 KB, per-column parameters 16.5 KB.  Supported: H <= 384, H % 64 == 0, F % 128 == 0 (MiniLM: H = 384, F = 1536);
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
constexpr int FF_EPI_WARPS = 16;
constexpr int FF_THREADS = 64 + FF_EPI_WARPS * 32;
constexpr int FF_RING_BYTES = 48 * 1024;           // weight ring: 3 x 16 KB boxes, or 6 x 8 KB with CTA pairs
constexpr int FF_MAX_H = 384;
constexpr int FF_MAX_F = 1536;
constexpr int FF_PARAM_BYTES = (2 * FF_MAX_F + 3 * FF_MAX_H) * 4;  // c1, d1, b2, gamma, beta staged in shared memory
constexpr int FF_MAX_STAGES = 8;
constexpr int FF_BOX_BYTES = 128 * FF_BK * 2;      // 16 KB: a TMA box of 128 rows x 64 bf16 (A slices, h slices)
constexpr int FF_HBUF_BYTES = FF_BM * FF_CH * 2;   // 32 KB: one h chunk = two K slices of 16 KB
constexpr int FF_HACC_COL = 384;                   // TMEM column of the chunk accumulator

// barrier slots (8 bytes each) after the ring's full/empty pairs
enum {
    FB_A_FULL = 2 * FF_MAX_STAGES,
    FB_A_EMPTY,
    FB_HACC_FULL,
    FB_HACC_EMPTY,
    FB_SH_FULL0,
    FB_SH_FULL1,
    FB_SH_EMPTY0,
    FB_SH_EMPTY1,
    FB_OUT_FULL,
    FB_OUT_EMPTY,
    FB_SH_LOCAL0,  // CTA pairs, peer CTA only: its epilogue warps' arrivals, relayed to the leader's FB_SH_FULL*
    FB_SH_LOCAL1,
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

// PAIR = 2: a cluster of two CTAs works on 256 rows with cta_group::2 MMAs (M = 256): each CTA keeps its own
// 128-row A tile, h buffers, TMEM and epilogue, but stages only half of every weight box (64 of the 128
// rows), which halves the weight traffic through shared memory - the bound of the single-CTA kernel, whose
// N = 128 MMAs move 12 KB of shared memory per 64 tensor cycles.  The leader (rank 0) issues all MMAs;
// completion is multicast to both CTAs' barriers, "consumed" arrivals go to the leader's barriers.
template <int PAIR>
__global__ void __launch_bounds__(FF_THREADS, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW1,
                 const __grid_constant__ CUtensorMap tmapW2, const __grid_constant__ CUtensorMap tmapC, int M, int H,
                 int F, const FfnArgs a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t sA = smem_u32(smem);                                  // A tile: H/64 slices of 16 KB
    const uint32_t sH = sA + (FF_MAX_H / FF_BK) * FF_BOX_BYTES;          // two h buffers (also the store staging)
    const uint32_t sW = sH + 2 * FF_HBUF_BYTES;                          // weight ring
    constexpr int WBOX_ROWS = 128 / PAIR;                                // weight rows staged per CTA and box
    constexpr int WBOX_BYTES = WBOX_ROWS * FF_BK * 2;
    constexpr int FF_STAGES = FF_RING_BYTES / WBOX_BYTES;
    const uint32_t cta_rank = PAIR == 2 ? cluster_ctarank() : 0u;
    // per-column parameters, read by every epilogue thread for every element: kept in shared memory (as
    // global loads they miss the small L1 left beside 228 KB of shared memory and each costs an L2 round trip
    // on the epilogue's critical path)
    const uint32_t sP = sW + FF_RING_BYTES;
    const uint32_t sC1 = sP, sD1 = sP + 4 * FF_MAX_F, sB2 = sP + 8 * FF_MAX_F, sG = sB2 + 4 * FF_MAX_H;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (FF_MAX_H / FF_BK) * FF_BOX_BYTES + 2 * FF_HBUF_BYTES +
                                                 FF_RING_BYTES + FF_PARAM_BYTES);
    const uint32_t bar0 = smem_u32(bars);
    auto bar = [&](int slot) { return bar0 + 8u * static_cast<uint32_t>(slot); };
    // the same barrier in the leader CTA (shared::cluster address); the leader's own for PAIR == 1
    auto lead_bar = [&](int slot) { return PAIR == 2 ? mapa_u32(bar(slot), 0) : bar(slot); };
    auto arrive_lead = [&](int slot) {
        if (PAIR == 2) mbar_arrive_cluster(lead_bar(slot)); else mbar_arrive(bar(slot));
    };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + FB_COUNT);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapA);
        tma_prefetch_desc(&tmapW1);
        tma_prefetch_desc(&tmapW2);
        tma_prefetch_desc(&tmapC);
        for (int s = 0; s < FF_STAGES; ++s) {
            mbar_init(bar(s), PAIR);                 // full: one arrival per CTA's producer (on the leader)
            mbar_init(bar(FF_MAX_STAGES + s), 1);    // empty
        }
        mbar_init(bar(FB_A_FULL), PAIR);
        mbar_init(bar(FB_A_EMPTY), 1 + FF_EPI_WARPS);  // the MMAs' last read + this CTA's epilogue warps (residual)
        mbar_init(bar(FB_HACC_FULL), 1);
        mbar_init(bar(FB_HACC_EMPTY), FF_EPI_WARPS * PAIR);
        mbar_init(bar(FB_SH_FULL0), FF_EPI_WARPS + (PAIR - 1));  // leader's warps + the peer's relay
        mbar_init(bar(FB_SH_FULL1), FF_EPI_WARPS + (PAIR - 1));
        mbar_init(bar(FB_SH_EMPTY0), 1);
        mbar_init(bar(FB_SH_EMPTY1), 1);
        mbar_init(bar(FB_OUT_FULL), 1);
        mbar_init(bar(FB_OUT_EMPTY), FF_EPI_WARPS * PAIR);
        mbar_init(bar(FB_SH_LOCAL0), FF_EPI_WARPS);
        mbar_init(bar(FB_SH_LOCAL1), FF_EPI_WARPS);
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < F; i += FF_THREADS) {
        sts_f32(sC1 + 4 * i, __ldg(a.c1 + i));
        sts_f32(sD1 + 4 * i, __ldg(a.d1 + i));
    }
    for (int i = threadIdx.x; i < H; i += FF_THREADS) {
        sts_f32(sB2 + 4 * i, __ldg(a.b2 + i) + __ldg(a.beta + i));  // b2 + beta: one constant per column
        sts_f32(sG + 4 * i, __ldg(a.gamma + i));
    }
    if (warp == 1) {
        if (PAIR == 2) {
            tmem_alloc_pair(smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
            tmem_relinquish_pair();
        } else {
            tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
            tmem_relinquish();
        }
    }
    if (threadIdx.x == 0) pdl_launch_dependents();
    tc_fence_before();
    if (PAIR == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // tmp / its statistics are complete; the parameter staging above overlapped the producer's tail

    // a tile = 128 * PAIR rows, one per CTA (pair); this CTA's rows are its rank's half
    const int num_tiles = (M + FF_BM * PAIR - 1) / (FF_BM * PAIR);
    const int tile0 = blockIdx.x / PAIR, tile_step = gridDim.x / PAIR;
    const int num_kb = H / FF_BK;   // K slices of G1
    const int nb_out = H / 128;     // 128-column blocks of OUT
    const int FC = F / FF_CH;       // chunks

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            uint32_t t = 0;
            // one weight box: rows [c1, c1 + 128) of the map; a CTA of a pair stages its half of them
            auto push = [&](const CUtensorMap* m, int c0, int c1) {
                mbar_wait(bar(FF_MAX_STAGES + stage), phase ^ 1);
                if (PAIR == 2) {
                    const uint32_t lf = lead_bar(stage);
                    if (cta_rank == 0) mbar_arrive_expect_tx(bar(stage), WBOX_BYTES * 2); else mbar_arrive_cluster(lf);
                    tma_load_2d_pair(sW + stage * WBOX_BYTES, m, lf, c0, c1 + static_cast<int>(cta_rank) * WBOX_ROWS);
                } else {
                    mbar_arrive_expect_tx(bar(stage), WBOX_BYTES);
                    tma_load_2d(sW + stage * WBOX_BYTES, m, bar(stage), c0, c1);
                }
                if (++stage == FF_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            };
            for (int tile = tile0; tile < num_tiles; tile += tile_step, ++t) {
                mbar_wait(bar(FB_A_EMPTY), (t & 1) ^ 1);
                const int arow = (tile * PAIR + static_cast<int>(cta_rank)) * FF_BM;
                if (PAIR == 2) {
                    const uint32_t lf = lead_bar(FB_A_FULL);
                    if (cta_rank == 0)
                        mbar_arrive_expect_tx(bar(FB_A_FULL), num_kb * FF_BOX_BYTES * 2);
                    else
                        mbar_arrive_cluster(lf);
                    for (int kb = 0; kb < num_kb; ++kb) tma_load_2d_pair(sA + kb * FF_BOX_BYTES, &tmapA, lf, kb * FF_BK, arow);
                } else {
                    mbar_arrive_expect_tx(bar(FB_A_FULL), num_kb * FF_BOX_BYTES);
                    for (int kb = 0; kb < num_kb; ++kb)
                        tma_load_2d(sA + kb * FF_BOX_BYTES, &tmapA, bar(FB_A_FULL), kb * FF_BK, arow);
                }
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
        // ===== MMA issuer (the pair's leader only) =====
        if (cta_rank == 0 && elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(FF_BM * PAIR, 128);
            auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t accumulate) {
                if (PAIR == 2) umma_bf16_pair(d, da, db, idesc, accumulate); else umma_bf16(d, da, db, idesc, accumulate);
            };
            auto commit = [&](int slot) {
                if (PAIR == 2) umma_commit_pair(bar(slot)); else umma_commit(bar(slot));
            };
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
                        const uint64_t db = umma_desc_sw128(sW + stage * WBOX_BYTES);
#pragma unroll
                        for (int k = 0; k < FF_BK / 16; ++k)
                            mma(tmem_base + nb * 128, da + 2 * k, db + 2 * k, (jj | ks | k) != 0 ? 1u : 0u);
                        commit(FF_MAX_STAGES + stage);
                        if (++stage == FF_STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                }
                commit(FB_SH_EMPTY0 + b);  // the h buffer may be rewritten when these MMAs retire
            };
            for (int tile = tile0; tile < num_tiles; tile += tile_step, ++t) {
                mbar_wait(bar(FB_A_FULL), t & 1);
                tc_fence_after();
                for (int j = 0; j < FC; ++j, ++g) {
                    mbar_wait(bar(FB_HACC_EMPTY), (g & 1) ^ 1);  // chunk accumulator has been read
                    tc_fence_after();
                    for (int kb = 0; kb < num_kb; ++kb) {
                        mbar_wait(bar(stage), phase);
                        tc_fence_after();
                        const uint64_t da = umma_desc_sw128(sA + kb * FF_BOX_BYTES);
                        const uint64_t db = umma_desc_sw128(sW + stage * WBOX_BYTES);
#pragma unroll
                        for (int k = 0; k < FF_BK / 16; ++k)
                            mma(tmem_base + FF_HACC_COL, da + 2 * k, db + 2 * k, (kb | k) != 0 ? 1u : 0u);
                        commit(FF_MAX_STAGES + stage);
                        if (++stage == FF_STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                    commit(FB_HACC_FULL);
                    if (j == FC - 1) commit(FB_A_EMPTY);  // last read of the A tile
                    if (j >= 1) g2(j - 1, g - 1);
                }
                g2(FC - 1, g - 1);
                commit(FB_OUT_FULL);
            }
        } else if (PAIR == 2 && cta_rank == 1 && elect_one()) {
            // ===== peer CTA: relay "h chunk written" to the leader =====
            // The peer's epilogue warps arrive on a local barrier; one releasing cluster-scope arrive per chunk
            // (a MEMBAR.ALL.GPU, kept off the epilogue warps) publishes their shared-memory writes to the
            // leader's MMA thread.
            uint32_t g = 0;
            for (int tile = tile0; tile < num_tiles; tile += tile_step) {
                for (int j = 0; j < FC; ++j, ++g) {
                    const uint32_t b = g & 1, u = g >> 1;
                    mbar_wait(bar(FB_SH_LOCAL0 + b), u & 1);
                    fence_proxy_async_smem();
                    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(
                                     lead_bar(FB_SH_FULL0 + b))
                                 : "memory");
                }
            }
        }
    } else {
        // ===== epilogue warps =====
        // Sixteen warps (four per scheduler: the epilogue is a chain of dependent TMEM / global loads, fp32
        // math and MUFU, and with two warps per scheduler it ran at an IPC of 0.13 per warp and bounded the
        // kernel).  Warp e owns TMEM lane quadrant (warp & 3) = 32 tile rows, thread = row, and column
        // quarter cq = e / 4: 32 of a chunk's 128 columns (half of K slice cq / 2 of the h buffer), and
        // H / 4 columns of OUT in blocks of 32 (TMA stores through 32x32 staging blocks, as in
        // gemm_tcgen05.cu; the staging occupies the h buffers, which are idle during the final epilogue).
        const int ew = warp - 2;
        const int quad = warp & 3;
        const int cq = ew >> 2;
        const int r = quad * 32 + lane;  // row inside the tile
        const uint32_t swz = static_cast<uint32_t>(r & 7);
        const uint32_t stg_addr = sH + ew * 4096;
        const uint32_t sw64 = (static_cast<uint32_t>(lane) >> 1) & 3;  // SWIZZLE_64B of the store staging
        const uint32_t tlane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        const int nch_out = H / 4 / 32;
        uint32_t t = 0, g = 0;
        int buf = 0;
        for (int tile = tile0; tile < num_tiles; tile += tile_step, ++t) {
            const int row0 = (tile * PAIR + static_cast<int>(cta_rank)) * FF_BM + quad * 32;
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
            uint32_t v[32];
            // ---- chunks: Hacc -> GELU -> h buffer ----
            for (int j = 0; j < FC; ++j, ++g) {
                const uint32_t b = g & 1, u = g >> 1;
                const int col0 = j * FF_CH + cq * 32;
                const uint32_t cp = sC1 + 4 * col0, dp = sD1 + 4 * col0;
                mbar_wait(bar(FB_HACC_FULL), g & 1);
                tc_fence_after();
                tmem_ld_32x32(tlane + FF_HACC_COL + cq * 32, v);
                mbar_wait(bar(FB_SH_EMPTY0 + b), (u & 1) ^ 1);  // G2 of two chunks ago has read this buffer
                const uint32_t hrow =
                    sH + b * FF_HBUF_BYTES + (cq >> 1) * FF_BOX_BYTES + static_cast<uint32_t>(r) * 128;
                tmem_ld_wait();
                // the load has landed: this warp's part of the accumulator may be overwritten by the next G1
                tc_fence_before();
                __syncwarp();
                if (lane == 0) arrive_lead(FB_HACC_EMPTY);
                const uint64_t rstd2 = pk2(rstd, rstd), nm2 = pk2(nm, nm);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 c0 = lds_f32x4(cp + 32 * i), c1 = lds_f32x4(cp + 32 * i + 16);
                    const float4 d0 = lds_f32x4(dp + 32 * i), d1 = lds_f32x4(dp + 32 * i + 16);
                    // h = gelu(rstd * acc + (nm * c1 + d1)), two elements per instruction
                    const uint64_t h0 = gelu2_pk(fma2(rstd2, pk2u(v[8 * i + 0], v[8 * i + 1]), fma2(nm2, pk2(c0.x, c0.y), pk2(d0.x, d0.y))));
                    const uint64_t h1 = gelu2_pk(fma2(rstd2, pk2u(v[8 * i + 2], v[8 * i + 3]), fma2(nm2, pk2(c0.z, c0.w), pk2(d0.z, d0.w))));
                    const uint64_t h2 = gelu2_pk(fma2(rstd2, pk2u(v[8 * i + 4], v[8 * i + 5]), fma2(nm2, pk2(c1.x, c1.y), pk2(d1.x, d1.y))));
                    const uint64_t h3 = gelu2_pk(fma2(rstd2, pk2u(v[8 * i + 6], v[8 * i + 7]), fma2(nm2, pk2(c1.z, c1.w), pk2(d1.z, d1.w))));
                    // K-major SWIZZLE_128B: 16-byte chunk index XOR (row & 7)
                    const uint32_t chunk = static_cast<uint32_t>((cq & 1) * 4 + i) ^ swz;
                    sts_u32x4(hrow + chunk * 16, pack_bf16x2_pk(h0), pack_bf16x2_pk(h1), pack_bf16x2_pk(h2), pack_bf16x2_pk(h3));
                }
                fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's smem reads
                __syncwarp();
                if (lane == 0) mbar_arrive(bar((PAIR == 2 && cta_rank == 1 ? FB_SH_LOCAL0 : FB_SH_FULL0) + b));
            }
            // ---- final: OUT + b2 + LN1(tmp) -> x, with row statistics ----
            const int col_base = cq * (H / 4);
            uint64_t osum2 = pk2(0.f, 0.f), osq2 = pk2(0.f, 0.f);
            // The residual IS this tile's A operand, still in shared memory (K-major SWIZZLE_128B slices): this
            // warp's 32 rows x H/4 columns are copied to registers now - before the accumulator is waited for - and
            // the A tile is released for the next tile's prefetch right after.  (As global loads, one block ahead,
            // the residual cost 45 us of the kernel's 300: three dependent L2 round trips per warp and tile with
            // the tensor pipe idle.)
            constexpr int MAXB = FF_MAX_H / 4 / 32;  // 32-column blocks per warp
            uint32_t res[MAXB][16];
#pragma unroll
            for (int c = 0; c < MAXB; ++c) {
                if (c < nch_out) {
                    const int col = col_base + c * 32;
                    const uint32_t arow = sA + (col >> 6) * FF_BOX_BYTES + static_cast<uint32_t>(r) * 128;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const uint32_t chunk = static_cast<uint32_t>(((col & 63) >> 3) + i) ^ swz;
                        asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                                     : "=r"(res[c][4 * i]), "=r"(res[c][4 * i + 1]), "=r"(res[c][4 * i + 2]), "=r"(res[c][4 * i + 3])
                                     : "r"(arow + chunk * 16)
                                     : "memory");
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(FB_A_EMPTY));
            mbar_wait(bar(FB_OUT_FULL), t & 1);
            tc_fence_after();
            const uint32_t tout = tlane + col_base;
#pragma unroll
            for (int c = 0; c < MAXB; ++c) {
                if (c >= nch_out) break;
                tmem_ld_32x32(tout + c * 32, v);
                const int col = col_base + c * 32;
                const uint32_t bp = sB2 + 4 * col, gp = sG + 4 * col;
                if (lane == 0) tma_store_wait_read<1>();  // the block filled two rounds ago has been read
                tmem_ld_wait();
                if (c + 1 == nch_out) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) arrive_lead(FB_OUT_EMPTY);  // OUT may be overwritten by the next tile
                }
                __syncwarp();
                const uint32_t blk = stg_addr + buf * 2048 + lane * 64;
                const uint64_t rstd2 = pk2(rstd, rstd), nm2 = pk2(nm, nm);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 b0 = lds_f32x4(bp + 32 * i), b1 = lds_f32x4(bp + 32 * i + 16);
                    const float4 g0 = lds_f32x4(gp + 32 * i), g1 = lds_f32x4(gp + 32 * i + 16);
                    const uint64_t bb[4] = {pk2(b0.x, b0.y), pk2(b0.z, b0.w), pk2(b1.x, b1.y), pk2(b1.z, b1.w)};
                    const uint64_t gg[4] = {pk2(g0.x, g0.y), pk2(g0.z, g0.w), pk2(g1.x, g1.y), pk2(g1.z, g1.w)};
                    uint32_t o[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        // x = acc + (b2 + beta) + gamma * (raw * rstd - mean * rstd), two columns at a time
                        const uint32_t w = res[c][4 * i + e];
                        const uint64_t raw = pk2u(w << 16, w & 0xffff0000u);
                        const uint64_t x = fma2(gg[e], fma2(raw, rstd2, nm2), add2(pk2u(v[8 * i + 2 * e], v[8 * i + 2 * e + 1]), bb[e]));
                        osum2 = add2(osum2, x);
                        osq2 = fma2(x, x, osq2);
                        o[e] = pack_bf16x2_pk(x);
                    }
                    sts_u32x4(blk + ((static_cast<uint32_t>(i) ^ sw64) * 16), o[0], o[1], o[2], o[3]);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&tmapC, stg_addr + buf * 2048, col, row0);
                    tma_store_commit();
                }
                buf ^= 1;
            }
            if (row_ok) {  // one partial per column quarter (p_out == 4)
                float s0, s1, q0, q1;
                upk2(osum2, s0, s1);
                upk2(osq2, q0, q1);
                reinterpret_cast<float2*>(a.stats_out)[static_cast<size_t>(grow) * a.p_out + cq] = make_float2(s0 + s1, q0 + q1);
            }
            // the staging blocks sit in the h buffers: every warp's stores must have been read before any warp
            // writes the next tile's chunks there
            if (lane == 0) tma_store_wait_read<0>();
            asm volatile("bar.sync 1, %0;" ::"n"(FF_EPI_WARPS * 32) : "memory");
        }
        if (lane == 0) tma_store_wait_all<0>();  // global writes complete before the kernel ends
    }

    tc_fence_before();
    if (PAIR == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (PAIR == 2) tmem_dealloc_pair(tmem_base, 512); else tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

bool ffn_fused_supported(int H, int F) {
    static const bool off = getenv("LSC_FFN_UNFUSED") != nullptr;  // debug: keep the two-GEMM path
    return !off && H > 0 && H <= FF_MAX_H && H % 128 == 0 && F > 0 && F <= FF_MAX_F && F % FF_CH == 0;
}

// LSC_FFN_PAIR=0 (debug) runs the fused kernel on single CTAs.
static bool ffn_pair_enabled() {
    static const bool on = [] {
        const char* e = getenv("LSC_FFN_PAIR");
        return e == nullptr || atoi(e) != 0;
    }();
    return on;
}

// Rows of the TMA boxes of the W1' / W2 maps: 128, or 64 with CTA pairs (each CTA stages half a box).
int ffn_fused_w_box_rows() { return ffn_pair_enabled() ? 64 : 128; }

template <int PAIR>
static int ffn_launch_impl(const CUtensorMap& tmapA, const CUtensorMap& tmapW1, const CUtensorMap& tmapW2,
                           const CUtensorMap& tmapC, int M, int H, int F, const FfnArgs& a, int num_sms,
                           cudaStream_t stream) {
    constexpr int SMEM = (FF_MAX_H / FF_BK) * FF_BOX_BYTES + 2 * FF_HBUF_BYTES + FF_RING_BYTES + FF_PARAM_BYTES + 256 + 1024;
    static_assert(SMEM <= 227 * 1024, "fused FFN: shared memory budget");
    static bool configured[64] = {};
    auto kern = ffn_fused_kernel<PAIR>;
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const int tiles = (M + FF_BM * PAIR - 1) / (FF_BM * PAIR);
    int grid = tiles * PAIR < num_sms ? tiles * PAIR : num_sms;
    grid -= grid % PAIR;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(grid);
    lc.blockDim = dim3(FF_THREADS);
    lc.dynamicSmemBytes = SMEM;
    lc.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PAIR;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see pdl_wait() in ptx.cuh
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = 2;
    LSC_CUDA(cudaLaunchKernelEx(&lc, kern, tmapA, tmapW1, tmapW2, tmapC, M, H, F, a));
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int ffn_fused_launch(const CUtensorMap& tmapA, const CUtensorMap& tmapW1, const CUtensorMap& tmapW2,
                     const CUtensorMap& tmapC, int M, int H, int F, const float* c1, const float* d1, const float* b2,
                     const float* gamma, const float* beta, const float* stats_in, int p_in, float* stats_out,
                     int p_out, const __nv_bfloat16* residual, float eps, int num_sms, cudaStream_t stream) {
    if (M <= 0) return 0;
    if (!ffn_fused_supported(H, F)) return fail(LSC_E_UNSUPPORTED, "fused FFN: unsupported hidden / ffn size");
    if (!c1 || !d1 || !b2 || !gamma || !beta || !stats_in || !stats_out || !residual || p_in < 1 || p_out != 4)
        return fail(LSC_E_INVALID, "fused FFN: missing argument");
    FfnArgs a;
    a.c1 = c1; a.d1 = d1; a.b2 = b2; a.gamma = gamma; a.beta = beta;
    a.stats_in = stats_in; a.stats_out = stats_out; a.residual = residual;
    a.p_in = p_in; a.p_out = p_out; a.inv_h = 1.0f / static_cast<float>(H); a.eps = eps;
    if (ffn_pair_enabled() && num_sms >= 2)
        return ffn_launch_impl<2>(tmapA, tmapW1, tmapW2, tmapC, M, H, F, a, num_sms, stream);
    return ffn_launch_impl<1>(tmapA, tmapW1, tmapW2, tmapC, M, H, F, a, num_sms, stream);
}

}  // namespace lsc
