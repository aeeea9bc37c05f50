// Dense contraction for the encoder (QKV, attention-output, FFN-up, FFN-down):
//   C[M,N] (bf16) = epilogue(A[M,K] (bf16) · W[N,K]^T (bf16) + bias[N] (fp32))
// replacing libtorch's Linear (+ gelu / + residual add) in the DJL forward pass
// (reference call site: AI/com/datastax/oss/streaming/ai/embeddings/
//  AbstractHuggingFaceEmbeddingService.java:194-196, predictor.predict).
//
// Blackwell-native structure: a persistent, warp-specialised kernel, one CTA per SM.
//   warp 0      TMA producer: 128x64 A tiles and BNx64 W tiles, SWIZZLE_128B, into a smem ring
//   warp 1      allocates TMEM and issues tcgen05.mma (one elected thread), fp32 accumulators in
//               TMEM, two accumulator stages so the epilogue of tile i overlaps the MMAs of tile i+1
//   warps 2-9   epilogue: tcgen05.ld (thread = row) -> bias (+GELU | +residual) in fp32 -> bf16 ->
//               32x32 blocks in swizzled smem -> TMA stores
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA) and TMEM full/empty (MMA <-> epilogue).
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>

#include "common.h"
#include "epilogue_math.cuh"
#include "ptx.cuh"

namespace lsc {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int TMEM_COLS = 512;
constexpr int ACC_STRIDE = 256;  // TMEM columns between the two accumulator stages

// Epilogue warps: BN = 192 (every encoder shape) uses twelve - three per TMEM lane quadrant and scheduler, each
// owning 64 of the tile's columns; the epilogue, not the main loop, bounds these kernels and is a chain of
// dependent loads and fp32 math, so warps are what hides its latency.  Other widths keep eight (BN / 2 columns).
template <int BN>
struct Epi {
    static constexpr int WARPS = BN == 192 ? 12 : 8;
    static constexpr int PARTS = WARPS / 4;              // column parts of a tile (one warp per quadrant and part)
    static constexpr int COLS = BN / PARTS;              // columns per warp
    static constexpr int NCH = COLS / 32;                // 32-column chunks per warp
    static constexpr int THREADS = 64 + WARPS * 32;      // warp 0 TMA, warp 1 MMA, the rest epilogue
    static constexpr int STAGING_BYTES = WARPS * 2 * 2048;   // per warp: two 32x32 bf16 blocks (TMA store sources)
    static constexpr int PARAM_BYTES = WARPS * COLS * 3 * 4;  // per warp: up to three fp32 vectors of its columns
};

// PAIR = 2: CTA pairs (cta_group::2).  A tile of the pair = 256 rows x BN; each CTA stages its own 128 rows
// of A and BN / 2 rows of W, so a stage is smaller and the ring deeper.
template <int BN, int PAIR = 1>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = (BN / PAIR) * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int FIXED_BYTES = Epi<BN>::STAGING_BYTES + Epi<BN>::PARAM_BYTES + 512 + 1024;  // + barriers + alignment
    static constexpr int FIT = (227 * 1024 - FIXED_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = FIT > 6 ? 6 : FIT;
    static constexpr int STAGING_OFFSET = STAGES * STAGE_BYTES;
    static constexpr int PARAM_OFFSET = STAGING_OFFSET + Epi<BN>::STAGING_BYTES;
    static constexpr int BAR_OFFSET = PARAM_OFFSET + Epi<BN>::PARAM_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + FIXED_BYTES;
    static_assert(STAGES >= 3, "gemm: shared memory budget");
};

__device__ __forceinline__ void sts_u32x4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint32_t pack_f16x2_pk(uint64_t v) {
    float a, b;
    upk2(v, a, b);
    return pack_f16x2(a, b);
}

template <int BN, int EPI, int PAIR>
__global__ void __launch_bounds__(Epi<BN>::THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW,
               const __grid_constant__ CUtensorMap tmapC, const float* __restrict__ bias,
               const __nv_bfloat16* __restrict__ residual, int M, int N, int K, int f16_from_col,
               const __grid_constant__ LnArgs ln) {
    using cfg = Cfg<BN, PAIR>;
    constexpr int STAGES = cfg::STAGES;
    const uint32_t cta_rank = PAIR == 2 ? cluster_ctarank() : 0u;
    extern __shared__ uint8_t smem_raw[];
    // aligned by pointer arithmetic on the shared array (an integer round trip would turn every access
    // through it into a generic-address one)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);
    constexpr int EPI_WARPS = Epi<BN>::WARPS;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + cfg::BAR_OFFSET);
    const uint32_t bar_full = smem_u32(bars);
    const uint32_t bar_empty = bar_full + 8 * STAGES;
    const uint32_t bar_tfull = bar_empty + 8 * STAGES;
    const uint32_t bar_tempty = bar_tfull + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + 2 * STAGES + 4);
    const uint32_t bar_res = bar_tempty + 16 + 16;  // [EPI_WARPS][2]: a warp's residual blocks have landed (TMA)

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapA);
        tma_prefetch_desc(&tmapW);
        tma_prefetch_desc(&tmapC);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar_full + 8 * s, PAIR);  // the leader's barrier: one arrival per CTA of the pair
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, EPI_WARPS * PAIR);  // one arrive per epilogue warp (of both CTAs)
        }
        for (int w = 0; w < 2 * EPI_WARPS; ++w) mbar_init(bar_res + 8 * w, 1);
        fence_mbar_init();
    }
    if (warp == 1) {
        if (PAIR == 2) {
            tmem_alloc_pair(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
            tmem_relinquish_pair();
        } else {
            tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
            tmem_relinquish();
        }
    }
    if (threadIdx.x == 0) pdl_launch_dependents();
    tc_fence_before();
    if (PAIR == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // the producer of A / residual / statistics has finished; everything above overlapped its tail

    // tiles are BM * PAIR rows x BN columns, one per CTA (pair); this CTA's rows are its rank's half
    const int m_tiles = (M + BM * PAIR - 1) / (BM * PAIR);
    const int n_tiles = N / BN;
    const int num_tiles = m_tiles * n_tiles;
    const int num_kb = K / BK;
    const int tile0 = blockIdx.x / PAIR, tile_step = gridDim.x / PAIR;

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = tile0; tile < num_tiles; tile += tile_step) {
                const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    const uint32_t sA = smem_base + stage * cfg::STAGE_BYTES;
                    const uint32_t sB = sA + cfg::A_BYTES;
                    if (PAIR == 2) {
                        // both CTAs' bytes are counted on the leader's barrier
                        const uint32_t lead_full = mapa_u32(bar_full + 8 * stage, 0);
                        if (cta_rank == 0)
                            mbar_arrive_expect_tx(bar_full + 8 * stage, cfg::STAGE_BYTES * 2);
                        else
                            mbar_arrive_cluster(lead_full);
                        tma_load_2d_pair(sA, &tmapA, lead_full, kb * BK, (m_blk * 2 + cta_rank) * BM);
                        tma_load_2d_pair(sB, &tmapW, lead_full, kb * BK, n_blk * BN + cta_rank * (BN / 2));
                    } else {
                        mbar_arrive_expect_tx(bar_full + 8 * stage, cfg::STAGE_BYTES);
                        tma_load_2d(sA, &tmapA, bar_full + 8 * stage, kb * BK, m_blk * BM);
                        tma_load_2d(sB, &tmapW, bar_full + 8 * stage, kb * BK, n_blk * BN);
                    }
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (the pair's leader only) =====
        if (cta_rank == 0 && elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(BM * PAIR, BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = tile0; tile < num_tiles; tile += tile_step) {
                mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * ACC_STRIDE;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after();
                    const uint32_t sA = smem_base + stage * cfg::STAGE_BYTES;
                    const uint64_t da = umma_desc_sw128(sA);
                    const uint64_t db = umma_desc_sw128(sA + cfg::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // +32 bytes per K=16 step inside the 128-byte swizzled row (encoded >>4)
                        if (PAIR == 2)
                            umma_bf16_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                        else
                            umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    // frees the smem slot (in both CTAs) when the MMAs retire
                    if (PAIR == 2) umma_commit_pair(bar_empty + 8 * stage); else umma_commit(bar_empty + 8 * stage);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                // accumulator ready for the epilogue (of both CTAs)
                if (PAIR == 2) umma_commit_pair(bar_tfull + 8 * acc); else umma_commit(bar_tfull + 8 * acc);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else {
        // ===== epilogue warps =====
        // Warp e owns TMEM lane quadrant (warp & 3) = 32 tile rows and column part e / 4 of the tile, in
        // chunks of 32 columns; thread = row.  Per chunk: tcgen05.ld -> epilogue math in packed fp32x2 ->
        // bf16 -> a 32x32 block in shared memory (SWIZZLE_64B, conflict-free 16-byte stores) -> one TMA
        // store issued by lane 0.  The stores bypass the LSU, clip the M tail in hardware and cost 2 KB of
        // smem write + 2 KB of smem read per chunk.  Everything with a long latency is issued ahead of its
        // use: the per-column vectors (bias, LN-fold c, gamma, beta) go to the warp's own shared-memory
        // cache at the top of the tile, before the accumulator is waited for; the TMEM load and the
        // residual rows of chunk c + 1 are in flight while chunk c is processed; two staging blocks per
        // warp let a TMA store drain while the next block is filled.
        constexpr int NCH = Epi<BN>::NCH;
        constexpr int COLS = Epi<BN>::COLS;
        constexpr bool kRes = EPI == EPI_BIAS_RESIDUAL || EPI == EPI_RES_LN;
        constexpr bool kLnIn = EPI == EPI_LNIN_BIAS || EPI == EPI_LNIN_GELU;
        constexpr bool kGelu = EPI == EPI_BIAS_GELU || EPI == EPI_LNIN_GELU;
        const int ew = warp - 2;
        const int quad = warp & 3;
        const int part = ew >> 2;
        const uint32_t stg_addr = smem_base + cfg::STAGING_OFFSET + ew * 4096;
        const uint32_t prm = smem_base + cfg::PARAM_OFFSET + ew * (COLS * 3 * 4);  // [3][COLS] fp32
        const uint32_t sw = (static_cast<uint32_t>(lane) >> 1) & 3;  // SWIZZLE_64B: chunk ^= (row >> 1) & 3
        int buf = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        // residual through TMA (see LnArgs::res_map): two chunks per tile = the warp's two staging blocks
        const bool tma_res = kRes && NCH == 2 && ln.has_res_map != 0;
        const uint32_t my_res_bar = bar_res + 16 * ew;
        uint32_t res_phase = 0;
        // TMEM "empty" arrivals go to the leader's barrier (the leader issues the MMAs)
        const uint32_t lead_tempty = PAIR == 2 ? mapa_u32(bar_tempty, 0) : bar_tempty;
        auto release_acc = [&](int which) {
            if (PAIR == 2) mbar_arrive_cluster(lead_tempty + 8 * which); else mbar_arrive(bar_tempty + 8 * which);
        };
        for (int tile = tile0; tile < num_tiles; tile += tile_step) {
            const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
            const int row0 = (m_blk * PAIR + static_cast<int>(cta_rank)) * BM + quad * 32;
            const int grow = row0 + lane;
            const bool row_ok = grow < M;
            const int col_base = n_blk * BN + part * COLS;
            const uint32_t tbase = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * ACC_STRIDE + part * COLS;
            if (EPI == EPI_DISCARD) {  // debug: measure the TMA + MMA main loop alone
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_after();
                __syncwarp();
                tc_fence_before();
                if (lane == 0) release_acc(acc);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
                continue;
            }
            // Tile set-up: every global load of the set-up is issued before any of them is used, so the warp
            // pays one memory latency per tile, not one per dependent step (the epilogue bounds these kernels;
            // a loop "load partial, add" over the row statistics alone cost p_in round trips).
            //  (1) the row's statistic partials, (2) the per-column vectors of this warp's columns,
            //  (3) the residual rows of chunk 0.
            constexpr int MAXP = 12;  // statistic partials per row loaded up front (12 = H 768 at BN 192); more: tail loop
            constexpr bool kStats = kLnIn || EPI == EPI_RES_LN;
            float2 spart[MAXP];
            if (kStats) {
                const float2* sp = reinterpret_cast<const float2*>(ln.stats_in) + static_cast<size_t>(row_ok ? grow : 0) * ln.p_in;
#pragma unroll
                for (int p = 0; p < MAXP; ++p) spart[p] = (row_ok && p < ln.p_in) ? __ldg(sp + p) : make_float2(0.f, 0.f);
            }
            constexpr int PPL = COLS / 32;  // columns per lane in the parameter cache
            float pv0[PPL], pv1[PPL], pv2[PPL];
#pragma unroll
            for (int i = 0; i < PPL; ++i) {
                const int k = lane + 32 * i;
                pv0[i] = __ldg(bias + col_base + k);
                pv1[i] = kLnIn ? __ldg(ln.c + col_base + k) : (EPI == EPI_RES_LN ? __ldg(ln.gamma + col_base + k) : 0.f);
                pv2[i] = EPI == EPI_RES_LN ? __ldg(ln.beta + col_base + k) : 0.f;
            }
            const __nv_bfloat16* rrow = kRes ? residual + static_cast<size_t>(row_ok ? grow : 0) * N + col_base : nullptr;
            uint32_t res[16], res_next[16];
            if (kRes && tma_res) {
                // both residual blocks of this tile -> the warp's staging blocks, in flight while the accumulator is
                // computed; the previous tile's stores must have read the blocks first
                if (lane == 0) {
                    tma_store_wait_read<0>();
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        mbar_arrive_expect_tx(my_res_bar + 8 * c, 2048);
                        tma_load_2d(stg_addr + c * 2048, &ln.res_map, my_res_bar + 8 * c, col_base + c * 32, row0);
                    }
                }
            } else if (kRes) {  // residual of chunk 0
                ldg256(rrow, res_next);
                ldg256(rrow + 16, res_next + 8);
            }
            // per-column vectors -> the warp's shared-memory cache (the previous tile's reads of the cache are
            // over: every lane is past its last chunk)
            __syncwarp();
#pragma unroll
            for (int i = 0; i < PPL; ++i) {
                const int k = lane + 32 * i;
                sts_f32(prm + 4 * k, pv0[i]);
                if (kStats) sts_f32(prm + 4 * (COLS + k), pv1[i]);
                if (EPI == EPI_RES_LN) sts_f32(prm + 4 * (2 * COLS + k), pv2[i]);
            }
            // LayerNorm folding: row statistics of the tensor being normalised (the A operand for
            // kLnIn, the residual for EPI_RES_LN) from its per-row partial (sum, sum of squares).
            float rstd = 1.f, nm = 0.f;  // nm = -mean * rstd
            if (kStats) {
                float sum = 0.f, sq = 0.f;
#pragma unroll
                for (int p = 0; p < MAXP; ++p) {
                    sum += spart[p].x;
                    sq += spart[p].y;
                }
                if (row_ok) {
                    const float2* sp = reinterpret_cast<const float2*>(ln.stats_in) + static_cast<size_t>(grow) * ln.p_in;
                    for (int p = MAXP; p < ln.p_in; ++p) {
                        const float2 t = __ldg(sp + p);
                        sum += t.x;
                        sq += t.y;
                    }
                }
                const float mean = sum * ln.inv_h;
                const float var = fmaxf(sq * ln.inv_h - mean * mean, 0.f);
                rstd = row_ok ? rsqrtf(var + ln.eps) : 1.f;
                nm = -mean * rstd;
            }
            const uint64_t rstd2 = pk2(rstd, rstd), nm2 = pk2(nm, nm);
            uint64_t osum2 = pk2(0.f, 0.f), osq2 = pk2(0.f, 0.f);  // EPI_RES_LN: statistics of the rows written
            __syncwarp();  // the parameter cache is complete
            mbar_wait(bar_tfull + 8 * acc, acc_phase);
            tc_fence_after();
            uint32_t va[32], vb[32];
            tmem_ld_32x32(tbase, va);
            if (EPI == EPI_LOADONLY) {  // debug: main loop + TMEM reads, nothing else
                uint32_t keep = 0;
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    uint32_t(&v)[32] = (c & 1) ? vb : va;
                    uint32_t(&vnext)[32] = (c & 1) ? va : vb;
                    tmem_ld_wait();
                    if (c + 1 < NCH) tmem_ld_32x32(tbase + (c + 1) * 32, vnext);
#pragma unroll
                    for (int i = 0; i < 32; ++i) keep ^= v[i];
                }
                if (keep == 0x12345678u && bias != nullptr) *const_cast<float*>(bias) = 0.f;  // never true in practice
                __syncwarp();
                tc_fence_before();
                if (lane == 0) release_acc(acc);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
                continue;
            }
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                uint32_t(&v)[32] = (c & 1) ? vb : va;
                uint32_t(&vnext)[32] = (c & 1) ? va : vb;
                if (kRes && tma_res) {
                    // this thread's row of the residual block: 64 bytes at the swizzled positions the output will take
                    mbar_wait(my_res_bar + 8 * c, res_phase);
                    const uint32_t rblk = stg_addr + c * 2048 + lane * 64;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                                     : "=r"(res[4 * i]), "=r"(res[4 * i + 1]), "=r"(res[4 * i + 2]), "=r"(res[4 * i + 3])
                                     : "r"(rblk + ((static_cast<uint32_t>(i) ^ sw) * 16))
                                     : "memory");
                } else if (kRes) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) res[i] = res_next[i];
                    if (c + 1 < NCH) {
                        ldg256(rrow + (c + 1) * 32, res_next);
                        ldg256(rrow + (c + 1) * 32 + 16, res_next + 8);
                    }
                }
                tmem_ld_wait();
                if (c + 1 < NCH) {
                    tmem_ld_32x32(tbase + (c + 1) * 32, vnext);
                } else {
                    // every TMEM read of this tile by this warp has been issued and has completed
                    __syncwarp();
                    tc_fence_before();
                    if (lane == 0) release_acc(acc);
                }
                const int col = col_base + c * 32;
                const uint32_t p0 = prm + 4 * (c * 32), p1 = p0 + 4 * COLS, p2 = p1 + 4 * COLS;
                // the block we are about to fill was the source of the TMA store issued two chunks ago
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
                const uint32_t blk = stg_addr + buf * 2048 + lane * 64;
                const bool as_f16 = col >= f16_from_col;  // V third of the QKV output
#pragma unroll
                for (int i = 0; i < 4; ++i) {  // 4 x 16 bytes = 8 columns each
                    const float4 b0 = lds_f32x4(p0 + 32 * i), b1 = lds_f32x4(p0 + 32 * i + 16);
                    const uint64_t bb[4] = {pk2(b0.x, b0.y), pk2(b0.z, b0.w), pk2(b1.x, b1.y), pk2(b1.z, b1.w)};
                    uint64_t q1[4], q2[4];  // second / third vector (LN-fold c | gamma, beta)
                    if (kLnIn || EPI == EPI_RES_LN) {
                        const float4 c0 = lds_f32x4(p1 + 32 * i), c1 = lds_f32x4(p1 + 32 * i + 16);
                        q1[0] = pk2(c0.x, c0.y); q1[1] = pk2(c0.z, c0.w); q1[2] = pk2(c1.x, c1.y); q1[3] = pk2(c1.z, c1.w);
                    }
                    if (EPI == EPI_RES_LN) {
                        const float4 e0 = lds_f32x4(p2 + 32 * i), e1 = lds_f32x4(p2 + 32 * i + 16);
                        q2[0] = pk2(e0.x, e0.y); q2[1] = pk2(e0.z, e0.w); q2[2] = pk2(e1.x, e1.y); q2[3] = pk2(e1.z, e1.w);
                    }
                    uint32_t o[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {  // two columns per instruction
                        const uint64_t a2 = pk2u(v[8 * i + 2 * e], v[8 * i + 2 * e + 1]);
                        uint64_t x;
                        if (kLnIn) {
                            // LN(a) . W^T + b  ==  rstd * (a . W'^T) - rstd * mean * c + d,  W' = W diag(gamma)
                            x = fma2(rstd2, a2, fma2(nm2, q1[e], bb[e]));
                        } else {
                            x = add2(a2, bb[e]);
                        }
                        if (kGelu) x = gelu2_pk(x);
                        if (kRes) {
                            const uint32_t w = res[4 * i + e];
                            const uint64_t raw = pk2u(w << 16, w & 0xffff0000u);
                            if (EPI == EPI_RES_LN) {
                                // + LayerNorm(raw) = gamma * (raw * rstd - mean * rstd) + beta, and the statistics
                                // of what is written (the next LayerNorm is folded into its consumer)
                                x = fma2(q1[e], fma2(raw, rstd2, nm2), add2(x, q2[e]));
                                osum2 = add2(osum2, x);
                                osq2 = fma2(x, x, osq2);
                            } else {
                                x = add2(x, raw);
                            }
                        }
                        o[e] = as_f16 ? pack_f16x2_pk(x) : pack_bf16x2_pk(x);
                    }
                    sts_u32x4(blk + ((static_cast<uint32_t>(i) ^ sw) * 16), o[0], o[1], o[2], o[3]);
                }
                fence_proxy_async_smem();  // make the generic-proxy writes visible to the TMA engine
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&tmapC, stg_addr + buf * 2048, col, row0);
                    tma_store_commit();
                }
                buf ^= 1;
            }
            if (EPI == EPI_RES_LN && row_ok) {
                // partial (sum, sum of squares) of this row over this warp's columns; the consumer adds the
                // PARTS * N / BN partials of a row
                float s0, s1, t0, t1;
                upk2(osum2, s0, s1);
                upk2(osq2, t0, t1);
                float2* so = reinterpret_cast<float2*>(ln.stats_out) +
                             static_cast<size_t>(grow) * (Epi<BN>::PARTS * n_tiles) + (n_blk * Epi<BN>::PARTS + part);
                *so = make_float2(s0 + s1, t0 + t1);
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
            res_phase ^= 1;
        }
        if (lane == 0) tma_store_wait_all<0>();  // global writes complete before the kernel ends
    }

    tc_fence_before();
    if (PAIR == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (PAIR == 2) tmem_dealloc_pair(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

template <int BN, int EPI, int PAIR>
int launch_pair(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const CUtensorMap& tmapC, const float* bias,
                const __nv_bfloat16* residual, int M, int N, int K, int f16_from_col, const LnArgs& ln, int num_sms,
                cudaStream_t stream) {
    using cfg = Cfg<BN, PAIR>;
    static bool configured[64] = {};  // per instantiation and per device
    auto kern = gemm_tc_kernel<BN, EPI, PAIR>;
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      cfg::SMEM_BYTES));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const int m_tiles = (M + BM * PAIR - 1) / (BM * PAIR);
    const int num_tiles = m_tiles * (N / BN);
    int grid = num_tiles * PAIR < num_sms ? num_tiles * PAIR : num_sms;
    grid -= grid % PAIR;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(grid);
    lc.blockDim = dim3(Epi<BN>::THREADS);
    lc.dynamicSmemBytes = cfg::SMEM_BYTES;
    lc.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see pdl_wait() in ptx.cuh
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
    if (PAIR == 2) {  // CTA pairs: clusters of two CTAs, scheduled on the two SMs of a TPC
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = PAIR;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    lc.attrs = attr;
    lc.numAttrs = na;
    LSC_CUDA(cudaLaunchKernelEx(&lc, kern, tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln));
    LSC_CUDA(cudaGetLastError());
    return 1;
}

// CTA pairs pay off when the tile row is long (QKV, FFN): +6 % on the main loop and a deeper ring; the
// narrow N = H GEMM (two tiles per row, epilogue-bound) is faster on single CTAs.  LSC_GEMM_PAIR=0 (debug)
// keeps every GEMM on single CTAs.
bool gemm_use_pair(int N, int bn) {
    static const bool on = [] {
        const char* e = getenv("LSC_GEMM_PAIR");
        return e == nullptr || atoi(e) != 0;
    }();
    return on && bn != 128 && bn > 0 && N / bn >= 4;
}

template <int BN, int EPI>
int launch_inst(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const CUtensorMap& tmapC, const float* bias,
                const __nv_bfloat16* residual, int M, int N, int K, int f16_from_col, const LnArgs& ln, int num_sms,
                cudaStream_t stream) {
    // the pair kernel wants W boxes of BN / 2 rows: see gemm_tc_w_box_rows()
    if (BN != 128 && gemm_use_pair(N, BN) && num_sms >= 2)
        return launch_pair<BN, EPI, 2>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
    return launch_pair<BN, EPI, 1>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
}

template <int BN>
int launch_bn(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const CUtensorMap& tmapC, const float* bias,
              const __nv_bfloat16* residual, int M, int N, int K, int epilogue, int f16_from_col, const LnArgs& ln,
              int num_sms, cudaStream_t stream) {
    switch (epilogue) {
        case EPI_BIAS:
            return launch_inst<BN, EPI_BIAS>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        case EPI_BIAS_GELU:
            return launch_inst<BN, EPI_BIAS_GELU>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        case EPI_BIAS_RESIDUAL:
            if (residual == nullptr) return fail(LSC_E_INVALID, "gemm: residual epilogue without residual");
            return launch_inst<BN, EPI_BIAS_RESIDUAL>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        case EPI_LNIN_BIAS:
            if (ln.c == nullptr || ln.stats_in == nullptr) return fail(LSC_E_INVALID, "gemm: LN-folded epilogue without statistics");
            return launch_inst<BN, EPI_LNIN_BIAS>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        case EPI_LNIN_GELU:
            if (ln.c == nullptr || ln.stats_in == nullptr) return fail(LSC_E_INVALID, "gemm: LN-folded epilogue without statistics");
            return launch_inst<BN, EPI_LNIN_GELU>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        case EPI_RES_LN:
            if (residual == nullptr || ln.stats_in == nullptr || ln.stats_out == nullptr || ln.gamma == nullptr || ln.beta == nullptr)
                return fail(LSC_E_INVALID, "gemm: residual-LN epilogue needs residual, statistics, gamma and beta");
            return launch_inst<BN, EPI_RES_LN>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        case EPI_DISCARD:
            return launch_inst<BN, EPI_DISCARD>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        case EPI_LOADONLY:
            return launch_inst<BN, EPI_LOADONLY>(tmapA, tmapW, tmapC, bias, residual, M, N, K, f16_from_col, ln, num_sms, stream);
        default:
            return fail(LSC_E_INVALID, "gemm: unknown epilogue");
    }
}

}  // namespace

int gemm_tc_block_n(int N) {
    // LSC_GEMM_BN (debug): force a tile width when it divides N
    static const int forced = [] {
        const char* e = getenv("LSC_GEMM_BN");
        return e ? atoi(e) : 0;
    }();
    if (forced > 0 && N % forced == 0 && (forced == 128 || forced == 192 || forced == 256)) return forced;
    if (N % 192 == 0) return 192;
    if (N % 256 == 0) return 256;
    if (N % 128 == 0) return 128;
    return 0;
}

// Rows of the TMA box the weight map of an N-column GEMM must have: BN on single CTAs, BN / 2 for CTA pairs
// (each CTA of a pair stages half of the tile's weight rows).
int gemm_tc_w_box_rows(int N) {
    const int bn = gemm_tc_block_n(N);
    if (bn == 0) return 0;
    return gemm_use_pair(N, bn) ? bn / 2 : bn;
}

// Number of per-row statistic partials an EPI_RES_LN GEMM with N output columns writes.
int gemm_tc_stats_parts(int N) {
    switch (gemm_tc_block_n(N)) {
        case 192: return Epi<192>::PARTS * (N / 192);
        case 256: return Epi<256>::PARTS * (N / 256);
        case 128: return Epi<128>::PARTS * (N / 128);
        default: return 0;
    }
}

int gemm_tc_launch(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const CUtensorMap& tmapC,
                   const float* bias, const __nv_bfloat16* residual, int M, int N, int K, int epilogue,
                   int f16_from_col, const LnArgs& ln, int num_sms, cudaStream_t stream) {
    if (M <= 0) return 0;
    if (K % BK != 0 || K <= 0) return fail(LSC_E_UNSUPPORTED, "gemm: K must be a positive multiple of 64");
    if (bias == nullptr) return fail(LSC_E_INVALID, "gemm: bias is required");
    switch (gemm_tc_block_n(N)) {
        case 192:
            return launch_bn<192>(tmapA, tmapW, tmapC, bias, residual, M, N, K, epilogue, f16_from_col, ln, num_sms, stream);
        case 256:
            return launch_bn<256>(tmapA, tmapW, tmapC, bias, residual, M, N, K, epilogue, f16_from_col, ln, num_sms, stream);
        case 128:
            return launch_bn<128>(tmapA, tmapW, tmapC, bias, residual, M, N, K, epilogue, f16_from_col, ln, num_sms, stream);
        default:
            return fail(LSC_E_UNSUPPORTED, "gemm: N must be a multiple of 128");
    }
}

// ---------------------------------------------------------------------------------------------
// CUDA-core validation GEMM (debug only; the tests use it to cross-check the tcgen05 kernel).
// ---------------------------------------------------------------------------------------------
namespace {

template <int EPI>
__global__ void __launch_bounds__(256)
gemm_ref_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ W,
                const float* __restrict__ bias, const __nv_bfloat16* __restrict__ residual,
                __nv_bfloat16* __restrict__ C, int M, int N, int K, int f16_from_col) {
    __shared__ float sA[16][64 + 1];
    __shared__ float sW[16][64 + 1];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += 16) {
        for (int i = threadIdx.x; i < 64 * 16; i += 256) {
            const int r = i >> 4, kk = i & 15;
            const int gm = m0 + r, gn = n0 + r, gk = k0 + kk;
            sA[kk][r] = (gm < M && gk < K) ? __bfloat162float(A[static_cast<size_t>(gm) * K + gk]) : 0.f;
            sW[kk][r] = (gn < N && gk < K) ? __bfloat162float(W[static_cast<size_t>(gn) * K + gk]) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float a[4], w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                a[i] = sA[kk][ty * 4 + i];
                w[i] = sW[kk][tx * 4 + i];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
        }
        __syncthreads();
    }
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            float x = acc[i][j] + bias[gn];
            if (EPI == EPI_BIAS_GELU) x = gelu_erf(x);
            if (EPI == EPI_BIAS_RESIDUAL) x += __bfloat162float(residual[static_cast<size_t>(gm) * N + gn]);
            if (gn >= f16_from_col)
                reinterpret_cast<__half*>(C)[static_cast<size_t>(gm) * N + gn] = __float2half_rn(x);
            else
                C[static_cast<size_t>(gm) * N + gn] = __float2bfloat16_rn(x);
        }
    }
}

}  // namespace

int gemm_ref_launch(const __nv_bfloat16* A, const __nv_bfloat16* W, const float* bias,
                    const __nv_bfloat16* residual, __nv_bfloat16* C, int M, int N, int K, int epilogue,
                    int f16_from_col, cudaStream_t stream) {
    if (M <= 0) return 0;
    dim3 grid((N + 63) / 64, (M + 63) / 64);
    switch (epilogue) {
        case EPI_BIAS:
            gemm_ref_kernel<EPI_BIAS><<<grid, 256, 0, stream>>>(A, W, bias, residual, C, M, N, K, f16_from_col);
            break;
        case EPI_BIAS_GELU:
            gemm_ref_kernel<EPI_BIAS_GELU><<<grid, 256, 0, stream>>>(A, W, bias, residual, C, M, N, K, f16_from_col);
            break;
        case EPI_BIAS_RESIDUAL:
            if (residual == nullptr) return fail(LSC_E_INVALID, "gemm: residual epilogue without residual");
            gemm_ref_kernel<EPI_BIAS_RESIDUAL><<<grid, 256, 0, stream>>>(A, W, bias, residual, C, M, N, K, f16_from_col);
            break;
        default:
            return fail(LSC_E_INVALID, "gemm: unknown epilogue");
    }
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
