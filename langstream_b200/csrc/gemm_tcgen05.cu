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
//   warps 2-9   epilogue: tcgen05.ld -> fp32 staging in swizzled smem -> bias (+GELU | +residual) in a
//               column-per-lane layout -> bf16 -> coalesced 16-byte global stores
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA) and TMEM full/empty (MMA <-> epilogue).
#include "common.h"
#include "ptx.cuh"

namespace lsc {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int EPI_WARPS = 8;
constexpr int GEMM_THREADS = 64 + EPI_WARPS * 32;  // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue
constexpr int TMEM_COLS = 512;
constexpr int ACC_STRIDE = 256;  // TMEM columns between the two accumulator stages
constexpr int STAGING_BYTES = EPI_WARPS * 32 * 128;  // per warp: 32 rows x 32 fp32 columns

template <int BN>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (BN <= 128) ? 5 : 4;
    static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + STAGING_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFFSET + 256 + 1024;  // + barriers + alignment slack
};

// ---- packed fp32x2 arithmetic (sm_100 FFMA2 / FMUL2): halves the issue slots of the epilogue ----
__device__ __forceinline__ uint64_t pk2(float a, float b) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& a, float& b) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// Exact-erf GELU, x * Phi(x), evaluated without erff():
//   gelu(x) = max(x,0) - |x| * tail(|x|),  tail(a) = Phi(-a) = 0.5 - a * P(a^2)  on a <= 4,
// P = 9-term minimax fit (Lawson, weighted for the absolute error of gelu), tail clamped to >= 0 and
// shifted so that tail(4) = 0.  Max abs error 1.3e-4 over all x (fp32 evaluation), i.e. an order of
// magnitude below the bf16 rounding of the result; two elements per FFMA2.
__device__ __forceinline__ void gelu2(float& x0, float& x1) {
    const float a0 = fabsf(x0), a1 = fabsf(x1);
    const float c0 = fminf(a0, 4.0f), c1 = fminf(a1, 4.0f);
    const uint64_t ac = pk2(c0, c1);
    const uint64_t u = mul2(ac, ac);
    uint64_t p = pk2(6.740397823e-11f, 6.740397823e-11f);
    p = fma2(p, u, pk2(-6.066908309e-09f, -6.066908309e-09f));
    p = fma2(p, u, pk2(2.440845606e-07f, 2.440845606e-07f));
    p = fma2(p, u, pk2(-5.860600595e-06f, -5.860600595e-06f));
    p = fma2(p, u, pk2(9.494025289e-05f, 9.494025289e-05f));
    p = fma2(p, u, pk2(-1.112856087e-03f, -1.112856087e-03f));
    p = fma2(p, u, pk2(9.815800935e-03f, 9.815800935e-03f));
    p = fma2(p, u, pk2(-6.632363051e-02f, -6.632363051e-02f));
    p = fma2(p, u, pk2(3.988810182e-01f, 3.988810182e-01f));
    // t = half - ac * p
    const uint64_t t = fma2(pk2(-c0, -c1), p, pk2(4.999701977e-01f, 4.999701977e-01f));
    float t0, t1;
    upk2(t, t0, t1);
    t0 = fmaxf(t0, 0.f);
    t1 = fmaxf(t1, 0.f);
    x0 = fmaf(-a0, t0, fmaxf(x0, 0.f));
    x1 = fmaf(-a1, t1, fmaxf(x1, 0.f));
}

__device__ __forceinline__ float gelu_erf(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
}

template <int BN, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW,
               const float* __restrict__ bias, const __nv_bfloat16* __restrict__ residual,
               __nv_bfloat16* __restrict__ C, int M, int N, int K) {
    using cfg = Cfg<BN>;
    constexpr int STAGES = cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    const uint32_t smem_base = smem_u32(smem);
    uint8_t* staging = smem + STAGES * cfg::STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + cfg::BAR_OFFSET);
    const uint32_t bar_full = smem_u32(bars);
    const uint32_t bar_empty = bar_full + 8 * STAGES;
    const uint32_t bar_tfull = bar_empty + 8 * STAGES;
    const uint32_t bar_tempty = bar_tfull + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapA);
        tma_prefetch_desc(&tmapW);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, EPI_WARPS);  // one arrive per epilogue warp
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int m_tiles = (M + BM - 1) / BM;
    const int n_tiles = N / BN;
    const int num_tiles = m_tiles * n_tiles;
    const int num_kb = K / BK;

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    const uint32_t sA = smem_base + stage * cfg::STAGE_BYTES;
                    const uint32_t sB = sA + cfg::A_BYTES;
                    mbar_arrive_expect_tx(bar_full + 8 * stage, cfg::STAGE_BYTES);
                    tma_load_2d(sA, &tmapA, bar_full + 8 * stage, kb * BK, m_blk * BM);
                    tma_load_2d(sB, &tmapW, bar_full + 8 * stage, kb * BK, n_blk * BN);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
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
                        umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(bar_empty + 8 * stage);  // frees the smem slot when the MMAs retire
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(bar_tfull + 8 * acc);  // accumulator ready for the epilogue
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else {
        // ===== epilogue warps =====
        // Warp e owns TMEM lane quadrant (warp & 3) = 32 tile rows and one half of the tile's columns,
        // in chunks of 32 columns.  Phase 1 (thread = row): tcgen05.ld -> raw fp32 accumulators into a
        // swizzled per-warp smem staging block.  Phase 2 (lane = 8 consecutive columns of one row, 8 rows
        // per instruction): bias / GELU / residual in fp32, bf16 pack, 16-byte coalesced stores.  All
        // per-column operands (bias) sit in registers in phase 2; the residual is read coalesced.
        constexpr int NCH = BN / 2 / 32;
        const int ew = warp - 2;
        const int quad = warp & 3;
        const int half = ew >> 2;
        uint8_t* stg = staging + ew * (32 * 128);
        const int piece = lane & 3;   // which 8-column piece of the 32-column chunk (phase 2)
        const int rsub = lane >> 2;   // row within a group of 8 (phase 2)
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
            const int row0 = m_blk * BM + quad * 32;
            const uint32_t tbase = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * ACC_STRIDE +
                                   half * (BN / 2);
            mbar_wait(bar_tfull + 8 * acc, acc_phase);
            tc_fence_after();
            uint32_t v[32];
            tmem_ld_32x32(tbase, v);
#pragma unroll 1
            for (int c = 0; c < NCH; ++c) {
                const int gcol = n_blk * BN + half * (BN / 2) + c * 32 + piece * 8;
                // operands of phase 2 that do not depend on the accumulator: issue their loads early
                const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + gcol));
                const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + gcol) + 1);
                uint4 res[4];
                if (EPI == EPI_BIAS_RESIDUAL) {
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const int grow = row0 + it * 8 + rsub;
                        res[it] = grow < M ? __ldg(reinterpret_cast<const uint4*>(
                                                 residual + static_cast<size_t>(grow) * N + gcol))
                                           : make_uint4(0, 0, 0, 0);
                    }
                }
                tmem_ld_wait();
                // ---- phase 1: registers -> swizzled staging (row = lane, 8 chunks of 4 floats) ----
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    *reinterpret_cast<uint4*>(stg + lane * 128 + ((i ^ (lane & 7)) * 16)) =
                        make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                }
                __syncwarp();
                if (c + 1 < NCH) tmem_ld_32x32(tbase + (c + 1) * 32, v);  // overlaps phase 2
                // ---- phase 2 ----
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int r = it * 8 + rsub;
                    const float4 f0 = *reinterpret_cast<const float4*>(stg + r * 128 + (((2 * piece) ^ (r & 7)) * 16));
                    const float4 f1 = *reinterpret_cast<const float4*>(stg + r * 128 + (((2 * piece + 1) ^ (r & 7)) * 16));
                    float x[8] = {f0.x + b0.x, f0.y + b0.y, f0.z + b0.z, f0.w + b0.w,
                                  f1.x + b1.x, f1.y + b1.y, f1.z + b1.z, f1.w + b1.w};
                    if (EPI == EPI_BIAS_GELU) {
#pragma unroll
                        for (int j = 0; j < 8; j += 2) gelu2(x[j], x[j + 1]);
                    }
                    if (EPI == EPI_BIAS_RESIDUAL) {
                        const uint32_t rw[4] = {res[it].x, res[it].y, res[it].z, res[it].w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            x[2 * j] += bf16_lo(rw[j]);
                            x[2 * j + 1] += bf16_hi(rw[j]);
                        }
                    }
                    const int grow = row0 + r;
                    if (grow < M) {
                        uint4 o;
                        o.x = pack_bf16x2(x[0], x[1]);
                        o.y = pack_bf16x2(x[2], x[3]);
                        o.z = pack_bf16x2(x[4], x[5]);
                        o.w = pack_bf16x2(x[6], x[7]);
                        *reinterpret_cast<uint4*>(C + static_cast<size_t>(grow) * N + gcol) = o;
                    }
                }
                __syncwarp();
            }
            // all of this warp's TMEM reads of the tile have completed (last tmem_ld_wait above)
            tc_fence_before();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

template <int BN, int EPI>
int launch_inst(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const float* bias,
                const __nv_bfloat16* residual, __nv_bfloat16* C, int M, int N, int K, int num_sms,
                cudaStream_t stream) {
    using cfg = Cfg<BN>;
    static bool configured[64] = {};  // per instantiation and per device
    auto kern = gemm_tc_kernel<BN, EPI>;
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      cfg::SMEM_BYTES));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const int m_tiles = (M + BM - 1) / BM;
    const int num_tiles = m_tiles * (N / BN);
    const int grid = num_tiles < num_sms ? num_tiles : num_sms;
    kern<<<grid, GEMM_THREADS, cfg::SMEM_BYTES, stream>>>(tmapA, tmapW, bias, residual, C, M, N, K);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

template <int BN>
int launch_bn(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const float* bias,
              const __nv_bfloat16* residual, __nv_bfloat16* C, int M, int N, int K, int epilogue,
              int num_sms, cudaStream_t stream) {
    switch (epilogue) {
        case EPI_BIAS:
            return launch_inst<BN, EPI_BIAS>(tmapA, tmapW, bias, residual, C, M, N, K, num_sms, stream);
        case EPI_BIAS_GELU:
            return launch_inst<BN, EPI_BIAS_GELU>(tmapA, tmapW, bias, residual, C, M, N, K, num_sms,
                                                  stream);
        case EPI_BIAS_RESIDUAL:
            if (residual == nullptr) return fail(LSC_E_INVALID, "gemm: residual epilogue without residual");
            return launch_inst<BN, EPI_BIAS_RESIDUAL>(tmapA, tmapW, bias, residual, C, M, N, K, num_sms,
                                                      stream);
        default:
            return fail(LSC_E_INVALID, "gemm: unknown epilogue");
    }
}

}  // namespace

int gemm_tc_block_n(int N) {
    if (N % 192 == 0) return 192;
    if (N % 256 == 0) return 256;
    if (N % 128 == 0) return 128;
    return 0;
}

int gemm_tc_launch(const CUtensorMap& tmapA, const CUtensorMap& tmapW, const float* bias,
                   const __nv_bfloat16* residual, __nv_bfloat16* C, int M, int N, int K, int epilogue,
                   int num_sms, cudaStream_t stream) {
    if (M <= 0) return 0;
    if (K % BK != 0 || K <= 0) return fail(LSC_E_UNSUPPORTED, "gemm: K must be a positive multiple of 64");
    if (bias == nullptr) return fail(LSC_E_INVALID, "gemm: bias is required");
    switch (gemm_tc_block_n(N)) {
        case 192:
            return launch_bn<192>(tmapA, tmapW, bias, residual, C, M, N, K, epilogue, num_sms, stream);
        case 256:
            return launch_bn<256>(tmapA, tmapW, bias, residual, C, M, N, K, epilogue, num_sms, stream);
        case 128:
            return launch_bn<128>(tmapA, tmapW, bias, residual, C, M, N, K, epilogue, num_sms, stream);
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
                __nv_bfloat16* __restrict__ C, int M, int N, int K) {
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
            C[static_cast<size_t>(gm) * N + gn] = __float2bfloat16_rn(x);
        }
    }
}

}  // namespace

int gemm_ref_launch(const __nv_bfloat16* A, const __nv_bfloat16* W, const float* bias,
                    const __nv_bfloat16* residual, __nv_bfloat16* C, int M, int N, int K, int epilogue,
                    cudaStream_t stream) {
    if (M <= 0) return 0;
    dim3 grid((N + 63) / 64, (M + 63) / 64);
    switch (epilogue) {
        case EPI_BIAS:
            gemm_ref_kernel<EPI_BIAS><<<grid, 256, 0, stream>>>(A, W, bias, residual, C, M, N, K);
            break;
        case EPI_BIAS_GELU:
            gemm_ref_kernel<EPI_BIAS_GELU><<<grid, 256, 0, stream>>>(A, W, bias, residual, C, M, N, K);
            break;
        case EPI_BIAS_RESIDUAL:
            if (residual == nullptr) return fail(LSC_E_INVALID, "gemm: residual epilogue without residual");
            gemm_ref_kernel<EPI_BIAS_RESIDUAL><<<grid, 256, 0, stream>>>(A, W, bias, residual, C, M, N, K);
            break;
        default:
            return fail(LSC_E_INVALID, "gemm: unknown epilogue");
    }
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
