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
//   warps 2-5   epilogue: tcgen05.ld -> bias (+GELU | +residual) in fp32 -> bf16 -> swizzled smem
//               staging -> coalesced 16-byte global stores
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA) and TMEM full/empty (MMA <-> epilogue).
#include "common.h"
#include "ptx.cuh"

namespace lsc {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 192;
constexpr int TMEM_COLS = 512;
constexpr int ACC_STRIDE = 256;  // TMEM columns between the two accumulator stages
constexpr int STAGING_BYTES = 4 * 32 * 128;

template <int BN>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (BN <= 128) ? 6 : (BN <= 192 ? 5 : 4);
    static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + STAGING_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFFSET + 256 + 1024;  // + barriers + alignment slack
};

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
            mbar_init(bar_tempty + 8 * a, 4);  // one arrive per epilogue warp
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
        const int ew = warp - 2;
        const int quad = warp & 3;  // TMEM lane quadrant this warp may read
        uint8_t* stg = staging + ew * (32 * 128);
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
            mbar_wait(bar_tfull + 8 * acc, acc_phase);
            tc_fence_after();
            const int row0 = m_blk * BM + quad * 32;
            const int my_row = row0 + lane;
#pragma unroll 1
            for (int c = 0; c < BN / 64; ++c) {
                uint32_t v0[32], v1[32];
                const uint32_t taddr =
                    tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * ACC_STRIDE + c * 64;
                tmem_ld_32x32(taddr, v0);
                tmem_ld_32x32(taddr + 32, v1);
                tmem_ld_wait();
                const int col0 = n_blk * BN + c * 64;
                uint4 res[8];
                if (EPI == EPI_BIAS_RESIDUAL) {
                    if (my_row < M) {
                        const uint4* rp = reinterpret_cast<const uint4*>(
                            residual + static_cast<size_t>(my_row) * N + col0);
#pragma unroll
                        for (int i = 0; i < 8; ++i) res[i] = __ldg(rp + i);
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i) res[i] = make_uint4(0, 0, 0, 0);
                    }
                }
                const float4* bp = reinterpret_cast<const float4*>(bias + col0);
#pragma unroll
                for (int i = 0; i < 8; ++i) {  // 8 chunks of 8 columns (16 bytes of bf16)
                    float x[8];
                    const float4 b0 = __ldg(bp + 2 * i), b1 = __ldg(bp + 2 * i + 1);
                    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int col = i * 8 + j;
                        const uint32_t raw = col < 32 ? v0[col] : v1[col - 32];
                        x[j] = __uint_as_float(raw) + bb[j];
                    }
                    if (EPI == EPI_BIAS_GELU) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) x[j] = gelu_erf(x[j]);
                    }
                    if (EPI == EPI_BIAS_RESIDUAL) {
                        const uint32_t rw[4] = {res[i].x, res[i].y, res[i].z, res[i].w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            x[2 * j] += bf16_lo(rw[j]);
                            x[2 * j + 1] += bf16_hi(rw[j]);
                        }
                    }
                    uint4 o;
                    o.x = pack_bf16x2(x[0], x[1]);
                    o.y = pack_bf16x2(x[2], x[3]);
                    o.z = pack_bf16x2(x[4], x[5]);
                    o.w = pack_bf16x2(x[6], x[7]);
                    *reinterpret_cast<uint4*>(stg + lane * 128 + ((i ^ (lane & 7)) * 16)) = o;
                }
                __syncwarp();
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int r = it * 4 + (lane >> 3);
                    const int ch = lane & 7;
                    const uint4 val =
                        *reinterpret_cast<const uint4*>(stg + r * 128 + ((ch ^ (r & 7)) * 16));
                    const int grow = row0 + r;
                    if (grow < M) {
                        *reinterpret_cast<uint4*>(C + static_cast<size_t>(grow) * N + col0 + ch * 8) =
                            val;
                    }
                }
                __syncwarp();
            }
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
