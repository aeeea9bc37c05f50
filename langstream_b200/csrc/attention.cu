// Packed variable-length multi-head self-attention (K3): ctx = softmax(Q K^T / sqrt(d)) V per
// (sequence, head), over the packed qkv [T, 3H] produced by the QKV GEMM.  Replaces libtorch's
// scaled-dot-product attention with an additive padding mask in the DJL forward pass
// (AI/com/datastax/oss/streaming/ai/embeddings/AbstractHuggingFaceEmbeddingService.java:194-196):
// packing makes the mask implicit — only a sequence's own tokens are ever keys.
//
// Layout of qkv: Q and K thirds bf16, V third fp16 (the QKV GEMM's epilogue writes it that way).
// One CTA per (sequence, head); K (bf16) and V (fp16) of that head are copied to shared memory with
// cp.async and stay there for the CTA's lifetime; each warp owns 16 query rows at a time and streams key blocks of 64 with an
// online (flash-style) softmax.  d_head is 32 or 64, so the contraction depth per tile is tiny and the
// kernel is bound by the exponentials and the instruction issue rate, not by the tensor pipe; it
// therefore uses warp-level mma.sync.m16n8k16 (fp32 accumulate) rather than tcgen05, and spends its
// effort on the softmax:
//   * P = exp2(...) is evaluated with ex2.approx.f16x2 — two probabilities per MUFU operation —
//     and the packed fp16 pair IS the A fragment of the P·V MMA (no separate convert/pack);
//   * the row sums come out of the tensor core too: one extra MMA per 16 keys against an all-ones
//     B fragment accumulates sum_j P[i][j] in fp32, so no per-element additions or shuffles;
//   * the key mask is applied only in a sequence's last, partial key block.
#include <cuda_fp16.h>

#include <cstdlib>

#include <cstring>

#include "common.h"
#include "ptx.cuh"

namespace lsc {

namespace {

constexpr int ATT_THREADS = 256;
constexpr int KEY_BLOCK = 64;

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                            uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1,
                                                  uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                               uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_f16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                              uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// exp2 of two fp32 arguments, evaluated as one packed fp16x2 MUFU operation.
__device__ __forceinline__ uint32_t exp2_f16x2(float lo, float hi) {
    uint32_t h, r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(hi), "f"(lo));  // first source -> upper half
    asm("ex2.approx.f16x2 %0, %1;" : "=r"(r) : "r"(h));
    return r;
}
__device__ __forceinline__ void cp_async_16(uint32_t smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int D>
__global__ void __launch_bounds__(ATT_THREADS, D == 32 ? 3 : 1)
attention_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ cu_seqlens, int H,
                 int heads, __nv_bfloat16* __restrict__ ctx, float scale_log2e) {
    constexpr int LD = D + 8;  // padded row (elements): conflict-free ldmatrix
    extern __shared__ __align__(16) uint8_t att_smem[];
    const int head = blockIdx.x % heads;  // heads of one sequence are adjacent blocks
    const int seq = blockIdx.x / heads;
    const int t0 = __ldg(cu_seqlens + seq);
    const int s = __ldg(cu_seqlens + seq + 1) - t0;
    if (s <= 0) return;
    const int s_pad = (s + KEY_BLOCK - 1) / KEY_BLOCK * KEY_BLOCK;
    __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(att_smem);
    __half* sV = reinterpret_cast<__half*>(sK + static_cast<size_t>(s_pad) * LD);
    const size_t row_stride = static_cast<size_t>(3) * H;
    const __nv_bfloat16* qbase = qkv + static_cast<size_t>(t0) * row_stride + head * D;

    // ---- stage K (bf16) and V (fp16, written as such by the QKV GEMM) with cp.async; rows past the
    //      sequence are zero-filled.  Two commit groups: the first QK^T only needs K. ----
    constexpr int CPR = D / 8;  // 16-byte chunks per row
    const uint32_t sK_addr = smem_u32(sK), sV_addr = smem_u32(sV);
    for (int i = threadIdx.x; i < s_pad * CPR; i += ATT_THREADS) {
        const int r = i / CPR, c = i % CPR;
        const uint32_t dst = sK_addr + (r * LD + c * 8) * 2;
        if (r < s) cp_async_16(dst, qbase + static_cast<size_t>(r) * row_stride + c * 8 + H);
        else *reinterpret_cast<uint4*>(sK + static_cast<size_t>(r) * LD + c * 8) = make_uint4(0, 0, 0, 0);
    }
    cp_async_commit();
    for (int i = threadIdx.x; i < s_pad * CPR; i += ATT_THREADS) {
        const int r = i / CPR, c = i % CPR;
        const uint32_t dst = sV_addr + (r * LD + c * 8) * 2;
        if (r < s) cp_async_16(dst, qbase + static_cast<size_t>(r) * row_stride + c * 8 + 2 * H);
        else *reinterpret_cast<uint4*>(sV + static_cast<size_t>(r) * LD + c * 8) = make_uint4(0, 0, 0, 0);
    }
    cp_async_commit();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, tq = lane & 3;
    const int lm = lane >> 3, lr = lane & 7;  // ldmatrix: matrix index / row within matrix
    constexpr uint32_t ONES = 0x3C003C00u;    // fp16x2 (1, 1)
    bool k_ready = false, v_ready = false;    // warp-uniform

    for (int qb = warp; qb * 16 < s; qb += ATT_THREADS / 32) {
        // ---- Q fragments (A operand), straight from global (overlaps the K/V copies) ----
        uint32_t qa[D / 16][4];
        const int r_lo = qb * 16 + g, r_hi = r_lo + 8;
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
            const int c = ks * 16 + 2 * tq;
            const uint32_t* plo = reinterpret_cast<const uint32_t*>(qbase + static_cast<size_t>(r_lo) * row_stride + c);
            const uint32_t* phi = reinterpret_cast<const uint32_t*>(qbase + static_cast<size_t>(r_hi) * row_stride + c);
            qa[ks][0] = r_lo < s ? __ldg(plo) : 0u;
            qa[ks][1] = r_hi < s ? __ldg(phi) : 0u;
            qa[ks][2] = r_lo < s ? __ldg(plo + 4) : 0u;
            qa[ks][3] = r_hi < s ? __ldg(phi + 4) : 0u;
        }
        if (!k_ready) {  // every warp passes each of the two barriers exactly once (here or after the loop)
            cp_async_wait<1>();
            __syncthreads();
            k_ready = true;
        }
        // o[0..D/8) = output accumulators, o[D/8] = row sums (every column of that tile holds the sum)
        float o[D / 8 + 1][4];
#pragma unroll
        for (int i = 0; i <= D / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
        float m_run[2] = {-INFINITY, -INFINITY};

        for (int kb0 = 0; kb0 < s; kb0 += KEY_BLOCK) {
            float sc[8][4];
#pragma unroll
            for (int i = 0; i < 8; ++i) sc[i][0] = sc[i][1] = sc[i][2] = sc[i][3] = 0.f;
            // ---- S = Q K^T ----
#pragma unroll
            for (int ks = 0; ks < D / 16; ++ks) {
#pragma unroll
                for (int nt = 0; nt < 8; nt += 2) {
                    const int key = kb0 + (nt + (lm >> 1)) * 8 + lr;
                    const int dof = ks * 16 + (lm & 1) * 8;
                    uint32_t b0, b1, b2, b3;
                    ldmatrix_x4(sK_addr + (key * LD + dof) * 2, b0, b1, b2, b3);
                    mma_bf16_16816(sc[nt], qa[ks], b0, b1);
                    mma_bf16_16816(sc[nt + 1], qa[ks], b2, b3);
                }
            }
            // ---- mask (last partial block only) and running max ----
            if (kb0 + KEY_BLOCK > s) {
#pragma unroll
                for (int nt = 0; nt < 8; ++nt)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (kb0 + nt * 8 + 2 * tq + (j & 1) >= s) sc[nt][j] = -INFINITY;
            }
            float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                mx[0] = fmaxf(mx[0], fmaxf(sc[nt][0], sc[nt][1]));
                mx[1] = fmaxf(mx[1], fmaxf(sc[nt][2], sc[nt][3]));
            }
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
                mx[h2] = fmaxf(mx[h2], __shfl_xor_sync(0xffffffffu, mx[h2], 1));
                mx[h2] = fmaxf(mx[h2], __shfl_xor_sync(0xffffffffu, mx[h2], 2));
            }
            float nms[2];  // -max * scale
            if (kb0 == 0) {
                m_run[0] = mx[0];
                m_run[1] = mx[1];
            } else {
                float corr[2];
#pragma unroll
                for (int h2 = 0; h2 < 2; ++h2) {
                    const float m_new = fmaxf(m_run[h2], mx[h2]);  // finite: every block has a valid key
                    corr[h2] = exp2f((m_run[h2] - m_new) * scale_log2e);
                    m_run[h2] = m_new;
                }
#pragma unroll
                for (int i = 0; i <= D / 8; ++i) {
                    o[i][0] *= corr[0];
                    o[i][1] *= corr[0];
                    o[i][2] *= corr[1];
                    o[i][3] *= corr[1];
                }
            }
            nms[0] = -m_run[0] * scale_log2e;
            nms[1] = -m_run[1] * scale_log2e;
            if (!v_ready) {
                cp_async_wait<0>();
                __syncthreads();
                v_ready = true;
            }
            // ---- P = exp2(S*scale - max*scale) as packed fp16 A fragments; O += P V; sums += P 1 ----
#pragma unroll
            for (int kk = 0; kk < KEY_BLOCK / 16; ++kk) {
                uint32_t pa[4];
                pa[0] = exp2_f16x2(fmaf(sc[2 * kk][0], scale_log2e, nms[0]), fmaf(sc[2 * kk][1], scale_log2e, nms[0]));
                pa[1] = exp2_f16x2(fmaf(sc[2 * kk][2], scale_log2e, nms[1]), fmaf(sc[2 * kk][3], scale_log2e, nms[1]));
                pa[2] = exp2_f16x2(fmaf(sc[2 * kk + 1][0], scale_log2e, nms[0]), fmaf(sc[2 * kk + 1][1], scale_log2e, nms[0]));
                pa[3] = exp2_f16x2(fmaf(sc[2 * kk + 1][2], scale_log2e, nms[1]), fmaf(sc[2 * kk + 1][3], scale_log2e, nms[1]));
#pragma unroll
                for (int nt = 0; nt < D / 8; nt += 2) {
                    const int key = kb0 + kk * 16 + (lm & 1) * 8 + lr;
                    const int dof = (nt + (lm >> 1)) * 8;
                    uint32_t b0, b1, b2, b3;
                    ldmatrix_x4_trans(sV_addr + (key * LD + dof) * 2, b0, b1, b2, b3);
                    mma_f16_16816(o[nt], pa, b0, b1);
                    mma_f16_16816(o[nt + 1], pa, b2, b3);
                }
                mma_f16_16816(o[D / 8], pa, ONES, ONES);
            }
        }
        // ---- normalise and store ----
        const float inv0 = 1.0f / o[D / 8][0], inv1 = 1.0f / o[D / 8][2];
        __nv_bfloat16* obase = ctx + static_cast<size_t>(t0) * H + head * D;
#pragma unroll
        for (int nt = 0; nt < D / 8; ++nt) {
            const int c = nt * 8 + 2 * tq;
            if (r_lo < s)
                *reinterpret_cast<uint32_t*>(obase + static_cast<size_t>(r_lo) * H + c) =
                    pack_bf16x2(o[nt][0] * inv0, o[nt][1] * inv0);
            if (r_hi < s)
                *reinterpret_cast<uint32_t*>(obase + static_cast<size_t>(r_hi) * H + c) =
                    pack_bf16x2(o[nt][2] * inv1, o[nt][3] * inv1);
        }
    }
    if (!k_ready) {  // warps without a query block still take part in the two barriers
        cp_async_wait<1>();
        __syncthreads();
    }
    if (!v_ready) {
        cp_async_wait<0>();
        __syncthreads();
    }
}

// ---- persistent, double-buffered variant (short sequences) ---------------------------------------
// One CTA loops over (sequence, head) items; while it computes item i out of one shared-memory buffer,
// cp.async fills the other buffer with Q, K and V of item i + gridDim.x, so the global-memory latency
// of every item but the first is hidden.  Q comes from shared memory through ldmatrix as well.
template <int D>
__global__ void __launch_bounds__(ATT_THREADS, D == 32 ? 3 : 2)
attention_pipelined_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ cu_seqlens, int H,
                           int heads, int n_items, int s_pad_max, __nv_bfloat16* __restrict__ ctx,
                           float scale_log2e) {
    constexpr int LD = D + 8;
    constexpr int CPR = D / 8;
    extern __shared__ __align__(16) uint8_t att_smem[];
    const uint32_t smem_addr = smem_u32(att_smem);
    const uint32_t mat_bytes = static_cast<uint32_t>(s_pad_max) * LD * 2;  // one of Q / K / V
    const size_t row_stride = static_cast<size_t>(3) * H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, tq = lane & 3;
    const int lm = lane >> 3, lr = lane & 7;
    constexpr uint32_t ONES = 0x3C003C00u;

    auto issue = [&](int item, int buf) {
        const int seq = item / heads, head = item % heads;
        const int t0 = __ldg(cu_seqlens + seq);
        const int s = __ldg(cu_seqlens + seq + 1) - t0;
        const int s_pad = (s + KEY_BLOCK - 1) / KEY_BLOCK * KEY_BLOCK;
        const __nv_bfloat16* base = qkv + static_cast<size_t>(t0) * row_stride + head * D;
        const uint32_t dst0 = smem_addr + buf * 3 * mat_bytes;
        for (int i = threadIdx.x; i < 3 * s_pad * CPR; i += ATT_THREADS) {
            const int which = i / (s_pad * CPR);  // 0 Q, 1 K, 2 V
            const int rem = i - which * s_pad * CPR;
            const int r = rem / CPR, c = rem % CPR;
            const uint32_t dst = dst0 + which * mat_bytes + (r * LD + c * 8) * 2;
            if (r < s) cp_async_16(dst, base + static_cast<size_t>(r) * row_stride + which * H + c * 8);
            else asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(dst), "r"(0) : "memory");
        }
    };

    int item = blockIdx.x;
    if (item < n_items) issue(item, 0);
    cp_async_commit();
    for (int buf = 0; item < n_items; item += gridDim.x, buf ^= 1) {
        const int next = item + gridDim.x;
        if (next < n_items) issue(next, buf ^ 1);
        cp_async_commit();   // possibly empty: keeps the group count uniform
        cp_async_wait<1>();  // the current item's copies have landed
        __syncthreads();
        const int seq = item / heads, head = item % heads;
        const int t0 = __ldg(cu_seqlens + seq);
        const int s = __ldg(cu_seqlens + seq + 1) - t0;
        const uint32_t sQ_addr = smem_addr + buf * 3 * mat_bytes;
        const uint32_t sK_addr = sQ_addr + mat_bytes, sV_addr = sK_addr + mat_bytes;
        for (int qb = warp; qb * 16 < s; qb += ATT_THREADS / 32) {
            uint32_t qa[D / 16][4];
            const int r_lo = qb * 16 + g, r_hi = r_lo + 8;
#pragma unroll
            for (int ks = 0; ks < D / 16; ++ks) {
                const int row = qb * 16 + (lm & 1) * 8 + lr;
                const int col = ks * 16 + (lm >> 1) * 8;
                ldmatrix_x4(sQ_addr + (row * LD + col) * 2, qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
            }
            // o[0..D/8) = output accumulators, o[D/8] = row sums (every column of that tile holds the sum)
            float o[D / 8 + 1][4];
    #pragma unroll
            for (int i = 0; i <= D / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
            float m_run[2] = {-INFINITY, -INFINITY};

            for (int kb0 = 0; kb0 < s; kb0 += KEY_BLOCK) {
                float sc[8][4];
    #pragma unroll
                for (int i = 0; i < 8; ++i) sc[i][0] = sc[i][1] = sc[i][2] = sc[i][3] = 0.f;
                // ---- S = Q K^T ----
    #pragma unroll
                for (int ks = 0; ks < D / 16; ++ks) {
    #pragma unroll
                    for (int nt = 0; nt < 8; nt += 2) {
                        const int key = kb0 + (nt + (lm >> 1)) * 8 + lr;
                        const int dof = ks * 16 + (lm & 1) * 8;
                        uint32_t b0, b1, b2, b3;
                        ldmatrix_x4(sK_addr + (key * LD + dof) * 2, b0, b1, b2, b3);
                        mma_bf16_16816(sc[nt], qa[ks], b0, b1);
                        mma_bf16_16816(sc[nt + 1], qa[ks], b2, b3);
                    }
                }
                // ---- mask (last partial block only) and running max ----
                if (kb0 + KEY_BLOCK > s) {
    #pragma unroll
                    for (int nt = 0; nt < 8; ++nt)
    #pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (kb0 + nt * 8 + 2 * tq + (j & 1) >= s) sc[nt][j] = -INFINITY;
                }
                float mx[2] = {-INFINITY, -INFINITY};
    #pragma unroll
                for (int nt = 0; nt < 8; ++nt) {
                    mx[0] = fmaxf(mx[0], fmaxf(sc[nt][0], sc[nt][1]));
                    mx[1] = fmaxf(mx[1], fmaxf(sc[nt][2], sc[nt][3]));
                }
    #pragma unroll
                for (int h2 = 0; h2 < 2; ++h2) {
                    mx[h2] = fmaxf(mx[h2], __shfl_xor_sync(0xffffffffu, mx[h2], 1));
                    mx[h2] = fmaxf(mx[h2], __shfl_xor_sync(0xffffffffu, mx[h2], 2));
                }
                float nms[2];  // -max * scale
                if (kb0 == 0) {
                    m_run[0] = mx[0];
                    m_run[1] = mx[1];
                } else {
                    float corr[2];
    #pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        const float m_new = fmaxf(m_run[h2], mx[h2]);  // finite: every block has a valid key
                        corr[h2] = exp2f((m_run[h2] - m_new) * scale_log2e);
                        m_run[h2] = m_new;
                    }
    #pragma unroll
                    for (int i = 0; i <= D / 8; ++i) {
                        o[i][0] *= corr[0];
                        o[i][1] *= corr[0];
                        o[i][2] *= corr[1];
                        o[i][3] *= corr[1];
                    }
                }
                nms[0] = -m_run[0] * scale_log2e;
                nms[1] = -m_run[1] * scale_log2e;
                // ---- P = exp2(S*scale - max*scale) as packed fp16 A fragments; O += P V; sums += P 1 ----
    #pragma unroll
                for (int kk = 0; kk < KEY_BLOCK / 16; ++kk) {
                    uint32_t pa[4];
                    pa[0] = exp2_f16x2(fmaf(sc[2 * kk][0], scale_log2e, nms[0]), fmaf(sc[2 * kk][1], scale_log2e, nms[0]));
                    pa[1] = exp2_f16x2(fmaf(sc[2 * kk][2], scale_log2e, nms[1]), fmaf(sc[2 * kk][3], scale_log2e, nms[1]));
                    pa[2] = exp2_f16x2(fmaf(sc[2 * kk + 1][0], scale_log2e, nms[0]), fmaf(sc[2 * kk + 1][1], scale_log2e, nms[0]));
                    pa[3] = exp2_f16x2(fmaf(sc[2 * kk + 1][2], scale_log2e, nms[1]), fmaf(sc[2 * kk + 1][3], scale_log2e, nms[1]));
    #pragma unroll
                    for (int nt = 0; nt < D / 8; nt += 2) {
                        const int key = kb0 + kk * 16 + (lm & 1) * 8 + lr;
                        const int dof = (nt + (lm >> 1)) * 8;
                        uint32_t b0, b1, b2, b3;
                        ldmatrix_x4_trans(sV_addr + (key * LD + dof) * 2, b0, b1, b2, b3);
                        mma_f16_16816(o[nt], pa, b0, b1);
                        mma_f16_16816(o[nt + 1], pa, b2, b3);
                    }
                    mma_f16_16816(o[D / 8], pa, ONES, ONES);
                }
            }
            // ---- normalise and store ----
            const float inv0 = 1.0f / o[D / 8][0], inv1 = 1.0f / o[D / 8][2];
            __nv_bfloat16* obase = ctx + static_cast<size_t>(t0) * H + head * D;
    #pragma unroll
            for (int nt = 0; nt < D / 8; ++nt) {
                const int c = nt * 8 + 2 * tq;
                if (r_lo < s)
                    *reinterpret_cast<uint32_t*>(obase + static_cast<size_t>(r_lo) * H + c) =
                        pack_bf16x2(o[nt][0] * inv0, o[nt][1] * inv0);
                if (r_hi < s)
                    *reinterpret_cast<uint32_t*>(obase + static_cast<size_t>(r_hi) * H + c) =
                        pack_bf16x2(o[nt][2] * inv1, o[nt][3] * inv1);
            }
        }
        __syncthreads();  // the buffer is refilled by the next iteration's prefetch
    }
    cp_async_wait<0>();
}

template <int D>
int launch_d(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int H, int heads,
             int max_seqlen, __nv_bfloat16* ctx, cudaStream_t stream) {
    const int s_pad = (max_seqlen + KEY_BLOCK - 1) / KEY_BLOCK * KEY_BLOCK;
    const float scale = 1.4426950408889634f / sqrtf(static_cast<float>(D));
    // short sequences: persistent CTAs with double-buffered Q/K/V (two or three CTAs per SM)
    const size_t smem_pipe = static_cast<size_t>(2) * 3 * s_pad * (D + 8) * sizeof(__nv_bfloat16);
    if (smem_pipe <= 112 * 1024) {
        static int sms_cache[64] = {};
        int dev = 0;
        LSC_CUDA(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && sms_cache[dev] == 0)
            LSC_CUDA(cudaDeviceGetAttribute(&sms_cache[dev], cudaDevAttrMultiProcessorCount, dev));
        const int sms = (dev >= 0 && dev < 64) ? sms_cache[dev] : 148;
        auto pk = attention_pipelined_kernel<D>;
        LSC_CUDA(cudaFuncSetAttribute(pk, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_pipe)));
        const int per_sm = smem_pipe <= 74 * 1024 ? 3 : 2;
        const long long n_items = static_cast<long long>(heads) * n_seq;
        const int grid_p = static_cast<int>(n_items < static_cast<long long>(sms) * per_sm ? n_items : static_cast<long long>(sms) * per_sm);
        pk<<<grid_p, ATT_THREADS, smem_pipe, stream>>>(qkv, cu_seqlens, H, heads, static_cast<int>(n_items), s_pad, ctx, scale);
        LSC_CUDA(cudaGetLastError());
        return 1;
    }
    const size_t smem = static_cast<size_t>(2) * s_pad * (D + 8) * sizeof(__nv_bfloat16);
    if (smem > 227 * 1024) return fail(LSC_E_UNSUPPORTED, "attention: sequence too long for shared memory");
    auto kern = attention_kernel<D>;
    if (smem > 48 * 1024) {
        LSC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    }
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(D));
    const unsigned grid = static_cast<unsigned>(heads) * static_cast<unsigned>(n_seq);
    kern<<<grid, ATT_THREADS, smem, stream>>>(qkv, cu_seqlens, H, heads, ctx, scale_log2e);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace

bool attention_wants_order(int n_seq, int T, int H, int heads, int max_seqlen) {
    static const bool no_ragged = getenv("LSC_ATTN_NO_RAGGED") != nullptr;  // debug: one full tile per item whatever its length
    static const bool legacy_only = getenv("LSC_ATTN_LEGACY") != nullptr;
    return !no_ragged && !legacy_only && n_seq > 0 && heads > 0 && H / heads == 32 && max_seqlen <= 128 &&
           static_cast<long long>(n_seq) * 128 != T;
}

int attention_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int T, int H, int heads,
                     int max_seqlen, __nv_bfloat16* ctx, cudaStream_t stream, const int32_t* tiles, const uint32_t* rowinfo,
                     const int32_t* order, const float* rel_tbl, const int32_t* blocks) {
    if (n_seq <= 0 || max_seqlen <= 0) return 0;
    const int d = H / heads;
    if (d * heads != H) return fail(LSC_E_INVALID, "attention: hidden not divisible by heads");
    static const bool legacy_only = getenv("LSC_ATTN_LEGACY") != nullptr;  // debug: force the mma.sync kernel
    if (rel_tbl != nullptr && !(d == 64 && attention_tc64_supported(H, heads)))
        return fail(LSC_E_UNSUPPORTED, "attention: a relative-position bias needs the d_head 64 tcgen05 kernel");
    if (d == 32 && max_seqlen <= 128 && !legacy_only) {
        static int sms_cache[64] = {};
        int dev = 0;
        LSC_CUDA(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && sms_cache[dev] == 0)
            LSC_CUDA(cudaDeviceGetAttribute(&sms_cache[dev], cudaDevAttrMultiProcessorCount, dev));
        const int sms = (dev >= 0 && dev < 64) ? sms_cache[dev] : 148;
        if (blocks != nullptr && attention_wants_order(n_seq, T, H, heads, max_seqlen))  // tiles of 32-row blocks
            return attention_tc_blocks_launch(qkv, blocks, n_seq, T, H, heads, max_seqlen, ctx, sms, stream);
        if (tiles != nullptr && rowinfo != nullptr)  // short sequences sharing 128-row tiles (attention_tc_packed.cu)
            return attention_tc_packed_launch(qkv, cu_seqlens, n_seq, T, H, heads, max_seqlen, tiles, rowinfo, ctx, sms, stream);
        if (attention_wants_order(n_seq, T, H, heads, max_seqlen))  // some sequence is shorter than its 128-row tile
            return attention_tc_ragged_launch(qkv, cu_seqlens, order, n_seq, T, H, heads, max_seqlen, ctx, sms, stream);
        return attention_tc_launch(qkv, cu_seqlens, n_seq, T, H, heads, max_seqlen, ctx, sms, stream);
    }
    if (d == 64 && attention_tc64_supported(H, heads)) {  // tcgen05, any sequence length
        static int sms_cache64[64] = {};
        int dev = 0;
        LSC_CUDA(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && sms_cache64[dev] == 0)
            LSC_CUDA(cudaDeviceGetAttribute(&sms_cache64[dev], cudaDevAttrMultiProcessorCount, dev));
        return attention_tc64_launch(qkv, cu_seqlens, n_seq, T, H, heads, ctx, (dev >= 0 && dev < 64) ? sms_cache64[dev] : 148,
                                     stream, rel_tbl);
    }
    if (d == 32) return launch_d<32>(qkv, cu_seqlens, n_seq, H, heads, max_seqlen, ctx, stream);
    if (d == 64) return launch_d<64>(qkv, cu_seqlens, n_seq, H, heads, max_seqlen, ctx, stream);
    return fail(LSC_E_UNSUPPORTED, "attention: head dimension must be 32 or 64");
}

}  // namespace lsc

extern "C" int lsc_debug_attention_dev(const void* d_qkv, const int32_t* d_cu_seqlens, int32_t n_seq,
                                       int32_t total_tokens, int32_t hidden, int32_t heads, int32_t max_seqlen,
                                       void* d_ctx, void* stream) {
    if (d_qkv == nullptr || d_cu_seqlens == nullptr || d_ctx == nullptr || n_seq < 0 || hidden <= 0 || heads <= 0)
        return lsc::fail(LSC_E_INVALID, "lsc_debug_attention_dev: bad argument");
    // LSC_DEBUG_ATTN_PACK=1 (tests): run with the packed tile list the encoder builds for batches of short sequences
    int32_t* tiles = nullptr;
    uint32_t* rowinfo = nullptr;
    const char* pack = getenv("LSC_DEBUG_ATTN_PACK");
    if (pack != nullptr && atoi(pack) != 0 && n_seq > 0 && n_seq <= lsc::attention_tiles_max_seqs() && hidden / heads == 32 &&
        max_seqlen <= 128) {
        if (cudaMalloc(reinterpret_cast<void**>(&tiles), static_cast<size_t>(2 * n_seq + 2 + total_tokens) * sizeof(int32_t)) != cudaSuccess)
            return lsc::fail(LSC_E_NOMEM, "lsc_debug_attention_dev: tile list");
        rowinfo = reinterpret_cast<uint32_t*>(tiles + 2 * n_seq + 2);
        const int trc = lsc::attention_tiles_launch(d_cu_seqlens, n_seq, total_tokens, tiles, rowinfo, static_cast<cudaStream_t>(stream));
        if (trc < 0) {
            cudaFree(tiles);
            return trc;
        }
    }
    // ragged batches: the block tile list, as the encoder builds it once per pass (LSC_ATTN_NO_BLOCKS=1: the older forms)
    int32_t* blocks = nullptr;
    if (tiles == nullptr && getenv("LSC_ATTN_NO_BLOCKS") == nullptr && lsc::attention_wants_order(n_seq, total_tokens, hidden, heads, max_seqlen)) {
        if (cudaMalloc(reinterpret_cast<void**>(&blocks), static_cast<size_t>(lsc::attention_blocks_scratch_ints(n_seq)) * sizeof(int32_t)) != cudaSuccess)
            return lsc::fail(LSC_E_NOMEM, "lsc_debug_attention_dev: block tile list");
        const int brc = lsc::attention_blocks_launch(d_cu_seqlens, n_seq, blocks, static_cast<cudaStream_t>(stream));
        if (brc < 0) {
            cudaFree(blocks);
            return brc;
        }
    }
    // ... or the sequences by descending length (LSC_DEBUG_ATTN_NO_ORDER=1: batch order)
    int32_t* order = nullptr;
    if (tiles == nullptr && blocks == nullptr && getenv("LSC_DEBUG_ATTN_NO_ORDER") == nullptr && lsc::attention_wants_order(n_seq, total_tokens, hidden, heads, max_seqlen)) {
        if (cudaMalloc(reinterpret_cast<void**>(&order), static_cast<size_t>(n_seq) * sizeof(int32_t)) != cudaSuccess)
            return lsc::fail(LSC_E_NOMEM, "lsc_debug_attention_dev: order list");
        const int orc = lsc::attention_order_launch(d_cu_seqlens, n_seq, order, static_cast<cudaStream_t>(stream));
        if (orc < 0) {
            cudaFree(order);
            return orc;
        }
    }
    int rc = lsc::attention_launch(static_cast<const __nv_bfloat16*>(d_qkv), d_cu_seqlens, n_seq, total_tokens, hidden,
                                   heads, max_seqlen, static_cast<__nv_bfloat16*>(d_ctx), static_cast<cudaStream_t>(stream), tiles, rowinfo,
                                   order, nullptr, blocks);
    if (tiles != nullptr || order != nullptr || blocks != nullptr) {
        cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
        if (tiles != nullptr) cudaFree(tiles);
        if (order != nullptr) cudaFree(order);
        if (blocks != nullptr) cudaFree(blocks);
    }
    return rc < 0 ? rc : LSC_OK;
}

extern "C" int lsc_debug_attention_blocks_dev(const int32_t* d_cu_seqlens, int32_t n_seq, int32_t* d_out, void* stream) {
    if (d_cu_seqlens == nullptr || d_out == nullptr || n_seq < 0)
        return lsc::fail(LSC_E_INVALID, "lsc_debug_attention_blocks_dev: bad argument");
    const int rc = lsc::attention_blocks_launch(d_cu_seqlens, n_seq, d_out, static_cast<cudaStream_t>(stream));
    return rc < 0 ? rc : LSC_OK;
}

extern "C" int lsc_debug_mpnet_bias_table(const float* weight, int32_t rel_buckets, int32_t rel_max_distance, int32_t heads,
                                          float* out) {
    if (weight == nullptr || out == nullptr || heads < 1 || rel_buckets < 4 || rel_buckets % 4 != 0 ||
        rel_max_distance <= rel_buckets / 4 || rel_max_distance > 128)
        return lsc::fail(LSC_E_INVALID, "lsc_debug_mpnet_bias_table: bad argument");
    std::vector<float> tbl;
    lsc::attention_rel_table(weight, rel_buckets, rel_max_distance, heads, &tbl);
    memcpy(out, tbl.data(), tbl.size() * sizeof(float));
    return LSC_OK;
}
