// Packed variable-length attention for d_head = 64 and any sequence length (the "base" shapes: BGE-base,
// 12 heads of 64, up to 512 tokens) on tcgen05 - the long-sequence sibling of attention_tc.cu, replacing the
// scaled-dot-product attention inside libtorch's BertSelfAttention (reference call site:
// AbstractHuggingFaceEmbeddingService.java:194-196, predictor.predict).
//
// Work item = (sequence, head).  Its queries are cut into tiles of 128 rows, its keys into tiles of 64; four
// softmax groups (four warps each, thread = query row = TMEM lane) work on four query tiles of the item at the
// same time and share every key / value tile, which is streamed through one shared-memory ring:
//   pass A  for every key tile:  S_g = Q_g . K^T (tcgen05, 128 x 64, K = 64)  ->  running row maximum
//   pass B  for every key tile:  S_g again -> P_g = exp2(S_g * c - max * c) as fp16 into a SWIZZLE_128B K-major
//           tile -> O_g += P_g . V (tcgen05, A = P_g from shared memory, B = V MN-major exactly as it lies in
//           memory), row sums in registers
//   then    O_g / sum -> bf16 -> ctx
// Computing S twice costs tensor time, which this kernel has to spare (it is bound by the exponentials); it
// buys an exact maximum before any exponential is taken, so O accumulates in TMEM with plain accumulation and
// never needs rescaling.  TMEM: per group 64 columns of S and 64 of O (4 x 128 = 512).
//
//   warp 0      TMA producer: the item's Q tiles (once per round of four query tiles), K / K+V tiles in the ring
//   warp 1      TMEM allocation, S issuer        warp 2      P.V issuer        warp 3      idle
//   warps 4-19  softmax groups 0-3
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "common.h"
#include "ptx.cuh"

namespace lsc {
namespace {

constexpr int A6_G = 4;                   // softmax groups = query tiles in flight
constexpr int A6_QT = 128;                // query rows per tile
constexpr int A6_KT = 64;                 // keys per tile
constexpr int A6_D = 64;
constexpr int A6_NST = 4;                 // ring stages
constexpr int A6_THREADS = 128 + A6_G * 128;
constexpr int A6_Q_BYTES = A6_QT * A6_D * 2;      // 16 KB
constexpr int A6_KV_BYTES = A6_KT * A6_D * 2;     // 8 KB: one K or V tile
constexpr int A6_STAGE_BYTES = 2 * A6_KV_BYTES;   // K | V
constexpr int A6_P_BYTES = A6_QT * A6_KT * 2;     // 16 KB: P tile, fp16 [128 x 64]
constexpr int A6_SMEM = A6_G * A6_Q_BYTES + A6_NST * A6_STAGE_BYTES + A6_G * A6_P_BYTES + 512 + 1024;
// BIAS variant (MPNet): the relative-position bias of every head, [heads][ATT_REL_W] fp32 behind the barriers.  Entry
// ATT_REL_HALF + r is the bias of (key position - query position) = r, times log2(e).  The buckets saturate at |r| >= 128,
// so a thread whose 64 keys of a tile all lie beyond that may read any 64 saturated entries: clamping the first index
// to [0, ATT_REL_W - 64] is exact and the table needs |r| <= 191 only.
constexpr int A6_REL_MAXHEADS = (227 * 1024 - A6_SMEM) / (ATT_REL_W * 4);

// barrier slots
enum {
    B6_FULL = 0,                       // ring full  [A6_NST]
    B6_EMPTY = B6_FULL + A6_NST,       // ring empty [A6_NST]
    B6_QFULL = B6_EMPTY + A6_NST,
    B6_QFREE,
    B6_SFULL,                          // [A6_G]
    B6_SFREE = B6_SFULL + A6_G,
    B6_PREADY = B6_SFREE + A6_G,
    B6_PFREE = B6_PREADY + A6_G,
    B6_OFULL = B6_PFREE + A6_G,
    B6_OFREE = B6_OFULL + A6_G,
    B6_COUNT = B6_OFREE + A6_G
};

// instruction descriptor: fp32 accumulate; fmt 0 = f16, 1 = bf16; B operand K-major or MN-major
__host__ __device__ constexpr uint32_t idesc64(uint32_t M, uint32_t N, uint32_t fmt, uint32_t b_mn_major) {
    return (1u << 4) | (fmt << 7) | (fmt << 10) | (b_mn_major << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <bool BIAS>
__global__ void __launch_bounds__(A6_THREADS, 1)
attention_tc64_kernel(const __grid_constant__ CUtensorMap tmapQ, const __grid_constant__ CUtensorMap tmapKV,
                      const int32_t* __restrict__ cu_seqlens, int H, int heads, int n_items,
                      __nv_bfloat16* __restrict__ ctx, float scale_log2e, const float* __restrict__ rel_tbl) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t sQ = smem_u32(smem);                       // [A6_G] query tiles
    const uint32_t sR = sQ + A6_G * A6_Q_BYTES;               // ring: [A6_NST] x (K | V)
    const uint32_t sP = sR + A6_NST * A6_STAGE_BYTES;         // [A6_G] P tiles
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A6_G * A6_Q_BYTES + A6_NST * A6_STAGE_BYTES + A6_G * A6_P_BYTES);
    const uint32_t bar0 = smem_u32(bars);
    auto bar = [&](int slot) { return bar0 + 8u * static_cast<uint32_t>(slot); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + B6_COUNT);
    const uint32_t sT = bar0 + 512;  // BIAS: [heads][ATT_REL_W] fp32

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if constexpr (BIAS) {  // model constants: nothing the previous kernel writes
        float* t = reinterpret_cast<float*>(smem + A6_G * A6_Q_BYTES + A6_NST * A6_STAGE_BYTES + A6_G * A6_P_BYTES + 512);
        for (int i = threadIdx.x; i < heads * ATT_REL_W; i += A6_THREADS) t[i] = __ldg(rel_tbl + i);
    }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapQ);
        tma_prefetch_desc(&tmapKV);
        for (int s = 0; s < A6_NST; ++s) {
            mbar_init(bar(B6_FULL + s), 1);
            mbar_init(bar(B6_EMPTY + s), 1);
        }
        mbar_init(bar(B6_QFULL), 1);
        mbar_init(bar(B6_QFREE), 1);
        for (int g = 0; g < A6_G; ++g) {
            mbar_init(bar(B6_SFULL + g), 1);
            mbar_init(bar(B6_SFREE + g), 4);    // one arrive per softmax warp of the group
            mbar_init(bar(B6_PREADY + g), 4);
            mbar_init(bar(B6_PFREE + g), 1);
            mbar_init(bar(B6_OFULL + g), 1);
            mbar_init(bar(B6_OFREE + g), 4);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
        tmem_relinquish();
    }
    if (threadIdx.x == 0) pdl_launch_dependents();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // the QKV GEMM has finished

    // item = (sequence, head); rounds of A6_G query tiles; every role walks the same items / rounds / tiles
    auto item_info = [&](int item, int& t0, int& head) {
        const int seq = item / heads;
        head = item - seq * heads;
        t0 = __ldg(cu_seqlens + seq);
        return __ldg(cu_seqlens + seq + 1) - t0;
    };

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int st = 0;
            uint32_t ph = 0, rounds = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int t0, head;
                const int s = item_info(item, t0, head);
                if (s <= 0) continue;
                const int nq = (s + A6_QT - 1) / A6_QT, nk = (s + A6_KT - 1) / A6_KT;
                for (int q0 = 0; q0 < nq; q0 += A6_G, ++rounds) {
                    const int act = nq - q0 < A6_G ? nq - q0 : A6_G;
                    mbar_wait(bar(B6_QFREE), (rounds & 1) ^ 1);  // the previous round's S MMAs have read their Q
                    mbar_arrive_expect_tx(bar(B6_QFULL), act * A6_Q_BYTES);
                    for (int g = 0; g < act; ++g)
                        tma_load_2d(sQ + g * A6_Q_BYTES, &tmapQ, bar(B6_QFULL), head * A6_D, t0 + (q0 + g) * A6_QT);
                    for (int pass = 0; pass < 2; ++pass) {
                        for (int kt = 0; kt < nk; ++kt) {
                            mbar_wait(bar(B6_EMPTY + st), ph ^ 1);
                            const uint32_t dst = sR + st * A6_STAGE_BYTES;
                            mbar_arrive_expect_tx(bar(B6_FULL + st), pass ? A6_STAGE_BYTES : A6_KV_BYTES);
                            tma_load_2d(dst, &tmapKV, bar(B6_FULL + st), H + head * A6_D, t0 + kt * A6_KT);
                            if (pass) tma_load_2d(dst + A6_KV_BYTES, &tmapKV, bar(B6_FULL + st), 2 * H + head * A6_D, t0 + kt * A6_KT);
                            if (++st == A6_NST) {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== S = Q K^T issuer =====
        if (elect_one()) {
            constexpr uint32_t idesc_s = idesc64(A6_QT, A6_KT, 1, 0);  // bf16 x bf16, both K-major
            int st = 0;
            uint32_t ph = 0, rounds = 0;
            uint32_t su[A6_G] = {0, 0, 0, 0};  // S tiles issued per group (barrier parities)
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int t0, head;
                const int s = item_info(item, t0, head);
                if (s <= 0) continue;
                const int nq = (s + A6_QT - 1) / A6_QT, nk = (s + A6_KT - 1) / A6_KT;
                for (int q0 = 0; q0 < nq; q0 += A6_G, ++rounds) {
                    const int act = nq - q0 < A6_G ? nq - q0 : A6_G;
                    mbar_wait(bar(B6_QFULL), rounds & 1);
                    for (int pass = 0; pass < 2; ++pass) {
                        for (int kt = 0; kt < nk; ++kt) {
                            mbar_wait(bar(B6_FULL + st), ph);
                            tc_fence_after();
                            const uint32_t sK = sR + st * A6_STAGE_BYTES;
#pragma unroll
                            for (int g = 0; g < A6_G; ++g) {
                                if (g >= act) break;
                                mbar_wait(bar(B6_SFREE + g), (su[g] & 1) ^ 1);  // the group has read its previous S
                                tc_fence_after();
                                const uint32_t s_tmem = tmem_base + g * 128;
#pragma unroll
                                for (int k = 0; k < A6_D / 16; ++k)
                                    umma_bf16(s_tmem, umma_desc_sw128(sQ + g * A6_Q_BYTES) + 2 * k, umma_desc_sw128(sK) + 2 * k,
                                              idesc_s, k != 0 ? 1u : 0u);
                                umma_commit(bar(B6_SFULL + g));
                                ++su[g];
                            }
                            if (pass == 0) umma_commit(bar(B6_EMPTY + st));  // pass B: the P.V issuer frees the slot
                            if (++st == A6_NST) {
                                st = 0;
                                ph ^= 1;
                            }
                        }
                    }
                    umma_commit(bar(B6_QFREE));  // every S MMA of the round has read its Q tile
                }
            }
        }
    } else if (warp == 2) {
        // ===== O += P V issuer =====
        if (elect_one()) {
            constexpr uint32_t idesc_o = idesc64(A6_QT, A6_D, 0, 1);  // fp16 x fp16, B (= V) MN-major
            int st = 0;
            uint32_t ph = 0;
            uint32_t pu[A6_G] = {0, 0, 0, 0}, ou[A6_G] = {0, 0, 0, 0};
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int t0, head;
                const int s = item_info(item, t0, head);
                if (s <= 0) continue;
                const int nq = (s + A6_QT - 1) / A6_QT, nk = (s + A6_KT - 1) / A6_KT;
                for (int q0 = 0; q0 < nq; q0 += A6_G) {
                    const int act = nq - q0 < A6_G ? nq - q0 : A6_G;
                    // pass A uses nk ring slots this warp has nothing to do with
                    for (int kt = 0; kt < nk; ++kt)
                        if (++st == A6_NST) {
                            st = 0;
                            ph ^= 1;
                        }
                    for (int kt = 0; kt < nk; ++kt) {
                        mbar_wait(bar(B6_FULL + st), ph);  // V of this key tile has landed
                        tc_fence_after();
                        const uint32_t sV = sR + st * A6_STAGE_BYTES + A6_KV_BYTES;
#pragma unroll
                        for (int g = 0; g < A6_G; ++g) {
                            if (g >= act) break;
                            mbar_wait(bar(B6_PREADY + g), pu[g] & 1);
                            if (kt == 0) mbar_wait(bar(B6_OFREE + g), (ou[g] & 1) ^ 1);  // previous O has been collected
                            tc_fence_after();
                            const uint32_t o_tmem = tmem_base + g * 128 + 64;
#pragma unroll
                            for (int k = 0; k < A6_KT / 16; ++k)  // 16 keys per step: 32 bytes along P's row, 16 rows of V
                                umma_bf16(o_tmem, umma_desc_sw128(sP + g * A6_P_BYTES) + 2 * k, umma_desc_sw128(sV + k * 2048),
                                          idesc_o, (kt | k) != 0 ? 1u : 0u);
                            umma_commit(bar(B6_PFREE + g));  // the P tile may be rewritten
                            ++pu[g];
                            if (kt == nk - 1) {
                                umma_commit(bar(B6_OFULL + g));
                                ++ou[g];
                            }
                        }
                        umma_commit(bar(B6_EMPTY + st));  // K (read by completed S MMAs) and V of this slot are free
                        if (++st == A6_NST) {
                            st = 0;
                            ph ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp >= 4) {
        // ===== softmax groups: group g = query tile q0 + g of the round; thread = query row = TMEM lane =====
        const int g = (warp - 4) >> 2;
        const int quad = warp & 3;
        const int row = quad * 32 + lane;  // row inside the query tile
        const uint32_t tlane = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + g * 128;
        const uint32_t prow = sP + g * A6_P_BYTES + static_cast<uint32_t>(row) * 128;  // this row of the P tile
        const uint32_t swz = static_cast<uint32_t>(row & 7);
        uint32_t su = 0, pu = 0, ou = 0;  // S tiles read / P tiles written / O tiles collected (barrier parities)
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int t0, head;
            const int s = item_info(item, t0, head);
            if (s <= 0) continue;
            const int nq = (s + A6_QT - 1) / A6_QT, nk = (s + A6_KT - 1) / A6_KT;
            for (int q0 = 0; q0 < nq; q0 += A6_G) {
                if (q0 + g >= nq) continue;  // this group has no query tile in this round (warp-uniform)
                // BIAS: shared-memory address of this thread's bias entry for key 0 of key tile kt (see A6_REL_MAXHEADS)
                auto rel_row = [&](int kt) {
                    const int first = kt * A6_KT - ((q0 + g) * A6_QT + row) + ATT_REL_HALF;
                    return sT + static_cast<uint32_t>(head * ATT_REL_W + min(max(first, 0), ATT_REL_W - A6_KT)) * 4u;
                };
                // ---- pass A: the row maximum over all keys ----
                float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
                for (int kt = 0; kt < nk; ++kt, ++su) {
                    mbar_wait(bar(B6_SFULL + g), su & 1);
                    tc_fence_after();
                    uint32_t va[32], vb[32];
                    tmem_ld_32x32(tlane, va);
                    tmem_ld_32x32(tlane + 32, vb);
                    tmem_ld_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(B6_SFREE + g));
                    const int kleft = s - kt * A6_KT;  // keys of this tile that exist (the rest belong to other sequences)
                    if constexpr (BIAS) {
                        // scores in the log2 domain with the bias added: the maximum is taken over what is exponentiated
                        const uint32_t tb = rel_row(kt);
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            const float xa = fmaf(__uint_as_float(va[c]), scale_log2e, lds_f32(tb + 4 * c));
                            const float xb = fmaf(__uint_as_float(vb[c]), scale_log2e, lds_f32(tb + 4 * (c + 32)));
                            m4[c & 3] = fmaxf(m4[c & 3], c < kleft ? xa : -INFINITY);
                            m4[c & 3] = fmaxf(m4[c & 3], c + 32 < kleft ? xb : -INFINITY);
                        }
                    } else if (kleft >= A6_KT) {
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            m4[c & 3] = fmaxf(m4[c & 3], __uint_as_float(va[c]));
                            m4[c & 3] = fmaxf(m4[c & 3], __uint_as_float(vb[c]));
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 32; ++c) {
                            m4[c & 3] = fmaxf(m4[c & 3], c < kleft ? __uint_as_float(va[c]) : -INFINITY);
                            m4[c & 3] = fmaxf(m4[c & 3], c + 32 < kleft ? __uint_as_float(vb[c]) : -INFINITY);
                        }
                    }
                }
                const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
                const float nms = BIAS ? -mx : -mx * scale_log2e;
                // ---- pass B: P = exp2(S * c - max * c) -> fp16 -> K-major SWIZZLE_128B tile (row r at r * 128,
                //      16-byte chunk c at (c ^ (r & 7)) * 16); the P.V issuer accumulates O in TMEM ----
                float s4[4] = {0.f, 0.f, 0.f, 0.f};
                for (int kt = 0; kt < nk; ++kt, ++su, ++pu) {
                    mbar_wait(bar(B6_SFULL + g), su & 1);
                    tc_fence_after();
                    uint32_t va[32], vb[32];
                    tmem_ld_32x32(tlane, va);
                    tmem_ld_32x32(tlane + 32, vb);
                    tmem_ld_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(B6_SFREE + g));
                    mbar_wait(bar(B6_PFREE + g), (pu & 1) ^ 1);  // the previous P.V has read the P tile
                    const int kleft = s - kt * A6_KT;
                    uint32_t tb = 0;
                    if constexpr (BIAS) tb = rel_row(kt);
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        uint32_t(&v)[32] = half ? vb : va;
#pragma unroll
                        for (int c8 = 0; c8 < 4; ++c8) {
                            float pe[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                float off = nms;
                                if constexpr (BIAS) off += lds_f32(tb + 4 * (half * 32 + c8 * 8 + e));
                                const float x = fmaf(__uint_as_float(v[c8 * 8 + e]), scale_log2e, off);
                                pe[e] = ex2_approx(x);
                                if (kleft < A6_KT) pe[e] = (half * 32 + c8 * 8 + e < kleft) ? pe[e] : 0.f;
                                s4[e & 3] += pe[e];
                            }
                            const uint32_t chunk = static_cast<uint32_t>(half * 4 + c8) ^ swz;
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow + chunk * 16),
                                         "r"(pack_f16x2(pe[0], pe[1])), "r"(pack_f16x2(pe[2], pe[3])),
                                         "r"(pack_f16x2(pe[4], pe[5])), "r"(pack_f16x2(pe[6], pe[7]))
                                         : "memory");
                        }
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(B6_PREADY + g));
                }
                // ---- O / sum -> bf16 -> ctx ----
                const float inv = 1.0f / ((s4[0] + s4[1]) + (s4[2] + s4[3]));
                mbar_wait(bar(B6_OFULL + g), ou & 1);
                ++ou;
                tc_fence_after();
                uint32_t oa[32], ob[32];
                tmem_ld_32x32(tlane + 64, oa);
                tmem_ld_32x32(tlane + 96, ob);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B6_OFREE + g));
                const int qrow = (q0 + g) * A6_QT + row;
                if (qrow < s) {  // rows past the sequence's end belong to other sequences
                    uint4* dst = reinterpret_cast<uint4*>(ctx + static_cast<size_t>(t0 + qrow) * H + head * A6_D);
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        uint32_t(&ov)[32] = half ? ob : oa;
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            uint4 o;
                            o.x = pack_bf16x2(__uint_as_float(ov[8 * c]) * inv, __uint_as_float(ov[8 * c + 1]) * inv);
                            o.y = pack_bf16x2(__uint_as_float(ov[8 * c + 2]) * inv, __uint_as_float(ov[8 * c + 3]) * inv);
                            o.z = pack_bf16x2(__uint_as_float(ov[8 * c + 4]) * inv, __uint_as_float(ov[8 * c + 5]) * inv);
                            o.w = pack_bf16x2(__uint_as_float(ov[8 * c + 6]) * inv, __uint_as_float(ov[8 * c + 7]) * inv);
                            dst[half * 4 + c] = o;
                        }
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

// MPNetEncoder.compute_position_bias / relative_position_bucket (HuggingFace transformers, the graph DJL runs for an MPNet
// model; call site AbstractHuggingFaceEmbeddingService.java:194-196): bias[h][q][k] = W[bucket(k - q)][h], the same for
// every layer.  Half of the buckets per direction; the first half of those exact, the rest logarithmic up to max_distance.
void attention_rel_table(const float* w, int buckets, int max_distance, int heads, std::vector<float>* out) {
    out->assign(static_cast<size_t>(heads) * ATT_REL_W, 0.f);
    const int nb = buckets / 2, max_exact = nb / 2;
    for (int r = -ATT_REL_HALF; r <= ATT_REL_HALF; ++r) {
        int n = -r, ret = 0;
        if (n < 0) {
            ret = nb;
            n = -n;
        }
        if (n < max_exact) {
            ret += n;
        } else {  // float32 arithmetic, truncation toward zero: as torch evaluates it
            const float v = logf(static_cast<float>(n) / static_cast<float>(max_exact)) /
                            static_cast<float>(log(static_cast<double>(max_distance) / max_exact)) * static_cast<float>(nb - max_exact);
            ret += std::min(max_exact + static_cast<int>(v), nb - 1);
        }
        for (int h = 0; h < heads; ++h)
            (*out)[static_cast<size_t>(h) * ATT_REL_W + (r + ATT_REL_HALF)] = w[static_cast<size_t>(ret) * heads + h] * 1.4426950408889634f;
    }
}

bool attention_tc64_supported(int H, int heads) {
    static const bool off = getenv("LSC_ATTN_LEGACY") != nullptr;  // debug: keep the mma.sync kernels
    return !off && heads > 0 && H / heads == A6_D && H % heads == 0;
}

// qkv: [T, 3H] with Q, K bf16 and V fp16; d_head == 64, any sequence length.
int attention_tc64_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int T, int H, int heads,
                          __nv_bfloat16* ctx, int num_sms, cudaStream_t stream, const float* rel_tbl) {
    if (n_seq <= 0 || T <= 0) return 0;
    if (H / heads != A6_D) return fail(LSC_E_UNSUPPORTED, "attention_tc64: needs d_head 64");
    if (rel_tbl != nullptr && heads > A6_REL_MAXHEADS)
        return fail(LSC_E_UNSUPPORTED, "attention_tc64: too many heads for the relative-position table in shared memory");
    CUtensorMap tmQ, tmKV;
    LSC_TRY(make_tmap_bf16_box(&tmQ, qkv, static_cast<uint64_t>(T), static_cast<uint64_t>(3) * H, A6_QT, A6_D, 128));
    LSC_TRY(make_tmap_bf16_box(&tmKV, qkv, static_cast<uint64_t>(T), static_cast<uint64_t>(3) * H, A6_KT, A6_D, 128));
    static bool configured[64] = {};
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(attention_tc64_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, A6_SMEM));
        LSC_CUDA(cudaFuncSetAttribute(attention_tc64_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      A6_SMEM + A6_REL_MAXHEADS * ATT_REL_W * 4));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const long long n_items = static_cast<long long>(heads) * n_seq;
    const int grid = static_cast<int>(n_items < num_sms ? n_items : num_sms);
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(A6_D));
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(grid);
    lc.blockDim = dim3(A6_THREADS);
    lc.dynamicSmemBytes = A6_SMEM + (rel_tbl != nullptr ? heads * ATT_REL_W * 4 : 0);
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see pdl_wait() in ptx.cuh
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    if (rel_tbl != nullptr)
        LSC_CUDA(cudaLaunchKernelEx(&lc, attention_tc64_kernel<true>, tmQ, tmKV, cu_seqlens, H, heads, static_cast<int>(n_items), ctx,
                                    scale_log2e, rel_tbl));
    else
        LSC_CUDA(cudaLaunchKernelEx(&lc, attention_tc64_kernel<false>, tmQ, tmKV, cu_seqlens, H, heads, static_cast<int>(n_items), ctx,
                                    scale_log2e, rel_tbl));
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
