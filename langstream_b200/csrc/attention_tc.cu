// tcgen05 attention for short sequences (s <= 128, d_head = 32): ctx = softmax(Q K^T / sqrt(d)) V per
// (sequence, head).  Same role as attention.cu (libtorch's scaled-dot-product attention inside the DJL
// forward pass, AbstractHuggingFaceEmbeddingService.java:194-196); the mma.sync kernel there is bound by
// the legacy HMMA rate, this one puts both contractions on the 5th-generation tensor cores and leaves the
// exponentials as the only per-element work.
//
// Persistent, one CTA (320 threads) per SM, items = (sequence, head) pairs:
//   warp 0      TMA: Q, K, V tiles [128 rows x 32] (64-byte rows, SWIZZLE_64B) straight out of the packed
//               qkv [T, 3H] tensor, 3-deep ring.  Rows past the sequence belong to the next sequences (or are
//               zero-filled past T): keys are masked in the softmax, surplus query rows are never stored.
//   warp 1      issues S = Q K^T (tcgen05.mma M=128, N=128, K=32, bf16 x bf16) into TMEM
//   warp 14     issues O = P V (M=128, N=32, K=128, fp16 x fp16, V consumed MN-major exactly as it lies in
//               memory); a separate warp so that waiting for a P tile never delays the next S
//   warps 2-13  three softmax groups of four warps taking items round-robin; thread = query row = TMEM lane, so max
//               and sum are thread-local.  P is written as fp16 into a SWIZZLE_128B K-major smem tile (the A
//               operand of the second MMA); O comes back through tcgen05.ld, is scaled by 1/sum and stored.
//               Each group is software-pipelined by one item: it collects O of its previous item after the
//               exponentials of the current one, so the P.V latency is off its critical path.
// TMEM: AT_G = 3 slots of 160 columns (S: 128, O: 32), one per softmax group.  mbarrier pipelines: load full/empty, S written / S read,
// P written, O written / O read.
#include <cstdlib>

#include "common.h"
#include "ptx.cuh"

namespace lsc {

namespace {

constexpr int AT_G = 3;                         // softmax groups = items in flight = TMEM slots
constexpr int AT_THREADS = 64 + AT_G * 128 + 32;  // TMA, S-MMA, AT_G x 4 softmax warps, PV-MMA
constexpr int AT_ROWS = 128;
constexpr int AT_D = 32;
constexpr int AT_NST = 4;
constexpr int AT_TILE_BYTES = AT_ROWS * AT_D * 2;      // 8 KB: one of Q / K / V
constexpr int AT_STAGE_BYTES = 3 * AT_TILE_BYTES;      // 24 KB
constexpr int AT_P_BYTES = AT_ROWS * 128 * 2;          // 32 KB: P tile, fp16 [128 x 128]
constexpr int AT_SLOT_COLS = 160;
constexpr int AT_OSTAGE_BYTES = 2048;                  // per softmax warp: a 32 x 32 bf16 block of ctx, source of a TMA store
constexpr int AT_SMEM = AT_NST * AT_STAGE_BYTES + AT_G * AT_P_BYTES + AT_G * 4 * AT_OSTAGE_BYTES + 512 + 1024;

// K-major operand with 64-byte rows written by TMA SWIZZLE_64B: 8-row atoms 512 B apart.
__device__ __forceinline__ uint64_t desc_sw64(uint32_t addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>(1) << 16;             // LBO (unused: one atom along the leading dimension)
    d |= static_cast<uint64_t>(512 >> 4) << 32;      // SBO = 512 B
    d |= static_cast<uint64_t>(1) << 46;             // version 1
    d |= static_cast<uint64_t>(4) << 61;             // SWIZZLE_64B
    return d;
}
// instruction descriptor: fp32 accumulate, formats/majors as given (format: 0 = f16, 1 = bf16)
__host__ __device__ constexpr uint32_t idesc(uint32_t M, uint32_t N, uint32_t fmt, uint32_t b_mn_major) {
    return (1u << 4) | (fmt << 7) | (fmt << 10) | (b_mn_major << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

__device__ __forceinline__ float ex2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// UNIFORM: every sequence has exactly AT_ROWS tokens (the headline shape): nothing is looked up, nothing is masked
template <bool UNIFORM>
__global__ void __launch_bounds__(AT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmapQKV, const __grid_constant__ CUtensorMap tmapCtx,
                    const int32_t* __restrict__ cu_seqlens, int H,
                    int heads, int n_items, __nv_bfloat16* __restrict__ ctx, float scale_log2e, int uniform_len) {
    extern __shared__ uint8_t smem_raw[];
    // aligned by pointer arithmetic on the shared array (an integer round trip would turn every access
    // through it into a generic-address one)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t sP = smem_base + AT_NST * AT_STAGE_BYTES;  // two 32 KB P tiles, 1024-byte aligned
    const uint32_t sO = sP + AT_G * AT_P_BYTES;  // ctx staging blocks, one per softmax warp
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AT_NST * AT_STAGE_BYTES + AT_G * AT_P_BYTES + AT_G * 4 * AT_OSTAGE_BYTES);
    const uint32_t bar_full = smem_u32(bars);            // [NST]
    const uint32_t bar_empty = bar_full + 8 * AT_NST;    // [NST]
    const uint32_t bar_sfull = bar_empty + 8 * AT_NST;     // [G] S accumulator written
    const uint32_t bar_sfree = bar_sfull + 8 * AT_G;       // [G] S accumulator no longer read
    const uint32_t bar_pready = bar_sfree + 8 * AT_G;      // [G] P tile written to smem
    const uint32_t bar_ofull = bar_pready + 8 * AT_G;      // [G] O accumulator written (P tile, V no longer read)
    const uint32_t bar_ofree = bar_ofull + 8 * AT_G;       // [G] O accumulator read into registers
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + 2 * AT_NST + 5 * AT_G);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapQKV);
        tma_prefetch_desc(&tmapCtx);
        for (int s = 0; s < AT_NST; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int j = 0; j < AT_G; ++j) {
            mbar_init(bar_sfull + 8 * j, 1);
            mbar_init(bar_sfree + 8 * j, 4);   // one arrive per softmax warp of the group
            mbar_init(bar_pready + 8 * j, 4);
            mbar_init(bar_ofull + 8 * j, 1);
            mbar_init(bar_ofree + 8 * j, 4);
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

    // uniform_len > 0: every sequence has that many tokens (the headline shape, 128 each) - nothing to look up
    auto item_len = [&](int item, int& t0) {
        const int seq = item / heads;
        if (UNIFORM) {
            t0 = seq * AT_ROWS;
            return AT_ROWS;
        }
        if (uniform_len > 0) {
            t0 = seq * uniform_len;
            return uniform_len;
        }
        t0 = __ldg(cu_seqlens + seq);
        return __ldg(cu_seqlens + seq + 1) - t0;
    };

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int i = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int t0;
                if (item_len(item, t0) <= 0) continue;
                const int head = item % heads;
                const int st = i % AT_NST;
                const uint32_t ph = (i / AT_NST) & 1;
                mbar_wait(bar_empty + 8 * st, ph ^ 1);
                const uint32_t dst = smem_base + st * AT_STAGE_BYTES;
                mbar_arrive_expect_tx(bar_full + 8 * st, AT_STAGE_BYTES);
                tma_load_2d(dst, &tmapQKV, bar_full + 8 * st, head * AT_D, t0);
                tma_load_2d(dst + AT_TILE_BYTES, &tmapQKV, bar_full + 8 * st, H + head * AT_D, t0);
                tma_load_2d(dst + 2 * AT_TILE_BYTES, &tmapQKV, bar_full + 8 * st, 2 * H + head * AT_D, t0);
                ++i;
            }
        }
    } else if (warp == 1) {
        // ===== S = Q K^T issuer =====
        if (elect_one()) {
            constexpr uint32_t idesc_s = idesc(128, 128, 1, 0);  // bf16 x bf16, both K-major
            int i = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int t0;
                if (item_len(item, t0) <= 0) continue;
                const int j = i % AT_G, st = i % AT_NST;
                mbar_wait(bar_full + 8 * st, (i / AT_NST) & 1);
                mbar_wait(bar_sfree + 8 * j, ((i / AT_G) & 1) ^ 1);
                tc_fence_after();
                const uint32_t sQ = smem_base + st * AT_STAGE_BYTES;
                const uint32_t sK = sQ + AT_TILE_BYTES;
                const uint32_t s_tmem = tmem_base + j * AT_SLOT_COLS;
#pragma unroll
                for (int k = 0; k < AT_D / 16; ++k)
                    umma_bf16(s_tmem, desc_sw64(sQ + k * 32), desc_sw64(sK + k * 32), idesc_s, k != 0 ? 1u : 0u);
                umma_commit(bar_sfull + 8 * j);
                ++i;
            }
        }
    } else if (warp == 2 + 4 * AT_G) {
        // ===== O = P V issuer (its own warp: never blocks the S issuer) =====
        if (elect_one()) {
            constexpr uint32_t idesc_o = idesc(128, 32, 0, 1);   // fp16 x fp16, B (= V) MN-major
            int i = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int t0;
                if (item_len(item, t0) <= 0) continue;
                const int j = i % AT_G, st = i % AT_NST;
                mbar_wait(bar_full + 8 * st, (i / AT_NST) & 1);           // V has landed (long ago)
                mbar_wait(bar_pready + 8 * j, (i / AT_G) & 1);
                mbar_wait(bar_ofree + 8 * j, ((i / AT_G) & 1) ^ 1);
                tc_fence_after();
                const uint32_t sV = smem_base + st * AT_STAGE_BYTES + 2 * AT_TILE_BYTES;
                const uint32_t o_tmem = tmem_base + j * AT_SLOT_COLS + 128;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    // A = P: SW128 K-major, 64 fp16 per 128-byte row-chunk, 16 KB per 64-wide k block
                    const uint64_t da = umma_desc_sw128(sP + j * AT_P_BYTES + (k >> 2) * 16384 + (k & 3) * 32);
                    // B = V [keys x d] as stored: MN-major, 16 keys = two 8-row atoms of 512 B
                    const uint64_t db = desc_sw64(sV + k * 1024);
                    umma_bf16(o_tmem, da, db, idesc_o, k != 0 ? 1u : 0u);
                }
                umma_commit(bar_ofull + 8 * j);
                umma_commit(bar_empty + 8 * st);  // Q, K, V of this item are no longer read
                ++i;
            }
        }
    } else {
        // ===== softmax groups (software-pipelined by one item) =====
        const int grp = (warp - 2) >> 2;       // this group takes local items i with i % AT_G == grp
        const int quad = warp & 3;             // TMEM lane quadrant of this warp
        const int row = quad * 32 + lane;      // query row owned by this thread
        const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
        const int j = grp;
        const uint32_t s_tmem = tmem_base + lane_off + j * AT_SLOT_COLS;
        const uint32_t prow_addr = sP + j * AT_P_BYTES + row * 128;
        const uint32_t sw = row & 7;
        // state of the previous item of this group, whose O is still to be collected
        bool have_prev = false;
        int prev_t0 = 0, prev_s = 0, prev_head = 0;
        float prev_sum = 1.f;
        uint32_t uses = 0;  // how many items this group has started (slot-use counter -> barrier parity)

        auto collect = [&](uint32_t use) {  // O of slot use `use`: TMEM -> registers -> scaled bf16 row of ctx
            mbar_wait(bar_ofull + 8 * j, use & 1);
            tc_fence_after();
            uint32_t ov[32];
            tmem_ld_32x32(s_tmem + 128, ov);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_ofree + 8 * j);
            const float inv = 1.0f / prev_sum;
            if (UNIFORM || quad * 32 + 32 <= prev_s) {
                // all 32 rows of this warp belong to the sequence: the 32 x 32 block goes out as ONE TMA store from a
                // swizzled staging block (per-thread 16-byte stores to rows 768 bytes apart cost a fifth of the kernel:
                // every store instruction touches 32 different lines)
                const uint32_t stage = sO + static_cast<uint32_t>(warp - 2) * AT_OSTAGE_BYTES;
                if (lane == 0) tma_store_wait_read<0>();  // the previous store has read the block
                __syncwarp();
                const uint32_t sw64 = (static_cast<uint32_t>(lane) >> 1) & 3;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stage + lane * 64 + ((static_cast<uint32_t>(c) ^ sw64) * 16)),
                                 "r"(pack_bf16x2(__uint_as_float(ov[8 * c]) * inv, __uint_as_float(ov[8 * c + 1]) * inv)),
                                 "r"(pack_bf16x2(__uint_as_float(ov[8 * c + 2]) * inv, __uint_as_float(ov[8 * c + 3]) * inv)),
                                 "r"(pack_bf16x2(__uint_as_float(ov[8 * c + 4]) * inv, __uint_as_float(ov[8 * c + 5]) * inv)),
                                 "r"(pack_bf16x2(__uint_as_float(ov[8 * c + 6]) * inv, __uint_as_float(ov[8 * c + 7]) * inv))
                                 : "memory");
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&tmapCtx, stage, prev_head * AT_D, prev_t0 + quad * 32);
                    tma_store_commit();
                }
            } else if (row < prev_s) {
                uint4* dst = reinterpret_cast<uint4*>(ctx + static_cast<size_t>(prev_t0 + row) * H + prev_head * AT_D);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 o;
                    o.x = pack_bf16x2(__uint_as_float(ov[8 * c]) * inv, __uint_as_float(ov[8 * c + 1]) * inv);
                    o.y = pack_bf16x2(__uint_as_float(ov[8 * c + 2]) * inv, __uint_as_float(ov[8 * c + 3]) * inv);
                    o.z = pack_bf16x2(__uint_as_float(ov[8 * c + 4]) * inv, __uint_as_float(ov[8 * c + 5]) * inv);
                    o.w = pack_bf16x2(__uint_as_float(ov[8 * c + 6]) * inv, __uint_as_float(ov[8 * c + 7]) * inv);
                    dst[c] = o;
                }
            }
        };

        int i = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int t0;
            const int s = item_len(item, t0);
            if (s <= 0) continue;
            if ((i % AT_G) != grp) {
                ++i;
                continue;
            }
            ++i;
            const uint32_t use = uses++;
            mbar_wait(bar_sfull + 8 * j, use & 1);
            tc_fence_after();
            // ---- pass 1 over the score row: the row maximum.  Full-length sequences (the common case) take the
            //      mask-free path.  The TMEM reads are double-buffered: a 32-column read occupies the warp's TMEM port for
            //      ~64 cycles, longer than the arithmetic on the previous part takes. ----
            const bool full = UNIFORM || s >= AT_ROWS;  // warp-uniform
            float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
            uint32_t va[32], vb[32];
            auto max_part = [&](const uint32_t (&v)[32], int part) {
                if (full) {
#pragma unroll
                    for (int c = 0; c < 32; ++c) m4[c & 3] = fmaxf(m4[c & 3], __uint_as_float(v[c]));
                } else {
#pragma unroll
                    for (int c = 0; c < 32; ++c)  // keys past the sequence are other sequences' rows
                        m4[c & 3] = fmaxf(m4[c & 3], (part * 32 + c < s) ? __uint_as_float(v[c]) : -INFINITY);
                }
            };
            tmem_ld_32x32(s_tmem, va);
            tmem_ld_wait();
            tmem_ld_32x32(s_tmem + 32, vb);
            max_part(va, 0);
            tmem_ld_wait();
            tmem_ld_32x32(s_tmem + 64, va);
            max_part(vb, 1);
            tmem_ld_wait();
            tmem_ld_32x32(s_tmem + 96, vb);
            max_part(va, 2);
            tmem_ld_wait();
            tmem_ld_32x32(s_tmem, va);  // first part of pass 2, in flight across the collection below
            max_part(vb, 3);
            const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
            const float nms = -mx * scale_log2e;
            // the previous item's P.V has finished by now: collect its O; this also guarantees that the
            // tensor core no longer reads this group's P tile
            if (have_prev) collect(use - 1);
            // ---- pass 2: P = exp2(S*scale - max*scale) -> fp16 -> K-major SW128 A tile of the P.V MMA:
            //      k block b (64 keys) at b*16 KB, row r at r*128, 16-byte chunk c at (c ^ (r & 7)) * 16 ----
            float s4[4] = {0.f, 0.f, 0.f, 0.f};
            auto exp_part = [&](const uint32_t (&v)[32], int part) {
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    float pe[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float x = fmaf(__uint_as_float(v[c8 * 8 + e]), scale_log2e, nms);
                        pe[e] = ex2_fast(x);
                        if (!full) pe[e] = (part * 32 + c8 * 8 + e < s) ? pe[e] : 0.f;
                        s4[e & 3] += pe[e];
                    }
                    const int chunk = (part & 1) * 4 + c8;
                    const uint32_t dst = prow_addr + (part >> 1) * 16384 + ((chunk ^ sw) * 16);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pack_f16x2(pe[0], pe[1])),
                                 "r"(pack_f16x2(pe[2], pe[3])), "r"(pack_f16x2(pe[4], pe[5])),
                                 "r"(pack_f16x2(pe[6], pe[7]))
                                 : "memory");
                }
            };
            tmem_ld_wait();
            tmem_ld_32x32(s_tmem + 32, vb);
            exp_part(va, 0);
            tmem_ld_wait();
            tmem_ld_32x32(s_tmem + 64, va);
            exp_part(vb, 1);
            tmem_ld_wait();
            tmem_ld_32x32(s_tmem + 96, vb);
            exp_part(va, 2);
            tmem_ld_wait();
            // last read of S: the slot's S region may be overwritten
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_sfree + 8 * j);
            exp_part(vb, 3);
            const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_pready + 8 * j);
            have_prev = true;
            prev_t0 = t0;
            prev_s = s;
            prev_head = item % heads;
            prev_sum = sum;
        }
        if (have_prev) collect(uses - 1);
        if (lane == 0) tma_store_wait_all<0>();  // ctx writes complete before the kernel ends
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

// qkv: [T, 3H] with Q, K bf16 and V fp16.  Only d_head == 32 and max_seqlen <= 128.
int attention_tc_launch(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int n_seq, int T, int H, int heads,
                        int max_seqlen, __nv_bfloat16* ctx, int num_sms, cudaStream_t stream) {
    // sequences of at most max_seqlen tokens that add up to n_seq * max_seqlen all have max_seqlen tokens
    static const bool no_uniform = getenv("LSC_ATTN_NO_UNIFORM") != nullptr;  // debug: always read cu_seqlens
    const int uniform_len = (!no_uniform && static_cast<long long>(n_seq) * max_seqlen == T) ? max_seqlen : 0;
    if (n_seq <= 0 || T <= 0) return 0;
    if (H / heads != AT_D || max_seqlen > AT_ROWS) return fail(LSC_E_UNSUPPORTED, "attention_tc: needs d_head 32 and s <= 128");
    CUtensorMap tm, tmc;
    LSC_TRY(make_tmap_bf16_box(&tm, qkv, static_cast<uint64_t>(T), static_cast<uint64_t>(3) * H, AT_ROWS, AT_D, 64));
    LSC_TRY(make_tmap_bf16_store32(&tmc, ctx, static_cast<uint64_t>(T), static_cast<uint64_t>(H)));
    static bool configured[64] = {};
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(attention_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        LSC_CUDA(cudaFuncSetAttribute(attention_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const long long n_items = static_cast<long long>(heads) * n_seq;
    const int grid = static_cast<int>(n_items < num_sms ? n_items : num_sms);
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(AT_D));
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(grid);
    lc.blockDim = dim3(AT_THREADS);
    lc.dynamicSmemBytes = AT_SMEM;
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // see pdl_wait() in ptx.cuh
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    if (uniform_len == AT_ROWS)
        LSC_CUDA(cudaLaunchKernelEx(&lc, attention_tc_kernel<true>, tm, tmc, cu_seqlens, H, heads, static_cast<int>(n_items), ctx,
                                    scale_log2e, uniform_len));
    else
        LSC_CUDA(cudaLaunchKernelEx(&lc, attention_tc_kernel<false>, tm, tmc, cu_seqlens, H, heads, static_cast<int>(n_items), ctx,
                                    scale_log2e, uniform_len));
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
