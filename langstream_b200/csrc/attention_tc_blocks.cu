// tcgen05 attention for RAGGED batches of short sequences (s <= 128, d_head = 32), third form: attention tiles are
// composed of 32-ROW BLOCKS.  A sequence of s tokens takes np = ceil(s / 32) consecutive blocks of a 128-row tile and a
// tile holds as many whole sequences as its four blocks take - sequences from ANYWHERE in the batch, because every block
// is fetched by its own TMA boxes (attention_tc_ragged.cu showed that a SWIZZLE_64B tile is four independent 2 KB
// blocks).  attention_tc_ragged.cu sizes the work of an item by its sequence but still spends one item, i.e. one trip
// through the load -> S -> softmax -> P.V -> store chain, per (sequence, head): 12,288 trips for 1024 sequences, and
// the chain, not the arithmetic, is what the kernel's time is made of.  Here a UniformInt[16,128] batch needs 0.69
// tiles per sequence and nearly every tile is full.
//   * tile list (attention_blocks_kernel, one CTA, once per pass): stable counting sort of the sequences by np, then
//     4 | 3+1 | 2+2 | 2+1+1 | 1+1+1+1 by index arithmetic; one 32-byte record per tile: the first global row of each
//     of its four blocks (-1: unused; used blocks are a prefix) and, per block, where its sequence's keys lie in the
//     tile (first key block, number of key blocks), the sequence length and the block's position in its sequence;
//   * S = Q K^T is ONE MMA over the whole tile (M = 128, N = 32 x used blocks): the blocks off the diagonal are other
//     sequences' scores and are never read;
//   * a softmax warp = one block = one sequence: it reads the np 32-column parts of ITS sequence, masks the tail of the
//     last one, and writes zeros into P for the other sequences' key blocks, so that O = P V over the whole tile
//     (K = 32 x used blocks) adds exact zeros for foreign keys;
//   * a warp whose block is unused only keeps the barrier protocol in step (see attention_tc_ragged.cu).
// A sequence's result does not depend on its tile-mates or on its place in the tile: its scores come from its own Q / K
// rows, foreign keys contribute +0 to sums that are otherwise added in the same order.
// Role layout, barriers and the one-item software pipeline of the softmax groups are those of attention_tc.cu.
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

__global__ void __launch_bounds__(AT_THREADS, 1)
attention_tc_blocks_kernel(const __grid_constant__ CUtensorMap tmapQKV, const __grid_constant__ CUtensorMap tmapCtx,
                    const int32_t* __restrict__ tiles, int H, int heads, __nv_bfloat16* __restrict__ ctx,
                    float scale_log2e) {
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

    // tile list: header (tiles[0] = number of tiles), then one record of 8 ints per tile (see attention_blocks_kernel)
    const int n_items = __ldg(tiles) * heads;
    const int4* const recs = reinterpret_cast<const int4*>(tiles + 8);
    auto used_blocks = [](const int4& r0) { return (r0.x >= 0) + (r0.y >= 0) + (r0.z >= 0) + (r0.w >= 0); };

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int i = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++i) {
                const int tile = item / heads, head = item - tile * heads;
                const int4 r0 = __ldg(recs + 2 * tile);
                const int rows[4] = {r0.x, r0.y, r0.z, r0.w};
                const int nb = used_blocks(r0);
                const int st = i % AT_NST;
                const uint32_t ph = (i / AT_NST) & 1;
                mbar_wait(bar_empty + 8 * st, ph ^ 1);
                const uint32_t dst = smem_base + st * AT_STAGE_BYTES;
                mbar_arrive_expect_tx(bar_full + 8 * st, 3 * 2048 * nb);
#pragma unroll
                for (int p = 0; p < 4; ++p) {  // 32-row boxes, 2 KB each: block p of the tile, wherever its sequence lies
                    if (p >= nb) break;
                    tma_load_2d(dst + p * 2048, &tmapQKV, bar_full + 8 * st, head * AT_D, rows[p]);
                    tma_load_2d(dst + AT_TILE_BYTES + p * 2048, &tmapQKV, bar_full + 8 * st, H + head * AT_D, rows[p]);
                    tma_load_2d(dst + 2 * AT_TILE_BYTES + p * 2048, &tmapQKV, bar_full + 8 * st, 2 * H + head * AT_D, rows[p]);
                }
            }
        }
    } else if (warp == 1) {
        // ===== S = Q K^T issuer =====
        if (elect_one()) {
            int i = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++i) {
                const int nb = used_blocks(__ldg(recs + 2 * (item / heads)));
                const uint32_t idesc_s = idesc(128, static_cast<uint32_t>(32 * nb), 1, 0);  // bf16 x bf16, both K-major; N = the used blocks
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
            }
        }
    } else if (warp == 2 + 4 * AT_G) {
        // ===== O = P V issuer (its own warp: never blocks the S issuer) =====
        if (elect_one()) {
            constexpr uint32_t idesc_o = idesc(128, 32, 0, 1);   // fp16 x fp16, B (= V) MN-major
            int i = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++i) {
                const int nk = used_blocks(__ldg(recs + 2 * (item / heads))) * 2;  // 16-key steps: K = the used blocks
                const int j = i % AT_G, st = i % AT_NST;
                mbar_wait(bar_full + 8 * st, (i / AT_NST) & 1);           // V has landed (long ago)
                mbar_wait(bar_pready + 8 * j, (i / AT_G) & 1);
                mbar_wait(bar_ofree + 8 * j, ((i / AT_G) & 1) ^ 1);
                tc_fence_after();
                const uint32_t sV = smem_base + st * AT_STAGE_BYTES + 2 * AT_TILE_BYTES;
                const uint32_t o_tmem = tmem_base + j * AT_SLOT_COLS + 128;
                for (int k = 0; k < nk; ++k) {
                    // A = P: SW128 K-major, 64 fp16 per 128-byte row-chunk, 16 KB per 64-wide k block
                    const uint64_t da = umma_desc_sw128(sP + j * AT_P_BYTES + (k >> 2) * 16384 + (k & 3) * 32);
                    // B = V [keys x d] as stored: MN-major, 16 keys = two 8-row atoms of 512 B
                    const uint64_t db = desc_sw64(sV + k * 1024);
                    umma_bf16(o_tmem, da, db, idesc_o, k != 0 ? 1u : 0u);
                }
                umma_commit(bar_ofull + 8 * j);
                umma_commit(bar_empty + 8 * st);  // Q, K, V of this item are no longer read
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
        // state of the previous item of this group, whose O is still to be collected: the first global row of this
        // warp's block (-1: the block was unused), how many of its 32 rows belong to the sequence, head, 1 / row sum
        bool have_prev = false;
        int prev_row0 = -1, prev_rows = 0, prev_head = 0;
        float prev_sum = 1.f;
        uint32_t uses = 0;  // how many items this group has started (slot-use counter -> barrier parity)

        auto collect = [&](uint32_t use) {  // O of slot use `use`: TMEM -> registers -> scaled bf16 row of ctx
            mbar_wait(bar_ofull + 8 * j, use & 1);
            tc_fence_after();
            uint32_t ov[32];
            if (prev_row0 >= 0) {  // warp-uniform: otherwise none of this warp's rows is stored
                tmem_ld_32x32(s_tmem + 128, ov);
                tmem_ld_wait();
                tc_fence_before();
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_ofree + 8 * j);
            if (prev_row0 < 0) return;
            const float inv = 1.0f / prev_sum;
            if (prev_rows >= 32) {
                // all 32 rows of this warp belong to the sequence: the 32 x 32 block goes out as ONE TMA store from a
                // swizzled staging block
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
                    tma_store_2d(&tmapCtx, stage, prev_head * AT_D, prev_row0);
                    tma_store_commit();
                }
            } else if (lane < prev_rows) {
                uint4* dst = reinterpret_cast<uint4*>(ctx + static_cast<size_t>(prev_row0 + lane) * H + prev_head * AT_D);
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

        const int32_t* const rec32 = tiles + 8;
        for (int item = blockIdx.x + grp * gridDim.x; item < n_items; item += AT_G * gridDim.x) {
            const int tile = item / heads;
            // this warp's block of the tile: its first global row and (first key block | key blocks << 4 | sequence
            // length << 8 | block within the sequence << 16 | used blocks of the tile << 20); the loads fly during the wait
            const int row0 = __ldg(rec32 + 8 * tile + quad);
            const uint32_t info = static_cast<uint32_t>(__ldg(rec32 + 8 * tile + 4 + quad));
            const uint32_t use = uses++;
            mbar_wait(bar_sfull + 8 * j, use & 1);
            tc_fence_after();
            const bool active = row0 >= 0;  // warp-uniform
            const int kb0 = info & 15, np = (info >> 4) & 15, s = (info >> 8) & 255, blk = (info >> 16) & 15;
            const int nb = (info >> 20) & 15;
            // ---- pass 1 over the sequence's own key blocks: the row maximum; only the last one can hold foreign keys ----
            float nms = 0.f;
            if (active) {
                float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll 1
                for (int p = 0; p < np; ++p) {
                    uint32_t v[32];
                    tmem_ld_32x32(s_tmem + (kb0 + p) * 32, v);
                    tmem_ld_wait();
                    const int lim = s - p * 32;  // keys of this part that belong to the sequence
                    if (lim >= 32) {
#pragma unroll
                        for (int c = 0; c < 32; ++c) m4[c & 3] = fmaxf(m4[c & 3], __uint_as_float(v[c]));
                    } else {
#pragma unroll
                        for (int c = 0; c < 32; ++c)  // the rest of the block: rows of whatever follows the sequence in memory
                            m4[c & 3] = fmaxf(m4[c & 3], (c < lim) ? __uint_as_float(v[c]) : -INFINITY);
                    }
                }
                nms = -fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * scale_log2e;
            }
            // the previous item's P.V has finished by now: collect its O; this also guarantees that the
            // tensor core no longer reads this group's P tile.  (Warps without a block in either item only keep the
            // barriers in step: every arrival below happens behind the same waits as in a working warp.)
            if (have_prev) collect(use - 1);
            // ---- pass 2: P = exp2(S*scale - max*scale) -> fp16 -> K-major SW128 A tile of the P.V MMA:
            //      k block b (64 keys) at b*16 KB, row r at r*128, 16-byte chunk c at (c ^ (r & 7)) * 16;
            //      zeros for the key blocks of the tile's other sequences ----
            float sum = 1.f;
            if (active) {
                float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
                for (int p = 0; p < np; ++p) {
                    const int part = kb0 + p;
                    uint32_t v[32];
                    tmem_ld_32x32(s_tmem + part * 32, v);
                    tmem_ld_wait();
                    if (p == np - 1) {  // last read of S: the slot's S region may be overwritten
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_sfree + 8 * j);
                    }
                    const int lim = s - p * 32;
                    const uint32_t dst0 = prow_addr + (part >> 1) * 16384;
                    const int chunk0 = (part & 1) * 4;
                    if (lim >= 32) {
#pragma unroll
                        for (int c8 = 0; c8 < 4; ++c8) {
                            float pe[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                pe[e] = ex2_fast(fmaf(__uint_as_float(v[c8 * 8 + e]), scale_log2e, nms));
                                s4[e & 3] += pe[e];
                            }
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst0 + (((chunk0 + c8) ^ sw) * 16)),
                                         "r"(pack_f16x2(pe[0], pe[1])), "r"(pack_f16x2(pe[2], pe[3])),
                                         "r"(pack_f16x2(pe[4], pe[5])), "r"(pack_f16x2(pe[6], pe[7]))
                                         : "memory");
                        }
                    } else {
#pragma unroll
                        for (int c8 = 0; c8 < 4; ++c8) {
                            float pe[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                pe[e] = ex2_fast(fmaf(__uint_as_float(v[c8 * 8 + e]), scale_log2e, nms));
                                pe[e] = (c8 * 8 + e < lim) ? pe[e] : 0.f;
                                s4[e & 3] += pe[e];
                            }
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst0 + (((chunk0 + c8) ^ sw) * 16)),
                                         "r"(pack_f16x2(pe[0], pe[1])), "r"(pack_f16x2(pe[2], pe[3])),
                                         "r"(pack_f16x2(pe[4], pe[5])), "r"(pack_f16x2(pe[6], pe[7]))
                                         : "memory");
                        }
                    }
                }
#pragma unroll 1
                for (int part = 0; part < nb; ++part) {  // the other sequences' keys: exact zeros
                    if (part >= kb0 && part < kb0 + np) continue;
                    const uint32_t dst0 = prow_addr + (part >> 1) * 16384;
                    const int chunk0 = (part & 1) * 4;
#pragma unroll
                    for (int c8 = 0; c8 < 4; ++c8)
                        asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(dst0 + (((chunk0 + c8) ^ sw) * 16)), "r"(0u) : "memory");
                }
                sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
                fence_proxy_async_smem();
            } else {
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_sfree + 8 * j);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_pready + 8 * j);
            have_prev = true;
            prev_row0 = row0;
            prev_rows = s - blk * 32;  // rows of this block that belong to the sequence (>= 32: all of them)
            prev_head = item - tile * heads;
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

// The tile list of a pass (one CTA, once per pass; `out` = 8 header ints, then 8 ints per tile, then n_seq ints of
// scratch).  Sequences are sorted by their number of 32-row blocks np = ceil(len / 32) with a STABLE counting sort
// (ballots + a scan over the warps, so the list does not depend on the timing of atomics), then tiles are formed by
// index arithmetic: every 4 alone, every 3 with a 1 while there are any, the 2s in pairs (an odd one with up to two
// 1s), the remaining 1s in fours.  Record of a tile: row0[4] = first global row of each block (-1: unused; the used
// blocks are a prefix), info[4] = first key block | key blocks << 4 | sequence length << 8 | block within the
// sequence << 16 | used blocks << 20.
__global__ void __launch_bounds__(1024) attention_blocks_kernel(const int32_t* __restrict__ cu_seqlens, int n_seq,
                                                                 int32_t* __restrict__ out) {
    __shared__ int cnt[5], base[5], run[5], wc[5][32];
    int32_t* const recs = out + 8;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    auto cls_of = [&](int i) {
        const int len = cu_seqlens[i + 1] - cu_seqlens[i];
        return len <= 0 ? 0 : (len > AT_ROWS ? 4 : (len + 31) >> 5);
    };
    if (tid < 5) cnt[tid] = 0, run[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n_seq; i += 1024) atomicAdd(&cnt[cls_of(i)], 1);
    __syncthreads();
    const int c1 = cnt[1], c2 = cnt[2], c3 = cnt[3], c4 = cnt[4];
    if (tid == 0) {  // sorted order: the 4s, the 3s, the 2s, the 1s (empty sequences take no part)
        base[4] = 0;
        base[3] = c4;
        base[2] = c4 + c3;
        base[1] = c4 + c3 + c2;
        base[0] = 0;
    }
    // the sorted list lives behind the records of the at most n_seq tiles
    int32_t* const order = recs + 8 * static_cast<size_t>(n_seq);
    for (int i0 = 0; i0 < n_seq; i0 += 1024) {
        const int i = i0 + tid;
        const int c = i < n_seq ? cls_of(i) : 0;
        int rank = 0;
#pragma unroll
        for (int k = 1; k <= 4; ++k) {
            const unsigned m = __ballot_sync(0xffffffffu, c == k);
            if (c == k) rank = __popc(m & ((1u << lane) - 1u));
            if (lane == 0) wc[k][wid] = __popc(m);
        }
        __syncthreads();
        if (c > 0) {
            int before = 0;
            for (int w = 0; w < wid; ++w) before += wc[c][w];
            order[base[c] + run[c] + before + rank] = i;
        }
        __syncthreads();
        if (tid >= 1 && tid <= 4) {
            int tot = 0;
            for (int w = 0; w < 32; ++w) tot += wc[tid][w];
            run[tid] += tot;
        }
        __syncthreads();
    }
    const int* const L4 = order;
    const int* const L3 = order + c4;
    const int* const L2 = order + c4 + c3;
    const int* const L1 = order + c4 + c3 + c2;
    const int m31 = min(c3, c1);            // 1s that ride with a 3
    const int pairs2 = c2 >> 1, odd2 = c2 & 1;
    const int with_odd = odd2 ? min(2, c1 - m31) : 0;  // 1s that ride with the odd 2
    const int ones0 = m31 + with_odd;       // first 1 of the tiles of four
    const int t1 = (c1 - ones0 + 3) >> 2;
    const int n_tiles = c4 + c3 + pairs2 + odd2 + t1;
    if (tid == 0) out[0] = n_tiles;
    for (int t = tid; t < n_tiles; t += 1024) {
        int seqs[4] = {-1, -1, -1, -1};
        int u = t;
        if (u < c4) {
            seqs[0] = L4[u];
        } else if ((u -= c4) < c3) {
            seqs[0] = L3[u];
            if (u < m31) seqs[1] = L1[u];
        } else if ((u -= c3) < pairs2) {
            seqs[0] = L2[2 * u];
            seqs[1] = L2[2 * u + 1];
        } else if ((u -= pairs2) < odd2) {
            seqs[0] = L2[c2 - 1];
            for (int x = 0; x < with_odd; ++x) seqs[1 + x] = L1[m31 + x];
        } else {
            u -= odd2;
            for (int x = 0; x < 4; ++x)
                if (ones0 + 4 * u + x < c1) seqs[x] = L1[ones0 + 4 * u + x];
        }
        int row0[4] = {-1, -1, -1, -1};
        unsigned info[4] = {0u, 0u, 0u, 0u};
        int kb = 0;
        for (int x = 0; x < 4; ++x) {
            if (seqs[x] < 0) continue;
            const int t0 = cu_seqlens[seqs[x]];
            int len = cu_seqlens[seqs[x] + 1] - t0;
            len = len > AT_ROWS ? AT_ROWS : len;  // the launcher refuses longer sequences
            const int np = (len + 31) >> 5;
            for (int b = 0; b < np; ++b) {
                row0[kb + b] = t0 + 32 * b;
                info[kb + b] = static_cast<unsigned>(kb) | static_cast<unsigned>(np) << 4 | static_cast<unsigned>(len) << 8 |
                               static_cast<unsigned>(b) << 16;
            }
            kb += np;
        }
        for (int b = 0; b < 4; ++b) {
            recs[8 * static_cast<size_t>(t) + b] = row0[b];
            recs[8 * static_cast<size_t>(t) + 4 + b] = static_cast<int32_t>(info[b] | static_cast<unsigned>(kb) << 20);
        }
    }
}

}  // namespace

// ints of scratch attention_blocks_launch needs for n_seq sequences
int64_t attention_blocks_scratch_ints(int n_seq) { return 8 + 9 * static_cast<int64_t>(n_seq); }

int attention_blocks_launch(const int32_t* cu_seqlens, int n_seq, int32_t* tiles, cudaStream_t stream) {
    if (n_seq <= 0) return 0;
    attention_blocks_kernel<<<1, 1024, 0, stream>>>(cu_seqlens, n_seq, tiles);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

// qkv: [T, 3H] with Q, K bf16 and V fp16.  Only d_head == 32 and max_seqlen <= 128; tiles: what attention_blocks_launch
// wrote for the same cu_seqlens.
int attention_tc_blocks_launch(const __nv_bfloat16* qkv, const int32_t* tiles, int n_seq, int T, int H, int heads,
                               int max_seqlen, __nv_bfloat16* ctx, int num_sms, cudaStream_t stream) {
    if (n_seq <= 0 || T <= 0) return 0;
    if (H / heads != AT_D || max_seqlen > AT_ROWS) return fail(LSC_E_UNSUPPORTED, "attention_tc_blocks: needs d_head 32 and s <= 128");
    CUtensorMap tm, tmc;
    LSC_TRY(make_tmap_bf16_box(&tm, qkv, static_cast<uint64_t>(T), static_cast<uint64_t>(3) * H, 32, AT_D, 64));
    LSC_TRY(make_tmap_bf16_store32(&tmc, ctx, static_cast<uint64_t>(T), static_cast<uint64_t>(H)));
    static bool configured[64] = {};
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(attention_tc_blocks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    const long long max_items = static_cast<long long>(heads) * n_seq;  // the tile count lives on the device
    const int grid = static_cast<int>(max_items < num_sms ? max_items : num_sms);
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
    LSC_CUDA(cudaLaunchKernelEx(&lc, attention_tc_blocks_kernel, tm, tmc, tiles, H, heads, ctx, scale_log2e));
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
