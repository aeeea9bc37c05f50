// Brute-force cosine top-k (K8/K9): the scan HerdDB runs for
//   SELECT … ORDER BY cosine_similarity(col, CAST(? AS FLOAT ARRAY)) DESC LIMIT k
// (reference call site AI/ai/langstream/ai/agents/datasource/impl/JdbcDataSourceProvider.java:103-124,
//  SQL at examples/applications/docker-chatbot/chatbot.yaml:44), with the cosine defined as in
// ReRankAgent.cosineSimilarity (AI/ai/langstream/ai/agents/rerank/ReRankAgent.java:262-273).
// Rows are stored L2-normalised in bf16, so cosine = dot product.
//
// topk_scan_kernel — persistent, one CTA per SM, HBM-bound for small query batches:
//   warp 0     TMA producer: the (<=128) queries once, then corpus tiles of 256 rows x 64 dims
//              (SWIZZLE_128B, L2 evict-first) through a 4-stage smem ring
//   warp 1     tcgen05.mma, M = 128 query rows (zero padded), N = 256 corpus rows, fp32 scores in
//              TMEM, two accumulator stages
//   warps 2-9  selection: thread q owns query q (TMEM lane q); the two warps of a lane quadrant each walk
//              one half (128 columns) of the tile's scores keeping their k best in a private sorted
//              list; a chunk of 32 scores is skipped with one max-reduction when nothing beats the
//              current k-th score; the two halves are merged through shared memory at the end
// Each CTA then writes its [128][k] candidates to scratch and topk_merge_kernel merges the per-CTA
// lists (a k-way merge over sorted heads) into the final [B][k] ids and scores.
#include <cstdlib>

#include "common.h"
#include "ptx.cuh"
#include "topk.h"

namespace lsc {

namespace {

// threads = 64 + EW * 32: warp 0 TMA, warp 1 MMA, EW selection warps (8 or 16)
constexpr int TK_BN = 256;
constexpr int TK_BK = 64;
constexpr int TK_ACC_STRIDE = 256;
constexpr int TK_TMEM_COLS = 512;
constexpr int TK_STAGE_BYTES = TK_BN * TK_BK * 2;  // 32 KB
constexpr int TK_QBLK_BYTES = TK_QROWS * TK_BK * 2;  // 16 KB per 64-dim slice of the query block
constexpr int TK_MAX_STAGES = 6;

// Per-query sorted (best first) list of the k best rows so far, in (L1-resident) local memory; empty
// slots hold (-inf, -1).  tie=false: the new row goes after every listed row with an equal score
// (the scan feeds rows in increasing order) — "ties by lower row" without index compares;
// tie=true compares row numbers explicitly (used where rows do not arrive in order).
__device__ __noinline__ void list_insert_mem(float* s, int32_t* i, int k, float v, int32_t idx, bool tie) {
    auto before = [&](int j) { return s[j] > v || (s[j] == v && (!tie || (i[j] >= 0 && i[j] < idx))); };
    if (before(k - 1)) return;
    int pos = k - 1;
    while (pos > 0 && !before(pos - 1)) {
        s[pos] = s[pos - 1];
        i[pos] = i[pos - 1];
        --pos;
    }
    s[pos] = v;
    i[pos] = idx;
}

template <int KCAP, int EW>
__global__ void __launch_bounds__(64 + EW * 32, 1)
topk_scan_kernel(const __grid_constant__ CUtensorMap tmapQ, const __grid_constant__ CUtensorMap tmapC,
                 int q_row0, int nq, int n_rows, int dim, int k, int stages,
                 float* __restrict__ cand_scores, int32_t* __restrict__ cand_idx) {
    extern __shared__ uint8_t smem_raw[];
    // aligned by pointer arithmetic on the shared array (an integer round trip would turn every access
    // through it into a generic-address one)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int num_kb = dim / TK_BK;
    const uint32_t sQ = smem_u32(smem);
    const uint32_t sC = sQ + num_kb * TK_QBLK_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + num_kb * TK_QBLK_BYTES + stages * TK_STAGE_BYTES);
    const uint32_t bar_full = smem_u32(bars);
    const uint32_t bar_empty = bar_full + 8 * TK_MAX_STAGES;
    const uint32_t bar_tfull = bar_empty + 8 * TK_MAX_STAGES;
    const uint32_t bar_tempty = bar_tfull + 16;
    const uint32_t bar_q = bar_tempty + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + 2 * TK_MAX_STAGES + 5);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapQ);
        tma_prefetch_desc(&tmapC);
        for (int s = 0; s < stages; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, EW);
        }
        mbar_init(bar_q, 1);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TK_TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int num_tiles = (n_rows + TK_BN - 1) / TK_BN;

    if (warp == 0) {
        if (elect_one()) {
            mbar_arrive_expect_tx(bar_q, num_kb * TK_QBLK_BYTES);
            for (int kb = 0; kb < num_kb; ++kb)
                tma_load_2d(sQ + kb * TK_QBLK_BYTES, &tmapQ, bar_q, kb * TK_BK, q_row0);
            const uint64_t pol = policy_evict_first();
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    mbar_arrive_expect_tx(bar_full + 8 * stage, TK_STAGE_BYTES);
                    tma_load_2d_hint(sC + stage * TK_STAGE_BYTES, &tmapC, bar_full + 8 * stage,
                                     kb * TK_BK, tile * TK_BN, pol);
                    if (++stage == stages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr uint32_t idesc = umma_idesc_bf16(TK_QROWS, TK_BN);
            mbar_wait(bar_q, 0);
            tc_fence_after();
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * TK_ACC_STRIDE;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after();
                    const uint64_t da = umma_desc_sw128(sQ + kb * TK_QBLK_BYTES);
                    const uint64_t db = umma_desc_sw128(sC + stage * TK_STAGE_BYTES);
#pragma unroll
                    for (int kk = 0; kk < TK_BK / 16; ++kk)
                        umma_bf16(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
                    umma_commit(bar_empty + 8 * stage);
                    if (++stage == stages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(bar_tfull + 8 * acc);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else {
        // Selection: EW/4 warps per TMEM lane quadrant, each walking its share of the columns of every
        // 256-row score tile for the same 32 queries; thread = query.
        constexpr int PARTS = EW / 4;
        constexpr int PCOLS = TK_BN / PARTS;
        const int quad = warp & 3;
        const int half = (warp - 2) >> 2;  // which column part
        const int q = quad * 32 + lane;        // query (TMEM lane) owned by this thread
        const bool active = q < nq;
        const bool warp_active = quad * 32 < nq;
        float ls[KCAP];
        int32_t li[KCAP];
        for (int j = 0; j < k; ++j) {
            ls[j] = -INFINITY;
            li[j] = -1;
        }
        float thr = -INFINITY;  // current k-th best score of this query
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            mbar_wait(bar_tfull + 8 * acc, acc_phase);
            tc_fence_after();
            if (warp_active) {
                const int col0 = half * PCOLS;
                const uint32_t tbase = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) +
                                       acc * TK_ACC_STRIDE + col0;
                uint32_t va[32], vb[32];
                tmem_ld_32x32(tbase, va);
#pragma unroll
                for (int c = 0; c < PCOLS / 32; ++c) {
                    uint32_t(&v)[32] = (c & 1) ? vb : va;
                    uint32_t(&vnext)[32] = (c & 1) ? va : vb;
                    tmem_ld_wait();
                    if (c + 1 < PCOLS / 32) tmem_ld_32x32(tbase + (c + 1) * 32, vnext);
                    const int r0 = tile * TK_BN + col0 + c * 32;
                    const int nvalid = n_rows - r0;  // rows past n_rows are not candidates
                    if (active && nvalid > 0) {
                        // max tree of the chunk: one comparison against the k-th best rejects it
                        float m[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            m[j] = fmaxf(j < nvalid ? __uint_as_float(v[j]) : -INFINITY,
                                         j + 16 < nvalid ? __uint_as_float(v[j + 16]) : -INFINITY);
#pragma unroll
                        for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], m[j + 8]);
#pragma unroll
                        for (int j = 0; j < 4; ++j) m[j] = fmaxf(m[j], m[j + 4]);
                        if (fmaxf(fmaxf(m[0], m[2]), fmaxf(m[1], m[3])) > thr) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const float sc = __uint_as_float(v[j]);
                                if (j < nvalid && sc > thr) {
                                    list_insert_mem(ls, li, k, sc, r0 + j, false);
                                    thr = ls[k - 1];
                                }
                            }
                        }
                    }
                }
            }
            __syncwarp();
            tc_fence_before();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
        // ---- combine the two column halves of each query inside the CTA, then write [128][k] ----
        // The smem ring is idle now (every TMA load has been consumed by an MMA that completed before
        // the last tfull arrival), so it serves as the exchange buffer: [128][k] (score, row) pairs.
        float* xs = reinterpret_cast<float*>(smem);  // queries and ring are both dead by now
        int32_t* xi = reinterpret_cast<int32_t*>(xs + (PARTS - 1) * TK_QROWS * k);
        asm volatile("bar.sync 1, %0;" ::"n"(EW * 32));  // all selection warps are past their last tile
        if (half > 0 && active) {
            for (int j = 0; j < k; ++j) {
                xs[((half - 1) * TK_QROWS + q) * k + j] = ls[j];
                xi[((half - 1) * TK_QROWS + q) * k + j] = li[j];
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(EW * 32));
        if (half == 0 && active) {
            for (int j = 0; j < (PARTS - 1) * k; ++j) {
                const int src = (j / k) * TK_QROWS * k + q * k + (j % k);
                const float sc = xs[src];
                const int32_t idx = xi[src];
                // the other half holds rows interleaved with ours: order equal scores by row number
                if (idx >= 0 && sc >= thr) {
                    list_insert_mem(ls, li, k, sc, idx, true);
                    thr = ls[k - 1];
                }
            }
            const size_t base = (static_cast<size_t>(blockIdx.x) * TK_QROWS + q) * k;
            for (int j = 0; j < k; ++j) {
                cand_scores[base + j] = ls[j];
                cand_idx[base + j] = li[j];
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TK_TMEM_COLS);
    }
}

// ---- k-way merge of sorted candidate lists -------------------------------------------------------
// One CTA per query, thread p holds the head of part p's list.  Each round a block-wide argmax picks
// the best head (score descending; ties by `tie` ascending), emits it and advances that list.
struct Head {
    float s;
    long long tie;
    int part;
};
__device__ __forceinline__ bool better(const Head& a, const Head& b) {
    if (a.s != b.s) return a.s > b.s;
    if (a.tie != b.tie) return a.tie < b.tie;
    return a.part < b.part;
}
__device__ __forceinline__ Head shfl_head(const Head& h, int o) {
    Head r;
    r.s = __shfl_xor_sync(0xffffffffu, h.s, o);
    r.tie = __shfl_xor_sync(0xffffffffu, h.tie, o);
    r.part = __shfl_xor_sync(0xffffffffu, h.part, o);
    return r;
}

// LAYOUT 0: scan scratch  [qblock][parts][128][k], idx = int32 storage slot, ids mapped through
//           `slot_ids`; ties by slot.
// LAYOUT 1: shard lists   [parts][B][k], idx = int64 id; ties by part (shards hold ascending
//           contiguous row ranges) then by position inside the part's list.
template <int LAYOUT>
__global__ void __launch_bounds__(256)
topk_merge_kernel(const float* __restrict__ sc, const void* __restrict__ ix, int parts, int B, int k,
                  const int64_t* __restrict__ slot_ids, int64_t* __restrict__ out_ids,
                  float* __restrict__ out_scores) {
    __shared__ Head red[8];
    __shared__ Head winner;
    const int q = blockIdx.x;
    const int p = threadIdx.x;
    size_t base = 0;
    if (p < parts) {
        if (LAYOUT == 0)
            base = ((static_cast<size_t>(q / TK_QROWS) * parts + p) * TK_QROWS + (q % TK_QROWS)) * k;
        else
            base = (static_cast<size_t>(p) * B + q) * k;
    }
    int pos = 0;
    auto load_head = [&](int at) {
        Head h;
        h.part = p;
        h.s = -INFINITY;
        h.tie = 0x7fffffffffffffffLL;
        if (p < parts && at < k) {
            long long id;
            if (LAYOUT == 0) id = static_cast<const int32_t*>(ix)[base + at];
            else id = static_cast<const int64_t*>(ix)[base + at];
            if (id >= 0) {
                h.s = sc[base + at];
                // LAYOUT 1: shard first, then position inside the shard's list = the global storage order
                // among equal scores (shards hold ascending row ranges), so duplicates come back in the same
                // order as from one unsharded index
                h.tie = LAYOUT == 0 ? id : ((static_cast<long long>(p) << 32) | static_cast<long long>(at));
            }
        }
        return h;
    };
    Head mine = load_head(0);
    for (int round = 0; round < k; ++round) {
        Head best = mine;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const Head other = shfl_head(best, o);
            if (better(other, best)) best = other;
        }
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = best;
        __syncthreads();
        if (threadIdx.x == 0) {
            Head b = red[0];
            for (int w = 1; w < 8; ++w)
                if (better(red[w], b)) b = red[w];
            winner = b;
        }
        __syncthreads();
        const Head w = winner;
        if (w.s == -INFINITY && w.tie == 0x7fffffffffffffffLL) {
            // every list exhausted: fewer than k rows exist
            if (threadIdx.x == 0) {
                for (int r = round; r < k; ++r) {
                    out_ids[static_cast<size_t>(q) * k + r] = -1;
                    out_scores[static_cast<size_t>(q) * k + r] = -INFINITY;
                }
            }
            return;
        }
        if (p == w.part) {
            long long id;
            if (LAYOUT == 0) {
                const int32_t slot = static_cast<const int32_t*>(ix)[base + pos];
                id = slot_ids != nullptr ? slot_ids[slot] : slot;
            } else {
                id = static_cast<const int64_t*>(ix)[base + pos];
            }
            out_ids[static_cast<size_t>(q) * k + round] = id;
            out_scores[static_cast<size_t>(q) * k + round] = w.s;
            ++pos;
            mine = load_head(pos);
        }
        // `winner`/`red` are rewritten only after the next round's first __syncthreads
    }
}

// ---- small row kernels -----------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// fp32 rows -> L2-normalised bf16 rows at dst_rows[i] (or row i when dst_rows == nullptr);
// rows in [n, n_pad) are written as zeros (query padding).  A zero vector stays zero.
__global__ void __launch_bounds__(256)
normalize_rows_kernel(const float* __restrict__ src, int n, int n_pad, int dim,
                      const int64_t* __restrict__ dst_rows, __nv_bfloat16* __restrict__ dst,
                      int32_t* __restrict__ gthr, int n_gthr, unsigned* __restrict__ done, int n_done) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    // state of the fused scan that follows in the stream (topk_fuse.cuh)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_gthr; i += gridDim.x * blockDim.x) gthr[i] = TKX_KEY_NEG_INF;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_done; i += gridDim.x * blockDim.x) done[i] = 0u;
    if (r >= n_pad) return;
    long long out_row = r;
    if (dst_rows != nullptr) {
        if (r >= n) return;
        out_row = dst_rows[r];
        if (out_row < 0) return;  // superseded duplicate inside one upsert batch
    }
    __nv_bfloat16* o = dst + static_cast<size_t>(out_row) * dim;
    if (r >= n) {
        for (int c = lane * 2; c < dim; c += 64) *reinterpret_cast<uint32_t*>(o + c) = 0u;
        return;
    }
    const float* s = src + static_cast<size_t>(r) * dim;
    float q = 0.f;
    for (int c = lane * 2; c < dim; c += 64) {
        const float2 v = *reinterpret_cast<const float2*>(s + c);
        q += v.x * v.x + v.y * v.y;
    }
    const float norm = sqrtf(warp_sum(q));
    const float inv = norm > 0.f ? 1.0f / norm : 0.f;
    for (int c = lane * 2; c < dim; c += 64) {
        const float2 v = *reinterpret_cast<const float2*>(s + c);
        *reinterpret_cast<uint32_t*>(o + c) = pack_bf16x2(v.x * inv, v.y * inv);
    }
}

// Debug / verification: exact fp32 dot products of every stored row with every (bf16) query.
__global__ void __launch_bounds__(256)
scores_naive_kernel(const __nv_bfloat16* __restrict__ rows, long long n, int dim,
                    const __nv_bfloat16* __restrict__ q, int B, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long r = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= n) return;
    const uint32_t* rp = reinterpret_cast<const uint32_t*>(rows + static_cast<size_t>(r) * dim);
    for (int b = 0; b < B; ++b) {
        const uint32_t* qp = reinterpret_cast<const uint32_t*>(q + static_cast<size_t>(b) * dim);
        float a = 0.f;
        for (int c = lane; c < dim / 2; c += 32) {
            const uint32_t x = __ldg(rp + c), y = __ldg(qp + c);
            a = fmaf(bf16_lo(x), bf16_lo(y), a);
            a = fmaf(bf16_hi(x), bf16_hi(y), a);
        }
        a = warp_sum(a);
        if (lane == 0) out[static_cast<size_t>(b) * n + r] = a;
    }
}

// Delete support: row (and id) moves[2j+1] -> slot moves[2j], one warp per move.  Sources lie past the new end
// of the slab and destinations in front of it, so no move reads what another one writes.
__global__ void __launch_bounds__(256)
move_rows_kernel(__nv_bfloat16* __restrict__ rows, int64_t* __restrict__ ids, int dim,
                 const int64_t* __restrict__ moves, int m) {
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (j >= m) return;
    const int64_t dst = moves[2 * j], src = moves[2 * j + 1];
    const uint4* s = reinterpret_cast<const uint4*>(rows + static_cast<size_t>(src) * dim);
    uint4* d = reinterpret_cast<uint4*>(rows + static_cast<size_t>(dst) * dim);
    for (int c = lane; c < dim / 8; c += 32) d[c] = s[c];
    if (lane == 0) ids[dst] = ids[src];
}

// Exchange + final merge for shard results that already lie in memory as [B][k] lists: one CTA per query.
__global__ void __launch_bounds__(TKX_THREADS)
topk_exchange_kernel(const int64_t* __restrict__ local_ids, const float* __restrict__ local_scores, int B, int k,
                     const __grid_constant__ TopkFuse fuse) {
    __shared__ TopkTailSmem sm;
    for (int q = blockIdx.x; q < B; q += gridDim.x) {
        if (threadIdx.x < k) {
            sm.res_s[threadIdx.x] = local_scores[static_cast<size_t>(q) * k + threadIdx.x];
            sm.res_i[threadIdx.x] = local_ids[static_cast<size_t>(q) * k + threadIdx.x];
        }
        tail_bar();
        tail_exchange(fuse, q, k, sm);
    }
}

__global__ void fill_empty_kernel(int64_t* ids, float* scores, size_t n) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) {
        ids[i] = -1;
        scores[i] = -INFINITY;
    }
}

template <int KCAP, int EW>
int scan_inst(const CUtensorMap& tmapQ, const CUtensorMap& tmapC, int q_row0, int nq, int n_rows,
              int dim, int k, int stages, size_t smem, int grid, float* cand_scores, int32_t* cand_idx,
              cudaStream_t stream) {
    auto kern = topk_scan_kernel<KCAP, EW>;
    static bool configured[64] = {};  // once per device: not on the launch path of a search whose peers may wait
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    kern<<<grid, 64 + EW * 32, smem, stream>>>(tmapQ, tmapC, q_row0, nq, n_rows, dim, k, stages, cand_scores,
                                             cand_idx);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace

// Load every kernel a sharded search can launch into the current device's context NOW.  The CUDA runtime loads a
// kernel lazily at its first launch, and that load can wait for the kernels already running on the device - one
// of which may be a peer shard's scan or exchange kernel that waits for this very launch.
int topk_preload_kernels() {
    cudaFuncAttributes a;
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_exchange_kernel));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_merge_kernel<0>));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_merge_kernel<1>));
    LSC_CUDA(cudaFuncGetAttributes(&a, normalize_rows_kernel));
    LSC_CUDA(cudaFuncGetAttributes(&a, fill_empty_kernel));
    LSC_CUDA(cudaFuncGetAttributes(&a, move_rows_kernel));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_kernel<16, 8>));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_kernel<32, 8>));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_kernel<128, 8>));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_kernel<16, 16>));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_kernel<32, 16>));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_kernel<128, 16>));
    return topk_rows_preload_kernels();
}

int topk_scan_grid(int n_rows, int num_sms) {
    const int tiles = (n_rows + TK_BN - 1) / TK_BN;
    return tiles < num_sms ? tiles : num_sms;
}

int topk_scan_launch(const CUtensorMap& tmapQ, const CUtensorMap& tmapC, int q_row0, int nq, int n_rows,
                     int dim, int k, int grid, float* cand_scores, int32_t* cand_idx,
                     cudaStream_t stream) {
    if (dim % TK_BK != 0 || dim <= 0 || dim > 512)
        return fail(LSC_E_UNSUPPORTED, "top-k: dim must be a multiple of 64 and <= 512");
    if (k < 1 || k > TK_KMAX) return fail(LSC_E_INVALID, "top-k: k must be in [1,128]");
    const int q_bytes = (dim / TK_BK) * TK_QBLK_BYTES;
    int stages = (227 * 1024 - 1024 - 256 - q_bytes) / TK_STAGE_BYTES;
    if (stages > TK_MAX_STAGES) stages = TK_MAX_STAGES;
    if (stages < 2) return fail(LSC_E_UNSUPPORTED, "top-k: dim too large for the shared-memory ring");
    const size_t smem = static_cast<size_t>(q_bytes) + static_cast<size_t>(stages) * TK_STAGE_BYTES + 256 + 1024;
    static const int ew = [] { const char* e = getenv("LSC_TOPK_WARPS"); return e ? atoi(e) : 8; }();
    if (ew == 16) {
        if (k <= 16) return scan_inst<16, 16>(tmapQ, tmapC, q_row0, nq, n_rows, dim, k, stages, smem, grid, cand_scores, cand_idx, stream);
        if (k <= 32) return scan_inst<32, 16>(tmapQ, tmapC, q_row0, nq, n_rows, dim, k, stages, smem, grid, cand_scores, cand_idx, stream);
        return scan_inst<128, 16>(tmapQ, tmapC, q_row0, nq, n_rows, dim, k, stages, smem, grid, cand_scores, cand_idx, stream);
    }
    if (k <= 16) return scan_inst<16, 8>(tmapQ, tmapC, q_row0, nq, n_rows, dim, k, stages, smem, grid, cand_scores, cand_idx, stream);
    if (k <= 32) return scan_inst<32, 8>(tmapQ, tmapC, q_row0, nq, n_rows, dim, k, stages, smem, grid, cand_scores, cand_idx, stream);
    return scan_inst<128, 8>(tmapQ, tmapC, q_row0, nq, n_rows, dim, k, stages, smem, grid, cand_scores, cand_idx, stream);
}

int topk_merge_scan_launch(const float* cand_scores, const int32_t* cand_idx, int parts, int B, int k,
                           const int64_t* slot_ids, int64_t* out_ids, float* out_scores,
                           cudaStream_t stream) {
    if (B <= 0) return 0;
    if (parts > 256) return fail(LSC_E_UNSUPPORTED, "top-k merge: more than 256 partial lists");
    topk_merge_kernel<0><<<B, 256, 0, stream>>>(cand_scores, cand_idx, parts, B, k, slot_ids, out_ids, out_scores);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int topk_merge_shards_launch(const float* scores, const int64_t* ids, int parts, int B, int k,
                             int64_t* out_ids, float* out_scores, cudaStream_t stream) {
    if (B <= 0) return 0;
    if (parts > 256) return fail(LSC_E_UNSUPPORTED, "top-k merge: more than 256 partial lists");
    topk_merge_kernel<1><<<B, 256, 0, stream>>>(scores, ids, parts, B, k, nullptr, out_ids, out_scores);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int normalize_rows_launch(const float* src, int n, int n_pad, int dim, const int64_t* dst_rows,
                          __nv_bfloat16* dst, cudaStream_t stream, int32_t* gthr, int n_gthr, unsigned* done,
                          int n_done) {
    if (n_pad <= 0) return 0;
    const int wpb = 8;
    normalize_rows_kernel<<<(n_pad + wpb - 1) / wpb, 256, 0, stream>>>(src, n, n_pad, dim, dst_rows, dst, gthr, n_gthr,
                                                                      done, n_done);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int scores_naive_launch(const __nv_bfloat16* rows, int64_t n, int dim, const __nv_bfloat16* q, int B,
                        float* out, cudaStream_t stream) {
    if (n <= 0 || B <= 0) return 0;
    const int wpb = 8;
    const int64_t grid = (n + wpb - 1) / wpb;
    scores_naive_kernel<<<static_cast<unsigned>(grid), 256, 0, stream>>>(rows, n, dim, q, B, out);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int move_rows_launch(__nv_bfloat16* rows, int64_t* ids, int dim, const int64_t* moves, int m, cudaStream_t stream) {
    if (m <= 0) return 0;
    const int wpb = 8;
    move_rows_kernel<<<(m + wpb - 1) / wpb, 256, 0, stream>>>(rows, ids, dim, moves, m);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

// One CTA per query, unless the shard shares its GPU with other shards (fuse.cooperative == 0): their waiting
// exchange CTAs must leave room for the scans they wait for, so a small grid walks the queries instead.
int topk_exchange_launch(const int64_t* local_ids, const float* local_scores, int B, int k, const TopkFuse& fuse,
                         cudaStream_t stream) {
    if (B <= 0) return 0;
    if (k > TKX_MAX_XK || k > fuse.xk || fuse.world > TKX_MAX_WORLD || fuse.world < 2)
        return fail(LSC_E_UNSUPPORTED, "top-k exchange: k <= 32 and 2..8 shards");
    const int grid = fuse.cooperative != 0 ? B : (B < 32 ? B : 32);
    topk_exchange_kernel<<<grid, TKX_THREADS, 0, stream>>>(local_ids, local_scores, B, k, fuse);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int fill_empty_launch(int64_t* ids, float* scores, size_t n, cudaStream_t stream) {
    if (n == 0) return 0;
    fill_empty_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(ids, scores, n);
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
