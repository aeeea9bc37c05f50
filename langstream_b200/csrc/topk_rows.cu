// Brute-force cosine top-k, row-major selection (k <= 16): the second scan kernel behind
// lsc_index_search*, same contract as topk_scan_kernel in topk.cu (HerdDB's
// `ORDER BY cosine_similarity(...) DESC LIMIT k`, JdbcDataSourceProvider.java:103-124).
//
// topk.cu puts the QUERIES on the TMEM lanes (thread = query): with 32 queries per warp, every lane has
// its own threshold and a hit in any lane sends the whole warp down the insertion path, which costs
// 1.6-1.8x the HBM time at B = 64.  Here the operands are swapped: the corpus tile is the A operand
// (M = 128 rows), the query block the B operand (N = queries), so the score tile has ROWS on lanes and
// queries on columns.  For one query the 32 lanes of a warp hold 32 different rows against ONE
// threshold, so the common case is branch-free: each lane builds the bit mask of queries its row beats
// (32 compares against broadcast thresholds) and one warp-wide OR says whether anything happened.
// Sixteen selection warps: warp (quadrant, chunk) looks at 32 rows x 32 queries of every tile, so a full
// query block keeps four warps on each scheduler (one warp per scheduler runs at the latency of every
// single instruction and cannot keep up with HBM).
// A hit is APPENDED by its own lane to the query's candidate buffer in shared memory (one shared-memory
// atomic for the slot; the score of the hit column is re-read from TMEM, which is how a run-time column
// index is done without local memory); nothing warp-wide happens per hit.  After each tile the selection
// warps meet at a named barrier, and every query whose buffer holds >= TR_TRIG candidates is compacted into
// its sorted top-k list by the warp that owns it: all-pairs ranking, every lane counts how many of the
// buffered candidates and list entries precede its own candidate / list entry and stores it at that
// position - no serial insertion.  The compaction also raises the query's threshold.  With k = 10 a query
// is compacted about 7 times per CTA over a 10M-row scan (each compaction lets the seen-row count grow
// 2.6x before the buffer refills), against ~100 list insertions when every hit is inserted at once.
// A buffer that overflows inside one tile (the first tiles, thresholds still at -inf) leaves the
// (row, query) pair pending: the tile is re-tested for pending pairs after the compaction, against the
// raised thresholds, until none is left.
// At the end the lists are written out as the same [128][k] per-CTA candidates topk_merge_kernel consumes.
#include <cstdlib>

#include "common.h"
#include "ptx.cuh"
#include "topk.h"
#include "topk_fuse.cuh"

namespace lsc {

namespace {

constexpr int TR_SEL_WARPS = 16;  // selection warp s: TMEM lane quadrant (s + 2) % 4, query chunk s / 4
constexpr int TR_THREADS = 64 + 32 * TR_SEL_WARPS;  // warp 0 TMA, warp 1 MMA, warps 2-17 selection
constexpr int TR_BM = 128;       // corpus rows per tile
constexpr int TR_BK = 64;
constexpr int TR_STAGE_BYTES = TR_BM * TR_BK * 2;      // 16 KB: one 64-dim slice of a corpus tile
constexpr int TR_QBLK_BYTES = TK_QROWS * TR_BK * 2;    // 16 KB: one 64-dim slice of the query block
constexpr int TR_MAX_STAGES = 8;
constexpr int TR_ACC_STRIDE = 128;                     // TMEM columns per accumulator stage (N <= 128)
constexpr int TR_BUF = 32;                             // candidate buffer entries per query (= one per lane)
constexpr int TR_TRIG = 16;                            // compact a query once its buffer holds this many

// One warp merges the n (<= 32) buffered candidates of a query into its sorted top-k list (k <= 16) by
// all-pairs ranking.  Order: higher score first, equal scores by lower row number (the reference's stable
// descending sort); rows are unique, so the positions are a permutation.  Lane t holds candidate t and, for
// t < k, list entry t; the others' values are broadcast reads from shared memory, all independent.
// ls / li / bs / bi / thr are 32-bit shared-window addresses; the caller owns the query.
// gthr_q (may be null): the query's threshold shared between the CTAs of the grid - a raised local k-th best
// score is published there with one atomic max (see topk_fuse.cuh).
__device__ __forceinline__ void compact_query(uint32_t ls, uint32_t li, uint32_t bs, uint32_t bi, uint32_t thr,
                                              int k, int n, int32_t* gthr_q) {
    const int lane = threadIdx.x & 31;
    const bool hc = lane < n, hl = lane < k;
    const float cs = hc ? lds_f32(bs + 4 * lane) : -INFINITY;
    const int32_t cr = hc ? lds_s32(bi + 4 * lane) : 0x7fffffff;
    const float es = hl ? lds_f32(ls + 4 * lane) : -INFINITY;
    const int32_t er = hl ? lds_s32(li + 4 * lane) : -1;
    int cpos = 0;     // candidates and valid list entries in front of my candidate
    int epos = lane;  // list entries in front of my list entry (their order is kept) + candidates in front
    for (int t = 0; t < n; ++t) {
        const float s = lds_f32(bs + 4 * t);
        const int32_t r = lds_s32(bi + 4 * t);
        cpos += (s > cs || (s == cs && r < cr)) ? 1 : 0;
        epos += (s > es || (s == es && r < er)) ? 1 : 0;  // an empty slot (-inf, -1) is behind everything
    }
    for (int t = 0; t < k; ++t) {
        const float s = lds_f32(ls + 4 * t);
        const int32_t r = lds_s32(li + 4 * t);
        cpos += (r >= 0 && (s > cs || (s == cs && r < cr))) ? 1 : 0;
    }
    __syncwarp();  // every lane has read the old list
    if (hc && cpos < k) {
        sts_f32(ls + 4 * cpos, cs);
        sts_s32(li + 4 * cpos, cr);
        if (cpos == k - 1) {
            sts_f32(thr, cs);
            if (gthr_q != nullptr) atomicMax(gthr_q, score_key(cs));
        }
    }
    if (hl && epos < k) {
        sts_f32(ls + 4 * epos, es);
        sts_s32(li + 4 * epos, er);
        if (epos == k - 1) {
            sts_f32(thr, er >= 0 ? es : -INFINITY);
            if (gthr_q != nullptr && er >= 0) atomicMax(gthr_q, score_key(es));
        }
    }
    __syncwarp();
}

// PAIR = 2 (query blocks of more than 128 queries): a cluster of two CTAs scans 256 corpus rows per tile
// against 256 queries with cta_group::2 MMAs (M = 256 rows, N = 256 queries).  Each CTA streams its own 128
// rows and keeps its own 128 queries resident - exactly the B-operand half a CTA pair stages - so one pass
// over the corpus serves twice the queries for the same HBM traffic and shared memory; each CTA selects for
// all 256 queries over its own rows (two 32-query chunks per warp, 16-entry candidate buffers).
template <int PAIR>
__global__ void __launch_bounds__(TR_THREADS, 1)
topk_scan_rows_kernel(const __grid_constant__ CUtensorMap tmapQ, const __grid_constant__ CUtensorMap tmapC,
                      int q_row0, int nq, int n_rows, int dim, int k, int stages, uint32_t idesc,
                      float* __restrict__ cand_scores, int32_t* __restrict__ cand_idx,
                      const __grid_constant__ TopkFuse fuse) {
    extern __shared__ uint8_t smem_raw[];
    // aligned by pointer arithmetic on the shared array (an integer round trip would turn every access
    // through it into a generic-address one)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int NQ = TK_QROWS * PAIR;             // queries this CTA selects for
    constexpr int BUF = PAIR == 2 ? 16 : TR_BUF;    // candidate buffer entries per query
    constexpr int TRIG = BUF / 2;
    constexpr int CPW = PAIR;                       // 32-query chunks per selection warp
    constexpr int ACC = TR_ACC_STRIDE * PAIR;       // TMEM columns per accumulator stage
    const uint32_t cta_rank = PAIR == 2 ? cluster_ctarank() : 0u;
    const int num_kb = dim / TR_BK;
    const uint32_t sQ = smem_u32(smem);
    const uint32_t sC = sQ + num_kb * TR_QBLK_BYTES;
    uint8_t* after_ring = smem + num_kb * TR_QBLK_BYTES + stages * TR_STAGE_BYTES;
    // per-query state shared by the CTA: sorted list [128][k] (scores, rows), candidate buffer [128][TR_BUF],
    // current k-th score, number of buffered candidates
    // (all held as 32-bit shared-window addresses, see ptx.cuh)
    const uint32_t lists_s = smem_u32(after_ring);
    const uint32_t lists_i = lists_s + 4 * NQ * k;
    const uint32_t buf_s = lists_i + 4 * NQ * k;
    const uint32_t buf_i = buf_s + 4 * NQ * BUF;
    const uint32_t thrs = buf_i + 4 * NQ * BUF;
    const uint32_t cnt = thrs + 4 * NQ;
    const uint32_t gthr_s = cnt + 4 * NQ;  // shared-memory copy of the 1 + k grid-wide threshold arrays (refreshed asynchronously)
    const int n_garr = 1 + k;              // array 0: best local k-th score; arrays 1..k: best score of each row class (row % k)
    uint64_t* bars = reinterpret_cast<uint64_t*>(after_ring + 8 * NQ * k + 8 * NQ * BUF + (12 + 4 * k) * NQ);
    const uint32_t bar_full = smem_u32(bars);
    const uint32_t bar_empty = bar_full + 8 * TR_MAX_STAGES;
    const uint32_t bar_tfull = bar_empty + 8 * TR_MAX_STAGES;
    const uint32_t bar_tempty = bar_tfull + 16;
    const uint32_t bar_q = bar_tempty + 16;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(bars + 2 * TR_MAX_STAGES + 5);
    const uint32_t again_epoch = smem_u32(bars + 2 * TR_MAX_STAGES + 6);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmapQ);
        tma_prefetch_desc(&tmapC);
        for (int s = 0; s < stages; ++s) {
            mbar_init(bar_full + 8 * s, PAIR);  // the leader's: one arrival per CTA's producer
            mbar_init(bar_empty + 8 * s, 1);
        }
        const int scanning_warps = 4 * (((nq + 31) / 32 + CPW - 1) / CPW);
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, scanning_warps * PAIR);  // one arrival per scanning warp (of both CTAs)
        }
        mbar_init(bar_q, PAIR);
        fence_mbar_init();
    }
    if (warp == 1) {
        if (PAIR == 2) {
            tmem_alloc_pair(smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
            tmem_relinquish_pair();
        } else {
            tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), 256);
            tmem_relinquish();
        }
    }
    for (int i = threadIdx.x; i < NQ * k; i += TR_THREADS) {
        sts_f32(lists_s + 4 * i, -INFINITY);
        sts_s32(lists_i + 4 * i, -1);
    }
    for (int i = threadIdx.x; i < NQ; i += TR_THREADS) {
        sts_f32(thrs + 4 * i, -INFINITY);
        sts_s32(cnt + 4 * i, 0);
        for (int a = 0; a < n_garr; ++a) sts_s32(gthr_s + 4 * (a * NQ + i), TKX_KEY_NEG_INF);
    }
    if (threadIdx.x == 0) sts_s32(again_epoch, -1);
    tc_fence_before();
    if (PAIR == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // a tile = 128 * PAIR corpus rows, one per CTA (pair); this CTA's rows are its rank's half
    const int num_tiles = (n_rows + TR_BM * PAIR - 1) / (TR_BM * PAIR);
    const int tile0 = blockIdx.x / PAIR, tile_step = gridDim.x / PAIR;
    // Grid-wide thresholds by ROW CLASS.  Array 1 + c of fuse.gthr holds, per query, the best score any CTA has seen
    // on a row with row % k == c (one atomic max per improving hit).  Rows of different classes are different rows,
    // so the minimum over the k classes is a score that k distinct rows of the corpus reach: nothing strictly below
    // it can make the top k.  It sits at about the (k ln k)-th best score of everything the GRID has seen so far -
    // the best local k-th score (array 0, raised at compactions) is only what ONE CTA's share of the rows gives,
    // two orders of magnitude further down the distribution on a short scan.
    // Tiles of this CTA (pair): tile0, tile0 + step, ...  With SAMPLING the first tile is visited twice: first
    // only looked at - every warp publishes the best of its 32 rows per query into that row's class - and at
    // the very end scanned like any other tile.  Without it every CTA starts from -inf and spends its first
    // tiles inserting rows that cannot make the top k (the fixed cost that dominated short scans: 148 CTAs x k
    // candidates per query); with it the first collecting tile already rejects all but a fraction of a percent.
    const int n_my = tile0 < num_tiles ? (num_tiles - 1 - tile0) / tile_step + 1 : 0;
    const bool sampling = fuse.gthr != nullptr && gridDim.x >= 4 && n_my >= 2;
    const int n_it = n_my + (sampling ? 1 : 0);
    auto tile_of = [=](int it) { return (sampling && it == n_it - 1) ? tile0 : tile0 + it * tile_step; };

    if (warp == 0) {
        if (elect_one()) {
            const uint32_t lead_q = PAIR == 2 ? mapa_u32(bar_q, 0) : bar_q;
            if (PAIR == 2) {
                // each CTA keeps its half of the 256 queries: rows [q_row0 + 128 * rank, + 128)
                if (cta_rank == 0) mbar_arrive_expect_tx(bar_q, num_kb * TR_QBLK_BYTES * 2); else mbar_arrive_cluster(lead_q);
                for (int kb = 0; kb < num_kb; ++kb)
                    tma_load_2d_pair(sQ + kb * TR_QBLK_BYTES, &tmapQ, lead_q, kb * TR_BK,
                                     q_row0 + static_cast<int>(cta_rank) * TK_QROWS);
            } else {
                mbar_arrive_expect_tx(bar_q, num_kb * TR_QBLK_BYTES);
                for (int kb = 0; kb < num_kb; ++kb)
                    tma_load_2d(sQ + kb * TR_QBLK_BYTES, &tmapQ, bar_q, kb * TR_BK, q_row0);
            }
            const uint64_t pol = policy_evict_first();
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < n_it; ++it) {
                const int tile = tile_of(it);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    if (PAIR == 2) {
                        const uint32_t lf = mapa_u32(bar_full + 8 * stage, 0);
                        if (cta_rank == 0) mbar_arrive_expect_tx(bar_full + 8 * stage, TR_STAGE_BYTES * 2); else mbar_arrive_cluster(lf);
                        tma_load_2d_pair_hint(sC + stage * TR_STAGE_BYTES, &tmapC, lf, kb * TR_BK,
                                              (tile * 2 + static_cast<int>(cta_rank)) * TR_BM, pol);
                    } else {
                        mbar_arrive_expect_tx(bar_full + 8 * stage, TR_STAGE_BYTES);
                        tma_load_2d_hint(sC + stage * TR_STAGE_BYTES, &tmapC, bar_full + 8 * stage, kb * TR_BK,
                                         tile * TR_BM, pol);
                    }
                    if (++stage == stages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (cta_rank == 0 && elect_one()) {  // the pair's leader issues all MMAs
            mbar_wait(bar_q, 0);
            tc_fence_after();
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int it = 0; it < n_it; ++it) {
                mbar_wait(bar_tempty + 8 * acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * ACC;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after();
                    const uint64_t da = umma_desc_sw128(sC + stage * TR_STAGE_BYTES);  // corpus rows: M
                    const uint64_t db = umma_desc_sw128(sQ + kb * TR_QBLK_BYTES);      // queries: N
#pragma unroll
                    for (int kk = 0; kk < TR_BK / 16; ++kk) {
                        if (PAIR == 2)
                            umma_bf16_pair(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
                        else
                            umma_bf16(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
                    }
                    if (PAIR == 2) umma_commit_pair(bar_empty + 8 * stage); else umma_commit(bar_empty + 8 * stage);
                    if (++stage == stages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                if (PAIR == 2) umma_commit_pair(bar_tfull + 8 * acc); else umma_commit(bar_tfull + 8 * acc);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else {
        // ===== selection: warp (quadrant, chunk group); lane = corpus row, columns = 32 queries per chunk =====
        const int quad = warp & 3;
        const int sw = warp - 2;   // selection warp 0..15
        const int cg = sw >> 2;    // chunk group: this warp scans chunks cg * CPW .. cg * CPW + CPW - 1
        const int nchunks = (nq + 31) / 32;
        const bool scans = cg * CPW < nchunks;
        // the queries this warp compacts and whose thresholds it maintains: `per` consecutive ones, lane l owns the l-th
        // (consecutive words in shared memory: the 1 + k threshold reads per tile are conflict-free)
        const int per = (nq + TR_SEL_WARPS - 1) / TR_SEL_WARPS;
        const int myq = sw * per + lane;
        const bool owner = lane < per && myq < nq;
        int32_t* const gthr = fuse.gthr != nullptr ? fuse.gthr + q_row0 : nullptr;
        const uint32_t lead_tempty = PAIR == 2 ? mapa_u32(bar_tempty, 0) : bar_tempty;
        int acc = 0;
        uint32_t acc_phase = 0;
        int epoch = 0;
        for (int it = 0; it < n_it; ++it) {
            const int tile = tile_of(it);
            const int32_t row = (tile * PAIR + static_cast<int>(cta_rank)) * TR_BM + quad * 32 + lane;
            const uint32_t tbase = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * ACC;
            if (sampling && it == 0) {
                // ---- sampling pass over the first tile: nothing is collected ----
                if (scans) {
                    mbar_wait(bar_tfull + 8 * acc, acc_phase);
                    tc_fence_after();
#pragma unroll
                    for (int cc = 0; cc < CPW; ++cc) {
                        const int c = cg * CPW + cc;
                        if (c >= nchunks) continue;
                        uint32_t v[32];
                        tmem_ld_32x32(tbase + c * 32, v);
                        tmem_ld_wait();
                        int32_t best = TKX_KEY_NEG_INF;  // lane j keeps the best score of this warp's 32 rows for query c*32+j ...
                        int best_lane = 0;               // ... and the lane (= row) that holds it
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int32_t key = row < n_rows ? score_key(__uint_as_float(v[j])) : TKX_KEY_NEG_INF;
                            const int32_t r = __reduce_max_sync(0xffffffffu, key);
                            const uint32_t who = __ballot_sync(0xffffffffu, key == r);
                            if (lane == j) {
                                best = r;
                                best_lane = __ffs(who) - 1;
                            }
                        }
                        if (c * 32 + lane < nq && best > TKX_KEY_NEG_INF) {
                            const int cls = (row - lane + best_lane) % k;
                            atomicMax(gthr + (1 + cls) * fuse.gstride + c * 32 + lane, best);
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (PAIR == 2) mbar_arrive_cluster(lead_tempty + 8 * acc); else mbar_arrive(bar_tempty + 8 * acc);
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(32 * TR_SEL_WARPS) : "memory");
                if (owner) {
                    // the other CTAs are at the same point: give their values a moment to arrive (bounded), so
                    // that the first collecting tile starts from a real threshold
                    const int32_t* const g0 = gthr + myq;
                    int32_t t_key = TKX_KEY_NEG_INF;
                    for (int tries = 0; tries < 48 && t_key == TKX_KEY_NEG_INF; ++tries) {
                        int32_t cv[TKX_MAX_K];  // all loads in flight together: one L2 round trip per try
#pragma unroll
                        for (int a = 0; a < TKX_MAX_K; ++a)
                            cv[a] = a < k ? ld_relaxed_gpu_s32(g0 + (1 + a) * fuse.gstride) : 0x7fffffff;
                        int32_t lo = ld_relaxed_gpu_s32(g0);
                        int32_t cl = 0x7fffffff;
#pragma unroll
                        for (int a = 0; a < TKX_MAX_K; ++a) cl = min(cl, cv[a]);
                        t_key = max(lo, cl);
                    }
                    if (t_key > TKX_KEY_NEG_INF) sts_f32(thrs + 4 * myq, key_score(t_key - 1));
                }
                asm volatile("bar.sync 1, %0;" ::"n"(32 * TR_SEL_WARPS) : "memory");
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
                continue;
            }
            uint32_t pend[CPW];  // (row, query) pairs whose append found the buffer full
#pragma unroll
            for (int cc = 0; cc < CPW; ++cc) pend[cc] = 0u;
            bool released = !scans;
            // The other CTAs' k-th best scores: every warp copies its slice of the grid-wide thresholds into
            // shared memory with cp.async (L2 -> shared, no register in between, nobody waits for it) and the
            // compaction phase reads whatever copy has landed.  A register load here costs an L2 round trip per
            // tile on the selection warps' critical path (measured: 2x the scan time at 64 queries); a stale
            // threshold only admits a few more candidates.
            // (every tile while the thresholds still move fast, every fourth one afterwards)
            if (gthr != nullptr && (it < 8 || (it & 3) == 0)) {
                constexpr int LPA = NQ / (TR_SEL_WARPS * 4);  // 16-byte pieces of this warp's slice of one array
                for (int p = lane; p < n_garr * LPA; p += 32) {
                    const int a = p / LPA;
                    const int i4 = sw * (NQ / TR_SEL_WARPS) + (p % LPA) * 4;
                    // only the pieces that hold queries of this batch: all CTAs read the same few lines every tile, and
                    // at one query 352 requests per CTA and tile on 44 hot lines slowed the whole scan by 15 %
                    if (i4 < nq)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(gthr_s + 4 * (a * NQ + i4)),
                                     "l"(gthr + a * fuse.gstride + i4)
                                     : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
            const int cls = row % k;  // this lane's row class
            if (scans) {
                mbar_wait(bar_tfull + 8 * acc, acc_phase);
                tc_fence_after();
            }
            for (int round = 0;; ++round, ++epoch) {
                // ---- append phase ----
                bool mine_pending = false;
#pragma unroll
                for (int cc = 0; cc < CPW; ++cc) {
                    const int c = cg * CPW + cc;
                    if (c >= nchunks) continue;
                    if (round > 0 && !__any_sync(0xffffffffu, pend[cc] != 0u)) continue;  // warp-uniform
                    const uint32_t tcol = tbase + c * 32;
                    const int qlim = nq - c * 32;  // queries of this chunk that exist
                    const uint32_t qmask = qlim >= 32 ? 0xffffffffu : ((1u << qlim) - 1u);
                    uint32_t v[32];
                    tmem_ld_32x32(tcol, v);
                    tmem_ld_wait();
                    const uint32_t tp = thrs + 4 * c * 32;  // broadcast reads
                    uint32_t m = 0;
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 t = lds_f32x4(tp + 16 * j4);
                        m |= (__uint_as_float(v[j4 * 4 + 0]) > t.x ? 1u : 0u) << (j4 * 4 + 0);
                        m |= (__uint_as_float(v[j4 * 4 + 1]) > t.y ? 1u : 0u) << (j4 * 4 + 1);
                        m |= (__uint_as_float(v[j4 * 4 + 2]) > t.z ? 1u : 0u) << (j4 * 4 + 2);
                        m |= (__uint_as_float(v[j4 * 4 + 3]) > t.w ? 1u : 0u) << (j4 * 4 + 3);
                    }
                    m &= row < n_rows ? qmask : 0u;
                    if (round > 0) m &= pend[cc];
                    uint32_t full = 0u;
                    uint32_t any = __reduce_or_sync(0xffffffffu, m);
                    while (any) {  // warp-uniform; rare once the lists have warmed up
                        const int j = __ffs(any) - 1;
                        any &= any - 1;
                        const uint32_t sc = tmem_ld_32x1(tcol + j);  // this lane's score against query c*32+j
                        tmem_ld_wait();
                        // (a warp-aggregated slot reservation - one atomic per warp and query - was tried: it costs
                        // more on the common single-hit iteration than it saves on the first, crowded tiles)
                        if ((m >> j) & 1u) {
                            const int q = c * 32 + j;
                            const int pos = atoms_add_s32(cnt + 4 * q, 1);
                            if (pos < BUF) {
                                sts_f32(buf_s + 4 * (q * BUF + pos), __uint_as_float(sc));
                                sts_s32(buf_i + 4 * (q * BUF + pos), row);
                            } else {
                                full |= 1u << j;
                            }
                            if (gthr != nullptr) {  // a new best row of its class: tell the grid (and remember having told it)
                                const int32_t key = score_key(__uint_as_float(sc));
                                const uint32_t gs = gthr_s + 4 * ((1 + cls) * NQ + q);
                                if (key > lds_s32(gs)) {
                                    sts_s32(gs, key);
                                    atomicMax(gthr + (1 + cls) * fuse.gstride + q, key);
                                }
                            }
                        }
                    }
                    pend[cc] = full;
                    mine_pending = mine_pending || full != 0u;
                }
                if (mine_pending) sts_s32(again_epoch, epoch);
                if (!released && !__any_sync(0xffffffffu, mine_pending)) {
                    // this warp is done with its part of the tile's accumulator
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (PAIR == 2) mbar_arrive_cluster(lead_tempty + 8 * acc); else mbar_arrive(bar_tempty + 8 * acc);
                    }
                    released = true;
                }
                asm volatile("bar.sync 1, %0;" ::"n"(32 * TR_SEL_WARPS) : "memory");
                // ---- compaction phase: buffers are stable until the second barrier ----
                const bool again = lds_s32(again_epoch) == epoch;
                const int n_my = (owner) ? lds_s32(cnt + 4 * myq) : 0;
                uint32_t need = __ballot_sync(0xffffffffu, n_my >= TRIG);
                while (need) {
                    const int l = __ffs(need) - 1;
                    need &= need - 1;
                    const int q = sw * per + l;
                    int n = __shfl_sync(0xffffffffu, n_my, l);
                    n = n < BUF ? n : BUF;
                    compact_query(lists_s + 4 * q * k, lists_i + 4 * q * k, buf_s + 4 * q * BUF, buf_i + 4 * q * BUF,
                                  thrs + 4 * q, k, n, gthr != nullptr ? gthr + q : nullptr);
                    if (lane == 0) sts_s32(cnt + 4 * q, 0);
                }
                int32_t gkey = TKX_KEY_NEG_INF;
                if (gthr != nullptr && owner) {
                    const uint32_t gp = gthr_s + 4 * myq;
                    int32_t cv[TKX_MAX_K];  // independent reads first, then the minimum
#pragma unroll
                    for (int a = 0; a < TKX_MAX_K; ++a) cv[a] = a < k ? lds_s32(gp + 4 * (1 + a) * NQ) : 0x7fffffff;
                    int32_t lo = 0x7fffffff;
#pragma unroll
                    for (int a = 0; a < TKX_MAX_K; ++a) lo = min(lo, cv[a]);
                    gkey = max(lds_s32(gp), lo);
                }
                if (gkey > TKX_KEY_NEG_INF) {
                    // k rows score >= key_score(gkey) somewhere in the grid: nothing strictly below it can make the
                    // top k.  One ulp under it, so that rows EQUAL to it still get through (ties go by row number,
                    // which only the merge can settle).
                    const float g = key_score(gkey - 1);
                    if (g > lds_f32(thrs + 4 * myq)) sts_f32(thrs + 4 * myq, g);
                }
                asm volatile("bar.sync 1, %0;" ::"n"(32 * TR_SEL_WARPS) : "memory");
                if (!again) break;  // CTA-uniform
            }
            ++epoch;
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
        // ---- flush what is left in the buffers, then the CTA's lists are final ----
        {
            const int n_my = (owner) ? lds_s32(cnt + 4 * myq) : 0;
            uint32_t need = __ballot_sync(0xffffffffu, n_my > 0);
            while (need) {
                const int l = __ffs(need) - 1;
                need &= need - 1;
                const int q = sw * per + l;
                int n = __shfl_sync(0xffffffffu, n_my, l);
                n = n < BUF ? n : BUF;
                compact_query(lists_s + 4 * q * k, lists_i + 4 * q * k, buf_s + 4 * q * BUF, buf_i + 4 * q * BUF,
                              thrs + 4 * q, k, n, nullptr);
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(32 * TR_SEL_WARPS) : "memory");
        // candidates as [query block][CTA][128][k] (the layout topk_merge_kernel reads)
        for (int i = sw * 32 + lane; i < nq * k; i += 32 * TR_SEL_WARPS) {
            const int q = i / k, pp = i % k;
            const size_t base =
                ((static_cast<size_t>(q >> 7) * gridDim.x + blockIdx.x) * TK_QROWS + static_cast<size_t>(q & 127)) * k;
            const int32_t r = lds_s32(lists_i + 4 * i);
            cand_scores[base + pp] = r >= 0 ? lds_f32(lists_s + 4 * i) : -INFINITY;
            cand_idx[base + pp] = r;
        }
    }

    tc_fence_before();
    if (PAIR == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (PAIR == 2) tmem_dealloc_pair(tmem_base, 512); else tmem_dealloc(tmem_base, 256);
    }
    // merge of the per-CTA lists (and, for a sharded corpus, the exchange with the peer GPUs) in this launch
    if (fuse.done != nullptr) {
        const unsigned parts = gridDim.x;
        // scratch of the tail: the candidate buffers (16 KB, dead since the lists were written out)
        static_assert(sizeof(TopkTailSmem) <= 4 * TK_QROWS * 16, "tail scratch must fit the candidate buffers");
        TopkTailSmem& tail_sm = *reinterpret_cast<TopkTailSmem*>(after_ring + 8 * NQ * k);
        topk_fused_tail(fuse, cand_scores, cand_idx, q_row0, nq, k, tail_sm, [parts, k](int p, int q) {
            return ((static_cast<size_t>(q >> 7) * parts + p) * TK_QROWS + static_cast<size_t>(q & 127)) * k;
        });
    }
}

}  // namespace

// fixed shared memory of the kernel: resident query block, per-query lists / buffers / thresholds / counters,
// barriers, alignment slack
static int rows_fixed_bytes(int dim, int k, int pair) {
    const int nq = TK_QROWS * pair, buf = pair == 2 ? 16 : TR_BUF;
    return (dim / TR_BK) * TR_QBLK_BYTES + nq * k * 8 + nq * buf * 8 + nq * (12 + 4 * k) + 256 + 1024;
}

int topk_rows_preload_kernels() {
    cudaFuncAttributes a;
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_rows_kernel<1>));
    LSC_CUDA(cudaFuncGetAttributes(&a, topk_scan_rows_kernel<2>));
    return LSC_OK;
}

int topk_scan_rows_grid(int n_rows, int num_sms) {
    const int tiles = (n_rows + TR_BM - 1) / TR_BM;
    return tiles < num_sms ? tiles : num_sms;
}

bool topk_scan_rows_supported(int dim, int k) {
    static const bool legacy = getenv("LSC_TOPK_LEGACY") != nullptr;  // debug: force the query-major kernel
    if (legacy || k > 16 || dim % TR_BK != 0 || dim <= 0 || dim > 512) return false;
    return (227 * 1024 - rows_fixed_bytes(dim, k, 1)) / TR_STAGE_BYTES >= 3;
}

// CTA pairs: 256 queries per pass.  LSC_TOPK_PAIR=0 (debug) keeps 128-query passes.
bool topk_scan_rows_pair_supported(int dim, int k, int num_sms) {
    static const bool off = [] {
        const char* e = getenv("LSC_TOPK_PAIR");
        return e != nullptr && atoi(e) == 0;
    }();
    if (off || num_sms < 2 || !topk_scan_rows_supported(dim, k)) return false;
    return (227 * 1024 - rows_fixed_bytes(dim, k, 2)) / TR_STAGE_BYTES >= 3;
}

int topk_scan_rows_pair_grid(int n_rows, int num_sms) {
    const int tile_pairs = (n_rows + 2 * TR_BM - 1) / (2 * TR_BM);
    const int clusters = tile_pairs < num_sms / 2 ? tile_pairs : num_sms / 2;
    return 2 * clusters;
}

template <int PAIR>
static int rows_launch(const CUtensorMap& tmapQ, const CUtensorMap& tmapC128, int q_row0, int nq, int n_rows, int dim,
                       int k, int grid, float* cand_scores, int32_t* cand_idx, const TopkFuse& fuse, cudaStream_t stream) {
    if (fuse.done != nullptr && (grid > TKX_THREADS || k > TKX_MAX_K))
        return fail(LSC_E_UNSUPPORTED, "top-k (row-major): fused merge needs grid <= 256 and k <= 16");
    const int fixed = rows_fixed_bytes(dim, k, PAIR);
    int stages = (227 * 1024 - fixed) / TR_STAGE_BYTES;
    if (stages > TR_MAX_STAGES) stages = TR_MAX_STAGES;
    if (stages < 3) return fail(LSC_E_UNSUPPORTED, "top-k (row-major): unsupported dim / k");
    const size_t smem = static_cast<size_t>(fixed) + static_cast<size_t>(stages) * TR_STAGE_BYTES;
    int n_mma = (nq + 15) / 16 * 16;
    if (n_mma < 16) n_mma = 16;
    if (PAIR == 2) n_mma = 256;  // each CTA supplies 128 query rows (zero rows past the end of the batch)
    const uint32_t idesc = umma_idesc_bf16(TR_BM * PAIR, static_cast<uint32_t>(n_mma));
    static bool configured[64] = {};
    auto kern = topk_scan_rows_kernel<PAIR>;
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev]) {
        LSC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(grid);
    lc.blockDim = dim3(TR_THREADS);
    lc.dynamicSmemBytes = smem;
    lc.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PAIR;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    // The fused tail waits for every CTA of the grid: a cooperative launch makes them all resident at once, also
    // when another handle's scan runs on the same GPU.  Shards that share one GPU on purpose (sharded_index.cu,
    // LSC_TOPK_MAX_CTAS) split the SMs between their grids instead and launch plainly, because their scans wait
    // for EACH OTHER's candidates and must run side by side.
    attr[1].id = cudaLaunchAttributeCooperative;
    attr[1].val.cooperative = 1;
    lc.attrs = attr;
    lc.numAttrs = (fuse.done != nullptr && fuse.cooperative != 0) ? 2 : 1;
    LSC_CUDA(cudaLaunchKernelEx(&lc, kern, tmapQ, tmapC128, q_row0, nq, n_rows, dim, k, stages, idesc, cand_scores, cand_idx,
                                fuse));
    LSC_CUDA(cudaGetLastError());
    return 1;
}

// tmapC here must have a box of 128 rows x 64 columns (the query-major kernel uses 256-row boxes).
int topk_scan_rows_launch(const CUtensorMap& tmapQ, const CUtensorMap& tmapC128, int q_row0, int nq, int n_rows,
                          int dim, int k, int grid, float* cand_scores, int32_t* cand_idx, const TopkFuse& fuse,
                          cudaStream_t stream) {
    if (!topk_scan_rows_supported(dim, k)) return fail(LSC_E_UNSUPPORTED, "top-k (row-major): unsupported dim / k");
    return rows_launch<1>(tmapQ, tmapC128, q_row0, nq, n_rows, dim, k, grid, cand_scores, cand_idx, fuse, stream);
}

// Up to 256 queries [q_row0, q_row0 + nq) per pass on CTA pairs; the query matrix behind tmapQ must have rows up
// to q_row0 + 256 (zeros past the batch) or end before them (TMA fills with zeros).  Candidates are written as
// two query blocks: cand + (qb * grid + cta) * 128 * k for qb = 0, 1.
int topk_scan_rows_pair_launch(const CUtensorMap& tmapQ, const CUtensorMap& tmapC128, int q_row0, int nq, int n_rows,
                               int dim, int k, int grid, float* cand_scores, int32_t* cand_idx, const TopkFuse& fuse,
                               cudaStream_t stream) {
    if (grid < 2 || grid % 2 != 0) return fail(LSC_E_INVALID, "top-k (row-major, pairs): grid must be even");
    return rows_launch<2>(tmapQ, tmapC128, q_row0, nq, n_rows, dim, k, grid, cand_scores, cand_idx, fuse, stream);
}

}  // namespace lsc
