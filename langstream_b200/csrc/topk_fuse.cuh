// What the row-major scan kernel does after its last corpus tile (topk_rows.cu), shared with the stand-alone
// exchange kernel (topk.cu):
//   1. thresholds shared between CTAs while scanning: every CTA publishes the k-th best score of each query it
//      has found so far with one atomic max per compaction, and reads the others' once per tile, so no CTA keeps
//      collecting rows that 147 other CTAs have already beaten;
//   2. the merge of the per-CTA candidate lists inside the scan kernel, behind a grid-wide arrival counter
//      (the kernel is launched cooperatively: all its CTAs are resident), instead of a second launch;
//   3. for a row-sharded corpus (one shard per GPU), the exchange of the per-shard top-k and the final merge
//      in the same kernel: the CTA that merged query q writes its k (score, id) records straight into every
//      peer GPU's receive buffer over NVLink (plain peer-to-peer stores), raises a flag there, waits for the
//      peers' flags for q and merges the `world` lists.  This replaces the NCCL all-gather + merge launch of the
//      unfused path: at 80 B - 80 KiB per rank the collective is pure latency, which a store and a flag beat.
//      Peer memory is only ever touched by ordinary st / ld instructions here (no TMA on it).
// There is no counterpart in the single-node reference (HerdDB answers the query in one process).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace lsc {

constexpr int TKX_MAX_WORLD = 8;    // GPUs of one NVSwitch domain
constexpr int TKX_MAX_K = 16;       // the fused merge serves the row-major scan (k <= 16)
constexpr int TKX_MAX_XK = 32;      // the exchange alone (behind the query-major scan) goes up to k = 32
constexpr int TKX_THREADS = 256;    // threads of a CTA that run the tail (warps 0-7)

// One (score, id) record of the exchange: 16 bytes, written with one vector store.
struct __align__(16) TopkRecord {
    float score;
    int32_t pad;
    int64_t id;
};

// Receive side of one GPU: records [2 (epoch parity)][world (source rank)][xB][xk] and one flag per
// (parity, source, query) holding the epoch of the records behind it.
struct TopkFuse {
    // Thresholds shared by the CTAs of a scan, as ordered-int keys; nullptr = off.  1 + TKX_MAX_K arrays of `gstride`
    // entries (one per query of the batch): [0] the best k-th score any single CTA has reached; [1 + c], c < k: the
    // best score seen anywhere in the grid on a row with row % k == c (topk_rows.cu: the minimum over the k classes
    // is reached by k distinct rows).
    int32_t* gthr = nullptr;
    int32_t gstride = 0;
    unsigned* done = nullptr;        // arrival counter of THIS launch (zeroed before it); nullptr = no fused tail
    const int64_t* slot_ids = nullptr;  // storage slot -> id (nullptr: the slot number is the id)
    int64_t* out_ids = nullptr;      // [B][k] final results of this GPU
    float* out_scores = nullptr;
    int32_t* status = nullptr;       // device status word: set non-zero when a wait gave up
    int32_t world = 1, rank = 0;     // world > 1: exchange with the peers below
    uint32_t epoch = 0;              // sequence number of this search within the peer group
    int32_t xB = 0, xk = 0;          // strides of the receive buffers
    int32_t cooperative = 1;         // host side only: launch the fused scan cooperatively (see rows_launch)
    TopkRecord* recv[TKX_MAX_WORLD] = {};  // recv[r]: rank r's receive buffer as addressable from this GPU
    uint32_t* flags[TKX_MAX_WORLD] = {};
};

// float <-> int key with the same ordering (for atomicMax on scores)
__device__ __forceinline__ int32_t score_key(float f) {
    const int32_t i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float key_score(int32_t k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }
constexpr int32_t TKX_KEY_NEG_INF = static_cast<int32_t>(0xff800000u) ^ 0x7fffffff;  // score_key(-inf)

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_relaxed_gpu_s32(const int32_t* p) {
    int32_t v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void tail_bar() { asm volatile("bar.sync 2, %0;" ::"n"(TKX_THREADS) : "memory"); }

constexpr uint64_t TKX_GRID_WAIT_NS = 4ull * 1000 * 1000 * 1000;    // own CTAs: all resident, seconds = broken
constexpr uint64_t TKX_PEER_WAIT_NS = 30ull * 1000 * 1000 * 1000;   // peers may be launched much later

// Shared-memory scratch of the tail (~4.5 KB; the scan kernel lends it a dead region of its dynamic shared memory).
struct TopkTailSmem {
    float red_s[2][8];
    int32_t red_tie[2][8];
    int32_t red_part[2][8];
    float res_s[TKX_MAX_XK];
    int32_t res_slot[TKX_MAX_K];
    int64_t res_i[TKX_MAX_XK];
    float x_s[TKX_MAX_WORLD * TKX_MAX_XK];
    int64_t x_i[TKX_MAX_WORLD * TKX_MAX_XK];
};

// k-way merge of `parts` sorted candidate lists of one query (scan scratch: scores + int32 slots, thread p owns
// list p at `base`), by the first TKX_THREADS threads of the CTA.  Order: score descending, ties by lower slot.
// Every thread first pulls its whole list (k <= 16 entries) into registers, so a round is one shuffle reduction,
// one barrier and a register shift in the winning thread - no memory access on the critical path.
// Result in sm.res_s / sm.res_i (ids mapped through slot_ids; -1 / -inf past the end).
__device__ __forceinline__ void tail_merge_lists(const float* __restrict__ sc, const int32_t* __restrict__ ix, size_t base,
                                                 bool has_list, int k, const int64_t* __restrict__ slot_ids,
                                                 TopkTailSmem& sm) {
    const int t = threadIdx.x;
    constexpr int32_t NONE = 0x7fffffff;
    float ls[TKX_MAX_K];
    int32_t li[TKX_MAX_K];
#pragma unroll
    for (int j = 0; j < TKX_MAX_K; ++j) {
        ls[j] = -INFINITY;
        li[j] = NONE;
        if (has_list && j < k) {
            const int32_t slot = __ldcg(ix + base + j);
            if (slot >= 0) {
                li[j] = slot;
                ls[j] = __ldcg(sc + base + j);
            }
        }
    }
    for (int round = 0; round < k; ++round) {
        float bs = ls[0];
        int32_t bt = li[0];
        int32_t bp = t;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float os = __shfl_xor_sync(0xffffffffu, bs, o);
            const int32_t ot = __shfl_xor_sync(0xffffffffu, bt, o);
            const int32_t op = __shfl_xor_sync(0xffffffffu, bp, o);
            if (os > bs || (os == bs && (ot < bt || (ot == bt && op < bp)))) {
                bs = os;
                bt = ot;
                bp = op;
            }
        }
        const int buf = round & 1;
        if ((t & 31) == 0) {
            sm.red_s[buf][t >> 5] = bs;
            sm.red_tie[buf][t >> 5] = bt;
            sm.red_part[buf][t >> 5] = bp;
        }
        tail_bar();
        float ws = sm.red_s[buf][0];
        int32_t wt = sm.red_tie[buf][0];
        int32_t wp = sm.red_part[buf][0];
#pragma unroll
        for (int w = 1; w < TKX_THREADS / 32; ++w) {
            const float os = sm.red_s[buf][w];
            const int32_t ot = sm.red_tie[buf][w];
            const int32_t op = sm.red_part[buf][w];
            if (os > ws || (os == ws && (ot < wt || (ot == wt && op < wp)))) {
                ws = os;
                wt = ot;
                wp = op;
            }
        }
        if (t == 0) {
            sm.res_s[round] = wt == NONE ? -INFINITY : ws;  // NONE: every list exhausted, fewer than k rows exist
            sm.res_slot[round] = wt == NONE ? -1 : wt;
        }
        if (wt != NONE && t == wp) {
#pragma unroll
            for (int j = 0; j + 1 < TKX_MAX_K; ++j) {
                ls[j] = ls[j + 1];
                li[j] = li[j + 1];
            }
            ls[TKX_MAX_K - 1] = -INFINITY;
            li[TKX_MAX_K - 1] = NONE;
        }
        // buffer `buf` is rewritten two rounds on, behind the next round's barrier
    }
    tail_bar();
    if (t < k) {
        const int32_t slot = sm.res_slot[t];
        sm.res_i[t] = slot < 0 ? -1 : (slot_ids != nullptr ? slot_ids[slot] : static_cast<int64_t>(slot));
    }
    tail_bar();
}

// Send this GPU's top-k of query `qg` (sm.res_*) to every peer, wait for theirs, merge and write the final
// result.  Equal scores keep the order (source rank, position in its list): shards hold ascending row ranges, so
// that is the storage order of an unsharded index.
__device__ __forceinline__ void tail_exchange(const TopkFuse& f, int qg, int k, TopkTailSmem& sm) {
    const int t = threadIdx.x;
    const int par = static_cast<int>(f.epoch & 1u);
    const int n_cand = f.world * k;
    if (t < n_cand) {
        const int d = t / k, j = t % k;
        TopkRecord r;
        r.score = sm.res_s[j];
        r.pad = 0;
        r.id = sm.res_i[j];
        TopkRecord* dst = f.recv[d] + ((static_cast<size_t>(par) * f.world + f.rank) * f.xB + qg) * f.xk + j;
        __stcg(reinterpret_cast<int4*>(dst), *reinterpret_cast<const int4*>(&r));
        __threadfence_system();
    }
    tail_bar();
    if (t < f.world) {
        st_release_sys_u32(f.flags[t] + (static_cast<size_t>(par) * f.world + f.rank) * f.xB + qg, f.epoch);
        const unsigned* mine = f.flags[f.rank] + (static_cast<size_t>(par) * f.world + t) * f.xB + qg;
        const uint64_t t0 = global_timer_ns();
        while (ld_acquire_sys_u32(mine) != f.epoch) {
            if (global_timer_ns() - t0 > TKX_PEER_WAIT_NS) {
                if (f.status != nullptr) atomicExch(f.status, 2);
                break;
            }
        }
    }
    tail_bar();
    if (t < n_cand) {
        const int s = t / k, j = t % k;
        const TopkRecord* src = f.recv[f.rank] + ((static_cast<size_t>(par) * f.world + s) * f.xB + qg) * f.xk + j;
        const int4 raw = __ldcv(reinterpret_cast<const int4*>(src));
        TopkRecord r;
        *reinterpret_cast<int4*>(&r) = raw;
        sm.x_s[t] = r.score;
        sm.x_i[t] = r.id;
    }
    tail_bar();
    if (t < n_cand) {
        const float ms = sm.x_s[t];
        int rank = 0;
        for (int u = 0; u < n_cand; ++u) {
            const float us = sm.x_s[u];
            rank += (us > ms || (us == ms && u < t)) ? 1 : 0;
        }
        if (rank < k) {
            f.out_ids[static_cast<size_t>(qg) * k + rank] = sm.x_i[t];
            f.out_scores[static_cast<size_t>(qg) * k + rank] = ms;
        }
    }
    tail_bar();
}

// Tail of a scan CTA: count this CTA in, and - if it owns queries - wait for all CTAs' candidates, merge the lists
// of its queries and finish them (write out, or exchange with the peers).  Called by ALL threads of the CTA after
// their candidate stores; threads past TKX_THREADS leave at once.  `cand_base_of(p, q)` must give the offset of
// CTA p's list for local query q.
template <typename BaseFn>
__device__ __forceinline__ void topk_fused_tail(const TopkFuse& f, const float* __restrict__ cand_scores,
                                                const int32_t* __restrict__ cand_idx, int q_row0, int nq, int k,
                                                TopkTailSmem& sm, BaseFn cand_base_of) {
    __threadfence();  // this thread's candidate stores, before the CTA is counted in
    __syncthreads();
    const int t = threadIdx.x;
    if (t == 0) atomicAdd(f.done, 1u);
    if (t >= TKX_THREADS || static_cast<int>(blockIdx.x) >= nq) return;
    if (t == 0) {
        const uint64_t t0 = global_timer_ns();
        while (ld_acquire_gpu_u32(f.done) < gridDim.x) {
            if (global_timer_ns() - t0 > TKX_GRID_WAIT_NS) {
                if (f.status != nullptr) atomicExch(f.status, 1);
                break;
            }
        }
    }
    tail_bar();
    const int parts = static_cast<int>(gridDim.x);
    for (int q = blockIdx.x; q < nq; q += gridDim.x) {
        const bool has_list = t < parts;
        const size_t base = has_list ? cand_base_of(t, q) : 0;
        tail_merge_lists(cand_scores, cand_idx, base, has_list, k, f.slot_ids, sm);
        const int qg = q_row0 + q;
        if (f.world <= 1) {
            if (t < k) {
                f.out_ids[static_cast<size_t>(qg) * k + t] = sm.res_i[t];
                f.out_scores[static_cast<size_t>(qg) * k + t] = sm.res_s[t];
            }
            tail_bar();
        } else {
            tail_exchange(f, qg, k, sm);
        }
    }
}

}  // namespace lsc
