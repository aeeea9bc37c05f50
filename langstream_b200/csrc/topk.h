// Launchers of the top-k kernels (topk.cu), used by the index host code (index.cu).
#pragma once
#include "common.h"
#include "topk_fuse.cuh"

namespace lsc {

constexpr int TK_QROWS = 128;  // queries per scan pass (UMMA M)
constexpr int TK_KMAX = 128;   // largest supported k

// Number of CTAs the scan uses for n_rows rows on a device with num_sms SMs.
int topk_scan_grid(int n_rows, int num_sms);

// Scan all rows for the query block [q_row0, q_row0 + nq) of the bf16 query matrix behind tmapQ.
// Scratch written: cand_scores / cand_idx [grid][128][k] (only rows q < nq).
int topk_scan_launch(const CUtensorMap& tmapQ, const CUtensorMap& tmapC, int q_row0, int nq, int n_rows,
                     int dim, int k, int grid, float* cand_scores, int32_t* cand_idx,
                     cudaStream_t stream);

// Row-major selection kernel (topk_rows.cu), k <= 16: same scratch format; tmapC128 must have 128-row boxes.
bool topk_scan_rows_supported(int dim, int k);
int topk_scan_rows_grid(int n_rows, int num_sms);
// `fuse` (topk_fuse.cuh): thresholds shared between the CTAs, and - when fuse.done is set - the merge of the
// per-CTA lists (plus the exchange with peer GPUs) inside the same launch; a default-constructed TopkFuse gives
// the plain scan that only writes the scratch.
int topk_scan_rows_launch(const CUtensorMap& tmapQ, const CUtensorMap& tmapC128, int q_row0, int nq, int n_rows,
                          int dim, int k, int grid, float* cand_scores, int32_t* cand_idx, const TopkFuse& fuse,
                          cudaStream_t stream);

// CTA-pair variant: up to 256 queries per pass (see topk_rows.cu); grid from topk_scan_rows_pair_grid.
bool topk_scan_rows_pair_supported(int dim, int k, int num_sms);
int topk_scan_rows_pair_grid(int n_rows, int num_sms);
int topk_scan_rows_pair_launch(const CUtensorMap& tmapQ, const CUtensorMap& tmapC128, int q_row0, int nq, int n_rows,
                               int dim, int k, int grid, float* cand_scores, int32_t* cand_idx, const TopkFuse& fuse,
                               cudaStream_t stream);

// Merge the scan scratch [qblocks][parts][128][k] into [B][k]; slots are mapped through slot_ids.
int topk_merge_scan_launch(const float* cand_scores, const int32_t* cand_idx, int parts, int B, int k,
                           const int64_t* slot_ids, int64_t* out_ids, float* out_scores,
                           cudaStream_t stream);

// Merge shard lists [parts][B][k] (int64 ids) into [B][k].
int topk_merge_shards_launch(const float* scores, const int64_t* ids, int parts, int B, int k,
                             int64_t* out_ids, float* out_scores, cudaStream_t stream);

// fp32 rows -> L2-normalised bf16 rows (see topk.cu).  When given, the launch also resets the state the fused
// scan needs: gthr[0 .. n_gthr) to "no threshold yet" and done[0 .. n_done) to zero.
int normalize_rows_launch(const float* src, int n, int n_pad, int dim, const int64_t* dst_rows,
                          __nv_bfloat16* dst, cudaStream_t stream, int32_t* gthr = nullptr, int n_gthr = 0,
                          unsigned* done = nullptr, int n_done = 0);

// Stand-alone form of the fused tail's exchange for shard results that already exist as [B][k] lists (the
// query-major scan, 16 < k <= 32): sends them to the peers, waits for theirs and writes the merged [B][k].
int topk_exchange_launch(const int64_t* local_ids, const float* local_scores, int B, int k, const TopkFuse& fuse,
                         cudaStream_t stream);

int scores_naive_launch(const __nv_bfloat16* rows, int64_t n, int dim, const __nv_bfloat16* q, int B,
                        float* out, cudaStream_t stream);

// rows[moves[2j]] = rows[moves[2j+1]] (and the ids likewise) for j < m: the batched form of "move the last
// row into the freed slot"; sources and destinations must not overlap.
int move_rows_launch(__nv_bfloat16* rows, int64_t* ids, int dim, const int64_t* moves, int m, cudaStream_t stream);

// Force-load the top-k kernels into the current device's context (see topk.cu).
int topk_preload_kernels();
int topk_rows_preload_kernels();

int fill_empty_launch(int64_t* ids, float* scores, size_t n, cudaStream_t stream);

}  // namespace lsc
