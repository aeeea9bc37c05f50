// lsc_topk_search_static_dev — the top-k search for callers that must not allocate device memory.
//
// Who needs this: an agent that shares its GPU with a framework which reserved (almost) all of HBM
// when it started — a JVM-side inference engine or a caching allocator in the same pod — so a
// cudaMalloc issued by this library at query time can fail with out-of-memory even though the index
// itself fits; and query threads on a latency budget, for which cudaMalloc/cudaFree are off limits
// because both synchronise the whole device and stall every stream that is running.
// The regular entry point, lsc_index_search_dev, grows its query and candidate buffers with
// cudaMalloc the first time a larger batch arrives.  This variant runs the same kernels
// (normalise queries -> the scan: topk_scan_rows_kernel for k <= 16, topk_scan_kernel above that ->
// topk_merge_kernel) but takes their scratch memory from a fixed array that is part of the library
// image, so it never calls cudaMalloc, cudaFree or any allocating library routine.  Caller-visible limits follow from the array's size: one query block
// (B <= 128), k <= 32, dim <= 512, at most 160 SMs.
//
// Things a caller must know:
//  * Every GPU that loads the library holds the array (about 5.4 MB) until the process exits.
//  * The first call on a device looks up the array's device address (cudaGetSymbolAddress) and sets
//    the kernels' shared-memory attribute; make that first call during start-up, before any stream
//    capture is begun on that device.
//  * All calls on a device share the one array and it keeps the previous call's contents.  Nothing
//    in it has to start at zero: each call rewrites the whole padded query block (rows past B as
//    zeros) and reads back only the candidate entries that the same call's scan wrote.
//  * Because of that sharing, two calls whose kernels overlap on the GPU (different streams, or
//    different host threads) race on the scratch and return garbage; keep calls on one stream or
//    order them yourself.  Host-side entry is serialised by a mutex, GPU-side execution is not.
#include <mutex>

#include "common.h"
#include "topk.h"

namespace lsc {
namespace {

constexpr int NA_MAX_B = TK_QROWS;   // 128 queries (one scan pass)
constexpr int NA_MAX_K = 32;
constexpr int NA_MAX_DIM = 512;
constexpr int NA_MAX_SMS = 160;
constexpr int NA_MAX_DEVICES = 64;

constexpr size_t NA_Q_BYTES = static_cast<size_t>(NA_MAX_B) * NA_MAX_DIM * sizeof(__nv_bfloat16);      // 128 KB
constexpr size_t NA_CAND = static_cast<size_t>(NA_MAX_SMS) * TK_QROWS * NA_MAX_K;                      // entries
constexpr size_t NA_SCORES_OFF = NA_Q_BYTES;
constexpr size_t NA_IDX_OFF = NA_SCORES_OFF + NA_CAND * sizeof(float);
constexpr size_t NA_TOTAL_BYTES = NA_IDX_OFF + NA_CAND * sizeof(int32_t);                              // ~5.4 MB

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress, never a
// device address: the kernels receive the looked-up pointer.
static __device__ __align__(256) unsigned char g_topk_static_scratch[NA_TOTAL_BYTES];

struct DeviceState {
    bool ready = false;
    unsigned char* scratch = nullptr;  // device address of g_topk_static_scratch on this device
    int num_sms = 0;
};
DeviceState g_state[NA_MAX_DEVICES];
std::mutex g_mu;

// Once per device, with that device current.
int setup_device(int dev, DeviceState** out) {
    if (dev < 0 || dev >= NA_MAX_DEVICES) return fail(LSC_E_UNSUPPORTED, "static top-k: device ordinal too large");
    DeviceState& s = g_state[dev];
    if (!s.ready) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_topk_static_scratch);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(g_topk_static_scratch)", __FILE__, __LINE__);
        if (p == nullptr) return fail(LSC_E_CUDA, "static top-k: scratch symbol has no device address");
        int sms = 0;
        LSC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        int major = 0;
        LSC_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major != 10) return fail(LSC_E_UNSUPPORTED, "langstream-cuda needs a Blackwell (sm_100) GPU; there is no fallback path");
        if (sms > NA_MAX_SMS) return fail(LSC_E_UNSUPPORTED, "static top-k: device has more SMs than the scratch array was sized for");
        s.scratch = static_cast<unsigned char*>(p);
        s.num_sms = sms;
        s.ready = true;
    }
    *out = &s;
    return LSC_OK;
}

}  // namespace
}  // namespace lsc

using namespace lsc;

extern "C" int lsc_topk_search_static_dev(const void* d_rows_bf16, const int64_t* d_row_ids, int64_t n_rows,
                                          int32_t dim, const float* d_queries, int32_t B, int32_t k,
                                          int64_t* d_out_ids, float* d_out_scores, void* stream) {
    if (d_rows_bf16 == nullptr || d_queries == nullptr || d_out_ids == nullptr || d_out_scores == nullptr)
        return fail(LSC_E_INVALID, "lsc_topk_search_static_dev: null pointer");
    if (B < 1 || B > NA_MAX_B)
        return fail(LSC_E_UNSUPPORTED, "lsc_topk_search_static_dev: B must be in [1,128] (fixed scratch); use lsc_index_search_dev for larger batches");
    if (k < 1 || k > NA_MAX_K)
        return fail(LSC_E_UNSUPPORTED, "lsc_topk_search_static_dev: k must be in [1,32] (fixed scratch); use lsc_index_search_dev for larger k");
    if (dim <= 0 || dim % 64 != 0 || dim > NA_MAX_DIM)
        return fail(LSC_E_UNSUPPORTED, "lsc_topk_search_static_dev: dim must be a multiple of 64 and <= 512");
    if (n_rows < 1 || n_rows > 0x7fffff00LL)
        return fail(LSC_E_UNSUPPORTED, "lsc_topk_search_static_dev: n_rows must be in [1, 2^31)");
    if ((reinterpret_cast<uintptr_t>(d_rows_bf16) & 15) != 0)
        return fail(LSC_E_INVALID, "lsc_topk_search_static_dev: rows must be 16-byte aligned");

    std::lock_guard<std::mutex> lock(g_mu);
    int dev = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    DeviceState* s = nullptr;
    LSC_TRY(setup_device(dev, &s));
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    __nv_bfloat16* d_q = reinterpret_cast<__nv_bfloat16*>(s->scratch);
    float* d_cand_scores = reinterpret_cast<float*>(s->scratch + NA_SCORES_OFF);
    int32_t* d_cand_idx = reinterpret_cast<int32_t*>(s->scratch + NA_IDX_OFF);

    // Tensor maps are plain host structs describing the caller's rows and the query block inside the
    // scratch array; encoding them touches no device memory.
    CUtensorMap tmapQ, tmapC;
    LSC_TRY(make_tmap_bf16_sw128(&tmapQ, d_q, TK_QROWS, static_cast<uint64_t>(dim), TK_QROWS));
    LSC_TRY(make_tmap_bf16_sw128(&tmapC, d_rows_bf16, static_cast<uint64_t>(n_rows), static_cast<uint64_t>(dim), 256));

    const int n = static_cast<int>(n_rows);
    const bool rows_kernel = topk_scan_rows_supported(dim, k);
    if (rows_kernel)  // the row-major scan reads the corpus through 128-row boxes
        LSC_TRY(make_tmap_bf16_sw128(&tmapC, d_rows_bf16, static_cast<uint64_t>(n_rows), static_cast<uint64_t>(dim), 128));
    const int grid = rows_kernel ? topk_scan_rows_grid(n, s->num_sms) : topk_scan_grid(n, s->num_sms);
    int rc = normalize_rows_launch(d_queries, B, TK_QROWS, dim, nullptr, d_q, st);  // rewrites all 128 rows
    if (rc < 0) return rc;
    const TopkFuse plain;  // the scan only writes its scratch; the merge below is a launch of its own
    rc = rows_kernel ? topk_scan_rows_launch(tmapQ, tmapC, 0, B, n, dim, k, grid, d_cand_scores, d_cand_idx, plain, st)
                     : topk_scan_launch(tmapQ, tmapC, 0, B, n, dim, k, grid, d_cand_scores, d_cand_idx, st);
    if (rc < 0) return rc;
    rc = topk_merge_scan_launch(d_cand_scores, d_cand_idx, grid, B, k, d_row_ids, d_out_ids, d_out_scores, st);
    if (rc < 0) return rc;
    return LSC_OK;
}
