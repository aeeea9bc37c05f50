// Library-level pieces of the C ABI: version, thread-local error text, tensor-map encoding and the
// debug GEMM hook.
#include <cstdio>
#include <cstring>
#include <mutex>

#include "common.h"

namespace lsc {

namespace {
thread_local std::string g_last_error;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
std::once_flag g_encode_once;

void load_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) g_encode = reinterpret_cast<EncodeTiledFn>(fn);
}
}  // namespace

void set_error(const std::string& msg) { g_last_error = msg; }

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

int fail_cuda(cudaError_t e, const char* what, const char* file, int line) {
    char buf[512];
    snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", static_cast<int>(e), cudaGetErrorString(e),
             file, line, what);
    g_last_error = buf;
    return e == cudaErrorMemoryAllocation ? LSC_E_NOMEM : LSC_E_CUDA;
}

static int make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                          uint32_t box_cols, CUtensorMapSwizzle swizzle) {
    std::call_once(g_encode_once, load_encode);
    if (g_encode == nullptr)
        return fail(LSC_E_CUDA, "cuTensorMapEncodeTiled is not available from the CUDA driver");
    if (rows == 0 || cols == 0 || cols % box_cols != 0 || box_rows == 0 || box_rows > 256)
        return fail(LSC_E_INVALID, "tensor map: bad shape");
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0)
        return fail(LSC_E_INVALID, "tensor map: base pointer must be 16-byte aligned");
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {cols * 2};  // bytes between rows
    const cuuint32_t box[2] = {box_cols, box_rows};
    const cuuint32_t estride[2] = {1, 1};
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride,
                          box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[128];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
        return fail(LSC_E_CUDA, buf);
    }
    return LSC_OK;
}

int make_tmap_bf16_sw128(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                         uint32_t box_rows) {
    return make_tmap_bf16(out, base, rows, cols, box_rows, 64, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_bf16_box(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                       uint32_t box_cols, int swizzle_bytes) {
    const CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                  : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                  : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
    return make_tmap_bf16(out, base, rows, cols, box_rows, box_cols, sw);
}

int make_tmap_bf16_store32(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols) {
    return make_tmap_bf16(out, base, rows, cols, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B);
}

}  // namespace lsc

using namespace lsc;

extern "C" {

const char* lsc_version(void) { return "langstream-cuda 0.1 sm_100a"; }

const char* lsc_last_error(void) { return g_last_error.c_str(); }

int lsc_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        fail_cuda(e, "cudaGetDeviceCount", __FILE__, __LINE__);
        return LSC_E_CUDA;
    }
    return n;
}

int lsc_debug_gemm_dev(const void* d_A, const void* d_W, const float* d_bias, const void* d_residual,
                       void* d_C, int32_t M, int32_t N, int32_t K, int32_t epilogue, int32_t impl,
                       void* stream) {
    if (d_A == nullptr || d_W == nullptr || d_bias == nullptr || d_C == nullptr || M <= 0 || N <= 0 || K <= 0)
        return fail(LSC_E_INVALID, "lsc_debug_gemm_dev: bad argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const __nv_bfloat16* A = static_cast<const __nv_bfloat16*>(d_A);
    const __nv_bfloat16* W = static_cast<const __nv_bfloat16*>(d_W);
    const __nv_bfloat16* R = static_cast<const __nv_bfloat16*>(d_residual);
    __nv_bfloat16* C = static_cast<__nv_bfloat16*>(d_C);
    if (impl == 1) {
        int rc = gemm_ref_launch(A, W, d_bias, R, C, M, N, K, epilogue, N, st);
        return rc < 0 ? rc : LSC_OK;
    }
    const int bn = gemm_tc_block_n(N);
    if (bn == 0) return fail(LSC_E_UNSUPPORTED, "gemm: N must be a multiple of 128");
    if (K % 64 != 0) return fail(LSC_E_UNSUPPORTED, "gemm: K must be a multiple of 64");
    int dev = 0, sms = 0;
    LSC_CUDA(cudaGetDevice(&dev));
    LSC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CUtensorMap ta, tw, tc;
    LSC_TRY(make_tmap_bf16_sw128(&ta, A, M, K, 128));
    LSC_TRY(make_tmap_bf16_sw128(&tw, W, N, K, gemm_tc_w_box_rows(N)));
    LSC_TRY(make_tmap_bf16_store32(&tc, C, M, N));
    int rc = gemm_tc_launch(ta, tw, tc, d_bias, R, M, N, K, epilogue, N, LnArgs(), sms, st);
    return rc < 0 ? rc : LSC_OK;
}

}  // extern "C"
