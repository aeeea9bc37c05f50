// HBM-bound row kernels of the encoder: embedding gather + LayerNorm (K1), LayerNorm over a
// residual-added row (the tail of K4/K6) and masked mean-pool + L2-normalise (K7).
// They replace libtorch's BertEmbeddings / LayerNorm and DJL's TextEmbeddingTranslator
// post-processing (mean pooling + normalisation; evidenced by the unit-norm golden vector at
// langstream-runtime/langstream-runtime-impl/src/test/java/ai/langstream/kafka/ComputeEmbeddingsIT.java:205-210).
// One warp per row, 16-byte / 8-byte vector accesses, fp32 statistics (two-pass in registers).
#include "common.h"
#include "ptx.cuh"

namespace lsc {

// <<<>>> with the programmatic-stream-serialization attribute (the kernels call pdl_wait() first thing)
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid;
    lc.blockDim = block;
    lc.dynamicSmemBytes = 0;
    lc.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = 1;
    return cudaLaunchKernelEx(&lc, kern, static_cast<KArgs>(args)...);
}


namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// VPL = float4 vectors per lane; H = 128 * VPL.
template <int VPL>
__global__ void __launch_bounds__(256)
embed_ln_kernel(const int32_t* __restrict__ token_ids, const int32_t* __restrict__ cu_seqlens,
                int n_seq, int T, int max_position, int vocab, const float* __restrict__ word,
                const float* __restrict__ pos, const float* __restrict__ type0,
                const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                __nv_bfloat16* __restrict__ out, int32_t* __restrict__ seq_of_token,
                float* __restrict__ stats_out, int stats_p) {
    constexpr int H = 128 * VPL;
    // programmatic dependent launch (ptx.cuh): x / its statistics may still be read by the previous pass
    pdl_launch_dependents();
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= T) return;
    // sequence of this token: binary search over cu_seqlens (n_seq+1 entries, L1/L2 resident)
    int lo = 0, hi = n_seq;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(cu_seqlens + mid) <= t) lo = mid; else hi = mid;
    }
    const int p = t - __ldg(cu_seqlens + lo);
    if (lane == 0 && seq_of_token != nullptr) seq_of_token[t] = lo;
    // The rows written here are already normalised; downstream GEMMs fold a LayerNorm over their input
    // from per-row partial (sum, sum of squares): emit partials that mean "mean 0, variance 1".
    if (stats_out != nullptr && lane < stats_p)
        reinterpret_cast<float2*>(stats_out)[static_cast<size_t>(t) * stats_p + lane] =
            make_float2(0.f, static_cast<float>(H) / static_cast<float>(stats_p));
    int id = __ldg(token_ids + t);
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    const int pp = p < max_position ? p : max_position - 1;
    const float4* w4 = reinterpret_cast<const float4*>(word + static_cast<size_t>(id) * H);
    const float4* p4 = reinterpret_cast<const float4*>(pos + static_cast<size_t>(pp) * H);
    const float4* t4 = reinterpret_cast<const float4*>(type0);
    float4 v[VPL];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = i * 32 + lane;
        const float4 a = __ldg(w4 + c), b = __ldg(p4 + c), d = __ldg(t4 + c);
        v[i] = make_float4(a.x + b.x + d.x, a.y + b.y + d.y, a.z + b.z + d.z, a.w + b.w + d.w);
        s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
    const float mean = warp_sum(s) * (1.0f / H);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
        q += a * a + b * b + c * c + d * d;
    }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / H) + eps);
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    const float4* b4 = reinterpret_cast<const float4*>(beta);
    uint2* o2 = reinterpret_cast<uint2*>(out + static_cast<size_t>(t) * H);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = i * 32 + lane;
        const float4 g = __ldg(g4 + c), b = __ldg(b4 + c);
        uint2 o;
        o.x = pack_bf16x2((v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y);
        o.y = pack_bf16x2((v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w);
        o2[c] = o;
    }
}

template <int VPL>
__global__ void __launch_bounds__(256)
layernorm_kernel(const __nv_bfloat16* __restrict__ in, const float* __restrict__ gamma,
                 const float* __restrict__ beta, float eps, __nv_bfloat16* __restrict__ out, int T) {
    constexpr int H = 128 * VPL;
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= T) return;
    const uint2* i2 = reinterpret_cast<const uint2*>(in + static_cast<size_t>(t) * H);
    float v[VPL][4];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const uint2 w = __ldg(i2 + i * 32 + lane);
        v[i][0] = bf16_lo(w.x);
        v[i][1] = bf16_hi(w.x);
        v[i][2] = bf16_lo(w.y);
        v[i][3] = bf16_hi(w.y);
        s += v[i][0] + v[i][1] + v[i][2] + v[i][3];
    }
    const float mean = warp_sum(s) * (1.0f / H);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float d = v[i][j] - mean;
            q += d * d;
        }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / H) + eps);
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    const float4* b4 = reinterpret_cast<const float4*>(beta);
    uint2* o2 = reinterpret_cast<uint2*>(out + static_cast<size_t>(t) * H);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = i * 32 + lane;
        const float4 g = __ldg(g4 + c), b = __ldg(b4 + c);
        uint2 o;
        o.x = pack_bf16x2((v[i][0] - mean) * rstd * g.x + b.x, (v[i][1] - mean) * rstd * g.y + b.y);
        o.y = pack_bf16x2((v[i][2] - mean) * rstd * g.z + b.z, (v[i][3] - mean) * rstd * g.w + b.w);
        o2[c] = o;
    }
}

// One CTA per sequence; thread j owns columns 2j, 2j+1 (blockDim = H/2); rows are read coalesced.
__global__ void pool_l2_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ cu_seqlens,
                               int H, float* __restrict__ out, const float* __restrict__ stats, int stats_p,
                               const float* __restrict__ gamma, const float* __restrict__ beta, float eps) {
    __shared__ float red[32];
    pdl_launch_dependents();
    pdl_wait();  // the last feed-forward kernel has finished (programmatic dependent launch, ptx.cuh)
    const int s = blockIdx.x;
    const int t0 = __ldg(cu_seqlens + s), t1 = __ldg(cu_seqlens + s + 1);
    const int j = threadIdx.x;
    float a0 = 0.f, a1 = 0.f;
    const uint32_t* xp = reinterpret_cast<const uint32_t*>(x) + j;
    const int H2 = H >> 1;
    int t = t0;
    if (stats != nullptr) {
        // x holds pre-LayerNorm rows; apply the final LayerNorm on the fly:
        // mean_t LN(x_t) = gamma * mean_t((x_t - mu_t) * rstd_t) + beta  (for a non-empty sequence).
        // Token statistics are finalised once per token into shared memory, in blocks of 128 tokens.
        __shared__ float2 tok[128];  // (mean * rstd, rstd)
        const float inv_h = 1.0f / static_cast<float>(H);
        for (int tb = t0; tb < t1; tb += 128) {
            const int nb = min(128, t1 - tb);
            __syncthreads();
            for (int i = threadIdx.x; i < nb; i += blockDim.x) {
                const float2* sp = reinterpret_cast<const float2*>(stats) + static_cast<size_t>(tb + i) * stats_p;
                float sum = 0.f, sq = 0.f;
                for (int p = 0; p < stats_p; ++p) {
                    const float2 v = __ldg(sp + p);
                    sum += v.x;
                    sq += v.y;
                }
                const float mean = sum * inv_h;
                const float rstd = rsqrtf(fmaxf(sq * inv_h - mean * mean, 0.f) + eps);
                tok[i] = make_float2(mean * rstd, rstd);
            }
            __syncthreads();
            int i = 0;
            for (; i + 4 <= nb; i += 4) {
                const uint32_t w0 = __ldg(xp + static_cast<size_t>(tb + i) * H2);
                const uint32_t w1 = __ldg(xp + static_cast<size_t>(tb + i + 1) * H2);
                const uint32_t w2 = __ldg(xp + static_cast<size_t>(tb + i + 2) * H2);
                const uint32_t w3 = __ldg(xp + static_cast<size_t>(tb + i + 3) * H2);
                const float2 s0 = tok[i], s1 = tok[i + 1], s2 = tok[i + 2], s3 = tok[i + 3];
                a0 += (fmaf(bf16_lo(w0), s0.y, -s0.x) + fmaf(bf16_lo(w1), s1.y, -s1.x)) +
                      (fmaf(bf16_lo(w2), s2.y, -s2.x) + fmaf(bf16_lo(w3), s3.y, -s3.x));
                a1 += (fmaf(bf16_hi(w0), s0.y, -s0.x) + fmaf(bf16_hi(w1), s1.y, -s1.x)) +
                      (fmaf(bf16_hi(w2), s2.y, -s2.x) + fmaf(bf16_hi(w3), s3.y, -s3.x));
            }
            for (; i < nb; ++i) {
                const uint32_t w0 = __ldg(xp + static_cast<size_t>(tb + i) * H2);
                const float2 s0 = tok[i];
                a0 += fmaf(bf16_lo(w0), s0.y, -s0.x);
                a1 += fmaf(bf16_hi(w0), s0.y, -s0.x);
            }
        }
        t = t1;
        const float cnt = static_cast<float>(t1 - t0);
        if (cnt > 0.f) {
            a0 = a0 * __ldg(gamma + 2 * j) + cnt * __ldg(beta + 2 * j);
            a1 = a1 * __ldg(gamma + 2 * j + 1) + cnt * __ldg(beta + 2 * j + 1);
        }
    }
    for (; t + 4 <= t1; t += 4) {
        const uint32_t w0 = __ldg(xp + static_cast<size_t>(t) * H2);
        const uint32_t w1 = __ldg(xp + static_cast<size_t>(t + 1) * H2);
        const uint32_t w2 = __ldg(xp + static_cast<size_t>(t + 2) * H2);
        const uint32_t w3 = __ldg(xp + static_cast<size_t>(t + 3) * H2);
        a0 += (bf16_lo(w0) + bf16_lo(w1)) + (bf16_lo(w2) + bf16_lo(w3));
        a1 += (bf16_hi(w0) + bf16_hi(w1)) + (bf16_hi(w2) + bf16_hi(w3));
    }
    for (; t < t1; ++t) {
        const uint32_t w0 = __ldg(xp + static_cast<size_t>(t) * H2);
        a0 += bf16_lo(w0);
        a1 += bf16_hi(w0);
    }
    // masked mean: sum / max(count, 1e-9)  (oracle: clamp(min=1e-9)); empty sequence -> zeros
    const float cnt = static_cast<float>(t1 - t0);
    const float inv = 1.0f / fmaxf(cnt, 1e-9f);
    a0 *= inv;
    a1 *= inv;
    float q = warp_sum(a0 * a0 + a1 * a1);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = q;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    float tot = 0.f;
    for (int i = 0; i < nw; ++i) tot += red[i];
    // F.normalize: x / max(||x||, eps), eps = 1e-12
    const float rn = 1.0f / fmaxf(sqrtf(tot), 1e-12f);
    out[static_cast<size_t>(s) * H + 2 * j] = a0 * rn;
    out[static_cast<size_t>(s) * H + 2 * j + 1] = a1 * rn;
}

}  // namespace

int embed_ln_launch(const int32_t* token_ids, const int32_t* cu_seqlens, int n_seq, int T, int H,
                    int max_position, int vocab, const float* word, const float* pos,
                    const float* type0, const float* gamma, const float* beta, float eps,
                    __nv_bfloat16* out, int32_t* seq_of_token, float* stats_out, int stats_p,
                    cudaStream_t stream) {
    if (T <= 0) return 0;
    if (stats_p > 32) return fail(LSC_E_UNSUPPORTED, "embed: more than 32 statistics partials per row");
    const int wpb = 8;
    const int grid = (T + wpb - 1) / wpb;
    switch (H) {
        case 128: LSC_CUDA(launch_pdl(embed_ln_kernel<1>, dim3(grid), dim3(256), stream, token_ids, cu_seqlens, n_seq, T, max_position, vocab, word, pos, type0, gamma, beta, eps, out, seq_of_token, stats_out, stats_p)); break;
        case 256: LSC_CUDA(launch_pdl(embed_ln_kernel<2>, dim3(grid), dim3(256), stream, token_ids, cu_seqlens, n_seq, T, max_position, vocab, word, pos, type0, gamma, beta, eps, out, seq_of_token, stats_out, stats_p)); break;
        case 384: LSC_CUDA(launch_pdl(embed_ln_kernel<3>, dim3(grid), dim3(256), stream, token_ids, cu_seqlens, n_seq, T, max_position, vocab, word, pos, type0, gamma, beta, eps, out, seq_of_token, stats_out, stats_p)); break;
        case 512: LSC_CUDA(launch_pdl(embed_ln_kernel<4>, dim3(grid), dim3(256), stream, token_ids, cu_seqlens, n_seq, T, max_position, vocab, word, pos, type0, gamma, beta, eps, out, seq_of_token, stats_out, stats_p)); break;
        case 768: LSC_CUDA(launch_pdl(embed_ln_kernel<6>, dim3(grid), dim3(256), stream, token_ids, cu_seqlens, n_seq, T, max_position, vocab, word, pos, type0, gamma, beta, eps, out, seq_of_token, stats_out, stats_p)); break;
        case 1024: LSC_CUDA(launch_pdl(embed_ln_kernel<8>, dim3(grid), dim3(256), stream, token_ids, cu_seqlens, n_seq, T, max_position, vocab, word, pos, type0, gamma, beta, eps, out, seq_of_token, stats_out, stats_p)); break;
        default: return fail(LSC_E_UNSUPPORTED, "hidden size must be one of 128,256,384,512,768,1024");
    }
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int layernorm_launch(const __nv_bfloat16* in, const float* gamma, const float* beta, float eps,
                     __nv_bfloat16* out, int T, int H, cudaStream_t stream) {
    if (T <= 0) return 0;
    const int wpb = 8;
    const int grid = (T + wpb - 1) / wpb;
    switch (H) {
        case 128: layernorm_kernel<1><<<grid, 256, 0, stream>>>(in, gamma, beta, eps, out, T); break;
        case 256: layernorm_kernel<2><<<grid, 256, 0, stream>>>(in, gamma, beta, eps, out, T); break;
        case 384: layernorm_kernel<3><<<grid, 256, 0, stream>>>(in, gamma, beta, eps, out, T); break;
        case 512: layernorm_kernel<4><<<grid, 256, 0, stream>>>(in, gamma, beta, eps, out, T); break;
        case 768: layernorm_kernel<6><<<grid, 256, 0, stream>>>(in, gamma, beta, eps, out, T); break;
        case 1024: layernorm_kernel<8><<<grid, 256, 0, stream>>>(in, gamma, beta, eps, out, T); break;
        default: return fail(LSC_E_UNSUPPORTED, "hidden size must be one of 128,256,384,512,768,1024");
    }
    LSC_CUDA(cudaGetLastError());
    return 1;
}

int pool_l2_launch(const __nv_bfloat16* x, const int32_t* cu_seqlens, int n_seq, int H, float* out,
                   const float* stats, int stats_p, const float* gamma, const float* beta, float eps,
                   cudaStream_t stream) {
    if (n_seq <= 0) return 0;
    if (H % 64 != 0 || H / 2 > 1024) return fail(LSC_E_UNSUPPORTED, "pool: bad hidden size");
    LSC_CUDA(launch_pdl(pool_l2_kernel, dim3(n_seq), dim3(H / 2), stream, x, cu_seqlens, H, out, stats, stats_p, gamma, beta, eps));
    LSC_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace lsc
