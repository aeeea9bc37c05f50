// Epilogue arithmetic shared by the GEMM kernels (gemm_tcgen05.cu, ffn_fused.cu): packed fp32x2 helpers,
// the 256-bit global load and the GELU evaluation.
#pragma once
#include <cstdint>

namespace lsc {

// ---- packed fp32x2 arithmetic (sm_100 FFMA2 / FMUL2): halves the issue slots of the epilogue ----
__device__ __forceinline__ uint64_t pk2(float a, float b) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& a, float& b) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t pk2u(uint32_t a, uint32_t b) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// 256-bit global load (sm_100: LDG.E.ENL2.256): one whole 32-byte sector per thread.
__device__ __forceinline__ void ldg256(const void* p, uint32_t* r) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}
// Exact-erf GELU, x * Phi(x), evaluated without erff():
//   Phi(x) = 0.5 * (1 + tanh(p(x))),  p(x) = atanh(erf(x / sqrt 2)),
// with p fitted by an odd degree-5 polynomial x * (a + b x^2 + c x^4) (minimax for the absolute error
// of gelu; 2.5e-5 with an exact tanh).  tanh is one MUFU.TANH (relative error 2^-11), so the result is
// within ~5e-4 absolute / 2.5e-4 relative of the erf form everywhere — an eighth of the bf16 rounding
// applied to it next.  x^2 is clamped at 64, beyond which tanh has saturated (and before the fitted
// polynomial turns over).  Two elements per FFMA2/FMUL2: ~5 issue slots per element instead of ~25
// for erff().  (The familiar 0.044715 "tanh GELU" is a different, 2-term fit of the same p with
// 4.7e-4 systematic error; this is not that approximation.)
__device__ __forceinline__ float tanh_approx(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void gelu2(float& x0, float& x1) {
    const uint64_t xx = pk2(x0, x1);
    float u0, u1;
    upk2(mul2(xx, xx), u0, u1);
    const uint64_t u = pk2(fminf(u0, 64.0f), fminf(u1, 64.0f));
    uint64_t p = fma2(pk2(-0.000351517994f, -0.000351517994f), u, pk2(0.0370056538f, 0.0370056538f));
    p = fma2(p, u, pk2(0.797507875f, 0.797507875f));
    float t0, t1;
    upk2(mul2(p, xx), t0, t1);
    const uint64_t th = pk2(tanh_approx(t0), tanh_approx(t1));
    const uint64_t hx = mul2(xx, pk2(0.5f, 0.5f));
    upk2(fma2(hx, th, hx), x0, x1);
}

// same, packed in / packed out
__device__ __forceinline__ uint64_t gelu2_pk(uint64_t xx) {
    float x0, x1;
    upk2(xx, x0, x1);
    gelu2(x0, x1);
    return pk2(x0, x1);
}
__device__ __forceinline__ uint32_t pack_bf16x2_pk(uint64_t v) {
    float a, b;
    upk2(v, a, b);
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}

__device__ __forceinline__ float gelu_erf(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
}

}  // namespace lsc
