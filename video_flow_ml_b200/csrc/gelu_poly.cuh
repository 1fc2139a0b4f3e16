// Packed fp32 pair arithmetic (Blackwell FFMA2 / FMUL2 / FADD2) and the polynomial GELU shared by the
// tensor-core GEMM epilogue and the tensor-core depthwise kernel.
#pragma once
#include <cuda_runtime.h>

namespace vfml {

__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("{\n.reg .b64 ra, rb, rc, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmov.b64 rc, {%6, %7};\n"
        "fma.rn.f32x2 rd, ra, rb, rc;\nmov.b64 {%0, %1}, rd;\n}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    float2 d;
    asm("{\n.reg .b64 ra, rb, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmul.rn.f32x2 rd, ra, rb;\nmov.b64 {%0, %1}, rd;\n}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return d;
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
    float2 d;
    asm("{\n.reg .b64 ra, rb, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nadd.rn.f32x2 rd, ra, rb;\nmov.b64 {%0, %1}, rd;\n}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return d;
}
__device__ __forceinline__ float2 splat(float v) { return make_float2(v, v); }

// gelu(x) = x * Phi(x);  Phi(x) - 1/2 = x * R(s),  s = x^2 * (2/25) - 1,  |x| clamped to 5.
// R: degree-12 Chebyshev fit converted to monomials in s (well conditioned on [-1, 1]); max abs error of
// gelu over all x: 2e-6 (tests/test_gelu_poly.py), far below the fp16 rounding of the stored result.
__device__ __forceinline__ float2 gelu_poly2(float2 x) {
    const float2 xc = make_float2(fminf(fmaxf(x.x, -5.0f), 5.0f), fminf(fmaxf(x.y, -5.0f), 5.0f));
    const float2 s = fma2(mul2(xc, xc), splat(0.08f), splat(-1.0f));
    float2 r = splat(7.353763795e-04f);
    r = fma2(r, s, splat(-1.676730928e-03f));
    r = fma2(r, s, splat(1.374596148e-03f));
    r = fma2(r, s, splat(-2.526916796e-03f));
    r = fma2(r, s, splat(6.766527425e-03f));
    r = fma2(r, s, splat(-1.130712498e-02f));
    r = fma2(r, s, splat(1.623608917e-02f));
    r = fma2(r, s, splat(-2.321312763e-02f));
    r = fma2(r, s, splat(3.147675842e-02f));
    r = fma2(r, s, splat(-4.045128077e-02f));
    r = fma2(r, s, splat(5.151792988e-02f));
    r = fma2(r, s, splat(-7.029590756e-02f));
    r = fma2(r, s, splat(1.413638145e-01f));
    return mul2(x, fma2(xc, r, splat(0.5f)));
}

// tanh-form evaluation of the SAME function: Phi(x) = (1 + tanh(g(x))) / 2 with g(x) = atanh(erf(x / sqrt 2)) fitted by
// the odd polynomial x * (c0 + c1 x^2 + c2 x^4) on |x| <= 7 (fit error of gelu: 3.1e-5 abs, tests/test_gelu_poly.py).
// tanh.approx.f32 runs on the MUFU pipe, so the FMA pipe sees 6 packed operations per pair instead of 17; its
// 2^-11 relative error gives |gelu error| <= 2.5e-4 |x|, the size of the fp16 rounding of the stored result.
__device__ __forceinline__ float tanh_approx(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float2 gelu_tanh2(float2 x) {
    // the fit holds for x^2 <= 49; beyond it the polynomial is frozen at its value there (1.79), so g = 1.79 x, |g| > 12.5
    // and tanh saturates to +-1 exactly as the true atanh(erf) does: one min per element instead of a two-sided clamp of x
    float2 x2 = mul2(x, x);
    x2 = make_float2(fminf(x2.x, 49.0f), fminf(x2.y, 49.0f));
    float2 q = fma2(splat(-3.41174769e-04f), x2, splat(3.69255454e-02f));
    q = fma2(q, x2, splat(7.97627597e-01f));
    const float2 t = mul2(x, q);
    const float2 th = make_float2(tanh_approx(t.x), tanh_approx(t.y));
    const float2 hx = mul2(x, splat(0.5f));
    return fma2(hx, th, hx);
}

// the GELU of the fp16 tensor-core tier's fused epilogues
#ifndef VFML_GELU_MUFU
#define VFML_GELU_MUFU 1
#endif
__device__ __forceinline__ float2 gelu_fast2(float2 x) {
#if VFML_GELU_MUFU
    return gelu_tanh2(x);
#else
    return gelu_poly2(x);
#endif
}

}  // namespace vfml
