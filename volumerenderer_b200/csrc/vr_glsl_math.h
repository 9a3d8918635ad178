// vr_glsl_math.h -- the "GLSL built-in function" layer, pinned so that host (g++) and device (nvcc,
// sm_100a) produce bit-identical FP32 results.
//
// GLSL leaves the precision of exp/sin/cos/acos/tan/normalize implementation-defined. The reference
// shaders (shaders/common_functions.h:21-25 randomDir, shaders/light_gen.comp:66 exp, ...) feed these
// results into branch decisions (scatter rolls, T cut-offs), so a bit-exact light list needs ONE
// definition of every built-in, executed with the same instruction sequence everywhere. This header
// is that definition. It is shared by the CUDA kernels and by the CPU checker
// under oracle/ (which uses it only as its "GLSL implementation").
//
// Rules: every FP32 operation is an individually rounded IEEE-754 binary32 op. On the device the
// __f*_rn intrinsics are used (never contracted into FMA); on the host the translation unit must be
// compiled with -ffp-contract=off (checked in the Makefiles). Polynomial coefficients are the public
// single-precision Cephes minimax sets.
#pragma once
#include <stdint.h>
#if !defined(__CUDACC__)
#include <math.h>
#include <string.h>
#endif

#if defined(__CUDACC__)
#define VR_HD __host__ __device__ __forceinline__
#else
#define VR_HD static inline
#endif

// ---------------------------------------------------------------- exactly-rounded primitive ops
VR_HD float vr_mul(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    return a * b;
#endif
}
VR_HD float vr_add(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    return a + b;
#endif
}
VR_HD float vr_sub(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fsub_rn(a, b);
#else
    return a - b;
#endif
}
VR_HD float vr_div(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
}
VR_HD float vr_sqrt(float a) {
#if defined(__CUDA_ARCH__)
    return __fsqrt_rn(a);
#else
    return sqrtf(a);
#endif
}
VR_HD float vr_floor(float a) { return floorf(a); }

VR_HD float vr_int_as_float(int32_t i) {
#if defined(__CUDA_ARCH__)
    return __int_as_float(i);
#else
    float f;
    memcpy(&f, &i, 4);
    return f;
#endif
}
VR_HD int32_t vr_float_as_int(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    int32_t i;
    memcpy(&i, &f, 4);
    return i;
#endif
}
// float(uint): round-to-nearest-even (GLSL conversion; Appendix B)
VR_HD float vr_u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint2float_rn(u);
#else
    return (float)u;
#endif
}
VR_HD float vr_i2f(int32_t i) {
#if defined(__CUDA_ARCH__)
    return __int2float_rn(i);
#else
    return (float)i;
#endif
}

// ---------------------------------------------------------------- GLSL min / max / clamp (literal definitions)
VR_HD float vr_min(float x, float y) { return (y < x) ? y : x; }
VR_HD float vr_max(float x, float y) { return (x < y) ? y : x; }
VR_HD float vr_clamp(float x, float lo, float hi) { return vr_min(vr_max(x, lo), hi); }

// ---------------------------------------------------------------- vectors
struct vr_f3 {
    float x, y, z;
};
VR_HD vr_f3 vr_make3(float x, float y, float z) {
    vr_f3 r;
    r.x = x;
    r.y = y;
    r.z = z;
    return r;
}
// dot: products summed left to right
VR_HD float vr_dot3(vr_f3 a, vr_f3 b) {
    return vr_add(vr_add(vr_mul(a.x, b.x), vr_mul(a.y, b.y)), vr_mul(a.z, b.z));
}
VR_HD float vr_length3(vr_f3 a) { return vr_sqrt(vr_dot3(a, a)); }
// normalize(v) = v / sqrt(dot(v,v)), IEEE divide per component
VR_HD vr_f3 vr_normalize3(vr_f3 a) {
    float l = vr_length3(a);
    return vr_make3(vr_div(a.x, l), vr_div(a.y, l), vr_div(a.z, l));
}
VR_HD vr_f3 vr_add3(vr_f3 a, vr_f3 b) { return vr_make3(vr_add(a.x, b.x), vr_add(a.y, b.y), vr_add(a.z, b.z)); }
VR_HD vr_f3 vr_sub3(vr_f3 a, vr_f3 b) { return vr_make3(vr_sub(a.x, b.x), vr_sub(a.y, b.y), vr_sub(a.z, b.z)); }
VR_HD vr_f3 vr_scale3(vr_f3 a, float s) { return vr_make3(vr_mul(a.x, s), vr_mul(a.y, s), vr_mul(a.z, s)); }

// ---------------------------------------------------------------- exp
// exp(x): k = rint(x*log2(e)); r = x - k*ln2 (two-constant Cody-Waite); degree-5 minimax on r; scale by 2^k.
VR_HD float vr_exp(float x) {
    if (x != x) return x;
    if (x > 88.7228317f) return vr_int_as_float(0x7f800000);
    if (x < -103.97f) return 0.0f;
    const float MAGIC = 12582912.0f;  // 1.5 * 2^23: add/sub rounds to nearest integer
    float fk = vr_sub(vr_add(vr_mul(x, 1.44269504088896341f), MAGIC), MAGIC);
    float r = vr_sub(x, vr_mul(fk, 0.693359375f));
    r = vr_sub(r, vr_mul(fk, -2.12194440e-4f));
    float r2 = vr_mul(r, r);
    float p = 1.9875691500e-4f;
    p = vr_add(vr_mul(p, r), 1.3981999507e-3f);
    p = vr_add(vr_mul(p, r), 8.3334519073e-3f);
    p = vr_add(vr_mul(p, r), 4.1665795894e-2f);
    p = vr_add(vr_mul(p, r), 1.6666665459e-1f);
    p = vr_add(vr_mul(p, r), 5.0000001201e-1f);
    p = vr_add(vr_add(vr_mul(p, r2), r), 1.0f);
    int32_t k = (int32_t)fk;
    if (k < -126) {  // result is subnormal or tiny: scale in two exact steps
        p = vr_mul(p, vr_int_as_float((k + 64 + 127) << 23));
        return vr_mul(p, vr_int_as_float((-64 + 127) << 23));
    }
    if (k > 127) {
        p = vr_mul(p, vr_int_as_float((k - 1 + 127) << 23));
        return vr_mul(p, 2.0f);
    }
    return vr_mul(p, vr_int_as_float((k + 127) << 23));
}

// ---------------------------------------------------------------- sin / cos
// q = rint(x * 2/pi); r = x - q*pi/2 (three-constant Cody-Waite); minimax polynomials on |r| <= pi/4.
VR_HD void vr_sincos_reduce(float x, float* r_out, int32_t* q_out) {
    const float MAGIC = 12582912.0f;
    float fq = vr_sub(vr_add(vr_mul(x, 0.636619772367581343f), MAGIC), MAGIC);
    float r = vr_sub(x, vr_mul(fq, 1.5703125f));
    r = vr_sub(r, vr_mul(fq, 4.837512969970703125e-4f));
    r = vr_sub(r, vr_mul(fq, 7.54978995489188216e-8f));
    *r_out = r;
    *q_out = (int32_t)fq;
}
VR_HD float vr_sin_poly(float r) {
    float z = vr_mul(r, r);
    float p = -1.9515295891e-4f;
    p = vr_add(vr_mul(p, z), 8.3321608736e-3f);
    p = vr_add(vr_mul(p, z), -1.6666654611e-1f);
    return vr_add(vr_mul(vr_mul(p, z), r), r);
}
VR_HD float vr_cos_poly(float r) {
    float z = vr_mul(r, r);
    float p = 2.443315711809948e-5f;
    p = vr_add(vr_mul(p, z), -1.388731625493765e-3f);
    p = vr_add(vr_mul(p, z), 4.166664568298827e-2f);
    p = vr_mul(vr_mul(p, z), z);
    return vr_add(vr_sub(p, vr_mul(0.5f, z)), 1.0f);
}
VR_HD float vr_sin(float x) {
    float r;
    int32_t q;
    vr_sincos_reduce(x, &r, &q);
    float v = (q & 1) ? vr_cos_poly(r) : vr_sin_poly(r);
    return (q & 2) ? -v : v;
}
VR_HD float vr_cos(float x) {
    float r;
    int32_t q;
    vr_sincos_reduce(x, &r, &q);
    float v = (q & 1) ? vr_sin_poly(r) : vr_cos_poly(r);
    return (((q + 1) & 2) != 0) ? -v : v;
}
VR_HD float vr_tan(float x) { return vr_div(vr_sin(x), vr_cos(x)); }

// ---------------------------------------------------------------- acos
VR_HD float vr_asin_core(float x) {  // |x| <= 0.5
    float z = vr_mul(x, x);
    float p = 4.2163199048e-2f;
    p = vr_add(vr_mul(p, z), 2.4181311049e-2f);
    p = vr_add(vr_mul(p, z), 4.5470025998e-2f);
    p = vr_add(vr_mul(p, z), 7.4953002686e-2f);
    p = vr_add(vr_mul(p, z), 1.6666752422e-1f);
    return vr_add(vr_mul(vr_mul(p, z), x), x);
}
VR_HD float vr_acos(float x) {
    const float PI_F = 3.14159265358979323846f;
    const float PIO2_F = 1.57079632679489661923f;
    if (x != x) return x;
    if (x > 1.0f || x < -1.0f) return vr_int_as_float(0x7fc00000);
    if (x > 0.5f) return vr_mul(2.0f, vr_asin_core(vr_sqrt(vr_mul(0.5f, vr_sub(1.0f, x)))));
    if (x < -0.5f) return vr_sub(PI_F, vr_mul(2.0f, vr_asin_core(vr_sqrt(vr_mul(0.5f, vr_add(1.0f, x))))));
    return vr_sub(PIO2_F, vr_asin_core(x));
}
