// lensrt — small vector maths shared by host (lens construction / autofocus) and device code.
//
// Two flavours of arithmetic live here:
//  * strict::*  — every operation individually rounded to nearest (no FMA contraction, IEEE div and
//    sqrt). Used by the RNG, the warps and the whole lens step so that lens ray-miss masks are
//    bit-identical to the pinned scalar build of the reference: Dr.Jit's scalar
//    fmadd(a,b,c) is round(round(a*b)+c) there, rcp(x) = 1/x, rsqrt(x) = 1/sqrt(x).
//    On the device these map to __fmul_rn/__fadd_rn/__fdiv_rn/__fsqrt_rn, which nvcc never
//    contracts; on the host the translation unit is compiled with -ffp-contract=off.
//  * plain float3 operators — used by traversal and shading, where nvcc is free to emit FFMA.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cmath>

#define LRT_HD __host__ __device__ __forceinline__
#define LRT_D __device__ __forceinline__

namespace lrt {

constexpr float kEpsilon = 5.9604644775390625e-08f;         // dr::Epsilon<float>
constexpr float kOneMinusEpsilon = 0.999999940395355225f;   // dr::OneMinusEpsilon<float>
constexpr float kRayEpsilon = 8.94069671630859375e-05f;     // math::RayEpsilon (core/math.h:18-22)
constexpr float kShadowEpsilon = 8.94069671630859375e-04f;  // math::ShadowEpsilon
constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kLargest = 3.402823466e+38f;

LRT_HD float inf_f() {
#ifdef __CUDA_ARCH__
    return __int_as_float(0x7f800000);
#else
    return INFINITY;
#endif
}
LRT_HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
LRT_HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

namespace strict {
#if defined(__CUDA_ARCH__) && defined(LRT_FAST_ARITH)
// the fast flavour (lrt_spot_fast.cu only): plain operators (nvcc contracts a * b + c into one FFMA) and the MUFU
// approximations (1-2 ulp) — the same functions, the same algorithm, no claim on the last bit
LRT_HD float mul(float a, float b) { return a * b; }
LRT_HD float add(float a, float b) { return a + b; }
LRT_HD float sub(float a, float b) { return a - b; }
LRT_HD float div(float a, float b) { return __fdividef(a, b); }
LRT_HD float sqrt(float a) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
#elif defined(__CUDA_ARCH__)
LRT_HD float mul(float a, float b) { return __fmul_rn(a, b); }
LRT_HD float add(float a, float b) { return __fadd_rn(a, b); }
LRT_HD float sub(float a, float b) { return __fsub_rn(a, b); }
LRT_HD float div(float a, float b) { return __fdiv_rn(a, b); }
LRT_HD float sqrt(float a) { return __fsqrt_rn(a); }
#else
LRT_HD float mul(float a, float b) { return a * b; }
LRT_HD float add(float a, float b) { return a + b; }
LRT_HD float sub(float a, float b) { return a - b; }
LRT_HD float div(float a, float b) { return a / b; }
LRT_HD float sqrt(float a) { return ::sqrtf(a); }
#endif
LRT_HD float sqr(float a) { return mul(a, a); }
// 1 / a: the correctly rounded reciprocal IS the correctly rounded quotient 1.f / a, so the shorter __frcp_rn sequence
// gives the same bits as the division the reference performs (dr::rcp in scalar mode, array_utils.h:212-214)
#if defined(__CUDA_ARCH__) && defined(LRT_FAST_ARITH)
LRT_HD float rcp(float a) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
LRT_HD float rsqrt(float a) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
#elif defined(__CUDA_ARCH__)
LRT_HD float rcp(float a) { return __frcp_rn(a); }
LRT_HD float rsqrt(float a) { return rcp(sqrt(a)); }
#else
LRT_HD float rcp(float a) { return div(1.f, a); }
LRT_HD float rsqrt(float a) { return rcp(sqrt(a)); }
#endif
LRT_HD float fmadd(float a, float b, float c) { return add(mul(a, b), c); }    // two roundings
LRT_HD float fmsub(float a, float b, float c) { return sub(mul(a, b), c); }
LRT_HD float fnmadd(float a, float b, float c) { return sub(c, mul(a, b)); }

struct V3 { float x, y, z; };
LRT_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
LRT_HD V3 add(V3 a, V3 b) { return v3(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)); }
LRT_HD V3 sub(V3 a, V3 b) { return v3(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
LRT_HD V3 neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
LRT_HD V3 mul(V3 a, float s) { return v3(mul(a.x, s), mul(a.y, s), mul(a.z, s)); }
LRT_HD float dot(V3 a, V3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
LRT_HD float squared_norm(V3 a) { return dot(a, a); }
LRT_HD float norm(V3 a) { return sqrt(squared_norm(a)); }
LRT_HD V3 normalize(V3 a) { return mul(a, rsqrt(squared_norm(a))); }
LRT_HD V3 at(V3 o, V3 d, float t) { return v3(fmadd(d.x, t, o.x), fmadd(d.y, t, o.y), fmadd(d.z, t, o.z)); }
}  // namespace strict

// ------------------------------------------------------------------ relaxed float3 helpers
LRT_HD float3 operator+(float3 a, float3 b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
LRT_HD float3 operator-(float3 a, float3 b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
LRT_HD float3 operator-(float3 a) { return make_float3(-a.x, -a.y, -a.z); }
LRT_HD float3 operator*(float3 a, float s) { return make_float3(a.x * s, a.y * s, a.z * s); }
LRT_HD float3 operator*(float s, float3 a) { return make_float3(a.x * s, a.y * s, a.z * s); }
LRT_HD float3 operator*(float3 a, float3 b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
LRT_HD float3 &operator+=(float3 &a, float3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
LRT_HD float3 &operator*=(float3 &a, float3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
LRT_HD float3 &operator*=(float3 &a, float s) { a.x *= s; a.y *= s; a.z *= s; return a; }
LRT_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
LRT_HD float3 cross(float3 a, float3 b) {
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
LRT_HD float3 normalize(float3 a) { return a * (1.f / sqrtf(dot(a, a))); }
LRT_HD float max3(float3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
LRT_HD float3 f3(strict::V3 v) { return make_float3(v.x, v.y, v.z); }

}  // namespace lrt
