// crt_math.cuh — float3 algebra with the reference's operation order.
//
// The reference's vec3 (CudaTest/src/core/vec3.h:7-240) is evaluated without
// FMA contraction by the host oracle (-ffp-contract=off). Every function in
// namespace crt::x ("exact") is built from __fmul_rn/__fadd_rn/__fdiv_rn/
// __fsqrt_rn, which nvcc never fuses and which ignore -use_fast_math, so the
// intersection, camera and scatter math is bit-identical to the oracle's.
// Box tests, where only conservativeness matters, use ordinary (fusable) math.
//
// All functions are __host__ __device__ so tests/emu can run the same code on
// the CPU (test-only; the product has no CPU path).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <float.h>

#define CRT_HD __host__ __device__ __forceinline__
#include "../../include/crt_rng.h"

namespace crt {

struct v3 { float x, y, z; };

namespace x {
CRT_HD float mul(float a, float b)
{
#ifdef __CUDA_ARCH__
    return __fmul_rn(a, b);
#else
    return a * b;
#endif
}
CRT_HD float add(float a, float b)
{
#ifdef __CUDA_ARCH__
    return __fadd_rn(a, b);
#else
    return a + b;
#endif
}
CRT_HD float sub(float a, float b)
{
#ifdef __CUDA_ARCH__
    return __fsub_rn(a, b);
#else
    return a - b;
#endif
}
CRT_HD float div(float a, float b)
{
#ifdef __CUDA_ARCH__
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
}
CRT_HD float sqrt(float a)
{
#ifdef __CUDA_ARCH__
    return __fsqrt_rn(a);
#else
    return sqrtf(a);
#endif
}
// float(1.0 / a): a double division rounded to float equals the correctly
// rounded float reciprocal (53 >= 2*24+2), used at triangle.h:80 and vec3.h:183.
CRT_HD float rcp(float a)
{
#ifdef __CUDA_ARCH__
    return __frcp_rn(a);
#else
    return 1.0f / a;
#endif
}
}  // namespace x

CRT_HD v3 V(float a, float b, float c) { v3 r; r.x = a; r.y = b; r.z = c; return r; }
CRT_HD v3 vadd(v3 a, v3 b) { return V(x::add(a.x, b.x), x::add(a.y, b.y), x::add(a.z, b.z)); }      // vec3.h:73
CRT_HD v3 vsub(v3 a, v3 b) { return V(x::sub(a.x, b.x), x::sub(a.y, b.y), x::sub(a.z, b.z)); }      // vec3.h:77
CRT_HD v3 vmul(v3 a, v3 b) { return V(x::mul(a.x, b.x), x::mul(a.y, b.y), x::mul(a.z, b.z)); }      // vec3.h:89
CRT_HD v3 vscale(float t, v3 a) { return V(x::mul(t, a.x), x::mul(t, a.y), x::mul(t, a.z)); }       // vec3.h:93-99
CRT_HD v3 vdivf(v3 a, float t) { return V(x::div(a.x, t), x::div(a.y, t), x::div(a.z, t)); }        // vec3.h:105
CRT_HD v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }                                                 // vec3.h:24
CRT_HD float vdot(v3 a, v3 b)                                                                         // vec3.h:109
{
    return x::add(x::add(x::mul(a.x, b.x), x::mul(a.y, b.y)), x::mul(a.z, b.z));
}
CRT_HD v3 vcross(v3 a, v3 b)                                                                          // vec3.h:113
{
    return V(x::sub(x::mul(a.y, b.z), x::mul(a.z, b.y)),
             -x::sub(x::mul(a.x, b.z), x::mul(a.z, b.x)),
             x::sub(x::mul(a.x, b.y), x::mul(a.y, b.x)));
}
CRT_HD float vlen2(v3 a) { return x::add(x::add(x::mul(a.x, a.x), x::mul(a.y, a.y)), x::mul(a.z, a.z)); }  // vec3.h:42
CRT_HD float vlen(v3 a) { return x::sqrt(vlen2(a)); }                                                // vec3.h:38
CRT_HD v3 vunit(v3 a) { return vdivf(a, vlen(a)); }                                                  // vec3.h:191
CRT_HD v3 vlerp(float t, v3 from, v3 to) { return vadd(from, vscale(t, vsub(to, from))); }           // vec3.h:196

// vec3.h:119-131 with the default int bounds 0 and 1.
CRT_HD float clip01(float f)
{
    if (f > 1.0f) return 1.0f;
    else if (f < 0.0f) return 0.0f;
    return f;
}

CRT_HD float u2f(uint32_t u)
{
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { uint32_t u; float f; } c; c.u = u; return c.f;
#endif
}
CRT_HD uint32_t f2u(float f)
{
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    union { uint32_t u; float f; } c; c.f = f; return c.u;
#endif
}

}  // namespace crt
