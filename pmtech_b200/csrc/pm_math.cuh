// pm_math.cuh — the arithmetic of the scene-update path, device side.
//
// Every float operation is an explicit round-to-nearest intrinsic (__fmul_rn / __fadd_rn /
// __fsub_rn / __fsqrt_rn), which nvcc never contracts into FMA, so the rounding sequence is
// exactly the one the reference's C++ source spells out at its call sites:
//   mat4 * mat4            core/put/source/ecs/ecs_scene.cpp:1470
//   transform_vector       ecs_scene.cpp:1505
//   min_union / max_union  ecs_scene.cpp:1507-1508,1524-1525,1548-1554
//   mag                    ecs_scene.cpp:1512
//   sgn / plane_distance / dot      core/put/source/ecs/ecs_cull.cpp:36-38
//   point_plane_distance   ecs_cull.cpp:68
// (the maths library's generic dot: `r = 0; r += a[i] * b[i]`, left to right).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pm {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }

// dot of 4 with the accumulator starting at +0 (0 + a0*b0 turns a -0 product into +0)
__device__ __forceinline__ float dot4(float a0, float b0, float a1, float b1, float a2, float b2, float a3, float b3)
{
    float r = add(0.0f, mul(a0, b0));
    r = add(r, mul(a1, b1));
    r = add(r, mul(a2, b2));
    r = add(r, mul(a3, b3));
    return r;
}

__device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2)
{
    float r = add(0.0f, mul(a0, b0));
    r = add(r, mul(a1, b1));
    r = add(r, mul(a2, b2));
    return r;
}

// min_union / max_union per component: a < b ? a : b  /  a > b ? a : b  (ties and NaN -> b)
__device__ __forceinline__ float min_union(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float max_union(float a, float b) { return a > b ? a : b; }

// ------------------------------------------------------------------------------------------------
// packed pairs (Blackwell FMUL2 / FFMA2: two independent binary32 operations per issue slot).
// Each half is rounded to nearest on its own, so a pair op gives the same bits as two scalar ops.
// The frame pass is bound by instruction issue, not by the FP32 lanes, which is why pairs pay.
//
// ptxas (12.9) contracts `mul.rn.f32x2` followed by `add.rn.f32x2` into one FFMA2 even under
// --fmad=false, which would change the rounding.  Sums are therefore written as
// fma.rn.f32x2(a, one, b) with `one` = (1.0f, 1.0f) read from kernel-parameter memory: a * 1 is
// exact, so the single rounding of the fma is exactly the rounding of a + b, and since ptxas
// cannot see the value of `one` it cannot fold the multiply away and contract a neighbouring
// product into the sum.
// ------------------------------------------------------------------------------------------------
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b, f32x2 one)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(one), "l"(b));
    return r;
}

// a * b + c with one rounding per half.  Only used where one of the two terms is an exact zero
// (the corner select of the plane tests), so it never merges two reference roundings into one.
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// 4x4 row-major matrix held in registers as 4 rows
struct mat4r
{
    float4 r[4];
};

__device__ __forceinline__ float4 mul_row(const float4 a, const mat4r& b)
{
    // row of (A * B): out[c] = dot(A.row, B.column(c))
    float4 o;
    o.x = dot4(a.x, b.r[0].x, a.y, b.r[1].x, a.z, b.r[2].x, a.w, b.r[3].x);
    o.y = dot4(a.x, b.r[0].y, a.y, b.r[1].y, a.z, b.r[2].y, a.w, b.r[3].y);
    o.z = dot4(a.x, b.r[0].z, a.y, b.r[1].z, a.z, b.r[2].z, a.w, b.r[3].z);
    o.w = dot4(a.x, b.r[0].w, a.y, b.r[1].w, a.z, b.r[2].w, a.w, b.r[3].w);
    return o;
}

// world = parent * local
__device__ __forceinline__ mat4r mat_mul(const mat4r& a, const mat4r& b)
{
    mat4r o;
    o.r[0] = mul_row(a.r[0], b);
    o.r[1] = mul_row(a.r[1], b);
    o.r[2] = mul_row(a.r[2], b);
    o.r[3] = mul_row(a.r[3], b);
    return o;
}

// quat::get_matrix (maths library, called at ecs_scene.cpp:1421 and :1293): normalise, then expand.
// quat::get_matrix = the quaternion normalises ITSELF (unless |q| is 0 or 1), then expands; the caller
// writes the normalised rotation back where the reference's in-place call would have left it
// (ecs_scene.cpp:1416-1421, :1291-1300).  One rounded operation per C operation of the library's
// expansion: C evaluates `1 - 2*y*y - 2*z*z` as ((1 - ((2*y)*y)) - ((2*z)*z)).  The expansion is pinned
// by the reference's own scene file: stored rotations expand to the stored local matrices.
__device__ __forceinline__ bool quat_normalise(float& x, float& y, float& z, float& w)
{
    const float mag2 = add(add(add(mul(w, w), mul(x, x)), mul(y, y)), mul(z, z));
    const float mag = __fsqrt_rn(mag2);
    if (mag != 0.0f && mag != 1.0f)
    {
        x = __fdiv_rn(x, mag);
        y = __fdiv_rn(y, mag);
        z = __fdiv_rn(z, mag);
        w = __fdiv_rn(w, mag);
        return true;
    }
    return false;
}

__device__ __forceinline__ mat4r quat_expand(float x, float y, float z, float w)
{
    const float x2 = mul(2.0f, x), y2 = mul(2.0f, y), z2 = mul(2.0f, z);
    mat4r       m;
    m.r[0] = make_float4(sub(sub(1.0f, mul(y2, y)), mul(z2, z)), sub(mul(x2, y), mul(z2, w)), add(mul(x2, z), mul(y2, w)), 0.0f);
    m.r[1] = make_float4(add(mul(x2, y), mul(z2, w)), sub(sub(1.0f, mul(x2, x)), mul(z2, z)), sub(mul(y2, z), mul(x2, w)), 0.0f);
    m.r[2] = make_float4(sub(mul(x2, z), mul(y2, w)), add(mul(y2, z), mul(x2, w)), sub(sub(1.0f, mul(x2, x)), mul(y2, y)), 0.0f);
    m.r[3] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
    return m;
}

// float atomics on raw IEEE bits (exact min/max, order independent; -0 sorts below +0)
__device__ __forceinline__ void atomic_min_float(float* addr, float v)
{
    if (!(__float_as_uint(v) >> 31))
        atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
    else
        atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_float(float* addr, float v)
{
    if (!(__float_as_uint(v) >> 31))
        atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
    else
        atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}

// order-preserving float -> uint key (for warp REDUX min/max); key_to_float inverts it
__device__ __forceinline__ uint32_t float_key(float f)
{
    uint32_t b = __float_as_uint(f);
    return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k)
{
    uint32_t b = k ^ ((k >> 31) ? 0x80000000u : 0xFFFFFFFFu);
    return __uint_as_float(b);
}

} // namespace pm
