// triangle.cuh -- watertight ray/triangle test with the reference's exact arithmetic.
//
// Hit ids are defined by WHICH candidates this predicate accepts, so the
// operation order and every fused multiply-add sit where the reference's sm_100 build of
// Triangle::intersect_triangle (shapes/triangle.cu:213-323, util/math.h:167-173) has them.
// Everything is written with explicit-rounding intrinsics so nvcc/ptxas can neither contract
// nor reassociate: the function computes the same bits at -O0 and -O3.
#pragma once

#include <cuda_runtime.h>

namespace b200pt {

// util/math.h:167-173 difference_of_products: a*b - c*d with one compensation step
__device__ __forceinline__ float dop(float a, float b, float c, float d) {
    const float cd = __fmul_rn(c, d);
    const float diff = __fmaf_rn(a, b, -cd);
    const float err = __fmaf_rn(-c, d, cd);
    return __fadd_rn(diff, err);
}

__device__ __forceinline__ float max3f(float a, float b, float c) {
    const float m = (a < b) ? b : a;
    return (m < c) ? c : m;
}

struct TriHit {
    float b0, b1, b2, t;
};

// gamma(n) = n*eps/(1 - n*eps), eps = 2^-24 (util/math.h:81-84), as folded by nvcc
#define B200PT_GAMMA2 __uint_as_float(0x34000001u)
#define B200PT_GAMMA3 __uint_as_float(0x34400002u)
#define B200PT_GAMMA5 __uint_as_float(0x34A00003u)

// :218 degenerate triangle: |(p2-p0) x (p1-p0)|^2 == 0.  Depends on the triangle alone: the accelerator build
// evaluates it once per triangle and stores the answer as a flag bit of the leaf record (bvh_build.cu).
__device__ __forceinline__ bool triangle_is_degenerate(float p0x, float p0y, float p0z, float p1x, float p1y, float p1z,
                                                       float p2x, float p2y, float p2z) {
    const float ax = __fsub_rn(p2x, p0x), ay = __fsub_rn(p2y, p0y), az = __fsub_rn(p2z, p0z);
    const float bx = __fsub_rn(p1x, p0x), by = __fsub_rn(p1y, p0y), bz = __fsub_rn(p1z, p0z);
    const float cx = dop(ay, bz, az, by);
    const float cy = dop(az, bx, ax, bz);
    const float cz = dop(ax, by, ay, bx);
    const float sq = __fmaf_rn(cz, cz, __fmaf_rn(cx, cx, __fmul_rn(cy, cy)));
    return sq == 0.0f;
}

// :229-249 the ray-only part of the test: the permutation that makes the largest |d| component z
// (vector3.h:103-105) and the shear constants -- three IEEE divides that depend on the direction alone, so the
// trace kernels evaluate them once per ray (at ray fetch) instead of once per (ray, triangle) pair.
struct RayShear {
    float Sx, Sy, Sz;
    int kz;
};

__device__ __forceinline__ RayShear ray_shear(float dirx, float diry, float dirz) {
    const float adx = fabsf(dirx), ady = fabsf(diry), adz = fabsf(dirz);
    RayShear r;
    r.kz = (adx > ady) ? ((adx > adz) ? 0 : 2) : ((ady > adz) ? 1 : 2);
    const float dx = r.kz == 0 ? diry : (r.kz == 1 ? dirz : dirx); // kx = kz + 1
    const float dy = r.kz == 0 ? dirz : (r.kz == 1 ? dirx : diry); // ky = kx + 1
    const float dz = r.kz == 0 ? dirx : (r.kz == 1 ? diry : dirz);
    r.Sx = __fdiv_rn(-dx, dz);
    r.Sy = __fdiv_rn(-dy, dz);
    r.Sz = __fdiv_rn(1.0f, dz);
    return r;
}

// the per-(ray, triangle) part; the caller has rejected degenerate triangles (leaf-record flag)
__device__ __forceinline__ bool intersect_triangle_sheared(float ox, float oy, float oz, const RayShear &rs, float t_max,
                                                           float p0x_, float p0y_, float p0z_, float p1x_, float p1y_,
                                                           float p1z_, float p2x_, float p2y_, float p2z_, TriHit &out) {
    // :224-226 translate to the ray origin
    const float q0x = __fsub_rn(p0x_, ox), q0y = __fsub_rn(p0y_, oy), q0z = __fsub_rn(p0z_, oz);
    const float q1x = __fsub_rn(p1x_, ox), q1y = __fsub_rn(p1y_, oy), q1z = __fsub_rn(p1z_, oz);
    const float q2x = __fsub_rn(p2x_, ox), q2y = __fsub_rn(p2y_, oy), q2z = __fsub_rn(p2z_, oz);

    // :229-238 permute (selects, no divergence: kz differs from lane to lane)
    const bool k0 = rs.kz == 0, k1 = rs.kz == 1;
    float p0x = k0 ? q0y : (k1 ? q0z : q0x), p0y = k0 ? q0z : (k1 ? q0x : q0y), p0z = k0 ? q0x : (k1 ? q0y : q0z);
    float p1x = k0 ? q1y : (k1 ? q1z : q1x), p1y = k0 ? q1z : (k1 ? q1x : q1y), p1z = k0 ? q1x : (k1 ? q1y : q1z);
    float p2x = k0 ? q2y : (k1 ? q2z : q2x), p2y = k0 ? q2z : (k1 ? q2x : q2y), p2z = k0 ? q2x : (k1 ? q2y : q2z);

    // :241-249 shear
    const float Sx = rs.Sx, Sy = rs.Sy, Sz = rs.Sz;
    p0x = __fmaf_rn(Sx, p0z, p0x);
    p0y = __fmaf_rn(Sy, p0z, p0y);
    p1x = __fmaf_rn(Sx, p1z, p1x);
    p1y = __fmaf_rn(Sy, p1z, p1y);
    p2x = __fmaf_rn(Sx, p2z, p2x);
    p2y = __fmaf_rn(Sy, p2z, p2y);

    // :252-254 edge functions
    float e0 = dop(p1x, p2y, p1y, p2x);
    float e1 = dop(p2x, p0y, p2y, p0x);
    float e2 = dop(p0x, p1y, p0y, p1x);

    // :257-267 double-precision fallback when an edge function is exactly zero
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        e0 = __double2float_rn(__dsub_rn(__dmul_rn((double)p2y, (double)p1x),
                                         __dmul_rn((double)p2x, (double)p1y)));
        e1 = __double2float_rn(__dsub_rn(__dmul_rn((double)p0y, (double)p2x),
                                         __dmul_rn((double)p0x, (double)p2y)));
        e2 = __double2float_rn(__dsub_rn(__dmul_rn((double)p1y, (double)p0x),
                                         __dmul_rn((double)p1x, (double)p0y)));
    }

    // :270-277 sign / determinant tests
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    const float det = __fadd_rn(__fadd_rn(e0, e1), e2);
    if (det == 0) return false;

    // :280-290 scaled hit distance, tested against t_max * det (accepts t == t_max)
    p0z = __fmul_rn(Sz, p0z);
    p1z = __fmul_rn(Sz, p1z);
    p2z = __fmul_rn(Sz, p2z);
    const float tScaled = __fmaf_rn(p2z, e2, __fmaf_rn(p0z, e0, __fmul_rn(p1z, e1)));
    const float tlim = __fmul_rn(t_max, det);
    if (det < 0 && (tScaled >= 0 || tScaled < tlim)) return false;
    if (det > 0 && (tScaled <= 0 || tScaled > tlim)) return false;

    // :293-295 barycentrics and t
    const float invDet = __fdiv_rn(1.0f, det);
    const float b0 = __fmul_rn(e0, invDet), b1 = __fmul_rn(e1, invDet), b2 = __fmul_rn(e2, invDet);
    const float t = __fmul_rn(tScaled, invDet);

    // :299-319 conservative t > delta_t test
    const float maxZt = max3f(fabsf(p0z), fabsf(p1z), fabsf(p2z));
    const float deltaZ = __fmul_rn(maxZt, B200PT_GAMMA3);
    const float maxXt = max3f(fabsf(p0x), fabsf(p1x), fabsf(p2x));
    const float maxYt = max3f(fabsf(p0y), fabsf(p1y), fabsf(p2y));
    const float deltaX = __fmul_rn(__fadd_rn(maxZt, maxXt), B200PT_GAMMA5);
    const float deltaY = __fmul_rn(__fadd_rn(maxZt, maxYt), B200PT_GAMMA5);
    const float s = __fmaf_rn(deltaX, maxYt,
                              __fmaf_rn(__fmul_rn(maxXt, B200PT_GAMMA2), maxYt, __fmul_rn(maxXt, deltaY)));
    const float deltaE = __fadd_rn(s, s);
    const float maxE = max3f(fabsf(e0), fabsf(e1), fabsf(e2));
    const float u = __fmaf_rn(deltaZ, maxE,
                              __fmaf_rn(maxZt, deltaE, __fmul_rn(maxZt, __fmul_rn(maxE, B200PT_GAMMA3))));
    const float deltaT = __fmul_rn(fabsf(invDet), __fmul_rn(u, 3.0f));
    if (t <= deltaT) return false;

    out.b0 = b0;
    out.b1 = b1;
    out.b2 = b2;
    out.t = t;
    return true;
}

// the whole predicate as the reference evaluates it per call (probe entry point, light sampling, per-thread walks)
__device__ __forceinline__ bool intersect_triangle(float ox, float oy, float oz, float dirx, float diry, float dirz,
                                                   float t_max, float p0x, float p0y, float p0z, float p1x, float p1y,
                                                   float p1z, float p2x, float p2y, float p2z, TriHit &out) {
    if (triangle_is_degenerate(p0x, p0y, p0z, p1x, p1y, p1z, p2x, p2y, p2z)) return false;
    const RayShear rs = ray_shear(dirx, diry, dirz);
    return intersect_triangle_sheared(ox, oy, oz, rs, t_max, p0x, p0y, p0z, p1x, p1y, p1z, p2x, p2y, p2z, out);
}

} // namespace b200pt
