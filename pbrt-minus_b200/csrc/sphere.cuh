// sphere.cuh -- ray/sphere intersection in object space with interval arithmetic, and the
// surface interaction of a sphere hit.
//   Sphere::basic_intersect              shapes/sphere.cu:35-129
//   Sphere::interaction_from_intersection shapes/sphere.h:71-125
//   Interval                             euclidean_space/interval.h (directed rounding: the reference's
//                                        device build uses __fadd_rd/__fadd_ru etc., util/bit_arithmetic.h)
//   Transform on Point3fi / Vector3fi    euclidean_space/transform.h:270-427
//   Disk::basic_intersect                shapes/disk.cu:99-131
// Spheres / disks are few (one in the reference's shipped scene) and are tested linearly beside the BVH.
#pragma once

#include "common.cuh"
#include "vecmath.cuh"

namespace b200pt {

struct Iv {
    float lo, hi;
};
__device__ __forceinline__ Iv iv(float v) { return Iv{v, v}; }
__device__ __forceinline__ Iv iv_err(float v, float e) {
    if (e == 0.f) return Iv{v, v};
    return Iv{__fsub_rd(v, e), __fadd_ru(v, e)};
}
__device__ __forceinline__ float iv_mid(Iv a) { return (a.lo + a.hi) / 2.f; }
__device__ __forceinline__ Iv operator+(Iv a, Iv b) { return Iv{__fadd_rd(a.lo, b.lo), __fadd_ru(a.hi, b.hi)}; }
__device__ __forceinline__ Iv operator-(Iv a, Iv b) { return Iv{__fsub_rd(a.lo, b.hi), __fsub_ru(a.hi, b.lo)}; }
__device__ __forceinline__ Iv operator-(Iv a) { return Iv{-a.hi, -a.lo}; }
__device__ __forceinline__ Iv operator*(Iv a, Iv b) {
    const float l0 = __fmul_rd(a.lo, b.lo), l1 = __fmul_rd(a.hi, b.lo), l2 = __fmul_rd(a.lo, b.hi), l3 = __fmul_rd(a.hi, b.hi);
    const float h0 = __fmul_ru(a.lo, b.lo), h1 = __fmul_ru(a.hi, b.lo), h2 = __fmul_ru(a.lo, b.hi), h3 = __fmul_ru(a.hi, b.hi);
    return Iv{fminf(fminf(l0, l1), fminf(l2, l3)), fmaxf(fmaxf(h0, h1), fmaxf(h2, h3))};
}
__device__ __forceinline__ Iv operator*(float f, Iv a) {
    if (f > 0.f) return Iv{__fmul_rd(f, a.lo), __fmul_ru(f, a.hi)};
    return Iv{__fmul_rd(f, a.hi), __fmul_ru(f, a.lo)};
}
__device__ __forceinline__ Iv operator/(Iv a, Iv b) {
    if (b.lo <= 0.f && b.hi >= 0.f) return Iv{-kInf, kInf};
    const float l0 = __fdiv_rd(a.lo, b.lo), l1 = __fdiv_rd(a.hi, b.lo), l2 = __fdiv_rd(a.lo, b.hi), l3 = __fdiv_rd(a.hi, b.hi);
    const float h0 = __fdiv_ru(a.lo, b.lo), h1 = __fdiv_ru(a.hi, b.lo), h2 = __fdiv_ru(a.lo, b.hi), h3 = __fdiv_ru(a.hi, b.hi);
    return Iv{fminf(fminf(l0, l1), fminf(l2, l3)), fmaxf(fmaxf(h0, h1), fmaxf(h2, h3))};
}
__device__ __forceinline__ Iv iv_sqr(Iv a) {
    float alow = fabsf(a.lo), ahigh = fabsf(a.hi);
    if (alow > ahigh) {
        const float t = alow;
        alow = ahigh;
        ahigh = t;
    }
    if (a.lo <= 0.f && a.hi >= 0.f) return Iv{0.f, __fmul_ru(ahigh, ahigh)};
    return Iv{__fmul_rd(alow, alow), __fmul_ru(ahigh, ahigh)};
}
__device__ __forceinline__ Iv iv_sqrt(Iv a) { return Iv{__fsqrt_rd(a.lo), __fsqrt_ru(a.hi)}; }

struct SphereHit {
    float t;
    V3 p_obj;
    float phi;
};

// rows 0..2 of a 4x4, applied to a point / a vector
__device__ __forceinline__ V3 xf_point(const float *m, V3 p) {
    return v3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
__device__ __forceinline__ V3 xf_vector(const float *m, V3 v) {
    return v3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
// normals transform by the inverse transpose: pass the INVERSE matrix
__device__ __forceinline__ V3 xf_normal(const float *inv, V3 n) {
    return v3(inv[0] * n.x + inv[4] * n.y + inv[8] * n.z, inv[1] * n.x + inv[5] * n.y + inv[9] * n.z,
              inv[2] * n.x + inv[6] * n.y + inv[10] * n.z);
}
__device__ __forceinline__ V3 xf_point_error(const float *m, V3 p, V3 in_err, bool exact) {
    const float g3 = gamma_n(3);
    V3 e;
    float *ev = &e.x;
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const float *row = m + 4 * r;
        const float a = fabsf(row[0] * p.x) + fabsf(row[1] * p.y) + fabsf(row[2] * p.z) + fabsf(row[3]);
        if (exact) ev[r] = g3 * a;
        else ev[r] = (g3 + 1.f) * (fabsf(row[0]) * in_err.x + fabsf(row[1]) * in_err.y + fabsf(row[2]) * in_err.z) + g3 * a;
    }
    return e;
}
__device__ __forceinline__ V3 xf_vector_error(const float *m, V3 v) {
    const float g3 = gamma_n(3);
    return v3(g3 * (fabsf(m[0] * v.x) + fabsf(m[1] * v.y) + fabsf(m[2] * v.z)), g3 * (fabsf(m[4] * v.x) + fabsf(m[5] * v.y) + fabsf(m[6] * v.z)),
              g3 * (fabsf(m[8] * v.x) + fabsf(m[9] * v.y) + fabsf(m[10] * v.z)));
}

// Transform::operator()(Point3<T>) with T = Interval (euclidean_space/transform.h:255-268): the reference has no dedicated
// Point3fi overload for the FORWARD direction, so the ray origin brought to object space and the hit point brought to render
// space go through plain interval arithmetic, term by term from the left.  The bounds this gives are tighter than the
// gamma(3) formula of apply_inverse; for a disk (zero thickness, zero pError) they decide whether a spawned ray clears the
// surface, so they are reproduced as they are.
__device__ __forceinline__ void xf_point_iv(const float *m, Iv px, Iv py, Iv pz, Iv out[3]) {
#pragma unroll
    for (int r = 0; r < 3; r++) out[r] = m[4 * r] * px + m[4 * r + 1] * py + m[4 * r + 2] * pz + iv(m[4 * r + 3]);
}
__device__ __forceinline__ V3 iv_mid3(const Iv v[3]) { return v3(iv_mid(v[0]), iv_mid(v[1]), iv_mid(v[2])); }
__device__ __forceinline__ V3 iv_err3(const Iv v[3]) { return v3((v[0].hi - v[0].lo) / 2.f, (v[1].hi - v[1].lo) / 2.f, (v[2].hi - v[2].lo) / 2.f); }

// Transform::apply_inverse(Ray, &t_max) (euclidean_space/transform.cu:36-54, transform.h:380-420): a render-space ray into the
// space of an object instance (TransformedPrimitive::intersect, primitives/transformed_primitive.cu:20-34).  The origin is an
// exact point, transformed with the (m00 x + m01 y) + (m02 z + m03) association and pushed along the ray to the edge of its
// error bound; t_max shrinks by the same dt.  The direction is NOT renormalised: t means the same along both rays (up to dt).
__device__ __forceinline__ void instance_ray(const float *inv, V3 o, V3 d, V3 *oi, V3 *di, float *dt) {
    const V3 p = v3((inv[0] * o.x + inv[1] * o.y) + (inv[2] * o.z + inv[3]), (inv[4] * o.x + inv[5] * o.y) + (inv[6] * o.z + inv[7]),
                    (inv[8] * o.x + inv[9] * o.y) + (inv[10] * o.z + inv[11]));
    const float g3 = gamma_n(3);
    const V3 e = v3(g3 * (fabsf(inv[0] * o.x) + fabsf(inv[1] * o.y) + fabsf(inv[2] * o.z)),
                    g3 * (fabsf(inv[4] * o.x) + fabsf(inv[5] * o.y) + fabsf(inv[6] * o.z)),
                    g3 * (fabsf(inv[8] * o.x) + fabsf(inv[9] * o.y) + fabsf(inv[10] * o.z)));
    const V3 dd = xf_vector(inv, d);
    // Point3fi(p, e): the interval [p - e, p + e] with outward rounding; its midpoint and half width are what the ray uses
    const Iv ix = iv_err(p.x, e.x), iy = iv_err(p.y, e.y), iz = iv_err(p.z, e.z);
    V3 po = v3(iv_mid(ix), iv_mid(iy), iv_mid(iz));
    const float l2 = len2(dd);
    *dt = 0.f;
    if (l2 > 0.f) {
        const V3 oerr = v3((ix.hi - ix.lo) / 2.f, (iy.hi - iy.lo) / 2.f, (iz.hi - iz.lo) / 2.f);
        *dt = dot(vabs(dd), oerr) / l2;
        // Point3fi += Vector3f * dt, then to_point3f(): interval arithmetic, midpoint
        const Iv jx = ix + iv(dd.x * *dt), jy = iy + iv(dd.y * *dt), jz = iz + iv(dd.z * *dt);
        po = v3(iv_mid(jx), iv_mid(jy), iv_mid(jz));
    }
    *oi = po;
    *di = dd;
}

__device__ __forceinline__ bool sphere_clip(const SphereRec &s, V3 &pHit, float &phi) {
    if (pHit.x == 0.f && pHit.y == 0.f) pHit.x = 1e-5f * s.radius;
    phi = atan2f(pHit.y, pHit.x);
    if (phi < 0.f) phi += 2.f * kPi;
    return (s.z_min > -s.radius && pHit.z < s.z_min) || (s.z_max < s.radius && pHit.z > s.z_max) || phi > s.phi_max;
}

// Disk::basic_intersect (shapes/disk.cu:99-131): the plane z = height in object space, from the MIDPOINTS of the transformed
// origin / direction intervals; inside [inner radius, radius] and phi <= phi_max
__device__ __forceinline__ bool disk_intersect(const SphereRec &s, V3 ro, V3 rd, float t_max, SphereHit &hit) {
    Iv oi[3];
    xf_point_iv(s.ofr, iv(ro.x), iv(ro.y), iv(ro.z), oi);
    const V3 dc = xf_vector(s.ofr, rd), de = xf_vector_error(s.ofr, rd);
    const V3 o = iv_mid3(oi);
    const V3 d = v3(iv_mid(iv_err(dc.x, de.x)), iv_mid(iv_err(dc.y, de.y)), iv_mid(iv_err(dc.z, de.z)));
    if (d.z == 0.f) return false;
    const float height = s.z_min, inner = s.z_max;
    const float t = (height - o.z) / d.z;
    if (t <= 0.f || t >= t_max) return false;
    const V3 pHit = o + t * d;
    const float dist2 = sqr(pHit.x) + sqr(pHit.y);
    if (dist2 > sqr(s.radius) || dist2 < sqr(inner)) return false;
    float phi = atan2f(pHit.y, pHit.x);
    if (phi < 0.f) phi += 2.f * kPi;
    if (phi > s.phi_max) return false;
    hit.t = t;
    hit.p_obj = pHit;
    hit.phi = phi;
    return true;
}

__device__ __forceinline__ bool sphere_intersect(const SphereRec &s, V3 ro, V3 rd, float t_max, SphereHit &hit) {
    if (s.kind == 1) return disk_intersect(s, ro, rd, t_max, hit);
    const V3 o = xf_point(s.ofr, ro), d = xf_vector(s.ofr, rd);
    const V3 oe = xf_point_error(s.ofr, ro, v3(0.f, 0.f, 0.f), true), de = xf_vector_error(s.ofr, rd);
    const Iv ox = iv_err(o.x, oe.x), oy = iv_err(o.y, oe.y), oz = iv_err(o.z, oe.z);
    const Iv dx = iv_err(d.x, de.x), dy = iv_err(d.y, de.y), dz = iv_err(d.z, de.z);
    const Iv a = iv_sqr(dx) + iv_sqr(dy) + iv_sqr(dz);
    const Iv b = 2.f * (dx * ox + dy * oy + dz * oz);
    const Iv c = iv_sqr(ox) + iv_sqr(oy) + iv_sqr(oz) - iv_sqr(iv(s.radius));
    const Iv f = b / (2.f * a);
    const Iv vx = ox - f * dx, vy = oy - f * dy, vz = oz - f * dz;
    const Iv length = iv_sqrt(iv_sqr(vx) + iv_sqr(vy) + iv_sqr(vz));
    const Iv discrim = 4.f * a * (iv(s.radius) + length) * (iv(s.radius) - length);
    if (discrim.lo < 0.f) return false;
    const Iv root = iv_sqrt(discrim);
    const Iv q = (iv_mid(b) < 0.f) ? -.5f * (b - root) : -.5f * (b + root);
    Iv t0 = q / a, t1 = c / q;
    if (t0.lo > t1.lo) {
        const Iv tmp = t0;
        t0 = t1;
        t1 = tmp;
    }
    if (t0.hi > t_max || t1.lo <= 0.f) return false;
    Iv ts = t0;
    bool is_t1 = false;
    if (ts.lo <= 0.f) {
        ts = t1;
        is_t1 = true;
        if (ts.hi > t_max) return false;
    }
    V3 pHit = o + iv_mid(ts) * d;
    pHit = pHit * (s.radius / len(pHit));
    float phi;
    if (sphere_clip(s, pHit, phi)) {
        if (is_t1) return false;
        if (t1.hi > t_max) return false;
        ts = t1;
        pHit = o + iv_mid(ts) * d;
        pHit = pHit * (s.radius / len(pHit));
        if (sphere_clip(s, pHit, phi)) return false;
    }
    hit.t = iv_mid(ts);
    hit.p_obj = pHit;
    hit.phi = phi;
    return true;
}

} // namespace b200pt
