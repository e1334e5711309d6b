// light.cuh -- emissive triangles: DiffuseAreaLight over Triangle, and the power light sampler.
//   DiffuseAreaLight::{l,sample_li,pdf_li}   lights/diffuse_area_light.cu:33-73
//   Triangle::sample(ctx,u) / pdf(ctx,wi)    shapes/triangle.cu:24-72,122-211
//   solid_angle / spherical_triangle_area    shapes/triangle.h:24-27,86-95
//   sample_spherical_triangle                util/sampling.h:166-256  (incl. the b1 /= b1+b2; b2 /= b1+b2 order)
//   InvertSphericalTriangleSample            util/sampling.cu:42-104
//   sample_bilinear / bilinear_pdf           util/sampling.h:88-116   (w[0] and w[1] both use wi[1], triangle.cu:63-66)
//   PowerLightSampler / AliasTable::sample   light_samplers/power_light_sampler.cu:33-63, distribution/alias_table.cu:70-87
// The reference reaches a light through Light* -> DiffuseAreaLight -> Shape* -> Triangle -> mesh
// and finds pmf(light) through a pointer hash map; here a light is one 64-byte record holding its
// vertices, and the hit triangle carries its light id, so pmf is an array lookup.
#pragma once

#include "common.cuh"
#include "sampler.cuh"
#include "sphere.cuh"
#include "triangle.cuh"
#include "vecmath.cuh"

namespace b200pt {

struct LightRec { // 64 B
    float p0[3], p1[3], p2[3]; // AREA: vertices.  DISTANT: p0 = render_from_light(0,0,1).  SPOT: p0 = position
    float scale;      // already divided by the photometric integral
    int32_t spectrum; // Lemit
    int32_t flags;    // bit 0: two_sided, bit 1: flip normal (reverse_orientation ^ swaps_handedness), bits 4..6: B200PT_LIGHT_*
    float pmf;        // power_light_sampler pdfs[idx]
    float area;
    int32_t prim;     // primitive (triangle) index of this light: its own shadow rays skip it (-1: none)
    int32_t ext;      // SPOT: index into the LightExt array
};
__device__ __host__ __forceinline__ int light_type(const LightRec &L) { return (L.flags >> 4) & 7; }

struct LightExt { // 48 B, spot lights only (lights/spot_light.h)
    float m[9];   // upper 3x3 of light_from_render
    float cos_start, cos_end, pad;
};

struct AliasBin {
    float p;
    int32_t first, second;
};

struct LightCtx { // LightSampleContext / ShapeSampleContext
    V3 p, err, n, ns;
};

constexpr float kMinSphericalSampleArea = 3e-4f;
constexpr float kMaxSphericalSampleArea = 6.22f;

__device__ __forceinline__ float spherical_triangle_area(V3 a, V3 b, V3 c) {
    return fabsf(2.f * atan2f(dot(a, cross(b, c)), 1.f + dot(a, b) + dot(a, c) + dot(b, c)));
}
__device__ __forceinline__ float tri_solid_angle(V3 p0, V3 p1, V3 p2, V3 p) {
    return spherical_triangle_area(normalize(p0 - p), normalize(p1 - p), normalize(p2 - p));
}
__device__ __forceinline__ float tri_area(V3 p0, V3 p1, V3 p2) { return 0.5f * len(cross(p1 - p0, p2 - p0)); }

__device__ __forceinline__ float sample_linear(float u, float a, float b) {
    if (u == 0.f && a == 0.f) return 0.f;
    const float x = u * (a + b) / (a + sqrtf(lerpf(u, sqr(a), sqr(b))));
    return fminf(x, kOneMinusEps);
}
__device__ __forceinline__ float2 sample_bilinear(float2 u, const float w[4]) {
    float2 p;
    p.y = sample_linear(u.y, w[0] + w[1], w[2] + w[3]);
    p.x = sample_linear(u.x, lerpf(p.y, w[0], w[2]), lerpf(p.y, w[1], w[3]));
    return p;
}
__device__ __forceinline__ float bilinear_pdf(float2 p, const float w[4]) {
    if (p.x < 0.f || p.x > 1.f || p.y < 0.f || p.y > 1.f) return 0.f;
    if (w[0] + w[1] + w[2] + w[3] == 0.f) return 1.f;
    return 4.f * ((1.f - p.x) * (1.f - p.y) * w[0] + p.x * (1.f - p.y) * w[1] + (1.f - p.x) * p.y * w[2] + p.x * p.y * w[3]) /
           (w[0] + w[1] + w[2] + w[3]);
}
__device__ __forceinline__ void sample_uniform_triangle(float2 u, float b[3]) {
    float b0, b1;
    if (u.x < u.y) {
        b0 = u.x / 2.f;
        b1 = u.y - b0;
    } else {
        b1 = u.y / 2.f;
        b0 = u.x - b1;
    }
    b[0] = b0;
    b[1] = b1;
    b[2] = 1.f - b0 - b1;
}

// returns false when pdf == 0 (degenerate spherical triangle)
__device__ __forceinline__ bool sample_spherical_triangle(const V3 v[3], V3 p, float2 u, float out[3], float *pdf) {
    *pdf = 0.f;
    V3 a = normalize(v[0] - p), b = normalize(v[1] - p), c = normalize(v[2] - p);
    V3 n_ab = cross(a, b), n_bc = cross(b, c), n_ca = cross(c, a);
    if (len2(n_ab) == 0.f || len2(n_bc) == 0.f || len2(n_ca) == 0.f) return false;
    n_ab = normalize(n_ab);
    n_bc = normalize(n_bc);
    n_ca = normalize(n_ca);
    const float alpha = angle_between(n_ab, -n_ca);
    const float beta = angle_between(n_bc, -n_ab);
    const float gamma = angle_between(n_ca, -n_bc);
    const float A_pi = alpha + beta + gamma;
    const float Ap_pi = lerpf(u.x, kPi, A_pi);
    const float A = A_pi - kPi;
    *pdf = (A <= 0.f) ? 0.f : 1.f / A;
    float cosAlpha, sinAlpha, sinAp, cosAp;
    sincosf(alpha, &sinAlpha, &cosAlpha);
    sincosf(Ap_pi, &sinAp, &cosAp);
    const float sinPhi = sinAp * cosAlpha - cosAp * sinAlpha;
    const float cosPhi = cosAp * cosAlpha + sinAp * sinAlpha;
    const float k1 = cosPhi + cosAlpha;
    const float k2 = sinPhi - sinAlpha * dot(a, b);
    float cosBp = (k2 + diff_of_products(k2, cosPhi, k1, sinPhi) * cosAlpha) / (sum_of_products(k2, sinPhi, k1, cosPhi) * sinAlpha);
    cosBp = clampf(cosBp, -1.f, 1.f);
    const float sinBp = safe_sqrt(1.f - sqr(cosBp));
    const V3 cp = cosBp * a + sinBp * normalize(gram_schmidt(c, a));
    const float cosTheta = 1.f - u.y * (1.f - dot(cp, b));
    const float sinTheta = safe_sqrt(1.f - sqr(cosTheta));
    const V3 w = cosTheta * b + sinTheta * normalize(gram_schmidt(cp, b));
    const V3 e1 = v[1] - v[0], e2 = v[2] - v[0];
    const V3 s1 = cross(w, e2);
    const float divisor = dot(s1, e1);
    if (divisor == 0.f) {
        out[0] = out[1] = out[2] = 1.f / 3.f;
        return *pdf != 0.f;
    }
    const float inv = 1.f / divisor;
    const V3 s = p - v[0];
    float b1 = dot(s, s1) * inv;
    float b2 = dot(w, cross(s, e1)) * inv;
    b1 = clampf(b1, 0.f, 1.f);
    b2 = clampf(b2, 0.f, 1.f);
    if (b1 + b2 > 1.f) {
        b1 /= b1 + b2;
        b2 /= b1 + b2; // uses the updated b1, as the reference does
    }
    out[0] = 1.f - b1 - b2;
    out[1] = b1;
    out[2] = b2;
    return *pdf != 0.f;
}

__device__ __forceinline__ float2 invert_spherical_triangle_sample(const V3 v[3], V3 p, V3 w) {
    V3 a = normalize(v[0] - p), b = normalize(v[1] - p), c = normalize(v[2] - p);
    V3 n_ab = cross(a, b), n_bc = cross(b, c), n_ca = cross(c, a);
    if (len2(n_ab) == 0.f || len2(n_bc) == 0.f || len2(n_ca) == 0.f) return make_float2(0.f, 0.f);
    n_ab = normalize(n_ab);
    n_bc = normalize(n_bc);
    n_ca = normalize(n_ca);
    const float alpha = angle_between(n_ab, -n_ca);
    const float beta = angle_between(n_bc, -n_ab);
    const float gamma = angle_between(n_ca, -n_bc);
    V3 cp = normalize(cross(cross(b, w), cross(c, a)));
    if (dot(cp, a + c) < 0.f) cp = -cp;
    float u0;
    if (dot(a, cp) > 0.99999847691f) {
        u0 = 0.f;
    } else {
        V3 n_cpb = cross(cp, b), n_acp = cross(a, cp);
        if (len2(n_cpb) == 0.f || len2(n_acp) == 0.f) return make_float2(0.5f, 0.5f);
        n_cpb = normalize(n_cpb);
        n_acp = normalize(n_acp);
        const float Ap = alpha + angle_between(n_ab, n_cpb) + angle_between(n_acp, -n_cpb) - kPi;
        const float A = alpha + beta + gamma - kPi;
        u0 = Ap / A;
    }
    const float u1 = (1.f - dot(w, b)) / (1.f - dot(cp, b));
    return make_float2(clampf(u0, 0.f, 1.f), clampf(u1, 0.f, 1.f));
}

struct LightSample {
    V3 wi;
    float pdf;        // solid-angle pdf of the shape sample (not yet times the light pmf)
    V3 p, err, n;     // sampled point, its error bound, the light's geometric normal
};

__device__ __forceinline__ void light_points(const LightRec &L, V3 v[3]) {
    v[0] = v3(L.p0[0], L.p0[1], L.p0[2]);
    v[1] = v3(L.p1[0], L.p1[1], L.p1[2]);
    v[2] = v3(L.p2[0], L.p2[1], L.p2[2]);
}

// cos-theta weights of the warp-product sampling; w[0] and w[1] both use wi[1] as the reference does
__device__ __forceinline__ void bilinear_weights(const V3 v[3], V3 rp, V3 ns, float w[4]) {
    const V3 w0 = normalize(v[0] - rp), w1 = normalize(v[1] - rp), w2 = normalize(v[2] - rp);
    w[0] = fmaxf(0.01f, absdot(ns, w1));
    w[1] = fmaxf(0.01f, absdot(ns, w1));
    w[2] = fmaxf(0.01f, absdot(ns, w0));
    w[3] = fmaxf(0.01f, absdot(ns, w2));
}

// Triangle::sample(ctx, u) + the DiffuseAreaLight::sample_li geometry (everything except Le)
__device__ __forceinline__ bool light_sample(const LightRec &L, const LightCtx &ctx, float2 u, LightSample &ls) {
    V3 v[3];
    light_points(L, v);
    V3 n = normalize(cross(v[1] - v[0], v[2] - v[0]));
    if (L.flags & 2) n = -n;
    const float sa = tri_solid_angle(v[0], v[1], v[2], ctx.p);
    float b[3];
    float pdf;
    if (sa < kMinSphericalSampleArea || sa > kMaxSphericalSampleArea) {
        sample_uniform_triangle(u, b);
        const V3 p = b[0] * v[0] + b[1] * v[1] + b[2] * v[2];
        V3 wi = p - ctx.p;
        if (len2(wi) == 0.f) return false;
        wi = normalize(wi);
        pdf = 1.f / tri_area(v[0], v[1], v[2]);
        pdf /= absdot(n, -wi) / len2(ctx.p - p);
        if (isinf(pdf)) return false;
    } else {
        pdf = 1.f;
        if (ctx.ns.x != 0.f || ctx.ns.y != 0.f || ctx.ns.z != 0.f) {
            float w[4];
            bilinear_weights(v, ctx.p, ctx.ns, w);
            u = sample_bilinear(u, w);
            pdf = bilinear_pdf(u, w);
        }
        float tri_pdf;
        if (!sample_spherical_triangle(v, ctx.p, u, b, &tri_pdf)) return false;
        pdf *= tri_pdf;
    }
    const V3 p = b[0] * v[0] + b[1] * v[1] + b[2] * v[2];
    const V3 abs_sum = vabs(b[0] * v[0]) + vabs(b[1] * v[1]) + vabs((1.f - b[0] - b[1]) * v[2]);
    ls.p = p;
    ls.err = gamma_n(6) * abs_sum;
    ls.n = n;
    ls.pdf = pdf;
    if (pdf == 0.f || len2(p - ctx.p) == 0.f) return false;
    ls.wi = normalize(p - ctx.p);
    return true;
}

// util/math.h:211-218
__device__ __forceinline__ float smooth_step(float x, float a, float b) {
    if (a == b) return (x < a) ? 0.f : 1.f;
    const float t = clampf((x - a) / (b - a), 0.f, 1.f);
    return t * t * (3.f - 2.f * t);
}
// util/sampling.h:259-266
__device__ __forceinline__ V3 sample_uniform_sphere(float2 u) {
    const float z = 1.f - 2.f * u.x;
    const float r = safe_sqrt(1.f - sqr(z));
    const float phi = 2.f * kPi * u.y;
    float sn, cs;
    sincosf(phi, &sn, &cs);
    return v3(r * cs, r * sn, z);
}
constexpr float kUniformSpherePdf = 0.07957747154594767f; // 1 / (4 pi)

// sample_li of the lights that are not emissive triangles.  Li = weight * scale * Lemit(lambda) / dist2;
// returns false when the sample carries no radiance.  ls.n = ls.err = 0: Interaction(p) has no normal and
// no error bounds, so spawn_ray_to leaves that end of the segment where it is (base/ray.h:51-58).
//   DistantLight::sample_li           lights/distant_light.cu:43-49   (delta)
//   UniformInfiniteLight::sample_li   lights/uniform_infinite_light.cu:39-48
//   SpotLight::sample_li / I          lights/spot_light.cu:43-60,110-114   (delta)
__device__ __forceinline__ bool light_sample_other(const LightRec &L, int type, const LightExt *ext, float scene_radius,
                                                   const LightCtx &ctx, float2 u, LightSample &ls, float *weight, float *dist2, bool *delta) {
    ls.err = v3(0.f, 0.f, 0.f);
    ls.n = v3(0.f, 0.f, 0.f);
    *weight = 1.f;
    *dist2 = 1.f;
    if (type == B200PT_LIGHT_DISTANT) {
        ls.wi = normalize(v3(L.p0[0], L.p0[1], L.p0[2]));
        ls.p = ctx.p + ls.wi * (2.f * scene_radius);
        ls.pdf = 1.f;
        *delta = true;
        return true;
    }
    if (type == B200PT_LIGHT_UNIFORM_INFINITE) {
        ls.wi = sample_uniform_sphere(u);
        ls.p = ctx.p + ls.wi * (2.f * scene_radius);
        ls.pdf = kUniformSpherePdf;
        *delta = false;
        return true;
    }
    // spot
    const V3 p = v3(L.p0[0], L.p0[1], L.p0[2]);
    ls.p = p;
    ls.wi = normalize(p - ctx.p);
    const LightExt e = ext[L.ext];
    const V3 w = -ls.wi;
    const V3 wl = normalize(v3(e.m[0] * w.x + e.m[1] * w.y + e.m[2] * w.z, e.m[3] * w.x + e.m[4] * w.y + e.m[5] * w.z,
                               e.m[6] * w.x + e.m[7] * w.y + e.m[8] * w.z));
    *weight = smooth_step(wl.z, e.cos_end, e.cos_start);
    *dist2 = len2(p - ctx.p);
    ls.pdf = 1.f;
    *delta = true;
    return *weight > 0.f;
}

// DiffuseAreaLight::l: one-sided unless two_sided
__device__ __forceinline__ bool light_emits_towards(const LightRec &L, V3 n, V3 w) { return (L.flags & 1) || dot(n, w) >= 0.f; }

// Triangle::pdf(ctx, wi) (solid-angle pdf of sampling direction wi towards this light from ctx)
__device__ __forceinline__ float light_pdf(const LightRec &L, const LightCtx &ctx, V3 wi) {
    V3 v[3];
    light_points(L, v);
    const float sa = tri_solid_angle(v[0], v[1], v[2], ctx.p);
    if (sa < kMinSphericalSampleArea || sa > kMaxSphericalSampleArea) {
        const V3 o = offset_ray_origin(ctx.p, ctx.err, ctx.n, wi);
        TriHit th;
        if (!intersect_triangle(o.x, o.y, o.z, wi.x, wi.y, wi.z, kInf, v[0].x, v[0].y, v[0].z, v[1].x, v[1].y, v[1].z, v[2].x,
                                v[2].y, v[2].z, th))
            return 0.f;
        const V3 ph = th.b0 * v[0] + th.b1 * v[1] + th.b2 * v[2];
        const V3 n = normalize(cross(v[0] - v[2], v[1] - v[2]));
        float pdf = len2(ctx.p - ph) / (tri_area(v[0], v[1], v[2]) * absdot(n, -wi));
        if (isinf(pdf)) pdf = 0.f;
        return pdf;
    }
    float pdf = 1.f / sa;
    if (ctx.ns.x != 0.f || ctx.ns.y != 0.f || ctx.ns.z != 0.f) {
        const float2 u = invert_spherical_triangle_sample(v, ctx.p, wi);
        float w[4];
        bilinear_weights(v, ctx.p, ctx.ns, w);
        pdf *= bilinear_pdf(u, w);
    }
    return pdf;
}

// ---- DiffuseAreaLight over Shape "sphere": Sphere::sample(ctx, u) / Sphere::pdf(ctx, wi) (shapes/sphere.cu:162-280) -----------
// LightRec of such a light: prim = n_triangles + sphere index (the shadow ray's own-primitive skip), ext = sphere index,
// area = Sphere::area() (shapes/sphere.h:23-25), flags bit 1 = reverse_orientation (the sampled normal flips on it alone).
// Like the reference, the cone branch treats `radius` as a render-space length (a scaled sphere light is sampled as if unscaled).

// Sphere::sample(u) (shapes/sphere.cu:131-160): uniform over the FULL sphere's area, used when the reference point is inside
__device__ __forceinline__ void sphere_sample_area(const SphereRec &S, bool reverse, float2 u, V3 *p, V3 *err, V3 *n) {
    V3 pObj = S.radius * sample_uniform_sphere(u);
    pObj = pObj * (S.radius / len(pObj));
    const V3 pObjErr = gamma_n(5) * vabs(pObj);
    V3 nr = normalize(xf_normal(S.ofr, pObj));
    if (reverse) nr = -nr;
    *p = xf_point(S.rfo, pObj);
    *err = xf_point_error(S.rfo, pObj, pObjErr, false);
    *n = nr;
}

__device__ __forceinline__ bool sphere_light_sample(const SphereRec &S, const LightRec &L, const LightCtx &ctx, float2 u, LightSample &ls) {
    const V3 pCenter = v3(S.rfo[3], S.rfo[7], S.rfo[11]); // render_from_object(0, 0, 0)
    const V3 pOrigin = offset_ray_origin(ctx.p, ctx.err, ctx.n, pCenter - ctx.p);
    const bool reverse = (L.flags & 2) != 0;
    if (len2(pOrigin - pCenter) <= sqr(S.radius)) {
        // inside: sample by area, convert to solid angle
        V3 p, err, n;
        sphere_sample_area(S, reverse, u, &p, &err, &n);
        V3 wi = p - ctx.p;
        if (len2(wi) == 0.f) return false;
        wi = normalize(wi);
        float pdf = 1.f / L.area;
        pdf *= len2(ctx.p - p) / absdot(n, -wi);
        if (isinf(pdf)) return false;
        ls.p = p; ls.err = err; ls.n = n; ls.pdf = pdf; ls.wi = wi;
        return pdf != 0.f;
    }
    // uniform inside the subtended cone
    const float sinThetaMax = S.radius / len(ctx.p - pCenter);
    const float sin2ThetaMax = sqr(sinThetaMax);
    const float cosThetaMax = safe_sqrt(1.f - sin2ThetaMax);
    float oneMinusCosThetaMax = 1.f - cosThetaMax;
    float cosTheta = (cosThetaMax - 1.f) * u.x + 1.f;
    float sin2Theta = 1.f - sqr(cosTheta);
    if (sin2ThetaMax < 0.00068523f) { // sin^2(1.5 deg): Taylor expansion for small angles
        sin2Theta = sin2ThetaMax * u.x;
        cosTheta = sqrtf(1.f - sin2Theta);
        oneMinusCosThetaMax = sin2ThetaMax / 2.f;
    }
    const float cosAlpha = sin2Theta / sinThetaMax + cosTheta * safe_sqrt(1.f - sin2Theta / sqr(sinThetaMax));
    const float sinAlpha = safe_sqrt(1.f - sqr(cosAlpha));
    const float phi = u.y * 2.f * kPi;
    float sp, cp;
    sincosf(phi, &sp, &cp);
    const V3 w = v3(clampf(sinAlpha, -1.f, 1.f) * cp, clampf(sinAlpha, -1.f, 1.f) * sp, clampf(cosAlpha, -1.f, 1.f)); // SphericalDirection
    Frame fr;
    fr.z = normalize(pCenter - ctx.p);
    coordinate_system(fr.z, &fr.x, &fr.y);
    V3 n = fr.from_local(-w);
    const V3 p = pCenter + S.radius * n;
    if (reverse) n = -n;
    ls.p = p;
    ls.err = gamma_n(5) * vabs(p);
    ls.n = n;
    ls.pdf = 1.f / (2.f * kPi * oneMinusCosThetaMax);
    if (ls.pdf == 0.f || len2(p - ctx.p) == 0.f) return false;
    ls.wi = normalize(p - ctx.p);
    return true;
}

// normal of a sphere hit as Sphere::interaction_from_intersection leaves it in Interaction::n (shapes/sphere.h:71-125):
// needed by Sphere::pdf's inside branch
__device__ __forceinline__ V3 sphere_hit_normal(const SphereRec &S, V3 pObj) {
    V3 n = normalize(xf_normal(S.ofr, pObj));
    return S.flip ? -n : n;
}

__device__ __forceinline__ float sphere_light_pdf(const SphereRec &S, const LightRec &L, const LightCtx &ctx, V3 wi) {
    const V3 pCenter = v3(S.rfo[3], S.rfo[7], S.rfo[11]);
    const V3 pOrigin = offset_ray_origin(ctx.p, ctx.err, ctx.n, pCenter - ctx.p);
    if (len2(pOrigin - pCenter) <= sqr(S.radius)) {
        const V3 o = offset_ray_origin(ctx.p, ctx.err, ctx.n, wi);
        SphereHit sh;
        if (!sphere_intersect(S, o, wi, kInf, sh)) return 0.f;
        const V3 ph = xf_point(S.rfo, sh.p_obj);
        const V3 n = sphere_hit_normal(S, sh.p_obj);
        float pdf = len2(ctx.p - ph) / (L.area * absdot(n, -wi));
        if (isinf(pdf)) pdf = 0.f;
        return pdf;
    }
    const float sin2ThetaMax = S.radius * S.radius / len2(ctx.p - pCenter);
    const float cosThetaMax = safe_sqrt(1.f - sin2ThetaMax);
    float oneMinusCosThetaMax = 1.f - cosThetaMax;
    if (sin2ThetaMax < 0.00068523f) oneMinusCosThetaMax = sin2ThetaMax / 2.f;
    return (float)(1.0 / (2.0 * (double)kPi * (double)oneMinusCosThetaMax));
}

// AliasTable::sample (distribution/alias_table.cu:70-87): the second variate is a hash of the first
__device__ __forceinline__ int alias_sample(const AliasBin *bins, int n, float u0) {
    int idx = (int)((double)u0 * (double)n);
    idx = idx < 0 ? 0 : (idx > n - 1 ? n - 1 : idx);
    const AliasBin b = bins[idx];
    if (b.second < 0) return b.first;
    const float u1 = hash_float_fi(u0, idx);
    return u1 <= b.p ? b.first : b.second;
}

} // namespace b200pt
