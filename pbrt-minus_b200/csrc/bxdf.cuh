// bxdf.cuh -- BSDF sample / eval / pdf for the three materials of the hot path, specialised per
// material at compile time (the reference dispatches a cuda::std::variant per call, base/bxdf.h).
//   DiffuseBxDF                    bxdfs/diffuse_bxdf.h:8-68
//   ConductorBxDF                  bxdfs/conductor_bxdf.cu:3-106
//   DielectricBxDF                 bxdfs/dielectric_bxdf.cu:4-230
//   TrowbridgeReitzDistribution    util/scattering.h:109-229
//   FrComplex / FrDielectric / refract / Reflect   util/scattering.h:11-106
//   BSDF frame wrapper             base/bsdf.cu:10-53, base/bsdf.h:16-17
// All functions work in the local shading frame (z = shading normal), TransportMode::Radiance.
#pragma once

#include "common.cuh"
#include "vecmath.cuh"

namespace b200pt {

enum : uint32_t {
    BXDF_REFLECTION = 1, BXDF_TRANSMISSION = 2, BXDF_DIFFUSE = 4, BXDF_GLOSSY = 8, BXDF_SPECULAR = 16
};

struct BSDFSample {
    S4 f;
    V3 wi;
    float pdf;
    uint32_t flags;
    float eta;
    bool pdf_is_proportional; // LayeredBxDF samples: pdf is only proportional to the true density (base/bxdf.h BSDFSample)
};
// BxDFReflTransFlags (bxdfs/bxdf_util.h): which lobes a sample / pdf call may use
constexpr uint32_t SF_REFLECTION = 1u, SF_TRANSMISSION = 2u, SF_ALL = 3u;

__device__ __forceinline__ bool same_hemisphere(V3 a, V3 b) { return a.z * b.z > 0.f; }
__device__ __forceinline__ float cos2_theta(V3 w) { return sqr(w.z); }
__device__ __forceinline__ float sin2_theta(V3 w) { return fmaxf(0.f, 1.f - cos2_theta(w)); }
__device__ __forceinline__ float tan2_theta(V3 w) { return sin2_theta(w) / cos2_theta(w); }
__device__ __forceinline__ float cos_phi(V3 w) {
    const float s = sqrtf(sin2_theta(w));
    return s == 0.f ? 1.f : clampf(w.x / s, -1.f, 1.f);
}
__device__ __forceinline__ float sin_phi(V3 w) {
    const float s = sqrtf(sin2_theta(w));
    return s == 0.f ? 0.f : clampf(w.y / s, -1.f, 1.f);
}

__device__ __forceinline__ float2 sample_disk_concentric(float2 u) {
    const float ox = 2.f * u.x - 1.f, oy = 2.f * u.y - 1.f;
    if (ox == 0.f && oy == 0.f) return make_float2(0.f, 0.f);
    float r, theta;
    if (fabsf(ox) > fabsf(oy)) {
        r = ox;
        theta = (kPi / 4.f) * (oy / ox);
    } else {
        r = oy;
        theta = (kPi / 2.f) - (kPi / 4.f) * (ox / oy);
    }
    float s, c;
    sincosf(theta, &s, &c);
    return make_float2(r * c, r * s);
}
__device__ __forceinline__ float2 sample_disk_polar(float2 u) {
    const float r = sqrtf(u.x);
    float s, c;
    sincosf(2.f * kPi * u.y, &s, &c);
    return make_float2(r * c, r * s);
}
__device__ __forceinline__ V3 sample_cosine_hemisphere(float2 u) {
    const float2 d = sample_disk_concentric(u);
    return v3(d.x, d.y, safe_sqrt(1.f - sqr(d.x) - sqr(d.y)));
}

// ---- Trowbridge-Reitz (GGX)
struct TRDist {
    float ax, ay;
    __device__ __forceinline__ bool smooth() const { return fmaxf(ax, ay) < 1e-3f; }
    __device__ __forceinline__ float D(V3 wm) const {
        const float t2 = tan2_theta(wm);
        if (isinf(t2)) return 0.f;
        const float c4 = sqr(cos2_theta(wm));
        if (c4 < 1e-16f) return 0.f;
        const float e = t2 * (sqr(cos_phi(wm) / ax) + sqr(sin_phi(wm) / ay));
        return 1.f / (kPi * ax * ay * c4 * sqr(1.f + e));
    }
    __device__ __forceinline__ float Lambda(V3 w) const {
        const float t2 = tan2_theta(w);
        if (isinf(t2)) return 0.f;
        const float a2 = sqr(cos_phi(w) * ax) + sqr(sin_phi(w) * ay);
        return (sqrtf(1.f + a2 * t2) - 1.f) / 2.f;
    }
    __device__ __forceinline__ float G1(V3 w) const { return 1.f / (1.f + Lambda(w)); }
    __device__ __forceinline__ float G(V3 wo, V3 wi) const { return 1.f / (1.f + Lambda(wo) + Lambda(wi)); }
    __device__ __forceinline__ float Dw(V3 w, V3 wm) const { return G1(w) / fabsf(w.z) * D(wm) * absdot(w, wm); }
    __device__ __forceinline__ float pdf(V3 w, V3 wm) const { return Dw(w, wm); }
    __device__ __forceinline__ V3 sample_wm(V3 w, float2 u) const {
        V3 wh = normalize(v3(ax * w.x, ay * w.y, w.z));
        if (wh.z < 0.f) wh = -wh;
        const V3 T1 = (wh.z < 0.99999f) ? normalize(cross(v3(0.f, 0.f, 1.f), wh)) : v3(1.f, 0.f, 0.f);
        const V3 T2 = cross(wh, T1);
        float2 p = sample_disk_polar(u);
        const float h = sqrtf(1.f - sqr(p.x));
        p.y = lerpf((1.f + wh.z) / 2.f, h, p.y);
        const float pz = sqrtf(fmaxf(0.f, 1.f - (sqr(p.x) + sqr(p.y))));
        const V3 nh = p.x * T1 + p.y * T2 + pz * wh;
        return normalize(v3(ax * nh.x, ay * nh.y, fmaxf(1e-6f, nh.z)));
    }
    __device__ __forceinline__ void regularize() {
        if (ax < 0.3f) ax = clampf(2.f * ax, 0.1f, 0.3f);
        if (ay < 0.3f) ay = clampf(2.f * ay, 0.1f, 0.3f);
    }
};
__device__ __forceinline__ TRDist make_trdist(float ax, float ay) {
    TRDist d{ax, ay};
    if (!d.smooth()) {
        d.ax = fmaxf(d.ax, 1e-4f);
        d.ay = fmaxf(d.ay, 1e-4f);
    }
    return d;
}

__device__ __forceinline__ V3 reflect(V3 wo, V3 n) { return -wo + 2.f * dot(wo, n) * n; }
__device__ __forceinline__ bool refract(V3 wi, V3 n, float eta, float *etap, V3 *wt) {
    float cos_i = dot(n, wi);
    if (cos_i < 0.f) {
        eta = 1.f / eta;
        cos_i = -cos_i;
        n = -n;
    }
    const float sin2_i = fmaxf(0.f, 1.f - sqr(cos_i));
    const float sin2_t = sin2_i / sqr(eta);
    if (sin2_t >= 1.f) return false;
    const float cos_t = sqrtf(1.f - sin2_t);
    *wt = -wi / eta + (cos_i / eta - cos_t) * n;
    if (etap) *etap = eta;
    return true;
}
__device__ __forceinline__ float fr_dielectric(float cos_i, float eta) {
    cos_i = clampf(cos_i, -1.f, 1.f);
    if (cos_i < 0.f) {
        eta = 1.f / eta;
        cos_i = -cos_i;
    }
    const float sin2_i = 1.f - sqr(cos_i);
    const float sin2_t = sin2_i / sqr(eta);
    if (sin2_t >= 1.f) return 1.f;
    const float cos_t = safe_sqrt(1.f - sin2_t);
    const float r_parl = (eta * cos_i - cos_t) / (eta * cos_i + cos_t);
    const float r_perp = (cos_i - eta * cos_t) / (cos_i + eta * cos_t);
    return (sqr(r_parl) + sqr(r_perp)) / 2.f;
}

// ---- complex arithmetic for FrComplex (util/complex.h)
struct Cx {
    float re, im;
};
__device__ __forceinline__ Cx cx(float re, float im = 0.f) { return Cx{re, im}; }
__device__ __forceinline__ Cx operator+(Cx a, Cx b) { return Cx{a.re + b.re, a.im + b.im}; }
__device__ __forceinline__ Cx operator-(Cx a, Cx b) { return Cx{a.re - b.re, a.im - b.im}; }
__device__ __forceinline__ Cx operator*(Cx a, Cx b) { return Cx{a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
__device__ __forceinline__ Cx operator/(Cx a, Cx b) {
    const float scale = 1.f / (b.re * b.re + b.im * b.im);
    return Cx{scale * (a.re * b.re + a.im * b.im), scale * (a.im * b.re - a.re * b.im)};
}
__device__ __forceinline__ float cx_norm(Cx a) { return a.re * a.re + a.im * a.im; }
__device__ __forceinline__ Cx cx_sqrt(Cx z) {
    const float n = sqrtf(cx_norm(z));
    const float t1 = sqrtf(.5f * (n + fabsf(z.re)));
    const float t2 = .5f * z.im / t1;
    if (n == 0.f) return Cx{0.f, 0.f};
    if (z.re >= 0.f) return Cx{t1, t2};
    return Cx{fabsf(t2), copysignf(t1, z.im)};
}
__device__ __forceinline__ float fr_complex(float cos_i, Cx eta) {
    cos_i = clampf(cos_i, 0.f, 1.f);
    const float sin2_i = 1.f - sqr(cos_i);
    const Cx sin2_t = cx(sin2_i) / (eta * eta);
    const Cx cos_t = cx_sqrt(cx(1.f) - sin2_t);
    const Cx r_parl = (eta * cx(cos_i) - cos_t) / (eta * cx(cos_i) + cos_t);
    const Cx r_perp = (cx(cos_i) - eta * cos_t) / (cx(cos_i) + eta * cos_t);
    return (cx_norm(r_parl) + cx_norm(r_perp)) / 2.f;
}
__device__ __forceinline__ S4 fr_complex4(float cos_i, S4 eta, S4 k) {
    return make_float4(fr_complex(cos_i, Cx{eta.x, k.x}), fr_complex(cos_i, Cx{eta.y, k.y}), fr_complex(cos_i, Cx{eta.z, k.z}),
                       fr_complex(cos_i, Cx{eta.w, k.w}));
}

// ---- the BSDF held in registers by a shade kernel.  MAT is the material type of the queue.
template <int MAT>
struct BSDF;

struct DiffuseCore {
    S4 r;
    __device__ __forceinline__ uint32_t flags() const { return is_positive(r) ? (BXDF_DIFFUSE | BXDF_REFLECTION) : 0u; }
    __device__ __forceinline__ void regularize() {}
    __device__ __forceinline__ S4 f(V3 wo, V3 wi, bool radiance = true) const {
        if (!same_hemisphere(wo, wi)) return s4(0.f);
        return r * kInvPi;
    }
    __device__ __forceinline__ float pdf(V3 wo, V3 wi, uint32_t sf = SF_ALL) const {
        if (!(sf & SF_REFLECTION) || !same_hemisphere(wo, wi)) return 0.f;
        return fabsf(wi.z) / kPi;
    }
    __device__ __forceinline__ bool sample(V3 wo, float uc, float2 u, BSDFSample &bs, bool radiance = true, uint32_t sf = SF_ALL) const {
        if (!(sf & SF_REFLECTION)) return false;
        bs.pdf_is_proportional = false;
        V3 wi = sample_cosine_hemisphere(u);
        if (wo.z < 0.f) wi.z *= -1.f;
        bs.f = r * kInvPi;
        bs.wi = wi;
        bs.pdf = fabsf(wi.z) / kPi;
        bs.flags = BXDF_DIFFUSE | BXDF_REFLECTION;
        bs.eta = 1.f;
        return true;
    }
};
template <>
struct BSDF<B200PT_MAT_DIFFUSE> : DiffuseCore {
    Frame frame;
};

// DiffuseTransmissionBxDF (bxdfs/diffuse_transmission_bxdf.h:8-103)
template <>
struct BSDF<B200PT_MAT_DIFFUSE_TRANSMISSION> {
    Frame frame;
    S4 r, t;
    __device__ __forceinline__ uint32_t flags() const {
        return (is_positive(r) ? (BXDF_DIFFUSE | BXDF_REFLECTION) : 0u) | (is_positive(t) ? (BXDF_DIFFUSE | BXDF_TRANSMISSION) : 0u);
    }
    __device__ __forceinline__ void regularize() {}
    __device__ __forceinline__ S4 f(V3 wo, V3 wi) const { return (same_hemisphere(wo, wi) ? r : t) * kInvPi; }
    __device__ __forceinline__ float pdf(V3 wo, V3 wi) const {
        const float pr = max_component(r), pt = max_component(t);
        if (pr == 0.f && pt == 0.f) return 0.f;
        return (same_hemisphere(wo, wi) ? pr : pt) / (pr + pt) * (fabsf(wi.z) / kPi);
    }
    __device__ __forceinline__ bool sample(V3 wo, float uc, float2 u, BSDFSample &bs) const {
        const float pr = max_component(r), pt = max_component(t);
        if (pr == 0.f && pt == 0.f) return false;
        bs.pdf_is_proportional = false;
        V3 wi = sample_cosine_hemisphere(u);
        if (uc < pr / (pr + pt)) {
            if (wo.z < 0.f) wi.z *= -1.f;
            bs.pdf = fabsf(wi.z) / kPi * pr / (pr + pt);
            bs.flags = BXDF_DIFFUSE | BXDF_REFLECTION;
        } else {
            if (wo.z > 0.f) wi.z *= -1.f;
            bs.pdf = fabsf(wi.z) / kPi * pt / (pr + pt);
            bs.flags = BXDF_DIFFUSE | BXDF_TRANSMISSION;
        }
        bs.f = f(wo, wi);
        bs.wi = wi;
        bs.eta = 1.f;
        return true;
    }
};

struct ConductorCore {
    TRDist d;
    S4 eta, k;
    __device__ __forceinline__ uint32_t flags() const { return d.smooth() ? (BXDF_SPECULAR | BXDF_REFLECTION) : (BXDF_GLOSSY | BXDF_REFLECTION); }
    __device__ __forceinline__ void regularize() { d.regularize(); }
    __device__ __forceinline__ S4 f(V3 wo, V3 wi, bool radiance = true) const {
        if (!same_hemisphere(wo, wi) || d.smooth()) return s4(0.f);
        const float co = fabsf(wo.z), ci = fabsf(wi.z);
        if (ci == 0.f || co == 0.f) return s4(0.f);
        V3 wm = wi + wo;
        if (len2(wm) == 0.f) return s4(0.f);
        wm = normalize(wm);
        const S4 F = fr_complex4(absdot(wo, wm), eta, k);
        return F * (d.D(wm) * d.G(wo, wi) / (4.f * ci * co));
    }
    __device__ __forceinline__ float pdf(V3 wo, V3 wi, uint32_t sf = SF_ALL) const {
        if (!(sf & SF_REFLECTION) || !same_hemisphere(wo, wi) || d.smooth()) return 0.f;
        V3 wm = wo + wi;
        if (len2(wm) == 0.f) return 0.f;
        wm = face_forward(normalize(wm), v3(0.f, 0.f, 1.f));
        return d.pdf(wo, wm) / (4.f * absdot(wo, wm));
    }
    __device__ __forceinline__ bool sample(V3 wo, float uc, float2 u, BSDFSample &bs, bool radiance = true, uint32_t sf = SF_ALL) const {
        if (!(sf & SF_REFLECTION)) return false;
        bs.eta = 1.f;
        bs.pdf_is_proportional = false;
        if (d.smooth()) {
            const V3 wi = v3(-wo.x, -wo.y, wo.z);
            bs.f = fr_complex4(fabsf(wi.z), eta, k) / fabsf(wi.z);
            bs.wi = wi;
            bs.pdf = 1.f;
            bs.flags = BXDF_SPECULAR | BXDF_REFLECTION;
            return true;
        }
        if (wo.z == 0.f) return false;
        const V3 wm = d.sample_wm(wo, u);
        const V3 wi = reflect(wo, wm);
        if (!same_hemisphere(wo, wi)) return false;
        bs.pdf = d.pdf(wo, wm) / (4.f * absdot(wo, wm));
        const float co = fabsf(wo.z), ci = fabsf(wi.z);
        if (ci == 0.f || co == 0.f) return false;
        const S4 F = fr_complex4(absdot(wo, wm), eta, k);
        bs.f = F * (d.D(wm) * d.G(wo, wi) / (4.f * ci * co));
        bs.wi = wi;
        bs.flags = BXDF_GLOSSY | BXDF_REFLECTION;
        return true;
    }
};
template <>
struct BSDF<B200PT_MAT_CONDUCTOR> : ConductorCore {
    Frame frame;
};

struct DielectricCore {
    TRDist d;
    float eta;
    __device__ __forceinline__ bool specular() const { return eta == 1.f || d.smooth(); }
    __device__ __forceinline__ uint32_t flags() const {
        const uint32_t fl = (eta == 1.f) ? BXDF_TRANSMISSION : (BXDF_REFLECTION | BXDF_TRANSMISSION);
        return fl | (d.smooth() ? BXDF_SPECULAR : BXDF_GLOSSY);
    }
    __device__ __forceinline__ void regularize() { d.regularize(); }
    // generalised half vector; false when the pair cannot scatter
    __device__ __forceinline__ bool half_vector(V3 wo, V3 wi, V3 &wm, float &etap, bool &refl) const {
        const float co = wo.z, ci = wi.z;
        refl = ci * co > 0.f;
        etap = 1.f;
        if (!refl) etap = co > 0.f ? eta : (1.f / eta);
        wm = wi * etap + wo;
        if (ci == 0.f || co == 0.f || len2(wm) == 0.f) return false;
        wm = normalize(face_forward(wm, v3(0.f, 0.f, 1.f)));
        if (dot(wm, wi) * ci < 0.f || dot(wm, wo) * co < 0.f) return false;
        return true;
    }
    __device__ __forceinline__ S4 f(V3 wo, V3 wi, bool radiance = true) const {
        if (specular()) return s4(0.f);
        V3 wm;
        float etap;
        bool refl;
        if (!half_vector(wo, wi, wm, etap, refl)) return s4(0.f);
        const float F = fr_dielectric(dot(wo, wm), eta);
        if (refl) return s4(d.D(wm) * d.G(wo, wi) * F / fabsf(4.f * wi.z * wo.z));
        const float denom = sqr(dot(wi, wm) + dot(wo, wm) / etap) * wi.z * wo.z;
        float ft = d.D(wm) * (1.f - F) * d.G(wo, wi) * fabsf(dot(wi, wm) * dot(wo, wm) / denom);
        if (radiance) ft /= sqr(etap);
        return s4(ft);
    }
    __device__ __forceinline__ float pdf(V3 wo, V3 wi, uint32_t sf = SF_ALL) const {
        if (specular()) return 0.f;
        V3 wm;
        float etap;
        bool refl;
        if (!half_vector(wo, wi, wm, etap, refl)) return 0.f;
        const float R0 = fr_dielectric(dot(wo, wm), eta), T0 = 1.f - R0;
        const float R = (sf & SF_REFLECTION) ? R0 : 0.f, T = (sf & SF_TRANSMISSION) ? T0 : 0.f;
        if (R == 0.f && T == 0.f) return 0.f;
        if (refl) return d.pdf(wo, wm) / (4.f * absdot(wo, wm)) * R / (R + T);
        const float denom = sqr(dot(wi, wm) + dot(wo, wm) / etap);
        const float dwm_dwi = absdot(wi, wm) / denom;
        return d.pdf(wo, wm) * dwm_dwi * T / (R + T);
    }
    // R0 / T0: Fresnel terms (they weight f); R / T: lobe selection probabilities under the sample flags
    __device__ __forceinline__ bool sample(V3 wo, float uc, float2 u, BSDFSample &bs, bool radiance = true, uint32_t sf = SF_ALL) const {
        bs.eta = 1.f;
        bs.pdf_is_proportional = false;
        if (specular()) {
            const float R0 = fr_dielectric(wo.z, eta), T0 = 1.f - R0;
            const float R = (sf & SF_REFLECTION) ? R0 : 0.f, T = (sf & SF_TRANSMISSION) ? T0 : 0.f;
            if (R == 0.f && T == 0.f) return false;
            if (uc < R / (R + T)) {
                const V3 wi = v3(-wo.x, -wo.y, wo.z);
                bs.f = s4(R0 / fabsf(wi.z));
                bs.wi = wi;
                bs.pdf = R / (R + T);
                bs.flags = BXDF_SPECULAR | BXDF_REFLECTION;
                return true;
            }
            V3 wi;
            float etap;
            if (!refract(wo, v3(0.f, 0.f, 1.f), eta, &etap, &wi)) return false;
            float ft = T0 / fabsf(wi.z);
            if (radiance) ft /= sqr(etap);
            bs.f = s4(ft);
            bs.wi = wi;
            bs.pdf = T / (R + T);
            bs.flags = BXDF_SPECULAR | BXDF_TRANSMISSION;
            bs.eta = etap;
            return true;
        }
        const V3 wm = d.sample_wm(wo, u);
        const float R0 = fr_dielectric(dot(wo, wm), eta), T0 = 1.f - R0;
        const float R = (sf & SF_REFLECTION) ? R0 : 0.f, T = (sf & SF_TRANSMISSION) ? T0 : 0.f;
        if (R == 0.f && T == 0.f) return false;
        if (uc < R / (R + T)) {
            const V3 wi = reflect(wo, wm);
            if (!same_hemisphere(wo, wi)) return false;
            bs.pdf = d.pdf(wo, wm) / (4.f * absdot(wo, wm)) * R / (R + T);
            bs.f = s4(d.D(wm) * d.G(wo, wi) * R0 / (4.f * wi.z * wo.z));
            bs.wi = wi;
            bs.flags = BXDF_GLOSSY | BXDF_REFLECTION;
            return true;
        }
        float etap;
        V3 wi = v3(0.f, 0.f, 0.f);
        const bool tir = !refract(wo, wm, eta, &etap, &wi);
        if (tir || same_hemisphere(wo, wi) || wi.z == 0.f) return false;
        const float denom = sqr(dot(wi, wm) + dot(wo, wm) / etap);
        const float dwm_dwi = absdot(wi, wm) / denom;
        bs.pdf = d.pdf(wo, wm) * dwm_dwi * T / (R + T);
        float ft = T0 * d.D(wm) * d.G(wo, wi) * fabsf(dot(wi, wm) * dot(wo, wm) / (wi.z * wo.z * denom));
        if (radiance) ft /= sqr(etap);
        bs.f = s4(ft);
        bs.wi = wi;
        bs.flags = BXDF_GLOSSY | BXDF_TRANSMISSION;
        bs.eta = etap;
        return true;
    }
};
template <>
struct BSDF<B200PT_MAT_DIELECTRIC> : DielectricCore {
    Frame frame;
};

__device__ __forceinline__ bool flags_non_specular(uint32_t fl) { return (fl & (BXDF_DIFFUSE | BXDF_GLOSSY)) != 0u; }

} // namespace b200pt
