// layered.cuh -- stochastic layered BxDF (coated diffuse / coated conductor) for the wavefront shade stage.
//   LayeredBxDF<Top, Bottom, twoSided = true>   bxdfs/layered_bxdf.h:9-501
//   TopOrBottomBxDF                             bxdfs/top_or_bottom_bxdf.h
//   CoatedDiffuseBxDF / CoatedConductorBxDF     bxdfs/coated_diffuse_bxdf.h, coated_conductor_bxdf.h
//   HGPhaseFunction, HenyeyGreenstein           medium/media_util.h:12-39, util/scattering.h:47-56
//   sample_henyey_greenstein                    util/sampling.cu:7-38
//   RNG(seq, offset), uniform<float>            util/rng.h:25-126
//   pbrt::hash(Vector3f), hash(uc, u)           util/hash.h:10-80 (MurmurHash64A over 12 bytes, seed 0)
// The random walk between the two interfaces draws from a PCG32 stream seeded by a hash of the BITS of the
// local directions, so an evaluation is a pure function of (wo, wi).  It is the same function as the reference's
// only where the directions carry the same BITS; this pipeline is ulp-close to the reference, not bit-identical
// (0 % identical pixels even where relMSE is 2e-9), so the walks differ and coated surfaces are checked as two
// unbiased estimators of the same BSDF (statistical gate, tests/test_render_gpu.py).  Both coated materials are
// two-sided: the walk always enters through the top (dielectric) interface.
#pragma once

#include <cfloat>

#include "bxdf.cuh"
#include "sampler.cuh"

namespace b200pt {

// MurmurHash64A over exactly 12 bytes (three 32-bit words), seed 0
__device__ __forceinline__ uint64_t hash_u32x3(uint32_t a, uint32_t b, uint32_t c) {
    const uint64_t m = 0xc6a4a7935bd1e995ULL;
    const int r = 47;
    uint64_t h = 0ull ^ (12ull * m);
    uint64_t k = (uint64_t)a | ((uint64_t)b << 32);
    k *= m;
    k ^= k >> r;
    k *= m;
    h ^= k;
    h *= m;
    h ^= (uint64_t)c; // 4-byte tail
    h *= m;
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return h;
}
__device__ __forceinline__ uint64_t hash_v3(V3 v) { return hash_u32x3(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z)); }

struct WalkRng { // util/rng.h: RNG(seqIndex, offset)
    uint64_t state, inc;
    __device__ __forceinline__ uint32_t next_u32() {
        const uint64_t old = state;
        state = old * PCG32_MULT + inc;
        const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    __device__ __forceinline__ WalkRng(uint64_t seq, uint64_t offset) {
        state = 0u;
        inc = (seq << 1u) | 1u;
        next_u32();
        state += offset;
        next_u32();
    }
    // r() of layered_bxdf.h:90: min(uniform<Real>(), OneMinusEpsilon)
    __device__ __forceinline__ float r() { return fminf(__fmul_rn((float)next_u32(), 0x1p-32f), kOneMinusEps); }
    // Point2f(r(), r()) / sample_f(w, r(), {r(), r()}, ...): drawn left to right.  (The reference leaves the order to
    // the compiler; it cannot matter for parity, see the header note: the seeds already differ.)
    __device__ __forceinline__ float2 pair() {
        const float a = r(), b = r();
        return make_float2(a, b);
    }
    __device__ __forceinline__ void uc_pair(float *uc, float2 *u) {
        *uc = r();
        *u = pair();
    }
};

// util/scattering.h:47-56.  1.0 / (4.0 * PI) promotes the whole expression to double (PI is a float constant).
__device__ __forceinline__ float henyey_greenstein(float cos_theta, float g) {
    g = clampf(g, -.99f, .99f);
    const float denom = 1.f + sqr(g) + 2.f * g * cos_theta;
    return (float)(1.0 / (4.0 * (double)kPi) * (double)(1.f - sqr(g)) / (double)(denom * safe_sqrt(denom)));
}
// util/sampling.cu:7-38
__device__ __forceinline__ V3 sample_henyey_greenstein(V3 wo, float g, float2 u, float *pdf) {
    g = clampf(g, -.99f, .99f);
    float cos_theta;
    if (fabsf(g) < 1e-3f) cos_theta = 1.f - 2.f * u.x;
    else cos_theta = -1.f / (2.f * g) * (1.f + sqr(g) - sqr((1.f - sqr(g)) / (1.f + g - 2.f * g * u.x)));
    const float sin_theta = safe_sqrt(1.f - sqr(cos_theta));
    const float phi = 2.f * kPi * u.y;
    V3 x, y;
    coordinate_system(wo, &x, &y); // Frame::from_z(wo)
    float sp, cp;
    sincosf(phi, &sp, &cp);
    const float st = clampf(sin_theta, -1.f, 1.f);
    const V3 l = v3(st * cp, st * sp, clampf(cos_theta, -1.f, 1.f)); // SphericalDirection
    *pdf = henyey_greenstein(cos_theta, g);
    return l.x * x + l.y * y + l.z * wo;
}

__device__ __forceinline__ bool sample_ok(bool have, const BSDFSample &bs) {
    return have && is_positive(bs.f) && bs.pdf != 0.f && bs.wi.z != 0.f;
}

template <class Top, class Bottom>
struct LayeredCore {
    Top top;
    Bottom bottom;
    float thickness, g;
    S4 albedo;
    int32_t max_depth, n_samples;

    __device__ __forceinline__ void regularize() {
        top.regularize();
        bottom.regularize();
    }
    // layered_bxdf.h:29-51
    __device__ __forceinline__ uint32_t flags() const {
        const uint32_t tf = top.flags(), bf = bottom.flags();
        uint32_t fl = BXDF_REFLECTION;
        if (tf & BXDF_SPECULAR) fl |= BXDF_SPECULAR;
        if ((tf & BXDF_DIFFUSE) || (bf & BXDF_DIFFUSE) || is_positive(albedo)) fl |= BXDF_DIFFUSE;
        else if ((tf & BXDF_GLOSSY) || (bf & BXDF_GLOSSY)) fl |= BXDF_GLOSSY;
        if ((tf & BXDF_TRANSMISSION) && (bf & BXDF_TRANSMISSION)) fl |= BXDF_TRANSMISSION;
        return fl;
    }
    static __device__ __forceinline__ float Tr(float dz, V3 w) {
        if (fabsf(dz) <= FLT_MIN) return 1.f;
        return expf(-fabsf(dz / w.z));
    }
    // TopOrBottomBxDF dispatch
    __device__ __forceinline__ S4 i_f(bool is_top, V3 wo, V3 wi, bool rad) const { return is_top ? top.f(wo, wi, rad) : bottom.f(wo, wi, rad); }
    __device__ __forceinline__ float i_pdf(bool is_top, V3 wo, V3 wi, uint32_t sf = SF_ALL) const {
        return is_top ? top.pdf(wo, wi, sf) : bottom.pdf(wo, wi, sf);
    }
    __device__ __forceinline__ bool i_sample(bool is_top, V3 wo, float uc, float2 u, BSDFSample &bs, bool rad, uint32_t sf = SF_ALL) const {
        return is_top ? top.sample(wo, uc, u, bs, rad, sf) : bottom.sample(wo, uc, u, bs, rad, sf);
    }
    __device__ __forceinline__ uint32_t i_flags(bool is_top) const { return is_top ? top.flags() : bottom.flags(); }

    // layered_bxdf.h:53-253 (twoSided: the walk enters through the top interface)
    __device__ S4 f(V3 wo, V3 wi, bool rad = true) const {
        S4 fsum = s4(0.f);
        if (wo.z < 0.f) {
            wo = -wo;
            wi = -wi;
        }
        const bool same = same_hemisphere(wo, wi);
        const bool exit_top = same; // same hemisphere: leave through the top again; else through the bottom
        const float exit_z = exit_top ? thickness : 0.f;
        if (same) fsum = float(n_samples) * top.f(wo, wi, rad);
        WalkRng rng(hash_v3(wo), hash_v3(wi));
        const bool exit_specular = (i_flags(exit_top) & BXDF_SPECULAR) != 0u;
        const bool nonexit_specular = (i_flags(!exit_top) & BXDF_SPECULAR) != 0u;
        const bool scatter = is_positive(albedo);
        for (int s = 0; s < n_samples; ++s) {
            float uc = rng.r();
            float2 u2 = rng.pair();
            float ux, uy;
            BSDFSample wos;
            if (!sample_ok(top.sample(wo, uc, u2, wos, rad, SF_TRANSMISSION), wos)) continue;
            uc = rng.r();
            u2 = rng.pair();
            BSDFSample wis;
            if (!sample_ok(i_sample(exit_top, wi, uc, u2, wis, !rad, SF_TRANSMISSION), wis)) continue;
            S4 beta = wos.f * fabsf(wos.wi.z) / wos.pdf;
            float z = thickness;
            V3 w = wos.wi;
            for (int depth = 0; depth < max_depth; ++depth) {
                if (depth > 3 && max_component(beta) < 0.25f) {
                    const float q = fmaxf(0.f, 1.f - max_component(beta));
                    if (rng.r() < q) break;
                    beta = beta / (1.f - q);
                }
                if (!scatter) {
                    z = (z == thickness) ? 0.f : thickness;
                    beta = beta * Tr(thickness, w);
                } else {
                    const float sigma_t = 1.f;
                    const float dz = -logf(1.f - rng.r()) / (sigma_t / fabsf(w.z)); // sample_exponential
                    const float zp = w.z > 0.f ? (z + dz) : (z - dz);
                    if (z == zp) continue;
                    if (0.f < zp && zp < thickness) {
                        // scattering event inside the slab: NEE through the exit interface along wis, then the phase sample
                        float wt = 1.f;
                        if (!exit_specular) wt = power_heuristic(wis.pdf, henyey_greenstein(dot(-w, -wis.wi), g));
                        fsum = fsum + beta * albedo * henyey_greenstein(dot(-w, -wis.wi), g) * wt * Tr(zp - exit_z, wis.wi) * wis.f / wis.pdf;
                        ux = rng.r();
                        uy = rng.r();
                        float ppdf;
                        const V3 pwi = sample_henyey_greenstein(-w, g, make_float2(ux, uy), &ppdf);
                        if (ppdf == 0.f || pwi.z == 0.f) continue;
                        beta = beta * (albedo * ppdf / ppdf);
                        w = pwi;
                        z = zp;
                        if (((z < exit_z && w.z > 0.f) || (z > exit_z && w.z < 0.f)) && !exit_specular) {
                            const S4 f_exit = i_f(exit_top, -w, wi, rad);
                            if (is_positive(f_exit)) {
                                const float exit_pdf = i_pdf(exit_top, -w, wi, SF_TRANSMISSION);
                                const float wt2 = power_heuristic(ppdf, exit_pdf);
                                fsum = fsum + beta * Tr(zp - exit_z, pwi) * f_exit * wt2;
                            }
                        }
                        continue;
                    }
                    z = clampf(zp, 0.f, thickness);
                }
                if (z == exit_z) {
                    uc = rng.r();
                    u2 = rng.pair();
                    BSDFSample bs;
                    if (!sample_ok(i_sample(exit_top, -w, uc, u2, bs, rad, SF_REFLECTION), bs)) break;
                    beta = beta * (bs.f * fabsf(bs.wi.z) / bs.pdf);
                    w = bs.wi;
                } else {
                    if (!nonexit_specular) {
                        float wt = 1.f;
                        if (!exit_specular) wt = power_heuristic(wis.pdf, i_pdf(!exit_top, -w, -wis.wi));
                        fsum = fsum + beta * i_f(!exit_top, -w, -wis.wi, rad) * fabsf(wis.wi.z) * wt * Tr(thickness, wis.wi) * wis.f / wis.pdf;
                    }
                    uc = rng.r();
                    u2 = rng.pair();
                    BSDFSample bs;
                    if (!sample_ok(i_sample(!exit_top, -w, uc, u2, bs, rad, SF_REFLECTION), bs)) break;
                    beta = beta * (bs.f * fabsf(bs.wi.z) / bs.pdf);
                    w = bs.wi;
                    if (!exit_specular) {
                        const S4 f_exit = i_f(exit_top, -w, wi, rad);
                        if (is_positive(f_exit)) {
                            float wt = 1.f;
                            if (!nonexit_specular) {
                                const float exit_pdf = i_pdf(exit_top, -w, wi, SF_TRANSMISSION);
                                wt = power_heuristic(bs.pdf, exit_pdf);
                            }
                            fsum = fsum + beta * Tr(thickness, bs.wi) * f_exit * wt;
                        }
                    }
                }
            }
        }
        return fsum / float(n_samples);
    }

    // layered_bxdf.h:255-372
    __device__ bool sample(V3 wo, float uc, float2 u, BSDFSample &out, bool rad = true, uint32_t sf = SF_ALL) const {
        bool flip_wi = false;
        if (wo.z < 0.f) {
            wo = -wo;
            flip_wi = true;
        }
        BSDFSample bs;
        if (!sample_ok(top.sample(wo, uc, u, bs, rad), bs)) return false;
        if (bs.flags & BXDF_REFLECTION) {
            if (flip_wi) bs.wi = -bs.wi;
            bs.pdf_is_proportional = true;
            out = bs;
            return true;
        }
        V3 w = bs.wi;
        bool specular_path = (bs.flags & BXDF_SPECULAR) != 0u;
        WalkRng rng(hash_v3(wo), hash_u32x3(__float_as_uint(uc), __float_as_uint(u.x), __float_as_uint(u.y)));
        S4 f = bs.f * fabsf(bs.wi.z);
        float pdf = bs.pdf;
        float z = thickness;
        const bool scatter = is_positive(albedo);
        for (int depth = 0; depth < max_depth; ++depth) {
            const float rr_beta = max_component(f) / pdf;
            if (depth > 3 && rr_beta < 0.25f) {
                const float q = fmaxf(0.f, 1.f - rr_beta);
                if (rng.r() < q) return false;
                pdf *= 1.f - q;
            }
            if (w.z == 0.f) return false;
            if (scatter) {
                const float sigma_t = 1.f;
                const float dz = -logf(1.f - rng.r()) / (sigma_t / fabsf(w.z));
                const float zp = w.z > 0.f ? (z + dz) : (z - dz);
                if (zp == z) return false;
                if (0.f < zp && zp < thickness) {
                    const float2 up = rng.pair();
                    float ppdf;
                    const V3 pwi = sample_henyey_greenstein(-w, g, up, &ppdf);
                    if (ppdf == 0.f || pwi.z == 0.f) return false;
                    f = f * (albedo * ppdf);
                    pdf *= ppdf;
                    specular_path = false;
                    w = pwi;
                    z = zp;
                    continue;
                }
                z = clampf(zp, 0.f, thickness);
            } else {
                z = (z == thickness) ? 0.f : thickness;
                f = f * Tr(thickness, w);
            }
            const bool at_top = !(z == 0.f);
            const float uc2 = rng.r();
            const float2 u2 = rng.pair();
            BSDFSample b2;
            if (!sample_ok(i_sample(at_top, -w, uc2, u2, b2, rad), b2)) return false;
            f = f * b2.f;
            pdf *= b2.pdf;
            specular_path = specular_path && (b2.flags & BXDF_SPECULAR) != 0u;
            w = b2.wi;
            if (b2.flags & BXDF_TRANSMISSION) {
                uint32_t fl = same_hemisphere(wo, w) ? BXDF_REFLECTION : BXDF_TRANSMISSION;
                fl |= specular_path ? BXDF_SPECULAR : BXDF_GLOSSY;
                if (flip_wi) w = -w;
                out.f = f;
                out.wi = w;
                out.pdf = pdf;
                out.flags = fl;
                out.eta = 1.f;
                out.pdf_is_proportional = true;
                return true;
            }
            f = f * fabsf(b2.wi.z);
        }
        return false;
    }

    // layered_bxdf.h:374-481
    __device__ float pdf(V3 wo, V3 wi, uint32_t sf = SF_ALL) const {
        if (wo.z < 0.f) {
            wo = -wo;
            wi = -wi;
        }
        WalkRng rng(hash_v3(wi), hash_v3(wo));
        const bool same = same_hemisphere(wo, wi);
        float pdf_sum = 0.f;
        if (same) pdf_sum += float(n_samples) * top.pdf(wo, wi, SF_REFLECTION);
        const bool top_non_specular = flags_non_specular(top.flags());
        const bool bottom_non_specular = flags_non_specular(bottom.flags());
        for (int s = 0; s < n_samples; ++s) {
            if (same) {
                // TRT term: rInterface = bottom, tInterface = top
                float uc;
                float2 u2;
                BSDFSample wos, wis;
                rng.uc_pair(&uc, &u2);
                const bool have_o = top.sample(wo, uc, u2, wos, true, SF_TRANSMISSION);
                rng.uc_pair(&uc, &u2);
                const bool have_i = top.sample(wi, uc, u2, wis, false, SF_TRANSMISSION);
                if (have_o && is_positive(wos.f) && wos.pdf > 0.f && have_i && is_positive(wis.f) && wis.pdf > 0.f) {
                    if (!top_non_specular) {
                        pdf_sum += bottom.pdf(-wos.wi, -wis.wi);
                    } else {
                        rng.uc_pair(&uc, &u2);
                        BSDFSample rs;
                        const bool have_r = bottom.sample(-wos.wi, uc, u2, rs, true);
                        if (have_r && is_positive(rs.f) && rs.pdf > 0.f) {
                            if (!bottom_non_specular) {
                                pdf_sum += top.pdf(-rs.wi, wi);
                            } else {
                                const float r_pdf = bottom.pdf(-wos.wi, -wis.wi);
                                float wt = power_heuristic(wis.pdf, r_pdf);
                                pdf_sum += wt * r_pdf;
                                const float t_pdf = top.pdf(-rs.wi, wi);
                                wt = power_heuristic(rs.pdf, t_pdf);
                                pdf_sum += wt * t_pdf;
                            }
                        }
                    }
                }
            } else {
                // TT term: toInterface = top, tiInterface = bottom
                float uc = rng.r();
                float2 u2 = rng.pair();
                BSDFSample wos;
                if (!sample_ok(top.sample(wo, uc, u2, wos, true), wos) || (wos.flags & BXDF_REFLECTION)) continue;
                uc = rng.r();
                u2 = rng.pair();
                BSDFSample wis;
                if (!sample_ok(bottom.sample(wi, uc, u2, wis, false), wis) || (wis.flags & BXDF_REFLECTION)) continue;
                if (top.flags() & BXDF_SPECULAR) pdf_sum += bottom.pdf(-wos.wi, wi);
                else if (bottom.flags() & BXDF_SPECULAR) pdf_sum += top.pdf(wo, -wis.wi);
                else pdf_sum += (top.pdf(wo, -wis.wi) + bottom.pdf(-wos.wi, wi)) / 2.f;
            }
        }
        // lerp(0.9, 1 / (4 pi), pdfSum / nSamples)
        return lerpf(0.9f, 1.f / (4.f * kPi), pdf_sum / float(n_samples));
    }
};

} // namespace b200pt
