// spectrum.cuh -- device-side spectra at the 4 sampled wavelengths.
//   Spectrum::sample                 base/spectrum.cu:161-201
//   RGBSigmoidPolynomial             spectrum_util/rgb_sigmoid_polynomial.h:15-39
//   RGBAlbedo/Unbounded/Illuminant   spectra/rgb_*_spectrum.*  (illuminant = scale * sigmoid * D65)
//   PiecewiseLinearSpectrum          spectra/piecewise_linear_spectrum.cu:82-103
//   DenselySampledSpectrum           spectra/densely_sampled_spectrum.h (1 nm table, 360..830)
//   sample_visible / pdf             spectrum_util/sampled_wavelengths.h:62-78, util/sampling.h:296-307
// The reference dispatches on {enum, void*}; here a spectrum is a 32-byte record + a slice of one
// float pool, and the RGB->sigmoid coefficients are resolved on the host at scene load.
#pragma once

#include "vecmath.cuh"

namespace b200pt {

struct SpecRec {
    int32_t type;   // B200PT_SPEC_*
    int32_t n;      // knots / samples
    int32_t offset; // into the float pool: PIECEWISE: lambda[n] then value[n]; DENSE: value[n]
    int32_t pad;
    float c[4];     // constant value | sigmoid c0,c1,c2 + scale
};

struct SpectrumTable {
    const SpecRec *recs;
    const float *pool;
    const float *illuminant; // dense 471 samples (colour-space illuminant, normalised), 360..830 nm
};

__device__ __forceinline__ float sigmoid_poly(const float *c, float lambda) {
    // s(x) = .5 + x / (2 sqrt(1 + x^2)), x = c0 l^2 + c1 l + c2 (evaluate_polynomial = Horner with fma)
    const float x = fmaf(lambda, fmaf(lambda, c[0], c[1]), c[2]);
    if (isinf(x)) return x > 0.f ? 1.f : 0.f;
    return 0.5f + x / (2.f * sqrtf(1.f + sqr(x)));
}

__device__ __forceinline__ float dense_lookup(const float *v, float lambda) {
    // DenselySampledSpectrum::operator() (spectra/densely_sampled_spectrum.h:46-55): lerp between the
    // integer-nm neighbours, 0 outside [360, 830].  The 1-nm piecewise-linear CIE tables and the 5-nm
    // D65 table resampled at 1 nm evaluate to the same numbers through this lookup.
    const float fl = floorf(lambda), ce = ceilf(lambda);
    if (fl < 360.f || ce > 830.f) return 0.f;
    return lerpf(lambda - fl, __ldg(v + (int)fl - 360), __ldg(v + (int)ce - 360));
}

__device__ __forceinline__ float piecewise_lookup(const float *lam, const float *val, int n, float lambda) {
    if (n == 0 || lambda < __ldg(lam) || lambda > __ldg(lam + n - 1)) return 0.f;
    // find_interval(n, lam[i] <= lambda)  (util/math.h:129-141)
    int size = n - 2, first = 1;
    while (size > 0) {
        const int half = size >> 1, middle = first + half;
        const bool pred = __ldg(lam + middle) <= lambda;
        first = pred ? middle + 1 : first;
        size = pred ? size - (half + 1) : half;
    }
    int o = first - 1;
    o = o < 0 ? 0 : (o > n - 2 ? n - 2 : o);
    const float l0 = __ldg(lam + o), l1 = __ldg(lam + o + 1);
    const float t = (lambda - l0) / (l1 - l0);
    return lerpf(t, __ldg(val + o), __ldg(val + o + 1));
}

__device__ __forceinline__ float spectrum_at(const SpectrumTable &tab, int id, float lambda) {
    const SpecRec r = tab.recs[id];
    switch (r.type) {
    case B200PT_SPEC_CONSTANT: return r.c[0];
    case B200PT_SPEC_RGB_ALBEDO: return sigmoid_poly(r.c, lambda);
    case B200PT_SPEC_RGB_UNBOUNDED: return r.c[3] * sigmoid_poly(r.c, lambda);
    case B200PT_SPEC_RGB_ILLUMINANT: return r.c[3] * sigmoid_poly(r.c, lambda) * dense_lookup(tab.illuminant, lambda);
    case B200PT_SPEC_PIECEWISE: return piecewise_lookup(tab.pool + r.offset, tab.pool + r.offset + r.n, r.n, lambda);
    case B200PT_SPEC_DENSE: return dense_lookup(tab.pool + r.offset, lambda);
    }
    return 0.f;
}

__device__ __forceinline__ S4 spectrum_sample(const SpectrumTable &tab, int id, S4 lambda) {
    return make_float4(spectrum_at(tab, id, lambda.x), spectrum_at(tab, id, lambda.y), spectrum_at(tab, id, lambda.z),
                       spectrum_at(tab, id, lambda.w));
}

__device__ __forceinline__ bool spectrum_is_constant(const SpectrumTable &tab, int id) { return tab.recs[id].type == B200PT_SPEC_CONSTANT; }

__device__ __forceinline__ float visible_wavelengths_pdf(float lambda) {
    if (lambda < 360.f || lambda > 830.f) return 0.f;
    return 0.0039398042f / sqr(coshf(0.0072f * (lambda - 538.f)));
}
__device__ __forceinline__ float sample_visible_wavelength(float u) { return 538.f - 138.888889f * atanhf(0.85691062f - 1.82750197f * u); }

__device__ __forceinline__ void sample_visible(float u, S4 *lambda, S4 *pdf) {
    float l[4], p[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float up = u + float(i) / 4.f;
        if (up > 1.f) up -= 1.f;
        l[i] = sample_visible_wavelength(up);
        p[i] = visible_wavelengths_pdf(l[i]);
    }
    *lambda = make_float4(l[0], l[1], l[2], l[3]);
    *pdf = make_float4(p[0], p[1], p[2], p[3]);
}

} // namespace b200pt
