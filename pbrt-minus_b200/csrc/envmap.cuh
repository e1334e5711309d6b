// envmap.cuh -- LightSource "infinite" with an image (lights/image_infinite_light.{h,cu}).
//   EqualAreaSphereToSquare / SquareToSphere   util/sampling.cu:104-175 (Clarberg's octahedral equal-area mapping)
//   GPUImage::bilerp with WrapMode::OctahedralSphere  textures/gpu_image.cu:19-51,276-294
//   RGBIlluminantSpectrum                      spectra/rgb_illuminant_spectrum.cu:5-40 (scale = 2 max(rgb); sigmoid of rgb / scale; times D65)
//   Distribution2D                             distribution/distribution_2d.cu: an alias table over the image COLUMNS (sums of
//                                              data[x][.]) and one alias table per column; a sample is a whole pixel, its (u, v)
//                                              the pixel's corner (x / W, y / H)
//   le / sample_li / pdf_li / phi              lights/image_infinite_light.cu:77-152
#pragma once

#include "light.cuh"
#include "texture.cuh"

namespace b200pt {

struct EnvMap {
    int32_t light;        // light id of the image infinite light, -1: the scene has none
    int32_t width, height;
    const float *rgb;     // [height][width][3]
    const AliasBin *xb;   // [width]
    const float *xp;      // [width]
    const AliasBin *yb;   // [width][height]
    const float *yp;      // [width][height]
    float rfl[9], lfr[9]; // upper 3x3 of render_from_light and of its inverse
    float scale;
};

__device__ __forceinline__ float2 equal_area_sphere_to_square(V3 d) {
    const float x = fabsf(d.x), y = fabsf(d.y), z = fabsf(d.z);
    const float r = safe_sqrt(1.f - z);
    const float a = fmaxf(x, y);
    float b = fminf(x, y);
    b = a == 0.f ? 0.f : b / a;
    const float t1 = 0.406758566246788489601959989e-5f, t2 = 0.636226545274016134946890922156f, t3 = 0.61572017898280213493197203466e-2f,
                t4 = -0.247333733281268944196501420480f, t5 = 0.881770664775316294736387951347e-1f, t6 = 0.419038818029165735901852432784e-1f,
                t7 = -0.251390972343483509333252996350e-1f;
    float phi = fmaf(b, fmaf(b, fmaf(b, fmaf(b, fmaf(b, fmaf(b, t7, t6), t5), t4), t3), t2), t1);
    if (x < y) phi = 1.f - phi;
    float v = phi * r;
    float u = r - v;
    if (d.z < 0.f) {
        const float t = u;
        u = v;
        v = t;
        u = 1.f - u;
        v = 1.f - v;
    }
    u = copysignf(u, d.x);
    v = copysignf(v, d.y);
    return make_float2(0.5f * (u + 1.f), 0.5f * (v + 1.f));
}

__device__ __forceinline__ V3 equal_area_square_to_sphere(float2 p) {
    const float u = 2.f * p.x - 1.f, v = 2.f * p.y - 1.f;
    const float up = fabsf(u), vp = fabsf(v);
    const float signed_distance = 1.f - (up + vp);
    const float d = fabsf(signed_distance);
    const float r = 1.f - d;
    const float phi = (r == 0.f ? 1.f : (vp - up) / r + 1.f) * kPi / 4.f;
    const float z = copysignf(1.f - sqr(r), signed_distance);
    const float cos_phi = copysignf(cosf(phi), u), sin_phi = copysignf(sinf(phi), v);
    return v3(cos_phi * r * safe_sqrt(2.f - sqr(r)), sin_phi * r * safe_sqrt(2.f - sqr(r)), z);
}

__device__ __forceinline__ V3 env_fetch(const EnvMap &E, int x, int y) {
    const int W = E.width, H = E.height;
    if (x < 0) {
        x = -x;
        y = H - 1 - y;
    } else if (x >= W) {
        x = 2 * W - 1 - x;
        y = H - 1 - y;
    }
    if (y < 0) {
        x = W - 1 - x;
        y = -y;
    } else if (y >= H) {
        x = W - 1 - x;
        y = 2 * H - 1 - y;
    }
    if (W == 1) x = 0;
    if (H == 1) y = 0;
    const float *p = E.rgb + 3 * ((size_t)y * W + x);
    return v3(__ldg(p), __ldg(p + 1), __ldg(p + 2));
}

__device__ __forceinline__ V3 env_bilerp(const EnvMap &E, float2 st) {
    const double x = (double)(st.x * (float)E.width) - 0.5, y = (double)(st.y * (float)E.height) - 0.5;
    const double xf = floor(x), yf = floor(y);
    const double dx = x - xf, dy = y - yf;
    const int xi = (int)xf, yi = (int)yf;
    const V3 a = env_fetch(E, xi, yi), b = env_fetch(E, xi + 1, yi), c = env_fetch(E, xi, yi + 1), d = env_fetch(E, xi + 1, yi + 1);
    const float w0 = (float)((1.0 - dx) * (1.0 - dy)), w1 = (float)(dx * (1.0 - dy)), w2 = (float)((1.0 - dx) * dy), w3 = (float)(dx * dy);
    return w0 * a + w1 * b + w2 * c + w3 * d;
}

// RGBIlluminantSpectrum(rgb, colour space).sample(lambda)
__device__ __forceinline__ S4 rgb_illuminant_sample(const TexTables &T, const SpectrumTable &spectra, V3 rgb, S4 lambda) {
    const float scale = (float)(2.0 * (double)fmaxf(rgb.x, fmaxf(rgb.y, rgb.z)));
    const V3 q = scale > 0.f ? rgb / scale : v3(0.f, 0.f, 0.f);
    float c[3];
    rgb_to_sigmoid(T, q, c);
    const S4 s = make_float4(scale * sigmoid_poly(c, lambda.x), scale * sigmoid_poly(c, lambda.y), scale * sigmoid_poly(c, lambda.z),
                             scale * sigmoid_poly(c, lambda.w));
    const S4 ill = make_float4(dense_lookup(spectra.illuminant, lambda.x), dense_lookup(spectra.illuminant, lambda.y),
                               dense_lookup(spectra.illuminant, lambda.z), dense_lookup(spectra.illuminant, lambda.w));
    return ill * s;
}

__device__ __forceinline__ V3 mul3(const float *m, V3 v) {
    return v3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z);
}

// ImageInfiniteLight::ImageLe
__device__ __forceinline__ S4 env_image_le(const EnvMap &E, const TexTables &T, const SpectrumTable &spectra, float2 uv, S4 lambda) {
    V3 rgb = env_bilerp(E, uv);
    rgb = v3(fmaxf(rgb.x, 0.f), fmaxf(rgb.y, 0.f), fmaxf(rgb.z, 0.f));
    return rgb_illuminant_sample(T, spectra, rgb, lambda) * E.scale;
}
// ImageInfiniteLight::le(ray)
__device__ __forceinline__ S4 env_le(const EnvMap &E, const TexTables &T, const SpectrumTable &spectra, V3 d, S4 lambda) {
    const V3 w = normalize(mul3(E.lfr, d));
    return env_image_le(E, T, spectra, equal_area_sphere_to_square(w), lambda);
}
// ImageInfiniteLight::pdf_li
__device__ __forceinline__ float env_pdf(const EnvMap &E, V3 w_render) {
    const V3 w = mul3(E.lfr, w_render);
    const float2 uv = equal_area_sphere_to_square(w);
    int xi = (int)(uv.x * (float)E.width), yi = (int)(uv.y * (float)E.height);
    xi = min(max(xi, 0), E.width - 1);
    yi = min(max(yi, 0), E.height - 1);
    const float pdf = E.xp[xi] * E.yp[(size_t)xi * E.height + yi] * (float)E.width * (float)E.height;
    return pdf / (4.f * kPi);
}
// ImageInfiniteLight::sample_li: direction, pdf and Le of one sample (the caller places the far point)
__device__ __forceinline__ void env_sample(const EnvMap &E, const TexTables &T, const SpectrumTable &spectra, float2 u, S4 lambda, V3 *wi,
                                           float *pdf, S4 *Le) {
    int xi = alias_sample(E.xb, E.width, u.x);
    xi = min(max(xi, 0), E.width - 1);
    const int yi = alias_sample(E.yb + (size_t)xi * E.height, E.height, u.y);
    const float pixel_pdf = E.xp[xi] * E.yp[(size_t)xi * E.height + yi];
    const float2 uv = make_float2((float)xi / (float)E.width, (float)yi / (float)E.height);
    const float map_pdf = pixel_pdf * (float)E.width * (float)E.height;
    *wi = mul3(E.rfl, equal_area_square_to_sphere(uv));
    *pdf = map_pdf / (4.f * kPi);
    *Le = env_image_le(E, T, spectra, uv, lambda);
}

} // namespace b200pt
