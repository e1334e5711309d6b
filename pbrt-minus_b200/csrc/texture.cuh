// texture.cuh -- spectrum / float textures evaluated at a hit (textures/, base/spectrum_texture.cu, base/float_texture.cu).
//   UVMapping::map                          textures/texture_mapping_2d.cu:114-124
//   GPUImage::bilerp / remap_pixel_coord     textures/gpu_image.cu:19-69,276-294  (MIPMap::filter is a plain bilerp, mipmap.cu:22-25)
//   SpectrumImageTexture::evaluate           textures/spectrum_image_texture.cu:17-32
//   FloatImageTexture::evaluate              textures/float_image_texture.h:14-24
//   Spectrum{Scaled,Mix,DirectionMix}Texture textures/spectrum_{scaled,mix,direction_mix}_texture.cu
//   FloatScaledTexture                       textures/float_scaled_texture.cu
//   RGBtoSpectrumTable::operator()           spectrum_util/rgb_to_spectrum_data.cu:388-426
// The reference walks a tree of {enum, void*} nodes recursively; a texture tree is linear in its leaves, so here it is walked
// with a small explicit stack of (node, weight) pairs and the leaves are accumulated.
#pragma once

#include "spectrum.cuh"
#include "vecmath.cuh"

namespace b200pt {

struct TexRec {
    int32_t type, a, b, c;
    float f[4];
    float map[4];
    int32_t mapping;
    float tfr[12];
    float v1[3], v2[3];
};
struct ImageRec {
    int32_t width, height;
    const float *rgb;
};
struct TexTables {
    const TexRec *tex;
    const ImageRec *images;
    const double *coeff;   // [3][64][64][64][3]
    const double *z_nodes; // [64]
};
struct TexCtx {
    float u, v;
    V3 n;
    V3 p; // render-space hit point (planar / spherical / cylindrical mappings)
};

__device__ __forceinline__ int tex_mod(int a, int b) {
    const int r = a - (a / b) * b;
    return r < 0 ? r + b : r;
}

__device__ __forceinline__ V3 image_fetch(const ImageRec &im, int x, int y) {
    if (x < 0 || x >= im.width) x = tex_mod(x, im.width);
    if (y < 0 || y >= im.height) y = tex_mod(y, im.height);
    const float *p = im.rgb + 3 * ((size_t)y * im.width + x);
    return v3(__ldg(p), __ldg(p + 1), __ldg(p + 2));
}

__device__ __forceinline__ V3 image_bilerp(const ImageRec &im, float s, float t) {
    const double x = (double)(s * (float)im.width) - 0.5, y = (double)(t * (float)im.height) - 0.5;
    const double xf = floor(x), yf = floor(y);
    const double dx = x - xf, dy = y - yf;
    const int xi = (int)xf, yi = (int)yf;
    const V3 v0 = image_fetch(im, xi, yi), v1 = image_fetch(im, xi + 1, yi), v2 = image_fetch(im, xi, yi + 1),
             v3_ = image_fetch(im, xi + 1, yi + 1);
    const float w0 = (float)((1.0 - dx) * (1.0 - dy)), w1 = (float)(dx * (1.0 - dy)), w2 = (float)((1.0 - dx) * dy), w3 = (float)(dx * dy);
    return w0 * v0 + w1 * v1 + w2 * v2 + w3 * v3_;
}

// image lookup at the mapped, t-flipped coordinates
__device__ __forceinline__ V3 image_lookup(const TexTables &T, const TexRec &r, const TexCtx &ctx) {
    float s, t;
    if (r.mapping == B200PT_MAP_UV) { // UVMapping::map (textures/texture_mapping_2d.cu:114-124)
        s = r.map[0] * ctx.u + r.map[2];
        t = r.map[1] * ctx.v + r.map[3];
    } else {
        // textureFromRender(ctx.p): Transform::operator()(Point3f)
        const float *m = r.tfr;
        const V3 pt = v3(m[0] * ctx.p.x + m[1] * ctx.p.y + m[2] * ctx.p.z + m[3], m[4] * ctx.p.x + m[5] * ctx.p.y + m[6] * ctx.p.z + m[7],
                         m[8] * ctx.p.x + m[9] * ctx.p.y + m[10] * ctx.p.z + m[11]);
        if (r.mapping == B200PT_MAP_PLANAR) { // PlanarMapping::map (:27-42)
            s = r.map[2] + dot(pt, v3(r.v1[0], r.v1[1], r.v1[2]));
            t = r.map[3] + dot(pt, v3(r.v2[0], r.v2[1], r.v2[2]));
        } else if (r.mapping == B200PT_MAP_SPHERICAL) { // SphericalMapping::map (:44-72): theta / pi, phi / 2 pi of the direction
            const V3 vec = normalize(pt);
            float phi = atan2f(vec.y, vec.x);
            if (phi < 0.f) phi += 2.f * kPi;
            s = acosf(clampf(vec.z, -1.f, 1.f)) * kInvPi;
            t = phi * (0.5f * kInvPi);
        } else { // CylindricalMapping::map (:4-25)
            s = (kPi + atan2f(pt.y, pt.x)) * (0.5f * kInvPi);
            t = pt.z;
        }
    }
    t = (float)(1.0 - (double)t);
    return image_bilerp(T.images[r.a], s, t);
}

__device__ __forceinline__ void rgb_to_sigmoid(const TexTables &T, V3 rgb, float c_out[3]) {
    const float v[3] = {rgb.x, rgb.y, rgb.z};
    if (v[0] == v[1] && v[1] == v[2]) {
        c_out[0] = 0.f;
        c_out[1] = 0.f;
        c_out[2] = (float)((double)(v[0] - 0.5f) / sqrt((double)v[0] * (1.0 - (double)v[0])));
        return;
    }
    const int m = (v[0] > v[1]) ? ((v[0] > v[2]) ? 0 : 2) : ((v[1] > v[2]) ? 1 : 2);
    const double z = v[m];
    const double x = v[(m + 1) % 3] * 63 / z, y = v[(m + 2) % 3] * 63 / z;
    const int xi = min((int)x, 62), yi = min((int)y, 62);
    // find_interval(64, z_nodes[i] < z)  (util/math.h:129-141)
    int size = 62, first = 1;
    while (size > 0) {
        const int half = size >> 1, middle = first + half;
        const bool pred = T.z_nodes[middle] < z;
        first = pred ? middle + 1 : first;
        size = pred ? size - (half + 1) : half;
    }
    const int zi = max(0, min(first - 1, 62));
    const double dx = x - xi, dy = y - yi, dz = (z - T.z_nodes[zi]) / (T.z_nodes[zi + 1] - T.z_nodes[zi]);
#pragma unroll
    for (int i = 0; i < 3; i++) {
        auto co = [&](int ddx, int ddy, int ddz) {
            return T.coeff[3 * ((((size_t)m * 64 + (zi + ddz)) * 64 + (yi + ddy)) * 64 + (xi + ddx)) + i];
        };
        // pbrt::lerp(Real, Real, Real) (util/math.h:74-76): the double offsets and table entries are rounded to fp32 at the call and
        // the interpolation runs in fp32 (spectrum_util/rgb_to_spectrum_data.cu:411-424)
        auto lerp = [](float t, float a, float b) { return (1 - t) * a + t * b; };
        c_out[i] = lerp((float)dz, lerp((float)dy, lerp((float)dx, (float)co(0, 0, 0), (float)co(1, 0, 0)), lerp((float)dx, (float)co(0, 1, 0), (float)co(1, 1, 0))),
                        lerp((float)dy, lerp((float)dx, (float)co(0, 0, 1), (float)co(1, 0, 1)), lerp((float)dx, (float)co(0, 1, 1), (float)co(1, 1, 1))));
    }
}

// float textures: constant, image (red channel), product of two
__device__ __forceinline__ float ftex_eval(const TexTables &T, int id, const TexCtx &ctx) {
    int stack[8];
    int sp = 0;
    stack[sp++] = id;
    float prod = 1.f;
    while (sp > 0) {
        const TexRec r = T.tex[stack[--sp]];
        if (r.type == B200PT_TEX_FLOAT_CONSTANT) prod *= r.f[0];
        else if (r.type == B200PT_TEX_FLOAT_IMAGE) {
            const float v = r.f[0] * image_lookup(T, r, ctx).x;
            prod *= (r.f[1] != 0.f) ? fmaxf(0.f, 1.f - v) : v;
        } else if (sp <= 6) { // FLOAT_SCALE
            stack[sp++] = r.b;
            stack[sp++] = r.a;
        }
    }
    return prod;
}

__device__ __forceinline__ S4 tex_eval(const TexTables &T, const SpectrumTable &spectra, int id, const TexCtx &ctx, S4 lambda) {
    int stack[8];
    float weight[8];
    int sp = 0;
    stack[sp] = id;
    weight[sp++] = 1.f;
    S4 acc = s4(0.f);
    while (sp > 0) {
        sp--;
        const TexRec r = T.tex[stack[sp]];
        const float w = weight[sp];
        switch (r.type) {
        case B200PT_TEX_CONSTANT: acc = acc + spectrum_sample(spectra, r.a, lambda) * w; break;
        case B200PT_TEX_IMAGE: {
            V3 rgb = r.f[0] * image_lookup(T, r, ctx);
            if (r.f[1] != 0.f) rgb = v3(1.f, 1.f, 1.f) - rgb;
            rgb = v3(clampf(fmaxf(rgb.x, 0.f), 0.f, 1.f), clampf(fmaxf(rgb.y, 0.f), 0.f, 1.f), clampf(fmaxf(rgb.z, 0.f), 0.f, 1.f));
            float c[3];
            rgb_to_sigmoid(T, rgb, c);
            acc = acc + make_float4(sigmoid_poly(c, lambda.x), sigmoid_poly(c, lambda.y), sigmoid_poly(c, lambda.z), sigmoid_poly(c, lambda.w)) * w;
            break;
        }
        case B200PT_TEX_SCALE:
            if (r.f[0] != 0.f && sp < 8) {
                stack[sp] = r.a;
                weight[sp++] = w * r.f[0];
            }
            break;
        case B200PT_TEX_MIX: {
            const float amt = r.c >= 0 ? ftex_eval(T, r.c, ctx) : r.f[0];
            if (amt != 1.f && sp < 8) {
                stack[sp] = r.a;
                weight[sp++] = w * (1.f - amt);
            }
            if (amt != 0.f && sp < 8) {
                stack[sp] = r.b;
                weight[sp++] = w * amt;
            }
            break;
        }
        case B200PT_TEX_DIRECTION_MIX: {
            const float amt = absdot(ctx.n, v3(r.f[0], r.f[1], r.f[2]));
            if (amt != 0.f && sp < 8) {
                stack[sp] = r.a;
                weight[sp++] = w * amt;
            }
            if (amt != 1.f && sp < 8) {
                stack[sp] = r.b;
                weight[sp++] = w * (1.f - amt);
            }
            break;
        }
        default: break;
        }
    }
    return acc;
}

} // namespace b200pt
