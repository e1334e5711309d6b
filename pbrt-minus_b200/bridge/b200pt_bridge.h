// b200pt_bridge.h -- the reference-side binding of libb200pt.so (include/b200pt.h).
//
// This is the code a pbrt-minus maintainer adds to the reference to switch its wavefront path integrator over to the
// B200-native core: it is #included at the end of (a generated copy of) scene/scene_builder.cu, and SceneBuilder::render
// calls B200ptBridge::render(*this) where the reference calls
//     wavefront_path_integrator->render(film, preview); film->write_to_png(output_filename);
// (scene/scene_builder.cu:1036-1043).  Everything up to that point is the reference's own, unmodified front-end: its lexer,
// its parser, its ParameterDictionary, its shape / material / light / camera / filter / film constructors.  The bridge only
// READS the objects those constructors built (the reference has no accessor API, so the generated copy of scene_builder.cu
// is compiled with the access specifiers opened), lays them out as a b200pt_scene_desc, and calls the C-ABI:
//     b200pt_scene_create -> b200pt_bvh_build           (replaces HLBVH::HLBVH,                    scene_builder.cu:912)
//     b200pt_film_create  -> b200pt_render              (replaces WavefrontPathIntegrator::render,  :1041)
//     b200pt_film_read_srgb8 / b200pt_film_read          (replaces Film::write_to_png's pixel loop,  base/film.cu:57-99)
// What it carries over: triangle meshes (with normals / uvs), spheres and disks, every material type (diffuse, conductor,
// dielectric, diffusetransmission, coateddiffuse, coatedconductor, mix, interface) with constant textures, every reference
// spectrum class, diffuse area lights on triangles and spheres, LightSource distant / uniform infinite / spot, homogeneous
// media with MediumInterface and the camera medium, object instances (triangle objects), the perspective camera, any pixel
// filter (tabulated through Filter::evaluate), the RGB film, image / scale / mix / directionmix / float textures with all four
// mappings, the image infinite light.  Anything else (a texture on a parameter the kernels take constants for) is refused with a message naming it: the Python front-end (pbrt-minus_b200/pbrt_scene.py)
// covers those.
//
// Built by oracle/ref_build/build_ref.py into oracle/_ref/pbrt-minus-b200 (it links the reference's objects, so it lives with
// them); exercised by tests/test_bridge_gpu.py.
#pragma once

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "b200pt.h"

#include <pbrt/base/camera.h>
#include <pbrt/base/film.h>
#include <pbrt/base/filter.h>
#include <pbrt/base/float_texture.h>
#include <pbrt/base/light.h>
#include <pbrt/base/material.h>
#include <pbrt/base/primitive.h>
#include <pbrt/base/shape.h>
#include <pbrt/base/spectrum.h>
#include <pbrt/base/spectrum_texture.h>
#include <pbrt/cameras/perspective.h>
#include <pbrt/integrators/wavefront_path.h>
#include <pbrt/distribution/alias_table.h>
#include <pbrt/distribution/distribution_1d.h>
#include <pbrt/light_samplers/power_light_sampler.h>
#include <pbrt/lights/diffuse_area_light.h>
#include <pbrt/lights/distant_light.h>
#include <pbrt/lights/image_infinite_light.h>
#include <pbrt/lights/spot_light.h>
#include <pbrt/lights/uniform_infinite_light.h>
#include <pbrt/materials/coated_conductor_material.h>
#include <pbrt/materials/coated_diffuse_material.h>
#include <pbrt/materials/conductor_material.h>
#include <pbrt/materials/dielectric_material.h>
#include <pbrt/materials/diffuse_material.h>
#include <pbrt/materials/diffuse_transmission_material.h>
#include <pbrt/materials/mix_material.h>
#include <pbrt/medium/homogeneous_medium.h>
#include <pbrt/medium/medium_interface.h>
#include <pbrt/primitives/geometric_primitive.h>
#include <pbrt/primitives/simple_primitive.h>
#include <pbrt/primitives/transformed_primitive.h>
#include <pbrt/accelerator/hlbvh.h>
#include <pbrt/shapes/disk.h>
#include <pbrt/shapes/sphere.h>
#include <pbrt/shapes/triangle.h>
#include <pbrt/shapes/triangle_mesh.h>
#include <pbrt/spectra/constant_spectrum.h>
#include <pbrt/spectra/densely_sampled_spectrum.h>
#include <pbrt/spectra/piecewise_linear_spectrum.h>
#include <pbrt/spectra/rgb_albedo_spectrum.h>
#include <pbrt/spectra/rgb_illuminant_spectrum.h>
#include <pbrt/spectra/rgb_unbounded_spectrum.h>
#include <pbrt/spectrum_util/global_spectra.h>
#include <pbrt/spectrum_util/rgb_color_space.h>
#include <pbrt/spectrum_util/rgb_to_spectrum_data.h>
#include <pbrt/textures/float_constant_texture.h>
#include <pbrt/textures/float_image_texture.h>
#include <pbrt/textures/float_scaled_texture.h>
#include <pbrt/textures/gpu_image.h>
#include <pbrt/textures/mipmap.h>
#include <pbrt/textures/spectrum_direction_mix_texture.h>
#include <pbrt/textures/spectrum_image_texture.h>
#include <pbrt/textures/spectrum_mix_texture.h>
#include <pbrt/textures/spectrum_scaled_texture.h>
#include <pbrt/textures/texture_mapping_2d.h>
#include <pbrt/textures/spectrum_constant_texture.h>
#include <ext/lodepng/lodepng.h>

// the reference's own Simpson tables of the RGB-to-spectrum fit (spectrum_util/rgb_to_spectrum_data.cu:34-40, 163-212): the
// struct lives in that .cu file, the function has external linkage
namespace RGBtoSpectrumData {
struct RGBtoSpectrumBuffer {
    double lambda_tbl[CIE_FINE_SAMPLES];
    double rgb_tbl[3][CIE_FINE_SAMPLES];
    double rgb_to_xyz[3][3];
    double xyz_to_rgb[3][3];
    double xyz_whitepoint[3];
};
void init_tables(RGBtoSpectrumBuffer *data, Gamut gamut);
} // namespace RGBtoSpectrumData

struct B200ptBridge {
    [[noreturn]] static void refuse(const char *what) {
        fprintf(stderr, "\nb200pt bridge: %s is not carried over by the C++ bridge (use the Python front-end for it)\n", what);
        exit(2);
    }
    static void check(int rc, const char *call) {
        if (rc != B200PT_OK) {
            fprintf(stderr, "\nb200pt bridge: %s failed (%d): %s\n", call, rc, b200pt_last_error());
            exit(1);
        }
    }
    static void matrix16(const Transform &t, bool inverse, float out[16]) {
        const auto &m = inverse ? t.inv_m : t.m;
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) out[4 * r + c] = (float)m.val[r][c];
    }

    // ---- spectra: every reference class has a one-to-one b200pt_spectrum type
    std::vector<b200pt_spectrum> spectra;
    std::vector<std::vector<float>> spectrum_store; // lambda / value arrays the records point into
    std::map<const Spectrum *, int> spectrum_ids;

    std::vector<float> dense(const Spectrum *s) {
        std::vector<float> v(471);
        for (int l = 360; l <= 830; l++) v[l - 360] = (float)(*s)((Real)l);
        return v;
    }
    int spectrum(const Spectrum *s) {
        if (!s) refuse("a null spectrum");
        auto it = spectrum_ids.find(s);
        if (it != spectrum_ids.end()) return it->second;
        b200pt_spectrum o = {};
        switch (s->type) {
        case Spectrum::Type::constant:
            o.type = B200PT_SPEC_CONSTANT;
            o.c[0] = (float)static_cast<const ConstantSpectrum *>(s->ptr)->c;
            break;
        case Spectrum::Type::rgb_albedo: {
            const auto *p = static_cast<const RGBAlbedoSpectrum *>(s->ptr);
            o.type = B200PT_SPEC_RGB_ALBEDO;
            o.c[0] = p->rsp.c0; o.c[1] = p->rsp.c1; o.c[2] = p->rsp.c2;
            break;
        }
        case Spectrum::Type::rgb_unbounded: {
            const auto *p = static_cast<const RGBUnboundedSpectrum *>(s->ptr);
            o.type = B200PT_SPEC_RGB_UNBOUNDED;
            o.c[0] = p->rsp.c0; o.c[1] = p->rsp.c1; o.c[2] = p->rsp.c2; o.c[3] = p->scale;
            break;
        }
        case Spectrum::Type::rgb_illuminant: {
            const auto *p = static_cast<const RGBIlluminantSpectrum *>(s->ptr);
            o.type = B200PT_SPEC_RGB_ILLUMINANT; // times the colour space's illuminant, which the film description carries
            o.c[0] = p->rsp.c0; o.c[1] = p->rsp.c1; o.c[2] = p->rsp.c2; o.c[3] = p->scale;
            break;
        }
        case Spectrum::Type::piecewise_linear: {
            const auto *p = static_cast<const PiecewiseLinearSpectrum *>(s->ptr);
            o.type = B200PT_SPEC_PIECEWISE;
            o.n = p->size;
            spectrum_store.emplace_back(p->lambdas, p->lambdas + p->size);
            spectrum_store.emplace_back(p->values, p->values + p->size);
            break;
        }
        default: // densely sampled, black body: tabulated at the integer wavelengths the reference's own dense tables use
            o.type = B200PT_SPEC_DENSE;
            o.n = 471;
            spectrum_store.push_back(dense(s));
        }
        spectra.push_back(o);
        spectrum_ids[s] = (int)spectra.size() - 1;
        return (int)spectra.size() - 1;
    }
    // pointers into spectrum_store are taken once the vectors no longer grow
    void fix_spectrum_pointers() {
        size_t k = 0;
        for (auto &o : spectra) {
            if (o.type == B200PT_SPEC_PIECEWISE) {
                o.lambda = spectrum_store[k].data();
                o.value = spectrum_store[k + 1].data();
                k += 2;
            } else if (o.type == B200PT_SPEC_DENSE) {
                o.value = spectrum_store[k].data();
                k += 1;
            }
        }
    }

    // ---- media (medium/homogeneous_medium.{h,cu}): sigma_a / sigma_s are DenselySampledSpectrum(spec, scale) already -- handed over
    // as dense tables with scale 1 (the device reads the two integer-wavelength entries and lerps, exactly as from spec * scale)
    std::vector<b200pt_medium> media;
    std::map<const Medium *, int> medium_ids;
    int dense_spectrum(const DenselySampledSpectrum *d) {
        b200pt_spectrum o = {};
        o.type = B200PT_SPEC_DENSE;
        o.n = 471;
        std::vector<float> v(471);
        for (int l = 360; l <= 830; l++) v[l - 360] = (float)(*d)((Real)l);
        spectrum_store.push_back(v);
        spectra.push_back(o);
        return (int)spectra.size() - 1;
    }
    int medium(const Medium *m) {
        if (!m) return -1;
        auto it = medium_ids.find(m);
        if (it != medium_ids.end()) return it->second;
        b200pt_medium o = {};
        o.sigma_a = dense_spectrum(m->sigma_a);
        o.sigma_s = dense_spectrum(m->sigma_s);
        o.scale = 1.f;
        o.g = (float)m->phase.g;
        media.push_back(o);
        medium_ids[m] = (int)media.size() - 1;
        return (int)media.size() - 1;
    }
    // the per-triangle medium word of b200pt_scene_desc.tri_medium
    uint32_t medium_code(bool geometric, const MediumInterface *mi) {
        if (!geometric) return 0u; // SimplePrimitive: the spawned ray keeps the incoming ray's medium
        if (mi && mi->interior != mi->exterior) return uint32_t(medium(mi->interior) + 1) | (uint32_t(medium(mi->exterior) + 1) << 16);
        return 0xffffffffu;        // GeometricPrimitive without a transition: no medium
    }
    int interface_material = -1;

    // ---- textures (base/spectrum_texture.cu, base/float_texture.cu): one b200pt_texture per node, children before parents
    std::vector<b200pt_texture> textures;
    std::map<const void *, int> texture_ids;
    std::vector<b200pt_image> images;
    std::vector<std::vector<float>> image_store;
    std::map<const GPUImage *, int> image_ids;
    int image(const GPUImage *im) {
        auto it = image_ids.find(im);
        if (it != image_ids.end()) return it->second;
        b200pt_image o = {};
        o.width = im->resolution.x;
        o.height = im->resolution.y;
        std::vector<float> px((size_t)o.width * o.height * 3);
        for (size_t k = 0; k < (size_t)o.width * o.height; k++) {
            px[3 * k] = (float)im->pixels[k].r;
            px[3 * k + 1] = (float)im->pixels[k].g;
            px[3 * k + 2] = (float)im->pixels[k].b;
        }
        image_store.push_back(std::move(px));
        images.push_back(o);
        image_ids[im] = (int)images.size() - 1;
        return (int)images.size() - 1;
    }
    static b200pt_texture blank_texture(int type) {
        b200pt_texture t = {};
        t.type = type;
        t.a = t.b = t.c = -1;
        t.map[0] = t.map[1] = 1.f;
        t.texture_from_render[0] = t.texture_from_render[5] = t.texture_from_render[10] = 1.f;
        t.v1[0] = 1.f;
        t.v2[1] = 1.f;
        return t;
    }
    int push_texture(const void *key, const b200pt_texture &t) {
        textures.push_back(t);
        if (key) texture_ids[key] = (int)textures.size() - 1;
        return (int)textures.size() - 1;
    }
    // image node: ImageTextureBase (mapping, MIPMap -> GPUImage, scale, invert)
    void image_node(const ImageTextureBase *b, b200pt_texture &t) {
        if (b->mipmap->wrap_mode != WrapMode::Repeat) refuse("a texture wrap mode other than repeat");
        t.a = image(b->mipmap->image);
        t.f[0] = (float)b->scale;
        t.f[1] = b->invert ? 1.f : 0.f;
        const TextureMapping2D *tm = b->texture_mapping;
        auto rows = [&](const Transform &x) {
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 4; c++) t.texture_from_render[4 * r + c] = (float)x.m.val[r][c];
        };
        switch (tm->type) {
        case TextureMapping2D::Type::uv: {
            const auto *m = static_cast<const UVMapping *>(tm->ptr);
            t.mapping = B200PT_MAP_UV;
            t.map[0] = (float)m->su; t.map[1] = (float)m->sv; t.map[2] = (float)m->du; t.map[3] = (float)m->dv;
            break;
        }
        case TextureMapping2D::Type::planar: {
            const auto *m = static_cast<const PlanarMapping *>(tm->ptr);
            t.mapping = B200PT_MAP_PLANAR;
            rows(m->textureFromRender);
            t.v1[0] = (float)m->vs.x; t.v1[1] = (float)m->vs.y; t.v1[2] = (float)m->vs.z;
            t.v2[0] = (float)m->vt.x; t.v2[1] = (float)m->vt.y; t.v2[2] = (float)m->vt.z;
            t.map[2] = (float)m->ds; t.map[3] = (float)m->dt;
            break;
        }
        case TextureMapping2D::Type::spherical:
            t.mapping = B200PT_MAP_SPHERICAL;
            rows(static_cast<const SphericalMapping *>(tm->ptr)->textureFromRender);
            break;
        default:
            t.mapping = B200PT_MAP_CYLINDRICAL;
            rows(static_cast<const CylindricalMapping *>(tm->ptr)->textureFromRender);
        }
    }
    int float_texture(const FloatTexture *t) {
        auto it = texture_ids.find(t);
        if (it != texture_ids.end()) return it->second;
        switch (t->type) {
        case FloatTexture::Type::constant: {
            b200pt_texture o = blank_texture(B200PT_TEX_FLOAT_CONSTANT);
            o.f[0] = (float)static_cast<const FloatConstantTexture *>(t->ptr)->value;
            return push_texture(t, o);
        }
        case FloatTexture::Type::image: {
            b200pt_texture o = blank_texture(B200PT_TEX_FLOAT_IMAGE);
            image_node((const ImageTextureBase *)static_cast<const FloatImageTexture *>(t->ptr), o); // (a C cast: the base is private)
            return push_texture(t, o);
        }
        default: { // scale: the product of two float textures (textures/float_scaled_texture.cu)
            const auto *p = static_cast<const FloatScaledTexture *>(t->ptr);
            b200pt_texture o = blank_texture(B200PT_TEX_FLOAT_SCALE);
            o.a = float_texture(p->scale);
            o.b = float_texture(p->texture);
            return push_texture(t, o);
        }
        }
    }
    int spectrum_texture(const SpectrumTexture *t) {
        auto it = texture_ids.find(t);
        if (it != texture_ids.end()) return it->second;
        switch (t->type) {
        case SpectrumTexture::Type::constant: {
            b200pt_texture o = blank_texture(B200PT_TEX_CONSTANT);
            o.a = spectrum(static_cast<const SpectrumConstantTexture *>(t->ptr)->value);
            return push_texture(t, o);
        }
        case SpectrumTexture::Type::image: {
            const auto *p = static_cast<const SpectrumImageTexture *>(t->ptr);
            if (p->spectrum_type != SpectrumType::Albedo) refuse("an image texture that is not an albedo (illuminant / unbounded image textures)");
            b200pt_texture o = blank_texture(B200PT_TEX_IMAGE);
            image_node((const ImageTextureBase *)p, o);
            return push_texture(t, o);
        }
        case SpectrumTexture::Type::scaled: {
            const auto *p = static_cast<const SpectrumScaledTexture *>(t->ptr);
            b200pt_texture o = blank_texture(B200PT_TEX_SCALE);
            o.a = spectrum_texture(p->texture);
            o.f[0] = (float)p->scale;
            return push_texture(t, o);
        }
        case SpectrumTexture::Type::mix: {
            const auto *p = static_cast<const SpectrumMixTexture *>(t->ptr);
            b200pt_texture o = blank_texture(B200PT_TEX_MIX);
            o.a = spectrum_texture(p->tex1);
            o.b = spectrum_texture(p->tex2);
            o.c = float_texture(p->amount);
            return push_texture(t, o);
        }
        case SpectrumTexture::Type::direction_mix: {
            const auto *p = static_cast<const SpectrumDirectionMixTexture *>(t->ptr);
            b200pt_texture o = blank_texture(B200PT_TEX_DIRECTION_MIX);
            o.a = spectrum_texture(p->tex1);
            o.b = spectrum_texture(p->tex2);
            o.f[0] = (float)p->dir.x; o.f[1] = (float)p->dir.y; o.f[2] = (float)p->dir.z;
            return push_texture(t, o);
        }
        default:
            refuse("a checkerboard texture (the reference itself renders NaN with it)");
        }
    }
    int constant_texture(const SpectrumTexture *t) {
        if (!t) return -1;
        if (t->type != SpectrumTexture::Type::constant) refuse("a non-constant spectrum texture on a parameter that takes none here");
        return spectrum(static_cast<const SpectrumConstantTexture *>(t->ptr)->value);
    }
    // a "reflectance" / "transmittance" parameter: the spectrum id of a constant texture, or a placeholder spectrum plus the
    // texture id + 1 (b200pt_material.*_tex1)
    int textured(const SpectrumTexture *t, int32_t *tex1) {
        *tex1 = 0;
        if (!t) return -1;
        if (t->type == SpectrumTexture::Type::constant) return spectrum(static_cast<const SpectrumConstantTexture *>(t->ptr)->value);
        *tex1 = spectrum_texture(t) + 1;
        if (placeholder_spectrum < 0) {
            b200pt_spectrum o = {};
            o.type = B200PT_SPEC_CONSTANT;
            o.c[0] = 0.5f;
            spectra.push_back(o);
            placeholder_spectrum = (int)spectra.size() - 1;
        }
        return placeholder_spectrum;
    }
    int placeholder_spectrum = -1;
    // a roughness parameter: the value of a constant float texture, or 0 plus the texture id + 1
    float roughness(const FloatTexture *t, int32_t *tex1) {
        *tex1 = 0;
        if (!t) return 0.f;
        if (t->type == FloatTexture::Type::constant) return (float)static_cast<const FloatConstantTexture *>(t->ptr)->value;
        *tex1 = float_texture(t) + 1;
        return 0.f;
    }
    float constant_float(const FloatTexture *t) {
        if (!t) return 0.f;
        if (t->type != FloatTexture::Type::constant) refuse("a non-constant float texture");
        return (float)static_cast<const FloatConstantTexture *>(t->ptr)->value;
    }

    // ---- materials (base/material.cu:97-131)
    std::vector<b200pt_material> materials;
    std::map<const Material *, int> material_ids;
    int material(const Material *m) {
        if (!m) { // Material "interface": no material at all, the surface only separates media (base/material.cu:34-36)
            if (interface_material < 0) {
                b200pt_material o = {};
                o.type = B200PT_MAT_INTERFACE;
                o.reflectance = o.eta = o.k = o.transmittance = o.albedo = o.c_eta = o.c_k = -1;
                o.remap_roughness = 1;
                materials.push_back(o);
                interface_material = (int)materials.size() - 1;
            }
            return interface_material;
        }
        auto it = material_ids.find(m);
        if (it != material_ids.end()) return it->second;
        b200pt_material o = {};
        o.reflectance = o.eta = o.k = o.transmittance = o.albedo = o.c_eta = o.c_k = -1;
        o.remap_roughness = 1;
        switch (m->type) {
        case Material::Type::diffuse:
            o.type = B200PT_MAT_DIFFUSE;
            o.reflectance = textured(static_cast<const DiffuseMaterial *>(m->ptr)->reflectance, &o.reflectance_tex1);
            break;
        case Material::Type::conductor: {
            const auto *p = static_cast<const ConductorMaterial *>(m->ptr);
            o.type = B200PT_MAT_CONDUCTOR;
            if (p->reflectance) o.reflectance = textured(p->reflectance, &o.reflectance_tex1);
            else {
                o.eta = textured(p->eta, &o.eta_tex1);   // Unbounded-kind graphs: constants under scale / mix / directionmix
                o.k = textured(p->k, &o.k_tex1);
            }
            o.u_roughness = roughness(p->uRoughness, &o.u_roughness_tex1);
            o.v_roughness = roughness(p->vRoughness, &o.v_roughness_tex1);
            o.remap_roughness = p->remapRoughness ? 1 : 0;
            break;
        }
        case Material::Type::dielectric: {
            const auto *p = static_cast<const DielectricMaterial *>(m->ptr);
            o.type = B200PT_MAT_DIELECTRIC;
            o.eta = spectrum(p->eta);
            o.u_roughness = roughness(p->uRoughness, &o.u_roughness_tex1);
            o.v_roughness = roughness(p->vRoughness, &o.v_roughness_tex1);
            o.remap_roughness = p->remapRoughness ? 1 : 0;
            break;
        }
        case Material::Type::diffuse_transmission: {
            const auto *p = static_cast<const DiffuseTransmissionMaterial *>(m->ptr);
            o.type = B200PT_MAT_DIFFUSE_TRANSMISSION;
            o.reflectance = textured(p->reflectance, &o.reflectance_tex1);
            o.transmittance = textured(p->transmittance, &o.transmittance_tex1);
            o.scale = (float)p->scale;
            break;
        }
        case Material::Type::coated_diffuse: { // materials/coated_diffuse_material.cu:9-51
            const auto *p = static_cast<const CoatedDiffuseMaterial *>(m->ptr);
            o.type = B200PT_MAT_COATED_DIFFUSE;
            o.reflectance = textured(p->reflectance, &o.reflectance_tex1);
            o.albedo = constant_texture(p->albedo);
            o.eta = spectrum(p->eta);
            o.u_roughness = constant_float(p->u_roughness);
            o.v_roughness = constant_float(p->v_roughness);
            o.thickness = constant_float(p->thickness);
            o.g = constant_float(p->g);
            o.remap_roughness = p->remapRoughness ? 1 : 0;
            o.max_depth = p->maxDepth;
            o.n_samples = p->nSamples;
            break;
        }
        case Material::Type::coated_conductor: { // materials/coated_conductor_material.cu:26-90
            const auto *p = static_cast<const CoatedConductorMaterial *>(m->ptr);
            o.type = B200PT_MAT_COATED_CONDUCTOR;
            o.eta = spectrum(p->interfaceEta);
            o.u_roughness = constant_float(p->interfaceURoughness);
            o.v_roughness = constant_float(p->interfaceVRoughness);
            o.thickness = constant_float(p->thickness);
            o.g = constant_float(p->g);
            o.albedo = constant_texture(p->albedo);
            o.c_u_roughness = constant_float(p->conductorURoughness);
            o.c_v_roughness = constant_float(p->conductorVRoughness);
            if (p->reflectance) o.reflectance = textured(p->reflectance, &o.reflectance_tex1);
            else {
                o.c_eta = textured(p->conductorEta, &o.eta_tex1);
                o.c_k = textured(p->k, &o.k_tex1);
            }
            o.remap_roughness = p->remapRoughness ? 1 : 0;
            o.max_depth = p->maxDepth;
            o.n_samples = p->nSamples;
            break;
        }
        case Material::Type::mix: { // materials/mix_material.cu: resolved per hit by a hash of the hit point
            const auto *p = static_cast<const MixMaterial *>(m->ptr);
            o.type = B200PT_MAT_MIX;
            o.scale = (float)p->amount;
            // the slot is taken before the two inputs are converted, so that material ids stay in creation order
            materials.push_back(o);
            const int self = (int)materials.size() - 1;
            material_ids[m] = self;
            const int a = material(p->materials[0]), bb = material(p->materials[1]);
            materials[self].mix_a = a;
            materials[self].mix_b = bb;
            return self;
        }
        default: refuse("a material type the bridge does not know");
        }
        materials.push_back(o);
        material_ids[m] = (int)materials.size() - 1;
        return (int)materials.size() - 1;
    }

    // ---- geometry
    std::vector<float> P, N, UV;
    std::vector<int32_t> indices, tri_material, tri_light;
    std::vector<uint8_t> tri_flip;
    std::vector<uint32_t> tri_medium;
    std::vector<b200pt_sphere> spheres;
    std::map<const Shape *, int> triangle_of_shape, sphere_of_shape;
    bool any_n = false, any_uv = false;

    void add_shape(const Shape *shape, const Material *mat, uint32_t mcode) {
        const int mid = material(mat);
        if (shape->type == Shape::Type::triangle) {
            const auto *tri = static_cast<const Triangle *>(shape->ptr);
            const TriangleMesh *mesh = tri->mesh;
            const int *v = mesh->vertex_indices + 3 * tri->triangle_idx;
            triangle_of_shape[shape] = (int)tri_material.size();
            for (int k = 0; k < 3; k++) { // one private vertex triple per triangle: positions are already in render space
                indices.push_back((int32_t)(P.size() / 3));
                const Point3f p = mesh->p[v[k]];
                P.push_back(p.x); P.push_back(p.y); P.push_back(p.z);
                if (mesh->n) { N.push_back(mesh->n[v[k]].x); N.push_back(mesh->n[v[k]].y); N.push_back(mesh->n[v[k]].z); }
                else { N.push_back(0); N.push_back(0); N.push_back(0); }
                if (mesh->uv) { UV.push_back(mesh->uv[v[k]].x); UV.push_back(mesh->uv[v[k]].y); }
                else { UV.push_back(0); UV.push_back(0); }
            }
            any_n |= mesh->n != nullptr;
            any_uv |= mesh->uv != nullptr;
            tri_material.push_back(mid);
            tri_medium.push_back(mcode);
            tri_light.push_back(-1);
            tri_flip.push_back((uint8_t)(((mesh->reverse_orientation ^ mesh->transform_swaps_handedness) ? 1 : 0) | (mesh->n ? 2 : 0) |
                                         (mesh->uv ? 4 : 0)));
            return;
        }
        b200pt_sphere s = {};
        s.material = mid;
        s.medium = mcode;
        s.light = -1;
        if (shape->type == Shape::Type::sphere) {
            const auto *q = static_cast<const Sphere *>(shape->ptr);
            matrix16(q->render_from_object, false, s.render_from_object);
            matrix16(q->object_from_render, false, s.object_from_render);
            s.radius = (float)q->radius; s.z_min = (float)q->z_min; s.z_max = (float)q->z_max; s.phi_max = (float)q->phi_max;
            s.reverse_orientation = q->reverse_orientation;
            s.transform_swaps_handedness = q->transform_swaps_handedness;
            s.kind = 0;
        } else {
            const auto *q = static_cast<const Disk *>(shape->ptr);
            matrix16(q->render_from_object, false, s.render_from_object);
            matrix16(q->object_from_render, false, s.object_from_render);
            s.radius = (float)q->radius; s.z_min = (float)q->height; s.z_max = (float)q->inner_radius; s.phi_max = (float)q->phi_max;
            s.reverse_orientation = q->reverse_orientation;
            s.transform_swaps_handedness = q->transform_swaps_handedness;
            s.kind = 1;
        }
        sphere_of_shape[shape] = (int)spheres.size();
        spheres.push_back(s);
    }

    static void render(const SceneBuilder &sb);
};

// ---- SceneBuilder::render's wavefront branch, through the C-ABI -----------------------------------------------------------
inline void B200ptBridge::render(const SceneBuilder &sb) {
    B200ptBridge b;
    // lights first: the light-sampler index of a light is its position in gpu_lights (scene_builder.cu:914-935)
    std::map<const Light *, int> light_ids;
    for (size_t i = 0; i < sb.gpu_lights.size(); i++) light_ids[sb.gpu_lights[i]] = (int)i;
    std::vector<b200pt_light> lights(sb.gpu_lights.size());

    std::vector<b200pt_instance> instances;
    std::vector<const HLBVH *> object_bvhs;
    std::map<const HLBVH *, int> object_ids;
    // primitives, in the order the parser created them (= the reference's primitive indices)
    for (const Primitive *prim : sb.gpu_primitives) {
        const Shape *shape = nullptr;
        const Material *mat = nullptr;
        const Light *area = nullptr;
        uint32_t mcode = 0u;
        if (prim->type == Primitive::Type::simple) {
            const auto *sp = static_cast<const SimplePrimitive *>(prim->ptr);
            shape = sp->shape;
            mat = sp->material;
        } else if (prim->type == Primitive::Type::geometric) {
            const auto *gp = static_cast<const GeometricPrimitive *>(prim->ptr);
            shape = gp->shape_ptr;
            mat = gp->material;
            area = gp->area_light;
            mcode = b.medium_code(true, gp->medium_interface);
        } else {
            // ObjectInstance (scene_builder.cu:845-876): the object's HLBVH itself (identity transform) or a TransformedPrimitive
            // around it.  Its triangles go in AFTER all top-level ones (below); here only the instance is noted.
            b200pt_instance in = {};
            const Primitive *target = prim;
            for (int k = 0; k < 4; k++) in.render_from_instance[5 * k] = in.instance_from_render[5 * k] = 1.f;
            if (prim->type == Primitive::Type::transformed) {
                const auto *tp = static_cast<const TransformedPrimitive *>(prim->ptr);
                matrix16(tp->render_from_primitive, false, in.render_from_instance);
                matrix16(tp->render_from_primitive, true, in.instance_from_render);
                target = tp->primitive;
            }
            if (target->type != Primitive::Type::bvh) refuse("an instance of something that is not an object's BVH");
            const auto *hl = static_cast<const HLBVH *>(target->ptr);
            auto it = object_ids.find(hl);
            if (it == object_ids.end()) {
                object_ids[hl] = (int)object_bvhs.size();
                in.object = (int)object_bvhs.size();
                object_bvhs.push_back(hl);
            } else {
                in.object = it->second;
            }
            instances.push_back(in);
            continue;
        }
        b.add_shape(shape, mat, mcode);
        if (area) {
            const int li = light_ids.at(area);
            if (area->type != Light::Type::diffuse_area_light) refuse("an area light that is not a DiffuseAreaLight");
            const auto *dal = static_cast<const DiffuseAreaLight *>(area->ptr);
            b200pt_light L = {};
            L.spectrum = b.spectrum(dal->Lemit);
            L.scale = (float)dal->scale; // already divided by the photometric integral (lights/diffuse_area_light.cu:17-26)
            L.two_sided = dal->two_sided ? 1 : 0;
            if (shape->type == Shape::Type::triangle) {
                L.type = B200PT_LIGHT_AREA;
                L.triangle = b.triangle_of_shape.at(shape);
                b.tri_light[L.triangle] = li;
            } else if (shape->type == Shape::Type::sphere) {
                L.type = B200PT_LIGHT_SPHERE_AREA;
                L.triangle = b.sphere_of_shape.at(shape);
                b.spheres[L.triangle].light = li;
            } else {
                refuse("an emissive disk");
            }
            lights[li] = L;
        }
    }
    // the instanced objects' triangles, each object once, after the top-level triangles (b200pt_object)
    std::vector<b200pt_object> objects;
    std::map<int, std::vector<std::pair<const Shape *, const Material *>>> object_quadrics;
    for (const HLBVH *hl : object_bvhs) {
        b200pt_object ob = {};
        ob.first_triangle = (int64_t)b.tri_material.size();
        for (int k = 0; k < hl->num_primitives; k++) {
            const Primitive *q = hl->primitives[k];
            const Shape *shape = nullptr;
            const Material *mat = nullptr;
            if (q->type == Primitive::Type::simple) {
                const auto *sp = static_cast<const SimplePrimitive *>(q->ptr);
                shape = sp->shape;
                mat = sp->material;
            } else if (q->type == Primitive::Type::geometric) { // not reachable: emissive / medium-transition shapes are refused inside objects
                refuse("an emissive or medium-transition shape inside an instanced object");
            } else {
                refuse("a nested object instance");
            }
            if (shape->type != Shape::Type::triangle) { // quadrics do not go into the object's tree: one copy per instance, below
                object_quadrics[(int)objects.size()].push_back({shape, mat});
                continue;
            }
            b.add_shape(shape, mat, 0u);
        }
        ob.n_triangles = (int64_t)b.tri_material.size() - ob.first_triangle;
        const Bounds3f hb = hl->bounds(); // what TransformedPrimitive::bounds() transforms (primitives/transformed_primitive.cu:10-12)
        ob.bounds[0] = (float)hb.p_min.x; ob.bounds[1] = (float)hb.p_min.y; ob.bounds[2] = (float)hb.p_min.z;
        ob.bounds[3] = (float)hb.p_max.x; ob.bounds[4] = (float)hb.p_max.y; ob.bounds[5] = (float)hb.p_max.z;
        objects.push_back(ob);
    }
    // spheres and disks of instanced objects ride in the flat quadric list under the composite transform
    // (render_from_instance * render_from_object); the orientation flag stays that of the quadric's own transform, because
    // Transform::operator()(SurfaceInteraction) carries the normal through the instance transform without a handedness flip
    for (const b200pt_instance &in : instances) {
        auto it = object_quadrics.find(in.object);
        if (it == object_quadrics.end()) continue;
        for (const auto &sm : it->second) {
            b.add_shape(sm.first, sm.second, 0u);
            b200pt_sphere &q = b.spheres.back();
            float rfo[16], ofr[16];
            for (int r = 0; r < 4; r++)
                for (int c = 0; c < 4; c++) {
                    double a = 0, bb = 0;
                    for (int k = 0; k < 4; k++) {
                        a += (double)in.render_from_instance[4 * r + k] * (double)q.render_from_object[4 * k + c];
                        bb += (double)q.object_from_render[4 * r + k] * (double)in.instance_from_render[4 * k + c];
                    }
                    rfo[4 * r + c] = (float)a;
                    ofr[4 * r + c] = (float)bb;
                }
            memcpy(q.render_from_object, rfo, sizeof(rfo));
            memcpy(q.object_from_render, ofr, sizeof(ofr));
        }
    }

    // LightSource lights (base/light.cu:22-75): the constructors have folded from / to / scale / power / illuminance into
    // render_from_light and `scale` already
    for (size_t i = 0; i < sb.gpu_lights.size(); i++) {
        const Light *lt = sb.gpu_lights[i];
        if (lt->type == Light::Type::diffuse_area_light) continue;
        b200pt_light L = {};
        L.triangle = -1;
        L.image = -1;
        if (lt->type == Light::Type::distant_light) { // lights/distant_light.cu:10-41: the light shines along -render_from_light(0,0,1)
            const auto *d = static_cast<const DistantLight *>(lt->ptr);
            L.type = B200PT_LIGHT_DISTANT;
            L.spectrum = b.spectrum(d->Lemit);
            L.scale = (float)d->scale;
            for (int r = 0; r < 3; r++) L.v[r] = (float)d->render_from_light.m.val[r][2];
        } else if (lt->type == Light::Type::uniform_infinite_light) { // lights/uniform_infinite_light.cu:7-32
            const auto *u = static_cast<const UniformInfiniteLight *>(lt->ptr);
            L.type = B200PT_LIGHT_UNIFORM_INFINITE;
            L.spectrum = b.spectrum(u->Lemit);
            L.scale = (float)u->scale;
        } else if (lt->type == Light::Type::spot_light) { // lights/spot_light.cu:8-41
            const auto *sp = static_cast<const SpotLight *>(lt->ptr);
            L.type = B200PT_LIGHT_SPOT;
            L.spectrum = b.spectrum(sp->i_emit);
            L.scale = (float)sp->scale;
            for (int r = 0; r < 3; r++) {
                L.v[r] = (float)sp->render_from_light.m.val[r][3];
                for (int c = 0; c < 3; c++) L.light_from_render[3 * r + c] = (float)sp->render_from_light.inv_m.val[r][c];
            }
            L.cos_falloff_start = (float)sp->cosFalloffStart;
            L.cos_falloff_end = (float)sp->cosFalloffEnd;
        } else { // lights/image_infinite_light.cu: the image in the equal-area octahedral parameterisation
            const auto *im = static_cast<const ImageInfiniteLight *>(lt->ptr);
            for (size_t k = 0; k < i; k++)
                if (lights[k].type == B200PT_LIGHT_IMAGE_INFINITE) refuse("a second image infinite light");
            L.type = B200PT_LIGHT_IMAGE_INFINITE;
            L.spectrum = -1;
            L.scale = (float)im->scale;
            L.image = b.image(im->image_ptr);
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++) {
                    L.light_from_render[3 * r + c] = (float)im->render_from_light.inv_m.val[r][c];
                    L.render_from_light[3 * r + c] = (float)im->render_from_light.m.val[r][c];
                }
        }
        lights[i] = L;
    }

    // camera (cameras/perspective.cu:7-63), pixel filter tabulated as FilterSampler's constructor does (filters/filter_sampler.cu:4-20)
    b200pt_camera cam = {};
    cam.medium = -1;
    cuda::std::visit([&](auto &x) { cam.medium = b.medium(x.camera_base.medium); }, *sb.integrator_base->camera);
    cuda::std::visit(
        [&](auto &x) {
            matrix16(x.camera_from_raster, false, cam.camera_from_raster);
            matrix16(x.camera_base.camera_transform.render_from_camera, false, cam.render_from_camera);
            cam.lens_radius = (float)x.lens_radius;
            cam.focal_distance = (float)x.focal_distance;
        },
        *sb.integrator_base->camera);
    const Filter &filter = *sb.integrator_base->filter;
    const auto radius = filter.get_radius();
    const Bounds2f domain(Point2f(-radius), Point2f(radius));
    const int nx = static_cast<int>(32 * radius.x), ny = static_cast<int>(32 * radius.y);
    std::vector<float> ftab((size_t)nx * ny);
    for (int y = 0; y < ny; ++y)
        for (int x = 0; x < nx; ++x) ftab[(size_t)y * nx + x] = (float)filter.evaluate(domain.lerp(Point2f((x + 0.5f) / nx, (y + 0.5f) / ny)));
    b200pt_filter flt = {(float)radius.x, (float)radius.y, nx, ny, ftab.data()};

    // film / sensor (film/rgb_film.cu:27-47): CIE XYZ sensor, imaging ratio 1, output = the colour space's rgb_from_xyz
    const auto res = sb.film->get_resolution();
    const GlobalSpectra *gs = sb.global_spectra;
    const std::vector<float> cie_x = b.dense(gs->cie_xyz[0]), cie_y = b.dense(gs->cie_xyz[1]), cie_z = b.dense(gs->cie_xyz[2]);
    const std::vector<float> illum = b.dense(gs->rgb_color_space->illuminant);
    b200pt_film_desc fd = {};
    fd.width = res.x;
    fd.height = res.y;
    fd.cie_x = cie_x.data();
    fd.cie_y = cie_y.data();
    fd.cie_z = cie_z.data();
    fd.illuminant = illum.data();
    fd.imaging_ratio = 1.f;
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) fd.output_rgb_from_sensor_rgb[3 * r + c] = (float)gs->rgb_color_space->rgb_from_xyz.val[r][c];

    b.fix_spectrum_pointers();
    b200pt_scene_desc d = {};
    d.n_vertices = (int64_t)(b.P.size() / 3);
    d.P = b.P.data();
    d.N = b.any_n ? b.N.data() : nullptr;
    d.uv = b.any_uv ? b.UV.data() : nullptr;
    d.n_triangles = (int64_t)b.tri_material.size();
    d.indices = b.indices.data();
    d.tri_material = b.tri_material.data();
    d.tri_light = b.tri_light.data();
    d.tri_flip = b.tri_flip.data();
    d.n_spheres = (int32_t)b.spheres.size();
    d.spheres = b.spheres.data();
    d.n_spectra = (int32_t)b.spectra.size();
    d.spectra = b.spectra.data();
    std::vector<double> fit;
    if (!b.images.empty()) {
        for (size_t i = 0; i < b.images.size(); i++) b.images[i].rgb = b.image_store[i].data();
        // [3 x 283 rgb_tbl | 3 white | 283 lambda01] from the reference's own init_tables
        auto *buf = new RGBtoSpectrumData::RGBtoSpectrumBuffer();
        RGBtoSpectrumData::init_tables(buf, RGBtoSpectrumData::Gamut::sRGB);
        const int F = RGBtoSpectrumData::CIE_FINE_SAMPLES;
        for (int k = 0; k < 3; k++)
            for (int i = 0; i < F; i++) fit.push_back(buf->rgb_tbl[k][i]);
        for (int k = 0; k < 3; k++) fit.push_back(buf->xyz_whitepoint[k]);
        for (int i = 0; i < F; i++)
            fit.push_back((buf->lambda_tbl[i] - RGBtoSpectrumData::CIE_LAMBDA_MIN) / (RGBtoSpectrumData::CIE_LAMBDA_MAX - RGBtoSpectrumData::CIE_LAMBDA_MIN));
        delete buf;
        d.rgb2spec_fit = fit.data();
    }
    d.n_textures = (int32_t)b.textures.size();
    d.textures = b.textures.data();
    d.n_images = (int32_t)b.images.size();
    d.images = b.images.data();
    d.n_objects = (int32_t)objects.size();
    d.objects = objects.data();
    d.n_instances = (int32_t)instances.size();
    d.instances = instances.data();
    d.n_media = (int32_t)b.media.size();
    d.media = b.media.data();
    d.tri_medium = b.media.empty() ? nullptr : b.tri_medium.data();
    d.n_materials = (int32_t)b.materials.size();
    d.materials = b.materials.data();
    if (getenv("B200PT_BRIDGE_DUMP")) { // what the bridge read out of the reference's objects, for comparisons with the Python front-end
        for (size_t i = 0; i < b.materials.size(); i++) {
            const b200pt_material &m = b.materials[i];
            fprintf(stderr, "material %zu: type %d refl %d eta %d k %d ur %g vr %g remap %d trans %d scale %g thick %g g %g albedo %d depth %d ns %d cur %g cvr %g ceta %d ck %d mix %d %d\n",
                    i, m.type, m.reflectance, m.eta, m.k, m.u_roughness, m.v_roughness, m.remap_roughness, m.transmittance, m.scale, m.thickness,
                    m.g, m.albedo, m.max_depth, m.n_samples, m.c_u_roughness, m.c_v_roughness, m.c_eta, m.c_k, m.mix_a, m.mix_b);
        }
        for (size_t i = 0; i < b.spectra.size(); i++)
            fprintf(stderr, "spectrum %zu: type %d c %g %g %g %g n %d\n", i, b.spectra[i].type, b.spectra[i].c[0], b.spectra[i].c[1], b.spectra[i].c[2],
                    b.spectra[i].c[3], b.spectra[i].n);
        for (int i = 0; i < 471; i++) {
            uint32_t bits;
            memcpy(&bits, &fd.illuminant[i], 4);
            fprintf(stderr, "[refillum] %d %08x\n", i, bits);
        }
        for (size_t i = 0; i < lights.size(); i++) {
            uint32_t bits;
            memcpy(&bits, &lights[i].scale, 4);
            fprintf(stderr, "[reflightrec] %zu tri %d spectrum %d scale %08x two_sided %d\n", i, lights[i].triangle, lights[i].spectrum, bits, lights[i].two_sided);
        }
        // the reference's own light sampler table, bin by bin, in the format of the library's B200PT_DEBUG_LIGHTS lines
        if (const PowerLightSampler *ls = sb.integrator_base->light_sampler; ls && ls->light_num > 0) {
            const AliasTable *at = ls->lights_power_distribution->alias_table;
            for (int i = 0; i < at->size && i < 4096; i++)
                fprintf(stderr, "[reflight] %d pdf %.9g bin p %.9g first %d second %d\n", i, (double)at->pdfs[i], (double)at->bins[i].p,
                        at->bins[i].first_idx, at->bins[i].second_idx);
        }
    }
    d.n_lights = (int32_t)lights.size();
    d.lights = lights.data();
    d.camera = &cam;
    d.filter = &flt;
    d.film = &fd;

    b200pt_scene *scene = nullptr;
    b200pt_film *film = nullptr;
    check(b200pt_scene_create(&d, 0, &scene), "b200pt_scene_create");
    b200pt_build_stats bs = {};
    check(b200pt_bvh_build(scene, &bs), "b200pt_bvh_build");
    printf("b200pt BVH building: %lld primitives, %.3f ms on the device (%lld BVH8 nodes)\n", (long long)bs.n_triangles, bs.ms_total,
           (long long)bs.n_bvh8_nodes);
    check(b200pt_film_create(scene, &film), "b200pt_film_create");
    b200pt_render_opts o = {};
    o.spp_total = sb.samples_per_pixel.value();
    o.sample_begin = 0;
    o.sample_end = o.spp_total;
    o.max_depth = sb.wavefront_path_integrator->max_depth;
    o.regularize = sb.wavefront_path_integrator->regularize ? 1 : 0;
    o.sampler = sb.sampler_type == "independent" ? 1 : 0;
    b200pt_render_stats st = {};
    check(b200pt_render(scene, film, &o, &st), "b200pt_render");
    printf("b200pt wavefront pathtracing: %.3f ms on the device, %lld paths, %lld + %lld rays, %lld passes\n", st.ms_total,
           (long long)st.paths, (long long)st.rays_closest, (long long)st.rays_shadow, (long long)st.passes);

    // Film::write_to_png (base/film.cu:57-99) + the float film next to it (as oracle/_ref/pbrt-minus-ref writes it)
    std::vector<unsigned char> rgba((size_t)res.x * res.y * 4);
    int32_t nan_pixels = 0;
    check(b200pt_film_read_srgb8(film, rgba.data(), &nan_pixels), "b200pt_film_read_srgb8");
    if (nan_pixels > 0) {
        fprintf(stderr, "b200pt bridge: %d pixels carry NAN components\n", nan_pixels);
        exit(1);
    }
    if (unsigned error = lodepng::encode(sb.output_filename, rgba, res.x, res.y); error) {
        fprintf(stderr, "lodepng::encoder error %u: %s\n", error, lodepng_error_text(error));
        exit(1);
    }
    std::vector<float> rgb((size_t)res.x * res.y * 3);
    check(b200pt_film_read(film, rgb.data()), "b200pt_film_read");
    if (FILE *fp = fopen((sb.output_filename + ".f32").c_str(), "wb")) {
        fwrite(rgb.data(), sizeof(float), rgb.size(), fp);
        fclose(fp);
    }
    b200pt_film_destroy(film);
    b200pt_scene_destroy(scene);
}
