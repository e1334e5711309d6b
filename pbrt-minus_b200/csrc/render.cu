// render.cu -- the wavefront path integrator: scene-side tables, SoA path pool, stage kernels,
// film, and the host loop that drives a pass as a fixed kernel sequence without host syncs.
//
// Replaces WavefrontPathIntegrator::render (integrators/wavefront_path.cu:594-735) and its kernels:
//   gpu_generate_new_path :245-287  -> k_generate     (sampler start, filter sample, wavelengths, camera ray)
//   ray_cast              :348-366  -> k_trace_wf<false> (trace.cu)
//   control_logic         :68-214   -> k_logic        (termination, RR, Le + MIS, film add, material routing)
//   gpu_evaluate_material :322-346 + sample_bsdf :438-475 + sample_Ld_volume (megakernel_path.cu:185-264)
//                                   -> k_shade<MAT>   (one kernel per material, BSDF in registers,
//                                                      NEE emitted as a shadow-queue entry)
//   inline fast_intersect / compute_transmittance (base/integrator_base.cu:12-71)
//                                   -> k_trace_wf<true> (any-hit on the shadow queue + resolve)
//   write_frame_buffer :216-234 + CPU std::sort :689-694 + RGBFilm::add_sample (film/rgb_film.cu:63-80)
//                                   -> fp32 vector atomics on the film inside k_logic
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <numeric>
#include <set>

#include "bxdf.cuh"
#include "layered.cuh"
#include "common.cuh"
#include "sampler.cuh"
#include "sphere.cuh"
#include "traverse.cuh"
#include "wavefront.cuh"

struct b200pt_film {
    b200pt_scene *scene = nullptr;
    int32_t width = 0, height = 0;
    float4 *d_accum = nullptr; // {sum_x, sum_y, sum_z, sum_w} (sensor XYZ * filter weight, filter weight)
};

namespace b200pt {

struct RenderData {
    bool have_render = false;
    std::vector<void *> allocs;
    // scene tables
    SpecRec *d_spec = nullptr;
    float *d_spec_pool = nullptr;
    float *d_illuminant = nullptr;
    MatRec *d_materials = nullptr;
    LightRec *d_lights = nullptr;
    AliasBin *d_alias = nullptr;
    LightExt *d_light_ext = nullptr;
    int n_lights = 0, n_materials = 0, n_spectra = 0;
    int n_infinite = 0, infinite[4] = {0, 0, 0, 0};
    int n_area_lights = 0; // emissive triangles (k_emit runs only when there are any)
    float scene_radius = 0.f;
    bool has_mat[WF_NMAT] = {};
    bool has_mix = false;
    TexTables tex = {};                // textures: records, images, the fitted RGB-to-spectrum table
    std::vector<ImageRec> h_images;    // device pointers of the uploaded images
    EnvMap env = {};                   // image infinite light (env.light = -1: none)
    MediumRec *d_media = nullptr;      // media scenes only
    uint32_t *d_tri_medium = nullptr;
    int n_media = 0, camera_medium = -1;
    FilterTab filter = {};
    CameraRec camera = {};
    float *d_cie[3] = {nullptr, nullptr, nullptr};
    int width = 0, height = 0;
    float imaging_ratio = 1.f;
    float out_rgb[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    std::vector<b200pt_light> h_lights;
    // path pool (allocated on first render / when the pool size changes)
    int pool_size = 0;
    PathPool pool = {};
    WFQueues q = {};
    WFCounters *d_counters = nullptr;
    WFCounters *h_counters = nullptr; // pinned
    std::vector<void *> pool_allocs;
    std::vector<b200pt_film *> films;  // live films of this scene (their accumulators die with the scene)
};

template <typename T>
static int dev_upload(RenderData *r, T **dst, const T *src, size_t count, cudaStream_t st) {
    *dst = nullptr;
    void *p = nullptr;
    CU_TRY(cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
    r->allocs.push_back(p);
    if (count) CU_TRY(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, st));
    *dst = (T *)p;
    return B200PT_OK;
}

int build_rgb2spec(b200pt_scene *s, const double *fit, double **d_coeff_out, double **d_znodes_out, std::vector<void *> &allocs); // rgb2spec.cu

void free_render(b200pt_scene *s) {
    if (!s->render) return;
    // films outlive neither their scene nor its device context: free the accumulators here and leave dead handles
    // behind, so a later b200pt_film_* call on one returns B200PT_ERR_STATE instead of touching freed memory
    for (b200pt_film *f : s->render->films) {
        if (f->d_accum) cudaFree(f->d_accum);
        f->d_accum = nullptr;
        f->scene = nullptr;
    }
    for (void *p : s->render->allocs) cudaFree(p);
    for (void *p : s->render->pool_allocs) cudaFree(p);
    if (s->render->d_counters) cudaFree(s->render->d_counters);
    if (s->render->h_counters) cudaFreeHost(s->render->h_counters);
    delete s->render;
    s->render = nullptr;
}

// PiecewiseConstant1D ctor (distribution/piecewise_constant_1d.h:14-41), in fp32 like the reference
static float build_pc1d(const float *f, int n, float mn, float mx, float *func, float *cdf) {
    for (int i = 0; i < n; i++) func[i] = std::fabs(f[i]);
    cdf[0] = 0.f;
    for (int i = 1; i < n + 1; i++) cdf[i] = cdf[i - 1] + func[i - 1] * (mx - mn) / n;
    const float func_int = cdf[n];
    if (func_int == 0.f) {
        for (int i = 1; i < n + 1; i++) cdf[i] = float(i) / float(n);
    } else {
        for (int i = 1; i < n + 1; i++) cdf[i] /= func_int;
    }
    return func_int;
}

__global__ void k_light_power(SpectrumTable tab, const LightRec *lights, const LightExt *ext, float scene_radius, int n, S4 lambda,
                              S4 pdf, float *out, EnvMap env, TexTables tex) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // light_samplers/power_light_sampler.cu:15-21: phi at the wavelengths of sample_visible(0.5), which the
    // reference evaluates on the HOST; the 4th wavelength sits within an ulp of the 830 nm cut-off of
    // visible_wavelengths_pdf, so lambda / pdf come from the host's libm (see b200pt_upload_render_data)
    const LightRec L = lights[i];
    const S4 Le = L.spectrum >= 0 ? spectrum_sample(tab, L.spectrum, lambda) * L.scale : s4(0.f);
    S4 phi;
    switch (light_type(L)) {
    case B200PT_LIGHT_DISTANT: phi = Le * sqr(scene_radius); break;                        // lights/distant_light.cu:52-55
    case B200PT_LIGHT_UNIFORM_INFINITE:                                                     // lights/uniform_infinite_light.cu:34-37
        phi = spectrum_sample(tab, L.spectrum, lambda) * (4.f * kPi * kPi * sqr(scene_radius) * L.scale);
        break;
    case B200PT_LIGHT_IMAGE_INFINITE: {                                                     // lights/image_infinite_light.cu:118-135
        S4 sum = s4(0.f);
        for (int x = 0; x < env.width; x++)
            for (int y = 0; y < env.height; y++) {
                V3 rgb = env_fetch(env, x, y);
                rgb = v3(fmaxf(rgb.x, 0.f), fmaxf(rgb.y, 0.f), fmaxf(rgb.z, 0.f));
                sum = sum + rgb_illuminant_sample(tex, tab, rgb, lambda);
            }
        phi = sum * (4.f * kPi * kPi * sqr(scene_radius) * env.scale) / (float)(env.width * env.height);
        break;
    }
    case B200PT_LIGHT_SPOT: {                                                               // lights/spot_light.cu:104-108
        const LightExt e = ext[L.ext];
        phi = Le * 2.f * kPi * ((1.f - e.cos_start) + (e.cos_start - e.cos_end) / 2.f);
        break;
    }
    default: phi = Le * (kPi * ((L.flags & 1) ? 2.f : 1.f) * L.area);                       // lights/diffuse_area_light.cu:130-134
    }
    out[i] = average(safe_div(phi, pdf));
}

} // namespace b200pt

using namespace b200pt;

// ---- the light sampler's weights as the reference's HOST constructor rounds them (light_samplers/power_light_sampler.cu:15-21) --
// The weights decide the ORDER of the alias-table bins, so two lights whose powers differ in the last bit -- or tie -- on the
// host must do the same here: host compilers do not contract a * b + c (the device does), hence 1 + x * x in the sigmoid and
// (1 - t) * a + t * b in the table lookups are restated in plain C++ (fma only where the reference writes FMA itself).
static float host_sigmoid_poly(const float *c, float lambda) { // spectrum_util/rgb_sigmoid_polynomial.h:14-37
    const float x = std::fma(lambda, std::fma(lambda, c[0], c[1]), c[2]);
    if (std::isinf(x)) return x > 0.f ? 1.f : 0.f;
    const float xx = x * x;
    const float root = std::sqrt(1.f + xx);
    return 0.5f + x / (2.f * root);
}
static float host_lerp(float t, float a, float b) { // util/math.h:74-76
    const float l = (1.f - t) * a, h = t * b;
    return l + h;
}
static float host_dense_lookup(const float *v, float lambda) { // spectra/densely_sampled_spectrum.h:46-55
    const int fl = (int)std::floor(lambda), ce = (int)std::ceil(lambda);
    if (fl < 360 || ce > 830) return 0.f;
    return host_lerp(lambda - (float)fl, v[fl - 360], v[ce - 360]);
}
static float host_piecewise_lookup(const float *lam, const float *val, int n, float lambda) { // spectra/piecewise_linear_spectrum.cu:93-103
    if (n == 0 || lambda < lam[0] || lambda > lam[n - 1]) return 0.f;
    int size = n - 2, first = 1;
    while (size > 0) { // find_interval (util/math.h:129-141)
        const int half = size >> 1, middle = first + half;
        const bool pred = lam[middle] <= lambda;
        first = pred ? middle + 1 : first;
        size = pred ? size - (half + 1) : half;
    }
    int o = first - 1;
    o = o < 0 ? 0 : (o > n - 2 ? n - 2 : o);
    const float t = (lambda - lam[o]) / (lam[o + 1] - lam[o]);
    return host_lerp(t, val[o], val[o + 1]);
}
static float host_spectrum_at(const std::vector<SpecRec> &spec, const std::vector<float> &pool, const float *illuminant, int id, float lambda) {
    const SpecRec &r = spec[id];
    switch (r.type) {
    case B200PT_SPEC_CONSTANT: return r.c[0];
    case B200PT_SPEC_RGB_ALBEDO: return host_sigmoid_poly(r.c, lambda);
    case B200PT_SPEC_RGB_UNBOUNDED: return r.c[3] * host_sigmoid_poly(r.c, lambda);
    case B200PT_SPEC_RGB_ILLUMINANT: {
        const float a = r.c[3] * host_sigmoid_poly(r.c, lambda);
        return a * host_dense_lookup(illuminant, lambda);
    }
    case B200PT_SPEC_PIECEWISE: return host_piecewise_lookup(pool.data() + r.offset, pool.data() + r.offset + r.n, r.n, lambda);
    case B200PT_SPEC_DENSE: return host_dense_lookup(pool.data() + r.offset, lambda);
    }
    return 0.f;
}
// phi(lambda).safe_div(pdf).average() of one light; false for the kinds that stay on the device (image infinite lights)
static bool host_light_weight(const LightRec &L, const std::vector<LightExt> &ext, const std::vector<SpecRec> &spec, const std::vector<float> &pool,
                              const float *illuminant, float scene_radius, const float *lambda, const float *pdf, float *out) {
    float f;
    bool scaled_le = true; // Le = sample(lambda) * scale first, then the geometric factor
    switch (light_type(L)) {
    case B200PT_LIGHT_IMAGE_INFINITE: return false;
    case B200PT_LIGHT_DISTANT: f = scene_radius * scene_radius; break;
    case B200PT_LIGHT_UNIFORM_INFINITE:
        f = 4.f * kPi * kPi * (scene_radius * scene_radius) * L.scale;
        scaled_le = false;
        break;
    case B200PT_LIGHT_SPOT: {
        const LightExt &e = ext[L.ext];
        f = 0.f; // applied in two steps below, as the device kernel does
        float sum = 0.f;
        for (int i = 0; i < 4; i++) {
            float le = L.spectrum >= 0 ? host_spectrum_at(spec, pool, illuminant, L.spectrum, lambda[i]) * L.scale : 0.f;
            le = le * 2.f * kPi * ((1.f - e.cos_start) + (e.cos_start - e.cos_end) / 2.f);
            sum += pdf[i] == 0.f ? 0.f : le / pdf[i];
        }
        *out = sum / 4.f;
        return true;
    }
    default: f = kPi * ((L.flags & 1) ? 2.f : 1.f) * L.area; // lights/diffuse_area_light.cu:130-134
    }
    float sum = 0.f;
    for (int i = 0; i < 4; i++) {
        float le = L.spectrum >= 0 ? host_spectrum_at(spec, pool, illuminant, L.spectrum, lambda[i]) : 0.f;
        if (scaled_le) le = L.spectrum >= 0 ? le * L.scale : 0.f;
        const float phi = le * f;
        sum += pdf[i] == 0.f ? 0.f : phi / pdf[i];
    }
    *out = sum / 4.f;
    return true;
}

// AliasTable ctor (distribution/alias_table.cu:9-68): repeatedly take the smallest bin, top it up
// from the largest.  The reference re-sorts the whole vector per step (O(n^2 log n)); an ordered
// set gives the same extraction order (up to ties) in O(n log n), which config C4 (100 k lights) needs.
//
// Ties decide the layout too: bins of EQUAL weight (the two triangles of a quad light, after both were topped up to exactly 1)
// leave the reference's std::sort in the order its vector held them, which an ordered set keyed on (p, index) does not
// reproduce -- and a different bin order picks a different light for the same variate.  With `replay_sort` the table is
// built the reference's way (re-sort the remaining bins with std::sort, take the front, top it up from the back), which
// gives the same order bin for bin because it IS the same libstdc++ sort on the same sequence; the callers ask for it
// while the O(n^2 log n) cost is negligible and keep the ordered set beyond (C4's 100 k lights: statistical parity).
static void build_alias(const std::vector<float> &values, std::vector<AliasBin> &bins, std::vector<float> &pdfs, bool replay_sort = false) {
    const int n = (int)values.size();
    if (replay_sort && n > 0) {
        double total = 0.0;
        for (float v : values) total += v;
        if (total != 0.0) {
            struct R {
                float p;
                int first, second;
                bool operator<(const R &o) const { return p < o.p; }
            };
            std::vector<R> rest(n);
            pdfs.assign(n, 0.f);
            for (int i = 0; i < n; i++) {
                rest[i] = R{float(values[i] * n / total), i, -1};
                pdfs[i] = float(values[i] / total);
            }
            bins.clear();
            bins.reserve(n);
            size_t front = 0;   // bins before `front` are done (the reference erases them; sorting the tail is the same sort)
            while (n - front > 1) {
                std::sort(rest.begin() + front, rest.end());
                R &lo = rest[front];
                if (lo.p < 1.f - 0.001f) {
                    lo.second = rest.back().first;
                    rest.back().p -= 1.f - lo.p;
                }
                bins.push_back(AliasBin{lo.p, lo.first, lo.second});
                front++;
            }
            bins.push_back(AliasBin{rest[front].p, rest[front].first, rest[front].second});
            return;
        }
    }
    bins.assign(n, AliasBin{1.f, 0, -1});
    pdfs.assign(n, 0.f);
    double sum = 0.0;
    for (float v : values) sum += v;
    if (sum == 0.0) {
        for (int i = 0; i < n; i++) {
            bins[i] = AliasBin{1.f, i, -1};
            pdfs[i] = float(i) / n;
        }
        return;
    }
    struct B {
        float p;
        int first, second, seq;
    };
    auto cmp = [](const B &a, const B &b) { return a.p < b.p || (a.p == b.p && a.seq < b.seq); };
    std::vector<B> all(n);
    for (int i = 0; i < n; i++) {
        all[i] = B{float(values[i] * n / sum), i, -1, i};
        pdfs[i] = float(values[i] / sum);
    }
    std::multiset<B, decltype(cmp)> todo(all.begin(), all.end(), cmp);
    const float tol = 0.001f;
    int out = 0;
    while (!todo.empty()) {
        if (todo.size() == 1) {
            const B b = *todo.begin();
            bins[out++] = AliasBin{b.p, b.first, b.second};
            break;
        }
        B lo = *todo.begin();
        todo.erase(todo.begin());
        if (lo.p < 1.f - tol) {
            const float under = 1.f - lo.p;
            auto it = std::prev(todo.end());
            B hi = *it;
            todo.erase(it);
            lo.second = hi.first;
            hi.p -= under;
            todo.insert(hi);
        }
        bins[out++] = AliasBin{lo.p, lo.first, lo.second};
    }
}

int b200pt_upload_render_data(b200pt_scene *s, const b200pt_scene_desc *desc) {
    RenderData *r = new RenderData();
    s->render = r;
    cudaStream_t st = s->stream;
    // ---- spheres
    s->n_spheres = desc->n_spheres;
    if (desc->n_spheres > 0) {
        if (!desc->spheres) {
            set_error("b200pt_scene_create: n_spheres > 0 but spheres == NULL");
            return B200PT_ERR_INVALID;
        }
        std::vector<SphereRec> recs(desc->n_spheres);
        for (int i = 0; i < desc->n_spheres; i++) {
            const b200pt_sphere &sp = desc->spheres[i];
            SphereRec &o = recs[i];
            for (int k = 0; k < 12; k++) {
                o.rfo[k] = sp.render_from_object[k];
                o.ofr[k] = sp.object_from_render[k];
            }
            // shapes/sphere.h ctor: z range clamped to the radius, theta range from z
            o.radius = sp.radius;
            o.z_min = std::min(std::max(std::min(sp.z_min, sp.z_max), -sp.radius), sp.radius);
            o.z_max = std::min(std::max(std::max(sp.z_min, sp.z_max), -sp.radius), sp.radius);
            o.theta_z_min = std::acos(std::min(std::max(std::min(sp.z_min, sp.z_max) / sp.radius, -1.f), 1.f));
            o.theta_z_max = std::acos(std::min(std::max(std::max(sp.z_min, sp.z_max) / sp.radius, -1.f), 1.f));
            o.phi_max = sp.phi_max;
            o.flip = (sp.reverse_orientation != 0) ^ (sp.transform_swaps_handedness != 0);
            o.material = sp.material;
            o.medium = sp.medium;
            o.kind = sp.kind;
            if (sp.kind == 1) { // Disk (shapes/disk.cu:7-20): height, inner radius as given
                o.z_min = sp.z_min;
                o.z_max = sp.z_max;
                o.theta_z_min = o.theta_z_max = 0.f;
            } else if (sp.kind != 0) {
                set_error("b200pt_scene_create: quadric %d has kind %d", i, sp.kind);
                return B200PT_ERR_INVALID;
            }
            {   // padded render-space box (Transform::operator()(Bounds3f), euclidean_space/transform.h:326-351)
                const float bmin[3] = {-sp.radius, -sp.radius, sp.kind == 1 ? sp.z_min : o.z_min},
                            bmax[3] = {sp.radius, sp.radius, sp.kind == 1 ? sp.z_min : o.z_max};
                float ext = 0.f, mag = 0.f;
                for (int a = 0; a < 3; a++) {
                    float mn = sp.render_from_object[4 * a + 3], mx = mn;
                    for (int k = 0; k < 3; k++) {
                        const float x = sp.render_from_object[4 * a + k] * bmin[k], y = sp.render_from_object[4 * a + k] * bmax[k];
                        mn += std::min(x, y);
                        mx += std::max(x, y);
                    }
                    o.box_lo[a] = mn;
                    o.box_hi[a] = mx;
                    ext = std::max(ext, mx - mn);
                    mag = std::max(mag, std::max(std::fabs(mn), std::fabs(mx)));
                }
                const float pad = 1e-3f * ext + 1e-5f * mag + 1e-30f;
                for (int a = 0; a < 3; a++) {
                    o.box_lo[a] -= pad;
                    o.box_hi[a] += pad;
                }
            }
            o.light = sp.light;
            if (sp.light >= 0) {
                if (sp.kind != 0) {
                    set_error("b200pt_scene_create: emissive disks are not supported");
                    return B200PT_ERR_UNSUPPORTED;
                }
                if (sp.light >= desc->n_lights || !desc->lights || desc->lights[sp.light].type != B200PT_LIGHT_SPHERE_AREA ||
                    desc->lights[sp.light].triangle != i) {
                    set_error("b200pt_scene_create: quadric %d: light %d is not a sphere area light on it", i, sp.light);
                    return B200PT_ERR_INVALID;
                }
            } else if (sp.light != -1) {
                set_error("b200pt_scene_create: quadric %d has light id %d", i, sp.light);
                return B200PT_ERR_INVALID;
            }
        }
        CU_TRY(cudaMalloc((void **)&s->d_spheres, recs.size() * sizeof(SphereRec)));
        CU_TRY(cudaMemcpyAsync(s->d_spheres, recs.data(), recs.size() * sizeof(SphereRec), cudaMemcpyHostToDevice, st));
        CU_TRY(cudaStreamSynchronize(st));
    }
    if (!desc->film || !desc->camera || !desc->filter) return B200PT_OK; // traversal-only scene
    const b200pt_film_desc &fd = *desc->film;
    if (fd.width <= 0 || fd.height <= 0 || !fd.cie_x || !fd.cie_y || !fd.cie_z || !fd.illuminant) {
        set_error("b200pt_scene_create: incomplete film description");
        return B200PT_ERR_INVALID;
    }
    // ---- array pointers the counts promise (a bad description comes back as a status code, never as a fault)
    if (desc->n_spectra < 0 || desc->n_materials < 0 || desc->n_lights < 0 || (desc->n_spectra > 0 && !desc->spectra) ||
        (desc->n_materials > 0 && !desc->materials) || (desc->n_lights > 0 && !desc->lights)) {
        set_error("b200pt_scene_create: a spectra / materials / lights count is > 0 but its array is NULL");
        return B200PT_ERR_INVALID;
    }
    // ---- spectra
    std::vector<SpecRec> spec(desc->n_spectra);
    std::vector<float> pool;
    for (int i = 0; i < desc->n_spectra; i++) {
        const b200pt_spectrum &sp = desc->spectra[i];
        SpecRec &o = spec[i];
        o.type = sp.type;
        o.n = sp.n;
        o.offset = (int)pool.size();
        o.pad = 0;
        for (int k = 0; k < 4; k++) o.c[k] = sp.c[k];
        if (sp.type == B200PT_SPEC_PIECEWISE) {
            if (sp.n < 2 || !sp.lambda || !sp.value) {
                set_error("b200pt_scene_create: spectrum %d: piecewise spectrum needs >= 2 knots", i);
                return B200PT_ERR_INVALID;
            }
            pool.insert(pool.end(), sp.lambda, sp.lambda + sp.n);
            pool.insert(pool.end(), sp.value, sp.value + sp.n);
        } else if (sp.type == B200PT_SPEC_DENSE) {
            if (sp.n != 471 || !sp.value) {
                set_error("b200pt_scene_create: spectrum %d: dense spectrum needs 471 samples", i);
                return B200PT_ERR_INVALID;
            }
            pool.insert(pool.end(), sp.value, sp.value + sp.n);
        } else if (sp.type < 0 || sp.type > B200PT_SPEC_DENSE) {
            set_error("b200pt_scene_create: spectrum %d: unknown type %d", i, sp.type);
            return B200PT_ERR_INVALID;
        }
    }
    r->n_spectra = desc->n_spectra;
    int rc;
    if ((rc = dev_upload(r, &r->d_spec, spec.data(), spec.size(), st))) return rc;
    if ((rc = dev_upload(r, &r->d_spec_pool, pool.data(), pool.size(), st))) return rc;
    if ((rc = dev_upload(r, &r->d_illuminant, fd.illuminant, 471, st))) return rc;
    if ((rc = dev_upload(r, &r->d_cie[0], fd.cie_x, 471, st))) return rc;
    if ((rc = dev_upload(r, &r->d_cie[1], fd.cie_y, 471, st))) return rc;
    if ((rc = dev_upload(r, &r->d_cie[2], fd.cie_z, 471, st))) return rc;
    // ---- materials
    std::vector<MatRec> mats(desc->n_materials);
    auto spec_ok = [&](int id) { return id >= 0 && id < desc->n_spectra; };
    for (int i = 0; i < desc->n_materials; i++) {
        const b200pt_material &m = desc->materials[i];
        mats[i] = MatRec{m.type, m.reflectance, m.eta, m.k, m.u_roughness, m.v_roughness, m.remap_roughness, m.transmittance, m.scale,
                         m.thickness, m.g, m.albedo, m.max_depth, m.n_samples, m.c_u_roughness, m.c_v_roughness, m.c_eta, m.c_k, m.mix_a, m.mix_b,
                         m.reflectance_tex1, m.transmittance_tex1, m.u_roughness_tex1, m.v_roughness_tex1, m.eta_tex1, m.k_tex1};
        if (m.reflectance_tex1 < 0 || m.reflectance_tex1 > desc->n_textures || m.transmittance_tex1 < 0 || m.transmittance_tex1 > desc->n_textures ||
            m.u_roughness_tex1 < 0 || m.u_roughness_tex1 > desc->n_textures || m.v_roughness_tex1 < 0 || m.v_roughness_tex1 > desc->n_textures ||
            m.eta_tex1 < 0 || m.eta_tex1 > desc->n_textures || m.k_tex1 < 0 || m.k_tex1 > desc->n_textures) {
            set_error("b200pt_scene_create: material %d names a texture that does not exist", i);
            return B200PT_ERR_INVALID;
        }
        bool ok = true;
        if (m.type >= 0 && m.type < WF_NMAT) r->has_mat[m.type] = true;
        auto refl_ok = [&]() { return spec_ok(m.reflectance) || m.reflectance_tex1 > 0; };
        if (m.type == B200PT_MAT_DIFFUSE) ok = refl_ok();
        else if (m.type == B200PT_MAT_DIFFUSE_TRANSMISSION) ok = refl_ok() && (spec_ok(m.transmittance) || m.transmittance_tex1 > 0);
        else if (m.type == B200PT_MAT_MIX) {
            // materials/mix_material.cu:6-16: amount in (0, 1); the two materials must not be mixes themselves
            // ("you should not see MixMaterial here", integrators/wavefront_path.cu:207-210)
            ok = m.scale > 0.f && m.scale < 1.f && m.mix_a >= 0 && m.mix_a < desc->n_materials && m.mix_b >= 0 && m.mix_b < desc->n_materials &&
                 desc->materials[m.mix_a].type != B200PT_MAT_MIX && desc->materials[m.mix_b].type != B200PT_MAT_MIX;
            r->has_mix = true;
        } else if (m.type == B200PT_MAT_COATED_DIFFUSE)
            ok = refl_ok() && spec_ok(m.eta) && spec_ok(m.albedo) && m.max_depth >= 0 && m.n_samples >= 1;
        else if (m.type == B200PT_MAT_COATED_CONDUCTOR)
            ok = spec_ok(m.eta) && spec_ok(m.albedo) && m.max_depth >= 0 && m.n_samples >= 1 &&
                 (((spec_ok(m.c_eta) || m.eta_tex1 > 0) && (spec_ok(m.c_k) || m.k_tex1 > 0)) || (m.c_eta < 0 && m.eta_tex1 == 0 && refl_ok()));
        else if (m.type == B200PT_MAT_CONDUCTOR)
            ok = ((spec_ok(m.eta) || m.eta_tex1 > 0) && (spec_ok(m.k) || m.k_tex1 > 0)) || (m.eta < 0 && m.eta_tex1 == 0 && refl_ok());
        else if (m.type == B200PT_MAT_DIELECTRIC) ok = spec_ok(m.eta);
        else if (m.type == B200PT_MAT_INTERFACE) ok = desc->n_media > 0; // only meaningful between media
        else ok = false;
        if (!ok) {
            set_error("b200pt_scene_create: material %d is inconsistent (type %d)", i, m.type);
            return B200PT_ERR_INVALID;
        }
    }
    r->n_materials = desc->n_materials;
    if ((rc = dev_upload(r, &r->d_materials, mats.data(), mats.size(), st))) return rc;
    if (desc->tri_material)
        for (int64_t i = 0; i < desc->n_triangles; i++)
            if (desc->tri_material[i] < 0 || desc->tri_material[i] >= desc->n_materials) {
                set_error("b200pt_scene_create: triangle %lld has material id %d", (long long)i, desc->tri_material[i]);
                return B200PT_ERR_INVALID;
            }
    if (desc->n_materials == 0) {
        set_error("b200pt_scene_create: a render scene needs at least one material");
        return B200PT_ERR_INVALID;
    }
    for (int i = 0; i < desc->n_spheres; i++)
        if (desc->spheres[i].material < 0 || desc->spheres[i].material >= desc->n_materials) {
            set_error("b200pt_scene_create: quadric %d has material id %d", i, desc->spheres[i].material);
            return B200PT_ERR_INVALID;
        }
    // every emissive triangle must name an area light that names it back (k_logic reads lights[tri_light[t]])
    if (desc->tri_light)
        for (int64_t t = 0; t < desc->n_triangles; t++) {
            const int32_t li = desc->tri_light[t];
            if (li == -1) continue;
            if (li < 0 || li >= desc->n_lights || desc->lights[li].type != B200PT_LIGHT_AREA || desc->lights[li].triangle != t) {
                set_error("b200pt_scene_create: tri_light[%lld] = %d is not an area light on that triangle", (long long)t, li);
                return B200PT_ERR_INVALID;
            }
        }
    // ---- media
    r->n_media = desc->n_media;
    r->camera_medium = desc->n_media > 0 ? desc->camera->medium : -1;
    if (desc->n_media > 0) {
        if (!desc->media || (desc->n_triangles > 0 && !desc->tri_medium)) {
            set_error("b200pt_scene_create: n_media > 0 needs media and tri_medium");
            return B200PT_ERR_INVALID;
        }
        std::vector<MediumRec> med(desc->n_media);
        for (int i = 0; i < desc->n_media; i++) {
            const b200pt_medium &m = desc->media[i];
            if (!spec_ok(m.sigma_a) || !spec_ok(m.sigma_s)) {
                set_error("b200pt_scene_create: medium %d: bad spectrum id", i);
                return B200PT_ERR_INVALID;
            }
            med[i] = MediumRec{m.sigma_a, m.sigma_s, m.scale, m.g};
        }
        auto med_ok = [&](uint32_t id1) { return id1 <= (uint32_t)desc->n_media; };
        for (int64_t i = 0; i < desc->n_triangles; i++) {
            const uint32_t c = desc->tri_medium[i];
            if (c != 0u && c != 0xffffffffu && !(med_ok(c & 0xffffu) && med_ok(c >> 16))) {
                set_error("b200pt_scene_create: triangle %lld has medium code 0x%x", (long long)i, c);
                return B200PT_ERR_INVALID;
            }
        }
        if (r->camera_medium < -1 || r->camera_medium >= desc->n_media) {
            set_error("b200pt_scene_create: camera medium %d", r->camera_medium);
            return B200PT_ERR_INVALID;
        }
        for (int i = 0; i < desc->n_spheres; i++) {
            const uint32_t c = desc->spheres[i].medium;
            if (c != 0u && c != 0xffffffffu && !(med_ok(c & 0xffffu) && med_ok(c >> 16))) {
                set_error("b200pt_scene_create: quadric %d has medium code 0x%x", i, c);
                return B200PT_ERR_INVALID;
            }
        }
        if ((rc = dev_upload(r, &r->d_media, med.data(), med.size(), st))) return rc;
        if ((rc = dev_upload(r, &r->d_tri_medium, desc->tri_medium, (size_t)desc->n_triangles, st))) return rc;
    }
    // ---- images (textures, the image infinite light) and the RGB-to-spectrum table their lookups need
    r->env.light = -1;
    if (desc->n_images > 0) {
        if (!desc->images || !desc->rgb2spec_fit) {
            set_error("b200pt_scene_create: n_images > 0 needs images and rgb2spec_fit");
            return B200PT_ERR_INVALID;
        }
        r->h_images.resize(desc->n_images);
        for (int i = 0; i < desc->n_images; i++) {
            const b200pt_image &im = desc->images[i];
            if (im.width <= 0 || im.height <= 0 || !im.rgb) {
                set_error("b200pt_scene_create: image %d is empty", i);
                return B200PT_ERR_INVALID;
            }
            float *d_rgb = nullptr;
            if ((rc = dev_upload(r, &d_rgb, im.rgb, (size_t)im.width * im.height * 3, st))) return rc;
            r->h_images[i] = ImageRec{im.width, im.height, d_rgb};
        }
        ImageRec *d_imgs = nullptr;
        if ((rc = dev_upload(r, &d_imgs, r->h_images.data(), r->h_images.size(), st))) return rc;
        r->tex.images = d_imgs;
        double *d_coeff = nullptr, *d_z = nullptr;
        if ((rc = build_rgb2spec(s, desc->rgb2spec_fit, &d_coeff, &d_z, r->allocs))) return rc;
        r->tex.coeff = d_coeff;
        r->tex.z_nodes = d_z;
    }
    // ---- textures
    if (desc->n_textures > 0) {
        if (!desc->textures) {
            set_error("b200pt_scene_create: n_textures > 0 but textures == NULL");
            return B200PT_ERR_INVALID;
        }
        std::vector<TexRec> tex(desc->n_textures);
        for (int i = 0; i < desc->n_textures; i++) {
            const b200pt_texture &t = desc->textures[i];
            TexRec &o = tex[i];
            o.type = t.type; o.a = t.a; o.b = t.b; o.c = t.c;
            for (int k = 0; k < 4; k++) { o.f[k] = t.f[k]; o.map[k] = t.map[k]; }
            o.mapping = t.mapping;
            for (int k = 0; k < 12; k++) o.tfr[k] = t.texture_from_render[k];
            for (int k = 0; k < 3; k++) { o.v1[k] = t.v1[k]; o.v2[k] = t.v2[k]; }
            if (t.mapping < B200PT_MAP_UV || t.mapping > B200PT_MAP_CYLINDRICAL) {
                set_error("b200pt_scene_create: texture %d: unknown mapping %d", i, t.mapping);
                return B200PT_ERR_INVALID;
            }
            // children must come BEFORE their parent (the front-end creates them in that order): no cycles, bounded walks
            auto child_ok = [&](int id) { return id >= 0 && id < i; };
            bool ok = true;
            switch (t.type) {
            case B200PT_TEX_CONSTANT: ok = spec_ok(t.a); break;
            case B200PT_TEX_IMAGE: case B200PT_TEX_FLOAT_IMAGE: ok = t.a >= 0 && t.a < desc->n_images; break;
            case B200PT_TEX_SCALE: ok = child_ok(t.a); break;
            case B200PT_TEX_MIX: ok = child_ok(t.a) && child_ok(t.b) && (t.c < 0 || child_ok(t.c)); break;
            case B200PT_TEX_DIRECTION_MIX: ok = child_ok(t.a) && child_ok(t.b); break;
            case B200PT_TEX_FLOAT_CONSTANT: break;
            case B200PT_TEX_FLOAT_SCALE: ok = child_ok(t.a) && child_ok(t.b); break;
            default: ok = false;
            }
            if (!ok) {
                set_error("b200pt_scene_create: texture %d (type %d) is inconsistent", i, t.type);
                return B200PT_ERR_INVALID;
            }
        }
        TexRec *d_tex = nullptr;
        if ((rc = dev_upload(r, &d_tex, tex.data(), tex.size(), st))) return rc;
        r->tex.tex = d_tex;
        CU_TRY(cudaStreamSynchronize(st)); // tex is a local
    }
    // ---- lights
    r->n_lights = desc->n_lights;
    std::vector<LightRec> lights(desc->n_lights);
    std::vector<LightExt> ext;
    bool any_other = false;
    for (int i = 0; i < desc->n_lights; i++) {
        const b200pt_light &l = desc->lights[i];
        if (l.type == B200PT_LIGHT_SPHERE_AREA) {
            if (l.triangle < 0 || l.triangle >= desc->n_spheres || !spec_ok(l.spectrum) || desc->spheres[l.triangle].light != i ||
                desc->spheres[l.triangle].kind != 0) {
                set_error("b200pt_scene_create: light %d is not consistent with the sphere it names", i);
                return B200PT_ERR_INVALID;
            }
            const b200pt_sphere &sp = desc->spheres[l.triangle];
            LightRec &o = lights[i];
            memset(&o, 0, sizeof(o));
            o.scale = l.scale;
            o.spectrum = l.spectrum;
            o.flags = (l.two_sided ? 1 : 0) | (sp.reverse_orientation ? 2 : 0) | (B200PT_LIGHT_SPHERE_AREA << 4);
            const float zmin = std::min(std::max(std::min(sp.z_min, sp.z_max), -sp.radius), sp.radius);
            const float zmax = std::min(std::max(std::max(sp.z_min, sp.z_max), -sp.radius), sp.radius);
            o.area = sp.phi_max * sp.radius * (zmax - zmin); // Sphere::area (shapes/sphere.h:23-25)
            o.prim = (int32_t)desc->n_triangles + l.triangle;  // what this light's own shadow rays skip
            o.ext = l.triangle;
            r->n_area_lights++;
            continue;
        }
        if (l.type != B200PT_LIGHT_AREA) {
            if (l.type < 0 || l.type > B200PT_LIGHT_IMAGE_INFINITE || (l.type != B200PT_LIGHT_IMAGE_INFINITE && !spec_ok(l.spectrum))) {
                set_error("b200pt_scene_create: light %d is inconsistent (type %d)", i, l.type);
                return B200PT_ERR_INVALID;
            }
            any_other = true;
            LightRec &o = lights[i];
            memset(&o, 0, sizeof(o));
            for (int k = 0; k < 3; k++) o.p0[k] = l.v[k];
            o.scale = l.scale;
            o.spectrum = l.spectrum;
            o.flags = l.type << 4;
            o.prim = -1;
            if (l.type == B200PT_LIGHT_SPOT) {
                LightExt e = {};
                for (int k = 0; k < 9; k++) e.m[k] = l.light_from_render[k];
                e.cos_start = l.cos_falloff_start;
                e.cos_end = l.cos_falloff_end;
                o.ext = (int)ext.size();
                ext.push_back(e);
            }
            if (l.type == B200PT_LIGHT_IMAGE_INFINITE) {
                // ImageInfiniteLight ctor (lights/image_infinite_light.cu:8-60): per-pixel luminance, pixels below 0.001 of the
                // maximum dropped, Distribution2D = an alias table over the column sums + one per column
                if (l.image < 0 || l.image >= desc->n_images || r->env.light >= 0) {
                    set_error("b200pt_scene_create: light %d: bad image id, or a second image infinite light", i);
                    return B200PT_ERR_INVALID;
                }
                const b200pt_image &im = desc->images[l.image];
                const int W = im.width, H = im.height;
                std::vector<float> lum((size_t)W * H); // [x][y]
                float max_lum = 0.f;
                for (int x = 0; x < W; x++)
                    for (int y = 0; y < H; y++) {
                        const float *px = im.rgb + 3 * ((size_t)y * W + x);
                        const float v = (std::max(px[0], 0.f) + std::max(px[1], 0.f) + std::max(px[2], 0.f)) / 3;
                        lum[(size_t)x * H + y] = v;
                        max_lum = std::max(max_lum, v);
                    }
                if (!(max_lum > 0.f)) {
                    set_error("b200pt_scene_create: light %d: the image is black", i);
                    return B200PT_ERR_INVALID;
                }
                const double threshold = 0.001 * max_lum;
                for (float &v : lum)
                    if (v < threshold) v = 0.f;
                std::vector<float> col(W), ypdf((size_t)W * H), xpdf;
                std::vector<AliasBin> ybins((size_t)W * H), xbins;
                // small maps get the reference's own tie order (see build_alias): ~W * H^2 log H / 2 comparisons
                const bool replay_env = (double)W * H * H * 4.0 <= 3e7;
                for (int x = 0; x < W; x++) {
                    float sum = 0.f;
                    for (int y = 0; y < H; y++) sum += lum[(size_t)x * H + y];
                    col[x] = sum;
                    std::vector<float> vals(lum.begin() + (size_t)x * H, lum.begin() + (size_t)(x + 1) * H), pd;
                    std::vector<AliasBin> bn;
                    build_alias(vals, bn, pd, replay_env);
                    std::copy(bn.begin(), bn.end(), ybins.begin() + (size_t)x * H);
                    std::copy(pd.begin(), pd.end(), ypdf.begin() + (size_t)x * H);
                }
                build_alias(col, xbins, xpdf, W <= 2048);
                AliasBin *d_xb = nullptr, *d_yb = nullptr;
                float *d_xp = nullptr, *d_yp = nullptr;
                if ((rc = dev_upload(r, &d_xb, xbins.data(), xbins.size(), st))) return rc;
                if ((rc = dev_upload(r, &d_xp, xpdf.data(), xpdf.size(), st))) return rc;
                if ((rc = dev_upload(r, &d_yb, ybins.data(), ybins.size(), st))) return rc;
                if ((rc = dev_upload(r, &d_yp, ypdf.data(), ypdf.size(), st))) return rc;
                CU_TRY(cudaStreamSynchronize(st));
                EnvMap &E = r->env;
                E.light = i;
                E.width = W;
                E.height = H;
                E.rgb = r->h_images[l.image].rgb;
                E.xb = d_xb; E.xp = d_xp; E.yb = d_yb; E.yp = d_yp;
                for (int k = 0; k < 9; k++) {
                    E.rfl[k] = l.render_from_light[k];
                    E.lfr[k] = l.light_from_render[k];
                }
                E.scale = l.scale;
            }
            if (l.type == B200PT_LIGHT_UNIFORM_INFINITE || l.type == B200PT_LIGHT_IMAGE_INFINITE) {
                if (r->n_infinite >= 4) {
                    set_error("b200pt_scene_create: at most 4 infinite lights");
                    return B200PT_ERR_UNSUPPORTED;
                }
                r->infinite[r->n_infinite++] = i;
            }
            continue;
        }
        if (l.triangle < 0 || l.triangle >= desc->n_triangles || !spec_ok(l.spectrum)) {
            set_error("b200pt_scene_create: light %d is inconsistent", i);
            return B200PT_ERR_INVALID;
        }
        if (!desc->tri_light || desc->tri_light[l.triangle] != i) {
            set_error("b200pt_scene_create: tri_light[%d] must be %d", l.triangle, i);
            return B200PT_ERR_INVALID;
        }
        LightRec &o = lights[i];
        r->n_area_lights++;
        const int32_t *v = desc->indices + 3 * (size_t)l.triangle;
        for (int k = 0; k < 3; k++) {
            o.p0[k] = desc->P[3 * (size_t)v[0] + k];
            o.p1[k] = desc->P[3 * (size_t)v[1] + k];
            o.p2[k] = desc->P[3 * (size_t)v[2] + k];
        }
        o.scale = l.scale;
        o.spectrum = l.spectrum;
        o.flags = (l.two_sided ? 1 : 0) | ((desc->tri_flip && (desc->tri_flip[l.triangle] & 1)) ? 2 : 0);
        o.pmf = 0.f;
        // Triangle::area (shapes/triangle.h:41-44) with the reference's compensated cross product
        // (euclidean_space/vector3.h:151-155, util/math.h:167-173): the alias-table layout depends on the
        // ORDER of nearly equal light powers, so the area must round the way the reference's does
        auto dop = [](float a, float b, float c, float d) {
            const float cd = c * d;
            const float diff = std::fma(a, b, -cd);
            const float err = std::fma(-c, d, cd);
            return diff + err;
        };
        const float e1[3] = {o.p1[0] - o.p0[0], o.p1[1] - o.p0[1], o.p1[2] - o.p0[2]};
        const float e2[3] = {o.p2[0] - o.p0[0], o.p2[1] - o.p0[1], o.p2[2] - o.p0[2]};
        const float cx = dop(e1[1], e2[2], e1[2], e2[1]), cy = dop(e1[2], e2[0], e1[0], e2[2]), cz = dop(e1[0], e2[1], e1[1], e2[0]);
        o.area = std::sqrt(cx * cx + cy * cy + cz * cz) * 0.5f;
        o.prim = l.triangle;
        o.ext = 0;
    }
    if (any_other) {
        // Bounds3::bounding_sphere of HLBVH::bounds() (scene_builder.cu:914-925, euclidean_space/bounds3.h:142-152):
        // union of the top-level triangles' vertices, of the spheres' transformed boxes (transform.h:326-351) and, per instance, of
        // the OBJECT's bounding box taken through the instance transform the same way (TransformedPrimitive::bounds: the box of
        // the box, not the transformed vertices) -- the radius scales the power of infinite and distant lights, i.e. the light
        // sampler's pmf: without the instances the sky was picked with a slightly different probability in instanced scenes
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        const int64_t n_top = desc->n_objects > 0 ? desc->objects[0].first_triangle : desc->n_triangles;
        for (int64_t t = 0; t < n_top * 3; t++) {
            const float *p = desc->P + 3 * (size_t)desc->indices[t];
            for (int k = 0; k < 3; k++) {
                lo[k] = std::min(lo[k], p[k]);
                hi[k] = std::max(hi[k], p[k]);
            }
        }
        std::vector<float> obox((size_t)desc->n_objects * 6);
        for (int i = 0; i < desc->n_objects; i++) {
            float *b = &obox[6 * (size_t)i];
            b[0] = b[1] = b[2] = INFINITY;
            b[3] = b[4] = b[5] = -INFINITY;
            const b200pt_object &ob = desc->objects[i];
            if (ob.bounds[0] <= ob.bounds[3]) { // given by the front-end (includes the definition's quadrics)
                for (int k = 0; k < 6; k++) b[k] = ob.bounds[k];
                continue;
            }
            for (int64_t t = 3 * ob.first_triangle; t < 3 * (ob.first_triangle + ob.n_triangles); t++) {
                const float *p = desc->P + 3 * (size_t)desc->indices[t];
                for (int k = 0; k < 3; k++) {
                    b[k] = std::min(b[k], p[k]);
                    b[3 + k] = std::max(b[3 + k], p[k]);
                }
            }
        }
        for (int j = 0; j < desc->n_instances; j++) {
            const b200pt_instance &in = desc->instances[j];
            if (in.object < 0 || in.object >= desc->n_objects || desc->objects[in.object].n_triangles <= 0) continue;
            const float *b = &obox[6 * (size_t)in.object];
            for (int a = 0; a < 3; a++) {
                float mn = in.render_from_instance[4 * a + 3], mx = mn;
                for (int k = 0; k < 3; k++) {
                    const float x = in.render_from_instance[4 * a + k] * b[k], y = in.render_from_instance[4 * a + k] * b[3 + k];
                    mn += std::min(x, y);
                    mx += std::max(x, y);
                }
                lo[a] = std::min(lo[a], mn);
                hi[a] = std::max(hi[a], mx);
            }
        }
        for (int i = 0; i < desc->n_spheres; i++) {
            const b200pt_sphere &sp = desc->spheres[i];
            const float zmin = std::min(std::max(std::min(sp.z_min, sp.z_max), -sp.radius), sp.radius);
            const float zmax = std::min(std::max(std::max(sp.z_min, sp.z_max), -sp.radius), sp.radius);
            // Sphere::bounds / Disk::bounds (shapes/disk.h:23-26: the square around the disk at z = height)
            const float bmin[3] = {-sp.radius, -sp.radius, sp.kind == 1 ? sp.z_min : zmin},
                        bmax[3] = {sp.radius, sp.radius, sp.kind == 1 ? sp.z_min : zmax};
            for (int a = 0; a < 3; a++) {
                float mn = sp.render_from_object[4 * a + 3], mx = mn;
                for (int k = 0; k < 3; k++) {
                    const float x = sp.render_from_object[4 * a + k] * bmin[k], y = sp.render_from_object[4 * a + k] * bmax[k];
                    mn += std::min(x, y);
                    mx += std::max(x, y);
                }
                lo[a] = std::min(lo[a], mn);
                hi[a] = std::max(hi[a], mx);
            }
        }
        r->scene_radius = 0.f;
        if (lo[0] <= hi[0]) {
            const float c[3] = {(lo[0] + hi[0]) / 2, (lo[1] + hi[1]) / 2, (lo[2] + hi[2]) / 2};
            const float d[3] = {c[0] - hi[0], c[1] - hi[1], c[2] - hi[2]};
            r->scene_radius = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        }
    }
    if ((rc = dev_upload(r, &r->d_lights, lights.data(), lights.size(), st))) return rc;
    if (!ext.empty() && (rc = dev_upload(r, &r->d_light_ext, ext.data(), ext.size(), st))) return rc;
    if (desc->n_lights > 0) {
        // power weights on the device (needs the spectra), alias table on the host
        float *d_w = nullptr;
        std::vector<float> w(desc->n_lights);
        if ((rc = dev_upload(r, &d_w, w.data(), w.size(), st))) return rc;
        SpectrumTable tab{r->d_spec, r->d_spec_pool, r->d_illuminant};
        // SampledWavelengths::sample_visible(0.5f) with the host's atanhf / coshf, as the reference's host
        // constructor computes it (spectrum_util/sampled_wavelengths.h:62-78, util/sampling.h:296-307)
        float hl[4], hp[4];
        for (int i = 0; i < 4; i++) {
            float up = 0.5f + float(i) / 4;
            if (up > 1) up -= 1;
            hl[i] = 538 - 138.888889f * std::atanh(0.85691062f - 1.82750197f * up);
            const float c = std::cosh(0.0072f * (hl[i] - 538));
            hp[i] = (hl[i] < 360 || hl[i] > 830) ? 0.f : 0.0039398042f / (c * c);
        }
        k_light_power<<<(desc->n_lights + 127) / 128, 128, 0, st>>>(tab, r->d_lights, r->d_light_ext, r->scene_radius, desc->n_lights, make_float4(hl[0], hl[1], hl[2], hl[3]),
                                                                  make_float4(hp[0], hp[1], hp[2], hp[3]), d_w, r->env, r->tex);
        CU_TRY(cudaMemcpyAsync(w.data(), d_w, w.size() * 4, cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        // every kind but the image infinite light (a sum over the map through the device's fit table) is re-evaluated with the
        // host's rounding, which is what the reference's constructor sees; B200PT_LIGHT_WEIGHTS=device keeps the kernel's values
        const char *lw = getenv("B200PT_LIGHT_WEIGHTS");
        if (!(lw && !strcmp(lw, "device"))) {
            int moved = 0;
            for (int i = 0; i < desc->n_lights; i++) {
                float hw;
                if (host_light_weight(lights[i], ext, spec, pool, fd.illuminant, r->scene_radius, hl, hp, &hw)) {
                    moved += hw != w[i];
                    w[i] = hw;
                }
            }
            if (getenv("B200PT_DEBUG_LIGHTS")) fprintf(stderr, "[lightdbg] %d of %d weights differ between host and device rounding\n", moved, desc->n_lights);
        }
        std::vector<AliasBin> bins;
        std::vector<float> pdfs;
        // B200PT_ALIAS_REPLAY_MAX moves the limit (the reference itself spends minutes there at 100 k lights)
        const char *rm = getenv("B200PT_ALIAS_REPLAY_MAX");
        build_alias(w, bins, pdfs, desc->n_lights <= (rm ? atoi(rm) : 4096));
        for (int i = 0; i < desc->n_lights; i++) lights[i].pmf = pdfs[i];
        if (getenv("B200PT_DEBUG_LIGHTS"))
            for (int i = 0; i < desc->n_lights && i < 4096; i++)
                fprintf(stderr, "[lightdbg] %d weight %.9g pdf %.9g bin p %.9g first %d second %d\n", i, (double)w[i], (double)pdfs[i],
                        (double)bins[i].p, bins[i].first, bins[i].second);
        CU_TRY(cudaMemcpyAsync(r->d_lights, lights.data(), lights.size() * sizeof(LightRec), cudaMemcpyHostToDevice, st));
        if ((rc = dev_upload(r, &r->d_alias, bins.data(), bins.size(), st))) return rc;
        CU_TRY(cudaStreamSynchronize(st));
    }
    // ---- pixel filter: FilterSampler + PiecewiseConstant2D (filters/filter_sampler.cu:4-20,
    //      distribution/piecewise_constant_2d.h:11-32)
    const b200pt_filter &fl = *desc->filter;
    if (fl.nx <= 0 || fl.ny <= 0 || !fl.f) {
        set_error("b200pt_scene_create: filter table missing");
        return B200PT_ERR_INVALID;
    }
    {
        const int nx = fl.nx, ny = fl.ny;
        std::vector<float> cfunc(nx * ny), ccdf(ny * (nx + 1)), cint(ny), mfunc(ny), mcdf(ny + 1);
        for (int v = 0; v < ny; v++) cint[v] = build_pc1d(fl.f + v * nx, nx, -fl.radius_x, fl.radius_x, &cfunc[v * nx], &ccdf[v * (nx + 1)]);
        const float mint = build_pc1d(cint.data(), ny, -fl.radius_y, fl.radius_y, mfunc.data(), mcdf.data());
        FilterTab &t = r->filter;
        t.nx = nx;
        t.ny = ny;
        t.min_x = -fl.radius_x;
        t.max_x = fl.radius_x;
        t.min_y = -fl.radius_y;
        t.max_y = fl.radius_y;
        t.marg_int = mint;
        float *p;
        if ((rc = dev_upload(r, &p, fl.f, (size_t)nx * ny, st))) return rc;
        t.f = p;
        if ((rc = dev_upload(r, &p, cfunc.data(), cfunc.size(), st))) return rc;
        t.cond_func = p;
        if ((rc = dev_upload(r, &p, ccdf.data(), ccdf.size(), st))) return rc;
        t.cond_cdf = p;
        if ((rc = dev_upload(r, &p, cint.data(), cint.size(), st))) return rc;
        t.cond_int = p;
        if ((rc = dev_upload(r, &p, mfunc.data(), mfunc.size(), st))) return rc;
        t.marg_func = p;
        if ((rc = dev_upload(r, &p, mcdf.data(), mcdf.size(), st))) return rc;
        t.marg_cdf = p;
        CU_TRY(cudaStreamSynchronize(st));
    }
    for (int k = 0; k < 16; k++) {
        r->camera.cfr[k] = desc->camera->camera_from_raster[k];
        r->camera.rfc[k] = desc->camera->render_from_camera[k];
    }
    r->camera.lens_radius = desc->camera->lens_radius;
    r->camera.focal_distance = desc->camera->focal_distance;
    r->width = fd.width;
    r->height = fd.height;
    r->imaging_ratio = fd.imaging_ratio;
    for (int k = 0; k < 9; k++) r->out_rgb[k] = fd.output_rgb_from_sensor_rgb[k];
    r->have_render = true;
    return B200PT_OK;
}

namespace b200pt {

constexpr int WF_BLOCK = 256;

// PiecewiseConstant1D::sample (distribution/piecewise_constant_1d.h:46-63)
__device__ __forceinline__ float pc1d_sample(const float *func, const float *cdf, float func_int, int size, float mn, float mx,
                                             float u, float *pdf, int *offset) {
    // find_interval(size + 1, cdf[i] <= u)
    int sz = size + 1 - 2, first = 1;
    while (sz > 0) {
        const int half = sz >> 1, middle = first + half;
        const bool pred = __ldg(cdf + middle) <= u;
        first = pred ? middle + 1 : first;
        sz = pred ? sz - (half + 1) : half;
    }
    int o = first - 1;
    o = o < 0 ? 0 : (o > size - 1 ? size - 1 : o);
    *offset = o;
    const float c0 = __ldg(cdf + o), c1 = __ldg(cdf + o + 1);
    float du = u - c0;
    if (c1 - c0 > 0.f) du /= c1 - c0;
    *pdf = func_int > 0.f ? __ldg(func + o) / func_int : 0.f;
    return lerpf((o + du) / size, mn, mx);
}

// FilterSampler::sample (filters/filter_sampler.h:23-30)
__device__ __forceinline__ void filter_sample(const FilterTab &t, float2 u, float2 *p, float *weight) {
    float pdf0, pdf1;
    int ox, oy;
    const float d1 = pc1d_sample(t.marg_func, t.marg_cdf, t.marg_int, t.ny, t.min_y, t.max_y, u.y, &pdf1, &oy);
    const float d0 = pc1d_sample(t.cond_func + oy * t.nx, t.cond_cdf + oy * (t.nx + 1), __ldg(t.cond_int + oy), t.nx, t.min_x,
                                 t.max_x, u.x, &pdf0, &ox);
    *p = make_float2(d0, d1);
    *weight = __ldg(t.f + oy * t.nx + ox) / (pdf0 * pdf1);
}

__device__ __forceinline__ Sampler load_sampler(const WF &wf, int slot, const int4 &ids) {
    Sampler s;
    s.kind = wf.sampler_kind;
    s.spp = wf.spp;
    s.sqrt_spp = wf.sqrt_spp;
    s.pixel_idx = ids.x;
    s.sample_idx = ids.y;
    s.dimension = ids.z;
    s.state = wf.pool.rng[slot];
    s.inc = (hash_i32(ids.x) << 1u) | 1u;
    return s;
}


__global__ void k_init_pool(WF wf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < wf.pool_size) {
        wf.q.free_slots[i] = i;
        wf.pool.state[i] = ST_IDLE;
    }
    if (i == 0) {
        WFCounters z = {};
        z.p[0].n_free = (unsigned)wf.pool_size; // pass 0 generates into every slot
        *wf.c = z;
    }
}

// RGBFilm::add_sample + PixelSensor::to_sensor_rgb (film/rgb_film.cu:63-80, film/pixel_sensor.h:19-26)
__device__ __forceinline__ void film_add(const WF &wf, int pixel, S4 L, S4 lambda, S4 lpdf, float weight) {
    if (weight == 0.f) return;
    const S4 sp = safe_div(L, lpdf);
    const S4 xb = make_float4(dense_lookup(wf.film.cie_x, lambda.x), dense_lookup(wf.film.cie_x, lambda.y),
                              dense_lookup(wf.film.cie_x, lambda.z), dense_lookup(wf.film.cie_x, lambda.w));
    const S4 yb = make_float4(dense_lookup(wf.film.cie_y, lambda.x), dense_lookup(wf.film.cie_y, lambda.y),
                              dense_lookup(wf.film.cie_y, lambda.z), dense_lookup(wf.film.cie_y, lambda.w));
    const S4 zb = make_float4(dense_lookup(wf.film.cie_z, lambda.x), dense_lookup(wf.film.cie_z, lambda.y),
                              dense_lookup(wf.film.cie_z, lambda.z), dense_lookup(wf.film.cie_z, lambda.w));
    const float X = wf.film.imaging_ratio * average(xb * sp), Y = wf.film.imaging_ratio * average(yb * sp),
                Z = wf.film.imaging_ratio * average(zb * sp);
    // one 128-bit vector reduction per finished path (sm_90+), instead of the reference's
    // sort-by-pixel on the CPU + serial per-pixel adds
    atomicAdd(wf.film.accum + pixel, make_float4(weight * X, weight * Y, weight * Z, weight));
}

// ---- gpu_generate_new_path (integrators/wavefront_path.cu:245-287)
// The free list is also where finished paths meet the film: k_logic only MARKS a finished path (state ST_FREE, slot pushed
// here); its radiance is added to the film by the thread that recycles the slot -- a dense pass over the free list instead
// of a few lanes of every warp of k_logic (RGBFilm::add_sample needs 12 table lookups per path).
__global__ void __launch_bounds__(WF_BLOCK) k_generate(WF wf) {
    const unsigned n = wf.c->p[wf.parity].n_free;
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned n_round = (n + 31u) & ~31u;
    const int lane = threadIdx.x & 31;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool have_slot = i < n;
        // one 64-bit atomic per warp claims consecutive path ids (the reference: one per path)
        const unsigned m = __ballot_sync(0xffffffffu, have_slot);
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&wf.c->next_path, (unsigned long long)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        const long long id = (long long)base + __popc(m & ((1u << lane) - 1u));
        const bool live = have_slot && id < wf.total_paths;
        int slot = -1;
        if (have_slot) {
            slot = wf.q.free_slots[i];
            if (wf.pool.state[slot] == ST_FREE) // write_frame_buffer (:216-234) for the path that finished in this slot
                film_add(wf, wf.pool.ids[slot].x, wf.pool.L[slot], wf.pool.lambda[slot], wf.pool.lpdf[slot], wf.pool.misc[slot].x);
            if (!live) wf.pool.state[slot] = ST_IDLE; // no paths left: the slot retires
        }
        if (live) {
            const int pixel = int(id % wf.n_pixels);
            const int sample = wf.sample_begin + int(id / wf.n_pixels);
            const int px = pixel % wf.film.width, py = pixel / wf.film.width;
            Sampler s;
            s.start(wf.sampler_kind, wf.spp, wf.sqrt_spp, pixel, sample, 0);
            // Sampler::get_camera_sample (base/sampler.cu:167-171)
            float2 fp;
            float fw;
            filter_sample(wf.filter, s.get_2d(), &fp, &fw);
            const float fx = float(px) + fp.x + 0.5f, fy = float(py) + fp.y + 0.5f;
            S4 lambda, lpdf;
            sample_visible(s.get_1d(), &lambda, &lpdf);
            // PerspectiveCamera::generate_ray (cameras/perspective.cu:65-86)
            const float *M = wf.camera.cfr;
            V3 pc = v3(M[0] * fx + M[1] * fy + M[3], M[4] * fx + M[5] * fy + M[7], M[8] * fx + M[9] * fy + M[11]);
            const float wp = M[12] * fx + M[13] * fy + M[15];
            if (wp != 1.f) pc = pc / wp;
            V3 o = v3(0.f, 0.f, 0.f), d = normalize(pc);
            if (wf.camera.lens_radius > 0.f) {
                const float2 pl = sample_disk_concentric(s.get_2d());
                const float ft = wf.camera.focal_distance / d.z;
                const V3 pfocus = ft * d;
                o = v3(wf.camera.lens_radius * pl.x, wf.camera.lens_radius * pl.y, 0.f);
                d = normalize(pfocus - o);
            }
            // Transform::operator()(Ray) (euclidean_space/transform.cu:19-33): origin pushed to the
            // edge of its transform error bound
            const float *R = wf.camera.rfc;
            const V3 ow = xf_point(R, o);
            const V3 oerr = xf_point_error(R, o, v3(0.f, 0.f, 0.f), true);
            const V3 dw = xf_vector(R, d);
            V3 orig = ow;
            const float l2 = len2(dw);
            if (l2 > 0.f) {
                const float dt = dot(vabs(dw), oerr) / l2;
                orig = ow + dt * dw;
            }
            wf.pool.ray_o[slot] = make_float4(orig.x, orig.y, orig.z, kInf);
            wf.pool.ray_d[slot] = make_float4(dw.x, dw.y, dw.z, 0.f);
            wf.pool.beta[slot] = s4(1.f);
            wf.pool.L[slot] = s4(0.f);
            wf.pool.lambda[slot] = lambda;
            wf.pool.lpdf[slot] = lpdf;
            wf.pool.misc[slot] = make_float4(fw, 0.f, 0.f, 0.f);
            wf.pool.rng[slot] = s.state;
            // depth 0, next_time_russian_roulette = 8 (util/russian_roulette.h:5)
            wf.pool.ids[slot] = make_int4(pixel, sample, s.dimension, 8 << 16);
            if (wf.has_media) { // PathState::reset_path (:508-518); camera rays start in the camera's medium (cameras/perspective.cu:85)
                wf.pool.mtp[slot] = s4(1.f);
                wf.pool.vol[slot] = make_float4(0.f, __int_as_float(wf.camera_medium), 0.f, 0.f);
            }
            wf.pool.state[slot] = ST_RAY;
        }
    }
}

// geometric normal of a leaf-record triangle as Triangle::interaction_from_intersection sets it
// (shapes/triangle.cu:388-394): normalize(dp02 x dp12), flipped by reverse_orientation ^ swaps_handedness
__device__ __forceinline__ V3 tri_normal(V3 p0, V3 p1, V3 p2, bool flip) {
    const V3 n = normalize(cross(p0 - p2, p1 - p2));
    return flip ? -n : n;
}

// what the shade stage needs of SurfaceInteraction (base/interaction.h:96-164)
struct Isect {
    V3 p, err;  // pi = p +- err
    V3 n, ns;   // geometric / shading normal
    V3 dpdu;    // shading.dpdu
    V3 wo;
    float u, v; // surface parameterisation at the hit (texture lookups)
};

// Triangle::interaction_from_intersection (shapes/triangle.cu:325-465) from the 48-byte leaf record; per-vertex normals and
// uvs (flag bits of v1.w) are fetched through the primitive id and the index buffer.  Default uv (0,0),(1,0),(1,1).
__device__ __forceinline__ void triangle_interaction(const WF &wf, V3 p0, V3 p1, V3 p2, int prim, uint32_t mw, float b1, float b2,
                                                     V3 wo, Isect &si) {
    const float b0 = 1.f - b1 - b2;
    const bool flip = (mw >> 31) != 0u, has_n = (mw & 0x20000000u) != 0u, has_uv = (mw & 0x40000000u) != 0u;
    int i0 = 0, i1 = 0, i2 = 0;
    if (has_n || has_uv) {
        i0 = wf.idx[3 * (size_t)prim];
        i1 = wf.idx[3 * (size_t)prim + 1];
        i2 = wf.idx[3 * (size_t)prim + 2];
    }
    float2 uv0 = make_float2(0.f, 0.f), uv1 = make_float2(1.f, 0.f), uv2 = make_float2(1.f, 1.f);
    if (has_uv) {
        uv0 = make_float2(wf.uv[2 * (size_t)i0], wf.uv[2 * (size_t)i0 + 1]);
        uv1 = make_float2(wf.uv[2 * (size_t)i1], wf.uv[2 * (size_t)i1 + 1]);
        uv2 = make_float2(wf.uv[2 * (size_t)i2], wf.uv[2 * (size_t)i2 + 1]);
    }
    const float du02 = uv0.x - uv2.x, dv02 = uv0.y - uv2.y, du12 = uv1.x - uv2.x, dv12 = uv1.y - uv2.y;
    const V3 dp02 = p0 - p2, dp12 = p1 - p2;
    const float det = diff_of_products(du02, dv12, dv02, du12);
    const bool degenerate = fabsf(det) < 1e-9f;
    V3 dpdu = v3(0.f, 0.f, 0.f), dpdv = dpdu;
    if (!degenerate) {
        const float invdet = 1.f / det;
        dpdu = v3(diff_of_products(dv12, dp02.x, dv02, dp12.x), diff_of_products(dv12, dp02.y, dv02, dp12.y),
                  diff_of_products(dv12, dp02.z, dv02, dp12.z)) * invdet;
        dpdv = v3(diff_of_products(du02, dp12.x, du12, dp02.x), diff_of_products(du02, dp12.y, du12, dp02.y),
                  diff_of_products(du02, dp12.z, du12, dp02.z)) * invdet;
    }
    if (degenerate || len2(cross(dpdu, dpdv)) == 0.f) {
        const V3 ng = cross(p2 - p0, p1 - p0);
        coordinate_system(normalize(ng), &dpdu, &dpdv);
    }
    si.p = b0 * p0 + b1 * p1 + b2 * p2;
    si.err = gamma_n(7) * (vabs(b0 * p0) + vabs(b1 * p1) + vabs(b2 * p2));
    si.u = b0 * uv0.x + b1 * uv1.x + b2 * uv2.x; // uvHit (shapes/triangle.cu:372)
    si.v = b0 * uv0.y + b1 * uv1.y + b2 * uv2.y;
    V3 n = normalize(cross(dp02, dp12));
    if (flip) n = -n;
    si.n = n;
    si.ns = n;
    si.dpdu = dpdu;
    si.wo = normalize(wo);
    if (has_n) {
        // shading normal + tangent (:388-463), then set_shading_geometry(.., orientationIsAuthoritative = true)
        // (base/interaction.cu:67-89): the geometric normal is flipped to the side of the shading normal
        const V3 n0 = v3(wf.N[3 * (size_t)i0], wf.N[3 * (size_t)i0 + 1], wf.N[3 * (size_t)i0 + 2]);
        const V3 n1 = v3(wf.N[3 * (size_t)i1], wf.N[3 * (size_t)i1 + 1], wf.N[3 * (size_t)i1 + 2]);
        const V3 n2 = v3(wf.N[3 * (size_t)i2], wf.N[3 * (size_t)i2 + 1], wf.N[3 * (size_t)i2 + 2]);
        V3 ns = b0 * n0 + b1 * n1 + b2 * n2;
        ns = len2(ns) > 0.f ? normalize(ns) : n;
        V3 ss = dpdu;
        V3 ts = cross(ns, ss);
        if (len2(ts) > 0.f) ss = cross(ts, ns);
        else coordinate_system(ns, &ss, &ts);
        si.ns = ns;
        si.n = face_forward(n, ns);
        while (len2(ss) > 1e16f || len2(ts) > 1e16f) {
            ss = ss / 1e8f;
            ts = ts / 1e8f;
        }
        si.dpdu = ss;
    }
}

// Sphere::interaction_from_intersection (shapes/sphere.h:71-125) + Transform::operator()(SurfaceInteraction)
// (euclidean_space/transform.cu:57-97)
// Disk::interaction_from_intersection (shapes/disk.cu:133-153): the hit point is recomputed as disk_intersect does it
// Transform::operator()(SurfaceInteraction) (euclidean_space/transform.cu:56-97): the interaction found in an object instance's
// space, brought to render space as TransformedPrimitive::intersect does (primitives/transformed_primitive.cu:31-33).  pi goes
// through plain interval arithmetic (the generic Point3<Interval> operator), normals through the inverse transpose, and the two
// normals are re-faced towards each other at the end.
__device__ __forceinline__ void instance_to_render(const InstRec &I, Isect &si) {
    Iv pi[3];
    xf_point_iv(I.rfi, iv_err(si.p.x, si.err.x), iv_err(si.p.y, si.err.y), iv_err(si.p.z, si.err.z), pi);
    si.p = iv_mid3(pi);
    si.err = iv_err3(pi);
    V3 n = normalize(xf_normal(I.ifr, si.n));
    V3 ns = normalize(xf_normal(I.ifr, si.ns));
    si.wo = normalize(xf_vector(I.rfi, si.wo));
    si.dpdu = xf_vector(I.rfi, si.dpdu);
    n = face_forward(n, ns);
    ns = face_forward(ns, n);
    si.n = n;
    si.ns = ns;
}

// the hit of slot `slot` on leaf-record triangle `prim`: interaction in render space, through the hit's instance if it has one
__device__ __forceinline__ void hit_triangle_interaction(const WF &wf, int slot, const float4 *tp, float4 hit, V3 ro, V3 rd, Isect &si) {
    const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
    const int inst = wf.n_instances > 0 ? wf.pool.inst[slot] : -1;
    if (inst < 0) {
        triangle_interaction(wf, v3(v0), v3(v1), v3(v2), __float_as_int(v0.w), __float_as_uint(v1.w), hit.z, hit.w, -rd, si);
        return;
    }
    const InstRec &I = wf.instances[inst];
    V3 oi, di;
    float dt;
    instance_ray(I.ifr, ro, rd, &oi, &di, &dt);
    triangle_interaction(wf, v3(v0), v3(v1), v3(v2), __float_as_int(v0.w), __float_as_uint(v1.w), hit.z, hit.w, -di, si);
    instance_to_render(I, si);
}

__device__ __forceinline__ void disk_interaction(const SphereRec &s, V3 ro, V3 rd, float t, V3 wo, Isect &si) {
    Iv oi[3];
    xf_point_iv(s.ofr, iv(ro.x), iv(ro.y), iv(ro.z), oi);
    const V3 dc = xf_vector(s.ofr, rd), de = xf_vector_error(s.ofr, rd);
    const V3 o = iv_mid3(oi);
    const V3 d = v3(iv_mid(iv_err(dc.x, de.x)), iv_mid(iv_err(dc.y, de.y)), iv_mid(iv_err(dc.z, de.z)));
    V3 pHit = o + t * d;
    const float rHit = sqrtf(sqr(pHit.x) + sqr(pHit.y));
    const V3 dpdu = v3(-s.phi_max * pHit.y, s.phi_max * pHit.x, 0.f);
    const V3 dpdv = v3(pHit.x, pHit.y, 0.f) * (s.z_max - s.radius) / rHit;
    {
        float phi = atan2f(pHit.y, pHit.x);
        if (phi < 0.f) phi += 2.f * kPi;
        si.u = phi / s.phi_max;
        si.v = (s.radius - rHit) / (s.radius - s.z_max);
    }
    pHit.z = s.z_min;
    V3 n = normalize(cross(dpdu, dpdv));
    if (s.flip) n = -n;
    Iv pi[3]; // Point3fi(pHit, pError = 0) through the interval transform
    xf_point_iv(s.rfo, iv(pHit.x), iv(pHit.y), iv(pHit.z), pi);
    si.p = iv_mid3(pi);
    si.err = iv_err3(pi);
    const V3 nr = normalize(xf_normal(s.ofr, n));
    si.n = nr;
    si.ns = nr;
    si.dpdu = xf_vector(s.rfo, dpdu);
    si.wo = normalize(wo);
}

__device__ __forceinline__ void sphere_interaction(const SphereRec &s, V3 ro, V3 rd, float t, V3 wo, Isect &si) {
    if (s.kind == 1) {
        disk_interaction(s, ro, rd, t, wo, si);
        return;
    }
    // recompute the refined object-space hit point exactly as basic_intersect does
    const V3 o = xf_point(s.ofr, ro), d = xf_vector(s.ofr, rd);
    V3 pHit = o + t * d;
    pHit = pHit * (s.radius / len(pHit));
    if (pHit.x == 0.f && pHit.y == 0.f) pHit.x = 1e-5f * s.radius;
    const float cosTheta = pHit.z / s.radius;
    {
        float phi = atan2f(pHit.y, pHit.x);
        if (phi < 0.f) phi += 2.f * kPi;
        si.u = phi / s.phi_max;
        si.v = (acosf(clampf(cosTheta, -1.f, 1.f)) - s.theta_z_min) / (s.theta_z_max - s.theta_z_min);
    }
    const float zRadius = sqrtf(sqr(pHit.x) + sqr(pHit.y));
    const float cosPhi = pHit.x / zRadius, sinPhi = pHit.y / zRadius;
    const V3 dpdu = v3(-s.phi_max * pHit.y, s.phi_max * pHit.x, 0.f);
    const float sinTheta = safe_sqrt(1.f - sqr(cosTheta));
    const V3 dpdv = (s.theta_z_max - s.theta_z_min) * v3(pHit.z * cosPhi, pHit.z * sinPhi, -s.radius * sinTheta);
    const V3 perr = gamma_n(5) * vabs(pHit);
    V3 n = normalize(cross(dpdu, dpdv));
    if (s.flip) n = -n;
    // to render space
    si.p = xf_point(s.rfo, pHit);
    si.err = xf_point_error(s.rfo, pHit, perr, false);
    V3 nr = normalize(xf_normal(s.ofr, n));
    V3 nsr = nr;
    nr = face_forward(nr, nsr);
    si.n = nr;
    si.ns = nsr;
    si.dpdu = xf_vector(s.rfo, dpdu);
    si.wo = normalize(wo);
}

// SurfaceInteraction::set_intersection_properties (base/interaction.cu:41-64): a mix material is resolved at hit time by
// u = hash_float(pi, wo) < amount (materials/mix_material.cu:18-21); pi is the Point3fi interval of the hit point
// (euclidean_space/point3fi.h:31-34, interval.h:23-33), wo the normalised outgoing direction -- MurmurHash64A over 36 bytes
__device__ __forceinline__ float mix_hash_float(V3 p, V3 err, V3 wo) {
    float w[9];
    const float pv[3] = {p.x, p.y, p.z}, ev[3] = {err.x, err.y, err.z};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        w[2 * k] = ev[k] == 0.f ? pv[k] : next_down(pv[k] - ev[k]);
        w[2 * k + 1] = ev[k] == 0.f ? pv[k] : next_up(pv[k] + ev[k]);
    }
    w[6] = wo.x; w[7] = wo.y; w[8] = wo.z;
    const uint64_t m = 0xc6a4a7935bd1e995ULL;
    const int r = 47;
    uint64_t h = 0ull ^ (36ull * m);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint64_t kk = (uint64_t)__float_as_uint(w[2 * k]) | ((uint64_t)__float_as_uint(w[2 * k + 1]) << 32);
        kk *= m;
        kk ^= kk >> r;
        kk *= m;
        h ^= kk;
        h *= m;
    }
    h ^= (uint64_t)__float_as_uint(w[8]);
    h *= m;
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return __fmul_rn((float)(uint32_t)h, 0x1p-32f);
}

// ImageInfiniteLight::sample_li (lights/image_infinite_light.cu:84-103): the far point sits 2 R along the (un-normalised) wi
__device__ __forceinline__ bool env_nee(const WF &wf, V3 p, float2 u, S4 lambda, LightSample &ls, S4 &Le) {
    V3 wi;
    float pdf;
    env_sample(wf.env, wf.tex, wf.spectra, u, lambda, &wi, &pdf, &Le);
    ls.wi = wi;
    ls.pdf = pdf;
    ls.p = p + wi * (2.f * wf.scene_radius);
    ls.err = v3(0.f, 0.f, 0.f);
    ls.n = v3(0.f, 0.f, 0.f);
    return true;
}

// escaping ray against the image infinite light: Le and pdf_li (only in the ENV instantiations of k_logic)
__device__ __forceinline__ void env_escape(const WF &wf, V3 d, S4 lambda, S4 *Le, float *pdf) {
    *Le = env_le(wf.env, wf.tex, wf.spectra, d, lambda);
    *pdf = env_pdf(wf.env, d);
}

// ---- media ------------------------------------------------------------------------------------------------------------
// DenselySampledSpectrum(spectrum, scale)::sample (spectra/densely_sampled_spectrum.h:10-14,46-66): the medium's sigma
// spectra are tabulated at integer nanometres as spectrum(l) * scale and looked up with a lerp; evaluated here on the fly
__device__ __forceinline__ float medium_sigma_at(const SpectrumTable &tab, int id, float scale, float lambda) {
    const float fl = floorf(lambda), ce = ceilf(lambda);
    if (fl < 360.f || ce > 830.f) return 0.f;
    return lerpf(lambda - fl, spectrum_at(tab, id, fl) * scale, spectrum_at(tab, id, ce) * scale);
}
__device__ __forceinline__ S4 medium_sigma(const SpectrumTable &tab, int id, float scale, S4 l) {
    return make_float4(medium_sigma_at(tab, id, scale, l.x), medium_sigma_at(tab, id, scale, l.y), medium_sigma_at(tab, id, scale, l.z),
                       medium_sigma_at(tab, id, scale, l.w));
}
// Medium::sample_sigma_t (medium/homogeneous_medium.cu:43-47)
__device__ __forceinline__ S4 medium_sigma_t(const WF &wf, int medium, S4 lambda) {
    const MediumRec M = wf.media[medium];
    return medium_sigma(wf.spectra, M.sigma_a, M.scale, lambda) + medium_sigma(wf.spectra, M.sigma_s, M.scale, lambda);
}
__device__ __forceinline__ S4 exp_s4(S4 a) { return make_float4(expf(a.x), expf(a.y), expf(a.z), expf(a.w)); }

// Interaction::get_medium(w) (base/interaction.h:60-66) for a hit on triangle `prim_id` (or a quadric, see below): a medium transition picks exterior / interior by the side of the GEOMETRIC normal, a SimplePrimitive
// hands on the medium of the ray that hit it, a GeometricPrimitive without transition has no medium at all
// (primitives/geometric_primitive.cu:27, base/interaction.cu:59-63)
__device__ __forceinline__ int surface_medium(const WF &wf, int prim_id, V3 n, V3 w, int ray_medium) {
    // prim_id <= -2: quadric number -2 - prim_id, which carries its own medium word
    if (prim_id == -1) return ray_medium;
    const uint32_t code = prim_id < 0 ? wf.spheres[-2 - prim_id].medium : wf.tri_medium[prim_id];
    if (code == 0u) return ray_medium;
    if (code == 0xffffffffu) return -1;
    return dot(n, w) > 0.f ? int(code >> 16) - 1 : int(code & 0xffffu) - 1;
}

constexpr int VOL_BLOCK = 128;

// IntegratorBase::compute_transmittance (base/integrator_base.cu:22-71) from a point INSIDE a medium (no normal, no error
// bounds) to the light sample: march the normalised shadow ray from hit to hit, attenuating by the medium of each segment;
// any surface with a material blocks, material-less interfaces are crossed (normal := ray direction, so the far side is
// the interface's exterior, or whatever a plain primitive hands on).
// Does any instanced object block the ray within t_max?  Per-thread form of k_trace_wf's instance walk for the transmittance
// walk below: the tree over the instance boxes is walked with a small local stack, each proxy whose own box the ray touches takes
// the ray into the instance's space (Transform::apply_inverse(Ray, &t_max)) and walks the object's tree for ANY hit.  Instanced
// surfaces always carry a material (the reference refuses medium transitions inside objects, scene_builder.cu:651-656), so any
// hit ends a transmittance walk with zero.
__device__ __forceinline__ bool instances_block(const WF &wf, V3 o, V3 d, float t_max, uint2 *sm_stack, unsigned *overflow) {
    if (wf.inst_root < 0) return false;
    const float idx = safe_rcp_dir(d.x), idy = safe_rcp_dir(d.y), idz = safe_rcp_dir(d.z);
    const uint32_t octinv = (idx < 0.f ? 0u : 4u) | (idy < 0.f ? 0u : 2u) | (idz < 0.f ? 0u : 1u);
    const uint32_t octinv4 = octinv * 0x01010101u;
    constexpr int DEPTH = 24;
    uint2 stack[DEPTH];
    int sp = 0;
    uint2 ng = make_uint2((uint32_t)wf.inst_root, 0x80000000u), tg = make_uint2(0u, 0u);
    while (true) {
        if (tg.y != 0u) {
            const int ti = __ffs(tg.y) - 1;
            tg.y &= tg.y - 1u;
            const int ni = __float_as_int(__ldg(&wf.bvh.tris[3 * size_t(tg.x + ti)].w)) - int(wf.bvh.n_tris);
            const InstRec &I = wf.instances[ni];
            if (ray_touches_instance(I, o.x, o.y, o.z, idx, idy, idz, t_max)) {
                V3 oi, di;
                float dt;
                instance_ray(I.ifr, o, d, &oi, &di, &dt);
                RayHit h;
                TraceCounters ctr;
                traverse_bvh8<true, false, true>(wf.bvh, oi.x, oi.y, oi.z, di.x, di.y, di.z, t_max - dt, sm_stack, VOL_BLOCK, h, ctr, overflow,
                                                 I.root_node);
                if (h.prim >= 0) return true;
            }
            continue;
        }
        if (ng.y <= 0x00ffffffu) {
            if (sp == 0) return false;
            ng = stack[--sp];
        }
        const uint32_t hits = ng.y;
        const int child_bit = 31 - __clz(hits);
        ng.y &= ~(1u << child_bit);
        if (ng.y > 0x00ffffffu) {
            if (sp < DEPTH) stack[sp++] = ng;
            else *overflow = 1u;
        }
        const uint32_t cslot = (uint32_t(child_bit) - 24u) ^ octinv;
        const uint32_t rel = __popc(hits & 0xffu & ((1u << cslot) - 1u));
        const float4 *np = wf.bvh.nodes + 5 * size_t(ng.x + rel);
        const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
        const uint32_t hitmask = intersect_node8(n0, n1, n2, n3, n4, o.x, o.y, o.z, idx, idy, idz, octinv4, t_max, 0x47000000u);
        ng.x = __float_as_uint(n1.x);
        tg.x = __float_as_uint(n1.y);
        ng.y = (hitmask & 0xff000000u) | (__float_as_uint(n0.w) >> 24);
        tg.y = hitmask & 0x00ffffffu;
    }
}

// p0: the scattering point (offset_ray_origin with n = 0 moves nothing, base/ray.h:30-49); pt: the light point offset towards
// p0 (it fixes the direction of the walk); lp: the light point itself (the distances).
template <bool INST>
__device__ __forceinline__ S4 volume_transmittance(const WF &wf, V3 p0, int medium, V3 pt, V3 lp, S4 lambda, uint2 *sm_stack,
                                                   unsigned *overflow) {
    const float epsilon_distance = len(p0 - lp) * 1e-4f;
    V3 o = p0;
    const V3 d = normalize(pt - p0);
    bool previous_self_intersection = false;
    S4 T = s4(1.f);
    for (int depth = 0; depth < 128; depth++) { // MAX_VOLUME_BOUNCES (base/integrator_base.h:21)
        const float distance_to_light = len(lp - o);
        const float t_max = (float)((1.0 - (double)1e-4f) * (double)distance_to_light);
        RayHit h;
        TraceCounters ctr;
        traverse_bvh8<false, false, true>(wf.bvh, o.x, o.y, o.z, d.x, d.y, d.z, t_max, sm_stack, VOL_BLOCK, h, ctr, overflow);
        bool sphere_closer = false;
        int sphere_k = -1;
        float sphere_t = h.t;
        for (int k = 0; k < wf.n_spheres; k++) { // the closest quadric in front of the triangle hit
            SphereHit sh;
            if (sphere_intersect(wf.spheres[k], o, d, sphere_t, sh)) {
                sphere_closer = true;
                sphere_k = k;
                sphere_t = sh.t;
            }
        }
        // a quadric with a material in front of the triangle hit ends the walk with 0 whatever its exact distance; a material-less
        // quadric is crossed like an interface triangle
        bool cross_sphere = sphere_k >= 0 && wf.materials[wf.spheres[sphere_k].material].type == B200PT_MAT_INTERFACE;
        // an instanced object in front of whatever was found (and would be crossed or is the end of the segment) ends the walk too
        if (INST && (!sphere_closer || cross_sphere) && instances_block(wf, o, d, sphere_closer ? sphere_t : h.t, sm_stack, overflow)) {
            sphere_closer = true;
            cross_sphere = false;
        }
        const bool hit = sphere_closer || h.prim >= 0;
        const float next_t = hit ? (cross_sphere ? sphere_t : h.t) : distance_to_light;
        if (medium >= 0) T = T * exp_s4(medium_sigma_t(wf, medium, lambda) * (-next_t));
        if (!hit) return T;
        if (cross_sphere) {
            Isect si;
            sphere_interaction(wf.spheres[sphere_k], o, d, sphere_t, -d, si);
            o = offset_ray_origin(si.p, si.err, d, d);
            medium = surface_medium(wf, -2 - sphere_k, d, d, medium);
            if (sphere_t < epsilon_distance) {
                if (previous_self_intersection) o = o + epsilon_distance * d;
                previous_self_intersection = true;
            } else {
                previous_self_intersection = false;
            }
            continue;
        }
        if (sphere_closer) return s4(0.f);
        const float4 *tp = wf.bvh.tris + 3 * (size_t)h.prim;
        const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
        const uint32_t mw = __float_as_uint(v1.w);
        if (wf.materials[mw & kTriMaterialMask].type != B200PT_MAT_INTERFACE) return s4(0.f);
        // interaction.n = d; shadow_ray = interaction.spawn_ray(d)
        const float b1 = h.b1, b2 = h.b2, b0 = 1.f - b1 - b2;
        const V3 q0 = v3(v0), q1 = v3(v1), q2 = v3(v2);
        const V3 p = b0 * q0 + b1 * q1 + b2 * q2;
        const V3 err = gamma_n(7) * (vabs(b0 * q0) + vabs(b1 * q1) + vabs(b2 * q2));
        o = offset_ray_origin(p, err, d, d);
        medium = surface_medium(wf, __float_as_int(v0.w), d, d, medium);
        if (h.t < epsilon_distance) {
            if (previous_self_intersection) o = o + epsilon_distance * d; // forced offset against self-intersection
            previous_self_intersection = true;
        } else {
            previous_self_intersection = false;
        }
    }
    return s4(0.f); // the reference aborts here (REPORT_FATAL_ERROR)
}

// ---- volume_scatter (integrators/wavefront_path.cu:368-436) + the volume branch of sample_Ld_volume
// (integrators/megakernel_path.cu:185-264) -- runs between the closest-hit trace and the logic stage, over every traced slot:
// ray_cast clears in_volume for all of them (:361), rays inside a medium then either pass through to their hit
// (multi_transmittance_pdf *= exp(-sigma_t t_max)) or scatter: the origin moves to the scattering point, direct light is
// gathered THROUGH interfaces with the transmittance walk above, and a new direction is drawn from the phase function.
// Three launches: k_volume<0> visits every traced slot, samples the free-flight distance and either lets the path pass or
// queues it; k_volume<1> runs the scattering work densely over that queue instead of in whichever lanes of a warp happened to
// scatter, and leaves the NEE sample of each scattering point as a transmittance REQUEST (position i of the shadow-queue
// arrays, which are empty at this point of a pass, plus two lanes of their own); k_transmit walks those segments from
// interface to interface and adds the surviving light.  Round 1 ran the walk inline: 163 registers, 17 % of the warp slots.
template <int PHASE>
__global__ void __launch_bounds__(VOL_BLOCK) k_volume(WF wf) {
    if (wf.c->finished >= (unsigned long long)wf.total_paths) return;
    const unsigned n = PHASE == 0 ? (unsigned)wf.pool_size : wf.c->p[wf.parity].n_vol;
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned n_round = (n + 31u) & ~31u;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        if (PHASE == 0) {
            bool scatter = false;
            if (i < n && wf.pool.state[i] == ST_RAY) {
                const int slot = (int)i;
                int4 ids = wf.pool.ids[slot];
                const bool was_in_volume = (ids.w & PF_IN_VOLUME) != 0;
                ids.w &= ~PF_IN_VOLUME;
                float4 vol = wf.pool.vol[slot];
                const int medium = __float_as_int(vol.y);
                const float4 ro4 = wf.pool.ray_o[slot];
                if (medium < 0 || ro4.w < 0.f) {
                    if (was_in_volume) wf.pool.ids[slot] = ids;
                } else {
                    const S4 lambda = wf.pool.lambda[slot];
                    const S4 sigma_t = medium_sigma_t(wf, medium, lambda);
                    const float4 hit = wf.pool.hit[slot];
                    const float t_max = __float_as_int(hit.x) >= 0 ? hit.y : kInf;
                    Sampler s = load_sampler(wf, slot, ids);
                    const float t = -logf(1.f - s.get_1d()) / average(sigma_t); // sample_exponential (util/sampling.h:67-69)
                    if (t >= t_max) {
                        // pass through the medium
                        wf.pool.mtp[slot] = wf.pool.mtp[slot] * exp_s4(sigma_t * (-t_max));
                    } else {
                        scatter = true;
                        vol.z = t;
                        wf.pool.vol[slot] = vol;
                    }
                    ids.z = s.dimension;
                    wf.pool.ids[slot] = ids;
                    wf.pool.rng[slot] = s.state;
                }
            }
            queue_push(&wf.c->p[wf.parity].n_vol, wf.q.vol, (int)i, scatter);
            continue;
        }
        if (i >= n) continue;
        const int slot = wf.q.vol[i];
        int4 ids = wf.pool.ids[slot];
        float4 vol = wf.pool.vol[slot];
        const int medium = __float_as_int(vol.y);
        const float t = vol.z;
        const V3 ro = v3(wf.pool.ray_o[slot]), rd = v3(wf.pool.ray_d[slot]);
        const S4 lambda = wf.pool.lambda[slot];
        const MediumRec M = wf.media[medium];
        const S4 sigma_s = medium_sigma(wf.spectra, M.sigma_s, M.scale, lambda);
        const S4 sigma_t = medium_sigma(wf.spectra, M.sigma_a, M.scale, lambda) + sigma_s;
        Sampler s = load_sampler(wf, slot, ids);
        bool requested = false;
        {
            const V3 o = ro + rd * t;
            S4 beta = wf.pool.beta[slot] * (sigma_s / sigma_t);
            // -- sample_Ld_volume(medium_interaction, nullptr, ..): ctx = the bare point
            const LightCtx ctx{o, v3(0.f, 0.f, 0.f), v3(0.f, 0.f, 0.f), v3(0.f, 0.f, 0.f)};
            const float ul = s.get_1d();
            const float2 up = s.get_2d();
            if (wf.n_lights > 0) {
                const int li = alias_sample(wf.alias, wf.n_lights, ul);
                const LightRec Lr = wf.lights[li];
                const int lt = light_type(Lr);
                LightSample ls;
                bool have, delta = false;
                float lw = 1.f, ld2 = 1.f;
                S4 Le_env = s4(0.f);
                const bool is_env = lt == B200PT_LIGHT_IMAGE_INFINITE;
                if (lt == B200PT_LIGHT_AREA) have = light_sample(Lr, ctx, up, ls) && light_emits_towards(Lr, ls.n, -ls.wi);
                    else if (lt == B200PT_LIGHT_SPHERE_AREA) have = sphere_light_sample(wf.spheres[Lr.ext], Lr, ctx, up, ls) && light_emits_towards(Lr, ls.n, -ls.wi);
                else if (is_env) have = env_nee(wf, ctx.p, up, lambda, ls, Le_env);
                else have = light_sample_other(Lr, lt, wf.light_ext, wf.scene_radius, ctx, up, ls, &lw, &ld2, &delta);
                if (have) {
                    S4 Le;
                    if (is_env) Le = Le_env;
                    else if (lt == B200PT_LIGHT_SPOT) Le = spectrum_sample(wf.spectra, Lr.spectrum, lambda) * (lw * Lr.scale) / ld2;
                    else Le = spectrum_sample(wf.spectra, Lr.spectrum, lambda) * Lr.scale;
                    if (is_positive(Le) && ls.pdf != 0.f) {
                        const float pdf_light = Lr.pmf * ls.pdf;
                        // the phase function is SAMPLED here, not evaluated towards the light (megakernel_path.cu:247-263)
                        float ppdf;
                        (void)sample_henyey_greenstein(-rd, M.g, s.get_2d(), &ppdf);
                        const V3 pt = offset_ray_origin(ls.p, ls.err, ls.n, o - ls.p);
                        wf.q.sh_o[i] = make_float4(o.x, o.y, o.z, __int_as_float(medium));
                        wf.q.sh_d[i] = make_float4(pt.x, pt.y, pt.z, pdf_light);
                        wf.q.sh_contrib[i] = make_float4(ls.p.x, ls.p.y, ls.p.z, delta ? -ppdf : ppdf);
                        wf.q.tr_a[i] = Le / pdf_light;
                        wf.q.tr_beta[i] = beta;
                        requested = true;
                    }
                }
            }
            float ppdf;
            const V3 wi = sample_henyey_greenstein(-rd, M.g, s.get_2d(), &ppdf);
            beta = beta * (ppdf / ppdf); // rho / pdf
            wf.pool.beta[slot] = beta;
            float4 misc = wf.pool.misc[slot];
            misc.y = ppdf;
            wf.pool.misc[slot] = misc;
            wf.pool.ctx0[slot] = make_float4(o.x, o.y, o.z, 0.f);
            wf.pool.ctx1[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
            wf.pool.ctx2[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
            wf.pool.ray_o[slot] = make_float4(o.x, o.y, o.z, kInf);
            wf.pool.ray_d[slot] = make_float4(wi.x, wi.y, wi.z, 0.f);
            wf.pool.mtp[slot] = s4(1.f);
            vol.x += 1.f;
            wf.pool.vol[slot] = vol;
            // not re-traced at once: the logic stage first applies russian roulette and the depth limit (:431-432)
            ids.w = (ids.w & ~PF_SPECULAR) | PF_ANY_NON_SPECULAR | PF_IN_VOLUME;
        }
        if (!requested) wf.q.sh_o[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(-2)); // medium -2: no request at this position
        ids.z = s.dimension;
        wf.pool.ids[slot] = ids;
        wf.pool.rng[slot] = s.state;
    }
}

// the transmittance requests of k_volume<1>, one thread per scattering point of this pass
template <bool INST>
__global__ void __launch_bounds__(VOL_BLOCK) k_transmit(WF wf) {
    __shared__ uint2 s_stack[SM_STACK * VOL_BLOCK];
    if (wf.c->finished >= (unsigned long long)wf.total_paths) return;
    const unsigned n = wf.c->p[wf.parity].n_vol;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 r0 = wf.q.sh_o[i];
        const int medium = __float_as_int(r0.w);
        if (medium < -1) continue; // no light sample at this scattering point
        const float4 r1 = wf.q.sh_d[i], r2 = wf.q.sh_contrib[i];
        const int slot = wf.q.vol[i];
        const S4 lambda = wf.pool.lambda[slot];
        const S4 T = volume_transmittance<INST>(wf, v3(r0), medium, v3(r1), v3(r2), lambda, s_stack + threadIdx.x, &wf.c->stack_overflow);
        const float pdf_light = r1.w;
        const bool delta = signbit(r2.w); // the phase pdf is positive: its sign carries "delta light, no MIS"
        const float ppdf = fabsf(r2.w);
        const S4 contribution = wf.q.tr_a[i] * T;
        S4 Ld;
        if (delta) Ld = contribution * ppdf;
        else Ld = contribution * (power_heuristic(pdf_light, ppdf * average(T)) * ppdf);
        wf.pool.L[slot] = wf.pool.L[slot] + wf.q.tr_beta[i] * Ld;
    }
}

// ---- control_logic (integrators/wavefront_path.cu:68-214)
template <bool MEDIA, bool ENV>
__global__ void __launch_bounds__(WF_BLOCK, 4) k_logic(WF wf) {
    // one thread per pool slot, in slot order: every lane access below is a coalesced 16-byte gather
    if (blockIdx.x == 0 && threadIdx.x == 0) wf.c->p[wf.parity ^ 1] = WFPass{}; // the next pass starts with empty queues
    if (wf.c->finished >= (unsigned long long)wf.total_paths) return;
    const unsigned n = (unsigned)wf.pool_size;
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned n_round = (n + WF_BLOCK - 1u) / WF_BLOCK * WF_BLOCK;
    unsigned n_finished = 0;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool act = i < n && wf.pool.state[i] == ST_RAY;
        const int slot = (int)i;
        int route = -1, route_prim = -1, route_mat = 0; // route: material type, or -1; the hit primitive and its (resolved) material
        bool terminate = false, stay = false; // stay: the slot keeps ST_RAY (scattered in a medium / crossed an interface)
        int emit = 0; // 1: the hit triangle is emissive (k_emit adds its Le), | 2: added raw (camera ray / specular bounce)
        if (act) {
            int4 ids = wf.pool.ids[slot];
            int next_rr = (ids.w >> 16) & 0xff;
            // depth is a float in the reference (integrators/wavefront_path.h:40-41) because a material-less interface counts
            // 0.3 of a bounce; without media it only ever holds small integers and lives in the flags word
            float4 vol = make_float4(0.f, 0.f, 0.f, 0.f);
            if (MEDIA) vol = wf.pool.vol[slot];
            float depth = MEDIA ? vol.x : float((ids.w >> 8) & 0xff);
            S4 beta = wf.pool.beta[slot];
            S4 L = wf.pool.L[slot];
            const float4 hit = wf.pool.hit[slot];
            const int prim = __float_as_int(hit.x);
            terminate = depth >= wf.max_depth || !is_positive(beta);
            bool dirty_L = false;
            if (!terminate && depth >= next_rr) {
                // russian_roulette (util/russian_roulette.h:12-29)
                next_rr += 1;
                Sampler s = load_sampler(wf, slot, ids);
                const float q = fminf(max_component(beta), 0.95f);
                if (s.get_1d() > q) {
                    beta = s4(0.f);
                    terminate = true;
                } else {
                    beta = beta / q;
                }
                wf.pool.beta[slot] = beta;
                wf.pool.rng[slot] = s.state;
                ids.z = s.dimension;
            }
            // a path that scattered inside a medium this pass (k_volume) has a new ray and a stale hit record: it goes
            // straight back to the trace stage (:99-103).  One that russian roulette just zeroed would be retired by the
            // next logic stage with the same L, so it is retired now.
            const bool in_volume = MEDIA && (ids.w & PF_IN_VOLUME) && !terminate;
            if (in_volume) stay = true;
            const float dir_pdf_scale = MEDIA ? average(wf.pool.mtp[slot]) : 1.f; // multi_transmittance_pdf.average() (:118,165)
            if (!terminate && !stay && prim < 0) {
                // escaped: infinite lights' Le, MIS-weighted against NEE (integrators/wavefront_path.cu:105-127).
                // pdf_li(.., allow_incomplete_pdf = true) of the uniform infinite light is 0
                // (lights/uniform_infinite_light.cu:51-58), so the weight is power_heuristic(dir_pdf, 0)
                for (int k = 0; k < wf.n_infinite; k++) {
                    const LightRec Lr = wf.lights[wf.infinite[k]];
                    S4 Le;
                    float pdf_li = 0.f;
                    if (ENV && light_type(Lr) == B200PT_LIGHT_IMAGE_INFINITE) env_escape(wf, v3(wf.pool.ray_d[slot]), wf.pool.lambda[slot], &Le, &pdf_li);
                    else Le = spectrum_sample(wf.spectra, Lr.spectrum, wf.pool.lambda[slot]) * Lr.scale;
                    if (depth == 0 || (ids.w & PF_SPECULAR)) {
                        L = L + beta * Le;
                    } else {
                        const float pdf_light = Lr.pmf * pdf_li;
                        const float dir_pdf = wf.pool.misc[slot].y * dir_pdf_scale;
                        L = L + beta * power_heuristic(dir_pdf, pdf_light) * Le;
                    }
                    dirty_L = true;
                }
                terminate = true;
            }
            if (!terminate && !stay && MEDIA && prim >= (int)wf.bvh.n_tris) {
                const int k = prim - (int)wf.bvh.n_tris;
                const SphereRec &sp = wf.spheres[k];
                if (wf.materials[sp.material].type == B200PT_MAT_INTERFACE) { // a material-less quadric: as for a triangle below
                    const V3 ro = v3(wf.pool.ray_o[slot]), rd = v3(wf.pool.ray_d[slot]);
                    Isect si;
                    sphere_interaction(sp, ro, rd, hit.y, -rd, si);
                    const V3 o = offset_ray_origin(si.p, si.err, si.n, rd);
                    wf.pool.ray_o[slot] = make_float4(o.x, o.y, o.z, kInf);
                    vol.x = depth + 0.3f;
                    vol.y = __int_as_float(surface_medium(wf, -2 - k, si.n, rd, __float_as_int(vol.y)));
                    wf.pool.vol[slot] = vol;
                    stay = true;
                }
            }
            if (!terminate && !stay && MEDIA && prim < (int)wf.bvh.n_tris) {
                const float4 *tp = wf.bvh.tris + 3 * (size_t)prim;
                const float4 v1 = __ldg(tp + 1);
                const uint32_t mw = __float_as_uint(v1.w);
                if (wf.materials[mw & kTriMaterialMask].type == B200PT_MAT_INTERFACE) {
                    // material-less interface (:146-151, gpu_evaluate_interface_material :303-320): 0.3 of a bounce, the ray
                    // carries on in the same direction from the far side, in the medium found there
                    const float4 v0 = __ldg(tp), v2 = __ldg(tp + 2);
                    const V3 rd = v3(wf.pool.ray_d[slot]);
                    Isect si;
                    triangle_interaction(wf, v3(v0), v3(v1), v3(v2), __float_as_int(v0.w), mw, hit.z, hit.w, -rd, si);
                    const V3 o = offset_ray_origin(si.p, si.err, si.n, rd);
                    wf.pool.ray_o[slot] = make_float4(o.x, o.y, o.z, kInf);
                    vol.x = depth + 0.3f;
                    vol.y = __int_as_float(surface_medium(wf, __float_as_int(v0.w), si.n, rd, __float_as_int(vol.y)));
                    wf.pool.vol[slot] = vol;
                    stay = true;
                }
            }
            if (!terminate && !stay) {
                int mat;
                if (prim >= (int)wf.bvh.n_tris) {
                    mat = wf.spheres[prim - (int)wf.bvh.n_tris].material;
                    if (wf.spheres[prim - (int)wf.bvh.n_tris].light >= 0) emit = 1 | ((depth == 0 || (ids.w & PF_SPECULAR)) ? 2 : 0);
                } else {
                    const float4 *tp = wf.bvh.tris + 3 * (size_t)prim;
                    const float4 v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                    const uint32_t mw = __float_as_uint(v1.w);
                    mat = int(mw & kTriMaterialMask);
                    // emitted radiance (:155-171) is added by k_emit, densely over the few slots that hit a light
                    if (__float_as_int(v2.w) >= 0) emit = 1 | ((depth == 0 || (ids.w & PF_SPECULAR)) ? 2 : 0);
                }
                if (wf.has_mix && wf.materials[mat].type == B200PT_MAT_MIX) {
                    const MatRec mr = wf.materials[mat];
                    const V3 ro = v3(wf.pool.ray_o[slot]), rd = v3(wf.pool.ray_d[slot]);
                    Isect si;
                    if (prim >= (int)wf.bvh.n_tris) {
                        sphere_interaction(wf.spheres[prim - (int)wf.bvh.n_tris], ro, rd, hit.y, -rd, si);
                    } else {
                        hit_triangle_interaction(wf, slot, wf.bvh.tris + 3 * (size_t)prim, hit, ro, rd, si);
                    }
                    mat = mix_hash_float(si.p, si.err, si.wo) < mr.scale ? mr.mix_a : mr.mix_b;
                    // the resolved material travels to the shade stage in the material queue entry
                }
                depth += 1.f;
                if (MEDIA) {
                    vol.x = depth;
                    wf.pool.vol[slot] = vol;
                }
                route = wf.materials[mat].type;
                route_prim = prim;
                route_mat = mat;
            }
            if (dirty_L) wf.pool.L[slot] = L;
            if (terminate) {
                n_finished++; // its radiance goes to the film when k_generate recycles the slot
            } else {
                ids.w = (ids.w & 0xff) | ((MEDIA ? 0 : (int)depth) << 8) | (next_rr << 16);
                wf.pool.ids[slot] = ids;
            }
        }
        const int cat = (act && terminate) ? WF_NMAT : route; // material queue, WF_NMAT = free list, -1 nothing (the slot stays a ray)
        if (act && !stay) wf.pool.state[slot] = terminate ? ST_FREE : uint8_t(ST_MAT0 + route);
        block_push(wf, cat, slot, make_int2(route_prim, route_mat));
        queue_push(&wf.c->p[wf.parity].n_emit, wf.q.emit, slot | ((emit & 2) ? int(0x80000000u) : 0), emit != 0);
    }
    // finished-path count: warp reduce, one atomic per warp
    for (int o = 16; o > 0; o >>= 1) n_finished += __shfl_xor_sync(0xffffffffu, n_finished, o);
    if ((threadIdx.x & 31) == 0 && n_finished) atomicAdd(&wf.c->finished, (unsigned long long)n_finished);
}

// ---- emitted radiance at a hit on an area light: SurfaceInteraction::le -> DiffuseAreaLight::l, MIS against NEE
// (integrators/wavefront_path.cu:155-171).  Runs over the emit queue of the pass, after k_logic (which found the hits)
// and before the shade kernels (which move beta on).
__global__ void __launch_bounds__(WF_BLOCK) k_emit(WF wf) {
    const unsigned n = wf.c->p[wf.parity].n_emit;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int e = wf.q.emit[i];
        const int slot = e & 0x7fffffff;
        const bool raw = e < 0;
        const float4 hit = wf.pool.hit[slot];
        const int prim = __float_as_int(hit.x);
        const V3 rd = v3(wf.pool.ray_d[slot]);
        const bool on_sphere = prim >= (int)wf.bvh.n_tris;
        LightRec Lr;
        V3 n_g;
        if (on_sphere) {
            const SphereRec &sp = wf.spheres[prim - (int)wf.bvh.n_tris];
            Lr = wf.lights[sp.light];
            Isect si;
            sphere_interaction(sp, v3(wf.pool.ray_o[slot]), rd, hit.y, -rd, si);
            n_g = si.n;
        } else {
            const float4 *tp = wf.bvh.tris + 3 * (size_t)prim;
            const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
            Lr = wf.lights[__float_as_int(v2.w)];
            n_g = tri_normal(v3(v0), v3(v1), v3(v2), (__float_as_uint(v1.w) >> 31) != 0u);
        }
        if (!light_emits_towards(Lr, n_g, -rd)) continue;
        const S4 Le = spectrum_sample(wf.spectra, Lr.spectrum, wf.pool.lambda[slot]) * Lr.scale;
        if (!is_positive(Le)) continue;
        const S4 beta = wf.pool.beta[slot];
        S4 L = wf.pool.L[slot];
        if (raw) {
            L = L + beta * Le;
        } else {
            const float4 c0 = wf.pool.ctx0[slot], c1 = wf.pool.ctx1[slot], c2 = wf.pool.ctx2[slot];
            const LightCtx ctx{v3(c0), v3(c0.w, c1.w, c2.w), v3(c1), v3(c2)};
            const float pdf_light = Lr.pmf * (on_sphere ? sphere_light_pdf(wf.spheres[Lr.ext], Lr, ctx, rd) : light_pdf(Lr, ctx, rd));
            const float dir_pdf_scale = wf.has_media ? average(wf.pool.mtp[slot]) : 1.f; // multi_transmittance_pdf.average() (:165)
            const float dir_pdf = wf.pool.misc[slot].y * dir_pdf_scale;
            L = L + beta * Le * power_heuristic(dir_pdf, pdf_light);
        }
#ifdef B200PT_DEBUG_PIXELS
        {
            const int4 ids = wf.pool.ids[slot];
            if (B200PT_DEBUG_PIXELS(ids.x)) {
                const S4 was = wf.pool.L[slot];
                printf("[emit] pixel %d sample %d raw %d add %.9g %.9g %.9g %.9g\n", ids.x, ids.y, raw ? 1 : 0, L.x - was.x, L.y - was.y, L.z - was.z, L.w - was.w);
            }
        }
#endif
        wf.pool.L[slot] = L;
    }
}

// "reflectance" / "transmittance" of a material: a spectrum, or a spectrum texture evaluated at the hit
// (`ext`: a compile-time constant of the calling kernel -- the plain shade kernels carry no texture code at all)
__device__ __forceinline__ S4 material_spectrum(const WF &wf, int spectrum, int tex1, const Isect &si, S4 lambda, bool ext) {
    if (ext && tex1 > 0) return tex_eval(wf.tex, wf.spectra, tex1 - 1, TexCtx{si.u, si.v, si.n, si.p}, lambda);
    return spectrum_sample(wf.spectra, spectrum, lambda);
}

template <int MAT>
__device__ __forceinline__ void make_bsdf(const WF &wf, const MatRec &m, const Isect &si, S4 lambda, S4 &lpdf, bool &lpdf_dirty, bool ext,
                                          BSDF<MAT> &b);

template <>
__device__ __forceinline__ void make_bsdf<B200PT_MAT_DIFFUSE>(const WF &wf, const MatRec &m, const Isect &si, S4 lambda, S4 &lpdf,
                                                              bool &lpdf_dirty, bool ext, BSDF<B200PT_MAT_DIFFUSE> &b) {
    // materials/diffuse_material.cu:17-22
    b.r = clamp_s4(material_spectrum(wf, m.reflectance, m.reflectance_tex1, si, lambda, ext), 0.f, 1.f);
}
template <>
__device__ __forceinline__ void make_bsdf<B200PT_MAT_CONDUCTOR>(const WF &wf, const MatRec &m, const Isect &si, S4 lambda, S4 &lpdf,
                                                                bool &lpdf_dirty, bool ext, BSDF<B200PT_MAT_CONDUCTOR> &b) {
    // materials/conductor_material.cu:52-75
    float ur = m.u_rough, vr = m.v_rough;
    if (ext && m.u_rough_tex1 > 0) ur = ftex_eval(wf.tex, m.u_rough_tex1 - 1, TexCtx{si.u, si.v, si.n, si.p});
    if (ext && m.v_rough_tex1 > 0) vr = ftex_eval(wf.tex, m.v_rough_tex1 - 1, TexCtx{si.u, si.v, si.n, si.p});
    if (m.remap) {
        ur = sqrtf(ur);
        vr = sqrtf(vr);
    }
    if (m.eta >= 0 || m.eta_tex1 > 0) {
        b.eta = material_spectrum(wf, m.eta, m.eta_tex1, si, lambda, ext);
        b.k = material_spectrum(wf, m.k, m.k_tex1, si, lambda, ext);
    } else {
        const S4 r = clamp_s4(material_spectrum(wf, m.reflectance, m.reflectance_tex1, si, lambda, ext), 0.f, 0.9999f);
        b.eta = s4(1.f);
        b.k = 2.f * sqrt_s4(r) / sqrt_s4(clamp_s4(s4(1.f) + r * -1.f, 0.f, kInf));
    }
    b.d = make_trdist(ur, vr);
}
template <>
__device__ __forceinline__ void make_bsdf<B200PT_MAT_DIELECTRIC>(const WF &wf, const MatRec &m, const Isect &si, S4 lambda, S4 &lpdf,
                                                                 bool &lpdf_dirty, bool ext, BSDF<B200PT_MAT_DIELECTRIC> &b) {
    // materials/dielectric_material.cu:39-66
    float eta = spectrum_at(wf.spectra, m.eta, lambda.x);
    if (!spectrum_is_constant(wf.spectra, m.eta)) {
        // SampledWavelengths::terminate_secondary (spectrum_util/sampled_wavelengths.h:120-131)
        if (!(lpdf.y == 0.f && lpdf.z == 0.f && lpdf.w == 0.f)) {
            lpdf = make_float4(lpdf.x / 4.f, 0.f, 0.f, 0.f);
            lpdf_dirty = true;
        }
    }
    if (eta == 0.f) eta = 1.f;
    float ur = m.u_rough, vr = m.v_rough;
    if (ext && m.u_rough_tex1 > 0) ur = ftex_eval(wf.tex, m.u_rough_tex1 - 1, TexCtx{si.u, si.v, si.n, si.p});
    if (ext && m.v_rough_tex1 > 0) vr = ftex_eval(wf.tex, m.v_rough_tex1 - 1, TexCtx{si.u, si.v, si.n, si.p});
    if (m.remap) {
        ur = sqrtf(ur);
        vr = sqrtf(vr);
    }
    b.eta = eta;
    b.d = make_trdist(ur, vr);
}

template <>
__device__ __forceinline__ void make_bsdf<B200PT_MAT_DIFFUSE_TRANSMISSION>(const WF &wf, const MatRec &m, const Isect &si, S4 lambda,
                                                                           S4 &lpdf, bool &lpdf_dirty, bool ext,
                                                                           BSDF<B200PT_MAT_DIFFUSE_TRANSMISSION> &b) {
    // materials/diffuse_transmission_material.cu:25-31
    b.r = clamp_s4(material_spectrum(wf, m.reflectance, m.reflectance_tex1, si, lambda, ext) * m.scale, 0.f, 1.f);
    b.t = clamp_s4(material_spectrum(wf, m.transmittance, m.transmittance_tex1, si, lambda, ext) * m.scale, 0.f, 1.f);
}

template <>
struct BSDF<B200PT_MAT_COATED_DIFFUSE> : LayeredCore<DielectricCore, DiffuseCore> {
    Frame frame;
};
template <>
struct BSDF<B200PT_MAT_COATED_CONDUCTOR> : LayeredCore<DielectricCore, ConductorCore> {
    Frame frame;
};

// the dielectric interface + slab parameters shared by the two coated materials
template <class L>
__device__ __forceinline__ void make_coat(const WF &wf, const MatRec &m, S4 lambda, S4 &lpdf, bool &lpdf_dirty, bool ext, L &b) {
    float ur = m.u_rough, vr = m.v_rough;
    if (m.remap) {
        ur = sqrtf(ur);
        vr = sqrtf(vr);
    }
    float eta = spectrum_at(wf.spectra, m.eta, lambda.x);
    if (!spectrum_is_constant(wf.spectra, m.eta)) {
        if (!(lpdf.y == 0.f && lpdf.z == 0.f && lpdf.w == 0.f)) {
            lpdf = make_float4(lpdf.x / 4.f, 0.f, 0.f, 0.f);
            lpdf_dirty = true;
        }
    }
    if (eta == 0.f) eta = 1.f;
    b.top.eta = eta;
    b.top.d = make_trdist(ur, vr);
    // LayeredBxDF ctor (layered_bxdf.h:17-21): thickness = max(thickness, FLT_MIN)
    b.thickness = fmaxf(m.thickness, FLT_MIN);
    b.albedo = clamp_s4(spectrum_sample(wf.spectra, m.albedo, lambda), 0.f, 1.f);
    b.g = clampf(m.g, -1.f, 1.f);
    b.max_depth = m.max_depth;
    b.n_samples = m.n_samples;
}
template <>
__device__ __forceinline__ void make_bsdf<B200PT_MAT_COATED_DIFFUSE>(const WF &wf, const MatRec &m, const Isect &si, S4 lambda, S4 &lpdf,
                                                                     bool &lpdf_dirty, bool ext, BSDF<B200PT_MAT_COATED_DIFFUSE> &b) {
    // materials/coated_diffuse_material.cu:53-86
    b.bottom.r = clamp_s4(material_spectrum(wf, m.reflectance, m.reflectance_tex1, si, lambda, ext), 0.f, 1.f);
    make_coat(wf, m, lambda, lpdf, lpdf_dirty, ext, b);
}
template <>
__device__ __forceinline__ void make_bsdf<B200PT_MAT_COATED_CONDUCTOR>(const WF &wf, const MatRec &m, const Isect &si, S4 lambda, S4 &lpdf,
                                                                       bool &lpdf_dirty, bool ext, BSDF<B200PT_MAT_COATED_CONDUCTOR> &b) {
    // materials/coated_conductor_material.cu:92-141
    make_coat(wf, m, lambda, lpdf, lpdf_dirty, ext, b);
    S4 ce, ck;
    if (m.c_eta >= 0 || m.eta_tex1 > 0) {
        ce = material_spectrum(wf, m.c_eta, m.eta_tex1, si, lambda, ext);
        ck = material_spectrum(wf, m.c_k, m.k_tex1, si, lambda, ext);
    } else {
        const S4 r = clamp_s4(material_spectrum(wf, m.reflectance, m.reflectance_tex1, si, lambda, ext), 0.f, 0.9999f);
        ce = s4(1.f);
        ck = 2.f * sqrt_s4(r) / sqrt_s4(clamp_s4(s4(1.f) + r * -1.f, 0.f, kInf));
    }
    b.bottom.eta = ce / b.top.eta;
    b.bottom.k = ck / b.top.eta;
    float ur = m.c_u_rough, vr = m.c_v_rough;
    if (m.remap) {
        ur = sqrtf(ur);
        vr = sqrtf(vr);
    }
    b.bottom.d = make_trdist(ur, vr);
}

// ---- gpu_evaluate_material + sample_bsdf + sample_Ld_volume, specialised per material
// EXT: the extended variant of a shade kernel -- media (when the scene has any) and textured materials; scenes with neither
// run the plain one, which carries none of that code
// OCC: resident CTAs per SM the register allocation is held to (0: 2 for the plain materials, 1 for the layered ones)
template <int MAT, bool EXT, int OCC = 0>
__global__ void __launch_bounds__(WF_BLOCK, OCC ? OCC : ((MAT >= B200PT_MAT_COATED_DIFFUSE) ? 1 : 2)) k_shade(WF wf) {
    const bool MEDIA = EXT && wf.has_media;
    const unsigned n = wf.c->p[wf.parity].n_mat[MAT];
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned n_round = (n + 31u) & ~31u;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool act = i < n;
        int slot = -1;
        bool shadow = false;
        float4 sh_o = make_float4(0, 0, 0, 0), sh_d = sh_o, sh_c = sh_o;
        if (act) {
            slot = wf.q.mat[MAT][i];
            // the logic stage left the primitive and the material (a mix already resolved) beside the slot: every load below
            // depends on this one queue read only
            const int2 pm = wf.q.mat_pm[MAT][i];
            const int prim = pm.x, mat = pm.y;
            const MatRec m = wf.materials[mat];
            const float4 ro4 = wf.pool.ray_o[slot], rd4 = wf.pool.ray_d[slot];
            const V3 ro = v3(ro4), rd = v3(rd4);
            const float4 hit = wf.pool.hit[slot];
            Isect si;
            int prim_id = -1;
            if (prim >= (int)wf.bvh.n_tris) {
                const SphereRec &sp = wf.spheres[prim - (int)wf.bvh.n_tris];
                sphere_interaction(sp, ro, rd, hit.y, -rd, si);
                prim_id = -2 - (prim - (int)wf.bvh.n_tris);
            } else {
                const float4 *tp = wf.bvh.tris + 3 * (size_t)prim;
                prim_id = __float_as_int(__ldg(&tp[0].w));
                hit_triangle_interaction(wf, slot, tp, hit, ro, rd, si);
            }
            float4 vol = make_float4(0.f, 0.f, 0.f, 0.f); // media scenes: .y = medium of the ray that arrived here
            if (MEDIA) vol = wf.pool.vol[slot];
            const S4 lambda = wf.pool.lambda[slot];
            S4 lpdf = wf.pool.lpdf[slot];
            bool lpdf_dirty = false;
            int4 ids = wf.pool.ids[slot];
            // All four sampler draws of this vertex are taken HERE, before the geometry and the BSDF are set up: each draw is a
            // long chain of dependent integer operations (MurmurHash of (pixel, dimension), Kensler permutation, a modulo by spp)
            // that depends on nothing but the path's ids, so the four chains overlap with one another and with the memory
            // latency of the triangle / material / light records instead of running one after the other between them.  The
            // consumption order is the reference's: NEE draws get_1d + get_2d before any early-out, then BSDF sampling draws
            // get_1d + get_2d (megakernel_path.cu:201-207, wavefront_path.cu:458-459); a specular BSDF skips NEE, and then
            // the first two draws ARE the BSDF's.
            Sampler s = load_sampler(wf, slot, ids);
            const float draw_a = s.get_1d();
            const float2 draw_b = s.get_2d();
            const uint64_t state_ab = s.state;
            const int dim_ab = s.dimension;
            const float draw_c = s.get_1d();
            const float2 draw_d = s.get_2d();
            BSDF<MAT> bsdf;
            // BSDF(ns, dpdus) (base/bsdf.h:16-17): Frame::from_xz(normalize(dpdus), ns)
            bsdf.frame = frame_from_xz(normalize(si.dpdu), si.ns);
            make_bsdf<MAT>(wf, m, si, lambda, lpdf, lpdf_dirty, EXT, bsdf);
            if (lpdf_dirty) wf.pool.lpdf[slot] = lpdf;
            if (wf.regularize && (ids.w & PF_ANY_NON_SPECULAR)) bsdf.regularize();
            const uint32_t flags = bsdf.flags();
            S4 beta = wf.pool.beta[slot];
            const bool nee = flags_non_specular(flags);
            if (!nee) { // the sampler stands after two draws
                s.state = state_ab;
                s.dimension = dim_ab;
            }
            const V3 wo = si.wo;
            const V3 wo_l = bsdf.frame.to_local(wo);

            // ---- next-event estimation (MegakernelPathIntegrator::sample_Ld_volume, megakernel_path.cu:185-264)
            if (nee) {
                LightCtx ctx{si.p, si.err, si.n, si.ns};
                if ((flags & BXDF_REFLECTION) && !(flags & BXDF_TRANSMISSION)) {
                    ctx.p = offset_ray_origin(si.p, si.err, si.n, wo);
                    ctx.err = v3(0.f, 0.f, 0.f);
                } else if ((flags & BXDF_TRANSMISSION) && !(flags & BXDF_REFLECTION)) {
                    ctx.p = offset_ray_origin(si.p, si.err, si.n, -wo);
                    ctx.err = v3(0.f, 0.f, 0.f);
                }
                const float ul = draw_a;
                const float2 up = draw_b; // both drawn before any early-out, as the reference does
                if (wf.n_lights > 0) {
                    const int li = alias_sample(wf.alias, wf.n_lights, ul);
                    const LightRec Lr = wf.lights[li];
                    const int lt = light_type(Lr);
                    LightSample ls;
                    bool have, delta = false;
                    float lw = 1.f, ld2 = 1.f;
                    S4 Le_env = s4(0.f);
                    const bool is_env = EXT && lt == B200PT_LIGHT_IMAGE_INFINITE; // such scenes run the extended kernels
                    if (lt == B200PT_LIGHT_AREA) have = light_sample(Lr, ctx, up, ls) && light_emits_towards(Lr, ls.n, -ls.wi);
                    else if (lt == B200PT_LIGHT_SPHERE_AREA) have = sphere_light_sample(wf.spheres[Lr.ext], Lr, ctx, up, ls) && light_emits_towards(Lr, ls.n, -ls.wi);
                    else if (is_env) have = env_nee(wf, ctx.p, up, lambda, ls, Le_env);
                    else have = light_sample_other(Lr, lt, wf.light_ext, wf.scene_radius, ctx, up, ls, &lw, &ld2, &delta);
                    if (have) {
                        S4 Le;
                        if (is_env) Le = Le_env;
                        else if (lt == B200PT_LIGHT_SPOT) Le = spectrum_sample(wf.spectra, Lr.spectrum, lambda) * (lw * Lr.scale) / ld2;
                        else Le = spectrum_sample(wf.spectra, Lr.spectrum, lambda) * Lr.scale;
                        if (is_positive(Le) && wo_l.z != 0.f) {
                            const V3 wi_l = bsdf.frame.to_local(ls.wi);
                            const S4 f = bsdf.f(wo_l, wi_l) * absdot(ls.wi, si.ns);
                            if (is_positive(f)) {
                                const float pdf_light = Lr.pmf * ls.pdf;
                                // visibility: Ray::spawn_ray_to (base/ray.h:51-58) from the surface point to the
                                // light point, both ends offset; the segment (0, 1] of the un-normalised ray covers
                                // the union of the reference's three visibility rays (integrator_base.cu:12-14,24-33)
                                const V3 pf = offset_ray_origin(si.p, si.err, si.n, ls.p - si.p);
                                const V3 pt = offset_ray_origin(ls.p, ls.err, ls.n, pf - ls.p);
                                const V3 sd = pt - pf;
                                // compute_transmittance (base/integrator_base.cu:22-71) for a surface point: `unoccluded` lets
                                // nothing through, interfaces included, so whenever the sample survives the walk was one
                                // segment from pf to the light point inside the medium on that side of the surface
                                float pdf_bsdf = delta ? 0.f : bsdf.pdf(wo_l, wi_l);
                                if (MEDIA) {
                                    const int sm = surface_medium(wf, prim_id, si.n, sd, __float_as_int(vol.y));
                                    if (sm >= 0) {
                                        const S4 T = exp_s4(medium_sigma_t(wf, sm, lambda) * (-len(ls.p - pf)));
                                        Le = Le * T;
                                        pdf_bsdf *= average(T);
                                    }
                                }
                                // delta lights: no MIS (megakernel_path.cu:236-239)
                                const float w = delta ? 1.f : power_heuristic(pdf_light, pdf_bsdf);
                                const S4 Ld = f * (Le / pdf_light) * w;
#ifdef B200PT_DEBUG_PIXELS // debug build only (tools/build_debug_lib.py): the NEE record of chosen pixels, one line per sample
                                if (B200PT_DEBUG_PIXELS(ids.x))
                                    printf("[nee] pixel %d sample %d depth %d mat %d contrib %.9g %.9g %.9g %.9g p %.9g %.9g %.9g err %.9g %.9g %.9g n %.9g %.9g %.9g lp %.9g %.9g %.9g "
                                           "lerr %.9g %.9g %.9g ln %.9g %.9g %.9g skip %d\n",
                                           ids.x, ids.y, (ids.w >> 16) & 0xff, MAT, (beta * Ld).x, (beta * Ld).y, (beta * Ld).z, (beta * Ld).w, si.p.x, si.p.y, si.p.z,
                                           si.err.x, si.err.y, si.err.z, si.n.x, si.n.y, si.n.z, ls.p.x, ls.p.y, ls.p.z, ls.err.x, ls.err.y, ls.err.z, ls.n.x, ls.n.y,
                                           ls.n.z, Lr.prim);
#endif
                                shadow = true;
                                sh_o = make_float4(pf.x, pf.y, pf.z, 1.f);
                                sh_d = make_float4(sd.x, sd.y, sd.z, __int_as_float(Lr.prim)); // .w: primitive to skip
                                sh_c = beta * Ld;
                            }
                        }
                    }
                }
            }

            // ---- BSDF sampling (WavefrontPathIntegrator::sample_bsdf, wavefront_path.cu:456-474)
            const float uc = nee ? draw_c : draw_a;
            const float2 u2 = nee ? draw_d : draw_b;
            BSDFSample bs;
            bool ok = wo_l.z != 0.f && flags != 0u && bsdf.sample(wo_l, uc, u2, bs);
            if (ok) ok = is_positive(bs.f) && bs.pdf != 0.f && bs.wi.z != 0.f;
            if (!ok) {
                wf.pool.beta[slot] = s4(0.f);
                wf.pool.ray_o[slot] = make_float4(ro.x, ro.y, ro.z, -1.f); // dead: the next logic stage retires it
            } else {
                const V3 wi = bsdf.frame.from_local(bs.wi);
                beta = beta * bs.f * (absdot(wi, si.ns) / bs.pdf);
                wf.pool.beta[slot] = beta;
                const bool spec = (bs.flags & BXDF_SPECULAR) != 0u;
                ids.w = (ids.w & ~PF_SPECULAR) | (spec ? PF_SPECULAR : PF_ANY_NON_SPECULAR);
                float4 misc = wf.pool.misc[slot];
                // prev_direction_pdf (wavefront_path.cu:466-467): layered samples carry a pdf that is only proportional, so
                // BSDF::pdf is evaluated -- on the render-space wi brought back to the local frame, as BSDF::pdf does
                misc.y = bs.pdf_is_proportional ? bsdf.pdf(wo_l, bsdf.frame.to_local(wi)) : bs.pdf;
                wf.pool.misc[slot] = misc;
                wf.pool.ctx0[slot] = make_float4(si.p.x, si.p.y, si.p.z, si.err.x);
                wf.pool.ctx1[slot] = make_float4(si.n.x, si.n.y, si.n.z, si.err.y);
                wf.pool.ctx2[slot] = make_float4(si.ns.x, si.ns.y, si.ns.z, si.err.z);
                // Interaction::spawn_ray (base/interaction.h:73-76)
                const V3 o = offset_ray_origin(si.p, si.err, si.n, wi);
                wf.pool.ray_o[slot] = make_float4(o.x, o.y, o.z, kInf);
                wf.pool.ray_d[slot] = make_float4(wi.x, wi.y, wi.z, 0.f);
                if (MEDIA) {
                    vol.y = __int_as_float(surface_medium(wf, prim_id, si.n, wi, __float_as_int(vol.y)));
                    wf.pool.vol[slot] = vol;
                    wf.pool.mtp[slot] = s4(1.f); // :463
                }
            }
            ids.z = s.dimension;
            wf.pool.ids[slot] = ids;
            wf.pool.rng[slot] = s.state;
            wf.pool.state[slot] = ST_RAY; // traced and judged by the next pass (dead paths: tmax < 0)
        }
        const int spos = queue_claim(&wf.c->p[wf.parity].n_shadow, shadow);
        if (spos >= 0) {
            wf.q.sh_o[spos] = sh_o;
            wf.q.sh_d[spos] = sh_d;
            wf.q.sh_contrib[spos] = sh_c;
            wf.q.sh_slot[spos] = slot;
        }
    }
}

} // namespace b200pt

namespace b200pt {

template <typename T>
static int pool_alloc(RenderData *r, T **out, size_t count) {
    void *p = nullptr;
    CU_TRY(cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
    r->pool_allocs.push_back(p);
    *out = (T *)p;
    return B200PT_OK;
}

static int ensure_pool(b200pt_scene *s, int pool_size) {
    RenderData *r = s->render;
    if (r->pool_size == pool_size) return B200PT_OK;
    CU_TRY(cudaStreamSynchronize(s->stream));
    for (void *p : r->pool_allocs) cudaFree(p);
    r->pool_allocs.clear();
    r->pool_size = 0;
    const size_t n = (size_t)pool_size;
    int rc = 0;
    PathPool &P = r->pool;
    float4 **lanes[] = {&P.ray_o, &P.ray_d, &P.hit, &P.beta, &P.L, &P.lambda, &P.lpdf, &P.ctx0, &P.ctx1, &P.ctx2, &P.misc};
    for (float4 **l : lanes)
        if ((rc = pool_alloc(r, l, n))) return rc;
    P.mtp = P.vol = nullptr;
    r->q.vol = nullptr;
    r->q.tr_a = r->q.tr_beta = nullptr;
    if (r->n_media > 0) { // two more lanes and the scatter queue, media scenes only
        if ((rc = pool_alloc(r, &P.mtp, n))) return rc;
        if ((rc = pool_alloc(r, &P.vol, n))) return rc;
        if ((rc = pool_alloc(r, &r->q.vol, n))) return rc;
        if ((rc = pool_alloc(r, &r->q.tr_a, n))) return rc;
        if ((rc = pool_alloc(r, &r->q.tr_beta, n))) return rc;
    }
    P.inst = nullptr;
    if (!s->instances.empty() && (rc = pool_alloc(r, &P.inst, n))) return rc;
    if ((rc = pool_alloc(r, &P.rng, n))) return rc;
    if ((rc = pool_alloc(r, &P.ids, n))) return rc;
    if ((rc = pool_alloc(r, &P.state, n))) return rc;
    WFQueues &Q = r->q;
    for (int m = 0; m < WF_NMAT; m++) {
        if ((rc = pool_alloc(r, &Q.mat[m], r->has_mat[m] ? n : 1))) return rc;
        if ((rc = pool_alloc(r, &Q.mat_pm[m], r->has_mat[m] ? n : 1))) return rc;
    }
    int32_t **qs[] = {&Q.free_slots, &Q.sh_slot, &Q.emit};
    for (int32_t **q : qs)
        if ((rc = pool_alloc(r, q, n))) return rc;
    if ((rc = pool_alloc(r, &Q.sh_o, n))) return rc;
    if ((rc = pool_alloc(r, &Q.sh_d, n))) return rc;
    if ((rc = pool_alloc(r, &Q.sh_contrib, n))) return rc;
    if (!r->d_counters) CU_TRY(cudaMalloc((void **)&r->d_counters, sizeof(WFCounters)));
    if (!r->h_counters) CU_TRY(cudaMallocHost((void **)&r->h_counters, 2 * sizeof(WFCounters)));
    r->pool_size = pool_size;
    return B200PT_OK;
}

} // namespace b200pt

B200PT_API int b200pt_probe_sampler(b200pt_scene *s, int kind, int spp, const int32_t *pixel, const int32_t *sample,
                                    int npairs, const int32_t *pattern, int ndraw, float *out);

B200PT_API int b200pt_film_create(b200pt_scene *s, b200pt_film **out) {
    if (!s || !out) {
        set_error("b200pt_film_create: null argument");
        return B200PT_ERR_INVALID;
    }
    *out = nullptr;
    if (!s->render || !s->render->have_render) {
        set_error("b200pt_film_create: scene has no film / camera / filter description");
        return B200PT_ERR_STATE;
    }
    CU_TRY(cudaSetDevice(s->device));
    b200pt_film *f = new b200pt_film();
    f->scene = s;
    f->width = s->render->width;
    f->height = s->render->height;
    const size_t bytes = (size_t)f->width * f->height * sizeof(float4);
    if (cudaMalloc((void **)&f->d_accum, bytes) != cudaSuccess) {
        delete f;
        set_error("b200pt_film_create: out of device memory");
        return B200PT_ERR_CUDA;
    }
    cudaMemsetAsync(f->d_accum, 0, bytes, s->stream);
    cudaStreamSynchronize(s->stream);
    s->render->films.push_back(f);
    *out = f;
    return B200PT_OK;
}

// a film whose scene has been destroyed is a dead handle: only b200pt_film_destroy may still be called on it
static int film_alive(const b200pt_film *film, const char *who) {
    if (!film) {
        set_error("%s: null film", who);
        return B200PT_ERR_INVALID;
    }
    if (!film->scene || !film->d_accum) {
        set_error("%s: the film's scene has been destroyed", who);
        return B200PT_ERR_STATE;
    }
    return B200PT_OK;
}

B200PT_API void b200pt_film_destroy(b200pt_film *film) {
    if (!film) return;
    if (film->scene) {
        cudaSetDevice(film->scene->device);
        if (film->scene->render) {
            std::vector<b200pt_film *> &v = film->scene->render->films;
            v.erase(std::remove(v.begin(), v.end(), film), v.end());
        }
        if (film->d_accum) cudaFree(film->d_accum);
    }
    delete film;
}

B200PT_API int b200pt_film_clear(b200pt_film *film) {
    if (int rc = film_alive(film, "b200pt_film_clear")) return rc;
    CU_TRY(cudaSetDevice(film->scene->device));
    CU_TRY(cudaMemsetAsync(film->d_accum, 0, (size_t)film->width * film->height * sizeof(float4), film->scene->stream));
    CU_TRY(cudaStreamSynchronize(film->scene->stream));
    return B200PT_OK;
}

B200PT_API int b200pt_render(b200pt_scene *s, b200pt_film *film, const b200pt_render_opts *opts, b200pt_render_stats *st_out) {
    if (!s || !film || !opts) {
        set_error("b200pt_render: null argument");
        return B200PT_ERR_INVALID;
    }
    if (int rc = film_alive(film, "b200pt_render")) return rc;
    if (!s->render || !s->render->have_render || film->scene != s) {
        set_error("b200pt_render: scene has no render description, or the film belongs to another scene");
        return B200PT_ERR_STATE;
    }
    if (!s->built) {
        set_error("b200pt_render: call b200pt_bvh_build first");
        return B200PT_ERR_STATE;
    }
    RenderData *r = s->render;
    const int spp = opts->spp_total;
    const int sqrt_spp = (int)std::lround(std::sqrt((double)spp));
    if (spp <= 0 || opts->sample_begin < 0 || opts->sample_end > spp || opts->sample_begin > opts->sample_end) {
        set_error("b200pt_render: bad sample range [%d, %d) of %d", opts->sample_begin, opts->sample_end, spp);
        return B200PT_ERR_INVALID;
    }
    if (opts->sampler == 0 && sqrt_spp * sqrt_spp != spp) {
        // samplers/stratified.h:11-16: fatal in the reference
        set_error("b200pt_render: the stratified sampler needs a perfect-square spp (got %d)", spp);
        return B200PT_ERR_INVALID;
    }
    if (opts->max_depth < 0 || opts->max_depth > 255) {
        set_error("b200pt_render: max_depth must be in [0, 255]");
        return B200PT_ERR_INVALID;
    }
    CU_TRY(cudaSetDevice(s->device));
    const long long n_pixels = (long long)r->width * r->height;
    const long long total = n_pixels * (opts->sample_end - opts->sample_begin);
    // default pool: 16 Mi slots (about 5 GB of the 180): passes over a bigger pool have shorter tails and fewer launches --
    // C3's shape 622 -> 683 Msamples/s against the 2 Mi of round 1, nothing more beyond 16 Mi (tools/render_probe.py --pool)
    int pool_size = opts->pool_size > 0 ? opts->pool_size : (1 << 24);
    pool_size = (int)std::min<long long>(pool_size, std::max<long long>(total, 1024));
    pool_size = (pool_size + 31) & ~31;
    int rc = ensure_pool(s, pool_size);
    if (rc) return rc;

    WF wf = {};
    wf.pool = r->pool;
    wf.q = r->q;
    wf.c = r->d_counters;
    wf.pool_size = pool_size;
    wf.bvh = s->bvh;
    wf.spheres = s->d_spheres;
    wf.n_spheres = s->n_spheres;
    wf.instances = s->d_instances;
    wf.n_instances = (int32_t)s->instances.size();
    wf.inst_root = s->inst_root_node;
    wf.N = s->d_N;
    wf.uv = s->d_uv;
    wf.idx = s->d_idx;
    wf.spectra = SpectrumTable{r->d_spec, r->d_spec_pool, r->d_illuminant};
    wf.materials = r->d_materials;
    wf.lights = r->d_lights;
    wf.alias = r->d_alias;
    wf.light_ext = r->d_light_ext;
    wf.n_infinite = r->n_infinite;
    for (int k = 0; k < 4; k++) wf.infinite[k] = r->infinite[k];
    wf.scene_radius = r->scene_radius;
    wf.has_mix = r->has_mix ? 1 : 0;
    wf.tex = r->tex;
    wf.env = r->env;
    wf.media = r->d_media;
    wf.tri_medium = r->d_tri_medium;
    wf.has_media = r->n_media > 0 ? 1 : 0;
    wf.camera_medium = r->camera_medium;
    // a warp of k_trace_wf fetches new rays once this many lanes are idle: 4 is the measured best on triangle-only scenes;
    // with spheres the per-ray start-up is longer and waiting for half a warp pays (C1: 62.0 -> 58.5 ms together with the
    // sphere phase)
    wf.refill_min = s->tune_wf_refill > 0 ? s->tune_wf_refill : (s->n_spheres > 0 ? 16 : 4);
    wf.tri_steps = s->tune_wf_tri_steps > 0 ? s->tune_wf_tri_steps : 2;
    const bool ext_shade = wf.has_media || r->tex.tex != nullptr || r->env.light >= 0;
    int n_shade_kernels = 0;
    for (int m = 0; m < WF_NMAT; m++) n_shade_kernels += r->has_mat[m] ? 1 : 0;
    wf.n_lights = r->n_lights;
    wf.filter = r->filter;
    wf.camera = r->camera;
    wf.film = FilmRec{r->width, r->height, r->d_cie[0], r->d_cie[1], r->d_cie[2], r->imaging_ratio, film->d_accum};
    wf.sampler_kind = opts->sampler;
    wf.spp = spp;
    wf.sqrt_spp = sqrt_spp;
    wf.max_depth = opts->max_depth;
    wf.regularize = opts->regularize;
    wf.sample_begin = opts->sample_begin;
    wf.total_paths = total;
    wf.n_pixels = n_pixels;

    cudaStream_t st = s->stream;
    cudaEvent_t ev0, ev1, ev_poll[2];
    CU_TRY(cudaEventCreate(&ev0));
    CU_TRY(cudaEventCreate(&ev1));
    CU_TRY(cudaEventCreateWithFlags(&ev_poll[0], cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&ev_poll[1], cudaEventDisableTiming));
    struct EvGuard {
        cudaEvent_t a, b, c, d;
        ~EvGuard() {
            cudaEventDestroy(a);
            cudaEventDestroy(b);
            cudaEventDestroy(c);
            cudaEventDestroy(d);
        }
    } guard{ev0, ev1, ev_poll[0], ev_poll[1]};

    const int shade_occ = s->tune_shade_occ;
    const int grid = s->sm_count * 8; // stage kernels: grid-stride over the pool / device-side queue counts
    CU_TRY(cudaEventRecord(ev0, st));
    k_init_pool<<<(pool_size + 255) / 256, 256, 0, st>>>(wf);
    long long passes = 0, launches = 1;
    const int batch = 4; // passes enqueued between two looks at the finished-path counter
    // The only host<->device traffic of the render loop is one 64-byte counter block per batch, read ONE
    // BATCH LATE: batch k+1 is already enqueued when the host waits for the counters of batch k, so the
    // GPU never idles on the poll.
    WFCounters *hc = r->h_counters; // [2], pinned
    int last = -1;
    bool done = total == 0;
    for (long long k = 0; !done; k++) {
        for (int b = 0; b < batch; b++) {
            wf.parity = int(passes & 1);
            if ((rc = launch_trace_wf(s, wf, false))) return rc;
            if (wf.has_media) {
                k_volume<0><<<grid, VOL_BLOCK, 0, st>>>(wf);
                k_volume<1><<<grid, VOL_BLOCK, 0, st>>>(wf);
                if (wf.n_instances > 0) k_transmit<true><<<grid, VOL_BLOCK, 0, st>>>(wf);
                else k_transmit<false><<<grid, VOL_BLOCK, 0, st>>>(wf);
            }
            if (r->env.light >= 0) {
                if (wf.has_media) k_logic<true, true><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_logic<false, true><<<grid, WF_BLOCK, 0, st>>>(wf);
            } else {
                if (wf.has_media) k_logic<true, false><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_logic<false, false><<<grid, WF_BLOCK, 0, st>>>(wf);
            }
            if (r->n_area_lights > 0) k_emit<<<grid, WF_BLOCK, 0, st>>>(wf);
            k_generate<<<grid, WF_BLOCK, 0, st>>>(wf);
            // one shade kernel per material type PRESENT in the scene (a type without materials has no queue)
            if (r->has_mat[B200PT_MAT_DIFFUSE]) {
                if (ext_shade) k_shade<B200PT_MAT_DIFFUSE, true><<<grid, WF_BLOCK, 0, st>>>(wf);
                else if (shade_occ == 3) k_shade<B200PT_MAT_DIFFUSE, false, 3><<<grid, WF_BLOCK, 0, st>>>(wf);
                else if (shade_occ == 4) k_shade<B200PT_MAT_DIFFUSE, false, 4><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_shade<B200PT_MAT_DIFFUSE, false><<<grid, WF_BLOCK, 0, st>>>(wf);
            }
            if (r->has_mat[B200PT_MAT_CONDUCTOR]) {
                if (ext_shade) k_shade<B200PT_MAT_CONDUCTOR, true><<<grid, WF_BLOCK, 0, st>>>(wf);
                else if (shade_occ == 3) k_shade<B200PT_MAT_CONDUCTOR, false, 3><<<grid, WF_BLOCK, 0, st>>>(wf);
                else if (shade_occ == 4) k_shade<B200PT_MAT_CONDUCTOR, false, 4><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_shade<B200PT_MAT_CONDUCTOR, false><<<grid, WF_BLOCK, 0, st>>>(wf);
            }
            if (r->has_mat[B200PT_MAT_DIELECTRIC]) {
                if (ext_shade) k_shade<B200PT_MAT_DIELECTRIC, true><<<grid, WF_BLOCK, 0, st>>>(wf);
                else if (shade_occ == 3) k_shade<B200PT_MAT_DIELECTRIC, false, 3><<<grid, WF_BLOCK, 0, st>>>(wf);
                else if (shade_occ == 4) k_shade<B200PT_MAT_DIELECTRIC, false, 4><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_shade<B200PT_MAT_DIELECTRIC, false><<<grid, WF_BLOCK, 0, st>>>(wf);
            }
            if (r->has_mat[B200PT_MAT_DIFFUSE_TRANSMISSION]) {
                if (ext_shade) k_shade<B200PT_MAT_DIFFUSE_TRANSMISSION, true><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_shade<B200PT_MAT_DIFFUSE_TRANSMISSION, false><<<grid, WF_BLOCK, 0, st>>>(wf);
            }
            if (r->has_mat[B200PT_MAT_COATED_DIFFUSE]) {
                if (ext_shade) k_shade<B200PT_MAT_COATED_DIFFUSE, true><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_shade<B200PT_MAT_COATED_DIFFUSE, false><<<grid, WF_BLOCK, 0, st>>>(wf);
            }
            if (r->has_mat[B200PT_MAT_COATED_CONDUCTOR]) {
                if (ext_shade) k_shade<B200PT_MAT_COATED_CONDUCTOR, true><<<grid, WF_BLOCK, 0, st>>>(wf);
                else k_shade<B200PT_MAT_COATED_CONDUCTOR, false><<<grid, WF_BLOCK, 0, st>>>(wf);
            }
            if ((rc = launch_trace_wf(s, wf, true))) return rc;
            passes++;
            launches += 4 + n_shade_kernels + (wf.has_media ? 3 : 0) + (r->n_area_lights > 0 ? 1 : 0);
        }
        CU_TRY(cudaGetLastError());
        const int cur = int(k & 1);
        CU_TRY(cudaMemcpyAsync(&hc[cur], r->d_counters, sizeof(WFCounters), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaEventRecord(ev_poll[cur], st));
        if (k >= 1) {
            CU_TRY(cudaEventSynchronize(ev_poll[cur ^ 1]));
            if (hc[cur ^ 1].stack_overflow) {
                set_error("b200pt_render: traversal stack overflow");
                return B200PT_ERR_STATE;
            }
            done = (long long)hc[cur ^ 1].finished >= total;
        }
        last = cur;
        if (passes > 100000000ll) {
            set_error("b200pt_render: did not terminate");
            return B200PT_ERR_STATE;
        }
    }
    if (last >= 0) {
        CU_TRY(cudaEventSynchronize(ev_poll[last]));
        if ((long long)hc[last].finished < total) { // the look-ahead batch was needed after all
            set_error("b200pt_render: internal error: %lld of %lld paths finished", (long long)hc[last].finished, total);
            return B200PT_ERR_STATE;
        }
    }
    CU_TRY(cudaEventRecord(ev1, st));
    CU_TRY(cudaEventSynchronize(ev1));
    if (st_out) {
        b200pt_render_stats o = {};
        o.paths = total;
        if (total > 0) {
            o.rays_closest = (int64_t)hc[last].rays_closest;
            o.rays_shadow = (int64_t)hc[last].rays_shadow;
        }
        o.passes = passes;
        o.kernel_launches = launches;
        CU_TRY(cudaEventElapsedTime(&o.ms_total, ev0, ev1));
        *st_out = o;
    }
    return B200PT_OK;
}

// ---- film read-out and reduce -------------------------------------------------------------------
namespace b200pt {
// RGBFilm::get_pixel_rgb (film/rgb_film.cu:107-122): rgb_sum / weight_sum, then output_rgb_from_sensor_rgb
__global__ void k_film_resolve(const float4 *accum, long long n, const float *M, float *rgb) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = accum[i];
    float x = a.x, y = a.y, z = a.z;
    if (a.w != 0.f) {
        x /= a.w;
        y /= a.w;
        z /= a.w;
    }
    rgb[3 * i + 0] = M[0] * x + M[1] * y + M[2] * z;
    rgb[3 * i + 1] = M[3] * x + M[4] * y + M[5] * z;
    rgb[3 * i + 2] = M[6] * x + M[7] * y + M[8] * z;
}

// Film::write_to_png's pixel loop on the device (base/film.cu:57-99): get_pixel_rgb, then SRGBColorEncoding::from_linear
// (spectrum_util/color_encoding.h:6-45) -- the reference's sRGB curve is the minimax rational approximation, evaluated on the
// HOST there (each pixel read through managed memory); rgba[4i..] = {r, g, b, 255}; nan_count += pixels with a NaN component
__device__ __forceinline__ float linear_to_srgb(float v) {
    if (v <= 0.0031308f) return 12.92f * v;
    const float s = safe_sqrt(v);
    float p = -0.016202083165206348f;
    p = fmaf(s, p, 0.7551545191665577f);
    p = fmaf(s, p, 2.0041169284241644f);
    p = fmaf(s, p, 0.7642611304733891f);
    p = fmaf(s, p, 0.03453868659826638f);
    p = fmaf(s, p, -0.0016829072605308378f);
    float q = 1.f;
    q = fmaf(s, q, 1.8970238036421054f);
    q = fmaf(s, q, 0.6085338522168684f);
    q = fmaf(s, q, 0.03467195408529984f);
    q = fmaf(s, q, -0.00004375359692957097f);
    q = fmaf(s, q, 4.178892964897981e-7f);
    return p / q * v;
}
__device__ __forceinline__ unsigned char srgb8_from_linear(float v) {
    if (!(v > 0.f)) return 0; // also NaN (the reference aborts on NaN pixels; here they are counted and written as 0)
    if (v >= 1.f) return 255;
    const double r = round(255.0 * (double)linear_to_srgb(v));
    return (unsigned char)(r < 0.0 ? 0.0 : (r > 255.0 ? 255.0 : r));
}
__global__ void k_film_srgb8(const float4 *accum, long long n, const float *M, uchar4 *rgba, unsigned *nan_count) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = accum[i];
    float x = a.x, y = a.y, z = a.z;
    if (a.w != 0.f) {
        x /= a.w;
        y /= a.w;
        z /= a.w;
    }
    const float r = M[0] * x + M[1] * y + M[2] * z, g = M[3] * x + M[4] * y + M[5] * z, b = M[6] * x + M[7] * y + M[8] * z;
    if (isnan(r) || isnan(g) || isnan(b)) atomicAdd(nan_count, 1u);
    rgba[i] = make_uchar4(srgb8_from_linear(r), srgb8_from_linear(g), srgb8_from_linear(b), 255);
}

__global__ void k_probe_sampler(int kind, int spp, int sqrt_spp, const int32_t *__restrict__ pixel,
                                const int32_t *__restrict__ sample, int npairs, const int32_t *__restrict__ pattern, int ndraw,
                                int per, float *__restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= npairs) return;
    Sampler s;
    s.start(kind, spp, sqrt_spp, pixel[k], sample[k], 0);
    float *o = out + (size_t)k * per;
    for (int i = 0; i < ndraw; i++) {
        if (pattern[i] == 1) {
            *o++ = s.get_1d();
        } else {
            const float2 v = s.get_2d();
            *o++ = v.x;
            *o++ = v.y;
        }
    }
}
} // namespace b200pt

B200PT_API int b200pt_film_read_raw(b200pt_film *film, float *xyzw_out) {
    if (!xyzw_out) return B200PT_ERR_INVALID;
    if (int rc = film_alive(film, "b200pt_film_read_raw")) return rc;
    b200pt_scene *s = film->scene;
    CU_TRY(cudaSetDevice(s->device));
    CU_TRY(cudaMemcpyAsync(xyzw_out, film->d_accum, (size_t)film->width * film->height * sizeof(float4), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}

B200PT_API int b200pt_film_read(b200pt_film *film, float *rgb_out) {
    if (!rgb_out) return B200PT_ERR_INVALID;
    if (int rc = film_alive(film, "b200pt_film_read")) return rc;
    b200pt_scene *s = film->scene;
    CU_TRY(cudaSetDevice(s->device));
    const long long n = (long long)film->width * film->height;
    int rc = ensure_scratch(s, (size_t)n * 12 + 64);
    if (rc) return rc;
    float *d_M = (float *)s->d_scratch;
    float *d_rgb = d_M + 16;
    CU_TRY(cudaMemcpyAsync(d_M, s->render->out_rgb, 9 * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    k_film_resolve<<<unsigned((n + 255) / 256), 256, 0, s->stream>>>(film->d_accum, n, d_M, d_rgb);
    CU_TRY(cudaMemcpyAsync(rgb_out, d_rgb, (size_t)n * 12, cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}

B200PT_API int b200pt_film_read_srgb8(b200pt_film *film, uint8_t *rgba_out, int32_t *nan_pixels_out) {
    if (!rgba_out) return B200PT_ERR_INVALID;
    if (int rc = film_alive(film, "b200pt_film_read_srgb8")) return rc;
    b200pt_scene *s = film->scene;
    CU_TRY(cudaSetDevice(s->device));
    const long long n = (long long)film->width * film->height;
    int rc = ensure_scratch(s, (size_t)n * 4 + 128);
    if (rc) return rc;
    float *d_M = (float *)s->d_scratch;
    unsigned *d_nan = (unsigned *)(d_M + 12);
    uchar4 *d_rgba = (uchar4 *)(d_M + 16);
    unsigned h_nan = 0;
    CU_TRY(cudaMemcpyAsync(d_M, s->render->out_rgb, 9 * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemsetAsync(d_nan, 0, sizeof(unsigned), s->stream));
    k_film_srgb8<<<unsigned((n + 255) / 256), 256, 0, s->stream>>>(film->d_accum, n, d_M, d_rgba, d_nan);
    CU_TRY(cudaMemcpyAsync(rgba_out, d_rgba, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaMemcpyAsync(&h_nan, d_nan, sizeof(unsigned), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    if (nan_pixels_out) *nan_pixels_out = (int32_t)h_nan;
    return B200PT_OK;
}

// ---- b200pt_film_read_srgb8_static: the sRGB8 conversion for callers that may not allocate -----------------------------------
// No counterpart in the reference.  b200pt_film_read_srgb8 stages its 8-bit image
// in the scene's scratch buffer, which ensure_scratch() grows with cudaMalloc.  This variant stages it in a file-scope device
// array instead: every GPU that loads the library holds the array (8 MiB + 256 B) until the process exits, every call shares it,
// and it keeps the last call's pixels -- the NaN counter is cleared per call, the pixels of the film are overwritten before
// they are copied out, and two calls that overlap on one GPU race.  The host never uses the array's name as an address: the
// device address comes from cudaGetSymbolAddress, once per device, with that device current.
namespace b200pt {
constexpr long long kStaticFilmMaxPixels = 1ll << 21; // 2,097,152 pixels: a 1920x1080 film fits
constexpr size_t kStaticFilmHeader = 256;             // [0, 4): NaN-pixel counter; the RGBA bytes start at 256
static __device__ __align__(256) unsigned char g_film_static_scratch[kStaticFilmHeader + 4 * kStaticFilmMaxPixels];

struct Mat3 {
    float m[9];
};

__global__ void k_film_srgb8_static(const float4 *accum, long long n, Mat3 M, unsigned char *scratch) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned *nan_count = (unsigned *)scratch;
    uchar4 *rgba = (uchar4 *)(scratch + kStaticFilmHeader);
    const float4 a = accum[i];
    float x = a.x, y = a.y, z = a.z;
    if (a.w != 0.f) {
        x /= a.w;
        y /= a.w;
        z /= a.w;
    }
    const float r = M.m[0] * x + M.m[1] * y + M.m[2] * z, g = M.m[3] * x + M.m[4] * y + M.m[5] * z, b = M.m[6] * x + M.m[7] * y + M.m[8] * z;
    if (isnan(r) || isnan(g) || isnan(b)) atomicAdd(nan_count, 1u);
    rgba[i] = make_uchar4(srgb8_from_linear(r), srgb8_from_linear(g), srgb8_from_linear(b), 255);
}

// device address of g_film_static_scratch on `device` (current): looked up once per device
static int film_static_scratch(int device, unsigned char **out) {
    constexpr int kMaxDevices = 64;
    static unsigned char *ptr[kMaxDevices] = {};
    static std::mutex lock;
    if (device < 0 || device >= kMaxDevices) {
        set_error("b200pt_film_read_srgb8_static: device %d out of range", device);
        return B200PT_ERR_INVALID;
    }
    std::lock_guard<std::mutex> g(lock);
    if (!ptr[device]) {
        unsigned char *p = nullptr;
        CU_TRY(cudaGetSymbolAddress((void **)&p, g_film_static_scratch));
        ptr[device] = p;
    }
    *out = ptr[device];
    return B200PT_OK;
}
} // namespace b200pt

B200PT_API int b200pt_film_read_srgb8_static(b200pt_film *film, uint8_t *rgba_out, int32_t *nan_pixels_out) {
    if (!rgba_out) return B200PT_ERR_INVALID;
    if (int rc = film_alive(film, "b200pt_film_read_srgb8_static")) return rc;
    b200pt_scene *s = film->scene;
    const long long n = (long long)film->width * film->height;
    if (n > kStaticFilmMaxPixels) {
        set_error("b200pt_film_read_srgb8_static: the film has %lld pixels, the static staging array holds %lld (use b200pt_film_read_srgb8)",
                  n, kStaticFilmMaxPixels);
        return B200PT_ERR_INVALID;
    }
    CU_TRY(cudaSetDevice(s->device));
    unsigned char *scratch = nullptr;
    if (int rc = film_static_scratch(s->device, &scratch)) return rc;
    Mat3 M;
    for (int k = 0; k < 9; k++) M.m[k] = s->render->out_rgb[k];
    unsigned h_nan = 0;
    CU_TRY(cudaMemsetAsync(scratch, 0, sizeof(unsigned), s->stream));
    if (n > 0) {
        k_film_srgb8_static<<<unsigned((n + 255) / 256), 256, 0, s->stream>>>(film->d_accum, n, M, scratch);
        CU_TRY(cudaGetLastError());
        CU_TRY(cudaMemcpyAsync(rgba_out, scratch + kStaticFilmHeader, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
    }
    CU_TRY(cudaMemcpyAsync(&h_nan, scratch, sizeof(unsigned), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    if (nan_pixels_out) *nan_pixels_out = (int32_t)h_nan;
    return B200PT_OK;
}

// ncclReduce(sum, fp32, root) of the 4*W*H accumulators.  NCCL is resolved at run
// time from the hosting process (torch.distributed has it loaded) or libnccl.so.2: libb200pt.so
// itself has no link-time dependency on it.
B200PT_API int b200pt_film_reduce(b200pt_film *film, void *comm, int root, void *stream) {
    if (!film || !comm) {
        set_error("b200pt_film_reduce: null argument");
        return B200PT_ERR_INVALID;
    }
    if (int rc = film_alive(film, "b200pt_film_reduce")) return rc;
    typedef int (*nccl_reduce_t)(const void *, void *, size_t, int, int, int, void *, cudaStream_t);
    static nccl_reduce_t fn = nullptr;
    if (!fn) {
        fn = (nccl_reduce_t)dlsym(RTLD_DEFAULT, "ncclReduce");
        if (!fn) {
            void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
            if (h) fn = (nccl_reduce_t)dlsym(h, "ncclReduce");
        }
        if (!fn) {
            set_error("b200pt_film_reduce: ncclReduce not found (load NCCL in the hosting process)");
            return B200PT_ERR_NCCL;
        }
    }
    b200pt_scene *s = film->scene;
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
    if (st != s->stream) CU_TRY(cudaStreamSynchronize(s->stream)); // the render that filled the film ran on the scene's stream
    const size_t count = (size_t)film->width * film->height * 4;
    // ncclFloat32 = 7, ncclSum = 0
    const int rc = fn(film->d_accum, film->d_accum, count, 7, 0, root, comm, st);
    if (rc != 0) {
        set_error("b200pt_film_reduce: ncclReduce failed with %d", rc);
        return B200PT_ERR_NCCL;
    }
    CU_TRY(cudaStreamSynchronize(st));
    return B200PT_OK;
}

B200PT_API void *b200pt_film_device_ptr(b200pt_film *film) { return (film && film->scene) ? (void *)film->d_accum : nullptr; }
B200PT_API int b200pt_film_dims(b200pt_film *film, int32_t *w, int32_t *h) {
    if (!film) return B200PT_ERR_INVALID;
    if (w) *w = film->width;
    if (h) *h = film->height;
    return B200PT_OK;
}

B200PT_API int b200pt_probe_sampler(b200pt_scene *s, int kind, int spp, const int32_t *pixel, const int32_t *sample,
                                    int npairs, const int32_t *pattern, int ndraw, float *out) {
    if (!s || npairs < 0 || ndraw < 0 || (npairs > 0 && (!pixel || !sample || !pattern || !out))) {
        set_error("b200pt_probe_sampler: null argument");
        return B200PT_ERR_INVALID;
    }
    if (npairs == 0 || ndraw == 0) return B200PT_OK;
    CU_TRY(cudaSetDevice(s->device));
    int per = 0;
    for (int i = 0; i < ndraw; i++) per += pattern[i] == 1 ? 1 : 2;
    int32_t *d_pixel = nullptr, *d_sample = nullptr, *d_pattern = nullptr;
    float *d_out = nullptr;
    auto cleanup = [&]() {
        cudaFree(d_pixel);
        cudaFree(d_sample);
        cudaFree(d_pattern);
        cudaFree(d_out);
    };
    cudaError_t e = cudaMalloc(&d_pixel, npairs * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_sample, npairs * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_pattern, ndraw * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, (size_t)npairs * per * 4);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pixel, pixel, npairs * 4, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_sample, sample, npairs * 4, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pattern, pattern, ndraw * 4, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) {
        const int sqrt_spp = (int)std::lround(std::sqrt((double)spp));
        k_probe_sampler<<<(npairs + 127) / 128, 128, 0, s->stream>>>(kind, spp, sqrt_spp, d_pixel, d_sample, npairs, d_pattern,
                                                                    ndraw, per, d_out);
        e = cudaMemcpyAsync(out, d_out, (size_t)npairs * per * 4, cudaMemcpyDeviceToHost, s->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    cleanup();
    if (e != cudaSuccess) return cuda_fail(e, "b200pt_probe_sampler", __FILE__, __LINE__);
    return B200PT_OK;
}
