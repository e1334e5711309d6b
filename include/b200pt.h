/*
 * b200pt.h -- C-ABI of libb200pt.so, the B200-native (sm_100a) wavefront path-tracing core that
 * replaces the reference's HLBVH accelerator and WavefrontPathIntegrator.
 *
 * The reference (w3ntao/pbrt-minus) has no FFI: its host front-end reaches the hot path at two
 * C++ call sites, and this header is what a binding at those sites calls instead:
 *
 *   scene/scene_builder.cu:912        allocator.create<HLBVH>(gpu_primitives, "for ROOT", ...)
 *                                       -> b200pt_scene_create + b200pt_bvh_build
 *   scene/scene_builder.cu:1036-1043  wavefront_path_integrator->render(film, preview);
 *                                     film->write_to_png(output_filename)
 *                                       -> b200pt_film_create + b200pt_render + b200pt_film_read
 *   accelerator/hlbvh.h:96-100        HLBVH::intersect / fast_intersect (per ray, device side)
 *                                       -> b200pt_trace / b200pt_trace_device (batched, ids out)
 *
 * Conventions: plain pointers and sizes only; host arrays in (copied, caller keeps ownership),
 * opaque handles out; one b200pt_scene per GPU, driven by one host thread; every call returns 0
 * or a negative b200pt_status and never exits, throws or falls back to the CPU (the reference
 * prints and exit(1)s, gpu/macro.h:23-48).  b200pt_last_error() gives the thread-local message.
 * Pointers whose name ends in _dev are DEVICE pointers on the scene's GPU.
 */
#ifndef B200PT_H
#define B200PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PT_ABI_VERSION 1

typedef enum {
    B200PT_OK = 0,
    B200PT_ERR_INVALID = -1,   /* bad argument / inconsistent description */
    B200PT_ERR_CUDA = -2,      /* a CUDA runtime call failed; message has the CUDA error string */
    B200PT_ERR_NO_DEVICE = -3, /* no usable sm_100 device: there is NO CPU fallback */
    B200PT_ERR_STATE = -4,     /* call order violated (e.g. trace before bvh_build) */
    B200PT_ERR_NCCL = -5,
    B200PT_ERR_UNSUPPORTED = -6
} b200pt_status;

typedef struct b200pt_scene b200pt_scene; /* owns all device memory of one GPU's scene replica */
typedef struct b200pt_film b200pt_film;   /* W*H*{sum_x,sum_y,sum_z,sum_w} fp32 on device     */

/* ---- spectra (base/spectrum.cu:122-201) ------------------------------------------------- */
enum {
    B200PT_SPEC_CONSTANT = 0,       /* c[0]                                  spectra/constant_spectrum.h */
    B200PT_SPEC_RGB_ALBEDO = 1,     /* sigmoid(c0 l^2 + c1 l + c2)           spectra/rgb_albedo_spectrum.h */
    B200PT_SPEC_RGB_UNBOUNDED = 2,  /* c[3] * sigmoid(...)                   spectra/rgb_unbounded_spectrum.h */
    B200PT_SPEC_RGB_ILLUMINANT = 3, /* c[3] * sigmoid(...) * illuminant(l)   spectra/rgb_illuminant_spectrum.cu */
    B200PT_SPEC_PIECEWISE = 4,      /* n (lambda,value) knots, lerp          spectra/piecewise_linear_spectrum.cu:82-103 */
    B200PT_SPEC_DENSE = 5           /* n = 471 values at 360..830 nm, 1 nm   spectra/densely_sampled_spectrum.h */
};
typedef struct {
    int32_t type;
    float c[4];           /* constant value, or sigmoid c0,c1,c2 and scale */
    int32_t n;            /* knots (PIECEWISE) or samples (DENSE) */
    const float *lambda;  /* [n] PIECEWISE only */
    const float *value;   /* [n] */
} b200pt_spectrum;

/* ---- materials (base/material.cu:97-131) ------------------------------------------------ */
enum { B200PT_MAT_DIFFUSE = 0, B200PT_MAT_CONDUCTOR = 1, B200PT_MAT_DIELECTRIC = 2, B200PT_MAT_DIFFUSE_TRANSMISSION = 3,
       B200PT_MAT_COATED_DIFFUSE = 4, B200PT_MAT_COATED_CONDUCTOR = 5, B200PT_MAT_COUNT = 6 /* BxDF-bearing types */,
       B200PT_MAT_MIX = 6 /* resolved to one of its two materials at hit time (materials/mix_material.cu) */,
       B200PT_MAT_INTERFACE = 7 /* Material "interface": no material, the surface only separates media (base/material.cu:34-36) */ };
typedef struct {
    int32_t type;
    int32_t reflectance; /* spectrum id: diffuse reflectance, or conductor "reflectance" (-1: use eta/k) */
    int32_t eta;         /* spectrum id: conductor eta / dielectric eta */
    int32_t k;           /* spectrum id: conductor k */
    float u_roughness, v_roughness;
    int32_t remap_roughness; /* materials/conductor_material.cu:52-75: alpha = sqrt(roughness) */
    int32_t transmittance;   /* spectrum id: diffuse-transmission "transmittance" (materials/diffuse_transmission_material.cu:8-31) */
    float scale;             /* diffuse-transmission "scale" */
    /* coated diffuse / coated conductor (materials/coated_diffuse_material.cu, coated_conductor_material.cu; LayeredBxDF):
     * u/v_roughness = the dielectric interface's roughness, eta = the interface's eta spectrum id,
     * reflectance = the diffuse base (coated diffuse) or the conductor's "reflectance" (-1: use c_eta / c_k) */
    float thickness, g;
    int32_t albedo;          /* spectrum id of the medium albedo between the layers */
    int32_t max_depth, n_samples;
    float c_u_roughness, c_v_roughness; /* coated conductor: the conductor's roughness */
    int32_t c_eta, c_k;                 /* coated conductor: the conductor's eta / k spectrum ids */
    int32_t mix_a, mix_b;               /* mix: material ids; `scale` = "amount" in (0, 1): u < amount picks mix_a */
    /* spectrum TEXTURES (base/spectrum_texture.cu) for "reflectance" / "transmittance": texture id + 1, 0 = none (the spectrum
     * id above is used).  diffuse, diffusetransmission, coateddiffuse, and conductor / coatedconductor in reflectance mode */
    int32_t reflectance_tex1, transmittance_tex1;
    /* FLOAT textures (base/float_texture.cu) for "uroughness" / "vroughness" ("roughness" sets both) of conductor and dielectric
     * materials (materials/conductor_material.cu:35-44, dielectric_material.cu:20-31): texture id + 1, 0 = the constants above */
    int32_t u_roughness_tex1, v_roughness_tex1;
    /* spectrum textures for the conductor's "eta" / "k" (coated conductor: "conductor.eta" / "conductor.k"): texture id + 1, 0 = the
     * spectrum ids above.  The reference evaluates these as Unbounded spectra (materials/conductor_material.cu:52-75); its image
     * textures abort for that type (textures/spectrum_image_texture.cu:23-33), so the graph holds constant / scale / mix /
     * directionmix nodes (mix amounts may be float IMAGE textures: a spatially varying blend of two metals) */
    int32_t eta_tex1, k_tex1;
} b200pt_material;

/* ---- textures (textures/, base/{spectrum,float}_texture.cu) ---------------------------------------------------------------
 * A texture is a node of a small expression tree over images and constants, evaluated at the hit's (u, v) and normal.
 * Image textures find (s, t) through a uv, planar, spherical or cylindrical mapping (textures/texture_mapping_2d.cu); images are looked up bilinearly with "repeat" wrapping
 * (MIPMap::filter = GPUImage::bilerp, textures/mipmap.cu:22-25), flipped in t, scaled, optionally inverted, clamped,
 * and turned into an RGBAlbedoSpectrum per lookup (textures/spectrum_image_texture.cu:17-32). */
enum { B200PT_TEX_CONSTANT = 0,       /* a: spectrum id */
       B200PT_TEX_IMAGE = 1,          /* a: image id; f[0] scale, f[1] invert (0/1); map = su sv du dv */
       B200PT_TEX_SCALE = 2,          /* a: texture; f[0] scale (textures/spectrum_scaled_texture.cu) */
       B200PT_TEX_MIX = 3,            /* a, b: textures; c: float texture of the amount or -1, then f[0] (spectrum_mix_texture.cu) */
       B200PT_TEX_DIRECTION_MIX = 4,  /* a, b: textures; f[0..2]: unit direction in render space (spectrum_direction_mix_texture.cu) */
       B200PT_TEX_FLOAT_CONSTANT = 16, /* f[0] */
       B200PT_TEX_FLOAT_IMAGE = 17,    /* a: image id (red channel... see float_image_texture.h); f[0] scale, f[1] invert */
       B200PT_TEX_FLOAT_SCALE = 18 };  /* a, b: float textures, product (textures/float_scaled_texture.cu) */
enum { B200PT_MAP_UV = 0, B200PT_MAP_PLANAR = 1, B200PT_MAP_SPHERICAL = 2, B200PT_MAP_CYLINDRICAL = 3 };
typedef struct {
    int32_t type, a, b, c;
    float f[4];
    float map[4]; /* UVMapping: su, sv, du, dv;  PlanarMapping: -, -, udelta, vdelta (textures/texture_mapping_2d.cu) */
    int32_t mapping;               /* B200PT_MAP_*: how an image texture finds (s, t) at a hit */
    float texture_from_render[12]; /* planar / spherical / cylindrical: rows 0..2 of the inverse of the Texture directive's CTM */
    float v1[3], v2[3];            /* PlanarMapping: s = udelta + p . v1, t = vdelta + p . v2 */
} b200pt_texture;

typedef struct {
    int32_t width, height;
    const float *rgb; /* [height*width*3] linear RGB, top row first (GPUImage::init_png, textures/gpu_image.cu:194-222) */
} b200pt_image;

/* ---- lights (base/light.cu:22-75), in the order the scene file creates them (= light-sampler index):
 *      one DiffuseAreaLight per emissive triangle; DistantLight (lights/distant_light.cu), UniformInfiniteLight
 *      (lights/uniform_infinite_light.cu) and SpotLight (lights/spot_light.cu) from "LightSource" */
enum { B200PT_LIGHT_AREA = 0, B200PT_LIGHT_DISTANT = 1, B200PT_LIGHT_UNIFORM_INFINITE = 2, B200PT_LIGHT_SPOT = 3,
       B200PT_LIGHT_IMAGE_INFINITE = 4, /* lights/image_infinite_light.cu: `image` in the equal-area octahedral parameterisation;
                                          needs rgb2spec_fit; at most one per scene */
       B200PT_LIGHT_SPHERE_AREA = 5 /* DiffuseAreaLight over Shape "sphere" (Sphere::sample / Sphere::pdf, shapes/sphere.cu:162-280):
                                       `triangle` holds the index into the spheres array, whose `light` names this light back */ };
typedef struct {
    int32_t triangle;  /* AREA: index into the triangle array; SPHERE_AREA: index into the spheres array; other types: -1 */
    int32_t spectrum;  /* Lemit / I spectrum id */
    float scale;       /* already divided by the photometric integral (and times illuminance / power factors) */
    int32_t two_sided;
    int32_t type;      /* B200PT_LIGHT_* */
    float v[3];        /* DISTANT: render_from_light(0,0,1) (not normalised); SPOT: render_from_light(0,0,0) */
    float light_from_render[9]; /* SPOT: upper 3x3 of the inverse transform, row-major (Transform::apply_inverse) */
    float cos_falloff_start, cos_falloff_end; /* SPOT */
    int32_t image;               /* IMAGE_INFINITE: image id */
    float render_from_light[9];  /* IMAGE_INFINITE: upper 3x3 of the transform (light_from_render holds its inverse) */
} b200pt_light;

/* ---- spheres (shapes/sphere.h:71-125); intersected in object space ----------------------- */
typedef struct {
    float render_from_object[16]; /* row-major 4x4 */
    float object_from_render[16];
    float radius, z_min, z_max, phi_max; /* phi_max in radians; theta range derived from z */
    int32_t reverse_orientation, transform_swaps_handedness;
    int32_t material; /* material id */
    int32_t light;    /* -1: not emissive; else the index of its B200PT_LIGHT_SPHERE_AREA light (spheres only, not disks) */
    int32_t kind;     /* 0 sphere (shapes/sphere.{h,cu}); 1 disk (shapes/disk.{h,cu}): z_min = height, z_max = inner radius */
    uint32_t medium;  /* the quadric's medium word, as b200pt_scene_desc.tri_medium has one per triangle: 0 = SimplePrimitive,
                       * 0xffffffff = GeometricPrimitive without a transition, else (interior + 1) | (exterior + 1) << 16 */
} b200pt_sphere;

/* ---- camera (cameras/perspective.cu:7-86) ------------------------------------------------ */
typedef struct {
    float camera_from_raster[16]; /* row-major 4x4 */
    float render_from_camera[16];
    float lens_radius, focal_distance;
    int32_t medium;               /* CameraBase::medium (cameras/camera_base.h:99): medium id of camera rays, -1 = vacuum */
} b200pt_camera;

/* ---- homogeneous media (medium/homogeneous_medium.{h,cu}): sigma_a / sigma_s spectra times "scale", HG phase function */
typedef struct {
    int32_t sigma_a, sigma_s; /* spectrum ids */
    float scale, g;
} b200pt_medium;

/* ---- pixel filter, tabulated for importance sampling (filters/filter_sampler.{h,cu}) ----- */
typedef struct {
    float radius_x, radius_y;
    int32_t nx, ny;     /* table resolution: 32*radius per axis */
    const float *f;     /* [ny*nx] filter values at the cell centres (signed) */
} b200pt_filter;

/* ---- film / sensor (film/rgb_film.cu:27-47, film/pixel_sensor.h:19-26) ------------------- */
typedef struct {
    int32_t width, height;
    const float *cie_x, *cie_y, *cie_z; /* 471 samples, 360..830 nm (sensor = CIE 1931 XYZ) */
    float imaging_ratio;
    float output_rgb_from_sensor_rgb[9]; /* row-major 3x3 */
    const float *illuminant;             /* 471 samples: colour-space illuminant (D65), normalised */
} b200pt_film_desc;

/* ---- object instancing (ObjectBegin / ObjectInstance, scene/scene_builder.cu:809-876; primitives/transformed_primitive.cu):
 *      an object is a contiguous range of the triangle array whose vertices are in the OBJECT's space; it gets its own BVH8, built
 *      once, and every instance reaches it through a ray transformed by the reference's Transform::apply_inverse(Ray, &t_max) and
 *      brings the hit back with Transform::operator()(SurfaceInteraction).  Triangles before the first object range are the
 *      top-level (render-space) triangles.  Object triangles cannot be emissive or sit on a medium transition
 *      (scene_builder.cu:651-656). */
typedef struct {
    int64_t first_triangle, n_triangles;
    /* the object's bounding box in its own space as the reference's per-object HLBVH has it (the union of its primitives' bounds,
     * quadrics of the definition included): lo xyz, hi xyz.  lo[0] > hi[0] = derive it from the triangles.  An instance enters
     * the scene bounds as the box of this box under its transform (TransformedPrimitive::bounds), and the scene radius scales
     * the power of infinite and distant lights, i.e. the light sampler's probabilities. */
    float bounds[6];
} b200pt_object;
typedef struct {
    int32_t object;
    float render_from_instance[16]; /* row-major 4x4 */
    float instance_from_render[16];
} b200pt_instance;

typedef struct {
    int64_t n_vertices;
    const float *P;            /* [n_vertices*3] render-space positions (object space for the triangles of an object) */
    const float *N;            /* optional [n_vertices*3] shading normals */
    const float *uv;           /* optional [n_vertices*2] */
    int64_t n_triangles;
    const int32_t *indices;    /* [n_triangles*3] */
    const int32_t *tri_material; /* optional [n_triangles]; default 0 */
    const int32_t *tri_light;    /* optional [n_triangles]; -1 = not emissive */
    const uint8_t *tri_flip;     /* optional [n_triangles]; bit 0: reverse_orientation ^ swaps_handedness, bit 1: the triangle's
                                  * mesh has shading normals (N), bit 2: it has uvs (shapes/triangle.cu:336-346,388-463) */
    int32_t n_spheres;
    const b200pt_sphere *spheres;
    int32_t n_spectra;
    const b200pt_spectrum *spectra;
    int32_t n_materials;
    const b200pt_material *materials;
    int32_t n_lights;
    const b200pt_light *lights;
    int32_t n_media;
    const b200pt_medium *media;
    /* optional [n_triangles], media scenes only: 0 = SimplePrimitive (a spawned ray keeps the medium of the incoming ray,
     * primitives/simple_primitive.h:28-37); 0xffffffff = GeometricPrimitive without a medium transition (spawned rays get no
     * medium, primitives/geometric_primitive.cu:20-29); else a MediumInterface: (interior + 1) | (exterior + 1) << 16, 0 = vacuum */
    const uint32_t *tri_medium;
    int32_t n_textures;
    const b200pt_texture *textures;
    int32_t n_images;
    const b200pt_image *images;
    /* needed when n_images > 0: the Simpson tables of the RGB-to-spectrum fit in fp64, [3*283 rgb_tbl | 3 white | 283 lambda01]
     * (spectrum_util/rgb_to_spectrum_data.cu:172-212); the coefficient table itself is fitted on the device */
    const double *rgb2spec_fit;
    const b200pt_camera *camera; /* may be NULL for traversal-only scenes */
    const b200pt_filter *filter;
    const b200pt_film_desc *film;
    int32_t n_objects;           /* instanced objects: ranges [first, first + n) tile the END of the triangle array, in order */
    const b200pt_object *objects;
    int32_t n_instances;
    const b200pt_instance *instances;
} b200pt_scene_desc;

typedef struct {
    int64_t n_triangles;
    int64_t n_bvh2_nodes;
    int64_t n_bvh8_nodes;
    int32_t key_bits;      /* Morton bits: 30 (N <= 16Mi, the reference's code) else 3*floor((64-ceil(log2 N))/3) */
    int32_t sort_passes;
    float ms_bounds_morton, ms_sort, ms_emit, ms_refit, ms_collapse, ms_total; /* device events */
    int64_t device_bytes;  /* resident accelerator + geometry bytes after the build */
    float ms_wall;         /* host wall clock of the whole call, allocations included */
    int32_t collapse_levels;
} b200pt_build_stats;

typedef struct {
    int32_t spp_total;      /* sampler is configured for the TOTAL spp (strata stay global) */
    int32_t sample_begin;   /* this GPU renders sample_idx in [sample_begin, sample_end) */
    int32_t sample_end;
    int32_t max_depth;      /* Integrator "integer maxdepth" (5), integrators/wavefront_path.cu:589 */
    int32_t regularize;     /* "bool regularize" (false) */
    int32_t sampler;        /* 0 = stratified, 1 = independent (scene/scene_builder.cu:392-407) */
    int32_t pool_size;      /* path-pool slots; 0 = default */
    int32_t reserved;
} b200pt_render_opts;

typedef struct {
    int64_t paths;          /* camera paths generated = W*H*(sample_end-sample_begin) */
    int64_t rays_closest;   /* closest-hit rays traced */
    int64_t rays_shadow;    /* any-hit (NEE visibility) rays traced */
    int64_t passes;         /* wavefront iterations */
    int64_t kernel_launches;
    float ms_total;         /* device events around the whole render */
    float ms_trace, ms_shade, ms_generate, ms_logic; /* per-stage sums (0 when not profiled) */
} b200pt_render_stats;

/* Library / device */
int b200pt_abi_version(void);
const char *b200pt_last_error(void);
int b200pt_device_count(void);

/* Scene + accelerator: replaces HLBVH::HLBVH (accelerator/hlbvh.cu:172-404) */
int b200pt_scene_create(const b200pt_scene_desc *desc, int device, b200pt_scene **out);
/* Debugging aid: writes every field of a scene description to `dir` (which must exist) as raw arrays, one file per field, so that
 * the descriptions of two front-ends -- the Python one and the reference's own through the bridge -- can be compared bit by bit
 * (tools/desc_diff.py).  Makes no CUDA call; b200pt_scene_create calls it when B200PT_DESC_DUMP names a directory. */
int b200pt_debug_dump_desc(const b200pt_scene_desc *desc, const char *dir);
void b200pt_scene_destroy(b200pt_scene *scene);
int b200pt_bvh_build(b200pt_scene *scene, b200pt_build_stats *stats_out);
/* Morton codes of the build (reference-compatible 30-bit codes, accelerator/hlbvh.cu:128-142):
 * sorted primitive order and keys, for parity tests of the build stages. */
int b200pt_bvh_debug_sorted(b200pt_scene *scene, int32_t *prim_order_out, uint64_t *keys_out);
/* Structural check of the built BVH8 (all trees of the scene): out[0] = child boxes that do NOT contain what they bound (the
 * vertices of a leaf child's triangles; an inner child's own child boxes, up to that child's grid step), out[1] = triangle
 * slots referenced by leaf children (must equal n_triangles, plus n_instances in a scene with instances: the leaf records of
 * the tree over the instance boxes are one proxy per instance), out[2] = nodes visited.  A conservative tree has out[0] == 0. */
int b200pt_bvh_validate(b200pt_scene *scene, int64_t out[3]);
/* Host walk statistics of the built BVH8 (node / triangle visit counts for a ray set are
 * accumulated on the device by b200pt_trace when counters_out is non-NULL). */

/* Traversal: replaces HLBVH::intersect / fast_intersect (accelerator/hlbvh.cu:407-516).
 * Host-pointer form (copies in/out inside the call) and device-pointer form.
 * rays: o_xyz[n*3], d_xyz[n*3], tmax[n].  Outputs (any may be NULL): prim_out[n] = primitive
 * index (triangle index, or n_triangles + sphere index; -1 = miss; any_hit: 1/0),
 * t_out[n] (inf on miss), b1b2_out[n*2] barycentrics b1,b2 of the hit. */
int b200pt_trace(b200pt_scene *scene, const float *o_xyz, const float *d_xyz, const float *tmax,
                 int64_t n, int any_hit, int32_t *prim_out, float *t_out, float *b1b2_out);
/* Several ray sets in ONE call (an integrator's closest-hit + shadow batches of a bounce): the sets share one
 * copy-in / traverse / copy-out pipeline, so the copy-in of a set overlaps the traversal and the copy-out of the
 * one before it and the pipeline fills and drains once per call instead of once per set.  Same per-set meaning as
 * b200pt_trace; outputs passed as NULL are not copied back (an any-hit caller usually wants prim_out only). */
typedef struct {
    const float *o_xyz, *d_xyz, *tmax; /* [n*3], [n*3], [n] */
    int64_t n;
    int32_t any_hit;
    int32_t reserved;
    int32_t *prim_out; /* [n] or NULL */
    float *t_out;      /* [n] or NULL */
    float *b1b2_out;   /* [n*2] or NULL */
} b200pt_ray_set;
int b200pt_trace_batch(b200pt_scene *scene, const b200pt_ray_set *sets, int n_sets);
/* Device form: rays_dev = n x {ox,oy,oz,tmax, dx,dy,dz,0} (32 B), hits_dev = n x
 * {prim (int bits), t, b1, b2} (16 B).  Asynchronous on the scene's stream unless sync != 0. */
int b200pt_trace_device(b200pt_scene *scene, const void *rays_dev, int64_t n, int any_hit,
                        void *hits_dev, int sync);
/* Several device-resident ray sets in one call (same record formats as b200pt_trace_device): the sets run on the scene's two
 * traversal streams in turn, so the tail of one set's persistent kernel overlaps the start of the next; ordered on the scene's
 * stream like one operation.  Asynchronous unless sync != 0. */
typedef struct {
    const void *rays_dev;
    int64_t n;
    int32_t any_hit, reserved;
    void *hits_dev;
} b200pt_device_ray_set;
int b200pt_trace_device_batch(b200pt_scene *scene, const b200pt_device_ray_set *sets, int32_t n_sets, int sync);
/* Per-launch traversal counters of the most recent trace with counting enabled:
 * out[0] = BVH8 nodes fetched, out[1] = triangles tested, out[2] = rays. */
int b200pt_trace_set_counting(b200pt_scene *scene, int enable);
int b200pt_trace_counters(b200pt_scene *scene, int64_t out[3]);
/* Scheduling knobs of the traversal kernels (they change speed, never results): key 0 = idle lanes before a warp
 * of k_trace refills (1/4/8/16), 1 = triangle tests per triangle phase (1..3), 2 = resident CTAs per SM cap
 * (0 = occupancy limit), 3 / 4 = the same two knobs for the wavefront integrator's trace kernels (0 = default),
 * 5 = resident CTAs per SM the plain shade kernels' register allocation is held to (0 = default, 3, 4),
 * 6 = k_trace with a one-deep triangle stash (0 / 1), 7 = MB of BVH-build scratch a scene keeps between builds (default 16384;
 * a rebuild reuses the previous build's scratch, node and leaf-record arrays while they fit; 0 = allocate and free per build,
 * and a value below what is held releases it at once). */
int b200pt_trace_tune(b200pt_scene *scene, int key, int value);
void *b200pt_scene_stream(b200pt_scene *scene); /* cudaStream_t the scene launches on */
int b200pt_scene_sync(b200pt_scene *scene);

/* Device buffers for callers without their own CUDA allocator (tests, bench, host front-end) */
int b200pt_device_alloc(b200pt_scene *scene, size_t bytes, void **out_dev);
int b200pt_device_free(b200pt_scene *scene, void *ptr_dev);
int b200pt_copy_to_device(b200pt_scene *scene, void *dst_dev, const void *src, size_t bytes);
int b200pt_copy_to_host(b200pt_scene *scene, void *dst, const void *src_dev, size_t bytes);

/* Wavefront path integrator: replaces WavefrontPathIntegrator::render
 * (integrators/wavefront_path.cu:594-735) and RGBFilm (film/rgb_film.cu:63-122). */
int b200pt_film_create(b200pt_scene *scene, b200pt_film **out);
void b200pt_film_destroy(b200pt_film *film);
int b200pt_film_clear(b200pt_film *film);
int b200pt_render(b200pt_scene *scene, b200pt_film *film, const b200pt_render_opts *opts,
                  b200pt_render_stats *stats_out);
/* Sum of the per-GPU films onto root: ncclReduce(sum, fp32) of 4*W*H floats.  nccl_comm is an
 * ncclComm_t, stream a cudaStream_t (NULL = the film's scene stream). */
int b200pt_film_reduce(b200pt_film *film, void *nccl_comm, int root, void *stream);
void *b200pt_film_device_ptr(b200pt_film *film); /* float4[W*H] accumulators (for torch.distributed) */
int b200pt_film_dims(b200pt_film *film, int32_t *width, int32_t *height);
/* rgb_out[W*H*3] = output_rgb_from_sensor_rgb * (xyz_sum / weight_sum) (film/rgb_film.cu:107-122) */
int b200pt_film_read(b200pt_film *film, float *rgb_out);
/* The 8-bit image Film::write_to_png encodes (base/film.cu:57-99): rgba_out[W*H*4] = SRGBColorEncoding::from_linear of the
 * output RGB (the reference's minimax sRGB curve, spectrum_util/color_encoding.h:6-45), alpha 255; converted on the device.
 * nan_pixels_out (may be NULL): pixels with a NaN component (the reference aborts on them; here they are written as 0). */
int b200pt_film_read_srgb8(b200pt_film *film, uint8_t *rgba_out, int32_t *nan_pixels_out);
/* No counterpart in the reference: added for callers that may not allocate device
 * memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as b200pt_film_read_srgb8,
 * but the 8-bit image is staged in a file-scope device array of the library instead of the scene's growable scratch buffer:
 * the call makes no cudaMalloc / cudaFree.  Limit: films of at most 2,097,152 pixels (1920x1080 fits); a larger film is refused
 * with B200PT_ERR_INVALID.  Every GPU that loads the library holds the array (8 MiB + 256 B) until the process exits; all calls
 * share it, so two calls that overlap on one GPU race -- serialise them. */
int b200pt_film_read_srgb8_static(b200pt_film *film, uint8_t *rgba_out, int32_t *nan_pixels_out);
/* raw accumulators: xyzw_out[W*H*4] */
int b200pt_film_read_raw(b200pt_film *film, float *xyzw_out);

/* Stage probes for parity tests (each runs the production device functions on small batches).
 * Sampler (samplers/stratified.h, samplers/independent.h): same contract as the oracle's
 * oracle_sampler_stream: for each (pixel, sample) emit draws following pattern[] (1 = get_1d,
 * 2 = get_2d). */
int b200pt_probe_sampler(b200pt_scene *scene, int kind, int spp, const int32_t *pixel,
                         const int32_t *sample, int npairs, const int32_t *pattern, int ndraw,
                         float *out);
/* Triangle test (shapes/triangle.cu:213-323) on explicit (ray, triangle) pairs:
 * o,d [n*3], tmax[n], tri[n*9]; hit_out[n] (0/1), bt_out[n*4] = b0,b1,b2,t */
int b200pt_probe_triangle(b200pt_scene *scene, const float *o, const float *d, const float *tmax,
                          const float *tri, int64_t n, int32_t *hit_out, float *bt_out);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */
