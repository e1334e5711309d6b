// wavefront.cuh -- device-resident state of the wavefront path integrator (shared by render.cu and
// the queue-driven trace kernels in trace.cu).
//
// Replaces WavefrontPathIntegrator::PathState / Queues (integrators/wavefront_path.h:28-134): the
// reference keeps 15 AoS arrays (~858 B/slot, a 256-B optional<ShapeIntersection>, a 124-B BSDF and
// a 264-B MISParameter per slot) in managed memory, pushes with one atomicAdd per thread and lets
// the HOST zero and read the queue counters between kernels.  Here the pool is SoA float4 lanes
// (208 B/slot), every queue push is warp-aggregated, and all counters live on the device: a pass
// is a fixed kernel sequence with no host synchronisation.
#pragma once

#include "common.cuh"
#include "light.cuh"
#include "spectrum.cuh"
#include "texture.cuh"
#include "envmap.cuh"

namespace b200pt {

struct MatRec {
    int32_t type, reflectance, eta, k;
    float u_rough, v_rough;
    int32_t remap, transmittance;
    float scale;
    float thickness, g;           // coated materials (layered.cuh)
    int32_t albedo, max_depth, n_samples;
    float c_u_rough, c_v_rough;
    int32_t c_eta, c_k;
    int32_t mix_a, mix_b;         // mix material: the two material ids; `scale` holds the amount
    int32_t reflectance_tex1, transmittance_tex1; // spectrum texture id + 1 (0: the spectrum id is used)
    int32_t u_rough_tex1, v_rough_tex1;           // float texture id + 1 for the roughnesses (0: the constants)
    int32_t eta_tex1, k_tex1;                     // spectrum texture id + 1 for the conductor's eta / k (0: the spectrum ids)
};
constexpr int WF_NMAT = B200PT_MAT_COUNT; // one shading queue per material type

struct MediumRec { // medium/homogeneous_medium.h: two DenselySampledSpectrum(spectrum, scale) + HGPhaseFunction(g)
    int32_t sigma_a, sigma_s;
    float scale, g;
};

struct FilterTab {
    int32_t nx, ny;
    float min_x, max_x, min_y, max_y;
    const float *f;         // [ny*nx] signed filter values
    const float *cond_func; // [ny*nx] |f|
    const float *cond_cdf;  // [ny*(nx+1)]
    const float *cond_int;  // [ny]
    const float *marg_func; // [ny]
    const float *marg_cdf;  // [ny+1]
    float marg_int;
};

struct CameraRec {
    float cfr[16]; // camera_from_raster (projective)
    float rfc[16]; // render_from_camera
    float lens_radius, focal_distance;
};

struct FilmRec {
    int32_t width, height;
    const float *cie_x, *cie_y, *cie_z; // dense 471
    float imaging_ratio;
    float4 *accum;
};

// flags word of a path (ids.w): bit 0 specular_bounce, bit 1 any_non_specular_bounces,
// bits 8..15 depth, bits 16..23 next_time_russian_roulette
constexpr int PF_SPECULAR = 1, PF_ANY_NON_SPECULAR = 2, PF_IN_VOLUME = 4;

// per-slot state byte: what the next stage has to do with the slot
enum : uint8_t { ST_IDLE = 0, ST_RAY = 1, ST_MAT0 = 2 /* + material type */, ST_FREE = ST_MAT0 + WF_NMAT };

struct PathPool {
    float4 *ray_o;  // o.xyz, tmax (< 0: dead path, skipped by the trace kernel)
    float4 *ray_d;  // d.xyz, -
    float4 *hit;    // {leaf-order triangle slot | n_tris + sphere (int bits), t, b1, b2}
    float4 *beta, *L, *lambda, *lpdf;
    float4 *ctx0;   // previous vertex for MIS: p.xyz, err.x
    float4 *ctx1;   //                          n.xyz, err.y
    float4 *ctx2;   //                          ns.xyz, err.z
    float4 *misc;   // filter_weight, prev_direction_pdf, -, -
    float4 *mtp;    // media scenes only: multi_transmittance_pdf (integrators/wavefront_path.h:37)
    float4 *vol;    // media scenes only: depth (a FLOAT: +0.3 per interface bounce), medium id of the ray (int bits), free-flight distance of a queued scatter, -
    uint64_t *rng;  // PCG32 state (inc is re-derived from the pixel index)
    int4 *ids;      // pixel, sample, sampler dimension, flags
    uint8_t *state; // ST_*
    int32_t *inst;  // scenes with object instances only: the instance the hit belongs to (-1: top level)
};

// Queue sizes and work tickets of ONE pass.  There are two sets, used by alternate passes (wf.parity): every count is
// produced and consumed inside its pass, so the logic stage of pass p clears the set of pass p + 1 and no separate
// "reset the counters" launches are needed between the stages.
struct WFPass {
    unsigned n_mat[WF_NMAT];
    unsigned n_free;
    unsigned n_shadow;
    unsigned n_emit; // hits on emissive triangles this pass (k_logic -> k_emit)
    unsigned n_vol;  // media: paths that scatter this pass (k_volume<0> -> k_volume<1>)
    unsigned ticket_closest, ticket_shadow;
};
struct WFCounters {
    WFPass p[2];
    unsigned long long next_path, finished, rays_closest, rays_shadow;
    unsigned stack_overflow;
};

struct WFQueues {
    int32_t *mat[WF_NMAT];
    int2 *mat_pm[WF_NMAT]; // {primitive, resolved material id} of the entry at the same position: the shade kernels start their
                           // pool, triangle and material loads together instead of chasing slot -> hit -> triangle -> material
    int32_t *free_slots;
    int32_t *vol; // media scenes only: slots that scatter inside a medium this pass
    S4 *tr_a, *tr_beta; // media scenes only: Le / pdf_light and beta of the transmittance request at the same queue position
    int32_t *emit; // slots whose ray hit an emissive triangle this pass
    // shadow rays: compact SoA by queue position
    float4 *sh_o, *sh_d, *sh_contrib;
    int32_t *sh_slot;
};

struct WF {
    PathPool pool;
    WFQueues q;
    WFCounters *c;
    int32_t parity; // which WFPass set this pass uses
    int32_t pool_size;
    // scene
    BVH8 bvh;
    const SphereRec *spheres;
    int32_t n_spheres;
    const InstRec *instances; // object instances: each reaches its object's BVH8 through a transformed ray
    int32_t n_instances;
    int32_t inst_root;       // root node of the BVH8 over the instance boxes (-1: no instances)
    // per-vertex shading normals / uvs of the meshes that have them (render space; indexed through the triangle's
    // primitive id and the index buffer); null when no mesh has any
    const float *N, *uv;
    const int32_t *idx;
    // media (null / 0 when the scene has none): table, per-triangle medium code by primitive id (b200pt.h: tri_medium),
    // camera medium
    const MediumRec *media;
    const uint32_t *tri_medium;
    int32_t has_media, camera_medium;
    TexTables tex;           // spectrum / float textures (all null when the scene has none)
    EnvMap env;              // the image infinite light (env.light < 0: none)
    SpectrumTable spectra;
    const MatRec *materials;
    const LightRec *lights;
    const AliasBin *alias;
    const LightExt *light_ext;
    int32_t n_lights;
    int32_t n_infinite;      // infinite lights looked up by escaping rays (integrators/wavefront_path.cu:105-127)
    int32_t infinite[4];     // their light ids
    int32_t refill_min;      // k_trace_wf: a warp fetches new rays once this many lanes are idle
    int32_t tri_steps;       // k_trace_wf: triangle tests a lane runs per triangle phase
    int32_t has_mix;         // some material is a mix (resolved in k_logic, materials/mix_material.cu)
    float scene_radius;      // Bounds3::bounding_sphere of the whole scene (scene_builder.cu:914-925)
    FilterTab filter;
    CameraRec camera;
    FilmRec film;
    // render options
    int32_t sampler_kind, spp, sqrt_spp, max_depth, regularize;
    int32_t sample_begin;
    long long total_paths; // W*H*(sample_end - sample_begin)
    long long n_pixels;
};

// warp-aggregated queue push: one atomicAdd per warp (the reference: one per thread,
// integrators/wavefront_path.h:80-83).  Must be called by all 32 lanes of a converged warp.
__device__ __forceinline__ void queue_push(unsigned *counter, int32_t *queue, int32_t value, bool pred) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0u) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) queue[base + __popc(m & ((1u << lane) - 1u))] = value;
}
// same, returning the claimed position (or -1)
__device__ __forceinline__ int queue_claim(unsigned *counter, bool pred) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0u) return -1;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return pred ? int(base + __popc(m & ((1u << lane) - 1u))) : -1;
}

// block-ordered push into the material queues (cat 0..WF_NMAT-1) or the free list (cat WF_NMAT): the entries of one
// 256-slot block land contiguously and in thread (= slot) order, with ONE atomicAdd per block and queue.  cat < 0: nothing
// to push.  Must be called by every thread of the block (it synchronises); blockDim.x <= 1024.  Queue pointers and counters
// are read from the kernel parameter block by index (constant bank), never through a local array.
__device__ __forceinline__ void block_push(const WF &wf, int cat, int32_t value, int2 pm) {
    constexpr int NCAT = WF_NMAT + 1;
    __shared__ unsigned s_warp[NCAT][32];
    __shared__ unsigned s_base[NCAT];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = (blockDim.x + 31) >> 5;
    unsigned my_ballot = 0;
#pragma unroll
    for (int c = 0; c < NCAT; c++) {
        const unsigned m = __ballot_sync(0xffffffffu, cat == c);
        if (lane == 0) s_warp[c][warp] = __popc(m);
        if (cat == c) my_ballot = m;
    }
    __syncthreads();
    if (threadIdx.x < NCAT) {
        unsigned total = 0;
        for (int w = 0; w < n_warps; w++) {
            const unsigned v = s_warp[threadIdx.x][w];
            s_warp[threadIdx.x][w] = total;
            total += v;
        }
        WFPass &pc = wf.c->p[wf.parity];
        unsigned *counter = threadIdx.x < WF_NMAT ? &pc.n_mat[threadIdx.x] : &pc.n_free;
        s_base[threadIdx.x] = total ? atomicAdd(counter, total) : 0u;
    }
    __syncthreads();
    if (cat >= 0) {
        int32_t *queue = cat < WF_NMAT ? wf.q.mat[cat] : wf.q.free_slots;
        const unsigned pos = s_base[cat] + s_warp[cat][warp] + __popc(my_ballot & ((1u << lane) - 1u));
        queue[pos] = value;
        if (cat < WF_NMAT) wf.q.mat_pm[cat][pos] = pm;
    }
    __syncthreads(); // the shared arrays are reused by the next call
}

int launch_trace_wf(b200pt_scene *s, const WF &wf, bool shadow);

} // namespace b200pt
