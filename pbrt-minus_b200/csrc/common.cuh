// common.cuh -- shared declarations of libb200pt.so (hand-written CUDA for sm_100a).
// No CPU fallback anywhere: every entry point needs a CUDA device and fails with a status code
// when there is none (the reference prints and exit(1)s instead, gpu/macro.h:23-48).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/b200pt.h"

#define B200PT_API extern "C" __attribute__((visibility("default")))

namespace b200pt {

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

#define CU_TRY(expr)                                                        \
    do {                                                                    \
        cudaError_t _e = (expr);                                            \
        if (_e != cudaSuccess) return ::b200pt::cuda_fail(_e, #expr, __FILE__, __LINE__); \
    } while (0)

constexpr float kInf = __builtin_huge_valf();

// ---- BVH8 node: 80 B = 5 x 128-bit loads ---------------------------------------------------
//  n0 = { p.x, p.y, p.z, bytes(ex, ey, ez, imask) }   origin + per-axis pow-2 scale + inner mask
//  n1 = { child_base, tri_base, meta[0..3], meta[4..7] }
//  n2 = { qlo_x[0..3], qlo_x[4..7], qlo_y[0..3], qlo_y[4..7] }
//  n3 = { qlo_z[0..3], qlo_z[4..7], qhi_x[0..3], qhi_x[4..7] }
//  n4 = { qhi_y[0..3], qhi_y[4..7], qhi_z[0..3], qhi_z[4..7] }
// meta (per child slot): 0 = empty; inner child: 0b001'11sss (low 5 bits = 24 + slot);
// leaf child: high 3 bits = unary triangle count (001, 011, 111), low 5 bits = offset of its
// first triangle from tri_base (0..23).  Child boxes are quantised conservatively (mins rounded
// down, maxes up) on the node's grid  p + q * 2^e.
struct BVH8 {
    const float4 *nodes = nullptr; // 5 float4 per node
    const float4 *tris = nullptr;  // 3 float4 per triangle, BVH8 leaf order; v0.w = primitive id
    int64_t n_nodes = 0;
    int64_t n_tris = 0;
};

// flag word of a leaf-order triangle record (v1.w): material id in the low 28 bits, then
//  bit 28 degenerate triangle (Triangle::intersect_triangle's first reject, shapes/triangle.cu:218, evaluated once at build),
//  bit 29 has per-vertex normals, bit 30 has per-vertex uvs, bit 31 flip the geometric normal
constexpr uint32_t kTriMaterialMask = 0x0fffffffu;
constexpr uint32_t kTriDegenerate = 0x10000000u;

// Device-side sphere record for the (few) quadrics of a scene: tested linearly beside the BVH.
struct SphereRec {
    float rfo[12]; // render_from_object, rows 0..2
    float ofr[12]; // object_from_render
    float radius, z_min, z_max, phi_max;
    float theta_z_min, theta_z_max;
    int flip; // reverse_orientation ^ transform_swaps_handedness
    int material;
    int kind; // 0 sphere; 1 disk: z_min = height, z_max = inner radius
    int light; // -1, or the light id of this emissive sphere
    uint32_t medium; // medium word (b200pt_sphere.medium)
    // render-space AABB of the quadric, padded: a ray that misses it cannot pass the interval-arithmetic test, so the
    // trace kernels skip that (long) test for most rays
    float box_lo[3], box_hi[3];
};

// Device-side record of one object instance: the transforms and where its object's BVH8 starts in the shared node array.
struct InstRec {
    float rfi[12]; // render_from_instance, rows 0..2
    float ifr[12]; // instance_from_render
    int32_t root_node, object;
    float box_lo[3], box_hi[3]; // render-space AABB of the transformed object box, padded
};
struct HostObject {
    int64_t first_triangle = 0, n_triangles = 0;
    unsigned root_node = 0;
    float box[6] = {0, 0, 0, 0, 0, 0}; // object-space bounds of its triangles (filled by the build)
};

struct RenderData; // render.cu

} // namespace b200pt

struct b200pt_scene {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    // b200pt_trace (host buffers): copy-in / copy-out streams and per-chunk events of the three-stage pipeline
    cudaStream_t stream_h2d = nullptr, stream_d2h = nullptr;
    cudaStream_t stream2 = nullptr; // second traversal stream: consecutive chunks alternate, so a chunk's CTAs fill the tail of the one before
    std::vector<cudaEvent_t> pipe_events;

    // geometry as uploaded (render space)
    int64_t n_vertices = 0, n_triangles = 0;
    float *d_P = nullptr;        // [nv*3]
    float *d_N = nullptr;        // [nv*3] or null
    float *d_uv = nullptr;       // [nv*2] or null
    int32_t *d_idx = nullptr;    // [nt*3]
    int32_t *d_tri_material = nullptr;
    int32_t *d_tri_light = nullptr;
    uint8_t *d_tri_flip = nullptr;
    int32_t n_spheres = 0;
    b200pt::SphereRec *d_spheres = nullptr;

    // instancing (b200pt_scene_desc.objects / instances)
    int64_t n_top_triangles = 0;
    unsigned n_top_nodes = 0, n_nodes_total = 0;
    int32_t inst_root_node = -1;   // root of the BVH8 over the instance boxes (proxy leaf records at [n_triangles, +n_instances))
    std::vector<float> inst_boxes; // padded render-space box of every instance, 6 floats each (finish_instances)
    std::vector<b200pt::HostObject> objects;
    std::vector<b200pt_instance> instances;
    b200pt::InstRec *d_instances = nullptr;

    // accelerator
    bool built = false;
    b200pt::BVH8 bvh;
    float4 *d_nodes = nullptr;
    float4 *d_tris = nullptr;
    // debug copies of the sorted build order (kept for parity tests of the build stages)
    uint64_t *d_sorted_keys = nullptr;
    uint32_t *d_sorted_prim = nullptr;
    // a rebuild reuses the node / leaf-record / sorted-order arrays and the build scratch of the previous build while they are
    // large enough: the six cudaFree + cudaMalloc pairs of a 10 M-triangle build cost more host time than the build's kernels run.
    // The scratch is kept up to tune_build_scratch_mb (b200pt_trace_tune key 7; 0 = allocate and free per build).
    size_t d_nodes_bytes = 0, d_tris_bytes = 0;
    int64_t sorted_cap = 0;
    char *build_scratch = nullptr;
    size_t build_scratch_bytes = 0;
    int tune_build_scratch_mb = 16384;
    int key_bits = 0;
    float scene_bounds[6] = {0, 0, 0, 0, 0, 0};

    // traversal
    unsigned long long *d_trace_counters = nullptr; // [0] ray ticket, [1] nodes, [2] tris, [3] err, [4..11] tickets of pipelined chunks
    bool counting = false;
    // k_trace tuning (b200pt_trace_tune): lanes idle before a warp refills, triangle tests per triangle phase, CTA cap per SM
    int tune_refill = 4, tune_tri_steps = 3, tune_ctas_per_sm = 0;
    int tune_wf_refill = 0, tune_wf_tri_steps = 0; // wavefront trace kernels; 0 = the built-in default
    int tune_stash = 0;                            // k_trace with a one-deep triangle stash (k_trace_stash)
    int tune_shade_occ = 0;                        // shade kernels: CTAs per SM the register allocation is held to (0 = default)
    int64_t last_counts[3] = {0, 0, 0};
    int64_t device_bytes = 0;

    b200pt::RenderData *render = nullptr;
    std::vector<void *> user_allocs;
    // grow-only scratch for the host-pointer entry points (no cudaMalloc on the steady-state path)
    void *d_scratch = nullptr;
    size_t scratch_bytes = 0;
};

namespace b200pt {

int build_bvh(b200pt_scene *s, b200pt_build_stats *stats);
int validate_bvh(b200pt_scene *s, int64_t out[3]);
int finish_instances(b200pt_scene *s); // after the build: instance records (object roots and boxes) to the device
// keep_status: leave the node / triangle / overflow counters as they are (a later chunk of the same batch); the work ticket is
// always reset
// caller-layout device arrays of one chunk (b200pt_trace_batch): o[n*3], d[n*3], tmax[n] in; prim[n], t[n], b12[n*2] out (any may be null)
struct TraceHostIO {
    const float *o, *d, *tmax;
    int32_t *prim;
    float *t, *b12;
};
// ticket_slot / stream: a pipelined chunk (b200pt_trace_batch) runs on its own stream with its own work ticket, so that it
// can be resident together with the chunk before it
int launch_trace(b200pt_scene *s, const void *rays_dev, int64_t n, int any_hit, void *hits_dev, bool keep_status = false,
                 const TraceHostIO *io = nullptr, int ticket_slot = -1, cudaStream_t stream = nullptr);
constexpr int kTraceCounterWords = 12, kTraceTicketRing = 8;
void free_render(b200pt_scene *s);
int ensure_scratch(b200pt_scene *s, size_t bytes);

} // namespace b200pt
