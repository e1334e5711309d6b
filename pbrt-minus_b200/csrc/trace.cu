// trace.cu -- closest-hit and any-hit traversal as persistent-thread kernels.
//
// Replaces ray_cast (integrators/wavefront_path.cu:348-366 -> HLBVH::intersect,
// accelerator/hlbvh.cu:459-516) and the inline HLBVH::fast_intersect calls of the shading
// kernels (accelerator/hlbvh.cu:407-456).
//
// Shape of the kernel (the reference launches one thread per ray with a local-memory stack):
//  * persistent grid sized to the machine (SMs x resident CTAs);
//  * warp-aggregated ray fetch: whenever REFILL or more lanes of a warp are idle, ONE atomicAdd
//    on the ray ticket claims a ray for each idle lane (ballot + popc), so finished lanes are
//    replaced instead of idling until the slowest ray of the warp is done;
//  * "while-while" traversal: all lanes first walk interior nodes until they hold triangle work,
//    re-converge, then test triangles together -- the two long code paths (node decode, watertight
//    triangle test) each run with as many active lanes as possible;
//  * short stack of (node-group, hit-mask) pairs in shared memory, overflow in local memory.
// Rays: n x {ox,oy,oz,tmax, dx,dy,dz,-} (2 x 128-bit loads); hits: n x {prim,t,b1,b2} (one
// 128-bit store).
#include <cstdlib>

#include "common.cuh"
#include "sphere.cuh"
#include "traverse.cuh"
#include "wavefront.cuh"

namespace b200pt {

struct TraceParams {
    BVH8 bvh;
    const float4 *rays;
    float4 *hits;
    long long n;
    unsigned long long *counters; // [1] nodes, [2] tris, [3] overflow flag
    unsigned long long *ticket;   // the work ticket of this launch
    uint32_t prmt_bias;           // 0x47000000 (32768.0f), passed as a parameter so PRMT keeps an immediate selector
    // caller-layout I/O (b200pt_trace / b200pt_trace_batch): when `o` is set the kernel reads the caller's o[n*3] / d[n*3] /
    // tmax[n] and writes prim[n] / t[n] / b12[n*2] itself -- no repacking passes on either side of the traversal
    const float *o, *d, *tmax;
    int32_t *prim_out;
    float *t_out, *b12_out;
};

// TRI_STEPS: triangle tests a lane runs per triangle phase (a lane whose leaf slot holds several triangles does not
// need one vote round per triangle)
template <bool ANY, bool COUNT, int REFILL, int TRI_STEPS>
__global__ void __launch_bounds__(TRACE_BLOCK) k_trace(TraceParams p) {
    __shared__ uint2 sm_stack_all[SM_STACK * TRACE_BLOCK];
    uint2 *const sm_stack = sm_stack_all + threadIdx.x;
    uint2 lstack[LOCAL_STACK];
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const bool have_tree = p.bvh.n_nodes > 0;

    long long ray = -1; // this lane's ray index, -1 = idle
    bool exhausted = false;
    float ox = 0, oy = 0, oz = 0, idx = 0, idy = 0, idz = 0;
    RayShear rs = {0.f, 0.f, 0.f, 2}; // ray-only part of the watertight test, evaluated once per ray
    uint32_t octinv4 = 0;
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u);
    int sp = 0;
    float ht = 0, hb1 = 0, hb2 = 0;
    int hprim = -1;
    unsigned n_nodes = 0, n_tris = 0, overflow = 0;

    while (true) {
        // ---- warp-aggregated fetch of new rays into idle lanes
        const unsigned idle = __ballot_sync(0xffffffffu, ray < 0);
        if (!exhausted && (__popc(idle) >= REFILL || idle == 0xffffffffu)) {
            const int n_idle = __popc(idle);
            const int leader = __ffs(idle) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(p.ticket, (unsigned long long)n_idle);
            base = __shfl_sync(0xffffffffu, base, leader);
            if ((long long)base + n_idle >= p.n) exhausted = true;
            if (ray < 0) {
                const long long r = (long long)base + __popc(idle & lt_mask);
                if (r < p.n) {
                    ray = r;
                    float4 ro, rd;
                    if (p.o) {
                        const float *po = p.o + 3 * r, *pd = p.d + 3 * r;
                        ro = make_float4(__ldg(po), __ldg(po + 1), __ldg(po + 2), __ldg(p.tmax + r));
                        rd = make_float4(__ldg(pd), __ldg(pd + 1), __ldg(pd + 2), 0.f);
                    } else {
                        ro = __ldg(p.rays + 2 * r);
                        rd = __ldg(p.rays + 2 * r + 1);
                    }
                    ox = ro.x; oy = ro.y; oz = ro.z;
                    idx = safe_rcp_dir(rd.x); idy = safe_rcp_dir(rd.y); idz = safe_rcp_dir(rd.z);
                    rs = ray_shear(rd.x, rd.y, rd.z);
                    octinv4 = ((idx < 0.f ? 0u : 4u) | (idy < 0.f ? 0u : 2u) | (idz < 0.f ? 0u : 1u)) * 0x01010101u;
                    ng = make_uint2(0u, have_tree ? 0x80000000u : 0u);
                    tg = make_uint2(0u, 0u);
                    sp = 0;
                    ht = ro.w;
                    hb1 = 0.f; hb2 = 0.f;
                    hprim = -1;
                }
            }
        }
        if (__ballot_sync(0xffffffffu, ray >= 0) == 0u) break;

        // ---- phase vote: lanes holding triangle work vs lanes needing node work; the larger
        //      group runs ONE step (one triangle test / one node decode), the other group waits,
        //      so each of the two long code paths always runs with at least half the live lanes.
        const bool act = ray >= 0;
        const bool want_tri = act && tg.y != 0u;
        const unsigned m_tri = __ballot_sync(0xffffffffu, want_tri);
        const unsigned m_node = __ballot_sync(0xffffffffu, act && !want_tri);
        if (__popc(m_tri) >= __popc(m_node)) {
            if (want_tri) {
#pragma unroll 1
                for (int step = 0; step < TRI_STEPS && tg.y != 0u; step++) {
                    const int ti = __ffs(tg.y) - 1;
                    tg.y &= tg.y - 1u;
                    const float4 *tp = p.bvh.tris + 3 * size_t(tg.x + ti);
                    const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                    if (COUNT) n_tris++;
                    TriHit th;
                    if (!(__float_as_uint(v1.w) & kTriDegenerate) &&
                        intersect_triangle_sheared(ox, oy, oz, rs, ht, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z, th)) {
                        ht = th.t;
                        hb1 = th.b1;
                        hb2 = th.b2;
                        hprim = __float_as_int(v0.w);
                        if (ANY) { // first accepted candidate ends the ray
                            tg.y = 0u;
                            ng.y = 0u;
                            sp = 0;
                        }
                    }
                }
            }
        } else if (act && !want_tri) {
            if (ng.y <= 0x00ffffffu && sp > 0) {
                sp--;
                if (sp < SM_STACK) ng = sm_stack[sp * TRACE_BLOCK];
                else if (sp < SM_STACK + LOCAL_STACK) ng = lstack[sp - SM_STACK];
                else ng = make_uint2(0u, 0u);
            }
            if (ng.y > 0x00ffffffu) {
                const uint32_t hits = ng.y;
                const int child_bit = 31 - __clz(hits);
                ng.y &= ~(1u << child_bit);
                if (ng.y > 0x00ffffffu) {
                    if (sp < SM_STACK) sm_stack[sp * TRACE_BLOCK] = ng;
                    else if (sp < SM_STACK + LOCAL_STACK) lstack[sp - SM_STACK] = ng;
                    else overflow = 1u;
                    sp++;
                }
                const uint32_t slot = (uint32_t(child_bit) - 24u) ^ (octinv4 & 0xffu);
                const uint32_t rel = __popc(hits & 0xffu & ((1u << slot) - 1u));
                const float4 *np = p.bvh.nodes + 5 * size_t(ng.x + rel);
                const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
                if (COUNT) n_nodes++;
                const uint32_t hitmask = intersect_node8(n0, n1, n2, n3, n4, ox, oy, oz, idx, idy, idz, octinv4, ht, p.prmt_bias);
                ng.x = __float_as_uint(n1.x);
                tg.x = __float_as_uint(n1.y);
                ng.y = (hitmask & 0xff000000u) | (__float_as_uint(n0.w) >> 24);
                tg.y = hitmask & 0x00ffffffu;
            } else { // nothing left: retire the ray
                if (p.o) {
                    if (p.prim_out) p.prim_out[ray] = ANY ? (hprim >= 0 ? 1 : 0) : hprim;
                    if (p.t_out) p.t_out[ray] = (ANY || hprim < 0) ? (ANY ? 0.f : kInf) : ht;
                    if (p.b12_out) *reinterpret_cast<float2 *>(p.b12_out + 2 * ray) = ANY ? make_float2(0.f, 0.f) : make_float2(hb1, hb2);
                } else if (ANY) {
                    p.hits[ray] = make_float4(__int_as_float(hprim >= 0 ? 1 : 0), 0.f, 0.f, 0.f);
                } else {
                    p.hits[ray] = make_float4(__int_as_float(hprim), hprim >= 0 ? ht : kInf, hb1, hb2);
                }
                ray = -1;
            }
        }
    }
    if (COUNT) {
        atomicAdd(p.counters + 1, (unsigned long long)n_nodes);
        atomicAdd(p.counters + 2, (unsigned long long)n_tris);
    }
    if (overflow) p.counters[3] = 1ull; // sticky: cleared when the host reads it (check_trace_status)
}

// k_trace with a one-deep triangle STASH.  In k_trace a lane that found triangle work waits for a triangle phase, and a triangle
// phase runs as soon as the triangle lanes outnumber the node lanes -- both phases run with about half the warp.  Here a lane
// keeps walking nodes with its triangle group set aside (one more group fits in the stash), and a triangle phase starts only
// when three quarters of the live lanes hold triangle work (or nobody can do node work): both long code paths run with most of
// the warp.  The price is a best-t that is up to two node steps stale when the next nodes are culled.  Same results: the
// triangle test and the argmin over t do not depend on the order of the tests (exact ties aside).
template <bool ANY, bool COUNT, int REFILL, int TRI_STEPS>
__global__ void __launch_bounds__(TRACE_BLOCK) k_trace_stash(TraceParams p) {
    __shared__ uint2 sm_stack_all[SM_STACK * TRACE_BLOCK];
    uint2 *const sm_stack = sm_stack_all + threadIdx.x;
    uint2 lstack[LOCAL_STACK];
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const bool have_tree = p.bvh.n_nodes > 0;

    long long ray = -1; // this lane's ray index, -1 = idle
    bool exhausted = false;
    float ox = 0, oy = 0, oz = 0, idx = 0, idy = 0, idz = 0;
    RayShear rs = {0.f, 0.f, 0.f, 2}; // ray-only part of the watertight test, evaluated once per ray
    uint32_t octinv4 = 0;
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u), ts = make_uint2(0u, 0u); // ts: the stashed triangle group
    int sp = 0;
    float ht = 0, hb1 = 0, hb2 = 0;
    int hprim = -1;
    unsigned n_nodes = 0, n_tris = 0, overflow = 0;

    while (true) {
        // ---- warp-aggregated fetch of new rays into idle lanes
        const unsigned idle = __ballot_sync(0xffffffffu, ray < 0);
        if (!exhausted && (__popc(idle) >= REFILL || idle == 0xffffffffu)) {
            const int n_idle = __popc(idle);
            const int leader = __ffs(idle) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(p.ticket, (unsigned long long)n_idle);
            base = __shfl_sync(0xffffffffu, base, leader);
            if ((long long)base + n_idle >= p.n) exhausted = true;
            if (ray < 0) {
                const long long r = (long long)base + __popc(idle & lt_mask);
                if (r < p.n) {
                    ray = r;
                    float4 ro, rd;
                    if (p.o) {
                        const float *po = p.o + 3 * r, *pd = p.d + 3 * r;
                        ro = make_float4(__ldg(po), __ldg(po + 1), __ldg(po + 2), __ldg(p.tmax + r));
                        rd = make_float4(__ldg(pd), __ldg(pd + 1), __ldg(pd + 2), 0.f);
                    } else {
                        ro = __ldg(p.rays + 2 * r);
                        rd = __ldg(p.rays + 2 * r + 1);
                    }
                    ox = ro.x; oy = ro.y; oz = ro.z;
                    idx = safe_rcp_dir(rd.x); idy = safe_rcp_dir(rd.y); idz = safe_rcp_dir(rd.z);
                    rs = ray_shear(rd.x, rd.y, rd.z);
                    octinv4 = ((idx < 0.f ? 0u : 4u) | (idy < 0.f ? 0u : 2u) | (idz < 0.f ? 0u : 1u)) * 0x01010101u;
                    ng = make_uint2(0u, have_tree ? 0x80000000u : 0u);
                    tg = make_uint2(0u, 0u);
                    ts = make_uint2(0u, 0u);
                    sp = 0;
                    ht = ro.w;
                    hb1 = 0.f; hb2 = 0.f;
                    hprim = -1;
                }
            }
        }
        if (__ballot_sync(0xffffffffu, ray >= 0) == 0u) break;

        // ---- phase vote
        const bool act = ray >= 0;
        const bool want_tri = act && tg.y != 0u;                          // tg is non-empty whenever ts is
        const bool has_node = act && (ng.y > 0x00ffffffu || sp > 0);
        const bool can_node = act && !(want_tri && ts.y != 0u) && (has_node || !want_tri); // incl. lanes that only have to retire
        const int c_live = __popc(__ballot_sync(0xffffffffu, act));
        const int c_tri = __popc(__ballot_sync(0xffffffffu, want_tri));
        const int c_node = __popc(__ballot_sync(0xffffffffu, can_node));
        if (c_node == 0 || 4 * c_tri >= 3 * c_live) {
            if (want_tri) {
#pragma unroll 1
                for (int step = 0; step < TRI_STEPS && tg.y != 0u; step++) {
                    const int ti = __ffs(tg.y) - 1;
                    tg.y &= tg.y - 1u;
                    const float4 *tp = p.bvh.tris + 3 * size_t(tg.x + ti);
                    const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                    if (COUNT) n_tris++;
                    TriHit th;
                    if (!(__float_as_uint(v1.w) & kTriDegenerate) &&
                        intersect_triangle_sheared(ox, oy, oz, rs, ht, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z, th)) {
                        ht = th.t;
                        hb1 = th.b1;
                        hb2 = th.b2;
                        hprim = __float_as_int(v0.w);
                        if (ANY) { // first accepted candidate ends the ray
                            tg.y = 0u;
                            ts.y = 0u;
                            ng.y = 0u;
                            sp = 0;
                        }
                    }
                }
                if (tg.y == 0u) { // the stashed group moves up
                    tg = ts;
                    ts.y = 0u;
                }
            }
        } else if (can_node) {
            if (ng.y <= 0x00ffffffu && sp > 0) {
                sp--;
                if (sp < SM_STACK) ng = sm_stack[sp * TRACE_BLOCK];
                else if (sp < SM_STACK + LOCAL_STACK) ng = lstack[sp - SM_STACK];
                else ng = make_uint2(0u, 0u);
            }
            if (ng.y > 0x00ffffffu) {
                const uint32_t hits = ng.y;
                const int child_bit = 31 - __clz(hits);
                ng.y &= ~(1u << child_bit);
                if (ng.y > 0x00ffffffu) {
                    if (sp < SM_STACK) sm_stack[sp * TRACE_BLOCK] = ng;
                    else if (sp < SM_STACK + LOCAL_STACK) lstack[sp - SM_STACK] = ng;
                    else overflow = 1u;
                    sp++;
                }
                const uint32_t slot = (uint32_t(child_bit) - 24u) ^ (octinv4 & 0xffu);
                const uint32_t rel = __popc(hits & 0xffu & ((1u << slot) - 1u));
                const float4 *np = p.bvh.nodes + 5 * size_t(ng.x + rel);
                const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
                if (COUNT) n_nodes++;
                const uint32_t hitmask = intersect_node8(n0, n1, n2, n3, n4, ox, oy, oz, idx, idy, idz, octinv4, ht, p.prmt_bias);
                ng.x = __float_as_uint(n1.x);
                ng.y = (hitmask & 0xff000000u) | (__float_as_uint(n0.w) >> 24);
                const uint2 tn = make_uint2(__float_as_uint(n1.y), hitmask & 0x00ffffffu);
                if (tn.y != 0u) {
                    if (tg.y == 0u) tg = tn;
                    else ts = tn; // can_node guarantees the stash was free
                }
            } else { // nothing left: retire the ray
                if (p.o) {
                    if (p.prim_out) p.prim_out[ray] = ANY ? (hprim >= 0 ? 1 : 0) : hprim;
                    if (p.t_out) p.t_out[ray] = (ANY || hprim < 0) ? (ANY ? 0.f : kInf) : ht;
                    if (p.b12_out) *reinterpret_cast<float2 *>(p.b12_out + 2 * ray) = ANY ? make_float2(0.f, 0.f) : make_float2(hb1, hb2);
                } else if (ANY) {
                    p.hits[ray] = make_float4(__int_as_float(hprim >= 0 ? 1 : 0), 0.f, 0.f, 0.f);
                } else {
                    p.hits[ray] = make_float4(__int_as_float(hprim), hprim >= 0 ? ht : kInf, hb1, hb2);
                }
                ray = -1;
            }
        }
    }
    if (COUNT) {
        atomicAdd(p.counters + 1, (unsigned long long)n_nodes);
        atomicAdd(p.counters + 2, (unsigned long long)n_tris);
    }
    if (overflow) p.counters[3] = 1ull; // sticky: cleared when the host reads it (check_trace_status)
}

// ---- queue-driven form used by the wavefront integrator (render.cu) ---------------------------
// Same traversal loop; what differs from k_trace: the ray count and the ray list come from DEVICE
// memory (queue + counter written by the previous stage, no host round trip), rays are gathered from
// the SoA path pool by slot, spheres are tested beside the BVH, and
//   SHADOW = false: closest hit, writes {leaf-order triangle slot | n_tris + sphere, t, b1, b2} to
//                   pool.hit[slot] (the shade stage reads vertices / material / light ids from the
//                   48-byte leaf record directly: no prim -> index -> vertex pointer chase);
//   SHADOW = true:  any hit on the compact shadow queue; an UNOCCLUDED ray adds its pre-weighted
//                   NEE contribution to pool.L[slot] on retirement (one shadow ray per slot per pass,
//                   so the add needs no atomic) -- the reference runs these traversals inline in the
//                   shading threads (integrators/megakernel_path.cu:217-233, base/integrator_base.cu:12-71).
// first quadric at or after k whose padded box the ray (o, 1/d, [0, t_max]) touches; n when there is none
__device__ __forceinline__ int next_sphere(const SphereRec *sph, int n, int k, float ox, float oy, float oz, float idx, float idy,
                                           float idz, float t_max) {
    for (; k < n; k++) {
        const float *lo = sph[k].box_lo, *hi = sph[k].box_hi;
        const float x0 = (__ldg(lo) - ox) * idx, x1 = (__ldg(hi) - ox) * idx;
        const float y0 = (__ldg(lo + 1) - oy) * idy, y1 = (__ldg(hi + 1) - oy) * idy;
        const float z0 = (__ldg(lo + 2) - oz) * idz, z1 = (__ldg(hi + 2) - oz) * idz;
        const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.f));
        const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), t_max));
        if (tn <= tf * 1.00001f + 1e-30f) break;
    }
    return k;
}

// INST: the scene has object instances (two-level traversal).  A ray first walks the top-level tree of the scene's own
// triangles (and the quadrics); when that is exhausted it walks the INSTANCE TREE -- a BVH8 over the instances' padded
// render-space boxes whose leaf records are proxies carrying the instance index -- with the same node steps.  A proxy the ray
// touches suspends that walk (its pending groups go onto the stack, base_sp marks them), the ray is taken into the instance's
// space with the reference's Transform::apply_inverse(Ray, &t_max) (t_max = best t so far minus the origin shift) and walks
// the OBJECT's tree -- built once, shared by all its instances -- from that tree's root; when the object tree is exhausted
// the render-space ray comes back and the instance tree carries on.  Hits keep the instance-space t, as
// TransformedPrimitive::intersect returns it (primitives/transformed_primitive.cu:20-34).
// cur_inst: -1 in the scene's tree, -2 in the instance tree, >= 0 in that instance's object tree.
template <bool SHADOW, bool SPH, bool INST>
__global__ void __launch_bounds__(TRACE_BLOCK) k_trace_wf(WF wf) {
    __shared__ uint2 sm_stack_all[SM_STACK * TRACE_BLOCK];
    uint2 *const sm_stack = sm_stack_all + threadIdx.x;
    uint2 lstack[LOCAL_STACK];
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const bool have_tree = wf.bvh.n_nodes > 0;
    // closest hit: every pool slot in slot order (coalesced gathers), live ones have state ST_RAY;
    // shadow: the compact shadow queue
    if (wf.c->finished >= (unsigned long long)wf.total_paths) return; // a look-ahead pass after the last path retired
    WFPass &pc = wf.c->p[wf.parity];
    const long long n = SHADOW ? (long long)pc.n_shadow : (long long)wf.pool_size;
    unsigned *const ticket = SHADOW ? &pc.ticket_shadow : &pc.ticket_closest;
    if (SHADOW && blockIdx.x == 0 && threadIdx.x == 0) wf.c->rays_shadow += (unsigned long long)n;
    unsigned n_live = 0;
    const int REFILL = wf.refill_min;
    const int tri_steps = wf.tri_steps;
    const uint32_t prmt_bias = 0x47000000u;

    long long ray = -1;
    int slot = -1;
    bool exhausted = false;
    float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 0, idx = 0, idy = 0, idz = 0;
    RayShear rs = {0.f, 0.f, 0.f, 2};
    uint32_t octinv4 = 0;
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u);
    int sp = 0;
    float ht = 0, hb1 = 0, hb2 = 0;
    int hprim = -1;
    int cur_inst = -1, hinst = -1; // INST: where the lane is (see above); instance of the best hit
    int base_sp = 0;               // INST: stack entries below this belong to the suspended instance-tree walk
    int sph_k = 0;      // next sphere to test for this lane's ray (n_spheres: none pending)
    int skip_prim = -1; // SHADOW: the sampled light's own triangle (the reference's three visibility rays cannot reach it)
    unsigned overflow = 0;

    while (true) {
        const unsigned idle = __ballot_sync(0xffffffffu, ray < 0);
        if (!exhausted && (__popc(idle) >= REFILL || idle == 0xffffffffu)) {
            const int n_idle = __popc(idle);
            const int leader = __ffs(idle) - 1;
            unsigned base = 0;
            if (lane == leader) base = atomicAdd(ticket, (unsigned)n_idle);
            base = __shfl_sync(0xffffffffu, base, leader);
            if ((long long)base + n_idle >= n) exhausted = true;
            if (ray < 0) {
                const long long r = (long long)base + __popc(idle & lt_mask);
                if (r < n && (SHADOW || wf.pool.state[r] == ST_RAY)) {
                    ray = r;
                    float4 ro, rd;
                    if (SHADOW) {
                        ro = __ldg(wf.q.sh_o + r);
                        rd = __ldg(wf.q.sh_d + r);
                        skip_prim = __float_as_int(rd.w);
                    } else {
                        slot = int(r);
                        ro = wf.pool.ray_o[slot];
                        rd = wf.pool.ray_d[slot];
                        n_live++;
                    }
                    ox = ro.x; oy = ro.y; oz = ro.z;
                    dx = rd.x; dy = rd.y; dz = rd.z;
                    idx = safe_rcp_dir(dx); idy = safe_rcp_dir(dy); idz = safe_rcp_dir(dz);
                    rs = ray_shear(dx, dy, dz);
                    octinv4 = ((idx < 0.f ? 0u : 4u) | (idy < 0.f ? 0u : 2u) | (idz < 0.f ? 0u : 1u)) * 0x01010101u;
                    const bool live = ro.w >= 0.f; // tmax < 0 marks a dead path: report a miss
                    ng = make_uint2(0u, (have_tree && live) ? 0x80000000u : 0u);
                    tg = make_uint2(0u, 0u);
                    sp = 0;
                    ht = ro.w;
                    hb1 = 0.f; hb2 = 0.f;
                    hprim = -1;
                    cur_inst = -1;
                    hinst = -1;
                    base_sp = 0;
                    // spheres are tested as a phase of their own below, and only those whose box the ray touches
                    sph_k = (SPH && live) ? next_sphere(wf.spheres, wf.n_spheres, 0, ox, oy, oz, idx, idy, idz, ht) : wf.n_spheres;
                }
            }
        }
        if (__ballot_sync(0xffffffffu, ray >= 0) == 0u) {
            if (exhausted) break;
            continue; // every fetched slot was inactive: fetch again
        }

        const bool act = ray >= 0;
        const bool want_sph = SPH && act && sph_k < wf.n_spheres;
        const bool want_tri = act && !want_sph && tg.y != 0u;
        const int c_sph = SPH ? __popc(__ballot_sync(0xffffffffu, want_sph)) : 0;
        const int c_tri = __popc(__ballot_sync(0xffffffffu, want_tri));
        const int c_node = __popc(__ballot_sync(0xffffffffu, act && !want_sph && !want_tri));
        if (c_sph > 0 && c_sph >= c_tri && c_sph >= c_node) {
            // sphere phase (Sphere::basic_intersect in interval arithmetic is long: run it with as many lanes as want it)
            if (want_sph) {
                SphereHit sh;
                // SHADOW: the sampled sphere light itself is skipped, like the sampled light's own triangle
                if ((!SHADOW || int(wf.bvh.n_tris) + sph_k != skip_prim) &&
                    sphere_intersect(wf.spheres[sph_k], v3(ox, oy, oz), v3(dx, dy, dz), ht, sh)) {
                    ht = sh.t;
                    hprim = int(wf.bvh.n_tris) + sph_k;
                    if (SHADOW) {
                        ng.y = 0u;
                        sph_k = wf.n_spheres;
                    }
                }
                sph_k = next_sphere(wf.spheres, wf.n_spheres, sph_k + 1, ox, oy, oz, idx, idy, idz, ht);
            }
        } else if (c_tri >= c_node) {
            if (INST && want_tri && cur_inst == -2) {
                // a proxy of the instance tree: enter the instance if the ray touches its own box
                const int ti = __ffs(tg.y) - 1;
                tg.y &= tg.y - 1u;
                const int ni = __float_as_int(__ldg(&wf.bvh.tris[3 * size_t(tg.x + ti)].w)) - int(wf.bvh.n_tris);
                const InstRec &I = wf.instances[ni];
                if (ray_touches_instance(I, ox, oy, oz, idx, idy, idz, ht)) {
                    // suspend the instance-tree walk: the rest of this leaf group and of the node group wait on the stack
                    if (tg.y != 0u) {
                        if (sp < SM_STACK) sm_stack[sp * TRACE_BLOCK] = tg;
                        else if (sp < SM_STACK + LOCAL_STACK) lstack[sp - SM_STACK] = tg;
                        else overflow = 1u;
                        sp++;
                    }
                    if (ng.y > 0x00ffffffu) {
                        if (sp < SM_STACK) sm_stack[sp * TRACE_BLOCK] = ng;
                        else if (sp < SM_STACK + LOCAL_STACK) lstack[sp - SM_STACK] = ng;
                        else overflow = 1u;
                        sp++;
                    }
                    base_sp = sp;
                    // the best t so far is a t along the ray of the space it was found in; like the reference, it is handed to the
                    // next primitive as it is (best_t = intersection->t_hit, accelerator/hlbvh.cu:499) and shifted by that ray's dt
                    V3 oi, di;
                    float dt;
                    instance_ray(I.ifr, v3(ox, oy, oz), v3(dx, dy, dz), &oi, &di, &dt);
                    ox = oi.x; oy = oi.y; oz = oi.z;
                    dx = di.x; dy = di.y; dz = di.z;
                    idx = safe_rcp_dir(di.x); idy = safe_rcp_dir(di.y); idz = safe_rcp_dir(di.z);
                    rs = ray_shear(di.x, di.y, di.z);
                    octinv4 = ((idx < 0.f ? 0u : 4u) | (idy < 0.f ? 0u : 2u) | (idz < 0.f ? 0u : 1u)) * 0x01010101u;
                    ht = ht - dt;
                    ng = make_uint2((uint32_t)I.root_node, 0x80000000u);
                    tg = make_uint2(0u, 0u);
                    cur_inst = ni;
                }
            } else if (want_tri) {
#pragma unroll 1
                for (int step = 0; step < tri_steps && tg.y != 0u; step++) {
                    const int ti = __ffs(tg.y) - 1;
                    tg.y &= tg.y - 1u;
                    const float4 *tp = wf.bvh.tris + 3 * size_t(tg.x + ti);
                    const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                    TriHit th;
                    if ((!SHADOW || __float_as_int(v0.w) != skip_prim) && !(__float_as_uint(v1.w) & kTriDegenerate) &&
                        intersect_triangle_sheared(ox, oy, oz, rs, ht, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z, th)) {
                        ht = th.t;
                        hb1 = th.b1;
                        hb2 = th.b2;
                        hprim = int(tg.x + ti);
                        if (INST) hinst = cur_inst;
                        if (SHADOW) {
                            tg.y = 0u;
                            ng.y = 0u;
                            sp = 0;
                        }
                    }
                }
            }
        } else if (act && !want_sph && !want_tri) {
            if (ng.y <= 0x00ffffffu && sp > (INST ? base_sp : 0)) {
                sp--;
                if (sp < SM_STACK) ng = sm_stack[sp * TRACE_BLOCK];
                else if (sp < SM_STACK + LOCAL_STACK) ng = lstack[sp - SM_STACK];
                else ng = make_uint2(0u, 0u);
                if (INST && ng.y <= 0x00ffffffu) { // a suspended group of proxies: back to the triangle phase
                    tg = ng;
                    ng = make_uint2(0u, 0u);
                }
            }
            if (INST && tg.y != 0u) {
                // nothing to do in this node phase
            } else if (ng.y > 0x00ffffffu) {
                const uint32_t hits = ng.y;
                const int child_bit = 31 - __clz(hits);
                ng.y &= ~(1u << child_bit);
                if (ng.y > 0x00ffffffu) {
                    if (sp < SM_STACK) sm_stack[sp * TRACE_BLOCK] = ng;
                    else if (sp < SM_STACK + LOCAL_STACK) lstack[sp - SM_STACK] = ng;
                    else overflow = 1u;
                    sp++;
                }
                const uint32_t cslot = (uint32_t(child_bit) - 24u) ^ (octinv4 & 0xffu);
                const uint32_t rel = __popc(hits & 0xffu & ((1u << cslot) - 1u));
                const float4 *np = wf.bvh.nodes + 5 * size_t(ng.x + rel);
                const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
                const uint32_t hitmask = intersect_node8(n0, n1, n2, n3, n4, ox, oy, oz, idx, idy, idz, octinv4, ht, prmt_bias);
                ng.x = __float_as_uint(n1.x);
                tg.x = __float_as_uint(n1.y);
                ng.y = (hitmask & 0xff000000u) | (__float_as_uint(n0.w) >> 24);
                tg.y = hitmask & 0x00ffffffu;
            } else {
                bool retire = true;
                if (INST && !(SHADOW && hprim >= 0)) {
                    if (cur_inst >= 0) {
                        // the object's tree is exhausted: the render-space ray comes back and the instance tree carries on
                        float4 ro4, rd4;
                        if (SHADOW) {
                            ro4 = __ldg(wf.q.sh_o + ray);
                            rd4 = __ldg(wf.q.sh_d + ray);
                        } else {
                            ro4 = wf.pool.ray_o[slot];
                            rd4 = wf.pool.ray_d[slot];
                        }
                        ox = ro4.x; oy = ro4.y; oz = ro4.z;
                        dx = rd4.x; dy = rd4.y; dz = rd4.z;
                        idx = safe_rcp_dir(dx); idy = safe_rcp_dir(dy); idz = safe_rcp_dir(dz);
                        rs = ray_shear(dx, dy, dz);
                        octinv4 = ((idx < 0.f ? 0u : 4u) | (idy < 0.f ? 0u : 2u) | (idz < 0.f ? 0u : 1u)) * 0x01010101u;
                        cur_inst = -2;
                        base_sp = 0;
                        retire = false;
                    } else if (cur_inst == -1 && wf.inst_root >= 0) {
                        // the scene's own tree is exhausted: on to the tree over the instances (dead paths skip it)
                        const float w = SHADOW ? __ldg(wf.q.sh_o + ray).w : wf.pool.ray_o[slot].w;
                        if (w >= 0.f) {
                            ng = make_uint2((uint32_t)wf.inst_root, 0x80000000u);
                            tg = make_uint2(0u, 0u);
                            sp = 0;
                            cur_inst = -2;
                            retire = false;
                        }
                    }
                }
                if (!retire) {
                    // carries on in the instance tree
                } else if (SHADOW) {
                    if (hprim < 0) {
                        const int s2 = __ldg(wf.q.sh_slot + ray);
                        const float4 c = __ldg(wf.q.sh_contrib + ray);
                        float4 L = wf.pool.L[s2];
                        L.x += c.x; L.y += c.y; L.z += c.z; L.w += c.w;
                        wf.pool.L[s2] = L;
                    }
                } else {
                    wf.pool.hit[slot] = make_float4(__int_as_float(hprim), hprim >= 0 ? ht : kInf, hb1, hb2);
                    if (INST) wf.pool.inst[slot] = hinst;
                }
                if (retire) ray = -1;
            }
        }
    }
    if (overflow) wf.c->stack_overflow = 1u;
    if (!SHADOW) {
        for (int o = 16; o > 0; o >>= 1) n_live += __shfl_xor_sync(0xffffffffu, n_live, o);
        if (lane == 0 && n_live) atomicAdd(&wf.c->rays_closest, (unsigned long long)n_live);
    }
}

int launch_trace_wf(b200pt_scene *s, const WF &wf, bool shadow) {
    typedef void (*K)(WF);
    static const K kern[2][2][2] = {{{k_trace_wf<false, false, false>, k_trace_wf<false, false, true>},
                                     {k_trace_wf<false, true, false>, k_trace_wf<false, true, true>}},
                                    {{k_trace_wf<true, false, false>, k_trace_wf<true, false, true>},
                                     {k_trace_wf<true, true, false>, k_trace_wf<true, true, true>}}};
    static int per_sm[2][2][2] = {{{0, 0}, {0, 0}}, {{0, 0}, {0, 0}}};
    const int sph = wf.n_spheres > 0 ? 1 : 0, inst = wf.n_instances > 0 ? 1 : 0;
    if (per_sm[shadow][sph][inst] == 0) {
        int v = 0;
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, kern[shadow][sph][inst], TRACE_BLOCK, 0));
        per_sm[shadow][sph][inst] = v < 1 ? 1 : v;
    }
    const unsigned blocks = unsigned(s->sm_count * per_sm[shadow][sph][inst]);
    kern[shadow][sph][inst]<<<blocks, TRACE_BLOCK, 0, s->stream>>>(wf);
    CU_TRY(cudaGetLastError());
    return B200PT_OK;
}

typedef void (*TraceKernel)(TraceParams);

template <int REFILL, int TRI_STEPS>
static TraceKernel pick_kernel2(bool any_hit, bool counting) {
    return any_hit ? (counting ? k_trace<true, true, REFILL, TRI_STEPS> : k_trace<true, false, REFILL, TRI_STEPS>)
                   : (counting ? k_trace<false, true, REFILL, TRI_STEPS> : k_trace<false, false, REFILL, TRI_STEPS>);
}
static TraceKernel pick_stash_kernel(bool any_hit, bool counting, int tri_steps) {
    if (tri_steps <= 1) return any_hit ? (counting ? k_trace_stash<true, true, 4, 1> : k_trace_stash<true, false, 4, 1>)
                                       : (counting ? k_trace_stash<false, true, 4, 1> : k_trace_stash<false, false, 4, 1>);
    return any_hit ? (counting ? k_trace_stash<true, true, 4, 3> : k_trace_stash<true, false, 4, 3>)
                   : (counting ? k_trace_stash<false, true, 4, 3> : k_trace_stash<false, false, 4, 3>);
}

template <int REFILL>
static TraceKernel pick_kernel(bool any_hit, bool counting, int tri_steps) {
    return tri_steps <= 1 ? pick_kernel2<REFILL, 1>(any_hit, counting)
         : tri_steps == 2 ? pick_kernel2<REFILL, 2>(any_hit, counting)
                          : pick_kernel2<REFILL, 3>(any_hit, counting);
}

int launch_trace(b200pt_scene *s, const void *rays_dev, int64_t n, int any_hit, void *hits_dev, bool keep_status,
                 const TraceHostIO *io, int ticket_slot, cudaStream_t stream) {
    if (!s->built) {
        set_error("b200pt_trace: call b200pt_bvh_build first");
        return B200PT_ERR_STATE;
    }
    if (n <= 0) return B200PT_OK;
    TraceParams p;
    p.bvh = s->bvh;
    p.rays = (const float4 *)rays_dev;
    p.hits = (float4 *)hits_dev;
    p.o = io ? io->o : nullptr;
    p.d = io ? io->d : nullptr;
    p.tmax = io ? io->tmax : nullptr;
    p.prim_out = io ? io->prim : nullptr;
    p.t_out = io ? io->t : nullptr;
    p.b12_out = io ? io->b12 : nullptr;
    p.n = n;
    p.counters = s->d_trace_counters;
    p.ticket = s->d_trace_counters + (ticket_slot >= 0 ? 4 + ticket_slot % kTraceTicketRing : 0);
    p.prmt_bias = 0x47000000u;
    if (!stream) stream = s->stream;
    // per launch: the work ticket and (unless this is a later chunk of one batch) the visit counters.  The overflow word
    // [3] is STICKY across launches -- back-to-back asynchronous launches must not wipe an earlier launch's report; it is
    // cleared when the host reads it (check_trace_status).
    if (ticket_slot >= 0) CU_TRY(cudaMemsetAsync(p.ticket, 0, sizeof(unsigned long long), stream));
    else CU_TRY(cudaMemsetAsync(s->d_trace_counters, 0, (keep_status ? 1 : 3) * sizeof(unsigned long long), stream));

    // tuning knobs (defaults are the measured best; b200pt_trace_tune overrides them per scene)
    const int refill = s->tune_refill, tri_steps = s->tune_tri_steps;
    TraceKernel kern = refill <= 1 ? pick_kernel<1>(any_hit, s->counting, tri_steps)
                     : refill <= 4 ? pick_kernel<4>(any_hit, s->counting, tri_steps)
                     : refill <= 8 ? pick_kernel<8>(any_hit, s->counting, tri_steps)
                                   : pick_kernel<16>(any_hit, s->counting, tri_steps);
    if (s->tune_stash) kern = pick_stash_kernel(any_hit, s->counting, tri_steps);
    int per_sm = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TRACE_BLOCK, 0));
    if (per_sm < 1) per_sm = 1;
    if (s->tune_ctas_per_sm > 0 && per_sm > s->tune_ctas_per_sm) per_sm = s->tune_ctas_per_sm;
    long long blocks = (long long)s->sm_count * per_sm;
    const long long needed = (n + TRACE_BLOCK - 1) / TRACE_BLOCK;
    if (blocks > needed) blocks = needed;
    kern<<<(unsigned)blocks, TRACE_BLOCK, 0, stream>>>(p);
    CU_TRY(cudaGetLastError());
    return B200PT_OK;
}

} // namespace b200pt
