// trace.cu -- closest-hit and any-hit traversal as persistent-thread kernels.
//
// Replaces ray_cast (integrators/wavefront_path.cu:348-366 -> HLBVH::intersect,
// accelerator/hlbvh.cu:459-516) and the inline HLBVH::fast_intersect calls of the shading
// kernels (accelerator/hlbvh.cu:407-456).  The grid is sized to the machine (SMs x resident
// CTAs); each warp claims 32 rays at a time with ONE atomicAdd on the ray ticket
// (warp-aggregated fetch), and the traversal keeps its short stack in shared memory.
// Rays: n x {ox,oy,oz,tmax, dx,dy,dz,-} (2 x 128-bit loads); hits: n x {prim,t,b1,b2} (one
// 128-bit store).
#include "common.cuh"
#include "traverse.cuh"

namespace b200pt {

struct TraceParams {
    BVH8 bvh;
    const float4 *rays;
    float4 *hits;
    long long n;
    unsigned long long *counters; // [0] ticket, [1] nodes, [2] tris, [3] overflow flag
};

template <bool ANY, bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK) k_trace(TraceParams p) {
    __shared__ uint2 sm_stack[SM_STACK * TRACE_BLOCK];
    const int lane = threadIdx.x & 31;
    TraceCounters ctr;
    unsigned overflow = 0;
    while (true) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(p.counters, 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if ((long long)base >= p.n) break;
        const long long r = (long long)base + lane;
        if (r < p.n) {
            const float4 ro = __ldg(p.rays + 2 * r), rd = __ldg(p.rays + 2 * r + 1);
            RayHit hit;
            traverse_bvh8<ANY, COUNT>(p.bvh, ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, ro.w, sm_stack + threadIdx.x, TRACE_BLOCK, hit,
                                      ctr, &overflow);
            if (ANY) p.hits[r] = make_float4(__int_as_float(hit.prim >= 0 ? 1 : 0), 0.f, 0.f, 0.f);
            else p.hits[r] = make_float4(__int_as_float(hit.prim), hit.prim >= 0 ? hit.t : kInf, hit.b1, hit.b2);
        }
        __syncwarp();
    }
    if (COUNT) {
        atomicAdd(p.counters + 1, (unsigned long long)ctr.nodes);
        atomicAdd(p.counters + 2, (unsigned long long)ctr.tris);
    }
    if (overflow) p.counters[3] = 1ull;
}

int launch_trace(b200pt_scene *s, const void *rays_dev, int64_t n, int any_hit, void *hits_dev) {
    if (!s->built) {
        set_error("b200pt_trace: call b200pt_bvh_build first");
        return B200PT_ERR_STATE;
    }
    if (n <= 0) return B200PT_OK;
    TraceParams p;
    p.bvh = s->bvh;
    p.rays = (const float4 *)rays_dev;
    p.hits = (float4 *)hits_dev;
    p.n = n;
    p.counters = s->d_trace_counters;
    CU_TRY(cudaMemsetAsync(s->d_trace_counters, 0, 4 * sizeof(unsigned long long), s->stream));

    void (*kern)(TraceParams) = any_hit ? (s->counting ? k_trace<true, true> : k_trace<true, false>)
                                        : (s->counting ? k_trace<false, true> : k_trace<false, false>);
    int per_sm = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TRACE_BLOCK, 0));
    if (per_sm < 1) per_sm = 1;
    long long blocks = (long long)s->sm_count * per_sm;
    const long long needed = (n + TRACE_BLOCK - 1) / TRACE_BLOCK;
    if (blocks > needed) blocks = needed;
    kern<<<(unsigned)blocks, TRACE_BLOCK, 0, s->stream>>>(p);
    CU_TRY(cudaGetLastError());
    return B200PT_OK;
}

} // namespace b200pt
