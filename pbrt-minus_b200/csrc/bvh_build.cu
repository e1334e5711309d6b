// bvh_build.cu -- HLBVH build for B200, entirely on the device.
//
// Replaces HLBVH::HLBVH (accelerator/hlbvh.cu:172-404): the reference interleaves 12+ kernels
// with host loops over managed memory, a CPU SAH over treelets and one-thread-per-node serial
// partitions.  Here:
//   k_centroid_bounds   triangle bounds + centroid box reduction         (ref :102-115, :206-209)
//   k_cubify            widen the minor axes to the major extent         (ref :210-225)
//   k_morton            30-bit reference-compatible codes (N <= 16 Mi) or 63-bit codes
//                                                                        (ref :128-142, util/math.h:183-208)
//   radix_sort_u64      LSD radix sort, 8-bit digits (sort.cuh)          (ref :56-82, :256-281)
//   k_emit_hierarchy    binary radix tree over the sorted Morton prefixes, one thread per
//                       interior node (replaces the CPU SAH :611-813 and the per-level
//                       midpoint splits :518-609, :815-844)
//   k_refit             bottom-up AABBs with arrival counters
//   k_collapse_level    work-queue emission of 8-wide treelets: each queue item turns one BVH2
//                       subtree root into one quantised 80-byte BVH8 node, allocates its child
//                       nodes / triangle slots with atomic counters, re-lays the triangles out
//                       in node order and pushes the interior children on the next queue.
#include <algorithm>
#include <chrono>
#include <vector>
#include <cmath>
#include <cstdlib>

#include "common.cuh"
#include "sort.cuh"
#include "triangle.cuh"

namespace b200pt {

// ------------------------------------------------------------------ orderable float <-> uint
__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f);
    return u ^ ((u >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float(u ^ ((u >> 31) ? 0x80000000u : 0xffffffffu));
}

struct TriVerts {
    float3 a, b, c;
};
__device__ __forceinline__ TriVerts load_tri(const float *__restrict__ P, const int32_t *__restrict__ idx, int64_t t) {
    const int64_t i0 = idx[3 * t], i1 = idx[3 * t + 1], i2 = idx[3 * t + 2];
    TriVerts v;
    v.a = make_float3(P[3 * i0], P[3 * i0 + 1], P[3 * i0 + 2]);
    v.b = make_float3(P[3 * i1], P[3 * i1 + 1], P[3 * i1 + 2]);
    v.c = make_float3(P[3 * i2], P[3 * i2 + 1], P[3 * i2 + 2]);
    return v;
}
__device__ __forceinline__ void tri_bounds(const TriVerts &v, float3 &lo, float3 &hi) {
    lo = make_float3(fminf(v.a.x, fminf(v.b.x, v.c.x)), fminf(v.a.y, fminf(v.b.y, v.c.y)), fminf(v.a.z, fminf(v.b.z, v.c.z)));
    hi = make_float3(fmaxf(v.a.x, fmaxf(v.b.x, v.c.x)), fmaxf(v.a.y, fmaxf(v.b.y, v.c.y)), fmaxf(v.a.z, fmaxf(v.b.z, v.c.z)));
}
// euclidean_space/bounds3.h:102-104  centroid = 0.5 * (p_min + p_max)
__device__ __forceinline__ float3 centroid_of(const float3 &lo, const float3 &hi) {
    return make_float3(__fmul_rn(0.5f, __fadd_rn(lo.x, hi.x)), __fmul_rn(0.5f, __fadd_rn(lo.y, hi.y)),
                       __fmul_rn(0.5f, __fadd_rn(lo.z, hi.z)));
}

struct GatherArgs {
    const float *P;
    const int32_t *idx;
    const int32_t *tri_material, *tri_light;
    const uint8_t *tri_flip;
    int32_t prim_offset; // global id of the range's first triangle
};

// the 48-byte leaf record {v0 | prim, v1 | material word, v2 | light} of triangle t of the range
//   v1.w = material id | degenerate << 28 | has-normals << 29 | has-uv << 30 | flip << 31 (b200pt.h: tri_flip bits 1, 2, 0)
__device__ __forceinline__ void write_leaf_record(const GatherArgs &g, const TriVerts &tv, int64_t t, float4 *__restrict__ rec) {
    const uint32_t flip = g.tri_flip ? uint32_t(g.tri_flip[t]) : 0u;
    const uint32_t mat = (g.tri_material ? (uint32_t)g.tri_material[t] : 0u) | ((flip & 1u) << 31) | ((flip & 6u) << 28) |
                         (triangle_is_degenerate(tv.a.x, tv.a.y, tv.a.z, tv.b.x, tv.b.y, tv.b.z, tv.c.x, tv.c.y, tv.c.z) ? kTriDegenerate : 0u);
    const int light = g.tri_light ? g.tri_light[t] : -1;
    rec[0] = make_float4(tv.a.x, tv.a.y, tv.a.z, __int_as_float((int)t + g.prim_offset));
    rec[1] = make_float4(tv.b.x, tv.b.y, tv.b.z, __uint_as_float(mat));
    rec[2] = make_float4(tv.c.x, tv.c.y, tv.c.z, __int_as_float(light));
}

// red[0..5] = centroid box (min xyz, max xyz), red[6..11] = scene box; all as orderable uints
// The pass also writes every triangle's leaf record in PRIMITIVE order (coalesced): the Morton pass reads those instead of
// chasing the index triples again, and the one random gather of the build (k_gather_sorted) then moves one contiguous
// 48-byte record per triangle instead of six scattered words.
__global__ void __launch_bounds__(256) k_centroid_bounds(GatherArgs g, int64_t n, uint32_t *__restrict__ red, float4 *__restrict__ rec_unsorted) {
    float v[12] = {kInf, kInf, kInf, -kInf, -kInf, -kInf, kInf, kInf, kInf, -kInf, -kInf, -kInf};
    auto take = [&](const TriVerts &tv, int64_t t) {
        write_leaf_record(g, tv, t, rec_unsorted + 3 * t);
        float3 lo, hi;
        tri_bounds(tv, lo, hi);
        const float3 c = centroid_of(lo, hi);
        v[0] = fminf(v[0], c.x); v[1] = fminf(v[1], c.y); v[2] = fminf(v[2], c.z);
        v[3] = fmaxf(v[3], c.x); v[4] = fmaxf(v[4], c.y); v[5] = fmaxf(v[5], c.z);
        v[6] = fminf(v[6], lo.x); v[7] = fminf(v[7], lo.y); v[8] = fminf(v[8], lo.z);
        v[9] = fmaxf(v[9], hi.x); v[10] = fmaxf(v[10], hi.y); v[11] = fmaxf(v[11], hi.z);
    };
    // two triangles per trip: both index -> vertex chains are in flight before either is used
    const int64_t stride = int64_t(gridDim.x) * blockDim.x;
    int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    for (; t + stride < n; t += 2 * stride) {
        const TriVerts ta = load_tri(g.P, g.idx, t), tb = load_tri(g.P, g.idx, t + stride);
        take(ta, t);
        take(tb, t + stride);
    }
    if (t < n) take(load_tri(g.P, g.idx, t), t);
#pragma unroll
    for (int k = 0; k < 12; k++) {
        const bool is_min = (k % 6) < 3;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float other = __shfl_xor_sync(0xffffffffu, v[k], o);
            v[k] = is_min ? fminf(v[k], other) : fmaxf(v[k], other);
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 12; k++) {
            if ((k % 6) < 3) atomicMin(&red[k], f2ord(v[k]));
            else atomicMax(&red[k], f2ord(v[k]));
        }
    }
}

__global__ void k_init_reduction(uint32_t *red) {
    const int k = threadIdx.x;
    if (k < 12) red[k] = ((k % 6) < 3) ? 0xffffffffu : 0u;
}

// accelerator/hlbvh.cu:210-225: the centroid box is widened on its two minor axes to the major
// extent, so all Morton cells are cubes.  boxes[0..5] = cubified centroid box, [6..11] scene box.
__global__ void k_cubify(const uint32_t *__restrict__ red, float *__restrict__ boxes) {
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) {
        lo[a] = ord2f(red[a]);
        hi[a] = ord2f(red[3 + a]);
    }
    const float dx = __fsub_rn(hi[0], lo[0]), dy = __fsub_rn(hi[1], lo[1]), dz = __fsub_rn(hi[2], lo[2]);
    const int md = (dx > dy && dx > dz) ? 0 : ((dy > dz) ? 1 : 2); // bounds3.h:124-133
    const float radius = __fdiv_rn(__fsub_rn(hi[md], lo[md]), 2.0f);
    for (int a = 0; a < 3; a++) {
        if (a == md) {
            boxes[a] = lo[a];
            boxes[3 + a] = hi[a];
        } else {
            const float center = __fmul_rn(0.5f, __fadd_rn(lo[a], hi[a]));
            boxes[a] = __fsub_rn(center, radius);
            boxes[3 + a] = __fadd_rn(center, radius);
        }
    }
    for (int a = 0; a < 6; a++) boxes[6 + a] = ord2f(red[6 + a]);
}

// util/math.h:183-208 (10 bits per axis, x lowest; the value 1024 is clamped to 1023)
__device__ __forceinline__ uint32_t left_shift3(uint32_t x) {
    if (x == (1u << 10)) --x;
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}
__device__ __forceinline__ uint64_t spread21(uint64_t x) {
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// KEY30: key = morton30 << 32 | prim (sorted key-only on bits 32..63; the reference's code, util/math.h:183-208).
// otherwise (N > 16 Mi): key = morton(3 * axis_bits) << prim_bits | prim, axis_bits = (64 - prim_bits) / 3 <= 21 --
// still one 8-byte key-only sort, over 3 * axis_bits bits instead of 63 bits + a 4-byte payload.
template <bool KEY30>
__global__ void __launch_bounds__(256) k_morton(const float4 *__restrict__ rec_unsorted, int64_t n,
                                                const float *__restrict__ boxes, uint64_t *__restrict__ keys, int axis_bits,
                                                int prim_bits) {
    const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (t >= n) return;
    const float4 r0 = rec_unsorted[3 * t], r1 = rec_unsorted[3 * t + 1], r2 = rec_unsorted[3 * t + 2];
    TriVerts tv;
    tv.a = make_float3(r0.x, r0.y, r0.z);
    tv.b = make_float3(r1.x, r1.y, r1.z);
    tv.c = make_float3(r2.x, r2.y, r2.z);
    float3 lo, hi;
    tri_bounds(tv, lo, hi);
    const float3 c = centroid_of(lo, hi);
    const float cc[3] = {c.x, c.y, c.z};
    float off[3];
#pragma unroll
    for (int a = 0; a < 3; a++) { // bounds3.h:106-122 offset()
        float o = __fsub_rn(cc[a], boxes[a]);
        if (boxes[3 + a] > boxes[a]) o = __fdiv_rn(o, __fsub_rn(boxes[3 + a], boxes[a]));
        off[a] = o;
    }
    if (KEY30) {
        const uint32_t qx = uint32_t(__fmul_rn(off[0], 1024.0f));
        const uint32_t qy = uint32_t(__fmul_rn(off[1], 1024.0f));
        const uint32_t qz = uint32_t(__fmul_rn(off[2], 1024.0f));
        const uint32_t m = (left_shift3(qz) << 2) | (left_shift3(qy) << 1) | left_shift3(qx);
        keys[t] = (uint64_t(m) << 32) | uint64_t(uint32_t(t));
    } else {
        const float cells = float(1u << axis_bits);
        uint64_t q[3];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float s = off[a] * cells;
            q[a] = s <= 0.0f ? 0ull : (s >= cells - 1.0f ? uint64_t(cells - 1.0f) : uint64_t(s));
        }
        const uint64_t m = (spread21(q[2]) << 2) | (spread21(q[1]) << 1) | spread21(q[0]);
        keys[t] = (m << prim_bits) | uint64_t(t);
    }
}

// The ONE random gather of the build: sorted position k -> primitive -> its 48-byte leaf record (written in primitive order by
// k_centroid_bounds), moved to SORTED order.  The refit reads its leaf boxes from
// these records and the collapse copies them into BVH8 leaf order, both as contiguous runs.
//   v1.w = material id | degenerate << 28 | has-normals << 29 | has-uv << 30 | flip << 31 (b200pt.h: tri_flip bits 1, 2, 0)
__global__ void __launch_bounds__(256) k_gather_sorted(const uint64_t *__restrict__ keys, int64_t n, uint32_t mask,
                                                       const float4 *__restrict__ rec_unsorted, uint32_t *__restrict__ prims,
                                                       float4 *__restrict__ rec) {
    const int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (k >= n) return;
    const uint32_t prim = uint32_t(keys[k]) & mask;
    prims[k] = prim;
    const float4 *src = rec_unsorted + 3 * size_t(prim);
    const float4 a = __ldg(src), b = __ldg(src + 1), c = __ldg(src + 2);
    rec[3 * k] = a;
    rec[3 * k + 1] = b;
    rec[3 * k + 2] = c;
}

// ------------------------------------------------------------------------ BVH2 (build only)
// Node ids: interior i in [0, n-1), leaf k -> (n-1) + k (k = position in sorted order).
// One 32-byte record per node, i.e. ONE memory sector: everything the refit and the collapse want to know about a node comes
// with its box.
//   rec[2i]     = { lo.x, lo.y, lo.z, A }    A = gamma | left-is-leaf << 30 | right-is-leaf << 31   (interior; leaves: -1)
//   rec[2i + 1] = { hi.x, hi.y, hi.z, B }    B = first | min(count, 4) - 1 << 30                    (first = first sorted position covered)
// gamma = split position of the radix tree: the children are gamma / gamma + 1, as interior ids or as leaves n - 1 + gamma (+ 1).
struct Bvh2 {
    float4 *rec;
    uint32_t *nodeA, *nodeB; // the A / B words as the emission writes them (coalesced); the refit composes the records
    int32_t *last;           // interior: last sorted position covered
    int32_t *parent;
    uint32_t *arrive;
    int32_t n; // leaves
};
__device__ __forceinline__ int bvh2_left(uint32_t A, int n) { return int(A & 0x3fffffffu) + ((A & 0x40000000u) ? n - 1 : 0); }
__device__ __forceinline__ int bvh2_right(uint32_t A, int n) { return int(A & 0x3fffffffu) + 1 + ((A & 0x80000000u) ? n - 1 : 0); }

// length of the common prefix of keys i and j.  The keys carry the primitive id in their low bits, so no two are equal and the
// usual index tie-break is not needed (the kernel is issue-bound: every instruction of this function counts ~13 times per node).
__device__ __forceinline__ int delta_fn(const uint64_t *__restrict__ keys, int n, int i, uint64_t ki, int j) {
    if ((unsigned)j >= (unsigned)n) return -1;
    return __clzll((long long)(ki ^ __ldg(keys + j)));
}

// Binary radix tree over sorted keys, one thread per interior node (every split is the highest
// differing Morton bit of its range -- the same splits the reference's midpoint recursion makes
// on a power-of-two grid, found without recursion or partitioning).
__global__ void __launch_bounds__(256) k_emit_hierarchy(const uint64_t *__restrict__ keys, Bvh2 t) {
    const int n = t.n;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const uint64_t ki = keys[i];
    const int d = (delta_fn(keys, n, i, ki, i + 1) - delta_fn(keys, n, i, ki, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta_fn(keys, n, i, ki, i - d);
    int lmax = 2;
    while (delta_fn(keys, n, i, ki, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int s = lmax >> 1; s >= 1; s >>= 1)
        if (delta_fn(keys, n, i, ki, i + (l + s) * d) > dmin) l += s;
    const int j = i + l * d;
    const int dnode = delta_fn(keys, n, i, ki, j);
    int s = 0;
    int step = l;
    do {
        step = (step + 1) >> 1;
        if (delta_fn(keys, n, i, ki, i + (s + step) * d) > dnode) s += step;
    } while (step > 1);
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int left = (first == gamma) ? (n - 1 + gamma) : gamma;
    const int right = (last == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
    const uint32_t A = uint32_t(gamma) | (first == gamma ? 0x40000000u : 0u) | (last == gamma + 1 ? 0x80000000u : 0u);
    const uint32_t B = uint32_t(first) | (uint32_t(min(last - first + 1, 4) - 1) << 30);
    t.nodeA[i] = A;
    t.nodeB[i] = B;
    t.last[i] = last;
    t.parent[left] = i;
    t.parent[right] = i;
    if (i == 0) t.parent[0] = -1;
}

// Bottom-up boxes.  A warp owns 32 consecutive sorted leaves and the 32 interior nodes of the same indices (interior node i
// covers a range that starts or ends at leaf i).  Every interior node whose range lies INSIDE the warp's 32 leaves is a union
// of a contiguous run of the warp's own leaf boxes: the warp builds a 6-level sparse table of those boxes with shuffles and
// each lane answers its node's range query with two table entries -- no atomics, no fences, no memory round trips for the
// bottom ~5 levels of the tree, which hold 15/16 of all nodes.  Only the roots of those in-warp subtrees (a handful per warp)
// climb on through the arrival counters: the first child to arrive stops, the second carries on with its box in registers.
__device__ __forceinline__ void refit_climb(const Bvh2 &t, int n, int me, float3 lo, float3 hi) {
    int cur = t.parent[me];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&t.arrive[cur], 1u) == 0u) return; // first arrival: the sibling finishes the job
        __threadfence();
        const uint32_t A = t.nodeA[cur], B = t.nodeB[cur];
        const int l = bvh2_left(A, n), r = bvh2_right(A, n);
        const int sib = (l == me) ? r : l;
        const float4 slo = __ldcg(&t.rec[2 * sib]), shi = __ldcg(&t.rec[2 * sib + 1]);
        lo = make_float3(fminf(lo.x, slo.x), fminf(lo.y, slo.y), fminf(lo.z, slo.z));
        hi = make_float3(fmaxf(hi.x, shi.x), fmaxf(hi.y, shi.y), fmaxf(hi.z, shi.z));
        t.rec[2 * cur] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(A));
        t.rec[2 * cur + 1] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(B));
        me = cur;
        cur = t.parent[cur];
    }
}

constexpr int REFIT_BLOCK = 512; // leaves (and interior nodes) per CTA (two 24-byte tables of that many entries fit in 48 KB)
struct Box6 {
    float lx, ly, lz, hx, hy, hz;
};
__device__ __forceinline__ Box6 box_union(const Box6 &a, const Box6 &b) {
    return Box6{fminf(a.lx, b.lx), fminf(a.ly, b.ly), fminf(a.lz, b.lz), fmaxf(a.hx, b.hx), fmaxf(a.hy, b.hy), fmaxf(a.hz, b.hz)};
}
__device__ __forceinline__ Box6 box_shfl(const Box6 &a, int src) {
    return Box6{__shfl_sync(0xffffffffu, a.lx, src), __shfl_sync(0xffffffffu, a.ly, src), __shfl_sync(0xffffffffu, a.lz, src),
                __shfl_sync(0xffffffffu, a.hx, src), __shfl_sync(0xffffffffu, a.hy, src), __shfl_sync(0xffffffffu, a.hz, src)};
}

// A CTA owns 512 consecutive sorted leaves and the interior nodes of the same indices.  The box of ANY interior node whose range
// lies inside those 512 leaves is the union of three pieces, each a table lookup:
//   the tail of the first warp's leaves (suffix unions), whole warps in between (a sparse table over the 16 warp boxes),
//   the head of the last warp's leaves (prefix unions);  a range inside one warp uses that warp's own sparse table (registers).
// That covers the bottom ~9 levels of the tree -- all but a few nodes in a thousand -- with shuffles and shared memory.
// Only the roots of those in-CTA subtrees climb on through the arrival counters (k_refit_climb).
__global__ void __launch_bounds__(REFIT_BLOCK) k_refit(const float4 *__restrict__ tri_sorted, Bvh2 t, int32_t *__restrict__ roots,
                                                       unsigned *__restrict__ n_roots) {
    __shared__ Box6 s_suffix[REFIT_BLOCK]; // [warp * 32 + j] = union of that warp's leaves j..31
    __shared__ Box6 s_prefix[REFIT_BLOCK]; //                  = union of that warp's leaves 0..j
    __shared__ Box6 s_warp[6][32];         // sparse table over the 32 whole-warp boxes: [k][w] = warps w .. w + 2^k - 1
    const int n = t.n;
    const int g = blockIdx.x * REFIT_BLOCK + threadIdx.x; // leaf g and interior node g
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cta_base = blockIdx.x * REFIT_BLOCK, base = g - lane;
    const bool have_leaf = g < n;
    Box6 leaf = {kInf, kInf, kInf, -kInf, -kInf, -kInf};
    if (have_leaf) {
        const float4 r0 = tri_sorted[3 * (size_t)g], r1 = tri_sorted[3 * (size_t)g + 1], r2 = tri_sorted[3 * (size_t)g + 2];
        TriVerts tv;
        tv.a = make_float3(r0.x, r0.y, r0.z);
        tv.b = make_float3(r1.x, r1.y, r1.z);
        tv.c = make_float3(r2.x, r2.y, r2.z);
        float3 lo, hi;
        tri_bounds(tv, lo, hi);
        leaf = Box6{lo.x, lo.y, lo.z, hi.x, hi.y, hi.z};
        const int id = n - 1 + g;
        t.rec[2 * id] = make_float4(lo.x, lo.y, lo.z, __int_as_float(-1));
        t.rec[2 * id + 1] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(uint32_t(g)));
    }
    // interior node g and its range
    const bool have_node = g < n - 1;
    uint32_t A = 0, B = 0;
    int first = 0, last = 0;
    if (have_node) {
        A = t.nodeA[g];
        B = t.nodeB[g];
        first = int(B & 0x3fffffffu);
        last = t.last[g];
    }
    const bool in_warp = have_node && first >= base && last <= base + 31;
    const bool in_cta = have_node && first >= cta_base && last <= cta_base + REFIT_BLOCK - 1;
    // ---- per-warp: sparse table in registers (answers in-warp ranges), prefix / suffix unions to shared memory
    Box6 node = {kInf, kInf, kInf, -kInf, -kInf, -kInf};
    {
        const int len = last - first + 1;
        const int klevel = in_warp ? 31 - __clz(len) : -1;
        const int qa = first - base, qb = last - base - (in_warp ? (1 << klevel) : 0) + 1;
        Box6 tab = leaf;
#pragma unroll
        for (int k = 0; k < 6; k++) { // a range of exactly 32 leaves sits at level 5
            const Box6 qa_box = box_shfl(tab, klevel == k ? qa : lane), qb_box = box_shfl(tab, klevel == k ? qb : lane);
            if (klevel == k) node = box_union(qa_box, qb_box);
            if (k == 5) break;
            const int step = 1 << k;
            const Box6 up = box_shfl(tab, (lane + step) & 31);
            if (lane + step < 32) tab = box_union(tab, up);
        }
        // inclusive prefix and suffix unions of the warp's 32 leaf boxes
        Box6 pre = leaf, suf = leaf;
#pragma unroll
        for (int step = 1; step < 32; step <<= 1) {
            const Box6 a = box_shfl(pre, (lane - step) & 31), b = box_shfl(suf, (lane + step) & 31);
            if (lane >= step) pre = box_union(pre, a);
            if (lane + step < 32) suf = box_union(suf, b);
        }
        s_prefix[threadIdx.x] = pre;
        s_suffix[threadIdx.x] = suf;
        if (lane == 0) s_warp[0][warp] = suf; // suffix from lane 0 = the whole warp
    }
    __syncthreads();
    if (warp == 0) { // sparse table over the warp boxes
        Box6 tab = lane < REFIT_BLOCK / 32 ? s_warp[0][lane] : Box6{kInf, kInf, kInf, -kInf, -kInf, -kInf};
#pragma unroll
        for (int k = 1; k < 6; k++) {
            const int step = 1 << (k - 1);
            const Box6 up = box_shfl(tab, (lane + step) & 31);
            if (lane + step < 32) tab = box_union(tab, up);
            s_warp[k][lane] = tab;
        }
    }
    __syncthreads();
    if (in_cta && !in_warp) { // spans warps: tail of the first warp, whole warps in between, head of the last warp
        const int fa = first - cta_base, la = last - cta_base;
        const int wf = fa >> 5, wl = la >> 5;
        node = box_union(s_suffix[fa], s_prefix[la]);
        if (wl - wf >= 2) {
            const int w0 = wf + 1, cnt = wl - wf - 1;
            const int k = 31 - __clz(cnt);
            node = box_union(node, box_union(s_warp[k][w0], s_warp[k][w0 + cnt - (1 << k)]));
        }
    }
    if (in_cta) {
        t.rec[2 * g] = make_float4(node.lx, node.ly, node.lz, __uint_as_float(A));
        t.rec[2 * g + 1] = make_float4(node.hx, node.hy, node.hz, __uint_as_float(B));
    }
    if (n <= 1) return;
    // roots of the in-CTA subtrees climb on: an element whose parent's range leaves the CTA's 512 leaves
    auto parent_inside = [&](int me) {
        const int p = t.parent[me];
        if (p < 0) return false;
        const int pf = int(t.nodeB[p] & 0x3fffffffu), pl = t.last[p];
        return pf >= cta_base && pl <= cta_base + REFIT_BLOCK - 1;
    };
    // (they are handed to k_refit_climb through a list: a CTA that waited for its few climbers would hold its 512 thread slots
    // for the length of a dependent chain of atomics, and the kernel would run at a fifth of its occupancy)
    if (have_leaf && !parent_inside(n - 1 + g)) roots[atomicAdd(n_roots, 1u)] = n - 1 + g;
    if (in_cta && !parent_inside(g)) roots[atomicAdd(n_roots, 1u)] = g;
}

// the upper part of the tree: one thread per in-CTA subtree root, arrival counters up to the root
__global__ void __launch_bounds__(128) k_refit_climb(Bvh2 t, const int32_t *__restrict__ roots, const unsigned *__restrict__ n_roots) {
    const unsigned count = *n_roots;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const int me = roots[i];
        const float4 lo = t.rec[2 * me], hi = t.rec[2 * me + 1];
        refit_climb(t, t.n, me, make_float3(lo.x, lo.y, lo.z), make_float3(hi.x, hi.y, hi.z));
    }
}

// ------------------------------------------------------------------- BVH8 work-queue collapse
struct CollapseArgs {
    Bvh2 t;
    const float4 *tri_sorted; // leaf records in sorted order (k_gather_sorted)
    float4 *nodes; // BVH8 out
    float4 *tris;  // re-laid-out triangles out
    const int2 *q_in; // (bvh2 node, bvh8 node index)
    int2 *q_out;
    // [0] bvh8 nodes allocated, [1] tris allocated, [2] overflow flag, [3] levels run,
    // [4 + parity] items in q_in, [4 + (parity ^ 1)] items pushed to q_out -- all on the device: the host
    // enqueues a batch of levels without looking and reads the block back once per batch
    unsigned int *counters;
    int parity;
    unsigned int node_capacity;
};

// 2^e as a double, e in the normal range (here [-127, 127])
__device__ __forceinline__ double pow2_double(int e) { return __longlong_as_double((long long)(e + 1023) << 52); }

__device__ __forceinline__ float half_area(const float4 &lo, const float4 &hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

// One thread per BVH8 node.  What the thread keeps per child while the treelet grows is three words (node id, area, child word):
// the child boxes are NOT kept -- 8 x 6 floats of dynamically indexed local memory per thread were 100 MB across the resident
// threads, as much as the L2 holds, and the biggest level moved 2.6 KB of DRAM reads per node for 0.7 KB of records and
// triangles (ncu, round 2).  The boxes are read again from the BVH2 records (cache hits: this thread fetched them a moment
// ago) when the node box, the slot order and the quantised bytes need them; the bytes are assembled in registers in slot order.
__device__ __forceinline__ void collapse_item(const CollapseArgs &a, int item) {
    const int n = a.t.n;
    const int2 work = a.q_in[item];
    const int root2 = work.x;

    int c[8];
    float area[8];
    uint32_t cA[8];
    int cnt;
    auto load_child = [&](int slot, int id) {
        const float4 lo = a.t.rec[2 * id], hi = a.t.rec[2 * id + 1];
        c[slot] = id;
        cA[slot] = __float_as_uint(lo.w);
        area[slot] = (id >= n - 1) ? -1.f : half_area(lo, hi);
    };
    if (root2 >= n - 1) { // single-leaf tree (n == 1)
        load_child(0, root2);
        cnt = 1;
    } else {
        const uint32_t A = __float_as_uint(a.t.rec[2 * root2].w);
        load_child(0, bvh2_left(A, n));
        load_child(1, bvh2_right(A, n));
        cnt = 2;
    }
    // open the interior child with the largest surface area until the treelet has 8 leaves
    while (cnt < 8) {
        int best = -1;
        float best_a = -1.f;
        for (int i = 0; i < cnt; i++)
            if (area[i] > best_a) {
                best_a = area[i];
                best = i;
            }
        if (best < 0) break;
        const uint32_t A = cA[best];
        load_child(best, bvh2_left(A, n));
        load_child(cnt, bvh2_right(A, n));
        cnt++;
    }

    // node box and triangle ranges (second read of the records)
    float nlo[3] = {kInf, kInf, kInf}, nhi[3] = {-kInf, -kInf, -kInf};
    int tfirst[8];
    unsigned tcount4 = 0; // 4 bits per child: 1..3 triangles, 4 = "more than 3": an interior child of the BVH8 node
    for (int i = 0; i < cnt; i++) {
        const float4 lo = a.t.rec[2 * c[i]], hi = a.t.rec[2 * c[i] + 1];
        nlo[0] = fminf(nlo[0], lo.x); nlo[1] = fminf(nlo[1], lo.y); nlo[2] = fminf(nlo[2], lo.z);
        nhi[0] = fmaxf(nhi[0], hi.x); nhi[1] = fmaxf(nhi[1], hi.y); nhi[2] = fmaxf(nhi[2], hi.z);
        int tc;
        if (c[i] >= n - 1) {
            tfirst[i] = c[i] - (n - 1);
            tc = 1;
        } else {
            const uint32_t B = __float_as_uint(hi.w);
            tfirst[i] = int(B & 0x3fffffffu);
            tc = int(B >> 30) + 1;
        }
        tcount4 |= unsigned(tc) << (4 * i);
    }

    // octant-ordered slots: slot s prefers the child farthest along (+-x, +-y, +-z) by the bits
    // of s (4 = +x, 2 = +y, 1 = +z); traversal visits slots in (slot ^ octant) order, so near
    // children of any ray octant come first.  The best score a child can get is the L1 norm of its offset from
    // the node centre (in its own octant's slot): children are taken in the order of that norm, largest first,
    // and each gets the best slot still free -- 64 score evaluations per node instead of the 512 of a greedy
    // search over all (child, slot) pairs, with the same outcome whenever the octants do not collide.
    unsigned child_in_slot4 = 0x88888888u; // 4 bits per slot: the child in it, 8 = empty
    {
        const float pcx = 0.5f * (nlo[0] + nhi[0]), pcy = 0.5f * (nlo[1] + nhi[1]), pcz = 0.5f * (nlo[2] + nhi[2]);
        float norm[8];
        for (int i = 0; i < cnt; i++) {
            const float4 lo = a.t.rec[2 * c[i]], hi = a.t.rec[2 * c[i] + 1];
            const float l1 = fabsf(0.5f * (lo.x + hi.x) - pcx) + fabsf(0.5f * (lo.y + hi.y) - pcy) + fabsf(0.5f * (lo.z + hi.z) - pcz);
            norm[i] = (l1 == l1) ? l1 : -1.f; // non-finite centroids go last
        }
        unsigned used_slots = 0, done = 0;
        for (int round = 0; round < cnt; round++) {
            int bi = 0;
            float bn = -kInf;
            for (int i = 0; i < cnt; i++)
                if (!(done & (1u << i)) && norm[i] > bn) {
                    bn = norm[i];
                    bi = i;
                }
            if (bn == -kInf) // every remaining norm is NaN-like: first child left
                for (int i = cnt - 1; i >= 0; i--)
                    if (!(done & (1u << i))) bi = i;
            done |= 1u << bi;
            const float4 lo = a.t.rec[2 * c[bi]], hi = a.t.rec[2 * c[bi] + 1];
            const float ax = 0.5f * (lo.x + hi.x) - pcx, ay = 0.5f * (lo.y + hi.y) - pcy, az = 0.5f * (lo.z + hi.z) - pcz;
            float best = -kInf;
            int bs = -1;
#pragma unroll
            for (int sl = 0; sl < 8; sl++) {
                const float cost = ((sl & 4) ? ax : -ax) + ((sl & 2) ? ay : -ay) + ((sl & 1) ? az : -az);
                if (!(used_slots & (1u << sl)) && cost > best) {
                    best = cost;
                    bs = sl;
                }
            }
            if (bs < 0) bs = __ffs(~used_slots & 0xffu) - 1; // non-finite scores: first free slot
            child_in_slot4 = (child_in_slot4 & ~(0xfu << (4 * bs))) | (unsigned(bi) << (4 * bs));
            used_slots |= 1u << bs;
        }
    }

    // allocate child nodes, triangle slots and queue entries
    unsigned imask = 0;
    int n_inner = 0, n_tris = 0;
#pragma unroll
    for (int s = 0; s < 8; s++) {
        const unsigned i = (child_in_slot4 >> (4 * s)) & 0xfu;
        if (i > 7u) continue;
        const int tc = int((tcount4 >> (4 * i)) & 0xfu);
        if (tc > 3) {
            imask |= 1u << s;
            n_inner++;
        } else {
            n_tris += tc;
        }
    }
    unsigned child_base = 0, tri_base = 0, q_base = 0;
    if (n_inner) {
        child_base = atomicAdd(&a.counters[0], (unsigned)n_inner);
        q_base = atomicAdd(&a.counters[4 + (a.parity ^ 1)], (unsigned)n_inner);
        if (child_base + n_inner > a.node_capacity) {
            a.counters[2] = 1;
            return;
        }
    }
    if (n_tris) tri_base = atomicAdd(&a.counters[1], (unsigned)n_tris);

    // quantisation grid: p = node min, 2^e >= extent / 255 per axis
    int e[3];
    double scale[3], inv_scale[3];
    for (int k = 0; k < 3; k++) {
        const double ext = (double)nhi[k] - (double)nlo[k];
        int ek = -100;
        if (ext > 0) {
            // smallest ek with 255 * 2^ek >= ext: ceil(log2(ext / 255)) from the binary exponent (frexp: x = m 2^k, m in [0.5, 1))
            int kx;
            const double m = frexp(ext / 255.0, &kx);
            ek = (m == 0.5) ? kx - 1 : kx;
            ek = max(-100, min(127, ek));
            while (ek < 127 && 255.0 * pow2_double(ek) < ext) ek++; // the division above rounds: make the bound exact
        }
        e[k] = ek;
        scale[k] = pow2_double(ek);
        inv_scale[k] = pow2_double(-ek);
    }

    // the node's byte planes, in slot order, straight into the words that are stored (third read of the records)
    uint32_t w_meta[2] = {0u, 0u}, w_lo[3][2] = {{0u, 0u}, {0u, 0u}, {0u, 0u}}, w_hi[3][2] = {{0u, 0u}, {0u, 0u}, {0u, 0u}};
    int tri_off = 0, inner_rank = 0;
#pragma unroll
    for (int s = 0; s < 8; s++) {
        const int h = s >> 2, sh = 8 * (s & 3);
        const unsigned i = (child_in_slot4 >> (4 * s)) & 0xfu;
        if (i > 7u) { // empty slot: inverted box never hit (meta 0, mins 255, maxes 0)
            for (int k = 0; k < 3; k++) w_lo[k][h] |= 255u << sh;
            continue;
        }
        const float4 lo = a.t.rec[2 * c[i]], hi = a.t.rec[2 * c[i] + 1];
        const float clo[3] = {lo.x, lo.y, lo.z}, chi[3] = {hi.x, hi.y, hi.z};
        for (int k = 0; k < 3; k++) {
            const double l = ((double)clo[k] - (double)nlo[k]) * inv_scale[k];
            const double hh = ((double)chi[k] - (double)nlo[k]) * inv_scale[k];
            int ql = (int)floor(l), qh = (int)ceil(hh);
            ql = max(0, min(255, ql));
            qh = max(0, min(255, qh));
            // exactness guard (all terms are exact in double): p + ql*2^e <= lo, p + qh*2^e >= hi
            while (ql > 0 && (double)nlo[k] + ql * scale[k] > (double)clo[k]) ql--;
            while (qh < 255 && (double)nlo[k] + qh * scale[k] < (double)chi[k]) qh++;
            w_lo[k][h] |= unsigned(ql) << sh;
            w_hi[k][h] |= unsigned(qh) << sh;
        }
        const int tc = int((tcount4 >> (4 * i)) & 0xfu);
        if (imask & (1u << s)) {
            w_meta[h] |= ((1u << 5) | (24u + s)) << sh;
            a.q_out[q_base + inner_rank] = make_int2(c[i], (int)(child_base + inner_rank));
            inner_rank++;
        } else {
            const unsigned unary = (1u << tc) - 1u; // 1, 3, 7
            w_meta[h] |= ((unary << 5) | (unsigned)tri_off) << sh;
            for (int k = 0; k < tc; k++) { // the leaf records exist already, in sorted order: a contiguous run moves over
                const float4 *src = a.tri_sorted + 3 * (size_t)(tfirst[i] + k);
                float4 *dst = a.tris + 3 * (size_t)(tri_base + tri_off + k);
                dst[0] = src[0];
                dst[1] = src[1];
                dst[2] = src[2];
            }
            tri_off += tc;
        }
    }

    auto f = [](uint32_t v) { return __uint_as_float(v); };
    const unsigned ebits = (unsigned)(e[0] + 127) | ((unsigned)(e[1] + 127) << 8) | ((unsigned)(e[2] + 127) << 16) | (imask << 24);
    float4 *dst = a.nodes + 5 * (size_t)work.y;
    dst[0] = make_float4(nlo[0], nlo[1], nlo[2], f(ebits));
    dst[1] = make_float4(f(child_base), f(tri_base), f(w_meta[0]), f(w_meta[1]));
    dst[2] = make_float4(f(w_lo[0][0]), f(w_lo[0][1]), f(w_lo[1][0]), f(w_lo[1][1]));
    dst[3] = make_float4(f(w_lo[2][0]), f(w_lo[2][1]), f(w_hi[0][0]), f(w_hi[0][1]));
    dst[4] = make_float4(f(w_hi[1][0]), f(w_hi[1][1]), f(w_hi[2][0]), f(w_hi[2][1]));
}

__global__ void __launch_bounds__(128, 8) k_collapse_level(CollapseArgs a) {
    const int count = (int)a.counters[4 + a.parity];
    for (int item = blockIdx.x * blockDim.x + threadIdx.x; item < count; item += gridDim.x * blockDim.x) collapse_item(a, item);
}

// between two levels: the consumed queue's counter becomes the empty out-counter of the level after next
__global__ void k_collapse_advance(unsigned int *counters, int parity) {
    if (counters[4 + parity]) counters[3] += 1;
    counters[4 + parity] = 0;
}

__global__ void k_seed_queue(int2 *q, unsigned int *counters, int root2, unsigned node_base, unsigned leaf_base) {
    q[0] = make_int2(root2, (int)node_base);
    counters[0] = node_base + 1; // the first node of the range = its root
    counters[1] = leaf_base;
    counters[2] = 0;
    counters[3] = 0;
    counters[4] = 1; // one item in queue 0
    counters[5] = 0;
}

// ------------------------------------------------------------------------------ structural validation
// Direct check of what the traversal relies on (b200pt_bvh_validate): every child box of every BVH8 node, dequantised the way
// intersect_node8 reads it (p + q * 2^e, mins from qlo, maxes from qhi), contains
//   - all vertices of the triangles of a leaf child, and
//   - the whole box of an inner child's node (the union of that node's own dequantised child boxes);
// and every triangle slot is referenced by exactly one leaf child.  out[0] = boxes that fail, out[1] = triangle references,
// out[2] = nodes visited.  One thread per node.
__global__ void __launch_bounds__(128) k_validate(const float4 *__restrict__ nodes, const float4 *__restrict__ tris, unsigned first_node,
                                                  unsigned n_nodes, unsigned long long *out) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const float4 *np = nodes + 5 * size_t(first_node + i);
    const float4 n0 = np[0], n1 = np[1], n2 = np[2], n3 = np[3], n4 = np[4];
    const uint32_t eb = __float_as_uint(n0.w);
    const double sx = __uint_as_float((eb & 0xffu) << 23), sy = __uint_as_float(((eb >> 8) & 0xffu) << 23), sz = __uint_as_float(((eb >> 16) & 0xffu) << 23);
    auto byte_of = [](float lo, float hi, int s) { return (__float_as_uint(s < 4 ? lo : hi) >> (8 * (s & 3))) & 0xffu; };
    unsigned long long bad = 0, refs = 0;
    const uint32_t child_base = __float_as_uint(n1.x), tri_base = __float_as_uint(n1.y), imask = eb >> 24;
    for (int s = 0; s < 8; s++) {
        const uint32_t meta = byte_of(n1.z, n1.w, s);
        if (meta == 0u) continue;
        // exact in double: a float plus an 8-bit multiple of a power of two
        const double lo[3] = {n0.x + byte_of(n2.x, n2.y, s) * sx, n0.y + byte_of(n2.z, n2.w, s) * sy, n0.z + byte_of(n3.x, n3.y, s) * sz};
        const double hi[3] = {n0.x + byte_of(n3.z, n3.w, s) * sx, n0.y + byte_of(n4.x, n4.y, s) * sy, n0.z + byte_of(n4.z, n4.w, s) * sz};
        bool ok = true;
        if (imask & (1u << s)) {
            if ((meta >> 5) != 1u || (meta & 0x1fu) != 24u + s) ok = false;
            const uint32_t rel = __popc(imask & ((1u << s) - 1u));
            const float4 *cp = nodes + 5 * size_t(child_base + rel);
            const float4 c0 = cp[0], c1 = cp[1], c2 = cp[2], c3 = cp[3], c4 = cp[4];
            const uint32_t ceb = __float_as_uint(c0.w);
            const double cs[3] = {__uint_as_float((ceb & 0xffu) << 23), __uint_as_float(((ceb >> 8) & 0xffu) << 23), __uint_as_float(((ceb >> 16) & 0xffu) << 23)};
            for (int t = 0; t < 8; t++) {
                if (byte_of(c1.z, c1.w, t) == 0u) continue;
                const double clo[3] = {c0.x + byte_of(c2.x, c2.y, t) * cs[0], c0.y + byte_of(c2.z, c2.w, t) * cs[1], c0.z + byte_of(c3.x, c3.y, t) * cs[2]};
                const double chi[3] = {c0.x + byte_of(c3.z, c3.w, t) * cs[0], c0.y + byte_of(c4.x, c4.y, t) * cs[1], c0.z + byte_of(c4.z, c4.w, t) * cs[2]};
                // mins: exact (the child's grid starts at its own box minimum, which this box bounds); maxes: the child rounds its
                // own boxes up on ITS grid, so they may pass the parent's bound by less than one child step
                for (int a = 0; a < 3; a++)
                    if (clo[a] < lo[a] || chi[a] >= hi[a] + cs[a]) ok = false;
            }
        } else {
            const int off = meta & 0x1f, count = __popc(meta >> 5);
            if (off + count > 24) ok = false;
            for (int k = 0; k < count; k++) {
                const float4 *tp = tris + 3 * size_t(tri_base + off + k);
                for (int v = 0; v < 3; v++) {
                    const float4 q = tp[v];
                    if (q.x < lo[0] || q.x > hi[0] || q.y < lo[1] || q.y > hi[1] || q.z < lo[2] || q.z > hi[2]) ok = false;
                }
                refs++;
            }
        }
        if (!ok) bad++;
    }
    if (bad) atomicAdd(out, bad);
    if (refs) atomicAdd(out + 1, refs);
    atomicAdd(out + 2, 1ull);
}

int validate_bvh(b200pt_scene *s, int64_t out[3]) {
    unsigned long long *d = nullptr;
    CU_TRY(cudaMalloc(&d, 3 * sizeof(unsigned long long)));
    CU_TRY(cudaMemsetAsync(d, 0, 3 * sizeof(unsigned long long), s->stream));
    const unsigned n_all = s->n_nodes_total; // the top-level tree and every object tree: node indices are dense
    if (n_all > 0) k_validate<<<(n_all + 127) / 128, 128, 0, s->stream>>>(s->d_nodes, s->d_tris, 0, n_all, d);
    unsigned long long h[3] = {0, 0, 0};
    CU_TRY(cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    cudaFree(d);
    for (int k = 0; k < 3; k++) out[k] = (int64_t)h[k];
    return B200PT_OK;
}

// ------------------------------------------------------------------------------ host driver
namespace {
// one cudaMalloc for all build scratch, carved into 256-byte aligned pieces (every cudaMalloc costs
// 0.2-0.4 ms of host time with the stream idle -- more than the whole Morton + sort phase at 1 M triangles)
struct Arena {
    char *base = nullptr;
    size_t size = 0, used = 0;
    bool borrowed = false; // the scene's kept scratch (b200pt_scene::build_scratch): not freed here
    ~Arena() {
        if (base && !borrowed) cudaFree(base);
    }
    static size_t pad(size_t b) { return (b + 255) & ~size_t(255); }
    template <typename T>
    void take(T **out, size_t count) {
        *out = reinterpret_cast<T *>(base + used);
        used += pad(std::max<size_t>(count, 1) * sizeof(T));
    }
};
} // namespace

// One BVH8 over the triangles [first, first + n) of the scene (their ids in the leaf records stay GLOBAL), written into the
// shared node / leaf-record arrays starting at node_base / leaf_base: the top-level tree of the scene, or the tree of one
// instanced object (scene_builder.cu:809-876 keeps one HLBVH per object definition too).
struct RangeOut {
    unsigned n_nodes = 0;
    float box[6] = {0, 0, 0, 0, 0, 0}; // bounds of the range's triangles
};

// proxies: the range is not scene geometry but the instance-box proxies (finish_instances): their own vertex / index buffers, no
// per-triangle attributes; `first` still gives the ids written into the leaf records
static int build_range(b200pt_scene *s, int64_t first, int64_t n, unsigned node_base, unsigned node_capacity, unsigned leaf_base,
                       bool keep_sorted, b200pt_build_stats &bs, RangeOut &out, const float *proxy_P = nullptr,
                       const int32_t *proxy_idx = nullptr) {
    cudaStream_t st = s->stream;
    const bool proxies = proxy_P != nullptr;
    const float *range_P = proxies ? proxy_P : s->d_P;
    const int32_t *range_idx = proxies ? proxy_idx : s->d_idx + 3 * first;
    // key layout: the reference's 30-bit code above a 32-bit primitive id (N <= 16 Mi), else the widest Morton
    // code that fits above ceil(log2 N) primitive-id bits; both sort as 8-byte keys with no payload
    const bool key30 = n <= (int64_t(1) << 24);
    int prim_bits = 32, axis_bits = 10;
    if (!key30) {
        prim_bits = 1;
        while ((int64_t(1) << prim_bits) < n) prim_bits++;
        axis_bits = std::min(21, (64 - prim_bits) / 3);
    }
    const int sort_lo = prim_bits, sort_hi = prim_bits + ((3 * axis_bits + 7) / 8) * 8;
    const int n_passes = (sort_hi - sort_lo) / 8;

    cudaEvent_t ev[7];
    for (auto &e : ev) CU_TRY(cudaEventCreate(&e));
    struct EvFree {
        cudaEvent_t *e;
        ~EvFree() { for (int i = 0; i < 7; i++) cudaEventDestroy(e[i]); }
    } evfree{ev};

    uint64_t *d_sorted_keys = nullptr;
    uint32_t *d_sorted_prim = nullptr;
    struct SortedFree { // the sorted order of the top-level range is kept for the build-stage parity tests; other ranges drop it
        uint64_t **k;
        uint32_t **p;
        bool keep;
        ~SortedFree() {
            if (!keep) {
                if (*k) cudaFree(*k);
                if (*p) cudaFree(*p);
            }
        }
    } sorted_free{&d_sorted_keys, &d_sorted_prim, keep_sorted};
    if (keep_sorted && s->d_sorted_keys && s->d_sorted_prim && s->sorted_cap >= n && s->sorted_cap <= 2 * n + 1024) {
        d_sorted_keys = s->d_sorted_keys; // the previous build's arrays
        d_sorted_prim = s->d_sorted_prim;
    } else {
        if (keep_sorted) {
            if (s->d_sorted_keys) cudaFree(s->d_sorted_keys);
            if (s->d_sorted_prim) cudaFree(s->d_sorted_prim);
            s->d_sorted_keys = nullptr;
            s->d_sorted_prim = nullptr;
            s->sorted_cap = 0;
        }
        CU_TRY(cudaMalloc(&d_sorted_keys, n * sizeof(uint64_t)));
        if (keep_sorted) s->d_sorted_keys = d_sorted_keys; // owned by the scene from here on, whatever happens next
        CU_TRY(cudaMalloc(&d_sorted_prim, n * sizeof(uint32_t)));
        if (keep_sorted) {
            s->d_sorted_prim = d_sorted_prim;
            s->sorted_cap = n;
        }
    }
    Arena tmp;
    {
        const size_t sz = Arena::pad(sort_onesweep_bytes(n, n_passes)) + Arena::pad(12 * 4) * 2 + Arena::pad(n * 8) + Arena::pad(256 * sort_tiles(n) * 4) +
                          Arena::pad(sort_scan_tmp_words(n) * 4) + Arena::pad(4 * n * sizeof(float4)) + 3 * Arena::pad(n * 4) + 2 * Arena::pad(3 * n * sizeof(float4)) +
                          Arena::pad(2 * n * 4) + Arena::pad(n * 4) + 2 * Arena::pad(n * sizeof(int2)) + Arena::pad(8 * 4);
        if (s->build_scratch && s->build_scratch_bytes >= sz) {
            tmp.base = s->build_scratch;
            tmp.borrowed = true;
        } else if (sz <= (size_t)s->tune_build_scratch_mb << 20) {
            if (s->build_scratch) cudaFree(s->build_scratch);
            s->build_scratch = nullptr;
            s->build_scratch_bytes = 0;
            CU_TRY(cudaMalloc(&tmp.base, sz));
            s->build_scratch = tmp.base;
            s->build_scratch_bytes = sz;
            tmp.borrowed = true;
        } else {
            CU_TRY(cudaMalloc(&tmp.base, sz));
        }
        tmp.size = sz;
    }
    uint32_t *d_red;
    float *d_boxes;
    tmp.take(&d_red, 12);
    tmp.take(&d_boxes, 12);
    SortBuffers sb;
    uint64_t *d_keys_tmp;
    tmp.take(&d_keys_tmp, n);
    // the sort ping-pongs between two buffers; the one the last pass writes is the kept output
    sb.keys[n_passes & 1] = d_sorted_keys;
    sb.keys[(n_passes & 1) ^ 1] = d_keys_tmp;
    tmp.take(&sb.tile_hist, 256 * sort_tiles(n));
    tmp.take(&sb.scan_tmp, sort_scan_tmp_words(n));
    {
        static const int onesweep = getenv("B200PT_SORT_ONESWEEP") ? atoi(getenv("B200PT_SORT_ONESWEEP")) : 1;
        unsigned long long *os = nullptr;
        tmp.take(&os, (sort_onesweep_bytes(n, n_passes) + 7) / 8);
        sb.onesweep = onesweep ? os : nullptr;
    }
    float4 *d_tri_sorted, *d_tri_unsorted;
    tmp.take(&d_tri_sorted, 3 * n);
    tmp.take(&d_tri_unsorted, 3 * n);
    Bvh2 t;
    t.n = int32_t(n);
    tmp.take(&t.rec, 4 * n);
    tmp.take(&t.nodeA, n);
    tmp.take(&t.nodeB, n);
    tmp.take(&t.last, n);
    tmp.take(&t.parent, 2 * n);
    tmp.take(&t.arrive, n);
    int2 *q[2];
    unsigned int *d_counters;
    tmp.take(&q[0], n);
    tmp.take(&q[1], n);
    tmp.take(&d_counters, 8);
    // the refit's root list lives in the collapse's second queue (2 n ints; the collapse starts after the refit is done)
    int32_t *d_roots = reinterpret_cast<int32_t *>(q[1]);
    unsigned *d_n_roots = d_counters + 7;
    if (tmp.used > tmp.size) {
        set_error("b200pt_bvh_build: scratch arena overrun");
        return B200PT_ERR_STATE;
    }

    const int tb = 256;
    const int grid_n = int((n + tb - 1) / tb);
    CU_TRY(cudaEventRecord(ev[0], st));
    // 1. centroid box + Morton keys
    k_init_reduction<<<1, 32, 0, st>>>(d_red);
    GatherArgs ga;
    ga.P = range_P;
    ga.idx = range_idx;
    ga.tri_material = (s->d_tri_material && !proxies) ? s->d_tri_material + first : nullptr;
    ga.tri_light = (s->d_tri_light && !proxies) ? s->d_tri_light + first : nullptr;
    ga.tri_flip = (s->d_tri_flip && !proxies) ? s->d_tri_flip + first : nullptr;
    ga.prim_offset = (int32_t)first;
    k_centroid_bounds<<<std::min<int64_t>(grid_n, s->sm_count * 16), tb, 0, st>>>(ga, n, d_red, d_tri_unsorted);
    k_cubify<<<1, 1, 0, st>>>(d_red, d_boxes);
    if (key30) k_morton<true><<<grid_n, tb, 0, st>>>(d_tri_unsorted, n, d_boxes, sb.keys[0], axis_bits, prim_bits);
    else k_morton<false><<<grid_n, tb, 0, st>>>(d_tri_unsorted, n, d_boxes, sb.keys[0], axis_bits, prim_bits);
    CU_TRY(cudaEventRecord(ev[1], st));
    // 2. LSD radix sort (key-only)
    const int res = radix_sort_u64(sb, n, sort_lo, sort_hi, st, &bs.sort_passes, s->sm_count);
    if (sb.keys[res] != d_sorted_keys) {
        set_error("b200pt_bvh_build: sort ended in the wrong buffer");
        return B200PT_ERR_STATE;
    }
    k_gather_sorted<<<grid_n, tb, 0, st>>>(d_sorted_keys, n, key30 ? 0xffffffffu : uint32_t((uint64_t(1) << prim_bits) - 1), d_tri_unsorted,
                                           d_sorted_prim, d_tri_sorted);
    CU_TRY(cudaEventRecord(ev[2], st));
    // 3. hierarchy emission
    CU_TRY(cudaMemsetAsync(t.arrive, 0, n * sizeof(uint32_t), st));
    if (n > 1) k_emit_hierarchy<<<int((n - 1 + tb - 1) / tb), tb, 0, st>>>(d_sorted_keys, t);
    CU_TRY(cudaEventRecord(ev[3], st));
    // 4. refit
    CU_TRY(cudaMemsetAsync(d_n_roots, 0, sizeof(unsigned), st));
    k_refit<<<int((n + REFIT_BLOCK - 1) / REFIT_BLOCK), REFIT_BLOCK, 0, st>>>(d_tri_sorted, t, d_roots, d_n_roots);
    k_refit_climb<<<s->sm_count * 8, 128, 0, st>>>(t, d_roots, d_n_roots);
    CU_TRY(cudaEventRecord(ev[4], st));
    // 5. work-queue collapse to BVH8: levels are enqueued in batches, queue sizes stay on the device
    k_seed_queue<<<1, 1, 0, st>>>(q[0], d_counters, 0 /* leaf id n-1+0 == 0 when n == 1 */, node_base, leaf_base);
    unsigned int h_counters[8] = {node_base + 1, leaf_base, 0, 0, 1, 0, 0, 0};
    int parity = 0, levels = 0;
    const int collapse_grid = s->sm_count * 16;
    // resident CTAs per SM are capped through a dynamic shared-memory request: each thread of k_collapse_level keeps ~600 B of
    // indexed per-child state in local memory, and too many resident threads push that state out of L1 into L2 / DRAM
    static const int collapse_smem = getenv("B200PT_COLLAPSE_SMEM") ? atoi(getenv("B200PT_COLLAPSE_SMEM")) : 0;
    if (collapse_smem > 48 * 1024) cudaFuncSetAttribute(k_collapse_level, cudaFuncAttributeMaxDynamicSharedMemorySize, collapse_smem);
    while (true) {
        const int batch = levels == 0 ? 12 : 6;
        for (int b = 0; b < batch; b++) {
            CollapseArgs a;
            a.t = t;
            a.tri_sorted = d_tri_sorted;
            a.nodes = s->d_nodes;
            a.tris = s->d_tris;
            a.q_in = q[parity];
            a.q_out = q[parity ^ 1];
            a.counters = d_counters;
            a.parity = parity;
            a.node_capacity = node_capacity;
            k_collapse_level<<<collapse_grid, 128, collapse_smem, st>>>(a);
            k_collapse_advance<<<1, 1, 0, st>>>(d_counters, parity);
            parity ^= 1;
            levels++;
        }
        CU_TRY(cudaMemcpyAsync(h_counters, d_counters, sizeof(h_counters), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        if (h_counters[2]) {
            set_error("b200pt_bvh_build: BVH8 node pool overflow (capacity %u)", node_capacity);
            return B200PT_ERR_STATE;
        }
        if (h_counters[4 + parity] == 0) break;
        if (levels > 4096) {
            set_error("b200pt_bvh_build: collapse did not terminate");
            return B200PT_ERR_STATE;
        }
    }
    CU_TRY(cudaEventRecord(ev[5], st));
    float boxes[12];
    CU_TRY(cudaMemcpyAsync(boxes, d_boxes, sizeof(boxes), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    for (int k = 0; k < 6; k++) out.box[k] = boxes[6 + k];
    out.n_nodes = h_counters[0] - node_base;
    if (int64_t(h_counters[1]) - leaf_base != n) {
        set_error("b200pt_bvh_build: %lld of %lld triangles placed", (long long)(int64_t(h_counters[1]) - leaf_base), (long long)n);
        return B200PT_ERR_STATE;
    }
    float ms[6];
    for (int k = 0; k < 5; k++) CU_TRY(cudaEventElapsedTime(&ms[k], ev[k], ev[k + 1]));
    CU_TRY(cudaEventElapsedTime(&ms[5], ev[0], ev[5]));
    bs.ms_bounds_morton += ms[0];
    bs.ms_sort += ms[1];
    bs.ms_emit += ms[2];
    bs.ms_refit += ms[3];
    bs.ms_collapse += ms[4];
    bs.ms_total += ms[5];
    bs.n_bvh2_nodes += 2 * n - 1;
    bs.collapse_levels = std::max(bs.collapse_levels, int32_t(h_counters[3]));
    bs.key_bits = std::max(bs.key_bits, 3 * axis_bits);
    return B200PT_OK;
}

int build_bvh(b200pt_scene *s, b200pt_build_stats *stats) {
    const int64_t n = s->n_triangles;
    b200pt_build_stats bs = {};
    bs.n_triangles = n;
    // traversals of the previous tree may still be in flight on the scene's streams (the cudaFree calls this replaces waited too)
    CU_TRY(cudaDeviceSynchronize());
    s->bvh = BVH8();
    s->built = false;
    if (n == 0) {
        if (s->d_nodes) { cudaFree(s->d_nodes); s->d_nodes = nullptr; s->d_nodes_bytes = 0; }
        if (s->d_tris) { cudaFree(s->d_tris); s->d_tris = nullptr; s->d_tris_bytes = 0; }
        if (s->d_sorted_keys) { cudaFree(s->d_sorted_keys); s->d_sorted_keys = nullptr; }
        if (s->d_sorted_prim) { cudaFree(s->d_sorted_prim); s->d_sorted_prim = nullptr; }
        s->sorted_cap = 0;
        s->built = true;
        if (stats) *stats = bs;
        return B200PT_OK;
    }
    if (n > (int64_t(1) << 30)) {
        set_error("b200pt_bvh_build: %lld triangles exceed the 2^30 limit", (long long)n);
        return B200PT_ERR_UNSUPPORTED;
    }
    const auto wall0 = std::chrono::steady_clock::now();
    // ranges: the top-level triangles [0, n_top), then one range per instanced object (b200pt_scene_desc.objects)
    const int64_t n_top = s->objects.empty() ? n : s->n_top_triangles;
    const int64_t n_inst = (int64_t)s->instances.size();
    size_t cap = size_t(n_top / 2 + 64);
    for (const auto &o : s->objects) cap += size_t(o.n_triangles / 2 + 64);
    if (n_inst > 0) cap += size_t(n_inst / 2 + 64); // the tree over the instance boxes
    const unsigned node_capacity = unsigned(cap);
    const size_t nodes_bytes = size_t(node_capacity) * 5 * sizeof(float4);
    const size_t tris_bytes = size_t(n + n_inst) * 3 * sizeof(float4); // + one proxy leaf record per instance
    if (!s->d_nodes || s->d_nodes_bytes < nodes_bytes || s->d_nodes_bytes > 2 * nodes_bytes + (1 << 20)) {
        if (s->d_nodes) cudaFree(s->d_nodes);
        s->d_nodes = nullptr;
        s->d_nodes_bytes = 0;
        CU_TRY(cudaMalloc(&s->d_nodes, nodes_bytes));
        s->d_nodes_bytes = nodes_bytes;
    }
    if (!s->d_tris || s->d_tris_bytes < tris_bytes || s->d_tris_bytes > 2 * tris_bytes + (1 << 20)) {
        if (s->d_tris) cudaFree(s->d_tris);
        s->d_tris = nullptr;
        s->d_tris_bytes = 0;
        CU_TRY(cudaMalloc(&s->d_tris, tris_bytes));
        s->d_tris_bytes = tris_bytes;
    }
    unsigned node_base = 0;
    RangeOut ro;
    if (n_top > 0) {
        const int rc = build_range(s, 0, n_top, 0, node_capacity, 0, true, bs, ro);
        if (rc) return rc;
        node_base = ro.n_nodes;
        for (int k = 0; k < 6; k++) s->scene_bounds[k] = ro.box[k];
    }
    s->n_top_nodes = node_base;
    for (auto &o : s->objects) {
        RangeOut oo;
        const int rc = build_range(s, o.first_triangle, o.n_triangles, node_base, node_capacity, unsigned(o.first_triangle), false, bs, oo);
        if (rc) return rc;
        o.root_node = node_base;
        for (int k = 0; k < 6; k++) o.box[k] = oo.box[k];
        node_base += oo.n_nodes;
    }
    // the instance tree: one proxy triangle per instance whose bounds are the instance's padded render-space box
    // ((lo, hi, (lo.x, hi.y, lo.z)) has exactly that AABB), built like any other range; its leaf records hold n + instance index
    s->inst_root_node = -1;
    if (int rc = finish_instances(s)) return rc;
    if (n_inst > 0) {
        std::vector<float> pv(size_t(n_inst) * 9);
        std::vector<int32_t> pi(size_t(n_inst) * 3);
        for (int64_t i = 0; i < n_inst; i++) {
            const float *b = &s->inst_boxes[6 * i];
            const float v[9] = {b[0], b[1], b[2], b[3], b[4], b[5], b[0], b[4], b[2]};
            for (int k = 0; k < 9; k++) pv[9 * i + k] = v[k];
            for (int k = 0; k < 3; k++) pi[3 * i + k] = int32_t(3 * i + k);
        }
        float *d_pv = nullptr;
        int32_t *d_pi = nullptr;
        CU_TRY(cudaMalloc(&d_pv, pv.size() * sizeof(float)));
        CU_TRY(cudaMalloc(&d_pi, pi.size() * sizeof(int32_t)));
        CU_TRY(cudaMemcpyAsync(d_pv, pv.data(), pv.size() * sizeof(float), cudaMemcpyHostToDevice, s->stream));
        CU_TRY(cudaMemcpyAsync(d_pi, pi.data(), pi.size() * sizeof(int32_t), cudaMemcpyHostToDevice, s->stream));
        RangeOut io;
        b200pt_build_stats ibs = {}; // the proxy tree's stage times are not part of the triangle build's figures
        const int rc = build_range(s, n, n_inst, node_base, node_capacity, unsigned(n), false, ibs, io, d_pv, d_pi);
        cudaFree(d_pv);
        cudaFree(d_pi);
        if (rc) return rc;
        s->inst_root_node = (int32_t)node_base;
        node_base += io.n_nodes;
    }
    bs.n_bvh8_nodes = node_base;
    s->n_nodes_total = node_base;
    bs.ms_wall = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    s->bvh.nodes = s->d_nodes;
    s->bvh.tris = s->d_tris;
    s->bvh.n_nodes = s->n_top_nodes; // the top-level tree; object trees are entered through their instances
    s->bvh.n_tris = n;
    s->key_bits = bs.key_bits;
    s->built = true;
    bs.device_bytes = int64_t(node_capacity) * 80 + n * 48 + n * 12;
    if (stats) *stats = bs;
    return B200PT_OK;
}

} // namespace b200pt
