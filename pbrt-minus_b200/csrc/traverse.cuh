// traverse.cuh -- per-ray BVH8 traversal (device function shared by the trace kernels and the
// wavefront stages).  Replaces HLBVH::intersect / fast_intersect (accelerator/hlbvh.cu:407-516):
// instead of a 128-entry local-memory stack of BVH2 nodes with a 7-deep pointer chase per leaf,
// a ray walks 80-byte quantised 8-wide nodes (5 x 128-bit loads), keeps a short stack of
// (node-group, hit-mask) pairs in shared memory, and tests 48-byte triangle records with the
// reference's exact watertight arithmetic (triangle.cuh), which is what defines the hit id.
#pragma once

#include "common.cuh"
#include "triangle.cuh"

namespace b200pt {

constexpr int TRACE_BLOCK = 128;
constexpr int SM_STACK = 10;    // entries per thread in shared memory
constexpr int LOCAL_STACK = 86; // overflow entries in local memory (rarely touched)

struct RayHit {
    int prim; // -1 = miss
    float t, b1, b2;
};

struct TraceCounters {
    unsigned nodes = 0, tris = 0;
};

__device__ __forceinline__ float safe_rcp_dir(float d) {
    // 1/d with |d| clamped away from zero so slab distances stay finite (no 0*inf NaNs)
    const float c = fabsf(d) > 1e-18f ? d : copysignf(1e-18f, d);
    return __fdiv_rn(1.0f, c);
}

// Quantised byte -> float without the quarter-rate I2F: PRMT drops byte `byte` of `word` into
// mantissa bits 8..15 of 32768.0f, giving exactly 32768 + q on the ALU pipe; the constant is folded
// into the per-node origin term (b - 32768*a), whose rounding (<= a/256) the slack term covers.
// `bias` = 0x47000000 arrives through a kernel parameter (constant bank) so the compiler keeps the
// selector as the instruction's immediate instead of materialising 48 selector registers.
__device__ __forceinline__ float byte_to_float_biased(uint32_t word, int byte, uint32_t bias) {
    return __uint_as_float(__byte_perm(word, bias, 0x7604u | (uint32_t(byte) << 4)));
}
constexpr float kByteBias = 32768.f;

struct NodeAxis {
    float a, lo, hi; // t_near = q_near*a + lo, t_far = q_far*a + hi (with the byte bias folded in)
};

// per-axis decode terms of one node: scale a = 2^e/d, origin b = (p - o)/d, widened by the decode
// error bound k*eps*(|b| + 256|a|) + |a|/128 (conservative: never culls a box the exact arithmetic hits)
__device__ __forceinline__ NodeAxis node_axis(float p, float o, float inv_d, uint32_t e_byte) {
    constexpr float kSlack = 0x1p-20f;
    NodeAxis r;
    r.a = __uint_as_float(e_byte << 23) * inv_d;
    const float b = (p - o) * inv_d;
    const float err = fmaf(kSlack, fabsf(b), 0x1p-6f * fabsf(r.a));
    const float shift = b - kByteBias * r.a;
    r.lo = shift - err;
    r.hi = shift + err;
    return r;
}

// One BVH8 node against one ray: returns the hit mask (bits 24..31 = interior children in
// octant-sorted order, bits 0..23 = triangles relative to the node's triangle base).
__device__ __forceinline__ uint32_t intersect_node8(const float4 &n0, const float4 &n1, const float4 &n2, const float4 &n3,
                                                    const float4 &n4, float ox, float oy, float oz, float idx, float idy,
                                                    float idz, uint32_t octinv4, float t_far_limit, uint32_t bias) {
    const uint32_t ebits = __float_as_uint(n0.w);
    const NodeAxis X = node_axis(n0.x, ox, idx, ebits & 0xffu);
    const NodeAxis Y = node_axis(n0.y, oy, idy, (ebits >> 8) & 0xffu);
    const NodeAxis Z = node_axis(n0.z, oz, idz, (ebits >> 16) & 0xffu);
    const bool negx = idx < 0.f, negy = idy < 0.f, negz = idz < 0.f;
    uint32_t hitmask = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t meta4 = __float_as_uint(h ? n1.w : n1.z);
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
        const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = __float_as_uint(h ? n2.y : n2.x), qloy = __float_as_uint(h ? n2.w : n2.z);
        const uint32_t qloz = __float_as_uint(h ? n3.y : n3.x), qhix = __float_as_uint(h ? n3.w : n3.z);
        const uint32_t qhiy = __float_as_uint(h ? n4.y : n4.x), qhiz = __float_as_uint(h ? n4.w : n4.z);
        const uint32_t nx = negx ? qhix : qlox, fx = negx ? qlox : qhix;
        const uint32_t ny = negy ? qhiy : qloy, fy = negy ? qloy : qhiy;
        const uint32_t nz = negz ? qhiz : qloz, fz = negz ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float tnx = fmaf(byte_to_float_biased(nx, j, bias), X.a, X.lo), tfx = fmaf(byte_to_float_biased(fx, j, bias), X.a, X.hi);
            const float tny = fmaf(byte_to_float_biased(ny, j, bias), Y.a, Y.lo), tfy = fmaf(byte_to_float_biased(fy, j, bias), Y.a, Y.hi);
            const float tnz = fmaf(byte_to_float_biased(nz, j, bias), Z.a, Z.lo), tfz = fmaf(byte_to_float_biased(fz, j, bias), Z.a, Z.hi);
            const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.f));
            const float tf = fminf(fminf(tfx, tfy), fminf(tfz, t_far_limit));
            if (tn <= tf) {
                const uint32_t cb = (child_bits4 >> (8 * j)) & 0xffu;
                const uint32_t bi = (bit_index4 >> (8 * j)) & 0xffu;
                hitmask |= cb << bi;
            }
        }
    }
    return hitmask;
}

// does the ray (o, 1/d, [0, t_max]) touch the instance's padded render-space box?  (a leaf child of the instance tree bounds up to
// three proxies with ONE box: each candidate is tested against its own)
__device__ __forceinline__ bool ray_touches_instance(const InstRec &I, float ox, float oy, float oz, float idx, float idy, float idz,
                                                     float t_max) {
    const float *lo = I.box_lo, *hi = I.box_hi;
    const float x0 = (__ldg(lo) - ox) * idx, x1 = (__ldg(hi) - ox) * idx;
    const float y0 = (__ldg(lo + 1) - oy) * idy, y1 = (__ldg(hi + 1) - oy) * idy;
    const float z0 = (__ldg(lo + 2) - oz) * idz, z1 = (__ldg(hi + 2) - oz) * idz;
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), t_max));
    return tn <= tf * 1.00001f + 1e-30f;
}

// sm_stack: this thread's column of the CTA's shared stack, addressed as sm_stack[depth * stride].
// hit.prim = the triangle's primitive id, or (SLOT) its leaf-order record index.
template <bool ANY, bool COUNT, bool SLOT = false>
__device__ __forceinline__ bool traverse_bvh8(const BVH8 &bvh, float ox, float oy, float oz, float dx, float dy, float dz,
                                              float t_max, uint2 *sm_stack, int stride, RayHit &hit, TraceCounters &ctr,
                                              unsigned *overflow_flag, int root = -1) {
    // root: -1 walks the scene's own tree (node 0, if there is one); otherwise the tree of an instanced object rooted at that node
    hit.prim = -1;
    hit.t = t_max;
    hit.b1 = 0.f;
    hit.b2 = 0.f;
    if (root < 0 && bvh.n_nodes == 0) return false;

    const float idx = safe_rcp_dir(dx), idy = safe_rcp_dir(dy), idz = safe_rcp_dir(dz);
    const bool negx = idx < 0.f, negy = idy < 0.f, negz = idz < 0.f; // sign of the clamped reciprocal (covers -0)
    const uint32_t octinv = (negx ? 0u : 4u) | (negy ? 0u : 2u) | (negz ? 0u : 1u);
    const uint32_t octinv4 = octinv * 0x01010101u;

    uint2 lstack[LOCAL_STACK];
    int sp = 0;
    uint2 ng = make_uint2(root < 0 ? 0u : (uint32_t)root, 0x80000000u); // node group: x = child base index, y = hit bits | imask
    uint2 tg = make_uint2(0u, 0u);          // triangle group: x = tri base, y = hit bits

    while (true) {
        if (ng.y > 0x00ffffffu) {
            const uint32_t hits = ng.y;
            const int child_bit = 31 - __clz(hits);
            ng.y &= ~(1u << child_bit);
            if (ng.y > 0x00ffffffu) { // siblings left: keep the group for later
                if (sp < SM_STACK) sm_stack[sp * stride] = ng;
                else if (sp < SM_STACK + LOCAL_STACK) lstack[sp - SM_STACK] = ng;
                else *overflow_flag = 1u;
                sp++;
            }
            const uint32_t slot = (uint32_t(child_bit) - 24u) ^ octinv;
            const uint32_t rel = __popc(hits & 0xffu & ((1u << slot) - 1u));
            const float4 *np = bvh.nodes + 5 * size_t(ng.x + rel);
            const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            if (COUNT) ctr.nodes++;

            const uint32_t ebits = __float_as_uint(n0.w);
            ng.x = __float_as_uint(n1.x);
            tg.x = __float_as_uint(n1.y);
            const uint32_t hitmask = intersect_node8(n0, n1, n2, n3, n4, ox, oy, oz, idx, idy, idz, octinv4, hit.t, 0x47000000u);
            ng.y = (hitmask & 0xff000000u) | (ebits >> 24);
            tg.y = hitmask & 0x00ffffffu;
        }

        while (tg.y != 0u) {
            const int ti = __ffs(tg.y) - 1;
            tg.y &= tg.y - 1u;
            const float4 *tp = bvh.tris + 3 * size_t(tg.x + ti);
            const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
            if (COUNT) ctr.tris++;
            TriHit th;
            if (intersect_triangle(ox, oy, oz, dx, dy, dz, hit.t, v0.x, v0.y, v0.z, v1.x, v1.y, v1.z, v2.x, v2.y, v2.z, th)) {
                hit.t = th.t;
                hit.b1 = th.b1;
                hit.b2 = th.b2;
                hit.prim = SLOT ? int(tg.x + ti) : __float_as_int(v0.w);
                if (ANY) return true;
            }
        }

        if (ng.y <= 0x00ffffffu) {
            if (sp == 0) break;
            sp--;
            if (sp < SM_STACK) ng = sm_stack[sp * stride];
            else if (sp < SM_STACK + LOCAL_STACK) ng = lstack[sp - SM_STACK];
            else ng = make_uint2(0u, 0u);
        }
    }
    return hit.prim >= 0;
}

} // namespace b200pt
