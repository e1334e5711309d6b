// vecmath.cuh -- small fp32 vector / spectrum helpers for the wavefront stages.
// Arithmetic conventions follow the reference where they matter for parity:
//   gamma(n)               util/math.h:81-84
//   next_float_up/down     util/bit_arithmetic.h (ulp stepping used by ray-origin offsetting, base/ray.h:30-49)
//   coordinate_system      euclidean_space/vector3.h:161-167
//   Frame::from_xz         euclidean_space/frame.h:31-34
//   angle_between          euclidean_space/vector3.h:254-260
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace b200pt {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kOneMinusEps = 0x1.fffffep-1f;
constexpr float kMachineEps = 0x1p-24f;

__host__ __device__ constexpr float gamma_n(int n) { return (float(n) * kMachineEps) / (1.0f - float(n) * kMachineEps); }

typedef float3 V3;

__device__ __forceinline__ V3 v3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ V3 v3(const float4 &a) { return make_float3(a.x, a.y, a.z); }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float absdot(V3 a, V3 b) { return fabsf(dot(a, b)); }
__device__ __forceinline__ float len2(V3 a) { return dot(a, a); }
__device__ __forceinline__ float len(V3 a) { return sqrtf(len2(a)); }
__device__ __forceinline__ V3 normalize(V3 a) { return a / len(a); }
__device__ __forceinline__ V3 vabs(V3 a) { return v3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
__device__ __forceinline__ float sqr(float x) { return x * x; }
__device__ __forceinline__ float safe_sqrt(float x) { return sqrtf(fmaxf(0.f, x)); }
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }
__device__ __forceinline__ float lerpf(float t, float a, float b) { return (1.f - t) * a + t * b; }
__device__ __forceinline__ float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// util/math.h:167-181 (FMA-compensated products; here only accuracy matters, not bit parity)
__device__ __forceinline__ float diff_of_products(float a, float b, float c, float d) {
    const float cd = c * d;
    const float diff = fmaf(a, b, -cd);
    const float err = fmaf(-c, d, cd);
    return diff + err;
}
__device__ __forceinline__ float sum_of_products(float a, float b, float c, float d) {
    const float cd = c * d;
    const float sum = fmaf(a, b, cd);
    const float err = fmaf(c, d, -cd);
    return sum + err;
}
// euclidean_space/vector3.h:141-155: cross product with compensated components
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return v3(diff_of_products(a.y, b.z, a.z, b.y), diff_of_products(a.z, b.x, a.x, b.z), diff_of_products(a.x, b.y, a.y, b.x));
}

__device__ __forceinline__ float next_up(float v) {
    if (isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t u = __float_as_uint(v);
    u = (v >= 0.f) ? u + 1u : u - 1u;
    return __uint_as_float(u);
}
__device__ __forceinline__ float next_down(float v) {
    if (isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t u = __float_as_uint(v);
    u = (v > 0.f) ? u - 1u : u + 1u;
    return __uint_as_float(u);
}

__device__ __forceinline__ void coordinate_system(V3 v, V3 *v2, V3 *v3o) {
    const float sign = copysignf(1.0f, v.z);
    const float a = -1.0f / (sign + v.z);
    const float b = v.x * v.y * a;
    *v2 = v3(1.f + sign * sqr(v.x) * a, sign * b, -sign * v.x);
    *v3o = v3(b, sign + sqr(v.y) * a, -v.y);
}

struct Frame {
    V3 x, y, z;
    __device__ __forceinline__ V3 to_local(V3 v) const { return v3(dot(v, x), dot(v, y), dot(v, z)); }
    __device__ __forceinline__ V3 from_local(V3 v) const { return v.x * x + v.y * y + v.z * z; }
};
__device__ __forceinline__ Frame frame_from_xz(V3 x, V3 z) { return Frame{x, cross(z, x), z}; }

__device__ __forceinline__ float safe_asin(float x) { return asinf(clampf(x, -1.f, 1.f)); }
__device__ __forceinline__ float angle_between(V3 a, V3 b) {
    if (dot(a, b) < 0.f) return kPi - 2.f * safe_asin(len(a + b) / 2.f);
    return 2.f * safe_asin(len(b - a) / 2.f);
}
__device__ __forceinline__ V3 gram_schmidt(V3 v, V3 w) { return v - dot(v, w) * w; }
__device__ __forceinline__ V3 face_forward(V3 n, V3 v) { return dot(n, v) < 0.f ? -n : n; }

// ---- 4-wavelength spectrum (spectrum_util/sampled_spectrum.h)
typedef float4 S4;
__device__ __forceinline__ S4 s4(float v) { return make_float4(v, v, v, v); }
__device__ __forceinline__ S4 operator+(S4 a, S4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ S4 operator*(S4 a, S4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ S4 operator*(S4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ S4 operator*(float s, S4 a) { return a * s; }
__device__ __forceinline__ S4 operator/(S4 a, float s) { return make_float4(a.x / s, a.y / s, a.z / s, a.w / s); }
__device__ __forceinline__ S4 operator/(S4 a, S4 b) { return make_float4(a.x / b.x, a.y / b.y, a.z / b.z, a.w / b.w); }
__device__ __forceinline__ bool is_positive(S4 a) { return a.x > 0.f || a.y > 0.f || a.z > 0.f || a.w > 0.f; }
__device__ __forceinline__ float average(S4 a) { return (a.x + a.y + a.z + a.w) / 4.f; }
__device__ __forceinline__ float max_component(S4 a) { return fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w)); }
__device__ __forceinline__ S4 safe_div(S4 a, S4 b) {
    return make_float4(b.x == 0.f ? 0.f : a.x / b.x, b.y == 0.f ? 0.f : a.y / b.y, b.z == 0.f ? 0.f : a.z / b.z,
                       b.w == 0.f ? 0.f : a.w / b.w);
}
__device__ __forceinline__ S4 clamp_s4(S4 a, float lo, float hi) {
    return make_float4(clampf(a.x, lo, hi), clampf(a.y, lo, hi), clampf(a.z, lo, hi), clampf(a.w, lo, hi));
}
__device__ __forceinline__ S4 sqrt_s4(S4 a) { return make_float4(sqrtf(a.x), sqrtf(a.y), sqrtf(a.z), sqrtf(a.w)); }

// util/sampling.h:324-329
__device__ __forceinline__ float power_heuristic(float f_pdf, float g_pdf) {
    const float f = f_pdf, g = g_pdf;
    if (isinf(sqr(f))) return 1.f;
    return sqr(f) / (sqr(f) + sqr(g));
}

// base/ray.h:30-49 Ray::offset_ray_origin: offset along n by |n|.err, away from the surface on the
// side of w, then one ulp further per component
__device__ __forceinline__ V3 offset_ray_origin(V3 p, V3 err, V3 n, V3 w) {
    const float d = dot(vabs(n), err);
    V3 off = d * n;
    if (dot(n, w) < 0.f) off = -off;
    V3 po = p + off;
    if (off.x > 0.f) po.x = next_up(po.x); else if (off.x < 0.f) po.x = next_down(po.x);
    if (off.y > 0.f) po.y = next_up(po.y); else if (off.y < 0.f) po.y = next_down(po.y);
    if (off.z > 0.f) po.z = next_up(po.z); else if (off.z < 0.f) po.z = next_down(po.z);
    return po;
}

} // namespace b200pt
