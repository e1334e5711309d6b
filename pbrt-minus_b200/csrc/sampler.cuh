// sampler.cuh -- the reference's sampler stack, bit-exact (integer work):
//   PCG32           util/rng.h:25-148
//   MixBits         util/rng.h:14-21
//   MurmurHash64A   util/hash.h:10-87  (pbrt::hash(int), hash(int,int), hash_float(float,int))
//   permutation     util/permutation.h:6-34
//   Stratified      samplers/stratified.h:8-70
//   Independent     samplers/independent.h:7-38
// The (pixel, sample) -> random stream map is a pure function, which is what lets the render
// shard by sample index across GPUs.  State is 24 B per path (rng state + inc are
// re-derivable, but kept: state 8, inc 8, dimension 4, pad) instead of the reference's 56.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace b200pt {

constexpr uint64_t PCG32_MULT = 0x5851f42d4c957f2dULL;

__host__ __device__ __forceinline__ uint64_t mix_bits(uint64_t v) {
    v ^= (v >> 31);
    v *= 0x7fb5d329728ea185ULL;
    v ^= (v >> 27);
    v *= 0x81dadef4bc2dd44dULL;
    v ^= (v >> 33);
    return v;
}

// MurmurHash64A over exactly 4 bytes (one int), seed 0
__host__ __device__ __forceinline__ uint64_t hash_i32(int32_t a) {
    const uint64_t m = 0xc6a4a7935bd1e995ULL;
    const int r = 47;
    uint64_t h = 0ull ^ (4ull * m);
    const uint32_t u = (uint32_t)a;
    // tail of 4 bytes: h ^= b3<<24 | b2<<16 | b1<<8 | b0; h *= m
    h ^= (uint64_t)u;
    h *= m;
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return h;
}

// MurmurHash64A over exactly 8 bytes (two 32-bit words packed back to back), seed 0
__host__ __device__ __forceinline__ uint64_t hash_u32x2(uint32_t lo, uint32_t hi) {
    const uint64_t m = 0xc6a4a7935bd1e995ULL;
    const int r = 47;
    uint64_t h = 0ull ^ (8ull * m);
    uint64_t k = (uint64_t)lo | ((uint64_t)hi << 32);
    k *= m;
    k ^= k >> r;
    k *= m;
    h ^= k;
    h *= m;
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return h;
}

__host__ __device__ __forceinline__ uint64_t hash_i32x2(int32_t a, int32_t b) { return hash_u32x2((uint32_t)a, (uint32_t)b); }

// util/hash.h:82-85 hash_float(float, int): uint32(hash) * 2^-32  (distribution/alias_table.cu:70-87)
__device__ __forceinline__ float hash_float_fi(float a, int32_t b) {
    return __fmul_rn((float)(uint32_t)hash_u32x2(__float_as_uint(a), (uint32_t)b), 0x1p-32f);
}

__host__ __device__ __forceinline__ int32_t permutation_element(uint32_t i, uint32_t l, uint32_t p) {
    uint32_t w = l - 1;
    w |= w >> 1;
    w |= w >> 2;
    w |= w >> 4;
    w |= w >> 8;
    w |= w >> 16;
    do {
        i ^= p;
        i *= 0xe170893d;
        i ^= p >> 16;
        i ^= (i & w) >> 4;
        i ^= p >> 8;
        i *= 0x0929eb3f;
        i ^= p >> 23;
        i ^= (i & w) >> 1;
        i *= 1 | p >> 27;
        i *= 0x6935fa69;
        i ^= (i & w) >> 11;
        i *= 0x74dcb303;
        i ^= (i & w) >> 2;
        i *= 0x9e501cc3;
        i ^= (i & w) >> 2;
        i *= 0xc860a3df;
        i &= w;
        i ^= i >> 5;
    } while (i >= l);
    return (int32_t)((i + p) % l);
}

struct Sampler {
    uint64_t state, inc;
    int32_t pixel_idx, sample_idx, dimension;
    int32_t kind, spp, sqrt_spp; // kind: 0 stratified, 1 independent

    __device__ __forceinline__ uint32_t next_u32() {
        const uint64_t old = state;
        state = old * PCG32_MULT + inc;
        const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    // util/rng.h:123-126: min(OneMinusEpsilon, u32 * 2^-32)
    __device__ __forceinline__ float next_f32() {
        const float v = __fmul_rn((float)next_u32(), 0x1p-32f);
        return fminf(v, 0x1.fffffep-1f);
    }
    __device__ __forceinline__ void advance(uint64_t delta) {
        uint64_t cur_mult = PCG32_MULT, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
        while (delta > 0) {
            if (delta & 1) {
                acc_mult *= cur_mult;
                acc_plus = acc_plus * cur_mult + cur_plus;
            }
            cur_plus = (cur_mult + 1) * cur_plus;
            cur_mult *= cur_mult;
            delta >>= 1;
        }
        state = acc_mult * state + acc_plus;
    }
    // samplers/stratified.h:24-31, samplers/independent.h:13-16
    __device__ __forceinline__ void start(int _kind, int _spp, int _sqrt_spp, int pixel, int sample, int dim) {
        kind = _kind;
        spp = _spp;
        sqrt_spp = _sqrt_spp;
        pixel_idx = pixel;
        sample_idx = sample;
        dimension = dim;
        const uint64_t seq = hash_i32(pixel);
        state = 0u;
        inc = (seq << 1u) | 1u;
        next_u32();
        state += mix_bits(seq);
        next_u32();
        advance((uint64_t)sample * 65536ull + (uint64_t)dim);
    }
    // samplers/stratified.h:33-43
    __device__ __forceinline__ float get_1d() {
        if (kind == 1) return next_f32();
        const uint64_t h = hash_i32x2(pixel_idx, dimension);
        const int stratum = permutation_element((uint32_t)sample_idx, (uint32_t)spp, (uint32_t)h);
        dimension += 1;
        const float delta = next_f32();
        return __fdiv_rn(__fadd_rn((float)stratum, delta), (float)spp);
    }
    // samplers/stratified.h:45-59
    __device__ __forceinline__ float2 get_2d() {
        if (kind == 1) {
            const float a = next_f32();
            const float b = next_f32();
            return make_float2(a, b);
        }
        const uint64_t h = hash_i32x2(pixel_idx, dimension);
        const int stratum = permutation_element((uint32_t)sample_idx, (uint32_t)spp, (uint32_t)h);
        dimension += 2;
        const int x = stratum % sqrt_spp, y = stratum / sqrt_spp;
        const float dx = next_f32();
        const float dy = next_f32();
        return make_float2(__fdiv_rn(__fadd_rn((float)x, dx), (float)sqrt_spp), __fdiv_rn(__fadd_rn((float)y, dy), (float)sqrt_spp));
    }
};

} // namespace b200pt
