/*
 * oracle.c -- CPU restatement of the reference's wavefront-path hot path (plain C).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is on the product path: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the checker.  The product (libb200pt.so) has no CPU fallback.
 *
 * PARITY PIN STATUS: the reference (w3ntao/pbrt-minus) ships NO tests, golden vectors or
 * known-answer fixtures ("parity unpinned" by the reference itself).  This
 * restatement is pinned instead against outputs of the reference itself:
 *   - integer paths (PCG32, MurmurHash64A, MixBits, permutation_element, stratified sampler,
 *     Morton encode): tests/golden/ref_probe_*.bin, produced in the build container by
 *     oracle/ref_build/ref_probe.cu, which calls the reference's own __host__ __device__
 *     functions on the CPU (script: oracle/ref_build/build_ref.py --only probe).
 *   - floating-point triangle test and HLBVH hit ids: tests/golden/ref_trace_*.bin, produced on
 *     the B200 box by oracle/ref_build/ref_trace.cu linked against the reference's objects.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -pthread).  -ffp-contract=off matters:
 * every fused multiply-add below is explicit (fmaf) and sits exactly where the reference's
 * sm_100 PTX has fma.rn.f32; nothing else may be contracted.
 *
 * Citations are to /root/reference/src/pbrt/.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * Minimal pthread parallel-for (the image's default $CC has no libgomp spec, so no OpenMP).
 * ------------------------------------------------------------------------------------------ */
#include <pthread.h>
#include <unistd.h>
typedef void (*pf_body)(int64_t i, void *ctx);
typedef struct {
    pf_body body;
    void *ctx;
    int64_t n, chunk;
    volatile int64_t *next;
} pf_task;
static void *pf_worker(void *arg) {
    pf_task *t = (pf_task *)arg;
    for (;;) {
        int64_t b = __atomic_fetch_add(t->next, t->chunk, __ATOMIC_RELAXED);
        if (b >= t->n) break;
        int64_t e = b + t->chunk < t->n ? b + t->chunk : t->n;
        for (int64_t i = b; i < e; i++) t->body(i, t->ctx);
    }
    return NULL;
}
static int g_threads = 0;
ORACLE_API void oracle_set_threads(int n) { g_threads = n; }
ORACLE_API int oracle_get_threads(void) {
    if (g_threads > 0) return g_threads;
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}
static void parallel_for(int64_t n, int64_t chunk, pf_body body, void *ctx) {
    int nt = oracle_get_threads();
    if (nt > 256) nt = 256;
    volatile int64_t next = 0;
    pf_task t = {body, ctx, n, chunk, &next};
    if (nt <= 1 || n <= chunk) {
        pf_worker(&t);
        return;
    }
    pthread_t th[256];
    int started = 0;
    for (int i = 0; i < nt - 1; i++)
        if (pthread_create(&th[started], NULL, pf_worker, &t) == 0) started++;
    pf_worker(&t);
    for (int i = 0; i < started; i++) pthread_join(th[i], NULL);
}

/* ------------------------------------------------------------------------------------------
 * util/math.h:81-84  gamma(n) = n*eps/(1-n*eps), eps = 2^-24 (MachineEpsilon = epsilon*0.5)
 * The values are the ones nvcc folded into the reference PTX (0f34000001, 0f34400002, 0f34A00003).
 * ------------------------------------------------------------------------------------------ */
static inline float gamma_f(int n) {
    const float eps = 0x1p-24f;
    return ((float)n * eps) / (1.0f - (float)n * eps);
}

/* util/math.h:167-173 difference_of_products: a*b - c*d with one compensation step. */
static inline float dop(float a, float b, float c, float d) {
    float cd = c * d;
    float diff = fmaf(a, b, -cd);
    float err = fmaf(-c, d, cd);
    return diff + err;
}

static inline float max3f(float a, float b, float c) {
    float m = (a < b) ? b : a;
    return (m < c) ? c : m;
}

/*
 * shapes/triangle.cu:213-323  Triangle::intersect_triangle (PBRT-v4 watertight test).
 * Operation order and fma placement follow the reference's sm_100 PTX for this function
 * (nvcc 12.9, default -fmad=true): see DESIGN.md "triangle arithmetic".
 * Returns 1 and fills out[4] = {b0, b1, b2, t} on a hit.
 */
ORACLE_API int oracle_intersect_triangle(const float o[3], const float dir[3], float t_max,
                                         const float p0[3], const float p1[3], const float p2[3],
                                         float out[4]) {
    /* :218 degenerate test: (p2 - p0).cross(p1 - p0).squared_length() == 0 */
    float ax = p2[0] - p0[0], ay = p2[1] - p0[1], az = p2[2] - p0[2];
    float bx = p1[0] - p0[0], by = p1[1] - p0[1], bz = p1[2] - p0[2];
    float cx = dop(ay, bz, az, by);
    float cy = dop(az, bx, ax, bz);
    float cz = dop(ax, by, ay, bx);
    float sq = fmaf(cz, cz, fmaf(cx, cx, cy * cy));
    if (sq == 0.0f) return 0;

    /* :224-226 translate */
    float q0[3] = {p0[0] - o[0], p0[1] - o[1], p0[2] - o[2]};
    float q1[3] = {p1[0] - o[0], p1[1] - o[1], p1[2] - o[2]};
    float q2[3] = {p2[0] - o[0], p2[1] - o[1], p2[2] - o[2]};

    /* :229-238 permute; euclidean_space/vector3.h:103-105 max_component_index on |d| */
    float adx = fabsf(dir[0]), ady = fabsf(dir[1]), adz = fabsf(dir[2]);
    int kz = (adx > ady) ? ((adx > adz) ? 0 : 2) : ((ady > adz) ? 1 : 2);
    int kx = (kz + 1) % 3, ky = (kz + 2) % 3;
    float dx = dir[kx], dy = dir[ky], dz = dir[kz];
    float p0x = q0[kx], p0y = q0[ky], p0z = q0[kz];
    float p1x = q1[kx], p1y = q1[ky], p1z = q1[kz];
    float p2x = q2[kx], p2y = q2[ky], p2z = q2[kz];

    /* :241-249 shear (div.rn, rcp.rn, then fma) */
    float Sx = -dx / dz;
    float Sy = -dy / dz;
    float Sz = 1.0f / dz;
    p0x = fmaf(Sx, p0z, p0x);
    p0y = fmaf(Sy, p0z, p0y);
    p1x = fmaf(Sx, p1z, p1x);
    p1y = fmaf(Sy, p1z, p1y);
    p2x = fmaf(Sx, p2z, p2x);
    p2y = fmaf(Sy, p2z, p2y);

    /* :252-254 edge functions */
    float e0 = dop(p1x, p2y, p1y, p2x);
    float e1 = dop(p2x, p0y, p2y, p0x);
    float e2 = dop(p0x, p1y, p0y, p1x);

    /* :257-267 double-precision fallback on an exact zero (products of floats are exact in double) */
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double p2txp1ty = (double)p2x * (double)p1y;
        double p2typ1tx = (double)p2y * (double)p1x;
        e0 = (float)(p2typ1tx - p2txp1ty);
        double p0txp2ty = (double)p0x * (double)p2y;
        double p0typ2tx = (double)p0y * (double)p2x;
        e1 = (float)(p0typ2tx - p0txp2ty);
        double p1txp0ty = (double)p1x * (double)p0y;
        double p1typ0tx = (double)p1y * (double)p0x;
        e2 = (float)(p1typ0tx - p1txp0ty);
    }

    /* :270-277 sign and determinant tests */
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return 0;
    float det = (e0 + e1) + e2;
    if (det == 0) return 0;

    /* :280-290 scaled hit distance; note the PTX order: z1*e1 first, then fma z0*e0, then fma z2*e2 */
    p0z = Sz * p0z;
    p1z = Sz * p1z;
    p2z = Sz * p2z;
    float tScaled = fmaf(p2z, e2, fmaf(p0z, e0, p1z * e1));
    if (det < 0 && (tScaled >= 0 || tScaled < t_max * det)) return 0;
    if (det > 0 && (tScaled <= 0 || tScaled > t_max * det)) return 0;

    /* :293-295 barycentrics */
    float invDet = 1.0f / det;
    float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    float t = tScaled * invDet;

    /* :299-319 conservative t > delta_t test */
    const float g2 = gamma_f(2), g3 = gamma_f(3), g5 = gamma_f(5);
    float maxZt = max3f(fabsf(p0z), fabsf(p1z), fabsf(p2z));
    float deltaZ = maxZt * g3;
    float maxXt = max3f(fabsf(p0x), fabsf(p1x), fabsf(p2x));
    float maxYt = max3f(fabsf(p0y), fabsf(p1y), fabsf(p2y));
    float deltaX = (maxZt + maxXt) * g5;
    float deltaY = (maxZt + maxYt) * g5;
    float s = fmaf(deltaX, maxYt, fmaf(maxXt * g2, maxYt, maxXt * deltaY));
    float deltaE = s + s;
    float maxE = max3f(fabsf(e0), fabsf(e1), fabsf(e2));
    float u = fmaf(deltaZ, maxE, fmaf(maxZt, deltaE, maxZt * (maxE * g3)));
    float deltaT = fabsf(invDet) * (u * 3.0f);
    if (t <= deltaT) return 0;

    out[0] = b0;
    out[1] = b1;
    out[2] = b2;
    out[3] = t;
    return 1;
}

/*
 * Brute-force closest hit with the semantics of accelerator/hlbvh.cu:489-502: candidates are
 * visited in primitive order, each tested with t_max = best t so far (so t == best_t passes and
 * the last visited wins an exact tie).  Also records, per ray, the smallest and second smallest
 * t over ALL triangles tested against the caller's t_max, so a harness can flag near-ties (rays
 * whose winner legitimately depends on visitation order, SURVEY.md section 7 hard part 2).
 *
 * P: nv*3 floats; idx: nt*3 ints; rays AoS xyz.  Outputs: prim[n] (-1 = miss), t[n], b12[2n],
 * t2[n] (runner-up t among other primitives; +inf if none).  t2 may be NULL.
 */
typedef struct {
    const float *P;
    const int32_t *idx;
    int64_t nt;
    const float *ro, *rd, *tmax;
    int32_t *prim;
    float *t_out, *b12, *t2_out;
    int32_t *occluded;
    const int32_t *prim_in;
    int32_t *ok;
    float *bt;
} hit_ctx;

static void closest_body(int64_t r, void *vc) {
    hit_ctx *c = (hit_ctx *)vc;
    const float *o = c->ro + 3 * r, *d = c->rd + 3 * r;
    float best_t = c->tmax[r];
    int32_t best = -1;
    float bb1 = 0, bb2 = 0;
    float m1 = INFINITY, m2 = INFINITY; /* two smallest independent t's */
    for (int64_t k = 0; k < c->nt; k++) {
        const float *a = c->P + 3 * (int64_t)c->idx[3 * k + 0];
        const float *b = c->P + 3 * (int64_t)c->idx[3 * k + 1];
        const float *cc = c->P + 3 * (int64_t)c->idx[3 * k + 2];
        float h[4];
        if (c->t2_out) {
            if (!oracle_intersect_triangle(o, d, c->tmax[r], a, b, cc, h))
                continue; /* fails against the loose bound => fails against best_t too */
            if (h[3] < m1) {
                m2 = m1;
                m1 = h[3];
            } else if (h[3] < m2) {
                m2 = h[3];
            }
        }
        if (oracle_intersect_triangle(o, d, best_t, a, b, cc, h)) {
            best_t = h[3];
            best = (int32_t)k;
            bb1 = h[1];
            bb2 = h[2];
        }
    }
    c->prim[r] = best;
    c->t_out[r] = (best >= 0) ? best_t : INFINITY;
    c->b12[2 * r] = bb1;
    c->b12[2 * r + 1] = bb2;
    if (c->t2_out) c->t2_out[r] = m2;
}

ORACLE_API void oracle_closest_hit(const float *P, const int32_t *idx, int64_t nt,
                                   const float *ro, const float *rd, const float *tmax,
                                   int64_t nrays, int32_t *prim, float *t_out, float *b12,
                                   float *t2_out) {
    hit_ctx c = {P, idx, nt, ro, rd, tmax, prim, t_out, b12, t2_out, NULL, NULL, NULL, NULL};
    parallel_for(nrays, 4, closest_body, &c);
}

/* Any-hit with the semantics of accelerator/hlbvh.cu:407-456 (true iff some primitive passes
 * intersect(ray, t_max)).  Triangle::fast_intersect == intersect_triangle(...).has_value()
 * (shapes/triangle.h). */
static void any_body(int64_t r, void *vc) {
    hit_ctx *c = (hit_ctx *)vc;
    int32_t hit = 0;
    float h[4];
    for (int64_t k = 0; k < c->nt && !hit; k++) {
        hit = oracle_intersect_triangle(c->ro + 3 * r, c->rd + 3 * r, c->tmax[r],
                                        c->P + 3 * (int64_t)c->idx[3 * k + 0],
                                        c->P + 3 * (int64_t)c->idx[3 * k + 1],
                                        c->P + 3 * (int64_t)c->idx[3 * k + 2], h);
    }
    c->occluded[r] = hit;
}

ORACLE_API void oracle_any_hit(const float *P, const int32_t *idx, int64_t nt, const float *ro,
                               const float *rd, const float *tmax, int64_t nrays,
                               int32_t *occluded) {
    hit_ctx c = {P, idx, nt, ro, rd, tmax, NULL, NULL, NULL, NULL, occluded, NULL, NULL, NULL};
    parallel_for(nrays, 4, any_body, &c);
}

/* Test one ray against one listed primitive each (used to re-verify full-size GPU results:
 * "the primitive you returned really is hit at the t you returned"). */
static void pairs_body(int64_t r, void *vc) {
    hit_ctx *c = (hit_ctx *)vc;
    float h[4] = {0, 0, 0, 0};
    c->ok[r] = 0;
    if (c->prim_in[r] >= 0) {
        int64_t k = c->prim_in[r];
        c->ok[r] = oracle_intersect_triangle(c->ro + 3 * r, c->rd + 3 * r, c->tmax[r],
                                             c->P + 3 * (int64_t)c->idx[3 * k + 0],
                                             c->P + 3 * (int64_t)c->idx[3 * k + 1],
                                             c->P + 3 * (int64_t)c->idx[3 * k + 2], h);
    }
    memcpy(c->bt + 4 * r, h, sizeof h);
}

ORACLE_API void oracle_intersect_pairs(const float *P, const int32_t *idx, const int32_t *prim,
                                       const float *ro, const float *rd, const float *tmax,
                                       int64_t nrays, int32_t *ok, float *bt) {
    hit_ctx c = {P, idx, 0, ro, rd, tmax, NULL, NULL, NULL, NULL, NULL, prim, ok, bt};
    parallel_for(nrays, 1024, pairs_body, &c);
}

/* ------------------------------------------------------------------------------------------
 * util/math.h:183-208 left_shift3 / encode_morton3 (10 bits per axis, x lowest; 1024 -> 1023)
 * accelerator/hlbvh.cu:128-142 + euclidean_space/bounds3.h:106-122 (offset in the centroid box)
 * ------------------------------------------------------------------------------------------ */
static inline uint32_t left_shift3(uint32_t x) {
    if (x == (1u << 10)) --x;
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

ORACLE_API uint32_t oracle_encode_morton3(uint32_t x, uint32_t y, uint32_t z) {
    return (left_shift3(z) << 2) | (left_shift3(y) << 1) | left_shift3(x);
}

ORACLE_API uint32_t oracle_morton_of_centroid(const float c[3], const float bmin[3],
                                              const float bmax[3]) {
    uint32_t q[3];
    for (int a = 0; a < 3; a++) {
        float o = c[a] - bmin[a];
        if (bmax[a] > bmin[a]) o /= (bmax[a] - bmin[a]);
        q[a] = (uint32_t)(o * 1024.0f);
    }
    return oracle_encode_morton3(q[0], q[1], q[2]);
}

/* ==========================================================================================
 * Sampler stack (integer, bit-exact): util/rng.h:14-148, util/hash.h:10-87,
 * util/permutation.h:6-34, samplers/stratified.h:8-70, samplers/independent.h:7-38
 * ========================================================================================== */
#define PCG32_MULT 0x5851f42d4c957f2dULL

/* util/rng.h:14-21 */
ORACLE_API uint64_t oracle_mix_bits(uint64_t v) {
    v ^= (v >> 31);
    v *= 0x7fb5d329728ea185ULL;
    v ^= (v >> 27);
    v *= 0x81dadef4bc2dd44dULL;
    v ^= (v >> 33);
    return v;
}

/* util/hash.h:10-56 MurmurHash64A */
ORACLE_API uint64_t oracle_murmur64a(const unsigned char *key, size_t len, uint64_t seed) {
    const uint64_t m = 0xc6a4a7935bd1e995ULL;
    const int r = 47;
    uint64_t h = seed ^ (len * m);
    const unsigned char *end = key + 8 * (len / 8);
    while (key != end) {
        uint64_t k;
        memcpy(&k, key, 8);
        key += 8;
        k *= m;
        k ^= k >> r;
        k *= m;
        h ^= k;
        h *= m;
    }
    switch (len & 7) {
    case 7: h ^= (uint64_t)key[6] << 48; /* fallthrough */
    case 6: h ^= (uint64_t)key[5] << 40; /* fallthrough */
    case 5: h ^= (uint64_t)key[4] << 32; /* fallthrough */
    case 4: h ^= (uint64_t)key[3] << 24; /* fallthrough */
    case 3: h ^= (uint64_t)key[2] << 16; /* fallthrough */
    case 2: h ^= (uint64_t)key[1] << 8;  /* fallthrough */
    case 1: h ^= (uint64_t)key[0]; h *= m;
    }
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return h;
}

/* util/hash.h:73-80 pbrt::hash(int) and pbrt::hash(int,int): args packed back to back, seed 0 */
ORACLE_API uint64_t oracle_hash_i(int32_t a) { return oracle_murmur64a((const unsigned char *)&a, 4, 0); }
ORACLE_API uint64_t oracle_hash_ii(int32_t a, int32_t b) {
    int32_t buf[2] = {a, b};
    return oracle_murmur64a((const unsigned char *)buf, 8, 0);
}
/* util/hash.h:82-85 hash_float(float, int) as used by distribution/alias_table.cu:70-87 */
ORACLE_API float oracle_hash_float_fi(float a, int32_t b) {
    unsigned char buf[8];
    memcpy(buf, &a, 4);
    memcpy(buf + 4, &b, 4);
    return (float)(uint32_t)oracle_murmur64a(buf, 8, 0) * 0x1p-32f;
}

/* util/permutation.h:6-34 */
ORACLE_API int32_t oracle_permutation_element(uint32_t i, uint32_t l, uint32_t p) {
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

typedef struct {
    uint64_t state, inc;
} oracle_rng;

/* util/rng.h:76-83 */
static inline uint32_t rng_u32(oracle_rng *g) {
    uint64_t old = g->state;
    g->state = old * PCG32_MULT + g->inc;
    uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
}
/* util/rng.h:115-121 */
static inline void rng_set_sequence(oracle_rng *g, uint64_t seq, uint64_t offset) {
    g->state = 0u;
    g->inc = (seq << 1u) | 1u;
    rng_u32(g);
    g->state += offset;
    rng_u32(g);
}
/* util/rng.h:133-147 */
static inline void rng_advance(oracle_rng *g, int64_t idelta) {
    uint64_t curMult = PCG32_MULT, curPlus = g->inc, accMult = 1u;
    uint64_t accPlus = 0u, delta = (uint64_t)idelta;
    while (delta > 0) {
        if (delta & 1) {
            accMult *= curMult;
            accPlus = accPlus * curMult + curPlus;
        }
        curPlus = (curMult + 1) * curPlus;
        curMult *= curMult;
        delta /= 2;
    }
    g->state = accMult * g->state + accPlus;
}
/* util/rng.h:123-126: min(OneMinusEpsilon, u32 * 2^-32) */
static inline float rng_f32(oracle_rng *g) {
    float v = (float)rng_u32(g) * 0x1p-32f;
    const float one_minus_eps = 0x1.fffffep-1f;
    return v < one_minus_eps ? v : one_minus_eps;
}

typedef struct {
    int kind; /* 0 = stratified, 1 = independent */
    int spp, sqrt_spp;
    oracle_rng rng;
    int pixel_idx, sample_idx, dimension;
} oracle_sampler;

/* samplers/stratified.h:24-31, samplers/independent.h:13-16 */
ORACLE_API void oracle_sampler_start(oracle_sampler *s, int kind, int spp, int pixel_idx,
                                     int sample_idx, int dimension) {
    s->kind = kind;
    s->spp = spp;
    s->sqrt_spp = (int)sqrt((double)spp);
    s->pixel_idx = pixel_idx;
    s->sample_idx = sample_idx;
    s->dimension = dimension;
    uint64_t seq = oracle_hash_i(pixel_idx);
    rng_set_sequence(&s->rng, seq, oracle_mix_bits(seq));
    rng_advance(&s->rng, (int64_t)((uint64_t)sample_idx * 65536ull + (uint64_t)dimension));
}

/* samplers/stratified.h:33-43 */
ORACLE_API float oracle_sampler_get_1d(oracle_sampler *s) {
    if (s->kind == 1) return rng_f32(&s->rng);
    uint64_t h = oracle_hash_ii(s->pixel_idx, s->dimension);
    int stratum = oracle_permutation_element((uint32_t)s->sample_idx, (uint32_t)s->spp, (uint32_t)h);
    s->dimension += 1;
    float delta = rng_f32(&s->rng);
    return ((float)stratum + delta) / (float)s->spp;
}

/* samplers/stratified.h:45-59 */
ORACLE_API void oracle_sampler_get_2d(oracle_sampler *s, float out[2]) {
    if (s->kind == 1) {
        out[0] = rng_f32(&s->rng);
        out[1] = rng_f32(&s->rng);
        return;
    }
    uint64_t h = oracle_hash_ii(s->pixel_idx, s->dimension);
    int stratum = oracle_permutation_element((uint32_t)s->sample_idx, (uint32_t)s->spp, (uint32_t)h);
    s->dimension += 2;
    int x = stratum % s->sqrt_spp, y = stratum / s->sqrt_spp;
    float dx = rng_f32(&s->rng);
    float dy = rng_f32(&s->rng);
    out[0] = ((float)x + dx) / (float)s->sqrt_spp;
    out[1] = ((float)y + dy) / (float)s->sqrt_spp;
}

/* Flat driver for tests: for each (pixel, sample) pair emit `ndraw` draws following pattern[]
 * (1 = get_1d, 2 = get_2d).  out is packed floats in draw order. */
ORACLE_API void oracle_sampler_stream(int kind, int spp, const int32_t *pixel, const int32_t *sample,
                                      int npairs, const int32_t *pattern, int ndraw, float *out) {
    int per = 0;
    for (int i = 0; i < ndraw; i++) per += pattern[i];
    for (int k = 0; k < npairs; k++) {
        oracle_sampler s;
        oracle_sampler_start(&s, kind, spp, pixel[k], sample[k], 0);
        float *o = out + (size_t)k * per;
        for (int i = 0; i < ndraw; i++) {
            if (pattern[i] == 1) {
                *o++ = oracle_sampler_get_1d(&s);
            } else {
                oracle_sampler_get_2d(&s, o);
                o += 2;
            }
        }
    }
}
