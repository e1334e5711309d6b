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

/* ------------------------------------------------------------------------------------------
 * Host BVH for the oracle: the reference's traversal loop on the CPU.
 *
 * The reference's HLBVH cannot run on the host (every array is cudaMallocManaged and the build
 * is kernels), so the oracle builds its OWN binary tree -- Morton-ordered triangles (the codes
 * of util/math.h:183-208 widened to 21 bits per axis), split where the highest differing code
 * bit changes, leaves of <= 4 triangles, boxes refitted bottom-up -- and walks it with the
 * reference's loop: pop, slab test against the current best t (Bounds3::fast_intersect,
 * euclidean_space/bounds3.h:155-203, incl. the 1 + 2*gamma(3) widening evaluated in double),
 * leaf: every primitive against best_t in stored order, interior: push far then near
 * (accelerator/hlbvh.cu:459-516; any hit: :407-456).  Hit ids are argmin t under the reference's
 * watertight triangle test, whatever the tree; exact ties go to the last visited candidate, so a
 * checker comparing two trees must treat rays with |dt| of a few ulp as near-ties.
 *
 * Used (a) as the checker at full workload sizes, where brute force does not finish, and (b) as
 * bench.py's cpu_baseline: a real multi-threaded CPU traversal of the same rays.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    float lo[3], hi[3];
    int32_t left;  /* interior: index of the left child, right = left + 1; leaf: first primitive slot */
    int32_t count; /* 0 = interior, else number of primitives */
    int32_t axis;  /* interior: axis of the split bit (child order by ray direction sign) */
} obvh_node;

typedef struct {
    const float *P;
    const int32_t *idx;
    int64_t nt;
    int32_t *order; /* triangle ids in Morton order */
    obvh_node *nodes;
    int64_t n_nodes;
} obvh;

static inline uint64_t spread21(uint64_t x) {
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

typedef struct {
    uint64_t key;
    int32_t prim;
} obvh_key;

static void obvh_sort(obvh_key *a, obvh_key *tmp, int64_t n) { /* LSD radix sort, 8 x 8-bit digits */
    for (int pass = 0; pass < 8; pass++) {
        int64_t count[257];
        memset(count, 0, sizeof count);
        const int sh = 8 * pass;
        for (int64_t i = 0; i < n; i++) count[((a[i].key >> sh) & 0xff) + 1]++;
        for (int k = 0; k < 256; k++) count[k + 1] += count[k];
        for (int64_t i = 0; i < n; i++) tmp[count[(a[i].key >> sh) & 0xff]++] = a[i];
        obvh_key *t = a;
        a = tmp;
        tmp = t;
    }
}

static inline void obvh_tri_box(const obvh *b, int32_t prim, float *lo, float *hi) {
    for (int k = 0; k < 3; k++) {
        lo[k] = INFINITY;
        hi[k] = -INFINITY;
    }
    for (int v = 0; v < 3; v++) {
        const float *p = b->P + 3 * (int64_t)b->idx[3 * (int64_t)prim + v];
        for (int k = 0; k < 3; k++) {
            if (p[k] < lo[k]) lo[k] = p[k];
            if (p[k] > hi[k]) hi[k] = p[k];
        }
    }
}

/* builds the subtree over sorted slots [first, last] into node `me`; children are allocated as an adjacent pair */
static void obvh_build_range(obvh *b, const obvh_key *keys, int64_t root_first, int64_t root_last) {
    typedef struct {
        int64_t first, last;
        int32_t node;
        int bit;
    } item;
    int64_t cap = 256, sp = 0;
    item *stack = (item *)malloc(cap * sizeof(item));
    stack[sp++] = (item){root_first, root_last, 0, 62};
    while (sp > 0) {
        item it = stack[--sp];
        obvh_node *nd = &b->nodes[it.node];
        const int64_t n = it.last - it.first + 1;
        int64_t split = -1;
        int bit = it.bit;
        if (n > 4) {
            /* highest bit at or below `bit` in which the first and last key differ */
            const uint64_t diff = keys[it.first].key ^ keys[it.last].key;
            while (bit >= 0 && !((diff >> bit) & 1)) bit--;
            if (bit >= 0) { /* first slot whose key has the bit set (keys are sorted) */
                int64_t lo = it.first, hi = it.last;
                while (lo < hi) {
                    const int64_t mid = (lo + hi) / 2;
                    if ((keys[mid].key >> bit) & 1) hi = mid;
                    else lo = mid + 1;
                }
                split = lo;
            } else if (n > 16) { /* identical codes: halve */
                split = it.first + n / 2;
                bit = 0;
            }
        }
        if (split < 0) {
            nd->left = (int32_t)it.first;
            nd->count = (int32_t)n;
            nd->axis = 0;
            continue;
        }
        const int32_t l = (int32_t)b->n_nodes;
        b->n_nodes += 2;
        nd->left = l;
        nd->count = 0;
        nd->axis = bit % 3; /* x lowest in each 3-bit group */
        if (sp + 2 > cap) {
            cap *= 2;
            stack = (item *)realloc(stack, cap * sizeof(item));
        }
        stack[sp++] = (item){it.first, split - 1, l, bit - 1};
        stack[sp++] = (item){split, it.last, l + 1, bit - 1};
    }
    free(stack);
}

ORACLE_API void oracle_bvh_free(void *h) {
    obvh *b = (obvh *)h;
    if (!b) return;
    free(b->order);
    free(b->nodes);
    free(b);
}

ORACLE_API void *oracle_bvh_build(const float *P, const int32_t *idx, int64_t nt) {
    obvh *b = (obvh *)calloc(1, sizeof(obvh));
    b->P = P;
    b->idx = idx;
    b->nt = nt;
    if (nt <= 0) return b;
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    float *cen = (float *)malloc(sizeof(float) * 3 * nt);
    for (int64_t t = 0; t < nt; t++) {
        float lo[3], hi[3];
        obvh_tri_box(b, (int32_t)t, lo, hi);
        for (int k = 0; k < 3; k++) {
            const float c = 0.5f * (lo[k] + hi[k]);
            cen[3 * t + k] = c;
            if (c < clo[k]) clo[k] = c;
            if (c > chi[k]) chi[k] = c;
        }
    }
    obvh_key *keys = (obvh_key *)malloc(sizeof(obvh_key) * nt), *tmp = (obvh_key *)malloc(sizeof(obvh_key) * nt);
    for (int64_t t = 0; t < nt; t++) {
        uint64_t code = 0;
        for (int k = 0; k < 3; k++) {
            const double ext = (double)chi[k] - (double)clo[k];
            double f = ext > 0 ? ((double)cen[3 * t + k] - (double)clo[k]) / ext : 0.0;
            int64_t q = (int64_t)(f * 2097152.0);
            if (q > 2097151) q = 2097151;
            if (q < 0) q = 0;
            code |= spread21((uint64_t)q) << k;
        }
        keys[t].key = code;
        keys[t].prim = (int32_t)t;
    }
    free(cen);
    obvh_sort(keys, tmp, nt); /* 8 passes: the sorted data ends up back in `keys` */
    free(tmp);
    b->nodes = (obvh_node *)malloc(sizeof(obvh_node) * (size_t)(2 * nt + 1));
    b->n_nodes = 1;
    obvh_build_range(b, keys, 0, nt - 1);
    b->order = (int32_t *)malloc(sizeof(int32_t) * nt);
    for (int64_t t = 0; t < nt; t++) b->order[t] = keys[t].prim;
    free(keys);
    /* refit: children always have larger indices than their parent, so one backward sweep does it */
    for (int64_t i = b->n_nodes - 1; i >= 0; i--) {
        obvh_node *nd = &b->nodes[i];
        if (nd->count > 0) {
            for (int k = 0; k < 3; k++) {
                nd->lo[k] = INFINITY;
                nd->hi[k] = -INFINITY;
            }
            for (int32_t s = 0; s < nd->count; s++) {
                float lo[3], hi[3];
                obvh_tri_box(b, b->order[nd->left + s], lo, hi);
                for (int k = 0; k < 3; k++) {
                    if (lo[k] < nd->lo[k]) nd->lo[k] = lo[k];
                    if (hi[k] > nd->hi[k]) nd->hi[k] = hi[k];
                }
            }
        } else {
            const obvh_node *l = &b->nodes[nd->left], *r = &b->nodes[nd->left + 1];
            for (int k = 0; k < 3; k++) {
                nd->lo[k] = l->lo[k] < r->lo[k] ? l->lo[k] : r->lo[k];
                nd->hi[k] = l->hi[k] > r->hi[k] ? l->hi[k] : r->hi[k];
            }
        }
    }
    return b;
}

/* euclidean_space/bounds3.h:155-203 */
static inline int obvh_slab(const obvh_node *nd, const float *o, float ray_t_max, const float *inv, const int *neg) {
    const double widen = 1.0 + 2.0 * (double)gamma_f(3);
    const float *bmin = nd->lo, *bmax = nd->hi;
    float t_min = ((neg[0] ? bmax[0] : bmin[0]) - o[0]) * inv[0];
    float t_max = ((neg[0] ? bmin[0] : bmax[0]) - o[0]) * inv[0];
    float ty_min = ((neg[1] ? bmax[1] : bmin[1]) - o[1]) * inv[1];
    float ty_max = ((neg[1] ? bmin[1] : bmax[1]) - o[1]) * inv[1];
    t_max = (float)((double)t_max * widen);
    ty_max = (float)((double)ty_max * widen);
    if (t_min > ty_max || ty_min > t_max) return 0;
    if (ty_min > t_min) t_min = ty_min;
    if (ty_max < t_max) t_max = ty_max;
    float tz_min = ((neg[2] ? bmax[2] : bmin[2]) - o[2]) * inv[2];
    float tz_max = ((neg[2] ? bmin[2] : bmax[2]) - o[2]) * inv[2];
    tz_max = (float)((double)tz_max * widen);
    if (t_min > tz_max || tz_min > t_max) return 0;
    if (tz_min > t_min) t_min = tz_min;
    if (tz_max < t_max) t_max = tz_max;
    return t_min < ray_t_max && t_max > 0.0f;
}

typedef struct {
    const obvh *b;
    const float *ro, *rd, *tmax;
    int32_t *prim;
    float *t_out, *b12;
    int32_t *occluded;
} obvh_ctx;

static void obvh_ray(const obvh_ctx *c, int64_t r, int any) {
    const obvh *b = c->b;
    const float *o = c->ro + 3 * r, *d = c->rd + 3 * r;
    float best_t = c->tmax[r], bb1 = 0, bb2 = 0;
    int32_t best = -1;
    if (b->nt > 0) {
        /* accelerator/hlbvh.cu:468-474: 1.0 / d in double, rounded to float */
        const float inv[3] = {(float)(1.0 / (double)d[0]), (float)(1.0 / (double)d[1]), (float)(1.0 / (double)d[2])};
        const int neg[3] = {inv[0] < 0.0f, inv[1] < 0.0f, inv[2] < 0.0f};
        int32_t stack[192];
        int sp = 0;
        stack[sp++] = 0;
        while (sp > 0) {
            const obvh_node *nd = &b->nodes[stack[--sp]];
            if (!obvh_slab(nd, o, best_t, inv, neg)) continue;
            if (nd->count > 0) {
                for (int32_t s = 0; s < nd->count; s++) {
                    const int32_t k = b->order[nd->left + s];
                    float h[4];
                    if (!oracle_intersect_triangle(o, d, best_t, b->P + 3 * (int64_t)b->idx[3 * (int64_t)k],
                                                   b->P + 3 * (int64_t)b->idx[3 * (int64_t)k + 1],
                                                   b->P + 3 * (int64_t)b->idx[3 * (int64_t)k + 2], h))
                        continue;
                    best_t = h[3];
                    best = k;
                    bb1 = h[1];
                    bb2 = h[2];
                    if (any) {
                        sp = 0;
                        break;
                    }
                }
                continue;
            }
            if (sp + 2 > 192) continue; /* cannot happen: depth <= 63 + log2(16) */
            if (neg[nd->axis]) {
                stack[sp++] = nd->left;
                stack[sp++] = nd->left + 1;
            } else {
                stack[sp++] = nd->left + 1;
                stack[sp++] = nd->left;
            }
        }
    }
    if (any) {
        c->occluded[r] = best >= 0;
    } else {
        c->prim[r] = best;
        c->t_out[r] = best >= 0 ? best_t : INFINITY;
        c->b12[2 * r] = bb1;
        c->b12[2 * r + 1] = bb2;
    }
}
static void obvh_closest_body(int64_t r, void *vc) { obvh_ray((const obvh_ctx *)vc, r, 0); }
static void obvh_any_body(int64_t r, void *vc) { obvh_ray((const obvh_ctx *)vc, r, 1); }

ORACLE_API void oracle_bvh_closest_hit(void *h, const float *ro, const float *rd, const float *tmax, int64_t nrays,
                                       int32_t *prim, float *t_out, float *b12) {
    obvh_ctx c = {(const obvh *)h, ro, rd, tmax, prim, t_out, b12, NULL};
    parallel_for(nrays, 256, obvh_closest_body, &c);
}

ORACLE_API void oracle_bvh_any_hit(void *h, const float *ro, const float *rd, const float *tmax, int64_t nrays,
                                   int32_t *occluded) {
    obvh_ctx c = {(const obvh *)h, ro, rd, tmax, NULL, NULL, NULL, occluded};
    parallel_for(nrays, 256, obvh_any_body, &c);
}

ORACLE_API int64_t oracle_bvh_node_count(void *h) { return h ? ((const obvh *)h)->n_nodes : 0; }
