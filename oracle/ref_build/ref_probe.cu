// ref_probe.cu -- TEST INFRASTRUCTURE ONLY (oracle/; never on the product path).
//
// Host-only probe that calls the REFERENCE's own __host__ __device__ functions on the CPU and
// prints golden vectors as JSON (integers as decimal strings, floats as their uint32 bit
// patterns).  Runs in the build container (no GPU needed); compiled against the reference
// headers where they lie under /root/reference by oracle/ref_build/build_ref.py --only probe,
// which also writes tests/golden/ref_probe.json.
//
// Only integer / bit-exact paths are probed here (hashing, PCG32, permutation, the stratified and
// independent samplers, Morton encoding).  Floating-point paths whose GPU code differs from the
// host code by FMA contraction (the triangle test) are probed on the GPU by ref_trace.cu.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include <pbrt/euclidean_space/bounds3.h>
#include <pbrt/samplers/independent.h>
#include <pbrt/samplers/stratified.h>
#include <pbrt/util/hash.h>
#include <pbrt/util/math.h>
#include <pbrt/util/permutation.h>
#include <pbrt/util/rng.h>

static uint32_t fbits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}

// splitmix64: the probe's own deterministic input generator (not reference code)
static uint64_t sm_state = 0x243F6A8885A308D3ull;
static uint64_t sm_next() {
    uint64_t z = (sm_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

int main() {
    printf("{\n");

    // --- pbrt::hash(int), pbrt::hash(int,int), MixBits, hash_float(float,int)
    printf("\"hash\": [");
    for (int k = 0; k < 64; k++) {
        int a = int(sm_next() % 2000000) - (k % 7 == 0 ? 1000000 : 0);
        int b = int(sm_next() % 4096);
        float f = float(sm_next() >> 40) * 0x1p-24f;
        printf("%s{\"a\":%d,\"b\":%d,\"f\":%u,\"h1\":\"%llu\",\"h2\":\"%llu\",\"mix\":\"%llu\",\"hf\":%u}",
               k ? ",\n  " : "\n  ", a, b, fbits(f), (unsigned long long)pbrt::hash(a),
               (unsigned long long)pbrt::hash(a, b), (unsigned long long)MixBits(pbrt::hash(a)),
               fbits(pbrt::hash_float(f, b)));
    }
    printf("],\n");

    // --- permutation_element
    printf("\"perm\": [");
    const uint32_t lens[] = {1, 4, 16, 64, 100, 1024, 4096};
    for (int k = 0; k < 128; k++) {
        uint32_t l = lens[k % 7];
        uint32_t i = uint32_t(sm_next() % l);
        uint32_t p = uint32_t(sm_next());
        printf("%s[%u,%u,%u,%d]", k ? "," : "", i, l, p, permutation_element(i, l, p));
    }
    printf("],\n");

    // --- raw PCG32: set_sequence(hash(pixel)) + advance + uniform<uint32_t>/<float>
    printf("\"rng\": [");
    for (int k = 0; k < 16; k++) {
        int pixel = int(sm_next() % (1920 * 1080));
        long long adv = (long long)(sm_next() % 4096) * 65536ll + (long long)(sm_next() % 64);
        RNG rng;
        rng.set_sequence(pbrt::hash(pixel));
        rng.advance(adv);
        printf("%s{\"pixel\":%d,\"advance\":\"%lld\",\"u32\":[", k ? ",\n  " : "\n  ", pixel, adv);
        for (int j = 0; j < 4; j++) printf("%s%u", j ? "," : "", rng.uniform<uint32_t>());
        printf("],\"f32\":[");
        for (int j = 0; j < 4; j++) printf("%s%u", j ? "," : "", fbits(rng.uniform<float>()));
        printf("]}");
    }
    printf("],\n");

    // --- sampler streams in the wavefront integrator's consumption order:
    //     pattern 2,1 (camera), then per bounce 1,2 (NEE) 1,2 (BSDF)
    const int pattern[] = {2, 1, 1, 2, 1, 2, 1, 2, 1, 2, 1, 2, 1, 2};
    const int npat = sizeof(pattern) / sizeof(int);
    printf("\"pattern\": [");
    for (int j = 0; j < npat; j++) printf("%s%d", j ? "," : "", pattern[j]);
    printf("],\n\"sampler\": [");
    const int spps[] = {16, 4096, 1};
    bool first = true;
    for (int kind = 0; kind < 2; kind++) {
        for (int spp : spps) {
            for (int k = 0; k < 6; k++) {
                int pixel = int(sm_next() % (1080 * 1080));
                int sample = int(sm_next() % spp);
                printf("%s{\"kind\":%d,\"spp\":%d,\"pixel\":%d,\"sample\":%d,\"draws\":[",
                       first ? "\n  " : ",\n  ", kind, spp, pixel, sample);
                first = false;
                bool f2 = true;
                auto emit = [&](float v) {
                    printf("%s%u", f2 ? "" : ",", fbits(v));
                    f2 = false;
                };
                if (kind == 0) {
                    StratifiedSampler s(spp);
                    s.start_pixel_sample(pixel, sample, 0);
                    for (int j = 0; j < npat; j++) {
                        if (pattern[j] == 1) {
                            emit(s.get_1d());
                        } else {
                            auto p = s.get_2d();
                            emit(p.x);
                            emit(p.y);
                        }
                    }
                } else {
                    IndependentSampler s(spp);
                    s.start_pixel_sample(pixel, sample, 0);
                    for (int j = 0; j < npat; j++) {
                        if (pattern[j] == 1) {
                            emit(s.get_1d());
                        } else {
                            auto p = s.get_2d();
                            emit(p.x);
                            emit(p.y);
                        }
                    }
                }
                printf("]}");
            }
        }
    }
    printf("],\n");

    // --- Morton: Bounds3f::offset + encode_morton3 as in hlbvh_compute_morton_code
    printf("\"morton\": [");
    for (int k = 0; k < 64; k++) {
        float lo[3], hi[3], c[3];
        for (int a = 0; a < 3; a++) {
            lo[a] = float(sm_next() >> 40) * 0x1p-24f * 10.f - 5.f;
            hi[a] = lo[a] + float(sm_next() >> 40) * 0x1p-24f * 7.f + (k % 9 == 0 && a == 1 ? 0.f : 0.01f);
            float u = float(sm_next() >> 40) * 0x1p-24f;
            if (k % 5 == 0) u = (a == 0) ? 1.0f : 0.0f; // exercise the 1024 -> 1023 clamp
            c[a] = lo[a] + u * (hi[a] - lo[a]);
            if (c[a] > hi[a]) c[a] = hi[a];
        }
        Bounds3f b(Point3f(lo[0], lo[1], lo[2]), Point3f(hi[0], hi[1], hi[2]));
        auto off = b.offset(Point3f(c[0], c[1], c[2]));
        auto sc = off * 1024;
        uint32_t code = encode_morton3(uint32_t(sc.x), uint32_t(sc.y), uint32_t(sc.z));
        printf("%s{\"lo\":[%u,%u,%u],\"hi\":[%u,%u,%u],\"c\":[%u,%u,%u],\"code\":%u}", k ? ",\n  " : "\n  ",
               fbits(lo[0]), fbits(lo[1]), fbits(lo[2]), fbits(hi[0]), fbits(hi[1]), fbits(hi[2]),
               fbits(c[0]), fbits(c[1]), fbits(c[2]), code);
    }
    printf("]\n}\n");
    return 0;
}
