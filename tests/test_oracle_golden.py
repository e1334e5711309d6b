"""CPU: the oracle restatement (oracle/oracle.c) against golden vectors produced by the
REFERENCE's own __host__ __device__ functions (oracle/ref_build/ref_probe.cu ->
tests/golden/ref_probe.json).  Integer paths: bit-exact."""
import json
import os

import numpy as np
import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ref_probe.json")))


def bits(x):
    return np.asarray(x, dtype=np.float32).view(np.uint32)


def f32(u):
    return np.asarray(u, dtype=np.uint32).view(np.float32)


def test_hash_mixbits(oracle):
    L = oracle.lib()
    for g in GOLD["hash"]:
        assert L.oracle_hash_i(g["a"]) == int(g["h1"])
        assert L.oracle_hash_ii(g["a"], g["b"]) == int(g["h2"])
        assert L.oracle_mix_bits(int(g["h1"])) == int(g["mix"])
        hf = L.oracle_hash_float_fi(float(f32([g["f"]])[0]), g["b"])
        assert int(bits([hf])[0]) == g["hf"]


def test_permutation_element(oracle):
    L = oracle.lib()
    for i, l, p, want in GOLD["perm"]:
        assert L.oracle_permutation_element(i, l, p) == want


def test_sampler_streams(oracle):
    pattern = GOLD["pattern"]
    for g in GOLD["sampler"]:
        got = oracle.sampler_stream(g["kind"], g["spp"], [g["pixel"]], [g["sample"]], pattern)[0]
        want = list(g["draws"])
        if g["kind"] == 1:
            # IndependentSampler::get_2d is `Point2f(rng.uniform(), rng.uniform())` (samplers/independent.h:27-29):
            # argument evaluation order is unspecified; the host probe (g++) draws y first, the reference's
            # DEVICE code draws x first (checked in its sm_100 PTX).  The oracle and the product follow the
            # device order, so the host golden's 2-D pairs are swapped back here.
            k = 0
            for p in pattern:
                if p == 2:
                    want[k], want[k + 1] = want[k + 1], want[k]
                k += p
        assert bits(got).tolist() == want, g


def test_rng_via_independent_sampler(oracle):
    # raw PCG32 goldens: set_sequence(hash(pixel)) + advance(a): the independent sampler started at
    # (pixel, sample = a >> 16, dim = a & 0xffff) exposes exactly that stream
    for g in GOLD["rng"]:
        adv = int(g["advance"])
        sample, dim = adv >> 16, adv & 0xFFFF
        L = oracle.lib()
        import ctypes as C

        class S(C.Structure):
            _fields_ = [("kind", C.c_int), ("spp", C.c_int), ("sqrt_spp", C.c_int), ("state", C.c_uint64),
                        ("inc", C.c_uint64), ("pixel", C.c_int), ("sample", C.c_int), ("dim", C.c_int)]
        s = S()
        L.oracle_sampler_start(C.byref(s), 1, 1, g["pixel"], sample, dim)
        L.oracle_sampler_get_1d.restype = C.c_float
        # the golden consumed 4 u32 then 4 floats from the same stream: skip 4 draws, compare floats
        for _ in range(4):
            L.oracle_sampler_get_1d(C.byref(s))
        got = [L.oracle_sampler_get_1d(C.byref(s)) for _ in range(4)]
        assert bits(got).tolist() == g["f32"]


def test_morton(oracle):
    L = oracle.lib()
    import ctypes as C
    for g in GOLD["morton"]:
        c, lo, hi = f32(g["c"]), f32(g["lo"]), f32(g["hi"])
        fp = C.POINTER(C.c_float)
        code = L.oracle_morton_of_centroid(c.ctypes.data_as(fp), lo.ctypes.data_as(fp), hi.ctypes.data_as(fp))
        assert code == g["code"]


def test_triangle_known_answers(oracle):
    # analytic cases of the watertight test (shapes/triangle.cu:213-323)
    p0, p1, p2 = [0, 0, 1], [1, 0, 1], [0, 1, 1]
    hit, bt = oracle.intersect_triangle([0.25, 0.25, 0], [0, 0, 1], np.inf, p0, p1, p2)
    assert hit and bt[3] == 1.0 and np.allclose(bt[:3], [0.5, 0.25, 0.25])
    assert not oracle.intersect_triangle([0.25, 0.25, 0], [0, 0, -1], np.inf, p0, p1, p2)[0]   # behind
    assert not oracle.intersect_triangle([0.75, 0.75, 0], [0, 0, 1], np.inf, p0, p1, p2)[0]    # outside
    assert not oracle.intersect_triangle([0.25, 0.25, 0], [0, 0, 1], 0.5, p0, p1, p2)[0]       # t > t_max
    assert oracle.intersect_triangle([0.25, 0.25, 0], [0, 0, 1], 1.0, p0, p1, p2)[0]           # t == t_max accepted
    assert not oracle.intersect_triangle([0.25, 0.25, 0], [0, 0, 1], np.inf, p0, p1, p1)[0]    # degenerate
    # shared edge: watertight => exactly one side or both report, never neither
    q0, q1, q2 = [1, 0, 1], [1, 1, 1], [0, 1, 1]
    a = oracle.intersect_triangle([0.5, 0.5, 0], [0, 0, 1], np.inf, p0, p1, p2)[0]
    b = oracle.intersect_triangle([0.5, 0.5, 0], [0, 0, 1], np.inf, q0, q1, q2)[0]
    assert a or b


def test_bruteforce_semantics(oracle):
    # later primitive wins an exact tie (hlbvh.cu:495-501 with triangle.cu:284-290)
    P = np.array([[0, 0, 1], [1, 0, 1], [0, 1, 1], [0, 0, 1], [1, 0, 1], [0, 1, 1], [0, 0, 2], [1, 0, 2], [0, 1, 2]], np.float32)
    idx = np.arange(9, dtype=np.int32).reshape(3, 3)
    prim, t, b, t2 = oracle.closest_hit(P, idx, [[0.25, 0.25, 0]], [[0, 0, 1]], np.inf, want_t2=True)
    assert prim[0] == 1 and t[0] == 1.0 and t2[0] == 1.0
    assert oracle.any_hit(P, idx, [[0.25, 0.25, 0]], [[0, 0, 1]], 0.5)[0] == False  # noqa: E712
    assert oracle.any_hit(P, idx, [[0.25, 0.25, 0]], [[0, 0, 1]], 1.5)[0] == True  # noqa: E712


def test_host_bvh_equals_bruteforce(oracle, b200pt):
    """the oracle's host BVH (the reference's traversal loop over the oracle's own tree) gives the brute-force answers:
    ids and t bit-exact for closest hit, the same flags for any hit -- on a soup, on coincident triangles (exact ties are
    near ties by definition) and on an empty scene"""
    wl = b200pt.workloads
    P, idx = wl.triangle_soup(30000, seed=3)
    o, d, tm = wl.primary_rays(120, 80)
    prim, t, b, t2 = oracle.closest_hit(P, idx, o, d, tm, want_t2=True)
    bvh = oracle.HostBVH(P, idx)
    assert bvh.node_count() > 1000
    p2, t2b, b2 = bvh.closest_hit(o, d, tm)
    mism, near = oracle.compare_closest(P, idx, o, d, tm, p2, t2b, prim, t)
    assert mism == 0
    same = p2 == prim
    assert same.sum() >= len(prim) - near
    assert np.array_equal(t[same].view(np.uint32), t2b[same].view(np.uint32))
    assert np.array_equal(b[same], b2[same])
    bo, bd, bt = wl.bounce_rays(P, idx, o, d, prim, t, seed=7)
    pb, tb, _ = oracle.closest_hit(P, idx, bo, bd, bt)
    pb2, tb2, _ = bvh.closest_hit(bo, bd, bt)
    assert oracle.compare_closest(P, idx, bo, bd, bt, pb2, tb2, pb, tb)[0] == 0
    so, sd, st = wl.shadow_rays(P, idx, o, d, prim, t, seed=11)
    assert np.array_equal(oracle.any_hit(P, idx, so, sd, st), bvh.any_hit(so, sd, st))
    # a finite t_max is honoured exactly as by brute force (the cut sits right at the first hit: both sides of the
    # inclusive `tScaled > t_max * det` reject, shapes/triangle.cu:284-290, get exercised)
    hit = prim >= 0
    for tm2 in (t[hit], np.nextafter(t[hit], np.float32(0)), t[hit] * np.float32(1.5)):
        pa, ta, _ = oracle.closest_hit(P, idx, o[hit], d[hit], tm2)
        pb3, tb3, _ = bvh.closest_hit(o[hit], d[hit], tm2)
        assert oracle.compare_closest(P, idx, o[hit], d[hit], tm2, pb3, tb3, pa, ta)[0] == 0
    # coincident triangles: whichever is reported, the comparison calls it a near tie, not a mismatch
    Pc = np.array([[0, 0, 1], [1, 0, 1], [0, 1, 1]] * 2 + [[0, 0, 2], [1, 0, 2], [0, 1, 2]], np.float32)
    ic = np.arange(9, dtype=np.int32).reshape(3, 3)
    bc = oracle.HostBVH(Pc, ic)
    pc, tc, _ = bc.closest_hit([[0.25, 0.25, 0]], [[0, 0, 1]], np.inf)
    assert pc[0] in (0, 1) and tc[0] == 1.0
    assert oracle.compare_closest(Pc, ic, [[0.25, 0.25, 0]], [[0, 0, 1]], np.inf, np.array([0]), np.array([1.0], np.float32), np.array([1]),
                                  np.array([1.0], np.float32)) == (0, 1)
    assert oracle.compare_closest(Pc, ic, [[0.25, 0.25, 0]], [[0, 0, 1]], np.inf, np.array([2]), np.array([2.0], np.float32), np.array([1]),
                                  np.array([1.0], np.float32)) == (1, 0)
    # empty scene
    be = oracle.HostBVH(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32))
    pe, te, _ = be.closest_hit(o[:5], d[:5], np.inf)
    assert (pe == -1).all() and np.isinf(te).all()
