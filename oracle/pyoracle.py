"""ctypes loader of the CPU oracle (oracle/oracle.c).  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_build", "liboracle.so")


def build():
    src = os.path.join(HERE, "oracle.c")
    if (not os.path.exists(SO)) or os.path.getmtime(SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-s"], check=True)
    return SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(SO)
        fp, ip, i64 = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_int64
        L.oracle_intersect_triangle.argtypes = [fp, fp, C.c_float, fp, fp, fp, fp]
        L.oracle_closest_hit.argtypes = [fp, ip, i64, fp, fp, fp, i64, ip, fp, fp, fp]
        L.oracle_closest_hit.restype = None
        L.oracle_any_hit.argtypes = [fp, ip, i64, fp, fp, fp, i64, ip]
        L.oracle_any_hit.restype = None
        L.oracle_intersect_pairs.argtypes = [fp, ip, ip, fp, fp, fp, i64, ip, fp]
        L.oracle_intersect_pairs.restype = None
        L.oracle_encode_morton3.argtypes = [C.c_uint32] * 3
        L.oracle_encode_morton3.restype = C.c_uint32
        L.oracle_morton_of_centroid.argtypes = [fp, fp, fp]
        L.oracle_morton_of_centroid.restype = C.c_uint32
        L.oracle_mix_bits.argtypes = [C.c_uint64]
        L.oracle_mix_bits.restype = C.c_uint64
        L.oracle_hash_i.argtypes = [C.c_int32]
        L.oracle_hash_i.restype = C.c_uint64
        L.oracle_hash_ii.argtypes = [C.c_int32, C.c_int32]
        L.oracle_hash_ii.restype = C.c_uint64
        L.oracle_hash_float_fi.argtypes = [C.c_float, C.c_int32]
        L.oracle_hash_float_fi.restype = C.c_float
        L.oracle_permutation_element.argtypes = [C.c_uint32] * 3
        L.oracle_permutation_element.restype = C.c_int32
        L.oracle_sampler_stream.argtypes = [C.c_int, C.c_int, ip, ip, C.c_int, ip, C.c_int, fp]
        L.oracle_sampler_stream.restype = None
        L.oracle_bvh_build.argtypes = [fp, ip, i64]
        L.oracle_bvh_build.restype = C.c_void_p
        L.oracle_bvh_free.argtypes = [C.c_void_p]
        L.oracle_bvh_free.restype = None
        L.oracle_bvh_closest_hit.argtypes = [C.c_void_p, fp, fp, fp, i64, ip, fp, fp]
        L.oracle_bvh_closest_hit.restype = None
        L.oracle_bvh_any_hit.argtypes = [C.c_void_p, fp, fp, fp, i64, ip]
        L.oracle_bvh_any_hit.restype = None
        L.oracle_bvh_node_count.argtypes = [C.c_void_p]
        L.oracle_bvh_node_count.restype = i64
        L.oracle_set_threads.argtypes = [C.c_int]
        L.oracle_get_threads.restype = C.c_int
        _lib = L
    return _lib


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def closest_hit(P, idx, o, d, tmax, want_t2=False):
    P, idx, o, d = _f(P), _i(idx), _f(o), _f(d)
    n = o.reshape(-1, 3).shape[0]
    tmax = _f(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
    prim = np.empty(n, np.int32)
    t = np.empty(n, np.float32)
    b = np.empty((n, 2), np.float32)
    t2 = np.empty(n, np.float32) if want_t2 else None
    lib().oracle_closest_hit(_fp(P), _ip(idx), idx.reshape(-1, 3).shape[0], _fp(o), _fp(d), _fp(tmax), n, _ip(prim), _fp(t),
                             _fp(b), _fp(t2) if want_t2 else None)
    return (prim, t, b, t2) if want_t2 else (prim, t, b)


def any_hit(P, idx, o, d, tmax):
    P, idx, o, d = _f(P), _i(idx), _f(o), _f(d)
    n = o.reshape(-1, 3).shape[0]
    tmax = _f(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
    occ = np.empty(n, np.int32)
    lib().oracle_any_hit(_fp(P), _ip(idx), idx.reshape(-1, 3).shape[0], _fp(o), _fp(d), _fp(tmax), n, _ip(occ))
    return occ.astype(bool)


def intersect_pairs(P, idx, prim, o, d, tmax):
    P, idx, o, d, prim = _f(P), _i(idx), _f(o), _f(d), _i(prim)
    n = o.reshape(-1, 3).shape[0]
    tmax = _f(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
    ok = np.empty(n, np.int32)
    bt = np.empty((n, 4), np.float32)
    lib().oracle_intersect_pairs(_fp(P), _ip(idx), _ip(prim), _fp(o), _fp(d), _fp(tmax), n, _ip(ok), _fp(bt))
    return ok.astype(bool), bt


def intersect_triangle(o, d, tmax, p0, p1, p2):
    out = np.zeros(4, np.float32)
    h = lib().oracle_intersect_triangle(_fp(_f(o)), _fp(_f(d)), float(tmax), _fp(_f(p0)), _fp(_f(p1)), _fp(_f(p2)), _fp(out))
    return bool(h), out


def sampler_stream(kind, spp, pixel, sample, pattern):
    pixel, sample, pattern = _i(pixel), _i(sample), _i(pattern)
    per = int(sum(1 if p == 1 else 2 for p in pattern))
    out = np.empty((len(pixel), per), np.float32)
    lib().oracle_sampler_stream(kind, spp, _ip(pixel), _ip(sample), len(pixel), _ip(pattern), len(pattern), _fp(out))
    return out


class HostBVH:
    """The oracle's own host BVH (oracle.c: Morton-split binary tree walked with the reference's loop and
    predicates).  Same answers as the brute-force functions above (argmin t; exact ties may resolve to a different
    candidate), at sizes where brute force does not finish."""

    def __init__(self, P, idx):
        self.P, self.idx = _f(P), _i(idx)          # the C side keeps pointers into these
        self.h = lib().oracle_bvh_build(_fp(self.P), _ip(self.idx), self.idx.reshape(-1, 3).shape[0])

    def closest_hit(self, o, d, tmax):
        o, d = _f(o), _f(d)
        n = o.reshape(-1, 3).shape[0]
        tmax = _f(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
        prim = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        b = np.empty((n, 2), np.float32)
        lib().oracle_bvh_closest_hit(self.h, _fp(o), _fp(d), _fp(tmax), n, _ip(prim), _fp(t), _fp(b))
        return prim, t, b

    def any_hit(self, o, d, tmax):
        o, d = _f(o), _f(d)
        n = o.reshape(-1, 3).shape[0]
        tmax = _f(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
        occ = np.empty(n, np.int32)
        lib().oracle_bvh_any_hit(self.h, _fp(o), _fp(d), _fp(tmax), n, _ip(occ))
        return occ.astype(bool)

    def node_count(self):
        return int(lib().oracle_bvh_node_count(self.h))

    def close(self):
        if getattr(self, "h", None):
            lib().oracle_bvh_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def compare_closest(P, idx, o, d, tmax, prim, t, ref_prim, ref_t, ulps=4):
    """Hit-id comparison of two closest-hit results over the same rays: returns (mismatches, near_ties).  A ray whose
    two answers differ is a NEAR TIE, not a mismatch, when the other side's primitive is itself accepted by the
    reference's triangle test at a t within `ulps` ulp of the checker's t (visit order decides those)."""
    prim, ref_prim = np.asarray(prim), np.asarray(ref_prim)
    diff = np.nonzero(prim != ref_prim)[0]
    if diff.size == 0:
        return 0, 0
    o, d = _f(o).reshape(-1, 3), _f(d).reshape(-1, 3)
    tmax = np.broadcast_to(np.asarray(tmax, dtype=np.float32), (len(prim),))
    near = 0
    cand = diff[(prim[diff] >= 0) & (ref_prim[diff] >= 0)]
    if cand.size:
        ok, bt = intersect_pairs(P, idx, prim[cand], o[cand], d[cand], tmax[cand])
        t_mine = bt[:, 3]
        du = np.abs(t_mine.view(np.int32).astype(np.int64) - np.asarray(ref_t)[cand].view(np.int32).astype(np.int64))
        near = int((ok & (du <= ulps)).sum())
    return int(diff.size - near), near
