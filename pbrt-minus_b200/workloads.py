"""Synthetic workloads of BASELINE.json's configs (seeded).

C2 "traversal-only bench": N independent triangles (random soup) + three ray classes
(primary through a pinhole camera, incoherent cosine-hemisphere bounce rays from the primary
hits, shadow segments from the primary hits to a small quad light).
"""
import numpy as np


def triangle_soup(n, seed=20251104):
    """n independent triangles: centre c ~ U[0,1)^3, vertices c + s*(u-0.5), s = 2*n^(-1/3)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c = rng.random((n, 1, 3), dtype=np.float32)
    u = rng.random((n, 3, 3), dtype=np.float32)
    s = np.float32(2.0 * n ** (-1.0 / 3.0))
    P = (c + s * (u - np.float32(0.5))).astype(np.float32).reshape(-1, 3)
    idx = np.arange(3 * n, dtype=np.int32).reshape(-1, 3)
    return P, idx


def primary_rays(width, height, fov_deg=40.0, eye=(0.5, 0.5, -1.5), jitter_index=0):
    """one ray per pixel, camera at `eye` looking +z.  jitter_index 0 = pixel centre; r > 0 = the centre
    of stratum r of a 4x4 sub-pixel grid (sample-index sharding across GPUs)."""
    aspect = width / height
    t = np.tan(np.radians(fov_deg) / 2.0)
    jx, jy = (0.5, 0.5) if jitter_index == 0 else (((jitter_index % 4) + 0.5) / 4.0, (((jitter_index // 4) % 4) + 0.5) / 4.0)
    xs = ((np.arange(width) + jx) / width * 2.0 - 1.0) * t * aspect
    ys = (1.0 - (np.arange(height) + jy) / height * 2.0) * t
    X, Y = np.meshgrid(xs, ys)
    d = np.stack([X, Y, np.ones_like(X)], axis=-1).reshape(-1, 3)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = np.broadcast_to(np.asarray(eye, dtype=np.float32), d.shape).copy()
    return o.astype(np.float32), d.astype(np.float32), np.full(d.shape[0], np.inf, dtype=np.float32)


def _hit_frames(P, idx, o, d, prim, t):
    """hit points and geometric normals (flipped against the incoming ray) for hit rays"""
    ok = prim >= 0
    tri = P[idx[prim[ok]]]  # (m,3,3)
    n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]).astype(np.float64)
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)
    din = d[ok].astype(np.float64)
    n = np.where((np.sum(n * din, axis=1) > 0)[:, None], -n, n)
    p = o[ok].astype(np.float64) + t[ok, None].astype(np.float64) * din
    return ok, p, n


def bounce_rays(P, idx, o, d, prim, t, seed=7):
    """incoherent diffuse-bounce set: cosine-hemisphere directions about the hit normal, origins
    offset along the normal (stand-in for spawn_ray, base/ray.h:30-49)"""
    ok, p, n = _hit_frames(P, idx, o, d, prim, t)
    m = p.shape[0]
    rng = np.random.Generator(np.random.PCG64(seed))
    u = rng.random((m, 2))
    r = np.sqrt(u[:, 0])
    phi = 2 * np.pi * u[:, 1]
    lx, ly, lz = r * np.cos(phi), r * np.sin(phi), np.sqrt(np.maximum(0.0, 1.0 - u[:, 0]))
    a = np.where(np.abs(n[:, :1]) > 0.9, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
    tx = np.cross(a, n)
    tx /= np.linalg.norm(tx, axis=1, keepdims=True)
    ty = np.cross(n, tx)
    w = lx[:, None] * tx + ly[:, None] * ty + lz[:, None] * n
    org = p + 1e-4 * n
    return org.astype(np.float32), w.astype(np.float32), np.full(m, np.inf, dtype=np.float32)


def shadow_rays(P, idx, o, d, prim, t, seed=11):
    """shadow set: un-normalised segments from the primary hits to uniform points on a 0.1 x 0.1 quad
    light at y = 1.5; any-hit with t_max = 1 - 1e-4"""
    ok, p, n = _hit_frames(P, idx, o, d, prim, t)
    m = p.shape[0]
    rng = np.random.Generator(np.random.PCG64(seed))
    u = rng.random((m, 2))
    q = np.stack([0.45 + 0.1 * u[:, 0], np.full(m, 1.5), 0.45 + 0.1 * u[:, 1]], axis=1)
    org = p + 1e-4 * n
    seg = q - org
    return org.astype(np.float32), seg.astype(np.float32), np.full(m, 1.0 - 1e-4, dtype=np.float32)


def write_soup(path, P, idx):
    with open(path, "wb") as f:
        np.array([P.shape[0], idx.shape[0]], dtype=np.int64).tofile(f)
        np.ascontiguousarray(P, dtype=np.float32).tofile(f)
        np.ascontiguousarray(idx, dtype=np.int32).tofile(f)


def write_rays(path, o, d, tmax):
    with open(path, "wb") as f:
        np.array([o.shape[0]], dtype=np.int64).tofile(f)
        np.ascontiguousarray(o, dtype=np.float32).tofile(f)
        np.ascontiguousarray(d, dtype=np.float32).tofile(f)
        np.ascontiguousarray(tmax, dtype=np.float32).tofile(f)


def read_hits(path, n):
    raw = np.fromfile(path, dtype=np.int32, count=n)
    t = np.fromfile(path, dtype=np.float32, count=n, offset=4 * n)
    return raw, t
