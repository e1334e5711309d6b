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


# ------------------------------------------------------------------------------------------------
# C4 "procedural scene": displaced parametric meshes (spheres / tori / terrain patches), mixed
# diffuse / conductor / dielectric materials, a ceiling grid of small emissive quads for the power
# light sampler.  Written as one .pbrt + binary little-endian PLY files so that the reference's CLI
# and this package's front-end read identical bits.
# C4 in miniature: the scene behind tests/golden/films/c4tiny_96x54_1024spp.f32 (tools/make_golden.py)
C4_TINY = dict(n_meshes=6, grid=48, n_light_quads=200, width=96, height=54, seed=7, name="c4tiny")


def _value_noise(u, v, rng, lattice=8, octaves=4):
    """periodic 2-D value noise on [0,1)^2: seeded lattice values, smoothstep interpolation, `octaves`
    octaves of halving amplitude; result in about [-1, 1]"""
    out = np.zeros_like(u, dtype=np.float64)
    amp, tot = 1.0, 0.0
    L = lattice
    for _ in range(octaves):
        g = rng.random((L, L)) * 2.0 - 1.0
        x, y = u * L, v * L
        x0, y0 = np.floor(x).astype(np.int64), np.floor(y).astype(np.int64)
        fx, fy = x - x0, y - y0
        fx, fy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
        x0, y0 = x0 % L, y0 % L
        x1, y1 = (x0 + 1) % L, (y0 + 1) % L
        out += amp * ((g[y0, x0] * (1 - fx) + g[y0, x1] * fx) * (1 - fy) + (g[y1, x0] * (1 - fx) + g[y1, x1] * fx) * fy)
        tot += amp
        amp *= 0.5
        L *= 2
    return out / tot


def _grid_indices(m):
    """two triangles per cell of an (m+1) x (m+1) vertex grid -> int32 (2*m*m, 3)"""
    i, j = np.meshgrid(np.arange(m, dtype=np.int64), np.arange(m, dtype=np.int64), indexing="ij")
    a = (i * (m + 1) + j).reshape(-1)
    b, c, d = a + 1, a + (m + 1), a + (m + 2)
    return np.stack([np.stack([a, c, b], 1), np.stack([b, c, d], 1)], 1).reshape(-1, 3).astype(np.int32)


def _displaced_mesh(kind, m, centre, size, rng):
    """(m+1)^2 vertices, 2 m^2 triangles.  kind 0: terrain patch (size x size, height-displaced),
    1: sphere of radius size/2, 2: torus (major 0.35*size, minor 0.14*size); value-noise displacement
    along the surface normal"""
    t = np.linspace(0.0, 1.0, m + 1)
    u, v = np.meshgrid(t, t, indexing="ij")
    nz = _value_noise(u % 1.0, v % 1.0, rng)
    if kind == 0:
        P = np.stack([(u - 0.5) * size, 0.12 * size * nz, (v - 0.5) * size], -1)
    elif kind == 1:
        th, ph = np.pi * v, 2 * np.pi * u
        r = 0.5 * size * (1.0 + 0.12 * nz * np.sin(th))       # no displacement at the poles (noise is not periodic there)
        P = np.stack([r * np.sin(th) * np.cos(ph), r * np.cos(th), r * np.sin(th) * np.sin(ph)], -1)
    else:
        ph, th = 2 * np.pi * u, 2 * np.pi * v
        R, r = 0.35 * size, 0.14 * size * (1.0 + 0.25 * nz)
        P = np.stack([(R + r * np.cos(th)) * np.cos(ph), r * np.sin(th), (R + r * np.cos(th)) * np.sin(ph)], -1)
    P = P.reshape(-1, 3) + np.asarray(centre, dtype=np.float64)
    return P.astype(np.float32), _grid_indices(m)


def write_ply(path, P, idx):
    """binary little-endian PLY, float x y z + uchar/int face lists (read by rply in the reference,
    shapes/tri_quad_mesh.cu, and by pbrt_scene.load_ply here)"""
    hdr = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
           "element face %d\nproperty list uchar int vertex_indices\nend_header\n" % (len(P), len(idx)))
    rec = np.empty(len(idx), dtype=np.dtype([("n", "u1"), ("v", "<i4", (3,))]))
    rec["n"] = 3
    rec["v"] = idx
    with open(path, "wb") as f:
        f.write(hdr.encode("ascii"))
        np.ascontiguousarray(P, dtype="<f4").tofile(f)
        rec.tofile(f)


def procedural_scene(out_dir, n_meshes=48, grid=721, n_light_quads=50000, width=1920, height=1080, seed=7,
                     light_levels=64, name="c4"):
    """Writes <out_dir>/<name>.pbrt + mesh PLYs; returns (scene path, dict of counts).
    Full size (defaults): 48 meshes x 2*721^2 = 49.90 M triangles + 100,000 emissive triangles = 50.0 M,
    inside a 10 x 4 x 10 box (y up).  Materials round-robin per mesh: diffuse (seeded rgb) / conductor
    (Au, Cu, Ag; roughness 0, 0.05, 0.3) / dielectric (eta 1.5; roughness 0, 0.1).  The emissive quads sit on a
    regular ceiling grid at y = 4 facing down, radiance L ~ U[1, 20] quantised to `light_levels` values
    (one AreaLightSource block per level keeps both parsers fast)."""
    import os
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.Generator(np.random.PCG64(seed))
    lines = ["LookAt 5 2.6 -7.5  5 1.2 5  0 1 0", 'Camera "perspective" "float fov" [42]',
             'Film "rgb" "string filename" ["%s.png"] "integer xresolution" [%d] "integer yresolution" [%d]' % (name, width, height),
             'Integrator "path" "integer maxdepth" [5]', "WorldBegin",
             'Material "diffuse" "rgb reflectance" [0.2 0.2 0.2]']
    metals = ("Au", "Cu", "Ag")
    n_tris = n_obj = 0
    cols = int(np.ceil(np.sqrt(n_meshes / 3.0)))          # terrain tiles per side; objects share the same grid
    cell = 10.0 / cols
    kinds = [k % 3 for k in range(n_meshes)]
    slot = {0: 0, 1: 0, 2: 0}
    for k in range(n_meshes):
        kind = kinds[k]
        s = slot[kind]
        slot[kind] += 1
        cx, cz = (s % cols + 0.5) * cell, ((s // cols) % cols + 0.5) * cell
        if kind == 0:
            centre, size = (cx, 0.0, cz), cell * 1.02
        elif kind == 1:
            centre, size = (cx - 0.18 * cell, 0.28 * cell + 0.3, cz + 0.15 * cell), 0.5 * cell
        else:
            centre, size = (cx + 0.2 * cell, 0.75 * cell + 0.5, cz - 0.15 * cell), 0.6 * cell
        P, idx = _displaced_mesh(kind, grid, centre, size, rng)
        ply = "%s_mesh_%02d.ply" % (name, k)
        write_ply(os.path.join(out_dir, ply), P, idx)
        n_tris += len(idx)
        if kind == 0:                                      # terrain stays diffuse so the floor carries the NEE load
            j = 0
        else:
            j = n_obj
            n_obj += 1
        if j % 3 == 1:
            metal = metals[(j // 3) % 3]
            mat = 'Material "conductor" "spectrum eta" ["metal-%s-eta"] "spectrum k" ["metal-%s-k"] "float roughness" [%g]' % (
                metal, metal, (0.0, 0.05, 0.3)[(j // 9) % 3])
        elif j % 3 == 2:
            mat = 'Material "dielectric" "float eta" [1.5] "float roughness" [%g]' % ((0.0, 0.1)[(j // 3) % 2])
        else:
            c = 0.15 + 0.7 * rng.random(3)
            mat = 'Material "diffuse" "rgb reflectance" [%.4f %.4f %.4f]' % tuple(c)
        lines += ["AttributeBegin", "  " + mat, '  Shape "plymesh" "string filename" ["%s"]' % ply, "AttributeEnd"]
    # ceiling grid of emissive quads (two triangles each), facing -y
    gx = int(np.ceil(np.sqrt(n_light_quads)))
    gz = (n_light_quads + gx - 1) // gx
    q = np.arange(n_light_quads)
    qx, qz = (q % gx + 0.5) * (10.0 / gx), (q // gx + 0.5) * (10.0 / gz)
    h = 0.25 * min(10.0 / gx, 10.0 / gz)
    level = rng.integers(0, light_levels, n_light_quads)
    for lv in range(light_levels):
        sel = np.nonzero(level == lv)[0]
        if len(sel) == 0:
            continue
        Lv = 1.0 + 19.0 * (lv + 0.5) / light_levels
        x, z = qx[sel], qz[sel]
        y = np.full_like(x, 4.0)
        Pq = np.stack([np.stack([x - h, y, z - h], 1), np.stack([x + h, y, z - h], 1),
                       np.stack([x + h, y, z + h], 1), np.stack([x - h, y, z + h], 1)], 1).reshape(-1, 3).astype(np.float32)
        base = 4 * np.arange(len(sel), dtype=np.int32)[:, None]
        iq = np.concatenate([base + np.array([[0, 1, 2]], np.int32), base + np.array([[2, 3, 0]], np.int32)], 1).reshape(-1, 3)
        n_tris += len(iq)
        lines += ["AttributeBegin", '  AreaLightSource "diffuse" "rgb L" [%.5f %.5f %.5f]' % (Lv, Lv, Lv),
                  '  Shape "trianglemesh" "integer indices" [%s] "point3 P" [%s]' % (
                      " ".join(map(str, iq.reshape(-1))), " ".join("%.9g" % v for v in Pq.reshape(-1))), "AttributeEnd"]
    path = os.path.join(out_dir, name + ".pbrt")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path, {"triangles": n_tris, "meshes": n_meshes, "emissive_triangles": 2 * n_light_quads, "width": width, "height": height}
