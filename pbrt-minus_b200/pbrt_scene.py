"""The `.pbrt` scene front-end: same directives and string keys as the reference's host front-end
(scene/lexer.h, scene/tokenizer.h, scene/parameter_dictionary.{h,cu}, scene/scene_builder.cu:287-469,
741-893), emitting the flat arrays the C-ABI takes (include/b200pt.h: b200pt_scene_desc) instead of a
graph of managed-memory objects.  Host code only: nothing here runs a kernel.

Supported (what the hot-path configs use; everything else raises SceneError like the reference's
REPORT_FATAL_ERROR): LookAt / Translate / Scale / Rotate / Transform / ConcatTransform / Identity /
CoordSysTransform / ReverseOrientation / AttributeBegin|End, Camera "perspective", Film "rgb",
PixelFilter box|gaussian|mitchell|triangle, Sampler stratified|independent, Integrator path|volpath,
Material / MakeNamedMaterial / NamedMaterial diffuse|conductor|dielectric|diffusetransmission|coateddiffuse|coatedconductor,
AreaLightSource "diffuse", LightSource infinite (uniform)|distant|spot, ObjectBegin|End / ObjectInstance (flattened),
Shape trianglemesh|sphere|plymesh (binary little-endian or ascii PLY, triangles and quads), Include.
"""
import math
import os
import re

import numpy as np

from . import api, spectral

F32 = np.float32


class SceneError(ValueError):
    pass


# ------------------------------------------------------------------------------------- lexing
_TYPES = ("bool", "blackbody", "float", "integer", "normal", "rgb", "point2", "point3", "string", "spectrum",
          "texture", "vector3", "vector2", "normal3", "point")
_TOKEN = re.compile(r'#[^\n]*|"([^"]*)"|\[|\]|([A-Za-z_]+)|([-+.0-9eE]+)')


def lex(text):
    """Directive stream: [(keyword, [args...])]; an arg is a str, a float, a list, or ('param', type, name)."""
    out, cur, lst = [], None, None
    for m in _TOKEN.finditer(text):
        tok = m.group(0)
        if tok.startswith("#"):
            continue
        if tok == "[":
            lst = []
            continue
        if tok == "]":
            cur[1].append(lst)
            lst = None
            continue
        if m.group(1) is not None:
            s = m.group(1)
            if lst is not None:
                lst.append(s)
            elif " " in s.strip() and s.split()[0] in _TYPES:
                t, n = s.split()[:2]
                cur[1].append(("param", t, n))
            else:
                cur[1].append(s)
            continue
        if m.group(2) is not None:
            w = m.group(2)
            if lst is not None:
                lst.append(w)            # bare true / false inside brackets
            elif w in ("true", "false"):
                cur[1].append(w)
            else:
                cur = (w, [])
                out.append(cur)
            continue
        v = float(tok)
        if lst is not None:
            lst.append(v)
        else:
            cur[1].append(v)
    return out


def params_of(args):
    """ParameterDictionary: {name: (type, values)} from the args after the type string."""
    d, i = {}, 0
    while i < len(args):
        a = args[i]
        if isinstance(a, tuple) and a[0] == "param":
            v = args[i + 1] if i + 1 < len(args) else []
            if not isinstance(v, list):
                v = [v]
            d[a[2]] = (a[1], v)
            i += 2
        else:
            i += 1
    return d


def p_float(p, name, default):
    # a parameter that names a texture ("texture roughness" ["name"]) is picked up by the caller's *_tex1 fields
    return float(p[name][1][0]) if name in p and p[name][0] != "texture" else default


def p_int(p, name, default):
    return int(p[name][1][0]) if name in p else default


def p_bool(p, name, default):
    if name not in p:
        return default
    v = p[name][1][0]
    return v is True or str(v).lower() == "true"


def p_str(p, name, default):
    return str(p[name][1][0]) if name in p else default


# ---------------------------------------------------------------------------------- transforms
# The reference's Transform carries (m, inv_m) and multiplies both (euclidean_space/transform.h:245-252)
# with an fp32 FMA chain per element (euclidean_space/squared_matrix.h:98-109).  The same is done here,
# operation for operation: vertex bits decide the order of nearly equal light areas (alias-table layout).
def _mm(a, b):
    r = np.zeros((4, 4), dtype=F32)
    for k in range(4):
        r = (a[:, k:k + 1].astype(np.float64) * b[k:k + 1, :].astype(np.float64) + r.astype(np.float64)).astype(F32)
    return r


def inverse32(m):
    """SquareMatrix<4>::inverse (euclidean_space/squared_matrix.cu:83-151): the Laplace expansion over 2x2 sub-determinants, in
    fp32.  Its difference_of_products / inner_product are error-compensated (util/math.h:167-181), i.e. the exact value rounded
    once to fp32 -- evaluated here in fp64 (products of fp32 numbers are exact there) and rounded; then s = 1 / det and one
    fp32 product per entry.  Camera and LookAt matrices come out of this, so it has to round as the reference's does."""
    m = np.asarray(m, dtype=F32).reshape(4, 4)
    d = m.astype(np.float64)

    def dop(a, b, c, e):
        return np.float64(F32(a * b - c * e))

    def ip(*t):
        return np.float64(F32(sum(t[i] * t[i + 1] for i in range(0, len(t), 2))))

    s0, s1, s2 = dop(d[0, 0], d[1, 1], d[1, 0], d[0, 1]), dop(d[0, 0], d[1, 2], d[1, 0], d[0, 2]), dop(d[0, 0], d[1, 3], d[1, 0], d[0, 3])
    s3, s4, s5 = dop(d[0, 1], d[1, 2], d[1, 1], d[0, 2]), dop(d[0, 1], d[1, 3], d[1, 1], d[0, 3]), dop(d[0, 2], d[1, 3], d[1, 2], d[0, 3])
    c0, c1, c2 = dop(d[2, 0], d[3, 1], d[3, 0], d[2, 1]), dop(d[2, 0], d[3, 2], d[3, 0], d[2, 2]), dop(d[2, 0], d[3, 3], d[3, 0], d[2, 3])
    c3, c4, c5 = dop(d[2, 1], d[3, 2], d[3, 1], d[2, 2]), dop(d[2, 1], d[3, 3], d[3, 1], d[2, 3]), dop(d[2, 2], d[3, 3], d[3, 2], d[2, 3])
    det = ip(s0, c5, -s1, c4, s2, c3, s3, c2, s5, c0, -s4, c1)
    if det == 0:
        return np.full((4, 4), np.nan, dtype=F32)
    s = F32(1.0 / det)
    rows = [[ip(d[1, 1], c5, d[1, 3], c3, -d[1, 2], c4), ip(-d[0, 1], c5, d[0, 2], c4, -d[0, 3], c3),
             ip(d[3, 1], s5, d[3, 3], s3, -d[3, 2], s4), ip(-d[2, 1], s5, d[2, 2], s4, -d[2, 3], s3)],
            [ip(-d[1, 0], c5, d[1, 2], c2, -d[1, 3], c1), ip(d[0, 0], c5, d[0, 3], c1, -d[0, 2], c2),
             ip(-d[3, 0], s5, d[3, 2], s2, -d[3, 3], s1), ip(d[2, 0], s5, d[2, 3], s1, -d[2, 2], s2)],
            [ip(d[1, 0], c4, d[1, 3], c0, -d[1, 1], c2), ip(-d[0, 0], c4, d[0, 1], c2, -d[0, 3], c0),
             ip(d[3, 0], s4, d[3, 3], s0, -d[3, 1], s2), ip(-d[2, 0], s4, d[2, 1], s2, -d[2, 3], s0)],
            [ip(-d[1, 0], c3, d[1, 1], c1, -d[1, 2], c0), ip(d[0, 0], c3, d[0, 2], c0, -d[0, 1], c1),
             ip(-d[3, 0], s3, d[3, 1], s1, -d[3, 2], s0), ip(d[2, 0], s3, d[2, 2], s0, -d[2, 1], s1)]]
    return (s * np.array(rows, dtype=F32)).astype(F32)


class Xf:
    def __init__(self, m, inv_m=None):
        self.m = np.asarray(m, dtype=F32).reshape(4, 4).copy()
        self.inv = inverse32(self.m) if inv_m is None else np.asarray(inv_m, dtype=F32).reshape(4, 4).copy()

    def __mul__(self, o):
        return Xf(_mm(self.m, o.m), _mm(o.inv, self.inv))

    def inverse(self):
        return Xf(self.inv, self.m)

    def copy(self):
        return Xf(self.m, self.inv)

    def is_identity(self):
        return np.array_equal(self.m, np.eye(4, dtype=F32)) and np.array_equal(self.inv, self.m)


def identity():
    return Xf(np.eye(4, dtype=F32), np.eye(4, dtype=F32))


def _m(rows):
    return np.array(rows, dtype=F32)


def translate(x, y, z):
    x, y, z = F32(x), F32(y), F32(z)
    return Xf(_m([[1, 0, 0, x], [0, 1, 0, y], [0, 0, 1, z], [0, 0, 0, 1]]), _m([[1, 0, 0, -x], [0, 1, 0, -y], [0, 0, 1, -z], [0, 0, 0, 1]]))


def scale(x, y, z):
    x, y, z = F32(x), F32(y), F32(z)
    return Xf(_m([[x, 0, 0, 0], [0, y, 0, 0], [0, 0, z, 0], [0, 0, 0, 1]]),
              _m([[F32(1) / x, 0, 0, 0], [0, F32(1) / y, 0, 0], [0, 0, F32(1) / z, 0], [0, 0, 0, 1]]))


def rotate(theta_deg, ax):
    """Transform::rotate (euclidean_space/transform.h:62-89): fp32 sin/cos of an fp32 angle, elements
    evaluated in double (the 1.0 literals promote) and stored as fp32; inverse = transpose."""
    theta = F32(np.float64(F32(np.pi)) / 180.0 * np.float64(F32(theta_deg)))
    s, c = np.float64(F32(np.sin(np.float64(theta)))), np.float64(F32(np.cos(np.float64(theta))))
    a = np.asarray(ax, dtype=F32)
    ln = F32(np.sqrt(F32(F32(a[0] * a[0]) + F32(a[1] * a[1])) + F32(a[2] * a[2])))
    a = (a / ln).astype(F32)
    x, y, z = a
    xx, yy, zz = np.float64(F32(x * x)), np.float64(F32(y * y)), np.float64(F32(z * z))
    xy, xz, yz = np.float64(F32(x * y)), np.float64(F32(x * z)), np.float64(F32(y * z))
    x64, y64, z64 = np.float64(x), np.float64(y), np.float64(z)
    m = _m([[xx + (1.0 - xx) * c, xy * (1.0 - c) - z64 * s, xz * (1.0 - c) + y64 * s, 0],
            [xy * (1.0 - c) + z64 * s, yy + (1.0 - yy) * c, yz * (1.0 - c) - x64 * s, 0],
            [xz * (1.0 - c) - y64 * s, yz * (1.0 - c) + x64 * s, zz + (1.0 - zz) * c, 0], [0, 0, 0, 1]])
    return Xf(m, m.T)


def lookat(pos, look, up):
    """Transform::lookat (euclidean_space/transform.h:172-207): Transform(inverse(world_from_camera), world_from_camera)."""
    pos, look, up = (np.asarray(v, dtype=F32) for v in (pos, look, up))

    def norm(v):
        return (v / F32(np.sqrt(F32(F32(v[0] * v[0]) + F32(v[1] * v[1])) + F32(v[2] * v[2])))).astype(F32)

    d = norm((look - pos).astype(F32))
    right = np.cross(norm(up).astype(np.float64), d.astype(np.float64)).astype(F32)
    if np.linalg.norm(right) == 0:
        raise SceneError("LookAt: up vector and viewing direction are parallel")
    right = norm(right)
    new_up = np.cross(d.astype(np.float64), right.astype(np.float64)).astype(F32)
    wfc = np.eye(4, dtype=F32)
    wfc[:3, 0], wfc[:3, 1], wfc[:3, 2], wfc[:3, 3] = right, new_up, d, pos
    return Xf(inverse32(wfc), wfc)


def perspective(fov, n, f):
    n, f = F32(n), F32(f)
    persp = Xf(_m([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, f / (f - n), -f * n / (f - n)], [0, 0, 1, 0]]))
    it = 1.0 / math.tan(float(F32(np.float64(F32(np.pi)) / 180.0 * fov)) / 2.0)
    return scale(it, it, 1.0) * persp


def _fma32(a, b, c):
    """fp32 fused multiply-add emulated in fp64 (the 48-bit product is exact; one rounding to 53 bits
    before the final rounding to 24 is the only difference from a hardware FMA, and it is vanishingly rare)."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(F32)


def xf_points(m, P):
    """Transform::operator()(Point3f) as the reference's gpu_transform<Point3f> kernel evaluates it on the
    device (shapes/triangle_mesh.cu:7-18; sm_100 SASS: FMUL m1*y, FFMA m0*x, FFMA m2*z, FADD m3), so mesh
    vertices get the same bits -- the alias-table layout depends on the order of nearly equal light areas."""
    P = np.asarray(P, dtype=F32).reshape(-1, 3)
    m = np.asarray(m, dtype=F32)
    x, y, z = P[:, 0], P[:, 1], P[:, 2]

    def row(r):
        t = (m[r, 1] * y).astype(F32)
        t = _fma32(m[r, 0], x, t)
        t = _fma32(m[r, 2], z, t)
        return (t + m[r, 3]).astype(F32)

    out = np.stack([row(0), row(1), row(2)], axis=1)
    w = row(3)
    if not np.all(w == 1):
        out = (out / w[:, None]).astype(F32)
    return out.astype(F32)


def xf_vectors(m, V):
    """Transform::operator()(Vector3f) (euclidean_space/transform.h:306-310), fp32 with the reference's FMA chain"""
    V = np.asarray(V, dtype=F32).reshape(-1, 3)
    out = np.zeros_like(V)
    for r in range(3):
        acc = (np.float64(m[r, 0]) * V[:, 0].astype(np.float64)).astype(F32)
        acc = (np.float64(m[r, 1]) * V[:, 1].astype(np.float64) + acc.astype(np.float64)).astype(F32)
        acc = (np.float64(m[r, 2]) * V[:, 2].astype(np.float64) + acc.astype(np.float64)).astype(F32)
        out[:, r] = acc
    return out


def xf_normals(inv_m, N):
    """Transform::operator()(Normal3f) (euclidean_space/transform.h:312-320): (inv_m^T) n, fp32"""
    N = np.asarray(N, dtype=F32).reshape(-1, 3)
    im = np.asarray(inv_m, dtype=F32)
    x, y, z = N[:, 0], N[:, 1], N[:, 2]
    cols = []
    for c in range(3):
        t = (im[0, c] * x).astype(F32)
        t = _fma32(im[1, c], y, t)
        t = _fma32(im[2, c], z, t)
        cols.append(t)
    return np.stack(cols, axis=1).astype(F32)


def swaps_handedness(m):
    return float(np.linalg.det(m[:3, :3].astype(np.float64))) < 0


# ---------------------------------------------------------------------------------------- PLY
def read_ply(path):
    """Triangles (quads split 0-1-2, 0-2-3) of an ascii / binary_little_endian PLY (shapes/tri_quad_mesh.cu)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header") + len(b"end_header")
    end = data.index(b"\n", end) + 1
    header = data[:end].decode("ascii", "replace").split("\n")
    fmt, elems, cur = None, [], None
    for line in header:
        t = line.split()
        if not t:
            continue
        if t[0] == "format":
            fmt = t[1]
        elif t[0] == "element":
            cur = {"name": t[1], "count": int(t[2]), "props": []}
            elems.append(cur)
        elif t[0] == "property":
            cur["props"].append(t[1:])
    np_t = {"char": "i1", "uchar": "u1", "int8": "i1", "uint8": "u1", "short": "i2", "ushort": "u2", "int16": "i2",
            "uint16": "u2", "int": "i4", "uint": "u4", "int32": "i4", "uint32": "u4", "float": "f4", "float32": "f4",
            "double": "f8", "float64": "f8"}
    body, off = data[end:], 0
    verts = faces = None
    if fmt == "ascii":
        toks = body.split()
        pos = 0
    for e in elems:
        if e["name"] == "vertex":
            names = [p[-1] for p in e["props"]]
            if fmt == "ascii":
                k = len(names)
                arr = np.array(toks[pos:pos + k * e["count"]], dtype=np.float64).reshape(e["count"], k)
                pos += k * e["count"]
                verts = {n: arr[:, i] for i, n in enumerate(names)}
            else:
                dt = np.dtype([(p[-1], "<" + np_t[p[0]]) for p in e["props"]])
                arr = np.frombuffer(body, dtype=dt, count=e["count"], offset=off)
                off += dt.itemsize * e["count"]
                verts = {n: arr[n] for n in names}
        elif e["name"] == "face":
            tris = []
            if fmt == "ascii":
                for _ in range(e["count"]):
                    k = int(toks[pos])
                    idx = [int(v) for v in toks[pos + 1:pos + 1 + k]]
                    pos += 1 + k
                    tris.append(idx[:3])
                    if k == 4:
                        tris.append([idx[0], idx[2], idx[3]])
                faces = np.array(tris, dtype=np.int32)
            else:
                lp = [p for p in e["props"] if p[0] == "list"][0]
                ct, it = "<" + np_t[lp[1]], "<" + np_t[lp[2]]
                cs, isz = np.dtype(ct).itemsize, np.dtype(it).itemsize
                first = int(np.frombuffer(body, dtype=ct, count=1, offset=off)[0])
                rec = np.dtype([("n", ct), ("v", it, (first,))])
                arr = np.frombuffer(body, dtype=rec, count=e["count"], offset=off)
                if len(e["props"]) != 1 or not np.all(arr["n"] == first):
                    raise SceneError("PLY: mixed polygon sizes are not supported")
                off += rec.itemsize * e["count"]
                v = arr["v"].astype(np.int32)
                faces = v if first == 3 else np.concatenate([v[:, [0, 1, 2]], v[:, [0, 2, 3]]]) if first == 4 else None
                if faces is None:
                    raise SceneError("PLY: only triangles and quads")
                del cs, isz
    P = np.stack([verts["x"], verts["y"], verts["z"]], axis=1).astype(F32)
    N = np.stack([verts["nx"], verts["ny"], verts["nz"]], axis=1).astype(F32) if "nx" in verts else None
    uv = None
    for a, b in (("u", "v"), ("s", "t"), ("texture_u", "texture_v"), ("texture_s", "texture_t")):   # shapes/tri_quad_mesh.cu
        if a in verts and b in verts:
            uv = np.stack([verts[a], verts[b]], axis=1).astype(F32)
            break
    return P, faces, N, uv


# ------------------------------------------------------------------------------ pixel filters
def filter_table(kind, p):
    """FilterSampler's table (filters/filter_sampler.cu:4-20): f at the cell centres of a 32*radius grid."""
    defaults = {"box": 0.5, "gaussian": 1.5, "mitchell": 2.0, "triangle": 2.0}
    if kind not in defaults:
        raise SceneError("PixelFilter '%s' not implemented" % kind)
    rx, ry = F32(p_float(p, "xradius", defaults[kind])), F32(p_float(p, "yradius", defaults[kind]))
    nx, ny = int(F32(32) * rx), int(F32(32) * ry)
    xs = (F32(-rx) * (F32(1) - (np.arange(nx, dtype=F32) + F32(0.5)) / F32(nx)) + (np.arange(nx, dtype=F32) + F32(0.5)) / F32(nx) * rx).astype(F32)
    ys = (F32(-ry) * (F32(1) - (np.arange(ny, dtype=F32) + F32(0.5)) / F32(ny)) + (np.arange(ny, dtype=F32) + F32(0.5)) / F32(ny) * ry).astype(F32)
    if kind == "mitchell":
        b, c = F32(p_float(p, "B", 1.0 / 3.0)), F32(p_float(p, "C", 1.0 / 3.0))

        def m1d(x):   # filters/mitchell.cu:25-39
            x = np.abs(x).astype(F32)
            a = ((12 - 9 * b - 6 * c) * x * x * x + (-18 + 12 * b + 6 * c) * x * x + (6 - 2 * b)) * F32(1.0 / 6.0)
            bb = ((-b - 6 * c) * x * x * x + (6 * b + 30 * c) * x * x + (-12 * b - 48 * c) * x + (8 * b + 24 * c)) * F32(1.0 / 6.0)
            return np.where(x <= 1, a, np.where(x <= 2, bb, 0)).astype(F32)

        f = np.outer(m1d(F32(2) * ys / ry), m1d(F32(2) * xs / rx))
    elif kind == "box":
        f = np.ones((ny, nx))
    elif kind == "triangle":
        f = np.outer(np.maximum(0, ry - np.abs(ys)), np.maximum(0, rx - np.abs(xs)))
    else:  # gaussian
        sigma = F32(p_float(p, "sigma", 0.5))

        def g(x, r):   # gaussian() and GaussianFilter::evaluate (filters/gaussian.cu:8-11,33-37), one fp32 operation at a time
            two_ss = F32(F32(F32(2) * sigma) * sigma)
            norm = F32(F32(1) / F32(np.sqrt(F32(F32(F32(F32(2) * F32(np.pi)) * sigma) * sigma))))

            def gs(v):
                v = np.asarray(v, dtype=F32)
                arg = (-(v * v).astype(F32) / two_ss).astype(F32)
                return (norm * np.exp(arg.astype(np.float64)).astype(F32)).astype(F32)   # expf: the exact value rounded once
            return np.maximum(F32(0), (gs(x) - gs(r)).astype(F32)).astype(F32)

        f = np.outer(g(ys, ry), g(xs, rx))
    return float(rx), float(ry), nx, ny, np.ascontiguousarray(f, dtype=F32)


# ------------------------------------------------------------------------------ scene assembly
class Scene:
    """Flat scene arrays + options, ready for b200pt_scene_create."""

    def __init__(self):
        self.P, self.idx, self.tri_material, self.tri_light, self.tri_flip = [], [], [], [], []
        self.N, self.uv, self.any_N, self.any_uv = [], [], False, False
        self.media, self.tri_medium, self.camera_medium = [], [], -1
        self.textures, self.images, self.named_spec_tex, self.named_float_tex, self._image_ids = [], [], {}, {}, {}
        # the reference instantiates every spectrum Texture once per SpectrumType (scene_builder.cu:679-708); here: the albedo graph
        # (reflectance / transmittance) in named_spec_tex and the unbounded one (conductor eta / k) in named_spec_tex_unbounded
        self.named_spec_tex_unbounded = {}
        self.n_verts = 0
        self.spheres, self.spectra, self.materials, self.lights = [], [], [], []
        self.camera = self.filter = self.film = None
        self.spp, self.max_depth, self.regularize, self.sampler = None, 5, False, "stratified"
        self.integrator = "path"
        self.filename = "pbrt.png"
        self._spec_cache = {}
        self._keep = []

    # ---- spectra
    def add_spectrum(self, key, rec):
        if key in self._spec_cache:
            return self._spec_cache[key]
        self.spectra.append(rec)
        self._spec_cache[key] = len(self.spectra) - 1
        return len(self.spectra) - 1

    def spec_constant(self, v):
        return self.add_spectrum(("c", float(v)), {"type": 0, "c": [float(v), 0, 0, 0]})

    def spec_rgb(self, rgb, kind):
        rgb = tuple(float(v) for v in rgb)
        if kind == "albedo":
            if max(rgb) > 1:
                raise SceneError("rgb albedo components must be <= 1 (base/spectrum.cu:67-69)")
            c = spectral.rgb_to_sigmoid(rgb)
            return self.add_spectrum(("a",) + rgb, {"type": 1, "c": [c[0], c[1], c[2], 1.0]})
        # RGBUnboundedSpectrum / RGBIlluminantSpectrum ctors (spectra/rgb_unbounded_spectrum.cu:5-14, rgb_illuminant_spectrum.cu:6-10):
        # the components are fp32 (Real), scale = 2 * max and rgb / scale are fp32 operations
        r32 = np.asarray(rgb, dtype=F32)
        sc = F32(F32(2) * r32.max())
        c = spectral.rgb_to_sigmoid((r32 / sc).astype(F32) if sc > 0 else (0, 0, 0))
        if kind == "unbounded":
            return self.add_spectrum(("u",) + rgb, {"type": 2, "c": [c[0], c[1], c[2], float(sc)]})
        return self.add_spectrum(("i",) + rgb, {"type": 3, "c": [c[0], c[1], c[2], float(sc)]})

    def spec_piecewise(self, key, lam, val):
        lam, val = np.ascontiguousarray(lam, dtype=F32), np.ascontiguousarray(val, dtype=F32)
        return self.add_spectrum(("p", key), {"type": 4, "c": [0, 0, 0, 0], "lambda": lam, "value": val})

    def spec_param(self, p, name, kind, scene_dir):
        """ParameterDictionary::get_spectrum (scene/parameter_dictionary.cu:127-176,198-221)."""
        if name not in p:
            return None
        t, v = p[name]
        if t == "texture":          # resolved by material_texture(); the spectrum slot gets a placeholder
            if name not in ("reflectance", "transmittance", "eta", "k", "conductor.eta", "conductor.k"):
                raise SceneError("a texture for '%s' is outside the hot path" % name)
            return self.spec_constant(0.5)
        if t == "rgb":
            return self.spec_rgb(v[:3], kind)
        if t == "spectrum":
            if isinstance(v[0], str):
                path = os.path.join(scene_dir, v[0])
                if os.path.exists(path):
                    nums = [float(x) for x in open(path).read().split() if not x.startswith("#")]
                    lam, val = spectral.interleaved_to_knots(nums)
                    return self.spec_piecewise(("file", path), lam, val)
                lam, val = spectral.named_spectrum(v[0])
                return self.spec_piecewise(("named", v[0]), lam, val)
            lam, val = spectral.interleaved_to_knots(v)
            return self.spec_piecewise(("inline", tuple(v)), lam, val)
        if t == "float":
            return self.spec_constant(v[0])
        raise SceneError("spectrum parameter '%s' of type '%s' not supported" % (name, t))

    # ---- textures (SceneBuilder::parse_texture, scene_builder.cu:679-708; {Spectrum,Float}Texture::create)
    def _tex(self, type_, a=-1, b=-1, c=-1, f=(0, 0, 0, 0), mapping=(1, 1, 0, 0), kind=0, tfr=None, v1=(1, 0, 0), v2=(0, 1, 0)):
        self.textures.append({"type": type_, "a": a, "b": b, "c": c, "f": [float(v) for v in f], "map": [float(v) for v in mapping],
                              "mapping": int(kind), "tfr": [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0] if tfr is None else [float(v) for v in tfr],
                              "v1": [float(v) for v in v1], "v2": [float(v) for v in v2]})
        return len(self.textures) - 1

    def load_image(self, path):
        """GPUImage::init_png (textures/gpu_image.cu:194-222): 8-bit RGBA through the sRGB decoding table, top row first"""
        if path in self._image_ids:
            return self._image_ids[path]
        if not path.lower().endswith(".png"):
            raise SceneError("image textures: only .png is read here ('%s'); .exr / .pfm / .tga are outside the hot path" % path)
        if not os.path.exists(path):
            raise SceneError("image '%s' not found" % path)
        from PIL import Image
        im = Image.open(path)
        if im.mode in ("I;16", "I;16B", "I"):
            raise SceneError("16-bit PNG textures are not supported")
        px = np.asarray(im.convert("RGBA"), dtype=np.uint8)[:, :, :3]
        lut = spectral.tables()["srgb_to_linear_lut"].astype(F32)   # SRGBColorEncoding::to_linear (spectrum_util/color_encoding.h:29-31,46)
        self.images.append(np.ascontiguousarray(lut[px], dtype=F32))
        self._image_ids[path] = len(self.images) - 1
        return self._image_ids[path]

    def spec_tex_table(self, kind):
        return self.named_spec_tex if kind == "albedo" else self.named_spec_tex_unbounded

    def spectrum_texture_param(self, p, key, scene_dir, default=None, kind="albedo"):
        """ParameterDictionary::get_spectrum_texture (parameter_dictionary.cu:276-330): a named texture, or a constant one"""
        if key in p:
            t, v = p[key]
            if t == "texture":
                table = self.spec_tex_table(kind)
                if str(v[0]) not in table:
                    raise SceneError("spectrum texture '%s' not defined" % v[0])
                if table[str(v[0])] is None:
                    raise SceneError("texture '%s' holds an image: the reference's image textures abort for %s spectra "
                                     "(textures/spectrum_image_texture.cu:23-33)" % (v[0], kind))
                return table[str(v[0])]
            return self._tex(0, a=self.spec_param(p, key, kind, scene_dir))
        return None if default is None else self._tex(0, a=self.spec_constant(default))

    def float_texture_param(self, p, key, default):
        if key in p:
            t, v = p[key]
            if t == "texture":
                if str(v[0]) not in self.named_float_tex:
                    raise SceneError("float texture '%s' not defined" % v[0])
                return self.named_float_tex[str(v[0])]
            return self._tex(16, f=(float(v[0]), 0, 0, 0))
        return self._tex(16, f=(default, 0, 0, 0))

    def _image_node(self, type_, p, scene_dir, render_from_texture=None):
        # TextureMapping2D::create (textures/texture_mapping_2d.cu:86-133)
        mapping = p_str(p, "mapping", "uv")
        kinds = {"uv": 0, "planar": 1, "spherical": 2, "cylindrical": 3}
        if mapping not in kinds:
            raise SceneError("mapping `%s` not implemented (texture_mapping_2d.cu:129-132)" % mapping)
        extra = {}
        if mapping != "uv":
            if render_from_texture is None:
                raise SceneError("texture mapping '%s' needs the Texture directive's transform" % mapping)
            extra = {"kind": kinds[mapping], "tfr": render_from_texture.inv[:3, :].reshape(-1)}
            if mapping == "planar":
                extra["v1"] = p["v1"][1][:3] if "v1" in p else (1, 0, 0)
                extra["v2"] = p["v2"][1][:3] if "v2" in p else (0, 1, 0)
        if p_str(p, "wrap", "repeat") != "repeat":
            raise SceneError("texture wrap '%s': the reference implements \"repeat\" only (gpu_image.cu:56-66)" % p_str(p, "wrap", "repeat"))
        if p_str(p, "filter", "bilinear") not in ("ewa", "bilinear", "trilinear", "point"):
            raise SceneError("unknown texture filter")       # every filter is a bilerp in the reference (mipmap.cu:22-25)
        img = self.load_image(os.path.join(scene_dir, p_str(p, "filename", "")))
        return self._tex(type_, a=img, f=(p_float(p, "scale", 1.0), 1.0 if p_bool(p, "invert", False) else 0.0, 0, 0),
                         mapping=(p_float(p, "uscale", 1.0), p_float(p, "vscale", 1.0), p_float(p, "udelta", 0.0), p_float(p, "vdelta", 0.0)),
                         **extra)

    def add_texture(self, name, color_type, tex_type, p, render_from_texture, scene_dir):
        if color_type == "float":
            if tex_type == "constant":
                tid = self._tex(16, f=(p_float(p, "value", 1.0), 0, 0, 0))
            elif tex_type == "imagemap":
                tid = self._image_node(17, p, scene_dir, render_from_texture)
            elif tex_type == "scale":
                tid = self._tex(18, a=self.float_texture_param(p, "scale", 1.0), b=self.float_texture_param(p, "tex", 1.0))
            else:
                raise SceneError("float texture '%s' is not a reference texture (base/float_texture.cu:7-35)" % tex_type)
            self.named_float_tex[name] = tid
            return
        if color_type != "spectrum":
            raise SceneError("Texture colour type '%s' (scene_builder.cu:706-707)" % color_type)
        self.named_spec_tex[name] = self._spectrum_texture(tex_type, p, render_from_texture, scene_dir, "albedo")
        try:    # the unbounded instantiation of the same directive (conductor eta / k); images and their users have none
            self.named_spec_tex_unbounded[name] = self._spectrum_texture(tex_type, p, render_from_texture, scene_dir, "unbounded")
        except SceneError:
            self.named_spec_tex_unbounded[name] = None

    def _spectrum_texture(self, tex_type, p, render_from_texture, scene_dir, kind):
        if tex_type == "constant":
            v = self.spec_param(p, "value", kind, scene_dir)
            tid = self._tex(0, a=self.spec_constant(1.0) if v is None else v)
        elif tex_type == "imagemap":
            if kind != "albedo":
                raise SceneError("image textures abort in the reference for %s spectra (textures/spectrum_image_texture.cu:23-33)" % kind)
            tid = self._image_node(1, p, scene_dir, render_from_texture)
        elif tex_type == "scale":
            t = self.spectrum_texture_param(p, "tex", scene_dir, kind=kind)
            if t is None:
                raise SceneError("scale texture without \"tex\"")
            tid = self._tex(2, a=t, f=(p_float(p, "scale", 1.0), 0, 0, 0))
        elif tex_type == "mix":
            t1 = self.spectrum_texture_param(p, "tex1", scene_dir, 0.0, kind=kind)
            t2 = self.spectrum_texture_param(p, "tex2", scene_dir, 1.0, kind=kind)
            tid = self._tex(3, a=t1, b=t2, c=self.float_texture_param(p, "amount", 0.5))
        elif tex_type == "directionmix":
            t1 = self.spectrum_texture_param(p, "tex1", scene_dir, 0.0, kind=kind)
            t2 = self.spectrum_texture_param(p, "tex2", scene_dir, 1.0, kind=kind)
            d = np.asarray(p["dir"][1][:3] if "dir" in p else (0, 1, 0), dtype=F32)
            d = xf_vectors(render_from_texture.m, d.reshape(1, 3))[0]
            d = (d / F32(np.sqrt(F32(F32(d[0] * d[0]) + F32(d[1] * d[1])) + F32(d[2] * d[2])))).astype(F32)
            tid = self._tex(4, a=t1, b=t2, f=(d[0], d[1], d[2], 0))
        elif tex_type == "checkerboard":
            # CameraBase::minPosDifferentialX/Y, minDirDifferentialX/Y (cameras/camera_base.h:96-97) are never assigned and
            # Vector3's default constructor fills NaN (euclidean_space/vector3.h:12), so approximate_dp_dxy hands NaN
            # differentials to checkerboard(): its weight is NaN and the reference aborts on the NaN pixels it writes
            raise SceneError("checkerboard texture: in the reference its filter width is NaN (uninitialised camera differentials), "
                             "every textured pixel is NaN and the render aborts; nothing to reproduce")
        else:
            raise SceneError("spectrum texture '%s' is not a reference texture (base/spectrum_texture.cu:9-58)" % tex_type)
        return tid

    def material_texture(self, p, key, kind="albedo"):
        """texture id + 1 when the material parameter names a texture, else 0"""
        if key in p and p[key][0] == "texture":
            return self.spectrum_texture_param(p, key, None, kind=kind) + 1
        return 0

    # ---- materials
    def add_material(self, kind, p, scene_dir, named=None):
        n_before = len(self.materials)
        self._add_material(kind, p, scene_dir, named)
        if len(self.materials) > n_before:
            m = self.materials[-1]
            m["reflectance_tex1"] = self.material_texture(p, "reflectance") if m["type"] in (0, 1, 3, 4, 5) else 0
            m["transmittance_tex1"] = self.material_texture(p, "transmittance") if m["type"] == 3 else 0
            if m["type"] == 1:        # conductor_material.cu:52-75: eta / k are Unbounded spectrum textures
                m["eta_tex1"], m["k_tex1"] = self.material_texture(p, "eta", "unbounded"), self.material_texture(p, "k", "unbounded")
            elif m["type"] == 5:      # coated_conductor_material.cu:26-90
                m["eta_tex1"] = self.material_texture(p, "conductor.eta", "unbounded")
                m["k_tex1"] = self.material_texture(p, "conductor.k", "unbounded")
            if m["type"] in (1, 2):   # conductor_material.cu:35-44, dielectric_material.cu:20-31: "uroughness" else "roughness"
                for fld, key in (("ur_tex1", "uroughness"), ("vr_tex1", "vroughness")):
                    k = key if key in p else "roughness"
                    m[fld] = self.float_texture_param(p, k, 0.0) + 1 if k in p and p[k][0] == "texture" else 0
        return len(self.materials) - 1

    def _add_material(self, kind, p, scene_dir, named=None):
        if kind == "diffuse":
            r = self.spec_param(p, "reflectance", "albedo", scene_dir)
            if r is None:
                r = self.spec_constant(0.5)
            self.materials.append({"type": 0, "reflectance": r, "eta": -1, "k": -1, "ur": 0.0, "vr": 0.0, "remap": 1})
        elif kind == "conductor":
            refl = self.spec_param(p, "reflectance", "albedo", scene_dir)
            eta = self.spec_param(p, "eta", "unbounded", scene_dir)
            k = self.spec_param(p, "k", "unbounded", scene_dir)
            if refl is not None and (eta is not None or k is not None):
                raise SceneError("conductor: both reflectance and (eta, k) given")
            if refl is None:
                if eta is None:
                    eta = self.spec_piecewise(("named", "metal-Cu-eta"), *spectral.named_spectrum("metal-Cu-eta"))
                if k is None:
                    k = self.spec_piecewise(("named", "metal-Cu-k"), *spectral.named_spectrum("metal-Cu-k"))
            ro = p_float(p, "roughness", 0.0)
            self.materials.append({"type": 1, "reflectance": -1 if refl is None else refl, "eta": -1 if eta is None else eta,
                                   "k": -1 if k is None else k, "ur": p_float(p, "uroughness", ro), "vr": p_float(p, "vroughness", ro),
                                   "remap": int(p_bool(p, "remaproughness", True))})
        elif kind == "dielectric":
            if "eta" in p and p["eta"][0] == "float":
                eta = self.spec_constant(p["eta"][1][0])
            else:
                eta = self.spec_param(p, "eta", "unbounded", scene_dir)
            if eta is None:
                eta = self.spec_constant(1.5)
            ro = p_float(p, "roughness", 0.0)
            self.materials.append({"type": 2, "reflectance": -1, "eta": eta, "k": -1, "ur": p_float(p, "uroughness", ro),
                                   "vr": p_float(p, "vroughness", ro), "remap": int(p_bool(p, "remaproughness", True))})
        elif kind == "diffusetransmission":
            # materials/diffuse_transmission_material.cu:8-23 (constant-float textures of 0.25 when absent)
            r = self.spec_param(p, "reflectance", "albedo", scene_dir)
            t = self.spec_param(p, "transmittance", "albedo", scene_dir)
            self.materials.append({"type": 3, "reflectance": self.spec_constant(0.25) if r is None else r, "eta": -1, "k": -1,
                                   "ur": 0.0, "vr": 0.0, "remap": 1, "transmittance": self.spec_constant(0.25) if t is None else t,
                                   "scale": p_float(p, "scale", 1.0)})
        elif kind == "interface":
            # base/material.cu:34-36: no material at all; the surface only separates media
            self.materials.append({"type": 7, "reflectance": -1, "eta": -1, "k": -1, "ur": 0.0, "vr": 0.0, "remap": 1})
        elif kind == "mix":
            # materials/mix_material.cu:6-16: two NAMED materials and an amount in (0, 1)
            amount = p_float(p, "amount", 0.5)
            names = [str(v) for v in p["materials"][1]] if "materials" in p else []
            if not (0.0 < amount < 1.0) or len(names) != 2 or any(n not in (named or {}) for n in names):
                raise SceneError("mix: needs \"amount\" in (0, 1) and two named materials")
            self.materials.append({"type": 6, "reflectance": -1, "eta": -1, "k": -1, "ur": 0.0, "vr": 0.0, "remap": 1,
                                   "scale": amount, "mix_a": named[names[0]], "mix_b": named[names[1]]})
        elif kind == "coateddiffuse":
            # materials/coated_diffuse_material.cu:9-51
            r = self.spec_param(p, "reflectance", "albedo", scene_dir)
            ro = p_float(p, "roughness", 0.0)
            self.materials.append({"type": 4, "reflectance": self.spec_constant(0.5) if r is None else r,
                                   "eta": self.spec_constant(p_float(p, "eta", 1.5)), "k": -1,
                                   "ur": p_float(p, "uroughness", ro), "vr": p_float(p, "vroughness", ro),
                                   "remap": int(p_bool(p, "remaproughness", True)), "thickness": p_float(p, "thickness", 0.01),
                                   "g": p_float(p, "g", 0.0), "albedo": self.spec_constant(p_float(p, "albedo", 0.0)),
                                   "max_depth": p_int(p, "maxdepth", 10), "n_samples": p_int(p, "nsamples", 1)})
        elif kind == "coatedconductor":
            # materials/coated_conductor_material.cu:26-90
            def rough(primary, secondary):
                return p_float(p, primary, p_float(p, secondary, 0.0))
            if "interface.eta" in p and p["interface.eta"][0] == "float":
                ieta = self.spec_constant(p["interface.eta"][1][0])
            else:
                ieta = self.spec_param(p, "interface.eta", "unbounded", scene_dir)
            if ieta is None:
                ieta = self.spec_constant(1.5)
            ceta = self.spec_param(p, "conductor.eta", "unbounded", scene_dir)
            ck = self.spec_param(p, "conductor.k", "unbounded", scene_dir)
            refl = self.spec_param(p, "reflectance", "albedo", scene_dir)
            if refl is not None and (ceta is not None or ck is not None):
                raise SceneError("coatedconductor: both reflectance and (eta, k) given")
            if refl is None:
                if ceta is None:
                    ceta = self.spec_piecewise(("named", "metal-Cu-eta"), *spectral.named_spectrum("metal-Cu-eta"))
                if ck is None:
                    ck = self.spec_piecewise(("named", "metal-Cu-k"), *spectral.named_spectrum("metal-Cu-k"))
            alb = self.spec_param(p, "albedo", "albedo", scene_dir)
            self.materials.append({"type": 5, "reflectance": -1 if refl is None else refl, "eta": ieta, "k": -1,
                                   "ur": rough("interface.uroughness", "interface.roughness"),
                                   "vr": rough("interface.vroughness", "interface.roughness"),
                                   "remap": int(p_bool(p, "remaproughness", True)), "thickness": p_float(p, "thickness", 0.01),
                                   "g": p_float(p, "g", 0.0), "albedo": self.spec_constant(0.0) if alb is None else alb,
                                   "max_depth": p_int(p, "maxdepth", 10), "n_samples": p_int(p, "nsamples", 1),
                                   "cur": rough("conductor.uroughness", "conductor.roughness"),
                                   "cvr": rough("conductor.vroughness", "conductor.roughness"),
                                   "c_eta": -1 if ceta is None else ceta, "c_k": -1 if ck is None else ck})
        else:
            raise SceneError("Material '%s' is outside the hot path (SURVEY.md 8f)" % kind)
        return len(self.materials) - 1

    # ---- media (Medium::create, medium/homogeneous_medium.cu:15-41)
    def add_medium(self, p, scene_dir):
        if p_str(p, "type", "homogeneous") != "homogeneous":
            raise SceneError("only homogeneous media (medium/homogeneous_medium.h)")
        if "Le" in p and p_float(p, "Lescale", 1.0) != 0:
            raise SceneError("illuminating volume not implemented (homogeneous_medium.cu:25-30)")
        sa = self.spec_param(p, "sigma_a", "unbounded", scene_dir)
        ss = self.spec_param(p, "sigma_s", "unbounded", scene_dir)
        sa, ss = self.spec_constant(1.0) if sa is None else sa, self.spec_constant(1.0) if ss is None else ss
        # Medium::Medium (medium/homogeneous_medium.cu:7-13) keeps DenselySampledSpectrum(sigma, scale): the spectrum times "scale"
        # tabulated at the integer wavelengths on the host, interpolated linearly in between at render time -- not the analytic curve
        scale = F32(p_float(p, "scale", 1.0))
        dense = []
        for sid in (sa, ss):
            v = spectral.dense_scaled32(self.spectra[sid], scale)
            dense.append(self.add_spectrum(("dense", v.tobytes()), {"type": 5, "c": [0, 0, 0, 0], "value": v}))
        self.media.append({"sigma_a": dense[0], "sigma_s": dense[1], "scale": 1.0, "g": p_float(p, "g", 0.0)})
        return len(self.media) - 1

    # ---- lights that are not emissive shapes (SceneBuilder::parse_light_source, scene_builder.cu:493-504)
    def add_light_source(self, kind, p, render_from_light, scene_dir):
        def illum(name):
            sid = self.spec_param(p, name, "illuminant", scene_dir)
            if sid is None:
                lam, val, _ = spectral.illuminant_d65()
                sid = self.spec_piecewise(("illuminant",), lam, val)
            return sid

        def vnorm(v):
            v = np.asarray(v, dtype=F32)
            return (v / F32(np.sqrt(F32(F32(v[0] * v[0]) + F32(v[1] * v[1])) + F32(v[2] * v[2])))).astype(F32)

        def coordinate_system(v):       # euclidean_space/vector3.h:161-167, all in fp32
            x, y, z = (F32(c) for c in v)
            sign = F32(np.copysign(1.0, z))
            a = F32(F32(-1.0) / F32(sign + z))
            b = F32(F32(x * y) * a)
            v2 = np.array([F32(F32(1) + F32(F32(sign * F32(x * x)) * a)), F32(sign * b), F32(-sign * x)], dtype=F32)
            v3 = np.array([b, F32(sign + F32(F32(y * y) * a)), F32(-y)], dtype=F32)
            return v2, v3

        rec = {"triangle": -1, "two_sided": 0, "v": [0.0, 0.0, 0.0], "lfr": [0.0] * 9, "cs": 0.0, "ce": 0.0}
        frm = np.asarray(p["from"][1][:3] if "from" in p else (0, 0, 0), dtype=F32)
        to = np.asarray(p["to"][1][:3] if "to" in p else (0, 0, 1), dtype=F32)
        sc = F32(p_float(p, "scale", 1.0))
        if kind == "distant":           # lights/distant_light.cu:10-41
            L = illum("L")
            w = vnorm((frm - to).astype(F32))
            v1, v2 = coordinate_system(w)
            m = np.eye(4, dtype=F32)
            m[:3, 0], m[:3, 1], m[:3, 2] = v1, v2, w
            final = render_from_light * Xf(m)
            sc = F32(sc / self.photometric(L))
            ev = p_float(p, "illuminance", -1.0)
            if ev > 0:
                sc = F32(sc * F32(ev))
            rec.update({"type": 1, "spectrum": L, "scale": float(sc), "v": [float(x) for x in final.m[:3, 2]]})
        elif kind == "infinite":        # lights/uniform_infinite_light.cu:7-32
            if "filename" in p:         # lights/image_infinite_light.cu:62-75 (base/light.cu:30-36)
                img = self.load_image(os.path.join(scene_dir, p_str(p, "filename", "")))
                d65 = self.spec_piecewise(("illuminant",), *spectral.illuminant_d65()[:2])
                sc = F32(sc / self.photometric(d65))
                m, inv = render_from_light.m, render_from_light.inv
                rec.update({"type": 4, "spectrum": -1, "scale": float(sc), "image": img,
                            "lfr": [float(x) for x in inv[:3, :3].reshape(-1)], "rfl": [float(x) for x in m[:3, :3].reshape(-1)]})
                self.lights.append(rec)
                return
            L = illum("L")
            sc = F32(sc / self.photometric(L))
            ev = p_float(p, "illuminance", -1.0)
            if ev > 0:
                sc = F32(np.float64(sc) * (np.float64(F32(ev)) / np.float64(F32(np.pi))))
            rec.update({"type": 2, "spectrum": L, "scale": float(sc)})
        elif kind == "spot":            # lights/spot_light.cu:8-41
            I = illum("I")
            cone, delta = F32(p_float(p, "coneangle", 30.0)), F32(p_float(p, "conedeltaangle", 5.0))
            z = vnorm((to - frm).astype(F32))
            x, y = coordinate_system(z)
            m = np.eye(4, dtype=F32)
            m[0, :3], m[1, :3], m[2, :3] = x, y, z
            dir_to_z = Xf(m)
            final = render_from_light * (translate(*frm) * dir_to_z.inverse())
            sc = F32(sc / self.photometric(I))
            rad = lambda d: F32(np.float64(F32(np.pi)) / 180.0 * np.float64(F32(d)))
            ce, cs = F32(np.cos(np.float64(rad(cone)))), F32(np.cos(np.float64(rad(F32(cone - delta)))))
            pw = p_float(p, "power", -1.0)
            if pw > 0:
                ke = 2 * np.float64(F32(np.pi)) * ((1 - np.float64(cs)) + (np.float64(cs) - np.float64(ce)) / 2)
                sc = F32(np.float64(sc) * (np.float64(F32(pw)) / ke))
            pos = xf_points(final.m, [[0, 0, 0]])[0]
            rec.update({"type": 3, "spectrum": I, "scale": float(sc), "v": [float(v) for v in pos],
                        "lfr": [float(v) for v in final.inv[:3, :3].reshape(-1)], "cs": float(cs), "ce": float(ce)})
        else:
            raise SceneError("LightSource '%s' is outside the hot path (SURVEY.md 8f)" % kind)
        self.lights.append(rec)

    # ---- geometry
    def add_mesh(self, P, idx, material, flip, light_params, scene_dir, N=None, uv=None, medium_code=0):
        P = np.asarray(P, dtype=F32).reshape(-1, 3)
        idx = np.asarray(idx, dtype=np.int32).reshape(-1, 3)
        nt = len(idx)
        # per-vertex shading normals / uvs ride in scene-wide arrays (zeros for meshes without); tri_flip bits 1 / 2 say
        # which triangles use them
        self.N.append((len(P), None if N is None else np.asarray(N, dtype=F32).reshape(-1, 3)))
        self.uv.append((len(P), None if uv is None else np.asarray(uv, dtype=F32).reshape(-1, 2)))
        self.any_N |= N is not None
        self.any_uv |= uv is not None
        attr = (2 if N is not None else 0) | (4 if uv is not None else 0)
        first_tri = sum(len(i) for i in self.idx)
        self.P.append(P)
        self.idx.append(idx + self.n_verts)
        self.n_verts += len(P)
        self.tri_material.append(np.full(nt, material, dtype=np.int32))
        self.tri_flip.append(np.full(nt, (1 if flip else 0) | attr, dtype=np.uint8))
        self.tri_medium.append(np.full(nt, medium_code, dtype=np.uint32))
        tl = np.full(nt, -1, dtype=np.int32)
        if light_params is not None:
            # one DiffuseAreaLight per triangle (base/light.cu:56-75, lights/diffuse_area_light.cu:9-31)
            L, sc, two = self.area_light_params(light_params, scene_dir)
            for k in range(nt):
                tl[k] = len(self.lights)
                self.lights.append({"triangle": first_tri + k, "spectrum": L, "scale": float(sc), "two_sided": int(two)})
        self.tri_light.append(tl)

    # ---- object instancing (two-level: one BVH8 per object, a transformed ray per instance)
    def add_instance(self, name, mesh_recs, rfi, scene_dir, quadric_recs=()):
        """ObjectInstance of an object's MESHES (scene_builder.cu:845-876, primitives/transformed_primitive.cu): the object's
        triangles are stored once, in the object's own space, at the end of the triangle array (finalize_objects); each
        instance is a transform.  `mesh_recs` are the shape records of the object definition."""
        if not hasattr(self, "objects"):
            self.objects, self.object_index, self.instances = [], {}, []
        if name not in self.object_index:
            self.object_index[name] = len(self.objects)
            self.objects.append({"name": name, "meshes": mesh_recs, "quadrics": list(quadric_recs)})
        self.instances.append({"object": self.object_index[name], "rfi": rfi})

    def finalize_objects(self, scene_dir):
        """append the instanced objects' triangles after the top-level ones and record their ranges"""
        for ob in getattr(self, "objects", []):
            first = sum(len(i) for i in self.idx)
            for rec in ob["meshes"]:
                Pr = rec["P"] if rec["rfo"].is_identity() else xf_points(rec["rfo"].m, rec["P"])
                Nr = rec.get("N")
                if Nr is not None and (not rec["rfo"].is_identity() or rec["ro"]):
                    Nr = xf_normals(rec["rfo"].inv, Nr)
                    if rec["ro"]:
                        Nr = (-Nr).astype(F32)
                flip = rec["ro"] ^ swaps_handedness(rec["rfo"].m)
                self.add_mesh(Pr, rec["idx"], rec["material"], flip, None, scene_dir, N=Nr, uv=rec.get("uv"), medium_code=0)
            ob["first"], ob["count"] = first, sum(len(i) for i in self.idx) - first
            # the object's own bounds, as its HLBVH in the reference has them: triangle vertices and the quadrics' boxes through
            # Transform::operator()(Bounds3f) (euclidean_space/transform.h:326-351), fp32
            lo, hi = np.full(3, np.inf, dtype=F32), np.full(3, -np.inf, dtype=F32)
            tris = np.concatenate(self.idx)[first:first + ob["count"]] if ob["count"] else np.zeros((0, 3), np.int32)
            if len(tris):
                Pall = np.concatenate(self.P)
                v = Pall[np.unique(tris)]
                lo, hi = np.minimum(lo, v.min(axis=0)), np.maximum(hi, v.max(axis=0))
            for q in ob.get("quadrics", ()):
                r = F32(q["radius"])
                if q["kind"] == "disk":
                    bmin, bmax = np.array([-r, -r, q["zmin"]], F32), np.array([r, r, q["zmin"]], F32)
                else:
                    zmin = min(max(min(F32(q["zmin"]), F32(q["zmax"])), -r), r)
                    zmax = min(max(max(F32(q["zmin"]), F32(q["zmax"])), -r), r)
                    bmin, bmax = np.array([-r, -r, zmin], F32), np.array([r, r, zmax], F32)
                m = q["rfo"].m
                qlo, qhi = m[:3, 3].astype(F32).copy(), m[:3, 3].astype(F32).copy()
                for i in range(3):
                    for k in range(3):
                        a, b = F32(m[i, k] * bmin[k]), F32(m[i, k] * bmax[k])
                        qlo[i] = F32(qlo[i] + min(a, b))
                        qhi[i] = F32(qhi[i] + max(a, b))
                lo, hi = np.minimum(lo, qlo), np.maximum(hi, qhi)
            ob["bounds"] = [float(x) for x in lo] + [float(x) for x in hi]

    def area_light_params(self, light_params, scene_dir):
        """DiffuseAreaLight ctor (lights/diffuse_area_light.cu:9-31): Lemit (default: the colour space's illuminant), scale
        divided by the photometric integral of Lemit, twosided"""
        L = self.spec_param(light_params, "L", "illuminant", scene_dir)
        if L is None:
            lam, val, _ = spectral.illuminant_d65()
            L = self.spec_piecewise(("illuminant",), lam, val)
        sc = F32(p_float(light_params, "scale", 1.0))
        sc = F32(sc / self.photometric(L))
        return L, sc, p_bool(light_params, "twosided", False)

    def add_sphere_light(self, sphere_index, light_params, scene_dir):
        """one DiffuseAreaLight over the whole sphere (base/light.cu:56-75: Shape "sphere" is a single shape)"""
        L, sc, two = self.area_light_params(light_params, scene_dir)
        self.lights.append({"triangle": sphere_index, "spectrum": L, "scale": float(sc), "two_sided": int(two), "type": 5})
        return len(self.lights) - 1

    def photometric(self, spec_id):
        """Spectrum::to_photometric (base/spectrum.cu:115-120): RGB illuminants use the illuminant's integral."""
        rec = self.spectra[spec_id]
        y = spectral.cie_xyz()[1]
        grid = np.arange(360, 831)
        # fp32 running sums over 360..830 nm, as the reference's inner_product loops round them
        if rec["type"] == 3:
            return spectral.inner_product32(spectral.illuminant_d65()[2], y)
        if rec["type"] == 4:
            return spectral.inner_product32(spectral.piecewise_eval32(rec["lambda"], rec["value"], grid), y)
        if rec["type"] == 0:
            return spectral.inner_product32(np.full(471, rec["c"][0], dtype=F32), y)
        raise SceneError("to_photometric: unsupported spectrum type")

    # ---- to the C-ABI
    def to_desc(self):
        keep = self._keep = []
        d = api.SceneDesc()
        P = np.ascontiguousarray(np.concatenate(self.P) if self.P else np.zeros((0, 3), F32), dtype=F32)
        idx = np.ascontiguousarray(np.concatenate(self.idx) if self.idx else np.zeros((0, 3), np.int32), dtype=np.int32)
        tm = np.ascontiguousarray(np.concatenate(self.tri_material) if self.idx else np.zeros(0, np.int32))
        tl = np.ascontiguousarray(np.concatenate(self.tri_light) if self.idx else np.zeros(0, np.int32))
        tf = np.ascontiguousarray(np.concatenate(self.tri_flip) if self.idx else np.zeros(0, np.uint8))
        keep += [P, idx, tm, tl, tf]
        d.n_vertices, d.P = len(P), api._fp(P)
        d.n_triangles, d.indices = len(idx), api._ip(idx)
        d.tri_material, d.tri_light = api._ip(tm), api._ip(tl)
        d.tri_flip = tf.ctypes.data_as(api.C.POINTER(api.C.c_uint8))
        if self.any_N:
            Nn = np.ascontiguousarray(np.concatenate([np.zeros((n, 3), F32) if a is None else a for n, a in self.N]), dtype=F32)
            keep.append(Nn)
            d.N = api._fp(Nn)
        if self.any_uv:
            uvn = np.ascontiguousarray(np.concatenate([np.zeros((n, 2), F32) if a is None else a for n, a in self.uv]), dtype=F32)
            keep.append(uvn)
            d.uv = api._fp(uvn)
        C = api.C
        sph = (api.Sphere * max(1, len(self.spheres)))()
        for i, s in enumerate(self.spheres):
            sph[i].render_from_object[:] = [float(v) for v in s["rfo"].reshape(-1)]
            sph[i].object_from_render[:] = [float(v) for v in s["ofr"].reshape(-1)]
            sph[i].radius, sph[i].z_min, sph[i].z_max, sph[i].phi_max = s["radius"], s["zmin"], s["zmax"], s["phimax"]
            sph[i].reverse_orientation, sph[i].transform_swaps_handedness = int(s["ro"]), int(s["swaps"])
            sph[i].material, sph[i].light, sph[i].kind = s["material"], s.get("light", -1), s.get("kind", 0)
            sph[i].medium = (s.get("medium", 0) & 0xFFFFFFFF) if self.media else 0
            if s.get("light", -1) >= 0 and self.media:
                raise SceneError("emissive spheres in scenes with media are outside the hot path")
        d.n_spheres, d.spheres = len(self.spheres), sph
        spec = (api.Spectrum * max(1, len(self.spectra)))()
        for i, s in enumerate(self.spectra):
            spec[i].type = s["type"]
            spec[i].c[:] = [float(v) for v in s["c"]]
            if "value" in s:
                spec[i].n = len(s["value"])
                spec[i].value = api._fp(s["value"])
                keep.append(s["value"])
                if "lambda" in s:
                    spec[i].lambda_ = api._fp(s["lambda"])
                    keep.append(s["lambda"])
        d.n_spectra, d.spectra = len(self.spectra), spec
        mats = (api.Material * max(1, len(self.materials)))()
        for i, m in enumerate(self.materials):
            mats[i].type, mats[i].reflectance, mats[i].eta, mats[i].k = m["type"], m["reflectance"], m["eta"], m["k"]
            mats[i].u_roughness, mats[i].v_roughness, mats[i].remap_roughness = m["ur"], m["vr"], m["remap"]
            mats[i].transmittance, mats[i].scale = m.get("transmittance", -1), m.get("scale", 1.0)
            mats[i].thickness, mats[i].g, mats[i].albedo = m.get("thickness", 0.01), m.get("g", 0.0), m.get("albedo", -1)
            mats[i].max_depth, mats[i].n_samples = m.get("max_depth", 10), m.get("n_samples", 1)
            mats[i].c_u_roughness, mats[i].c_v_roughness = m.get("cur", 0.0), m.get("cvr", 0.0)
            mats[i].c_eta, mats[i].c_k = m.get("c_eta", -1), m.get("c_k", -1)
            mats[i].mix_a, mats[i].mix_b = m.get("mix_a", -1), m.get("mix_b", -1)
            mats[i].reflectance_tex1, mats[i].transmittance_tex1 = m.get("reflectance_tex1", 0), m.get("transmittance_tex1", 0)
            mats[i].u_roughness_tex1, mats[i].v_roughness_tex1 = m.get("ur_tex1", 0), m.get("vr_tex1", 0)
            mats[i].eta_tex1, mats[i].k_tex1 = m.get("eta_tex1", 0), m.get("k_tex1", 0)
        d.n_materials, d.materials = len(self.materials), mats
        lights = (api.Light * max(1, len(self.lights)))()
        for i, l in enumerate(self.lights):
            lights[i].triangle, lights[i].spectrum, lights[i].scale, lights[i].two_sided = l["triangle"], l["spectrum"], l["scale"], l["two_sided"]
            lights[i].type = l.get("type", 0)
            if lights[i].type == 5:
                continue                       # sphere area light: `triangle` is the sphere index
            if lights[i].type:
                lights[i].v[:] = l["v"]
                lights[i].light_from_render[:] = l["lfr"]
                lights[i].cos_falloff_start, lights[i].cos_falloff_end = l["cs"], l["ce"]
                if lights[i].type == 4:
                    lights[i].image = l["image"]
                    lights[i].render_from_light[:] = l["rfl"]
        d.n_lights, d.lights = len(self.lights), lights
        keep += [sph, spec, mats, lights]
        objs = [o for o in getattr(self, "objects", []) if o.get("count", 0) > 0]
        if objs:
            oa = (api.Object * len(objs))()
            for i, o in enumerate(objs):
                oa[i].first_triangle, oa[i].n_triangles = o["first"], o["count"]
                oa[i].bounds[:] = o.get("bounds", [1.0, 0, 0, 0.0, 0, 0])
            remap = {id(o): i for i, o in enumerate(objs)}
            live = [n for n in self.instances if id(self.objects[n["object"]]) in remap]
            ia = (api.Instance * max(1, len(live)))()
            for i, n in enumerate(live):
                ia[i].object = remap[id(self.objects[n["object"]])]
                ia[i].render_from_instance[:] = [float(v) for v in n["rfi"].m.reshape(-1)]
                ia[i].instance_from_render[:] = [float(v) for v in n["rfi"].inv.reshape(-1)]
            d.n_objects, d.objects, d.n_instances, d.instances = len(objs), oa, len(live), ia
            keep += [oa, ia]
        if self.media:
            med = (api.Medium * len(self.media))()
            for i, m in enumerate(self.media):
                med[i].sigma_a, med[i].sigma_s, med[i].scale, med[i].g = m["sigma_a"], m["sigma_s"], m["scale"], m["g"]
            tmed = np.ascontiguousarray(np.concatenate(self.tri_medium) if self.idx else np.zeros(0, np.uint32), dtype=np.uint32)
            d.n_media, d.media = len(self.media), med
            d.tri_medium = tmed.ctypes.data_as(api.C.POINTER(api.C.c_uint32))
            keep += [med, tmed]
        if self.images:
            imgs = (api.Image * len(self.images))()
            for i, im in enumerate(self.images):
                imgs[i].height, imgs[i].width = im.shape[0], im.shape[1]
                imgs[i].rgb = api._fp(im)
            rgb_tbl, white, lam01 = spectral._fit_tables()
            fit = np.ascontiguousarray(np.concatenate([np.asarray(rgb_tbl, np.float64).ravel(), np.asarray(white, np.float64).ravel(),
                                                       np.asarray(lam01, np.float64).ravel()]))
            d.n_images, d.images = len(self.images), imgs
            d.rgb2spec_fit = fit.ctypes.data_as(api.C.POINTER(api.C.c_double))
            keep += [imgs, fit]
        if self.textures:
            tex = (api.Texture * len(self.textures))()
            for i, t in enumerate(self.textures):
                tex[i].type, tex[i].a, tex[i].b, tex[i].c = t["type"], t["a"], t["b"], t["c"]
                tex[i].f[:] = t["f"]
                tex[i].map[:] = t["map"]
                tex[i].mapping = t["mapping"]
                tex[i].texture_from_render[:] = t["tfr"]
                tex[i].v1[:], tex[i].v2[:] = t["v1"], t["v2"]
            d.n_textures, d.textures = len(self.textures), tex
            keep.append(tex)
        if self.camera is not None:
            cam = api.Camera()
            cam.medium = self.camera_medium
            cam.camera_from_raster[:] = [float(v) for v in self.camera["cfr"].reshape(-1)]
            cam.render_from_camera[:] = [float(v) for v in self.camera["rfc"].reshape(-1)]
            cam.lens_radius, cam.focal_distance = self.camera["lens_radius"], self.camera["focal_distance"]
            flt = api.Filter()
            flt.radius_x, flt.radius_y, flt.nx, flt.ny, ftab = self.filter
            flt.f = api._fp(ftab)
            film = api.FilmDesc()
            film.width, film.height = self.film["width"], self.film["height"]
            cx, cy, cz = spectral.cie_xyz()
            ill = spectral.illuminant_d65()[2]
            film.cie_x, film.cie_y, film.cie_z, film.illuminant = api._fp(cx), api._fp(cy), api._fp(cz), api._fp(ill)
            film.imaging_ratio = 1.0
            M = spectral.srgb_color_space()["rgb_from_xyz"]
            film.output_rgb_from_sensor_rgb[:] = [float(v) for v in M.reshape(-1)]
            d.camera, d.filter, d.film = C.pointer(cam), C.pointer(flt), C.pointer(film)
            keep += [cam, flt, film, ftab, cx, cy, cz, ill]
        return d


# ------------------------------------------------------------------------------------ parsing
class _GState:
    def __init__(self):
        self.ctm = identity()
        self.reverse_orientation = False
        self.material = 0
        self.area_light = None
        self.medium_interface = None   # (interior id, exterior id), -1 = vacuum

    def copy(self):
        g = _GState()
        g.ctm, g.reverse_orientation, g.material, g.area_light = self.ctm.copy(), self.reverse_orientation, self.material, self.area_light
        g.medium_interface = self.medium_interface
        return g


def load(path=None, text=None, spp=None, integrator=None, scene_dir=None, flatten_instances=False):
    """SceneBuilder::parse_file + preprocess (scene/scene_builder.cu:741-966) -> Scene.  `scene_dir`: where relative file names
    (PLY meshes, spectra, images) of a scene given as text are looked up.  `flatten_instances`: copy every ObjectInstance's
    triangles into the flat render-space scene instead of the two-level form (one BVH8 per object + a transform per instance)."""
    if text is None:
        text = open(path).read()
    if scene_dir is None:
        scene_dir = os.path.dirname(os.path.abspath(path)) if path else os.getcwd()
    sc = Scene()
    g, stack = _GState(), []
    named_cs, named_mat = {}, {}
    pre = {"Camera": None, "Film": None, "PixelFilter": None, "Sampler": None, "Integrator": None}
    cam_ctm = None
    render_from_world = identity()
    sc.add_material("diffuse", {}, scene_dir)   # default material = diffuse 0.5 (scene_builder.cu:140-141)

    def directives(txt, base):
        for kw, args in lex(txt):
            if kw == "Include":
                inc = os.path.join(base, args[0])
                yield from directives(open(inc).read(), os.path.dirname(inc))
            else:
                yield kw, args

    def world_begin():
        nonlocal render_from_world
        # build_filter, build_film, build_camera (scene_builder.cu:778-781)
        fa = pre["PixelFilter"]
        sc.filter = filter_table(fa[0] if fa else "mitchell", params_of(fa[1:]) if fa else {})
        fp = params_of(pre["Film"][1:]) if pre["Film"] else {}
        sc.film = {"width": p_int(fp, "xresolution", 1280), "height": p_int(fp, "yresolution", 720)}
        sc.filename = os.path.splitext(p_str(fp, "filename", "pbrt.png"))[0] + ".png"
        if pre["Camera"] is None or pre["Camera"][0] != "perspective":
            raise SceneError("Camera 'perspective' is required")
        cp = params_of(pre["Camera"][1:])
        world_from_camera = cam_ctm.inverse()
        named_cs["camera"] = world_from_camera
        pos = xf_points(world_from_camera.m, [[0, 0, 0]])[0]
        world_from_render = translate(*pos)            # CameraWorldCoordSystem (cameras/camera_base.h:34-39)
        render_from_world = world_from_render.inverse()
        rfc = (render_from_world * world_from_camera).m
        W, H = sc.film["width"], sc.film["height"]
        aspect = F32(W) / F32(H)
        if aspect > 1:
            smin, smax = (-aspect, -1.0), (aspect, 1.0)
        else:
            smin, smax = (-1.0, -1.0 / aspect), (1.0, 1.0 / aspect)
        ndc_from_screen = scale(1.0 / (smax[0] - smin[0]), 1.0 / (smax[1] - smin[1]), 1.0) * translate(-smin[0], -smax[1], 0.0)
        raster_from_ndc = scale(W, -H, 1.0)
        raster_from_screen = raster_from_ndc * ndc_from_screen
        screen_from_camera = perspective(p_float(cp, "fov", 90.0), 1e-2, 1000.0)
        # Transform::inverse() swaps (m, inv_m) (euclidean_space/transform.h:244-246): raster_from_screen's inverse is the product
        # of the analytic inverses of its scale / translate factors, the perspective matrix's the numeric SquareMatrix::inverse
        # (cameras/perspective.cu:34-40)
        cfr = (screen_from_camera.inverse() * raster_from_screen.inverse()).m
        sc.camera = {"cfr": cfr, "rfc": rfc, "lens_radius": p_float(cp, "lensradius", 0.0),
                     "focal_distance": p_float(cp, "focaldistance", 1e6)}
        if g.medium_interface is not None:      # build_camera (scene_builder.cu:166-171)
            sc.camera_medium = g.medium_interface[1]
        sp = pre["Sampler"]
        if sp:
            sc.sampler = sp[0] if sp[0] in ("independent", "stratified") else "stratified"
            sc.spp = p_int(params_of(sp[1:]), "pixelsamples", None)
        ip = pre["Integrator"]
        if ip:
            sc.integrator = ip[0]
            ipp = params_of(ip[1:])
            sc.max_depth = p_int(ipp, "maxdepth", 5)
            sc.regularize = p_bool(ipp, "regularize", False)

    instances, instance_def = {}, None
    named_medium = {}

    def place_shape(rec, render_from_instance, area_light):
        """One shape into the flat scene.  Instances (primitives/transformed_primitive.cu:20-34) are flattened: the reference
        keeps a per-instance BVH and transforms the ray; here the instance's triangles are transformed once, by the shape's
        render_from_object and then by render_from_instance, which gives the same surface up to fp rounding.  The orientation
        flag of a triangle uses the handedness of the composite transform (the normal of the transformed triangle flips with
        det, the reference's transformed normal does not); a sphere keeps the flag of its own transform, because
        Transform::operator()(SurfaceInteraction) carries n through the instance transform without a handedness flip."""
        rfo_x = rec["rfo"] if render_from_instance is None else render_from_instance * rec["rfo"]
        flip = rec["ro"] ^ swaps_handedness(rfo_x.m)
        if rec["kind"] == "mesh":
            Pr = rec["P"] if rec["rfo"].is_identity() else xf_points(rec["rfo"].m, rec["P"])   # shapes/triangle_mesh.cu:49-78
            Nr = rec.get("N")
            if Nr is not None and (not rec["rfo"].is_identity() or rec["ro"]):
                # gpu_transform<Normal3f> (shapes/triangle_mesh.cu:6-18,66-77): Transform::operator()(Normal3f) = inv_m^T n,
                # not renormalised, negated under ReverseOrientation
                Nr = xf_normals(rec["rfo"].inv, Nr)
                if rec["ro"]:
                    Nr = (-Nr).astype(F32)
            if render_from_instance is not None:
                Pr = xf_points(render_from_instance.m, Pr)
                if Nr is not None:
                    Nr = xf_normals(render_from_instance.inv, Nr)
            sc.add_mesh(Pr, rec["idx"], rec["material"], flip, area_light, scene_dir, N=Nr, uv=rec.get("uv"), medium_code=rec.get("mcode", 0))
        else:
            if area_light is not None and rec["kind"] == "disk":
                raise SceneError("emissive disks are outside the hot path")
            sc.spheres.append({"rfo": rfo_x.m, "ofr": rfo_x.inv, "radius": rec["radius"], "zmin": rec["zmin"], "zmax": rec["zmax"],
                               "phimax": rec["phimax"], "ro": rec["ro"], "swaps": swaps_handedness(rec["rfo"].m), "material": rec["material"],
                               "kind": 1 if rec["kind"] == "disk" else 0, "medium": int(rec.get("mcode", 0))})
            if area_light is not None:
                sc.spheres[-1]["light"] = sc.add_sphere_light(len(sc.spheres) - 1, area_light, scene_dir)

    for kw, args in directives(text, scene_dir):
        if kw in pre:
            if kw == "Camera":
                cam_ctm = g.ctm.copy()
            pre[kw] = args
        elif kw == "WorldBegin":
            world_begin()
            g.ctm = identity()
            named_cs["world"] = g.ctm.copy()
        elif kw == "AttributeBegin":
            stack.append(g.copy())
        elif kw == "AttributeEnd":
            g = stack.pop()
        elif kw == "Identity":
            g.ctm = identity()
        elif kw == "Translate":
            g.ctm = g.ctm * translate(*args[:3])
        elif kw == "Scale":
            g.ctm = g.ctm * scale(*args[:3])
        elif kw == "Rotate":
            g.ctm = g.ctm * rotate(args[0], args[1:4])
        elif kw == "LookAt":
            g.ctm = g.ctm * lookat(args[0:3], args[3:6], args[6:9])
        elif kw == "Transform":
            g.ctm = Xf(np.array(args[0], dtype=F32).reshape(4, 4).T)
        elif kw == "ConcatTransform":
            g.ctm = g.ctm * Xf(np.array(args[0], dtype=F32).reshape(4, 4).T)
        elif kw == "CoordSysTransform":
            if args[0] not in named_cs:   # fatal in the reference too (scene/scene_builder.cu:309-314)
                raise SceneError("CoordSysTransform: unknown coordinate system %r" % (args[0],))
            g.ctm = named_cs[args[0]].copy()
        elif kw == "ReverseOrientation":
            g.reverse_orientation = not g.reverse_orientation
        elif kw == "Material":
            g.material = sc.add_material(args[0], params_of(args[1:]), scene_dir, named_mat)
        elif kw == "Texture":
            if len(args) < 3:
                raise SceneError("Texture needs a name, a colour type and a class")
            sc.add_texture(str(args[0]), str(args[1]), str(args[2]), params_of(args[3:]), render_from_world * g.ctm, scene_dir)
        elif kw == "MakeNamedMaterial":
            p = params_of(args[1:])
            named_mat[args[0]] = sc.add_material(p_str(p, "type", "diffuse"), p, scene_dir, named_mat)
        elif kw == "NamedMaterial":
            if args[0] not in named_mat:
                raise SceneError("NamedMaterial '%s' not defined" % args[0])
            g.material = named_mat[args[0]]
        elif kw == "AreaLightSource":
            if args[0] != "diffuse":
                raise SceneError("AreaLightSource: only 'diffuse' (scene_builder.cu:604-616)")
            g.area_light = params_of(args[1:])
        elif kw == "Shape":
            p = params_of(args[1:])
            rfo_x = render_from_world * g.ctm
            if args[0] in ("trianglemesh", "plymesh"):
                if args[0] == "plymesh":
                    P, idx, N, uv = read_ply(os.path.join(scene_dir, p_str(p, "filename", "")))
                else:
                    P = np.array(p["P"][1], dtype=F32).reshape(-1, 3)
                    idx = np.array(p["indices"][1], dtype=np.int32).reshape(-1, 3) if "indices" in p else np.array([[0, 1, 2]], np.int32)
                    N = np.array(p["N"][1], dtype=F32).reshape(-1, 3) if "N" in p else None
                    uv = np.array(p["uv"][1], dtype=F32).reshape(-1, 2) if "uv" in p else None
                if N is not None and len(N) != len(P):
                    N = None
                if uv is not None and len(uv) != len(P):
                    uv = None
                rec = {"kind": "mesh", "P": P, "idx": idx, "N": N, "uv": uv}
            elif args[0] == "loopsubdiv":        # base/shape.cu:65-74: the limit mesh enters as a triangle mesh with normals
                from . import loop_subdiv
                if "indices" not in p or "P" not in p:
                    raise SceneError("loopsubdiv needs indices and P")
                levels = int(p["levels"][1][0]) if "levels" in p else 3
                try:
                    P, idx, N = loop_subdiv.loop_subdivide(levels, p["indices"][1], p["P"][1])
                except ValueError as ex:
                    raise SceneError(str(ex))
                rec = {"kind": "mesh", "P": P, "idx": idx, "N": N, "uv": None}
            elif args[0] == "sphere":
                r = p_float(p, "radius", 1.0)
                rec = {"kind": "sphere", "radius": r, "zmin": p_float(p, "zmin", -r), "zmax": p_float(p, "zmax", r),
                       "phimax": math.radians(min(max(p_float(p, "phimax", 360.0), 0.0), 360.0))}
            elif args[0] == "disk":              # Disk::Disk (shapes/disk.cu:7-20); rides in the sphere record: zmin = height, zmax = inner radius
                rec = {"kind": "disk", "radius": p_float(p, "radius", 1.0), "zmin": p_float(p, "height", 0.0),
                       "zmax": p_float(p, "innerradius", 0.0),
                       "phimax": math.radians(min(max(p_float(p, "phimax", 360.0), 0.0), 360.0))}
            else:
                raise SceneError("Shape '%s' is not a reference shape (base/shape.cu:9-79)" % args[0])
            # parse_shape (scene_builder.cu:631-676): SimplePrimitive unless the shape is emissive or sits on a medium transition
            mi = g.medium_interface
            transition = mi is not None and mi[0] != mi[1]
            if g.area_light is None and not transition:
                mcode = 0
            elif transition:
                mcode = (mi[0] + 1) | ((mi[1] + 1) << 16)
            else:
                mcode = 0xFFFFFFFF
            rec.update({"rfo": rfo_x, "ro": g.reverse_orientation, "material": g.material, "mcode": mcode})
            if instance_def is not None:
                if g.area_light is not None:   # scene_builder.cu:653-656
                    raise SceneError("area lights are not supported with object instancing")
                instance_def["shapes"].append(rec)
            else:
                place_shape(rec, None, g.area_light)
        elif kw == "ObjectBegin":              # scene_builder.cu:809-823
            if instance_def is not None:
                raise SceneError("ObjectBegin called inside of instance definition")
            stack.append(g.copy())
            instance_def = {"name": args[0], "shapes": []}
        elif kw == "ObjectEnd":
            if instance_def is None:
                raise SceneError("ObjectEnd called before an instance defined")
            instances[instance_def["name"]] = instance_def["shapes"]
            instance_def = None
            g = stack.pop()
        elif kw == "ObjectInstance":           # scene_builder.cu:845-876
            if args[0] not in instances:
                raise SceneError("ObjectInstance '%s' not found" % args[0])
            rfi = (render_from_world * g.ctm) * render_from_world.inverse()
            meshes = [rec for rec in instances[args[0]] if rec["kind"] == "mesh"]
            if meshes and not flatten_instances:
                sc.add_instance(args[0], meshes, rfi, scene_dir,      # two-level: the object's tree is shared by its instances
                                quadric_recs=[rec for rec in instances[args[0]] if rec["kind"] != "mesh"])
            for rec in instances[args[0]]:
                if rec["kind"] != "mesh" or flatten_instances:       # quadrics ride in the flat list under the composite transform
                    place_shape(rec, None if rfi.is_identity() else rfi, None)
        elif kw == "LightSource":
            sc.add_light_source(args[0], params_of(args[1:]), render_from_world * g.ctm, scene_dir)
        elif kw == "MakeNamedMedium":           # scene_builder.cu:362-365
            named_medium[args[0]] = sc.add_medium(params_of(args[1:]), scene_dir)
        elif kw == "MediumInterface":           # scene_builder.cu:434-464
            names = [a for a in args if isinstance(a, str)]
            if len(names) != 2:
                raise SceneError("MediumInterface takes two names")
            for nm in names:
                if nm and nm not in named_medium:
                    raise SceneError("medium '%s' not defined" % nm)
            g.medium_interface = tuple(named_medium[nm] if nm else -1 for nm in names)
        elif kw in ("Texture",):
            raise SceneError("directive '%s' is outside the hot path (SURVEY.md 8f)" % kw)
        elif kw in ("ColorSpace", "Option", "Attribute"):
            continue
        else:
            raise SceneError("unknown directive '%s'" % kw)   # scene_builder.cu:467-468
    if sc.camera is None:
        raise SceneError("no WorldBegin")
    sc.finalize_objects(scene_dir)
    # spp: CLI --spp > Sampler "integer pixelsamples" > 4 (scene_builder.cu:249-251,402-404)
    sc.spp = spp if spp is not None else (sc.spp if sc.spp is not None else 4)
    if integrator is not None:
        sc.integrator = integrator
    if sc.integrator == "volpath":
        sc.integrator = "path"
    return sc
