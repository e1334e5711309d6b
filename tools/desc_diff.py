#!/usr/bin/env python3
"""Compare two scene-description dumps (b200pt_debug_dump_desc / B200PT_DESC_DUMP) field by field, bit by bit.

The usual pair: the reference's own front-end through the bridge (dumped on the GPU box) and the Python front-end (dumped
anywhere: the dump makes no CUDA call).  Spectrum, material and texture ids are compared by CONTENT (the two front-ends number
them in their own order).

  python tools/desc_diff.py --python-scene tests/scenes/normals.pbrt --spp 64 --ref gpurun_out/desc/normals_96x64_64spp
  python tools/desc_diff.py DIR_A DIR_B
"""
import argparse
import ctypes as C
import importlib
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("pbrt-minus_b200")
api = pkg.api


FILES = ("P.f32", "N.f32", "uv.f32", "indices.i32", "tri_material.i32", "tri_light.i32", "tri_flip.u8", "tri_medium.u32", "camera.bin", "filter_hdr.f32",
         "filter_f.f32", "film.f32", "images_dims.i32", "images_rgb.f32", "spectra.i32", "spectra_pool.f32", "spheres.bin", "materials.bin", "lights.bin",
         "textures.bin", "media.bin", "objects.bin", "instances.bin")


def pack(d, out_path):
    """a dump directory as one compressed .npz (the committed fixtures under tests/golden/desc/)"""
    z = {f.replace(".", "__"): np.fromfile(os.path.join(d, f), dtype=np.uint8) for f in FILES if os.path.exists(os.path.join(d, f))}
    sp = os.path.join(d, "scene.pbrt")
    if os.path.exists(sp):
        z["scene__pbrt"] = np.frombuffer(open(sp, "rb").read(), dtype=np.uint8)
    np.savez_compressed(out_path, **z)


def load(d):
    packed = dict(np.load(d)) if os.path.isfile(d) else None

    def arr(name, dt):
        if packed is not None:
            raw = packed.get(name.replace(".", "__"), np.zeros(0, np.uint8))
            return np.frombuffer(raw.tobytes(), dtype=dt).copy()
        p = os.path.join(d, name)
        return np.fromfile(p, dtype=dt) if os.path.exists(p) else np.zeros(0, dt)
    out = {"P": arr("P.f32", np.float32).reshape(-1, 3), "N": arr("N.f32", np.float32).reshape(-1, 3), "uv": arr("uv.f32", np.float32).reshape(-1, 2),
           "indices": arr("indices.i32", np.int32).reshape(-1, 3), "tri_material": arr("tri_material.i32", np.int32), "tri_light": arr("tri_light.i32", np.int32),
           "tri_flip": arr("tri_flip.u8", np.uint8), "tri_medium": arr("tri_medium.u32", np.uint32),
           "camera": arr("camera.bin", np.uint32), "filter_hdr": arr("filter_hdr.f32", np.float32), "filter_f": arr("filter_f.f32", np.float32),
           "film": arr("film.f32", np.float32), "images_dims": arr("images_dims.i32", np.int32), "images_rgb": arr("images_rgb.f32", np.float32),
           "spectra": arr("spectra.i32", np.int32).reshape(-1, 6), "pool": arr("spectra_pool.f32", np.float32)}
    for name, st in (("spheres", api.Sphere), ("materials", api.Material), ("lights", api.Light), ("textures", api.Texture), ("media", api.Medium),
                     ("objects", api.Object), ("instances", api.Instance)):
        raw = arr(name + ".bin", np.uint8)
        out[name] = [st.from_buffer_copy(raw[i * C.sizeof(st):(i + 1) * C.sizeof(st)].tobytes()) for i in range(len(raw) // C.sizeof(st))]
    # spectra by content
    specs, off = [], 0
    for t, c0, c1, c2, c3, n in out["spectra"]:
        c = np.array([c0, c1, c2, c3], np.int32).view(np.uint32)
        k = 2 * n if t == 4 else (n if t == 5 else 0)
        used = {0: c[:1], 1: c[:3], 2: c, 3: c}.get(int(t), c[:0])     # the words of c[] the type reads
        specs.append((int(t), tuple(int(x) for x in used), out["pool"][off:off + k].view(np.uint32).tobytes()))
        off += k
    out["spec_content"] = specs
    import hashlib
    imgs, off = [], 0
    for w, h in out["images_dims"].reshape(-1, 2):
        px = out["images_rgb"][off:off + int(w) * int(h) * 3]
        imgs.append((int(w), int(h), hashlib.sha1(px.tobytes()).hexdigest()))
        off += int(w) * int(h) * 3
    out["image_content"] = imgs
    return out


def fields(st):
    return [f[0] for f in st._fields_]


def bits(v):
    if isinstance(v, float):
        return int(np.float32(v).view(np.uint32))
    if hasattr(v, "__len__"):
        return tuple(bits(x) for x in v)
    return int(v)


SPEC_FIELDS = {"Material": ("reflectance", "eta", "k", "transmittance", "albedo", "c_eta", "c_k"), "Light": ("spectrum",), "Medium": ("sigma_a", "sigma_s")}
TEX_FIELDS = {"Material": ("reflectance_tex1", "transmittance_tex1", "u_roughness_tex1", "v_roughness_tex1", "eta_tex1", "k_tex1")}


def texture_content(d, i, depth=0):
    t = d["textures"][i]
    ty = t.type
    used = {0: (), 1: ("f", "map", "mapping", "texture_from_render", "v1", "v2"), 17: ("f", "map", "mapping", "texture_from_render", "v1", "v2"),
            2: ("f",), 3: ("f",) if t.c < 0 else (), 4: ("f",), 16: ("f",), 18: ()}.get(ty, ())
    rec = {"type": ty}
    for f in used:
        v = bits(getattr(t, f))
        if f == "f":
            v = v[:{1: 2, 17: 2, 2: 1, 3: 1, 4: 3, 16: 1}[ty]]
        if f in ("texture_from_render", "v1", "v2") and t.mapping == 0:
            continue                                      # uv mapping reads none of them
        if f in ("v1", "v2") and t.mapping != 1:
            continue
        rec[f] = v
    if ty == 0:
        rec["a"] = d["spec_content"][t.a]
    elif ty in (1, 17):
        rec["a"] = ("image", d["image_content"][t.a])      # by content: one front-end shares an image between textures, the other loads it per texture
    elif ty in (2,):
        rec["a"] = texture_content(d, t.a, depth + 1)
    elif ty in (3, 4, 18):
        rec["a"] = texture_content(d, t.a, depth + 1)
        rec["b"] = texture_content(d, t.b, depth + 1)
        if ty == 3:
            rec["c"] = texture_content(d, t.c, depth + 1) if t.c >= 0 else None
    return tuple(sorted((k, repr(v)) for k, v in rec.items()))


# the fields a material type reads (include/b200pt.h); a front-end may leave the others at any value
MATERIAL_FIELDS = {0: ("reflectance",), 1: ("reflectance", "eta", "k", "u_roughness", "v_roughness", "remap_roughness"),
                   2: ("eta", "u_roughness", "v_roughness", "remap_roughness"), 3: ("reflectance", "transmittance", "scale"),
                   4: ("reflectance", "albedo", "eta", "u_roughness", "v_roughness", "thickness", "g", "remap_roughness", "max_depth", "n_samples"),
                   5: ("reflectance", "albedo", "eta", "u_roughness", "v_roughness", "thickness", "g", "remap_roughness", "max_depth", "n_samples",
                       "c_eta", "c_k", "c_u_roughness", "c_v_roughness"),
                   6: ("scale", "mix_a", "mix_b"), 7: ()}
TEX_OF = {"reflectance": "reflectance_tex1", "transmittance": "transmittance_tex1", "u_roughness": "u_roughness_tex1", "v_roughness": "v_roughness_tex1",
          "eta": "eta_tex1", "k": "k_tex1", "c_eta": "eta_tex1", "c_k": "k_tex1"}


def material_content(d, r):
    """a material by the parameters its type reads; a parameter is its spectrum / value, or its texture graph -- a CONSTANT
    texture node counts as the spectrum it holds (one front-end may keep the node, the other fold it)"""
    out = {"type": r.type}
    for f in MATERIAL_FIELDS.get(r.type, ()):
        v = getattr(r, f)
        tex1 = getattr(r, TEX_OF[f]) if f in TEX_OF and not (r.type == 5 and f in ("eta", "k")) and not (r.type != 5 and f in ("c_eta", "c_k")) else 0
        if f in ("mix_a", "mix_b"):
            out[f] = material_content(d, d["materials"][v])
        elif tex1 > 0:
            t = d["textures"][tex1 - 1]
            if t.type == 0:
                out[f] = d["spec_content"][t.a]
            elif t.type == 16:
                out[f] = bits(t.f[0])
            else:
                out[f] = texture_content(d, tex1 - 1)
        elif f in SPEC_FIELDS["Material"]:
            out[f] = d["spec_content"][v] if v >= 0 else None
        else:
            out[f] = bits(v)
    return out


def record_content(d, r):
    name = type(r).__name__
    if name == "Material":
        return material_content(d, r)
    out = {}
    for f in fields(type(r)):
        if name == "Light" and f == "image" and r.type != 4:
            continue
        v = getattr(r, f)
        if f in SPEC_FIELDS.get(name, ()):
            out[f] = d["spec_content"][v] if v >= 0 else None
        elif f in TEX_FIELDS.get(name, ()):
            out[f] = texture_content(d, v - 1) if v > 0 else None
        elif name == "Material" and f in ("mix_a", "mix_b"):
            out[f] = record_content(d, d["materials"][v]) if v >= 0 and r.type == 6 else None
        else:
            out[f] = bits(v)
    return out


def cmp_arrays(name, a, b, report):
    if a.shape != b.shape:
        report.append("%-14s shapes differ: %s vs %s" % (name, a.shape, b.shape))
        return
    if a.size == 0:
        return
    ua, ub = a.view(np.uint32 if a.dtype.itemsize == 4 else a.dtype), b.view(np.uint32 if b.dtype.itemsize == 4 else b.dtype)
    bad = ua != ub
    if bad.any():
        idx = np.argwhere(bad)[:4]
        ex = ["%s: %r vs %r" % (tuple(int(x) for x in i), a[tuple(i)], b[tuple(i)]) for i in idx]
        if a.dtype == np.float32:
            ulp = np.abs(ua.astype(np.int64) - ub.astype(np.int64))[bad]
            report.append("%-14s %d of %d words differ (max %d ulp, median %d): %s" % (name, int(bad.sum()), a.size, int(ulp.max()), int(np.median(ulp)), "; ".join(ex)))
        else:
            report.append("%-14s %d of %d entries differ: %s" % (name, int(bad.sum()), a.size, "; ".join(ex)))


def diff(A, B):
    report = []
    # geometry: per triangle vertex positions / normals / uvs (the vertex arrays may be laid out differently)
    if A["indices"].shape != B["indices"].shape:
        report.append("triangle counts differ: %d vs %d" % (len(A["indices"]), len(B["indices"])))
    else:
        cmp_arrays("tri vertices", A["P"][A["indices"]], B["P"][B["indices"]], report)
        if len(A["N"]) and len(B["N"]):
            cmp_arrays("tri normals", A["N"][A["indices"]], B["N"][B["indices"]], report)
        elif len(A["N"]) != len(B["N"]):
            report.append("normals present in one description only")
        if len(A["uv"]) and len(B["uv"]):
            cmp_arrays("tri uvs", A["uv"][A["indices"]], B["uv"][B["indices"]], report)
        for k in ("tri_light", "tri_flip", "tri_medium"):
            cmp_arrays(k, A[k], B[k], report)
        if len(A["tri_material"]) == len(B["tri_material"]) and len(A["tri_material"]):
            ma = [record_content(A, A["materials"][i]) for i in range(len(A["materials"]))]
            mb = [record_content(B, B["materials"][i]) for i in range(len(B["materials"]))]
            seen = set()
            for t, (i, j) in enumerate(zip(A["tri_material"], B["tri_material"])):
                if (i, j) in seen:
                    continue
                seen.add((i, j))
                if ma[i] != mb[j]:
                    d = {k: (ma[i][k], mb[j][k]) for k in ma[i] if ma[i][k] != mb[j][k]}
                    report.append("material of triangle %d differs (ids %d / %d): %s" % (t, i, j, str(d)[:600]))
    for k in ("camera", "filter_hdr", "filter_f", "film"):
        cmp_arrays(k, A[k], B[k], report)
    if set(A["image_content"]) != set(B["image_content"]):
        report.append("images differ: %s vs %s" % (sorted(set(A["image_content"]) - set(B["image_content"])), sorted(set(B["image_content"]) - set(A["image_content"]))))
    for name in ("lights", "spheres", "media", "instances", "objects"):
        if len(A[name]) != len(B[name]):
            report.append("%s: %d vs %d" % (name, len(A[name]), len(B[name])))
            continue
        for i, (ra, rb) in enumerate(zip(A[name], B[name])):
            ca, cb = record_content(A, ra), record_content(B, rb)
            if name == "spheres" and not (len(A["media"]) or len(B["media"])):
                ca.pop("medium", None)
                cb.pop("medium", None)
            if name == "spheres":   # material ids by content
                ca["material"] = record_content(A, A["materials"][ra.material]) if ra.material >= 0 else None
                cb["material"] = record_content(B, B["materials"][rb.material]) if rb.material >= 0 else None
            if ca != cb:
                d = {k: (ca[k], cb[k]) for k in ca if ca[k] != cb[k]}
                report.append("%s[%d] differs: %s" % (name, i, str(d)[:700]))
                if len(report) > 40:
                    return report
    return report


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("dirs", nargs="*")
    ap.add_argument("--python-scene")
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--ref")
    ap.add_argument("--pack", nargs=2, metavar=("DIR", "NPZ"), help="write a dump directory as one compressed fixture")
    args = ap.parse_args()
    if args.pack:
        pack(*args.pack)
        return 0
    if args.python_scene:
        sc = pkg.pbrt_scene.load(args.python_scene, spp=args.spp, scene_dir=os.path.join(ROOT, "tests", "scenes"))
        tmp = tempfile.mkdtemp()
        desc = sc.to_desc()
        api._check(api.load_library().b200pt_debug_dump_desc(C.byref(desc), tmp.encode()))
        a, b = tmp, args.ref
    else:
        a, b = args.dirs
    rep = diff(load(a), load(b))
    print("\n".join(rep) if rep else "descriptions are bit-identical")
    return 1 if rep else 0


if __name__ == "__main__":
    sys.exit(main())
