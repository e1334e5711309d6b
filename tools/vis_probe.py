#!/usr/bin/env python3
"""Which of the reference's three visibility rays (base/integrator_base.cu:12-14,24-33) blocks the NEE samples that make our film
brighter than the reference's in a few pixels of the smooth-shaded scenes?  Renders with the debug library
(tools/build_debug_lib.py: prints the NEE record of chosen pixels), rebuilds the three rays of every printed sample on the host
and traces them, and our single ray, through the library's any-hit entry point.  GPU box only."""
import importlib, os, re, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
F = np.float32
CASES = [("loopsubdiv_96x64_64spp", 42 * 96 + 49), ("normals_96x64_64spp", 59 * 96 + 80)]
CHILD = r'''
import importlib, sys, tempfile
sys.path.insert(0, %r); sys.path.insert(0, %r)
pkg = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T
m, ref = T.golden(%r)
T.render(pkg, T.scene_text(m["scene"], None, None, [tuple(e) for e in m.get("edits", [])]), m["spp"], tempfile.mkdtemp())
'''


def next_up(v):
    return np.nextafter(F(v), F(np.inf)) if v > 0 or v == 0 else np.nextafter(F(v), F(np.inf))


def offset_ray_origin(p, err, n, w):   # base/ray.h:30-49
    p, err, n, w = (np.asarray(a, dtype=F) for a in (p, err, n, w))
    d = F(np.dot(np.abs(n), err))
    off = (d * n).astype(F)
    if np.dot(w.astype(np.float64), n.astype(np.float64)) < 0:
        off = -off
    po = (p + off).astype(F)
    for i in range(3):
        if off[i] > 0:
            po[i] = np.nextafter(po[i], F(np.inf))
        elif off[i] < 0:
            po[i] = np.nextafter(po[i], F(-np.inf))
    return po


def main():
    pkg = importlib.import_module("pbrt-minus_b200")
    import test_render_gpu as T
    for name, pixel in CASES:
        env = dict(os.environ, B200PT_LIB=os.path.join(ROOT, "pbrt-minus_b200", "libb200pt_dbg.so"))
        r = subprocess.run([sys.executable, "-c", CHILD % (ROOT, os.path.join(ROOT, "tests"), name)], capture_output=True, text=True, env=env, timeout=900)
        nee = [[float(x) for x in re.findall(r"-?(?:\d+\.?\d*(?:e[-+]?\d+)?|inf|nan)", l.split("pixel", 1)[1])] for l in r.stdout.splitlines() if l.startswith("[nee] pixel %d " % pixel)]
        emit = [l for l in r.stdout.splitlines() if l.startswith("[emit] pixel %d " % pixel)]
        print("== %s pixel %d: %d NEE records, %d emission records; child rc %d" % (name, pixel, len(nee), len(emit), r.returncode))
        for l in sorted(emit, key=lambda l: -max(float(x) for x in l.split("add")[1].split())):
            print("  ", l)
        if not nee:
            print(r.stdout[-500:], r.stderr[-500:]); continue
        if "ObjectInstance" in T.scene_text(T.golden(name)[0]["scene"]):
            for v in sorted(nee, key=lambda v: -max(v[4:8]))[:6]:
                print("   nee sample %d contrib %.4g (instanced scene: rays not traced through b200pt_trace)" % (int(v[1]), max(v[4:8])))
            continue
        m, ref = T.golden(name)
        text = T.scene_text(m["scene"], None, None, [tuple(e) for e in m.get("edits", [])])
        p = os.path.join(tempfile.mkdtemp(), "s.pbrt"); open(p, "w").write(text)
        sc = pkg.pbrt_scene.load(p, spp=m["spp"], scene_dir=os.path.join(ROOT, "tests", "scenes"))
        integ = pkg.WavefrontPathIntegrator(sc, device=0)
        acc = integ.accel
        nee.sort(key=lambda v: -max(v[4:8]))
        for v in nee[:10]:
            pix, sample, depth, mat = (int(x) for x in v[:4])
            contrib, P, E, N, LP, LE, LN = v[4:8], v[8:11], v[11:14], v[14:17], v[17:20], v[20:23], v[23:26]
            P, E, N, LP, LE, LN = (np.asarray(a, dtype=F) for a in (P, E, N, LP, LE, LN))
            pf = offset_ray_origin(P, E, N, LP - P); pt = offset_ray_origin(LP, LE, LN, pf - LP)
            d1 = (pt - pf).astype(F)
            pf2 = offset_ray_origin(LP, LE, LN, P - LP); pt2 = offset_ray_origin(P, E, N, pf2 - P)
            d2 = (pt2 - pf2).astype(F)
            d3 = (d1 / F(np.sqrt(F(F(d1[0] * d1[0] + d1[1] * d1[1]) + d1[2] * d1[2])))).astype(F)
            t3 = F((1.0 - 1e-4) * float(np.linalg.norm((LP - pf).astype(np.float64))))
            o = np.stack([pf, pf2, pf, pf]); d = np.stack([d1, d2, d3, d1]); tm = np.array([0.6, 0.6, t3, 1.0], dtype=F)
            occ = acc.fast_intersect(o, d, tm)
            print("  sample %2d depth %d mat %d contrib %-10.4g ray1[0,.6] %d  ray2[0,.6] %d  ray3 normalised %d | ours (0,1] %d   dist %.4g" % (
                sample, depth, mat, max(contrib), int(occ[0]), int(occ[1]), int(occ[2]), int(occ[3]), float(np.linalg.norm(d1))))
        integ.close()


if __name__ == "__main__":
    main()
