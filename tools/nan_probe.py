#!/usr/bin/env python3
"""Count non-finite pixels of a render, ours and (optionally) the reference's, on the same scene text.
  python tools/nan_probe.py --scene tests/scenes/media2.pbrt --res 960 640 --spp 64 [--reference] [--lib PATH]"""
import argparse
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def bad_pixels(film):
    bad = ~np.isfinite(film).all(axis=2)
    ys, xs = np.nonzero(bad)
    return int(bad.sum()), [(int(x), int(y)) for x, y in zip(xs[:12], ys[:12])], bad


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", required=True)
    ap.add_argument("--res", type=int, nargs=2, default=None)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--reference", action="store_true")
    ap.add_argument("--lib", default="")
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    if args.lib:
        pkg.api.LIB_PATH = os.path.abspath(args.lib)
    text = open(args.scene).read()
    if args.res:
        text = re.sub(r'("integer xresolution"\s*\[\s*)\d+', r"\g<1>%d" % args.res[0], text)
        text = re.sub(r'("integer yresolution"\s*\[\s*)\d+', r"\g<1>%d" % args.res[1], text)
    scene_dir = os.path.dirname(os.path.abspath(args.scene))
    sc = pkg.pbrt_scene.load(text=text, spp=args.spp, scene_dir=scene_dir)
    integ = pkg.WavefrontPathIntegrator(sc, device=0)
    film = integ.create_film()
    per_sample = []
    for k in range(args.spp):          # one sample index at a time: which sample brings the first non-finite value
        integ.render(film, sample_begin=k, sample_end=k + 1)
        n, _, _ = bad_pixels(film.get_pixel_rgb())
        per_sample.append(n)
    ours = film.get_pixel_rgb()
    n, where, bad_o = bad_pixels(ours)
    out = {"scene": os.path.basename(args.scene), "res": args.res, "spp": args.spp, "ours_bad_pixels": n, "ours_first": where,
           "ours_bad_after_sample": per_sample, "ours_mean_finite": float(np.nanmean(np.where(np.isfinite(ours), ours, np.nan)))}
    film.close()
    integ.close()
    if args.reference:
        exe = os.path.join(ROOT, "oracle", "_ref", "pbrt-minus-ref")
        tmp = tempfile.mkdtemp()
        for f in os.listdir(scene_dir):        # relative file names of the scene
            src = os.path.join(scene_dir, f)
            if not os.path.exists(os.path.join(tmp, f)):
                os.symlink(src, os.path.join(tmp, f))
        sp = os.path.join(tmp, "probe.pbrt")
        open(sp, "w").write(text)
        png = os.path.join(tmp, "probe.png")
        integ_name = "volpath" if "volpath" in text else None
        r = subprocess.run([exe, sp, "--spp", str(args.spp), "--outfile", png], capture_output=True, text=True, cwd=tmp, timeout=3000)
        out["reference_rc"] = r.returncode
        out["reference_tail"] = (r.stdout + r.stderr)[-300:]
        if os.path.exists(png + ".f32"):
            H, W = ours.shape[:2]
            ref = np.fromfile(png + ".f32", dtype=np.float32).reshape(H, W, 3)
            m, wh, bad_r = bad_pixels(ref)
            out.update({"reference_bad_pixels": m, "reference_first": wh, "bad_in_both": int((bad_o & bad_r).sum()),
                        "reference_mean_finite": float(np.nanmean(np.where(np.isfinite(ref), ref, np.nan)))})
    print(json.dumps(out))


if __name__ == "__main__":
    main()
