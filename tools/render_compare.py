#!/usr/bin/env python3
"""Render a scene with libb200pt and (when oracle/_ref/pbrt-minus-ref is present) with the
reference's own CUDA build, and compare the linear float films.  GPU box only.

  python tools/render_compare.py [--scene PATH] [--res W H] [--spp N] [--out DIR] [--no-ref]
"""
import argparse
import importlib
import json
import os
import re
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CORNELL = os.path.join(ROOT, "tests", "scenes", "cornell-box-specular.pbrt")


def with_resolution(text, w, h):
    text = re.sub(r'("integer xresolution"\s*\[\s*)\d+', r"\g<1>%d" % w, text)
    return re.sub(r'("integer yresolution"\s*\[\s*)\d+', r"\g<1>%d" % h, text)


def rel_mse(a, b):
    return float(np.mean((a - b) ** 2 / (b ** 2 + 1e-4)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=CORNELL)
    ap.add_argument("--res", type=int, nargs=2, default=None)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "render"))
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--integrator", default=None)
    ap.add_argument("--tag", default="cmp")
    args = ap.parse_args()
    args.out = os.path.abspath(args.out)
    os.makedirs(args.out, exist_ok=True)
    pkg = importlib.import_module("pbrt-minus_b200")
    text = open(args.scene).read()
    if args.res:
        text = with_resolution(text, *args.res)
    scene_path = os.path.join(args.out, args.tag + ".pbrt")
    open(scene_path, "w").write(text)

    t0 = time.time()
    sc = pkg.pbrt_scene.load(scene_path, spp=args.spp)
    t_load = time.time() - t0
    integ = pkg.WavefrontPathIntegrator(sc, device=0)
    film = integ.create_film()
    integ.render(film)          # warm-up (allocations, module load)
    film.clear()
    st = integ.render(film)
    ours = film.get_pixel_rgb()
    W, H = film.width, film.height
    ours.tofile(os.path.join(args.out, args.tag + "_ours.f32"))
    film.write_to_png(os.path.join(args.out, args.tag + "_ours.png"))
    res = {"scene": os.path.basename(args.scene), "width": W, "height": H, "spp": args.spp, "load_s": round(t_load, 3),
           "ours": {"ms": st["ms_total"], "Msamples_s": W * H * args.spp / st["ms_total"] / 1e3,
                    "Mrays_s": (st["rays_closest"] + st["rays_shadow"]) / st["ms_total"] / 1e3, "stats": st,
                    "mean_rgb": [float(v) for v in ours.reshape(-1, 3).mean(0)], "nan": int(np.isnan(ours).sum())}}
    exe = os.path.join(ROOT, "oracle", "_ref", "pbrt-minus-ref")
    if not args.no_ref and os.path.exists(exe):
        out_png = os.path.join(args.out, args.tag + "_ref.png")
        cmd = [exe, scene_path, "--spp", str(args.spp), "--outfile", out_png]
        if args.integrator:
            cmd += ["--integrator", args.integrator]
        t0 = time.time()
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=3000, cwd=args.out)
        wall = time.time() - t0
        open(os.path.join(args.out, args.tag + "_ref.log"), "w").write(r.stdout + r.stderr)
        dump = out_png + ".f32"
        if r.returncode == 0 and os.path.exists(dump):
            ref = np.fromfile(dump, dtype=np.float32).reshape(H, W, 3)
            m = re.search(r"rendering took\s+([0-9.]+)", r.stdout)
            ref_s = float(m.group(1)) if m else None
            d = ours - ref
            lum = ref.mean(axis=2)
            res["reference"] = {"render_s": ref_s, "wall_s": round(wall, 2),
                                "Msamples_s": (W * H * args.spp / ref_s / 1e6) if ref_s else None,
                                "mean_rgb": [float(v) for v in ref.reshape(-1, 3).mean(0)]}
            res["compare"] = {"relMSE": rel_mse(ours, ref), "mean_abs_diff": float(np.abs(d).mean()),
                              "mean_rel_diff_of_means": float(abs(ours.mean() - ref.mean()) / ref.mean()),
                              "p50_abs": float(np.percentile(np.abs(d), 50)), "p99_abs": float(np.percentile(np.abs(d), 99)),
                              "max_abs": float(np.abs(d).max()), "frac_pixels_gt_1pct": float((np.abs(d).max(axis=2) > 0.01 * (lum + 1e-3)).mean())}
            if ref_s:
                res["speedup_samples_per_s"] = res["ours"]["Msamples_s"] / res["reference"]["Msamples_s"]
        else:
            res["reference"] = {"error": (r.stdout + r.stderr)[-500:]}
    print(json.dumps(res))
    open(os.path.join(args.out, args.tag + ".json"), "w").write(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
