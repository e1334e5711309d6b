#!/usr/bin/env python3
"""Statistical image gate (SURVEY.md 8d "image gates"): ours vs the reference's own CUDA build at the same spp,
for scenes where the two cannot trace identical paths (many lights: the reference's alias table is built with a
re-sort per step whose tie order is libstdc++'s introsort state, distribution/alias_table.cu:43-62, so the
light picked for a given random number differs while the estimator's expectation is the same).

ours is rendered in K disjoint sample ranges of the same N-spp stratified pattern; the K partial films give a
per-pixel variance of the N-spp mean (var_k / K).  Gates:
  (i)  relMSE(ours, ref) = mean_p (ours - ref)^2 / (ref^2 + 1e-4)  <=  tau = 2 * mean_p sigma_p^2 / (mu_p^2 + 1e-4) * slack
  (ii) fraction of pixels with |ours - ref| > 3 * sqrt(2) * sigma_p   (two-sided 3 sigma of a difference of two
       independent N-spp estimators; 0.27 % expected for Gaussian noise, more for heavy-tailed path-traced pixels)
GPU box only.

  python tools/render_stat_compare.py --scene /tmp/c4s/c4.pbrt --spp 1024 --parts 16 --out /tmp/c4s
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", required=True)
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--parts", type=int, default=16)
    ap.add_argument("--out", required=True)
    ap.add_argument("--tag", default="stat")
    ap.add_argument("--slack", type=float, default=1.5)
    ap.add_argument("--ref-film", default=None, help="reuse a reference float film instead of running the reference")
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    sc = pkg.pbrt_scene.load(args.scene, spp=args.spp)
    integ = pkg.WavefrontPathIntegrator(sc, device=0)
    film = integ.create_film()
    W, H = film.width, film.height
    parts = []
    ms = 0.0
    for k in range(args.parts):
        b, e = pkg.sample_range(args.spp, k, args.parts)
        film.clear()
        st = integ.render(film, sample_begin=b, sample_end=e)
        ms += st["ms_total"]
        parts.append(film.raw().astype(np.float64))
    film.clear()
    st = integ.render(film)
    ours = film.get_pixel_rgb().astype(np.float64)
    raw = np.stack(parts)                                   # [K, H, W, 4] sensor XYZ sums + weight sums
    w = raw[..., 3:4]
    xyz_k = np.where(np.abs(w) > 0, raw[..., :3] / np.where(w == 0, 1, w), 0.0)
    M = np.asarray(pkg.spectral.srgb_color_space()["rgb_from_xyz"], dtype=np.float64)
    rgb_k = xyz_k @ M.T
    var_mean = rgb_k.var(axis=0, ddof=1) / args.parts       # variance of the N-spp mean, per pixel and channel
    res = {"scene": os.path.basename(args.scene), "width": W, "height": H, "spp": args.spp, "parts": args.parts,
           "ours": {"ms": st["ms_total"], "Msamples_s": W * H * args.spp / st["ms_total"] / 1e3,
                    "mean_rgb": [float(v) for v in ours.reshape(-1, 3).mean(0)]}}
    ref = None
    if args.ref_film:
        ref = np.fromfile(args.ref_film, dtype=np.float32).reshape(H, W, 3).astype(np.float64)
    else:
        exe = os.path.join(ROOT, "oracle", "_ref", "pbrt-minus-ref")
        out_png = os.path.join(args.out, args.tag + "_ref.png")
        t0 = time.time()
        r = subprocess.run([exe, os.path.abspath(args.scene), "--spp", str(args.spp), "--outfile", out_png], capture_output=True,
                           text=True, timeout=3000, cwd=os.path.dirname(os.path.abspath(args.scene)))
        wall = time.time() - t0
        if r.returncode == 0 and os.path.exists(out_png + ".f32"):
            ref = np.fromfile(out_png + ".f32", dtype=np.float32).reshape(H, W, 3).astype(np.float64)
            m = re.search(r"rendering took\s+([0-9.]+)", r.stdout)
            rs = float(m.group(1)) if m else None
            res["reference"] = {"render_s": rs, "wall_s": round(wall, 1), "Msamples_s": W * H * args.spp / rs / 1e6 if rs else None,
                                "mean_rgb": [float(v) for v in ref.reshape(-1, 3).mean(0)]}
        else:
            res["reference"] = {"error": (r.stdout + r.stderr)[-400:]}
    if ref is not None:
        d = ours - ref
        finite = np.isfinite(ref).all(axis=2) & (np.abs(ref).max(axis=2) < 1e4)      # the reference's own fireflies / NaNs
        rel = (d ** 2 / (ref ** 2 + 1e-4))[finite]
        tau = 2.0 * (var_mean / (ours ** 2 + 1e-4))[finite]
        sig = np.sqrt(2.0 * var_mean)
        lit = finite & (var_mean.max(axis=2) > 0)
        out3 = (np.abs(d) > 3.0 * sig + 1e-6).any(axis=2) & lit
        z = (np.abs(d) / np.maximum(sig, 1e-12))[lit]
        res["compare"] = {"relMSE": float(rel.mean()), "tau_expected": float(tau.mean()), "slack": args.slack,
                          "relMSE_over_tau": float(rel.mean() / max(tau.mean(), 1e-30)),
                          "gate_relMSE": bool(rel.mean() <= args.slack * tau.mean()),
                          "pixels_excluded_reference_nonfinite_or_gt_1e4": int((~finite).sum()),
                          # the excluded pixels, listed: (x, y, reference rgb, ours rgb); the reference's own fireflies / NaNs
                          "excluded_pixels": [{"x": int(x), "y": int(y), "ref": [float(v) for v in ref[y, x]], "ours": [float(v) for v in ours[y, x]]}
                                              for y, x in list(zip(*np.nonzero(~finite)))[:64]],
                          "mean_rgb_ref_without_excluded": [float(v) for v in ref[finite].reshape(-1, 3).mean(0)],
                          "mean_rgb_ours_without_excluded": [float(v) for v in ours[finite].reshape(-1, 3).mean(0)],
                          "max_abs_diff_without_excluded": float(np.abs(d[finite]).max()),
                          "frac_pixels_beyond_3sigma": float(out3.sum() / max(1, lit.sum())),
                          "z_rms": float(np.sqrt((z ** 2).mean())), "z_p50": float(np.median(z)),
                          "identical_pixels_frac": float((np.abs(d).max(axis=2) == 0).mean()),
                          "mean_rel_diff_of_means": float(abs(ours[finite].mean() - ref[finite].mean()) / max(ref[finite].mean(), 1e-12))}
    print(json.dumps(res))
    os.makedirs(args.out, exist_ok=True)
    open(os.path.join(args.out, args.tag + ".json"), "w").write(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
