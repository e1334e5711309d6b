#!/usr/bin/env python3
"""Render one scene a few times and print device times; optional sweep of the wavefront trace knobs.
  python tools/render_probe.py [--scene PATH] [--res W H] [--spp S] [--sweep] [--reps R]"""
import argparse
import importlib
import json
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=os.path.join(ROOT, "tests", "scenes", "cornell-box-specular.pbrt"))
    ap.add_argument("--res", type=int, nargs=2, default=None)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--pool", type=int, default=0)
    ap.add_argument("--occ", default="0", help="comma list of shade_occ values to try")
    ap.add_argument("--lib", default="", help="A/B runs: another build of libb200pt.so to load instead of the in-tree one")
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    if args.lib:
        pkg.api.LIB_PATH = os.path.abspath(args.lib)
    text = open(args.scene).read()
    if args.res:
        text = re.sub(r'("integer xresolution"\s*\[\s*)\d+', r"\g<1>%d" % args.res[0], text)
        text = re.sub(r'("integer yresolution"\s*\[\s*)\d+', r"\g<1>%d" % args.res[1], text)
    sc = pkg.pbrt_scene.load(text=text, spp=args.spp, scene_dir=os.path.dirname(os.path.abspath(args.scene)))
    integ = pkg.WavefrontPathIntegrator(sc, device=0)
    film = integ.create_film()

    def run():
        ms = []
        for _ in range(args.reps):
            film.clear()
            st = integ.render(film, pool_size=args.pool)
            ms.append(st["ms_total"])
        return float(np.median(ms)), st

    integ.render(film, pool_size=args.pool)
    configs = [(0, 0)]
    if args.sweep:
        configs = [(r, t) for r in (4, 8, 16, 24) for t in (1, 2, 3)]
    for r, t, occ in [(r, t, int(o)) for r, t in configs for o in args.occ.split(",")]:
        integ.accel.tune(wf_refill=r, wf_tri_steps=t, shade_occ=occ)
        ms, st = run()
        print(json.dumps({"lib": os.path.basename(args.lib) or "libb200pt.so", "mean": round(float(film.get_pixel_rgb().mean()), 6), "wf_refill": r, "wf_tri_steps": t, "shade_occ": occ, "ms": round(ms, 3), "Msamples_s": round(st["paths"] / ms / 1e3, 1),
                          "Mrays_s": round((st["rays_closest"] + st["rays_shadow"]) / ms / 1e3, 1), "passes": st["passes"],
                          "launches": st["kernel_launches"]}), flush=True)
    film.close()
    integ.close()


if __name__ == "__main__":
    main()
