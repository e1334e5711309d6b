#!/usr/bin/env python3
"""Config C4: write the procedural scene (workloads.procedural_scene) as .pbrt + binary PLY files.

  python tools/make_c4.py --out /tmp/c4 [--meshes 48] [--grid 721] [--light-quads 50000] [--res 1920 1080]
then   python tools/render_compare.py --scene /tmp/c4/c4.pbrt --out /tmp/c4 --spp 16 --tag c4_16
"""
import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--meshes", type=int, default=48)
    ap.add_argument("--grid", type=int, default=721)
    ap.add_argument("--light-quads", type=int, default=50000)
    ap.add_argument("--res", type=int, nargs=2, default=[1920, 1080])
    ap.add_argument("--seed", type=int, default=7)
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    t0 = time.time()
    path, info = pkg.workloads.procedural_scene(args.out, n_meshes=args.meshes, grid=args.grid, n_light_quads=args.light_quads,
                                                width=args.res[0], height=args.res[1], seed=args.seed)
    info.update({"scene": path, "generate_s": round(time.time() - t0, 2)})
    print(json.dumps(info))


if __name__ == "__main__":
    main()
