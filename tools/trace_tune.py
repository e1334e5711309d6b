#!/usr/bin/env python3
"""Sweep the scheduling knobs of k_trace (b200pt_trace_tune) on config C2 and check that the hit ids do not move.

  python tools/trace_tune.py [--triangles N] [--reps R] [--out profiles/xxx.json]

One scene build, then for every (refill, tri_steps, ctas_per_sm) combination: device-resident trace time of the three
ray sets (CUDA events on the scene's stream, median of R launches after a warm-up launch) and a comparison of the hit
records with those of the default configuration (the knobs must change speed only)."""
import argparse
import importlib
import itertools
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--triangles", type=int, default=10_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--refill", default="1,4,8,16")
    ap.add_argument("--tri-steps", default="1,2,3")
    ap.add_argument("--ctas", default="0")
    ap.add_argument("--stash", default="0,1")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    P, idx = pkg.workloads.triangle_soup(args.triangles)
    scene = pkg.HLBVH(P, idx, device=0, build=True)
    sets = bench.make_workload(pkg, args, scene, 0)
    stream = torch.cuda.ExternalStream(scene.stream, device=torch.device("cuda", 0))
    dev = []
    for name, any_hit, (o, d, tm) in sets:
        rays = scene.pack_rays(o, d, tm)
        rb, hb = scene.alloc(rays.nbytes), scene.alloc(len(o) * 16)
        rb.upload(rays)
        dev.append((name, any_hit, len(o), rb, hb))

    def run(cfg):
        scene.tune(refill=cfg[0], tri_steps=cfg[1], ctas_per_sm=cfg[2], stash=cfg[3] if len(cfg) > 3 else 0)
        res = {}
        hits = {}
        for name, any_hit, n, rb, hb in dev:
            scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=True)
            ts = []
            for _ in range(args.reps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=False)
                e1.record(stream)
                scene.sync()
                ts.append(e0.elapsed_time(e1))
            res[name] = float(np.median(ts))
            hits[name] = hb.download(np.uint32, n * 4).copy()
        return res, hits

    base_cfg = (4, 3, 0, 0)
    _, base_hits = run(base_cfg)
    rows = []
    total_rays = sum(n for _, _, n, _, _ in dev)
    for cfg in itertools.product([int(x) for x in args.refill.split(",")], [int(x) for x in args.tri_steps.split(",")],
                                 [int(x) for x in args.ctas.split(",")], [int(x) for x in args.stash.split(",")]):
        res, hits = run(cfg)
        same = all(np.array_equal(hits[k], base_hits[k]) for k in hits)
        ms = sum(res.values())
        row = {"refill": cfg[0], "tri_steps": cfg[1], "ctas_per_sm": cfg[2], "stash": cfg[3], "ms": {k: round(v, 4) for k, v in res.items()},
               "ms_total": round(ms, 4), "Mrays_s": round(total_rays / ms / 1e3, 1), "hits_identical_to_default": bool(same)}
        rows.append(row)
        print(json.dumps(row), flush=True)
    scene.tune(refill=4, tri_steps=3, ctas_per_sm=0, stash=0)
    best = max(rows, key=lambda r: r["Mrays_s"])
    print("best:", json.dumps(best))
    if args.out:
        json.dump({"triangles": args.triangles, "rays": total_rays, "rows": rows, "best": best}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
