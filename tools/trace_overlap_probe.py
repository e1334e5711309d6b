#!/usr/bin/env python3
"""C2's three device-resident ray sets: back to back on one stream (b200pt_trace_device x 3) against b200pt_trace_device_batch
(two traversal streams).  Prints ms per step and Mrays/s for both, and checks the hit records are identical."""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--triangles", type=int, default=10_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--steps", type=int, default=30)
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
        dev.append((rb, len(o), hb, any_hit))
    total = sum(n for _, n, _, _ in dev)

    def seq():
        for rb, n, hb, any_hit in dev:
            scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=False)

    def bat():
        scene.trace_device_batch(dev, sync=False)

    out = {}
    hits = {}
    for name, fn in (("sequential", seq), ("batch", bat), ("sequential2", seq), ("batch2", bat)):
        for _ in range(3):
            fn()
        scene.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            fn()
        e1.record(stream)
        scene.sync()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        out[name] = {"ms_per_step": round(ms, 4), "Mrays_s": round(total / ms / 1e3, 1)}
        hits[name] = [hb.download(np.uint32, n * 4).copy() for _, n, hb, _ in dev]
    out["identical"] = all(np.array_equal(a, b) for a, b in zip(hits["sequential"], hits["batch"]))
    scene.counters()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
