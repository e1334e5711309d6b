#!/usr/bin/env python3
"""Would sorting the rays pay?  Trace the C2 ray sets in the caller's order and in a few coherent orders (host-sorted, so only
the traversal is timed) and print the device times.  python tools/ray_order_probe.py [--triangles N]"""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def part1by2(x):
    x = x.astype(np.uint64) & 0x3ff
    x = (x | (x << 16)) & 0x30000ff
    x = (x | (x << 8)) & 0x300f00f
    x = (x | (x << 4)) & 0x30c30c3
    x = (x | (x << 2)) & 0x9249249
    return x


def morton3(p, lo, hi, bits=10):
    q = np.clip(((p - lo) / np.maximum(hi - lo, 1e-30) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    return (part1by2(q[:, 0]) << 2) | (part1by2(q[:, 1]) << 1) | part1by2(q[:, 2])


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--triangles", type=int, default=10_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--reps", type=int, default=7)
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    P, idx = pkg.workloads.triangle_soup(args.triangles)
    scene = pkg.HLBVH(P, idx, device=0, build=True)
    sets = bench.make_workload(pkg, args, scene, 0)
    stream = torch.cuda.ExternalStream(scene.stream, device=torch.device("cuda", 0))
    lo, hi = P.min(axis=0), P.max(axis=0)

    def timed(o, d, tm, any_hit):
        rays = scene.pack_rays(o, d, tm)
        rb, hb = scene.alloc(rays.nbytes), scene.alloc(len(o) * 16)
        rb.upload(rays)
        scene.trace_device(rb, len(o), hb, any_hit=bool(any_hit), sync=True)
        ts = []
        for _ in range(args.reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            scene.trace_device(rb, len(o), hb, any_hit=bool(any_hit), sync=False)
            e1.record(stream)
            scene.sync()
            ts.append(e0.elapsed_time(e1))
        hits = hb.download(np.uint32, len(o) * 4).reshape(-1, 4).copy()
        rb.free()
        hb.free()
        return float(np.median(ts)), hits

    for name, any_hit, (o, d, tm) in sets:
        n = len(o)
        base_ms, base_hits = timed(o, d, tm, any_hit)
        orders = {}
        octant = ((d[:, 0] < 0).astype(np.uint64) << 2) | ((d[:, 1] < 0).astype(np.uint64) << 1) | (d[:, 2] < 0).astype(np.uint64)
        mo = morton3(o, lo, hi)
        orders["octant+origin_morton"] = np.argsort((octant << 30) | mo, kind="stable")
        orders["origin_morton"] = np.argsort(mo, kind="stable")
        # where the ray is heading: the point a quarter of the scene diagonal along it
        diag = float(np.linalg.norm(hi - lo))
        dn = d / np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-30)
        orders["octant+target_morton"] = np.argsort((octant << 30) | morton3(o + dn * 0.1 * diag, lo, hi), kind="stable")
        if name == "primary" and n == args.width * args.height:
            y, x = np.divmod(np.arange(n), args.width)
            orders["pixel_tiles_8x4"] = np.lexsort((x % 8, y % 4, x // 8, y // 4))
            orders["pixel_morton"] = np.argsort((part1by2(x) << 1) | part1by2(y), kind="stable")
        row = {"set": name, "rays": n, "caller_order_ms": round(base_ms, 4)}
        for k, perm in orders.items():
            ms, hits = timed(o[perm], d[perm], None if tm is None else (tm[perm] if np.ndim(tm) else tm), any_hit)
            same = np.array_equal(hits[:, 0], base_hits[perm, 0])
            row[k + "_ms"] = round(ms, 4)
            row[k + "_same_ids"] = bool(same)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
