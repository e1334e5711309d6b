#!/usr/bin/env python3
"""Config C5: HLBVH build throughput sweep, 1 M .. 100 M synthetic triangles on one B200.

For every N: soup as C2 (seed = N, s = 2*N^(-1/3)); timed = b200pt_bvh_build, all phases, device events
(first build = with allocation of the build scratch, then --repeat rebuilds, median reported);
algorithmic bytes = 428 B/triangle (30-bit keys, N <= 16 Mi) or 372 + 16 * sort passes (packed keys, N > 16 Mi).
--reference additionally times the reference's own HLBVH constructor on the same soup file
(oracle/_ref/ref_trace with no ray jobs; its "BVH building" line, host cores + managed memory), for N <= --ref-max.
Validation per N (--check): a sample of primary rays traced through the fresh tree must hit the same
primitive as the tree built at the previous repeat (determinism) and every triangle must be placed once
(the build itself fails otherwise).

  python tools/build_sweep.py --sizes 1 2 5 10 20 50 100 --reference --out gpurun_out/c5.json
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", type=float, nargs="+", default=[1, 2, 5, 10, 20, 50, 100], help="millions of triangles")
    ap.add_argument("--repeat", type=int, default=5)
    ap.add_argument("--reference", action="store_true")
    ap.add_argument("--ref-max", type=float, default=20, help="largest N (millions) handed to the reference build")
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    wl = pkg.workloads
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    rows = []
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_trace")
    for m in args.sizes:
        n = int(round(m * 1e6))
        t0 = time.time()
        P, idx = wl.triangle_soup(n, seed=n)
        t_gen = time.time() - t0
        scene = pkg.HLBVH(P, idx, device=0, build=False)
        first = scene.build()
        reps = [scene.build() for _ in range(args.repeat)]
        keys = ("ms_bounds_morton", "ms_sort", "ms_emit", "ms_refit", "ms_collapse", "ms_total")
        med = {k: float(np.median([r[k] for r in reps])) for k in keys}
        # bounds+key 56, sort 4|8 + 16 P, emit+refit 116|120, collapse 92, re-layout 96
        bpt = 428 if first["key_bits"] == 30 else 372 + 16 * first["sort_passes"]
        row = {"n_triangles": n, "key_bits": first["key_bits"], "sort_passes": first["sort_passes"],
               "n_bvh8_nodes": first["n_bvh8_nodes"], "first_build_ms": round(first["ms_total"], 3), "first_build_wall_ms": round(first["ms_wall"], 3),
               "median_wall_ms": round(float(np.median([r["ms_wall"] for r in reps])), 3), "collapse_levels": first["collapse_levels"],
               "median_ms": {k: round(v, 3) for k, v in med.items()},
               "Mtris_s": round(n / med["ms_total"] / 1e3, 1), "algorithmic_bytes_per_tri": bpt,
               "achieved_GBs": round(n * bpt / med["ms_total"] / 1e6, 1),
               "hbm_peak_GBs": hbm, "frac": round(n * bpt / med["ms_total"] / 1e6 / hbm, 4) if hbm else None,
               "generate_s": round(t_gen, 2)}
        if args.check:
            o, d, tm = wl.primary_rays(480, 270)
            p1, t1, _ = scene.intersect(o, d, tm)
            scene.build()
            p2, t2, _ = scene.intersect(o, d, tm)
            row["check"] = {"rays": len(o), "hits": int((p1 >= 0).sum()), "rebuild_identical": bool(np.array_equal(p1, p2) and np.array_equal(t1, t2))}
        del scene
        if args.reference and os.path.exists(exe) and m <= args.ref_max:
            with tempfile.TemporaryDirectory(dir="/tmp") as td:
                soup = os.path.join(td, "soup.bin")
                wl.write_soup(soup, P, idx)
                try:
                    t0 = time.time()
                    r = subprocess.run([exe, soup], capture_output=True, text=True, timeout=1200)
                    wall = time.time() - t0
                    info = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
                    line = [l for l in r.stdout.splitlines() if l.startswith("BVH building")]
                    row["reference"] = {"bvh_build_s": info["bvh_build_s"], "load_s": info["load_s"], "line": line[-1] if line else None,
                                        "wall_s": round(wall, 1), "host_cores": os.cpu_count()}
                    row["speedup_vs_reference_build"] = round(info["bvh_build_s"] * 1e3 / med["ms_total"], 1)
                except Exception as ex:  # noqa: BLE001
                    row["reference"] = {"unavailable": str(ex)[-200:]}
        print(json.dumps(row), flush=True)
        rows.append(row)
        del P, idx
    if args.out:
        json.dump({"config": "C5 HLBVH build sweep", "rows": rows}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
