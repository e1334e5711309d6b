#!/usr/bin/env python3
"""Measure relMSE / fraction-of-close-pixels of every exact-parity golden film (tests/golden/films) on the GPU box and
write tests/golden/films/tolerances.json: per golden, the relMSE bound the tests enforce = 10 x the measured value
(floor 1e-9), and the close-pixel fraction bound = measured - 0.01.  python tools/measure_golden.py [--write]"""
import importlib
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
STATISTICAL = ("coated_", "mix_", "metals_coated_")   # compared as two unbiased estimators (tests: statistical_gate)


def main():
    import test_render_gpu as T
    import test_bridge_gpu as B
    import numpy as np
    pkg = importlib.import_module("pbrt-minus_b200")
    out = {}
    with tempfile.TemporaryDirectory() as td:
        for name, m in T.MANIFEST.items():
            if name.startswith(STATISTICAL) or m["integrator"] != "path":
                continue
            _, ref = T.golden(name)
            res = (m["width"], m["height"]) if m["scene"].startswith("cornell") else None
            sampler = m["sampler"] if m["sampler"] != "stratified" else None
            if m["scene"] == "@c4tiny":       # the procedural many-lights scene (workloads.procedural_scene)
                path, _ = pkg.workloads.procedural_scene(td, **pkg.workloads.C4_TINY)
                sc = pkg.pbrt_scene.load(path, spp=m["spp"])
                integ = pkg.WavefrontPathIntegrator(sc, device=0)
                film = integ.create_film()
                st = integ.render(film)
                ours = film.get_pixel_rgb()
                film.close()
                integ.close()
            else:
                text = T.scene_text(m["scene"], res, sampler, [tuple(e) for e in m.get("edits", [])])
                ours, st = T.render(pkg, text, m["spp"], td)
            r, f = T.rel_mse(ours, ref), T.frac_close(ours, ref)
            out[name] = {"relMSE_measured": r, "frac_close_measured": f, "relMSE_tol": max(10.0 * r, 1e-9),
                         "frac_close_tol": round(max(0.0, f - 0.01), 4), "ms": round(st["ms_total"], 2)}
            # the same scene through the reference's own front-end + the bridge (tests/test_bridge_gpu.py): its own measured
            # distance to the golden and to the Python path's film (the two descriptions differ in a few last bits)
            if m["scene"] != "@c4tiny" and m["spp"] <= 64 and os.path.exists(B.EXE):
                import pathlib
                tmp = pathlib.Path(tempfile.mkdtemp())
                rb, outp = B.run_bridge(text, m["spp"], tmp)
                if rb.returncode == 0:
                    film = np.fromfile(outp + ".f32", dtype=np.float32).reshape(m["height"], m["width"], 3)
                    br, bc = T.rel_mse(film, ref), T.rel_mse(film, ours)
                    out[name].update({"bridge_relMSE_measured": br, "bridge_relMSE_tol": max(10.0 * br, 1e-9),
                                      "bridge_vs_python_measured": bc, "bridge_vs_python_tol": max(10.0 * bc, 1e-9)})
            print(name, json.dumps(out[name]), flush=True)
    dst = os.path.join(ROOT, "tests", "golden", "films", "tolerances.json") if "--write" in sys.argv else os.path.join(ROOT, "gpurun_out", "tolerances.json")
    json.dump(out, open(dst, "w"), indent=1)


if __name__ == "__main__":
    main()
