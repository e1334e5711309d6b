#!/bin/bash
# c4tiny (400 emissive triangles, many equal weights) against the reference's golden with the alias table replaying the reference's
# sort and the light weights rounded as on the host; then every golden re-measured against the committed tolerances
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
for mode in host device; do
B200PT_LIGHT_WEIGHTS=$mode B200PT_DEBUG_LIGHTS=1 python - <<'PY' 2>&1 | grep -v 'lightdbg\] [0-9]* weight' | tail -4
import sys, os, numpy as np, pathlib, tempfile, importlib
sys.path.insert(0, "tests")
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T
tmp = tempfile.mkdtemp()
m, ref = T.golden("c4tiny_96x54_1024spp")
path, info = b200pt.workloads.procedural_scene(tmp, **b200pt.workloads.C4_TINY)
sc = b200pt.pbrt_scene.load(path, spp=m["spp"])
integ = b200pt.WavefrontPathIntegrator(sc, device=0)
film = integ.create_film(); integ.render(film); ours = film.get_pixel_rgb()
print(os.environ["B200PT_LIGHT_WEIGHTS"], "c4tiny relMSE", T.rel_mse(ours, ref), "frac_close", T.frac_close(ours, ref), "identical", float((ours == ref).all(axis=2).mean()), ours.mean(), ref.mean())
PY
done
python tools/measure_golden.py > gpurun_out/av_measure.log 2>&1; tail -2 gpurun_out/av_measure.log
python - <<'PY'
import json
old=json.load(open("tests/golden/films/tolerances.json")); new=json.load(open("gpurun_out/tolerances.json"))
for k,v in new.items():
    o=old.get(k)
    if o and (v["relMSE_measured"] > o["relMSE_tol"] or v["frac_close_measured"] < o["frac_close_tol"]): print("WORSE", k, o, v)
    elif o and v["relMSE_measured"] < 0.5*o["relMSE_measured"]: print("better", k, o["relMSE_measured"], "->", v["relMSE_measured"])
    elif o and v["relMSE_measured"] > 2*o["relMSE_measured"]: print("worse within tol", k, o["relMSE_measured"], "->", v["relMSE_measured"])
PY
