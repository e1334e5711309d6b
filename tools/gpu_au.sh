#!/bin/bash
# goldens for tests/scenes/metals.pbrt from the reference, merged into tests/golden/films on the box, then measured; the bridge's
# film for the same scene against both
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python tools/make_golden.py metals_96x64_64spp metals_96x64_1024spp metals_coated_96x64_1024spp > gpurun_out/au_make.log 2>&1; tail -3 gpurun_out/au_make.log
python - <<'PY'
import json, shutil, os
g="gpurun_out/golden"; dst="tests/golden/films"
new=json.load(open(os.path.join(g,"manifest_new.json")))
m=json.load(open(os.path.join(dst,"manifest.json")))
for k,v in new.items():
    m[k]=v
    shutil.copy(os.path.join(g,k+".f32"), os.path.join(dst,k+".f32"))
json.dump(m,open(os.path.join(dst,"manifest.json"),"w"),indent=1)
json.dump(m,open("gpurun_out/manifest_merged.json","w"),indent=1)
print(sorted(new))
PY
python tools/measure_golden.py > gpurun_out/au_measure.log 2>&1; grep metals gpurun_out/au_measure.log
python - <<'PY'
import json
old=json.load(open("tests/golden/films/tolerances.json")); new=json.load(open("gpurun_out/tolerances.json"))
for k,v in new.items():
    o=old.get(k)
    if o and (v["relMSE_measured"] > o["relMSE_tol"] or v["frac_close_measured"] < o["frac_close_tol"]): print("WORSE", k, o, v)
    if o and v["relMSE_measured"] < 0.1*o["relMSE_measured"]: print("better", k, o["relMSE_measured"], "->", v["relMSE_measured"])
    if not o: old[k]=v
json.dump(old, open("tests/golden/films/tolerances.json","w"), indent=1)
json.dump(old, open("gpurun_out/tolerances_merged.json","w"), indent=1)
PY
ls gpurun_out/golden | grep metals
python - <<'PY' 2>&1 | tail -20
import sys, os, numpy as np, pathlib, tempfile
sys.path.insert(0, "tests")
import importlib
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T, test_bridge_gpu as B
tmp = pathlib.Path(tempfile.mkdtemp())
m, ref = T.golden("metals_96x64_64spp")
text = T.scene_text(m["scene"], None, None, [])
r, out = B.run_bridge(text, m["spp"], tmp)
print("bridge rc", r.returncode, r.stdout[-600:], r.stderr[-600:])
film = np.fromfile(out + ".f32", dtype=np.float32).reshape(m["height"], m["width"], 3)
ours, _ = T.render(b200pt, text, m["spp"], tmp)
print("bridge vs golden", T.rel_mse(film, ref), "bridge vs ctypes", T.rel_mse(film, ours), "ctypes vs golden", T.rel_mse(ours, ref))
d = np.abs(film - ours).max(axis=2); print("pixels differing bridge/ctypes", int((d > 1e-4 * (1 + ours.max(axis=2))).sum()), "of", d.size)
d = np.abs(ours - ref).max(axis=2); print("pixels differing ctypes/golden", int((d > 1e-4 * (1 + ref.max(axis=2))).sum()), "means", ours.mean(), ref.mean(), film.mean())
PY
python -m pytest tests -q -m gpu -k "texture or bridge or material or coated or light or envmap or many" 2>&1 | tail -8
