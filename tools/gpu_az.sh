#!/bin/bash
# hashed materials (coated*, mix) through the BRIDGE (the reference's own scene objects) against the reference's films: how close
# do the films get when the scene description is the reference's to the bit
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python - <<'PY' 2>&1 | tail -12
import sys, os, numpy as np, pathlib, tempfile, importlib
sys.path.insert(0, "tests")
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T, test_bridge_gpu as B
for name in ("coated_96x64_64spp", "coated_96x64_1024spp", "mix_96x64_1024spp", "metals_coated_96x64_1024spp", "c4tiny_96x54_1024spp"):
    if name not in T.MANIFEST: print(name, "no golden"); continue
    m, ref = T.golden(name)
    tmp = pathlib.Path(tempfile.mkdtemp())
    if m["scene"] == "@c4tiny":
        path, _ = b200pt.workloads.procedural_scene(str(tmp), **b200pt.workloads.C4_TINY)
        text = open(path).read()
        import shutil
        for f in os.listdir(tmp):
            pass
        r, out = B.run_bridge(text, m["spp"], tmp)
    else:
        text = T.scene_text(m["scene"], None, None, [tuple(e) for e in m.get("edits", [])])
        r, out = B.run_bridge(text, m["spp"], tmp)
    if r.returncode: print(name, "bridge failed", r.stderr[-300:]); continue
    film = np.fromfile(out + ".f32", dtype=np.float32).reshape(m["height"], m["width"], 3)
    ours, _ = T.render(b200pt, text, m["spp"], tmp) if m["scene"] != "@c4tiny" else (None, None)
    d = np.abs(film - ref).max(axis=2)
    print(name, "bridge vs golden relMSE %.3e identical %.3f close %.4f" % (T.rel_mse(film, ref), float((d == 0).mean()), T.frac_close(film, ref)),
          "" if ours is None else "| python vs golden %.3e | bridge vs python %.3e" % (T.rel_mse(ours, ref), T.rel_mse(film, ours)))
PY
