#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python - <<'PY' 2>&1 | tail -8
import sys, os, numpy as np, tempfile, importlib
sys.path.insert(0, "tests")
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T
for name in ("textures_96x64_64spp", "textures_96x64_1024spp", "textures2_96x64_64spp", "textures2_96x64_1024spp", "envmap_96x64_64spp", "envmap_96x64_1024spp", "metals_96x64_64spp"):
    m, ref = T.golden(name)
    text = T.scene_text(m["scene"], None, None, [tuple(e) for e in m.get("edits", [])])
    ours, _ = T.render(b200pt, text, m["spp"], tempfile.mkdtemp())
    print("%-28s %.3e (was %.3e)" % (name, T.rel_mse(ours, ref), T.TOLERANCES[name]["relMSE_measured"]))
PY
