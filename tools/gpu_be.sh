#!/bin/bash
# our films of the goldens with the largest residuals, saved for a pixel-by-pixel look in the container
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/films
export B200PT_REQUIRE_GPU=1
python - <<'PY' 2>&1 | tail -8
import sys, os, numpy as np, tempfile, importlib
sys.path.insert(0, "tests")
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T
for name in ("normals_96x64_1024spp", "loopsubdiv_96x64_1024spp", "materials_96x64_1024spp", "disk_96x64_1024spp"):
    m, ref = T.golden(name)
    text = T.scene_text(m["scene"], None, None, [tuple(e) for e in m.get("edits", [])])
    ours, _ = T.render(b200pt, text, m["spp"], tempfile.mkdtemp())
    ours.astype(np.float32).tofile("gpurun_out/films/%s.f32" % name)
    print(name, T.rel_mse(ours, ref))
PY
