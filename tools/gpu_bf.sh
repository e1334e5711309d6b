#!/bin/bash
# instanced scenes rendered two-level (default) and with the instances flattened by the front-end, each against the reference's film
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/films
export B200PT_REQUIRE_GPU=1
python - <<'PY' 2>&1 | tail -8
import sys, os, numpy as np, tempfile, importlib
sys.path.insert(0, "tests")
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T
for name in ("normals_96x64_64spp", "instances_96x64_64spp", "normals_96x64_1024spp"):
    m, ref = T.golden(name)
    text = T.scene_text(m["scene"], None, None, [tuple(e) for e in m.get("edits", [])])
    for flat in (False, True):
        p = os.path.join(tempfile.mkdtemp(), "s.pbrt"); open(p, "w").write(text)
        sc = b200pt.pbrt_scene.load(p, spp=m["spp"], scene_dir=os.path.join("tests", "scenes"), flatten_instances=flat)
        integ = b200pt.WavefrontPathIntegrator(sc, device=0); film = integ.create_film(); integ.render(film); ours = film.get_pixel_rgb(); film.close(); integ.close()
        d = (ours - ref).sum(axis=2); big = np.abs(ours - ref).max(axis=2) > 1e-3 * (np.abs(ref).max(axis=2) + 1e-3)
        print("%-24s %-9s relMSE %.3e  differing %3d: brighter %3d darker %3d  sum %+.4f" % (name, "flattened" if flat else "two-level", T.rel_mse(ours, ref), big.sum(), (d[big] > 0).sum(), (d[big] < 0).sum(), d[big].sum()))
PY
