#!/bin/bash
# every golden scene through the BRIDGE with B200PT_DESC_DUMP: the reference-exact scene descriptions come back under
# gpurun_out/desc/<golden>/ for tools/desc_diff.py, with the bridge's and the Python front-end's relMSE against the golden
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/desc
export B200PT_REQUIRE_GPU=1
python - <<'PY' 2>&1 | tail -40
import sys, os, numpy as np, pathlib, tempfile, importlib, subprocess, shutil
sys.path.insert(0, "tests")
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T, test_bridge_gpu as B
names = ["cornell_64_16spp", "materials_96x64_64spp", "materials_96x64_64spp_lens", "lights_96x64_64spp", "instances_96x64_64spp", "normals_96x64_64spp",
         "media_96x64_64spp", "media2_96x64_64spp", "loopsubdiv_96x64_64spp", "disk_96x64_64spp", "textures_96x64_64spp", "envmap_96x64_64spp",
         "spherelight_96x64_64spp", "textures2_96x64_64spp", "occluders_96x64_64spp", "mediaball_96x64_64spp", "smokeball_96x64_64spp", "metals_96x64_64spp",
         "coated_96x64_64spp", "mix_96x64_1024spp", "c4tiny_96x54_1024spp"]
for name in names:
    m, ref = T.golden(name)
    tmp = pathlib.Path(tempfile.mkdtemp())
    dump = os.path.abspath(os.path.join("gpurun_out", "desc", name)); os.makedirs(dump, exist_ok=True)
    if m["scene"] == "@c4tiny":
        path, _ = b200pt.workloads.procedural_scene(str(tmp), **b200pt.workloads.C4_TINY)
        text = open(path).read()
    else:
        res = (m["width"], m["height"]) if m["scene"].startswith("cornell") else None
        text = T.scene_text(m["scene"], res, None, [tuple(e) for e in m.get("edits", [])])
    for f in os.listdir(os.path.join("tests", "scenes")):          # PLY meshes and the like, relative to the scene
        src = os.path.join("tests", "scenes", f)
        if os.path.isfile(src) and not f.endswith(".pbrt"): shutil.copy(src, tmp / f)
    os.environ["B200PT_DESC_DUMP"] = dump
    r, out = B.run_bridge(text, m["spp"], tmp)
    os.environ.pop("B200PT_DESC_DUMP")
    open(os.path.join(dump, "scene.pbrt"), "w").write(text)
    if r.returncode: print(name, "bridge failed", (r.stdout + r.stderr)[-300:]); continue
    film = np.fromfile(out + ".f32", dtype=np.float32).reshape(m["height"], m["width"], 3)
    if m["scene"] == "@c4tiny":
        print("%-32s bridge vs golden %.3e" % (name, T.rel_mse(film, ref))); continue
    ours, _ = T.render(b200pt, text, m["spp"], tmp)
    print("%-32s bridge vs golden %.3e | python vs golden %.3e | bridge vs python %.3e" % (name, T.rel_mse(film, ref), T.rel_mse(ours, ref), T.rel_mse(film, ours)), flush=True)
PY
du -sh gpurun_out/desc
# back in the container: for n in $(ls gpurun_out/desc); do python tools/desc_diff.py --pack gpurun_out/desc/$n tests/golden/desc/$n.npz; done
