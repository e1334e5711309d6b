#!/bin/bash
# C4 at 782 k triangles / 4000 emissive triangles, 480x270, 1024 spp against the reference's own render, with the alias
# table replayed beyond the default limit (B200PT_ALIAS_REPLAY_MAX); then the tests the light-sampler change touches
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out /tmp/c4m
export B200PT_REQUIRE_GPU=1
python tools/make_c4.py --out /tmp/c4m --meshes 12 --grid 180 --light-quads 2000 --res 480 270 | tail -1
B200PT_ALIAS_REPLAY_MAX=100000 python tools/render_stat_compare.py --scene /tmp/c4m/c4.pbrt --spp 1024 --parts 16 --out $PWD/gpurun_out --tag c4_4000lights_replay > gpurun_out/ax_replay.log 2>&1
python - <<'PY'
import json
r=json.load(open("gpurun_out/c4_4000lights_replay.json"))
print("replay:", {k: r["compare"][k] for k in ("relMSE","tau_expected","identical_pixels_frac","frac_pixels_beyond_3sigma","mean_rel_diff_of_means","max_abs_diff_without_excluded")}, r["ours"]["ms"], r["reference"])
PY
python tools/render_stat_compare.py --scene /tmp/c4m/c4.pbrt --spp 1024 --parts 16 --out $PWD/gpurun_out --tag c4_4000lights_default --ref-film $PWD/gpurun_out/c4_4000lights_replay_ref.png.f32 > gpurun_out/ax_default.log 2>&1
python - <<'PY'
import json
r=json.load(open("gpurun_out/c4_4000lights_default.json"))
print("default (ordered set beyond 2048):", {k: r["compare"][k] for k in ("relMSE","tau_expected","identical_pixels_frac","frac_pixels_beyond_3sigma","mean_rel_diff_of_means")})
PY
rm -f gpurun_out/*.png gpurun_out/*.f32
true
