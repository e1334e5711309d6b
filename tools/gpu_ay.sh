#!/bin/bash
# C4 proper (50 M triangles, 100 k emissive triangles, 1920x1080, 256 spp) against the reference's own render with the alias
# table replayed at full size (B200PT_ALIAS_REPLAY_MAX): the reference's O(n^2 log n) construction, paid once here too
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out /tmp/c4
export B200PT_REQUIRE_GPU=1
python tools/make_c4.py --out /tmp/c4 | tail -1
date +%s > /tmp/t0
B200PT_ALIAS_REPLAY_MAX=1000000 timeout 1300 python tools/render_stat_compare.py --scene /tmp/c4/c4.pbrt --spp 256 --parts 4 --out $PWD/gpurun_out --tag c4_50M_256spp_replay > gpurun_out/ay_replay.log 2>&1
echo "compare took $(( $(date +%s) - $(cat /tmp/t0) )) s"
rm -f gpurun_out/*.png gpurun_out/*.f32
python - <<'PY'
import json
r=json.load(open("gpurun_out/c4_50M_256spp_replay.json"))
r["compare"].pop("excluded_pixels", None)
print(json.dumps(r))
PY
tail -3 gpurun_out/ay_replay.log | cut -c1-400
