#!/bin/bash
# goldens for the occluder scene from the reference, merged into tests/golden/films on the box, then measured
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python tools/make_golden.py smokeball_96x64_64spp smokeball_96x64_1024spp > gpurun_out/ah_make.log 2>&1; tail -3 gpurun_out/ah_make.log
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
python tools/measure_golden.py > gpurun_out/ah_measure.log 2>&1; grep smokeball gpurun_out/ah_measure.log
ls gpurun_out/golden | grep smokeball
