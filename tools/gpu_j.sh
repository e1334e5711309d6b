#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python tools/make_golden.py textures2_96x64_64spp textures2_96x64_1024spp > gpurun_out/j_golden.log 2>&1; tail -3 gpurun_out/j_golden.log | cut -c1-300
cp gpurun_out/golden/textures2_96x64_64spp.f32 gpurun_out/golden/textures2_96x64_1024spp.f32 tests/golden/films/ 2>/dev/null
python - <<'PY'
import json
new=json.load(open('gpurun_out/golden/manifest_new.json'))
man=json.load(open('tests/golden/films/manifest.json'))
man.update(new)
json.dump(man, open('tests/golden/films/manifest.json','w'), indent=1)
PY
timeout 900 python -m pytest tests/test_render_gpu.py -m gpu -x -q -k "mappings or srgb8 or textures" > gpurun_out/j_pytest.log 2>&1; tail -15 gpurun_out/j_pytest.log
python tools/measure_golden.py > gpurun_out/j_measure.log 2>&1; grep textures2 gpurun_out/j_measure.log | cut -c1-220
