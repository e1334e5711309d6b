#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 300 python -m pytest tests/test_trace_gpu.py -m gpu -x -q > gpurun_out/ak_pytest_trace.log 2>&1; echo "trace tests rc=$?"; tail -3 gpurun_out/ak_pytest_trace.log
timeout 300 python tools/build_sweep.py --sizes 1 10 100 --repeat 5 --check > gpurun_out/ak_build.log 2>&1; echo "rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/ak_build.log"):
    if l.startswith("{"):
        d=json.loads(l); print(d["n_triangles"], d["sort_passes"], d["median_ms"], d["n_bvh8_nodes"], d.get("check"))
PY
