#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 1200 python -m pytest tests/test_trace_gpu.py -m gpu -x -q > gpurun_out/l_pytest_trace.log 2>&1; tail -5 gpurun_out/l_pytest_trace.log
timeout 600 python tools/build_sweep.py --sizes 1 10 100 --check --out gpurun_out/l_build_sweep.json > gpurun_out/l_build_sweep.log 2>&1
python - <<'PY'
import json
for r in json.load(open('gpurun_out/l_build_sweep.json'))["rows"] if isinstance(json.load(open('gpurun_out/l_build_sweep.json')), dict) else json.load(open('gpurun_out/l_build_sweep.json')):
    print(r["n_triangles"], r["median_ms"], r.get("check"))
PY
tail -3 gpurun_out/l_build_sweep.log | cut -c1-500
