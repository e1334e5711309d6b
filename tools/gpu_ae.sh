#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 300 python -m pytest tests/test_trace_gpu.py -m gpu -x -q > gpurun_out/ae_pytest_trace.log 2>&1; echo "trace tests rc=$?"; tail -5 gpurun_out/ae_pytest_trace.log
for v in 0 1; do
B200PT_SORT_ONESWEEP=$v timeout 300 python tools/build_sweep.py --sizes 1 10 100 --repeat 5 --check > gpurun_out/ae_build_os$v.log 2>&1; echo "os$v rc=$?"
done
python - <<'PY'
import json
for f in ("os0","os1"):
    for l in open("gpurun_out/ae_build_%s.log"%f):
        if l.startswith("{"):
            d=json.loads(l); print(f, d["n_triangles"], d["sort_passes"], d["median_ms"], d["n_bvh8_nodes"], d.get("check"))
PY
