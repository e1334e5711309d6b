#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 900 python -m pytest tests/test_trace_gpu.py -m gpu -x -q > gpurun_out/ab_pytest_trace.log 2>&1; echo "trace tests rc=$?"; tail -5 gpurun_out/ab_pytest_trace.log
B200PT_COLLAPSE_COOP=0 timeout 600 python tools/build_sweep.py --sizes 1 10 100 --repeat 5 --check > gpurun_out/ab_build_seq.log 2>&1; echo "seq rc=$?"
B200PT_COLLAPSE_COOP=1 timeout 600 python tools/build_sweep.py --sizes 1 10 100 --repeat 5 --check > gpurun_out/ab_build_coop.log 2>&1; echo "coop rc=$?"
python - <<'PY'
import json
for f in ("seq","coop"):
    for l in open("gpurun_out/ab_build_%s.log"%f):
        if l.startswith("{"):
            d=json.loads(l); print(f, d["n_triangles"], d["median_ms"], d["n_bvh8_nodes"], d.get("check"))
PY
