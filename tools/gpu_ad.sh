#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
for v in ca cb cc; do
B200PT_LIB=$PWD/pbrt-minus_b200/libb200pt_$v.so timeout 600 python tools/build_sweep.py --sizes 1 10 100 --repeat 5 --check > gpurun_out/ad_build_$v.log 2>&1; echo "$v rc=$?"
done
python - <<'PY'
import json
for f in ("ca","cb","cc"):
    for l in open("gpurun_out/ad_build_%s.log"%f):
        if l.startswith("{"):
            d=json.loads(l); print(f, d["n_triangles"], d["median_ms"], d["n_bvh8_nodes"], d.get("check"))
PY
B200PT_LIB=$PWD/pbrt-minus_b200/libb200pt_ca.so timeout 900 python -m pytest tests/test_trace_gpu.py -m gpu -x -q > gpurun_out/ad_pytest_trace.log 2>&1; tail -3 gpurun_out/ad_pytest_trace.log
