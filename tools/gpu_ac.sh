#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
for pf in 0 1; do
B200PT_COLLAPSE_PREFETCH=$pf timeout 600 python tools/build_sweep.py --sizes 1 10 100 --repeat 5 --check > gpurun_out/ac_build_pf$pf.log 2>&1; echo "pf$pf rc=$?"
done
python - <<'PY'
import json
for f in ("pf0","pf1"):
    for l in open("gpurun_out/ac_build_%s.log"%f):
        if l.startswith("{"):
            d=json.loads(l); print(f, d["n_triangles"], d["median_ms"], d["n_bvh8_nodes"], d.get("check"))
PY
