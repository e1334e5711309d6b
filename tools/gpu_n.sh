#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for sm in 0 24576 36864 49152 65536 98304; do
  echo "smem $sm" >> gpurun_out/n_collapse_smem.log
  B200PT_COLLAPSE_SMEM=$sm python tools/build_sweep.py --sizes 10 --out gpurun_out/n_tmp.json 2>&1 | python -c "import sys,json; [print(json.loads(l)['median_ms']) for l in sys.stdin if l.startswith('{')]" >> gpurun_out/n_collapse_smem.log
done
cat gpurun_out/n_collapse_smem.log
