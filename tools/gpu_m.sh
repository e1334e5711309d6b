#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
ncu --set full --clock-control none -k regex:k_refit --launch-skip 1 --launch-count 1 -o gpurun_out/m_refit python bench.py --steps 1 --warmup 1 --no-render --no-cpu-baseline > gpurun_out/m_ncu.log 2>&1
ncu -i gpurun_out/m_refit.ncu-rep --page raw --csv > gpurun_out/m_refit_raw.csv 2>/dev/null
ncu -i gpurun_out/m_refit.ncu-rep --page details > gpurun_out/m_refit_details.txt 2>/dev/null
rm -f gpurun_out/m_refit.ncu-rep
tail -3 gpurun_out/m_ncu.log
