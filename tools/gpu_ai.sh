#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:'k_trace_wf|k_generate|k_logic' --launch-skip 40 --launch-count 4 -o gpurun_out/ai_wf python tools/render_probe.py --reps 1 > gpurun_out/ai_ncu.log 2>&1
ncu -i gpurun_out/ai_wf.ncu-rep --page source --csv --print-source sass > gpurun_out/ai_wf_source_sass.csv 2>/dev/null
ncu -i gpurun_out/ai_wf.ncu-rep --page raw --csv > gpurun_out/ai_wf_raw.csv 2>/dev/null
rm -f gpurun_out/ai_wf.ncu-rep
ls -la gpurun_out/ai_*
