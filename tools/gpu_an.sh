#!/bin/bash
# launch list + full capture of one steady-state pass of the C1 render with the 16 Mi pool
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/render_probe.py --reps 1 > gpurun_out/an_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/an_render_launches.csv python tools/render_probe.py --reps 1 > gpurun_out/an_ncu1.log 2>&1
ncu --set full --clock-control none -k regex:'k_logic|k_shade|k_trace_wf|k_generate|k_emit' --launch-skip 40 --launch-count 8 -o gpurun_out/an_render_full python tools/render_probe.py --reps 1 > gpurun_out/an_ncu2.log 2>&1
ncu -i gpurun_out/an_render_full.ncu-rep --page raw --csv > gpurun_out/an_render_full_raw.csv 2>/dev/null
rm -f gpurun_out/an_render_full.ncu-rep
tail -1 gpurun_out/an_plain.log
