#!/bin/bash
# source-level ncu of the diffuse shade kernel and the closest-hit wavefront trace on a steady-state pass of C1
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/render_probe.py --reps 1 > gpurun_out/x_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_shade' --launch-skip 30 --launch-count 3 -o gpurun_out/x_shade python tools/render_probe.py --reps 1 > gpurun_out/x_ncu.log 2>&1
ncu -i gpurun_out/x_shade.ncu-rep --page source --csv --print-source sass > gpurun_out/x_shade_source_sass.csv 2>/dev/null
ncu -i gpurun_out/x_shade.ncu-rep --page source --csv --print-source cuda > gpurun_out/x_shade_source_cuda.csv 2>/dev/null
ncu -i gpurun_out/x_shade.ncu-rep --page raw --csv > gpurun_out/x_shade_raw.csv 2>/dev/null
ls -la gpurun_out/x_*
rm -f gpurun_out/x_shade.ncu-rep
tail -3 gpurun_out/x_ncu.log
