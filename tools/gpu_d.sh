#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/render_probe.py --occ 0,3,4 > gpurun_out/d_occ.log 2>&1
python tools/render_probe.py --occ 0,3,4 --res 1920 1080 --spp 64 >> gpurun_out/d_occ.log 2>&1
# build kernels + k_trace: full capture (bench without render / cpu leg; 2 steps)
ncu --set full --clock-control none --import-source on -k regex:'k_collapse_level|k_refit|k_sort_downsweep|k_sort_upsweep|k_emit_hierarchy|k_morton|k_centroid' --launch-skip 0 --launch-count 40 -o gpurun_out/d_build_full python bench.py --steps 1 --warmup 1 --no-render --no-cpu-baseline > gpurun_out/d_ncu_build.log 2>&1
ncu -i gpurun_out/d_build_full.ncu-rep --page raw --csv > gpurun_out/d_build_full_raw.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:'k_trace' --launch-skip 12 --launch-count 3 -o gpurun_out/d_trace_full python bench.py --steps 2 --warmup 3 --no-render --no-cpu-baseline > gpurun_out/d_ncu_trace.log 2>&1
ncu -i gpurun_out/d_trace_full.ncu-rep --page raw --csv > gpurun_out/d_trace_full_raw.csv 2>/dev/null
cat gpurun_out/d_occ.log
