#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
ncu --set full --clock-control none -k regex:'k_sort_onesweep|k_sort_histograms|k_gather_sorted|k_centroid_bounds|k_morton|k_emit_hierarchy|k_refit|k_collapse_level' --launch-skip 24 --launch-count 24 -o gpurun_out/aj_build python tools/build_sweep.py --sizes 10 --repeat 1 > gpurun_out/aj_ncu.log 2>&1
ncu -i gpurun_out/aj_build.ncu-rep --page raw --csv > gpurun_out/aj_build_raw.csv 2>/dev/null
rm -f gpurun_out/aj_build.ncu-rep
ncu --set full --clock-control none -k regex:'k_transmit|k_volume' --launch-skip 30 --launch-count 3 -o gpurun_out/aj_media python tools/render_probe.py --scene tests/scenes/media2.pbrt --res 480 320 --spp 16 --reps 1 > gpurun_out/aj_ncu2.log 2>&1
ncu -i gpurun_out/aj_media.ncu-rep --page raw --csv > gpurun_out/aj_media_raw.csv 2>/dev/null
rm -f gpurun_out/aj_media.ncu-rep
ls -la gpurun_out/aj_*
