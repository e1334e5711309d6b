#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/render_probe.py --reps 5 > gpurun_out/r_render.log 2>&1
python tools/render_probe.py --reps 3 --res 1920 1080 --spp 64 >> gpurun_out/r_render.log 2>&1
cat gpurun_out/r_render.log
