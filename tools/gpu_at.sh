#!/bin/bash
# C4 at size with the current code (ours only: the reference's 256-spp film of this scene was compared in profiles/r02_c4_50M_256spp_vs_reference.json)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python tools/make_c4.py --out /tmp/c4 > gpurun_out/at_c4_make.log 2>&1; tail -2 gpurun_out/at_c4_make.log
python tools/render_probe.py --scene /tmp/c4/c4.pbrt --res 1920 1080 --spp 256 --reps 2 > gpurun_out/at_c4.log 2>&1; tail -2 gpurun_out/at_c4.log
