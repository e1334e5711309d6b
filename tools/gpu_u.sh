#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
: > gpurun_out/u_media.log
for scene in media.pbrt media2.pbrt; do
for lib in prev v4 v5 v6; do
python tools/render_probe.py --scene tests/scenes/$scene --res 960 640 --spp 64 --reps 3 --lib pbrt-minus_b200/libb200pt_$lib.so >> gpurun_out/u_media.log 2>&1
done; done
cat gpurun_out/u_media.log
