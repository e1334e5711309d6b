#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
ROOT=$(pwd)
python tools/make_c4.py --out /tmp/c4 > gpurun_out/i_c4_make.log 2>&1
timeout 1500 python tools/render_stat_compare.py --scene /tmp/c4/c4.pbrt --spp 256 --parts 16 --out $ROOT/gpurun_out --tag i_c4_50M_256spp > gpurun_out/i_c4.log 2>&1
rm -f gpurun_out/*_ref.png gpurun_out/*.f32
tail -c 3000 gpurun_out/i_c4.log
