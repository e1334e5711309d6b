#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 600 python -m pytest tests/test_trace_gpu.py -m gpu -x -q -k "division" > gpurun_out/z_pytest_div.log 2>&1; tail -3 gpurun_out/z_pytest_div.log
python tools/render_probe.py --reps 5 > gpurun_out/z_render.log 2>&1
python tools/render_probe.py --reps 3 --res 1920 1080 --spp 64 >> gpurun_out/z_render.log 2>&1
grep '^{' gpurun_out/z_render.log
