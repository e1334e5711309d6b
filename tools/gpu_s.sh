#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 600 python -m pytest tests/test_trace_gpu.py -m gpu -x -q -k "conservative" > gpurun_out/s_pytest_validate.log 2>&1; tail -5 gpurun_out/s_pytest_validate.log
timeout 600 python -m pytest tests/test_render_gpu.py -m gpu -x -q -k "small_scene or cornell or spherelight or pool_size" > gpurun_out/s_pytest_small.log 2>&1; tail -5 gpurun_out/s_pytest_small.log
python tools/render_probe.py --reps 5 --small-max 0,64 > gpurun_out/s_render.log 2>&1
python tools/render_probe.py --reps 3 --res 1920 1080 --spp 64 --small-max 0,64 >> gpurun_out/s_render.log 2>&1
cat gpurun_out/s_render.log
