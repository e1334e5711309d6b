#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 600 python -m pytest tests/test_trace_gpu.py -m gpu -x -q -k "division" > gpurun_out/y_pytest_div.log 2>&1; tail -8 gpurun_out/y_pytest_div.log
python tools/render_probe.py --reps 5 --lib pbrt-minus_b200/libb200pt_prev.so > gpurun_out/y_render.log 2>&1
python tools/render_probe.py --reps 5 >> gpurun_out/y_render.log 2>&1
python tools/render_probe.py --reps 3 --res 1920 1080 --spp 64 --lib pbrt-minus_b200/libb200pt_prev.so >> gpurun_out/y_render.log 2>&1
python tools/render_probe.py --reps 3 --res 1920 1080 --spp 64 >> gpurun_out/y_render.log 2>&1
cat gpurun_out/y_render.log
timeout 1500 python -m pytest tests/test_render_gpu.py tests/test_bridge_gpu.py -m gpu -x -q > gpurun_out/y_pytest_render.log 2>&1; tail -8 gpurun_out/y_pytest_render.log
