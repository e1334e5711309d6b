#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 900 python -m pytest tests/test_render_gpu.py -m gpu -x -q -k "instanc or normals or cornell_16spp" > gpurun_out/g_pytest_inst.log 2>&1; tail -25 gpurun_out/g_pytest_inst.log
python tools/measure_golden.py > gpurun_out/g_measure.log 2>&1; grep -E "instances|normals" gpurun_out/g_measure.log | cut -c1-200
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/g_pytest.log 2>&1; tail -5 gpurun_out/g_pytest.log
