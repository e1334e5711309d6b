#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
timeout 600 python tools/measure_golden.py > gpurun_out/b_measure.log 2>&1; echo "measure rc=$?" >> gpurun_out/b_measure.log
tail -15 gpurun_out/b_pytest.log; tail -3 gpurun_out/b_measure.log
