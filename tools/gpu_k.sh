#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 900 python -m pytest tests/test_bridge_gpu.py -m gpu -x -q > gpurun_out/k_pytest_bridge.log 2>&1; tail -40 gpurun_out/k_pytest_bridge.log
