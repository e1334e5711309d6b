#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/t_pytest_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/t_pytest_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/t_pytest_all.log 2>&1; echo "all rc=$?"; tail -5 gpurun_out/t_pytest_all.log
