#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/trace_tune.py --refill 4 --tri-steps 1,3 --stash 0,1 --out gpurun_out/q_trace_tune.json > gpurun_out/q_trace_tune.log 2>&1; cat gpurun_out/q_trace_tune.log | cut -c1-330
