#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 900 python bench.py --impl reference > gpurun_out/o_bench_ref.json 2> gpurun_out/o_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/o_bench_ours.json 2> gpurun_out/o_bench_ours.err; echo "ours rc=$?"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/o_pytest.log 2>&1; tail -3 gpurun_out/o_pytest.log
python __graft_entry__.py --smoke > gpurun_out/o_smoke.log 2>&1; tail -2 gpurun_out/o_smoke.log
cut -c1-400 gpurun_out/o_bench_ours.json
