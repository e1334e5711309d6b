#!/bin/bash
# round-2 GPU batch A: regression tests, k_trace knob sweep, both bench arms, new goldens from the reference
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
nproc >> gpurun_out/a_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
timeout 600 python tools/trace_tune.py --out gpurun_out/a_trace_tune.json > gpurun_out/a_trace_tune.log 2>&1; echo "tune rc=$?" >> gpurun_out/a_trace_tune.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench_ours.json 2> gpurun_out/a_bench_ours.err; echo "bench rc=$?" >> gpurun_out/a_bench_ours.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/a_bench_ref.json 2> gpurun_out/a_bench_ref.err; echo "ref rc=$?" >> gpurun_out/a_bench_ref.err
timeout 900 python tools/make_golden.py round2 > gpurun_out/a_golden.log 2>&1; echo "golden rc=$?" >> gpurun_out/a_golden.log
tail -3 gpurun_out/a_pytest.log; tail -2 gpurun_out/a_trace_tune.log; cat gpurun_out/a_bench_ours.json | cut -c1-600; tail -2 gpurun_out/a_golden.log
