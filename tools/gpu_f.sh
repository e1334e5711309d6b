#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 1200 python -m pytest tests/test_trace_gpu.py -m gpu -x -q > gpurun_out/f_pytest_trace.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest_trace.log
tail -3 gpurun_out/f_pytest_trace.log
timeout 600 python tools/build_sweep.py --sizes 1 10 100 --check --out gpurun_out/f_build_sweep.json > gpurun_out/f_build_sweep.log 2>&1
tail -5 gpurun_out/f_build_sweep.log | cut -c1-600
timeout 300 python -m pytest tests/test_render_gpu.py -m gpu -x -q -k "world2 or cornell_16spp" > gpurun_out/f_pytest_mp.log 2>&1; tail -3 gpurun_out/f_pytest_mp.log
# C4 at size: 50 M triangles / 100 k lights, 256 spp, statistical gate with the reference's outlier pixels excluded and listed
python tools/make_c4.py --out /tmp/c4 > gpurun_out/f_c4_make.log 2>&1
timeout 1500 python tools/render_stat_compare.py --scene /tmp/c4/c4.pbrt --spp 256 --parts 16 --out gpurun_out --tag f_c4_50M_256spp > gpurun_out/f_c4.log 2>&1
rm -f gpurun_out/*_ref.png gpurun_out/*.f32
tail -c 1500 gpurun_out/f_c4.log
