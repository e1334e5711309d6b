#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 900 python -m pytest tests/test_render_gpu.py -m gpu -x -q -k "instanc" > gpurun_out/aa_pytest_inst.log 2>&1; tail -12 gpurun_out/aa_pytest_inst.log
python tools/render_probe.py --scene tools/inst_many.pbrt --res 960 640 --spp 16 --reps 3 --lib pbrt-minus_b200/libb200pt_prev.so > gpurun_out/aa_inst.log 2>&1
python tools/render_probe.py --scene tools/inst_many.pbrt --res 960 640 --spp 16 --reps 3 >> gpurun_out/aa_inst.log 2>&1
grep '^{' gpurun_out/aa_inst.log || tail -5 gpurun_out/aa_inst.log
