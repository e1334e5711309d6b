#!/bin/bash
# rebuilds that reuse the previous build's buffers: the new test, the build sweep (wall time per rebuild), the build / trace tests
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python -m pytest tests/test_trace_gpu.py -q -m gpu -s -k "rebuild or conservative or packed or morton or instance" 2>&1 | grep -v '^$' | tail -6
python tools/build_sweep.py --sizes 1 10 100 --check > gpurun_out/ba_sweep.log 2>&1; grep -o '"n_triangles": [0-9]*\|"first_build_wall_ms": [0-9.]*\|"median_wall_ms": [0-9.]*\|"ms_total": [0-9.]*' gpurun_out/ba_sweep.log | paste -sd' ' | fold -w 200; tail -2 gpurun_out/ba_sweep.log | cut -c1-300
python -m pytest tests/test_render_gpu.py -q -m gpu -k "instance or cornell_16spp or multi" 2>&1 | tail -2
