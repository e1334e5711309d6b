#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python tools/nan_probe.py --scene tests/scenes/media2.pbrt --res 960 640 --spp 64 --reference > gpurun_out/v_nan_media2.json 2> gpurun_out/v_nan_media2.err
python tools/nan_probe.py --scene tests/scenes/media2.pbrt --res 960 640 --spp 64 --lib pbrt-minus_b200/libb200pt_prev.so > gpurun_out/v_nan_media2_prev.json 2>> gpurun_out/v_nan_media2.err
python tools/nan_probe.py --scene tests/scenes/media.pbrt --res 960 640 --spp 64 > gpurun_out/v_nan_media.json 2>> gpurun_out/v_nan_media2.err
cat gpurun_out/v_nan_*.json; tail -3 gpurun_out/v_nan_media2.err
timeout 900 python -m pytest tests/test_render_gpu.py -m gpu -x -q -k "media or volume" > gpurun_out/v_pytest_media.log 2>&1; tail -3 gpurun_out/v_pytest_media.log
