#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python tools/render_probe.py > gpurun_out/e_render.log 2>&1
python tools/render_probe.py --res 1920 1080 --spp 64 >> gpurun_out/e_render.log 2>&1
timeout 900 python -m pytest tests/test_render_gpu.py -m gpu -x -q > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest.log
for c in 65536 131072 262144 524288 1048576; do
  B200PT_TRACE_CHUNK=$c python bench.py --steps 10 --warmup 3 --no-render --no-cpu-baseline 2>/dev/null | python -c "import sys,json; o=json.loads(sys.stdin.read()); print('chunk',$c,'value',o['value'],'e2e',o['e2e']['value'], o['e2e']['ids_equal_device_resident'])" >> gpurun_out/e_chunk_sweep.log 2>&1
done
B200PT_TRACE_ONE_STREAM=1 python bench.py --steps 10 --warmup 3 --no-render --no-cpu-baseline 2>/dev/null | python -c "import sys,json; o=json.loads(sys.stdin.read()); print('one-stream default chunk value',o['value'],'e2e',o['e2e']['value'])" >> gpurun_out/e_chunk_sweep.log 2>&1
ncu --set full --clock-control none -k regex:'k_collapse_level|k_refit|k_sort_downsweep|k_sort_upsweep|k_emit_hierarchy|k_morton|k_centroid' --launch-skip 23 --launch-count 23 -o gpurun_out/e_build_full python bench.py --steps 1 --warmup 1 --no-render --no-cpu-baseline > gpurun_out/e_ncu_build.log 2>&1
ncu -i gpurun_out/e_build_full.ncu-rep --page raw --csv > gpurun_out/e_build_full_raw.csv 2>/dev/null
rm -f gpurun_out/e_build_full.ncu-rep
ncu --set full --clock-control none -k regex:'k_trace' --launch-skip 12 --launch-count 3 -o gpurun_out/e_trace_full python bench.py --steps 2 --warmup 3 --no-render --no-cpu-baseline > gpurun_out/e_ncu_trace.log 2>&1
ncu -i gpurun_out/e_trace_full.ncu-rep --page raw --csv > gpurun_out/e_trace_full_raw.csv 2>/dev/null
ncu -i gpurun_out/e_trace_full.ncu-rep --page details > gpurun_out/e_trace_full_details.txt 2>/dev/null
rm -f gpurun_out/e_trace_full.ncu-rep
cat gpurun_out/e_render.log gpurun_out/e_chunk_sweep.log; tail -3 gpurun_out/e_pytest.log; ls -la gpurun_out
