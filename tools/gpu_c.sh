#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/render_probe.py --sweep > gpurun_out/c_render_sweep.log 2>&1
for c in 131072 262144 524288 1048576 4194304; do
  B200PT_TRACE_CHUNK=$c python bench.py --steps 10 --warmup 3 --no-render --no-cpu-baseline 2>/dev/null | python -c "import sys,json; o=json.loads(sys.stdin.read()); print('chunk',$c,'value',o['value'],'e2e',o['e2e']['value'])" >> gpurun_out/c_chunk_sweep.log 2>&1
done
# launch list + full capture of one steady-state pass of the C1 render
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c_render_launches.csv python tools/render_probe.py --reps 1 > gpurun_out/c_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_logic|k_shade|k_trace_wf|k_generate|k_emit' --launch-skip 200 --launch-count 8 -o gpurun_out/c_render_full python tools/render_probe.py --reps 1 > gpurun_out/c_ncu2.log 2>&1
ncu -i gpurun_out/c_render_full.ncu-rep --page raw --csv > gpurun_out/c_render_full_raw.csv 2>/dev/null
cat gpurun_out/c_render_sweep.log gpurun_out/c_chunk_sweep.log
