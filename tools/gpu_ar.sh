#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
timeout 300 python -m pytest tests/test_trace_gpu.py tests/test_abi.py -m gpu -x -q -k "device_batch or abi or header" > gpurun_out/ar_pytest.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/ar_pytest.log
python bench.py --steps 30 --warmup 3 --no-render > gpurun_out/ar_bench.json 2> gpurun_out/ar_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/ar_bench.err
python -c "
import json
d=json.loads(open('gpurun_out/ar_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['ms_per_step_one_stream'], d['per_set_ms'], d['e2e']['value'], d['roofline']['frac'], d.get('parity'))
"
