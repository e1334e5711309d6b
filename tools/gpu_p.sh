#!/bin/bash
# N-GPU runs: bench.py under torchrun and the e2e scaling probe ($1 = N)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=$1
nvidia-smi topo -m > gpurun_out/p_topo_n$N.txt 2>&1; nproc >> gpurun_out/p_topo_n$N.txt; numactl -H >> gpurun_out/p_topo_n$N.txt 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/p_bench_n$N.json 2> gpurun_out/p_bench_n$N.err; echo "bench rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 tools/e2e_scaling_probe.py > gpurun_out/p_e2e_probe_n$N.json 2> gpurun_out/p_e2e_probe_n$N.err; echo "probe rc=$?"
tail -1 gpurun_out/p_e2e_probe_n$N.json | cut -c1-1500
python - <<PY
import json
o=json.loads(open('gpurun_out/p_bench_n$N.json').read().strip().splitlines()[-1])
print('value',o['value'],'e2e',o['e2e']['value'],'render',o['render']['value'],o['render']['ms_per_step'],o['render']['reduce_ms'],o['render']['film_check']['relMSE'])
PY
