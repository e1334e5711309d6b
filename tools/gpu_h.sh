#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/h_bench_n2.json 2> gpurun_out/h_bench_n2.err; echo "rc=$?" >> gpurun_out/h_bench_n2.err
tail -5 gpurun_out/h_bench_n2.err; cat gpurun_out/h_bench_n2.json | cut -c1-300
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/h_bench_ref_n2.json 2> gpurun_out/h_bench_ref_n2.err; echo "rc=$?" >> gpurun_out/h_bench_ref_n2.err; tail -2 gpurun_out/h_bench_ref_n2.err; cut -c1-200 gpurun_out/h_bench_ref_n2.json
