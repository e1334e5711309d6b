#!/usr/bin/env python3
"""What limits the host-buffer traversal call (bench.py's `e2e`) when N ranks run it at once?  Under torchrun, every rank measures,
with all ranks active at the same moment: pinned H2D bandwidth on the bytes of one C2 step, pinned D2H bandwidth, both directions
together, the device-resident trace time, and the b200pt_trace_batch call.  Rank 0 prints one JSON line with the per-rank table.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/e2e_scaling_probe.py
"""
import argparse
import ctypes as C
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--triangles", type=int, default=10_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("pbrt-minus_b200")
    lib = pkg.load_library()
    P, idx = pkg.workloads.triangle_soup(args.triangles)
    scene = pkg.HLBVH(P, idx, device=local, build=True)
    sets = bench.make_workload(pkg, args, scene, rank)
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    host = []
    for name, any_hit, (o, d, tm) in sets:
        n = len(o)
        bufs = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (o, d, tm)]
        outs = [torch.empty(n, dtype=torch.int32).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory(),
                torch.empty(2 * n, dtype=torch.float32).pin_memory()]
        host.append((any_hit, n, bufs, outs))
    batch = (pkg.api.RaySet * len(host))()
    for i, (any_hit, n, bufs, outs) in enumerate(host):
        b = batch[i]
        b.o_xyz, b.d_xyz, b.tmax = (C.cast(x.data_ptr(), fp) for x in bufs)
        b.n, b.any_hit = n, any_hit
        b.prim_out = C.cast(outs[0].data_ptr(), ip)
        b.t_out = None if any_hit else C.cast(outs[1].data_ptr(), fp)
        b.b1b2_out = None if any_hit else C.cast(outs[2].data_ptr(), fp)
    dev_in = [[torch.empty_like(x, device="cuda") for x in bufs] for _, _, bufs, _ in host]
    dev_out = [[torch.empty_like(x, device="cuda") for x in outs] for _, _, _, outs in host]
    h2d_bytes = sum(x.numel() * x.element_size() for _, _, bufs, _ in host for x in bufs)
    d2h_bytes = sum(outs[0].numel() * 4 if any_hit else sum(x.numel() * x.element_size() for x in outs) for any_hit, _, _, outs in host)
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn):
        fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / args.reps
        barrier()
        return dt

    def h2d():
        with torch.cuda.stream(s_in):
            for (_, _, bufs, _), dv in zip(host, dev_in):
                for a, b in zip(dv, bufs):
                    a.copy_(b, non_blocking=True)

    def d2h():
        with torch.cuda.stream(s_out):
            for (any_hit, _, _, outs), dv in zip(host, dev_out):
                for k, (a, b) in enumerate(zip(outs, dv)):
                    if any_hit and k > 0:
                        continue
                    a.copy_(b, non_blocking=True)

    def both():
        h2d()
        d2h()

    rays = scene.pack_rays
    dev = []
    for name, any_hit, (o, d, tm) in sets:
        r = scene.pack_rays(o, d, tm)
        rb, hb = scene.alloc(r.nbytes), scene.alloc(len(o) * 16)
        rb.upload(r)
        dev.append((any_hit, len(o), rb, hb))

    def trace_dev():
        for any_hit, n, rb, hb in dev:
            scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=False)
        scene.sync()

    def call():
        rc = lib.b200pt_trace_batch(scene.handle, batch, len(host))
        assert rc == 0, lib.b200pt_last_error()

    t_h2d, t_d2h, t_both, t_dev, t_call = timed(h2d), timed(d2h), timed(both), timed(trace_dev), timed(call)
    row = [h2d_bytes / t_h2d / 1e9, d2h_bytes / t_d2h / 1e9, (h2d_bytes + d2h_bytes) / t_both / 1e9, t_dev * 1e3, t_call * 1e3,
           sum(n for _, n, _, _ in host) / t_call / 1e6, float(len(os.sched_getaffinity(0)))]
    t = torch.tensor(row, device="cuda", dtype=torch.float64)
    rows = [torch.zeros_like(t) for _ in range(world)]
    if world > 1:
        dist.all_gather(rows, t)
    else:
        rows = [t]
    if rank == 0:
        keys = ["h2d_GBs", "d2h_GBs", "both_directions_GBs", "trace_device_ms", "call_ms", "call_Mrays_s", "cpus_in_affinity_mask"]
        table = [{k: round(float(v), 2) for k, v in zip(keys, r.tolist())} for r in rows]
        out = {"n_gpus": world, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes, "per_rank": table,
               "sum_h2d_GBs": round(sum(r["h2d_GBs"] for r in table), 1), "sum_both_GBs": round(sum(r["both_directions_GBs"] for r in table), 1),
               "job_call_Mrays_s": round(sum(n for _, n, _, _ in host) * world / max(r["call_ms"] for r in table) / 1e3, 1)}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
