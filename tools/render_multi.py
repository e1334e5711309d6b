#!/usr/bin/env python3
"""Sample-partitioned multi-GPU render (configs C1 / C3): one process per GPU under torchrun, each rank renders
its sample range of every pixel, one ncclReduce of the film at the end (b200pt_film_reduce).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/render_multi.py [--res W H] [--spp S] [--scene PATH] [--strong]
Weak scaling by default (every rank renders --spp samples; total = spp * N); --strong splits --spp across ranks.
Rank 0 prints one JSON line; time = max over ranks of the device time of render, + the reduce.
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from render_compare import with_resolution  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=os.path.join(ROOT, "tests", "scenes", "cornell-box-specular.pbrt"))
    ap.add_argument("--res", type=int, nargs=2, default=[1920, 1080])
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--strong", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("pbrt-minus_b200")
    total_spp = args.spp if args.strong else args.spp * world
    r = int(round(total_spp ** 0.5))
    assert r * r == total_spp, "total spp %d must be a perfect square for the stratified sampler" % total_spp
    text = with_resolution(open(args.scene).read(), *args.res)
    # the scene is loaded from a file beside the original so that relative PLY / spectrum paths resolve
    sp = os.path.join(os.path.dirname(os.path.abspath(args.scene)), ".render_multi_rank%d.pbrt" % rank)
    try:
        open(sp, "w").write(text)
    except OSError:
        sp = None
    sc = pkg.pbrt_scene.load(sp, spp=total_spp) if sp else pkg.pbrt_scene.load(text=text, spp=total_spp)
    if sp:
        os.remove(sp)
    integ = pkg.WavefrontPathIntegrator(sc, device=local)
    film = integ.create_film()
    b, e = pkg.sample_range(total_spp, rank, world)
    comm = pkg.scene.NcclComm(rank, world, local) if world > 1 else None
    integ.render(film, sample_begin=b, sample_end=min(e, b + 1))     # warm-up (pool allocation, module load)
    if comm is not None:
        film.reduce(comm.comm, root=0)                               # warm-up (NCCL connection set-up)
    film.clear()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    st = integ.render(film, sample_begin=b, sample_end=e)
    t_render = time.perf_counter() - t0
    t1 = time.perf_counter()
    if comm is not None:
        film.reduce(comm.comm, root=0)
    t_reduce = time.perf_counter() - t1
    times = torch.tensor([st["ms_total"], t_render * 1e3, t_reduce * 1e3], device="cuda", dtype=torch.float64)
    rays = torch.tensor([float(st["rays_closest"] + st["rays_shadow"]), float(st["paths"])], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays)
    if rank == 0:
        W, H = film.width, film.height
        img = film.get_pixel_rgb()
        ms = float(times[0] + times[2])
        out = {"scene": os.path.basename(args.scene), "width": W, "height": H, "spp_total": total_spp, "n_gpus": world,
               "scaling": "strong" if args.strong else "weak", "ms_render_max": float(times[0]), "ms_wall_render_max": float(times[1]),
               "ms_film_reduce": float(times[2]), "Msamples_s": float(rays[1]) / ms / 1e3, "Mrays_s": float(rays[0]) / ms / 1e3,
               "film_bytes": W * H * 16, "mean_rgb": [float(v) for v in img.reshape(-1, 3).mean(0)], "nan": int(np.isnan(img).sum())}
        print(json.dumps(out))
        if args.out:
            img.tofile(args.out)
    if comm is not None:
        comm.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
