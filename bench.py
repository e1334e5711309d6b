#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on its config C2 (the largest single-GPU configuration):
traversal-only bench, synthetic 10M-triangle random soup, 1920x1080 primary + incoherent
diffuse-bounce + shadow rays.  One step = one pass of the hot path (closest-hit on the primary set,
closest-hit on the bounce set, any-hit on the shadow set) over the materialised ray sets.

  python bench.py --gpus N --steps K --warmup W           our arm (libb200pt.so through the C-ABI)
  python bench.py --impl reference ...                    the reference's own HLBVH traversal
                                                          (oracle/_ref/ref_trace: unmodified
                                                          reference objects, its CUDA build --
                                                          the reference has NO CPU path)
Prints ONE JSON line (rank 0).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s (closest-hit + any-hit rays traced per second, C2 traversal-only bench)"
UNIT = "Mrays/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--triangles", type=int, default=10_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip the extra config-C1 render line item")
    ap.add_argument("--cpu-rays", type=int, default=96, help="rays of the bounded CPU-oracle sample")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_workload(pkg, args, scene, rank):
    """materialised ray sets of C2; rank r jitters the primary rays to sample r of a 4x4 stratum
    grid (sample-index sharding: every GPU traces a disjoint sample of every pixel)."""
    wl = pkg.workloads
    o, d, tmax = wl.primary_rays(args.width, args.height, jitter_index=rank)
    prim, t, _ = scene.intersect(o, d, tmax)
    bo, bd, bt = wl.bounce_rays(scene_P(scene), scene_idx(scene), o, d, prim, t, seed=7 + rank)
    so, sd, stm = wl.shadow_rays(scene_P(scene), scene_idx(scene), o, d, prim, t, seed=11 + rank)
    return [("primary", 0, (o, d, tmax)), ("bounce", 0, (bo, bd, bt)), ("shadow", 1, (so, sd, stm))]


def scene_P(scene):
    return scene._keep[1]


def scene_idx(scene):
    return scene._keep[2]


CORNELL = os.path.join(ROOT, "tests", "scenes", "cornell-box-specular.pbrt")


def render_c1_ours(pkg, local, rank, world, spp=16):
    """config C1 (cornell-box-specular, 1080x1080, 16 spp per GPU) through WavefrontPathIntegrator.render:
    an extra line item beside the C2 headline.  N > 1: sample-index sharding, spp_total = 16 * N when that is a
    perfect square (stratified sampler), else the ranks render the same 16-sample image independently."""
    total = spp * world
    r = int(round(total ** 0.5))
    sharded = r * r == total
    sc = pkg.pbrt_scene.load(CORNELL, spp=total if sharded else spp)
    integ = pkg.WavefrontPathIntegrator(sc, device=local)
    film = integ.create_film()
    b, e = pkg.sample_range(total, rank, world) if sharded else (0, spp)
    integ.render(film, sample_begin=b, sample_end=e)      # warm-up
    film.clear()
    st = integ.render(film, sample_begin=b, sample_end=e)
    img = film.get_pixel_rgb()
    out = {"scene": "cornell-box-specular.pbrt 1080x1080", "spp_per_gpu": e - b, "ms": round(st["ms_total"], 3),
           "Msamples_s_per_gpu": round(st["paths"] / st["ms_total"] / 1e3, 2),
           "Mrays_s_per_gpu": round((st["rays_closest"] + st["rays_shadow"]) / st["ms_total"] / 1e3, 1),
           "passes": st["passes"], "kernel_launches": st["kernel_launches"], "mean_rgb": [round(float(v), 6) for v in img.reshape(-1, 3).mean(0)]}
    film.close()
    integ.close()
    return out


def render_c1_reference(td, spp=16):
    """the reference's own CLI (oracle/_ref/pbrt-minus-ref, unmodified hot path) on the same scene"""
    import re
    exe = os.path.join(ROOT, "oracle", "_ref", "pbrt-minus-ref")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/pbrt-minus-ref not prebuilt"}
    t0 = time.time()
    r = subprocess.run([exe, CORNELL, "--spp", str(spp), "--outfile", os.path.join(td, "ref.png")], capture_output=True, text=True,
                       timeout=3000, cwd=td)
    wall = time.time() - t0
    if r.returncode != 0:
        return {"unavailable": (r.stdout + r.stderr)[-200:].replace("\n", " ")}
    m = re.search(r"rendering took\s+([0-9.]+)", r.stdout)
    bvh = re.search(r"BVH building:[^\n]*", r.stdout)
    sec = float(m.group(1)) if m else None
    return {"scene": "cornell-box-specular.pbrt 1080x1080", "spp": spp, "render_s": sec, "wall_s": round(wall, 2),
            "Msamples_s": round(1080 * 1080 * spp / sec / 1e6, 2) if sec else None, "bvh_line": bvh.group(0) if bvh else None}


def run_ours(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module("pbrt-minus_b200")
    lib = pkg.load_library()
    if lib.b200pt_device_count() <= local:
        raise SystemExit("bench.py: no CUDA device; libb200pt has no CPU fallback")
    torch.cuda.set_device(local)

    t0 = time.time()
    P, idx = pkg.workloads.triangle_soup(args.triangles)
    t_gen = time.time() - t0
    t0 = time.time()
    scene = pkg.HLBVH(P, idx, device=local, build=False)
    t_upload = time.time() - t0
    scene.build()            # warm-up build (allocator, module load)
    build = scene.build()    # timed by device events inside the library
    sets = make_workload(pkg, args, scene, rank)
    total_rays = sum(len(s[2][0]) for s in sets)

    # ---- device-resident form: rays already in HBM when the timed region starts ----------------
    stream = torch.cuda.ExternalStream(scene.stream, device=torch.device("cuda", local))
    dev = []
    for name, any_hit, (o, d, tm) in sets:
        rays = scene.pack_rays(o, d, tm)
        rb, hb = scene.alloc(rays.nbytes), scene.alloc(len(o) * 16)
        rb.upload(rays)
        dev.append((name, any_hit, len(o), rb, hb))

    def step_device():
        for name, any_hit, n, rb, hb in dev:
            scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=False)

    def barrier():
        scene.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # visit counts of one pass (untimed, counting kernels) -> algorithmic bytes of the roofline
    scene.set_counting(True)
    counts = {}
    for name, any_hit, n, rb, hb in dev:
        scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=True)
        counts[name] = scene.counters()
    scene.set_counting(False)

    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 + 2 * len(dev) * args.steps)]
    ev[0].record(stream)
    k = 2
    for _ in range(args.steps):
        for name, any_hit, n, rb, hb in dev:   # per-kernel events on the launching stream
            ev[k].record(stream)
            scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=False)
            ev[k + 1].record(stream)
            k += 2
    ev[1].record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = ev[0].elapsed_time(ev[1])
    per_set_ms = {name: 0.0 for name, *_ in dev}
    k = 2
    for _ in range(args.steps):
        for name, *_ in dev:
            per_set_ms[name] += ev[k].elapsed_time(ev[k + 1]) / args.steps
            k += 2
    if world > 1:
        tt = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
        tr = torch.tensor([float(total_rays)], device="cuda", dtype=torch.float64)
        dist.all_reduce(tr)
        job_rays = float(tr.item())
    else:
        job_rays = float(total_rays)
    ms_per_step = ms_total / args.steps
    value = job_rays / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: the reference-facing call (host buffers in pinned memory, copies inside) -------------
    host = []
    for name, any_hit, (o, d, tm) in sets:
        n = len(o)
        bufs = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (o, d, tm)]
        outs = [torch.empty(n, dtype=torch.int32).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory(),
                torch.empty(2 * n, dtype=torch.float32).pin_memory()]
        host.append((any_hit, n, bufs, outs))
    import ctypes as C
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)

    def step_host():
        for any_hit, n, bufs, outs in host:
            rc = lib.b200pt_trace(scene.handle, C.cast(bufs[0].data_ptr(), fp), C.cast(bufs[1].data_ptr(), fp),
                                  C.cast(bufs[2].data_ptr(), fp), n, any_hit, C.cast(outs[0].data_ptr(), ip),
                                  C.cast(outs[1].data_ptr(), fp), C.cast(outs[2].data_ptr(), fp))
            if rc:
                raise RuntimeError(lib.b200pt_last_error())

    for _ in range(max(1, args.warmup)):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()   # each call ends with a stream synchronize (results are on the host)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt.item())
    e2e_value = job_rays * args.steps / e2e_s / 1e6
    h2d = sum(n * 28 for _, n, _, _ in host)
    d2h = sum(n * 16 for _, n, _, _ in host)

    # ---- roofline of the dominant kernel (closest-hit on the bounce set) ------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    dom = max(per_set_ms, key=per_set_ms.get)
    c = counts[dom]
    alg_bytes = 48 * c["rays"] + 80 * c["nodes"] + 48 * c["tris"]
    achieved = alg_bytes / (per_set_ms[dom] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "k_trace<closest> on the %s set" % dom, "achieved": round(achieved, 1),
                "peak": peak, "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback", "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": None,
                "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": round(per_set_ms[dom], 4),
                "nodes_per_ray": round(c["nodes"] / max(1, c["rays"]), 2), "tris_per_ray": round(c["tris"] / max(1, c["rays"]), 2)}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            roofline["traffic"] = json.load(open(prof)).get("k_trace_dram_bytes_per_launch")
        except Exception:
            pass

    out = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (+u32 ids; f64 edge fallback)", "data": "synthetic",
        "config": {"workload": "C2 traversal-only: %d-triangle random soup, %dx%d primary + diffuse-bounce + shadow rays"
                               % (args.triangles, args.width, args.height),
                   "rays_per_step_per_gpu": total_rays, "ray_sets": {n: c for n, _, c, _, _ in dev},
                   "l2": "inputs larger than L2 (BVH8 nodes + triangles %.0f MB, rays %.0f MB)"
                         % ((build["n_bvh8_nodes"] * 80 + args.triangles * 48) / 1e6, total_rays * 32 / 1e6),
                   "parallelism": "replicated scene, ray sets sharded by sample index, no data-path collective"},
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": 3 * args.steps,
        "clocks": clocks,
        "roofline": roofline,
        "per_set_ms": {k2: round(v, 4) for k2, v in per_set_ms.items()},
        "per_set_mrays": {n: round(cn / per_set_ms[n] / 1e3, 1) for n, _, cn, _, _ in dev},
        "bvh_build": {k2: (round(v, 3) if isinstance(v, float) else v) for k2, v in build.items()},
        "setup_s": {"generate": round(t_gen, 2), "upload": round(t_upload, 2)},
    }
    if not args.no_render:
        rc1 = render_c1_ours(pkg, local, rank, world)
        if world > 1:
            tt = torch.tensor([rc1["ms"]], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            rc1["ms_max_over_ranks"] = round(float(tt.item()), 3)
            rc1["Msamples_s_job"] = round(1080 * 1080 * rc1["spp_per_gpu"] * world / float(tt.item()) / 1e3, 2)
        out["render_c1"] = rc1
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args, P, idx, sets)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(args, P, idx, sets):
    """the oracle port (brute-force closest hit with the reference's triangle arithmetic) on the host
    cores, on a bounded sample of the same workload"""
    from oracle import pyoracle
    n = args.cpu_rays
    o, d, tm = sets[0][2]
    pick = np.linspace(0, len(o) - 1, n).astype(np.int64)
    cores = pyoracle.lib().oracle_get_threads()
    t0 = time.perf_counter()
    pyoracle.closest_hit(P, idx, o[pick], d[pick], tm[pick])
    dt = time.perf_counter() - t0
    return {"value": round(n / dt / 1e6, 9), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d primary rays of the same set, brute force over all %d triangles (oracle/oracle.c), %.1f s"
                      % (n, len(idx), dt)}


def run_reference(args):
    """The reference's own implementation of this path, unmodified: its HLBVH build + traversal
    (oracle/_ref/ref_trace links the reference objects compiled for sm_100).  The reference has no
    CPU path (every allocation is cudaMallocManaged, rendering is reachable only through kernels),
    so its CUDA build on this box's B200 is the baseline; its BVH build runs on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_trace")
    base = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    if not os.path.exists(exe):
        base["unavailable"] = "oracle/_ref/ref_trace not prebuilt (run __graft_entry__.build() where /root/reference exists)"
        print(json.dumps(base))
        return
    pkg = importlib.import_module("pbrt-minus_b200")   # workload generators only (numpy); no kernels of ours
    wl = pkg.workloads
    P, idx = wl.triangle_soup(args.triangles)
    cores = os.cpu_count()
    with tempfile.TemporaryDirectory(dir="/tmp") as td:
        soup = os.path.join(td, "soup.bin")
        wl.write_soup(soup, P, idx)
        o, d, tmax = wl.primary_rays(args.width, args.height)
        wl.write_rays(os.path.join(td, "primary.rays"), o, d, tmax)
        # pass 1: the reference's own primary hits define the bounce / shadow sets
        r = subprocess.run([exe, soup, "closest:%s:%s" % (os.path.join(td, "primary.rays"), os.path.join(td, "primary.out"))],
                           capture_output=True, text=True, timeout=3000)
        if r.returncode != 0:
            base["unavailable"] = "ref_trace failed: " + (r.stdout + r.stderr)[-300:].replace("\n", " ")
            print(json.dumps(base))
            return
        prim, t = wl.read_hits(os.path.join(td, "primary.out"), len(o))
        bo, bd, bt = wl.bounce_rays(P, idx, o, d, prim, t, seed=7)
        so, sd, stm = wl.shadow_rays(P, idx, o, d, prim, t, seed=11)
        wl.write_rays(os.path.join(td, "bounce.rays"), bo, bd, bt)
        wl.write_rays(os.path.join(td, "shadow.rays"), so, sd, stm)
        t0 = time.time()
        r = subprocess.run([exe, soup, "--warmup", str(args.warmup), "--steps", str(args.steps),
                            "closest:%s:-" % os.path.join(td, "primary.rays"), "closest:%s:-" % os.path.join(td, "bounce.rays"),
                            "any:%s:-" % os.path.join(td, "shadow.rays")], capture_output=True, text=True, timeout=3000)
        wall = time.time() - t0
        ref_render = None if args.no_render else render_c1_reference(td)
    if r.returncode != 0:
        base["unavailable"] = "ref_trace failed: " + (r.stdout + r.stderr)[-300:].replace("\n", " ")
        print(json.dumps(base))
        return
    info = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    build_line = [l for l in r.stdout.splitlines() if l.startswith("BVH building")]
    ms_per_step = sum(s["ms"] for s in info["sets"])
    rays = sum(s["rays"] for s in info["sets"])
    value = rays / (ms_per_step * 1e-3) / 1e6
    base.update({
        "value": round(value, 2), "ms_per_step": round(ms_per_step, 3),
        "config": {"workload": "C2 traversal-only: %d-triangle random soup, %dx%d primary + diffuse-bounce + shadow rays"
                               % (args.triangles, args.width, args.height), "rays_per_step_per_gpu": rays,
                   "ray_sets": {n: s["rays"] for n, s in zip(("primary", "bounce", "shadow"), info["sets"])},
                   "parallelism": "single GPU (the reference is single-device)"},
        "e2e": {"value": round(value, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "cpu_baseline": {"value": round(value, 2), "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "full workload; the reference has no CPU path, so this is its own CUDA build (sm_100) on GPU 0 "
                                   "with its BVH build on the %d host cores" % cores},
        "per_set_ms": {n: round(s["ms"], 3) for n, s in zip(("primary", "bounce", "shadow"), info["sets"])},
        "reference_bvh_build_s": info["bvh_build_s"], "reference_load_s": info["load_s"],
        "reference_build_line": build_line[-1] if build_line else None, "wall_s": round(wall, 1),
        "render_c1": ref_render,
    })
    print(json.dumps(base))


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
