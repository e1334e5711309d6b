#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric ("Msamples/s and Mrays/s") on its configs.

Headline line (`value`, Mrays/s): config C2, the largest single-GPU configuration -- traversal-only bench, synthetic
10M-triangle random soup, 1920x1080 primary + incoherent diffuse-bounce + shadow rays.  One step = one pass of the hot
path (closest-hit on the primary set, closest-hit on the bounce set, any-hit on the shadow set) over the materialised ray
sets, issued as ONE b200pt_trace_device_batch call (consecutive sets on two traversal streams); the per-kernel durations of the
roofline come from a second pass with one launch per set.  `parity` = hit ids of that step against the oracle's host BVH over
the FULL ray sets.

Second timed leg (`render`, Msamples/s): config C3's shape -- cornell-box-specular at 1920x1080, sample-partitioned across
the N GPUs with ONE ncclReduce of the film inside the timed region (strong scaling), its own roofline and e2e.

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
    ap.add_argument("--no-render", action="store_true", help="skip the render leg (C3 sample-partitioned cornell + film reduce)")
    ap.add_argument("--cpu-rays", type=int, default=0, help="cap on the rays per set of the CPU leg (0 = the full ray sets)")
    ap.add_argument("--render-width", type=int, default=1920)
    ap.add_argument("--render-height", type=int, default=1080)
    ap.add_argument("--render-spp", type=int, default=256, help="total samples per pixel of the render leg (perfect square)")
    ap.add_argument("--render-steps", type=int, default=3)
    ap.add_argument("--render-warmup", type=int, default=1)
    ap.add_argument("--ref-render-spp", type=int, default=16, help="reference arm: bounded sample of the render leg")
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
WORKLOAD = "C2 traversal-only: %d-triangle random soup, %dx%d primary + diffuse-bounce + shadow rays"
RENDER_WORKLOAD = "C3 cornell-box-specular at %dx%d, %d spp in total, sample-partitioned across the GPUs, one NCCL film reduce"


def c2_config(args, ray_sets):
    """the SAME dict in both arms (the driver compares them): workload and ray counts only"""
    return {"workload": WORKLOAD % (args.triangles, args.width, args.height), "rays_per_step_per_gpu": int(sum(ray_sets.values())),
            "ray_sets": {k: int(v) for k, v in ray_sets.items()}}


def cornell_text(width, height):
    import re
    text = open(CORNELL).read()
    text = re.sub(r'("integer xresolution"\s*\[\s*)\d+', r"\g<1>%d" % width, text)
    return re.sub(r'("integer yresolution"\s*\[\s*)\d+', r"\g<1>%d" % height, text)


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "MEASURED_PEAKS.json"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def render_bytes(st):
    """algorithmic bytes of one render: generate 144 B/path, logic + shade 448 B per
    path segment (= closest-hit ray), trace 48 B per ray of either kind, film 32 B per finished path"""
    return 144 * st["paths"] + 448 * st["rays_closest"] + 48 * (st["rays_closest"] + st["rays_shadow"]) + 32 * st["paths"]


def render_leg_ours(pkg, args, local, rank, world, dist, torch):
    """The north-star multi-GPU path: cornell at 1920x1080, --render-spp samples per pixel in total, rank r renders
    sample_idx in sample_range(total, r, N) of every pixel (sampler configured for the total: strata stay global),
    then ONE ncclReduce of the film onto rank 0 -- inside the timed region -- and rank 0 reads the film.  Strong scaling."""
    W, H, spp = args.render_width, args.render_height, args.render_spp
    sc = pkg.pbrt_scene.load(text=cornell_text(W, H), spp=spp)
    integ = pkg.WavefrontPathIntegrator(sc, device=local)
    film = integ.create_film()
    b, e = pkg.sample_range(spp, rank, world)
    comm = pkg.scene.NcclComm(rank, world, local) if world > 1 else None
    stream = torch.cuda.ExternalStream(integ.accel.stream, device=torch.device("cuda", local))

    def barrier():
        integ.accel.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step():
        film.clear()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(stream)
        st = integ.render(film, sample_begin=b, sample_end=e)
        e1.record(stream)
        if comm is not None:
            film.reduce(comm.comm, root=0)       # ncclReduce(sum) of 16*W*H bytes on the scene's stream
        e2.record(stream)
        integ.accel.sync()
        return st, e0.elapsed_time(e1), e1.elapsed_time(e2)

    for _ in range(max(1, args.render_warmup)):
        step()
    barrier()
    ms_sum = red_sum = e2e_s = 0.0
    st = None
    img = None
    for _ in range(args.render_steps):
        barrier()
        t0 = time.perf_counter()
        st, ms_r, ms_red = step()
        if rank == 0:
            img = film.get_pixel_rgb()           # e2e: the film comes back to the host (D2H of the resolved RGB)
        t1 = time.perf_counter() - t0
        vals = torch.tensor([ms_r + ms_red, ms_red, t1], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        ms_sum += float(vals[0])
        red_sum += float(vals[1])
        e2e_s += float(vals[2])
    K = args.render_steps
    stats = torch.tensor([float(st["paths"]), float(st["rays_closest"]), float(st["rays_shadow"])], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(stats)
    job = {"paths": float(stats[0]), "rays_closest": float(stats[1]), "rays_shadow": float(stats[2])}
    ms = ms_sum / K
    pk, pk_src = peaks()
    hbm = float(pk.get("hbm_gbs", 6650.0))
    gbs = render_bytes(job) / (ms * 1e-3) / 1e9 / world      # per GPU
    out = {
        "metric": "Msamples/s (camera paths per second: render + film reduce, scene load and BVH build excluded)",
        "value": round(job["paths"] / ms / 1e3, 2), "unit": "Msamples/s", "scaling": "strong", "n_gpus": world,
        "steps": K, "warmup": max(1, args.render_warmup), "ms_per_step": round(ms, 3), "reduce_ms": round(red_sum / K, 3),
        "config": {"workload": RENDER_WORKLOAD % (W, H, spp), "spp_total": spp, "maxdepth": int(sc.max_depth), "sampler": sc.sampler},
        "spp_per_gpu": e - b, "Mrays_s": round((job["rays_closest"] + job["rays_shadow"]) / ms / 1e3, 1),
        "passes": st["passes"], "kernel_launches_per_step": st["kernel_launches"],
        "collective": ("ncclReduce(sum, fp32) of %d bytes inside the timed region" % (W * H * 16)) if world > 1 else "none at N = 1",
        "e2e": {"value": round(job["paths"] * K / e2e_s / 1e6, 2), "unit": "Msamples/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": W * H * 12, "note": "render + reduce + b200pt_film_read on rank 0, wall clock, max over ranks"},
        "roofline": {"bound": "hbm", "kernel": "the whole wavefront pass sequence (generate, trace, logic, emit, shade, shadow trace)",
                     "achieved": round(gbs, 1), "peak": hbm, "peak_source": pk_src, "unit": "GB/s", "frac": round(gbs / hbm, 4),
                     "traffic": None, "algorithmic_bytes_per_step": int(render_bytes(job)),
                     "note": "per GPU; 144 B/path + 448 B/segment + 48 B/ray + 32 B/path (SURVEY 8d)"},
    }
    # ---- in-run check: the sample-partitioned (and reduced) film equals the one-GPU film -------------------------------
    if world > 1:
        if rank == 0:
            film2 = integ.create_film()
            integ.render(film2, sample_begin=0, sample_end=spp)
            ref = film2.get_pixel_rgb()
            film2.close()
            out["film_check"] = film_compare(img, ref, "N-GPU reduced film vs the same render on rank 0 alone")
    else:
        film.clear()
        half = pkg.sample_range(spp, 0, 2)[1]
        integ.render(film, sample_begin=0, sample_end=half)
        integ.render(film, sample_begin=half, sample_end=spp)
        out["film_check"] = film_compare(film.get_pixel_rgb(), img, "two sample ranges accumulated into one film vs one render")
    film.close()
    integ.close()
    if comm is not None:
        comm.close()
    return out


def film_compare(a, b, what):
    rel = float(np.mean((a.astype(np.float64) - b) ** 2 / (b.astype(np.float64) ** 2 + 1e-4)))
    return {"what": what, "relMSE": rel, "max_abs": float(np.abs(a - b).max()), "mean_a": [round(float(v), 6) for v in a.reshape(-1, 3).mean(0)],
            "mean_b": [round(float(v), 6) for v in b.reshape(-1, 3).mean(0)], "ok": bool(rel <= 1e-8)}


def render_c1_ours(pkg, local):
    """config C1 (the reference's shipped scene as it is: 1080x1080, 16 spp) on one GPU, beside the reference's own run"""
    sc = pkg.pbrt_scene.load(CORNELL, spp=16)
    integ = pkg.WavefrontPathIntegrator(sc, device=local)
    film = integ.create_film()
    integ.render(film)      # warm-up
    ms = []
    for _ in range(3):
        film.clear()
        st = integ.render(film)
        ms.append(st["ms_total"])
    m = float(np.median(ms))
    out = {"scene": "cornell-box-specular.pbrt 1080x1080", "spp": 16, "ms": round(m, 3), "Msamples_s": round(st["paths"] / m / 1e3, 2),
           "Mrays_s": round((st["rays_closest"] + st["rays_shadow"]) / m / 1e3, 1), "passes": st["passes"],
           "kernel_launches": st["kernel_launches"], "hbm_frac": round(render_bytes(st) / (m * 1e-3) / 1e9 / float(peaks()[0].get("hbm_gbs", 6650.0)), 4)}
    film.close()
    integ.close()
    return out


def render_reference(td, text, name, spp, npix):
    """the reference's own CLI (oracle/_ref/pbrt-minus-ref, unmodified hot path) on the same scene text"""
    import re
    exe = os.path.join(ROOT, "oracle", "_ref", "pbrt-minus-ref")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/pbrt-minus-ref not prebuilt"}
    scene = os.path.join(td, name + ".pbrt")
    open(scene, "w").write(text)
    t0 = time.time()
    r = subprocess.run([exe, scene, "--spp", str(spp), "--outfile", os.path.join(td, name + ".png")], capture_output=True, text=True,
                       timeout=3000, cwd=td)
    wall = time.time() - t0
    if r.returncode != 0:
        return {"unavailable": (r.stdout + r.stderr)[-200:].replace("\n", " ")}
    # the integrator's own Timer table (integrators/wavefront_path.cu:730-734, util/timer.h:31-50) prints the loop time with
    # 0.01 s resolution; "rendering took" (scene_builder.cu:1059) has 0.1 s and includes the PNG write
    m = re.search(r"wavefront pathtracing benchmark:\s*total time:\s*([0-9.]+)", r.stdout)
    m2 = re.search(r"rendering took\s+([0-9.]+)", r.stdout)
    bvh = re.search(r"BVH building:[^\n]*", r.stdout)
    sec = float(m.group(1)) if m else (float(m2.group(1)) if m2 else None)
    stages = dict(re.findall(r"(stage \d [a-z_]+\(\)): ([0-9.]+) sec", r.stdout))
    return {"scene": name, "spp": spp, "render_s": sec, "render_s_source": "integrator Timer total" if m else "rendering took",
            "rendering_took_s": float(m2.group(1)) if m2 else None, "wall_s": round(wall, 2),
            "Msamples_s": round(npix * spp / sec / 1e6, 3) if sec else None, "stages_s": {k: float(v) for k, v in stages.items()},
            "bvh_line": bvh.group(0) if bvh else None}


def run_ours(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
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

    dev_sets = [(rb, n, hb, any_hit) for _, any_hit, n, rb, hb in dev]

    def step_device():
        # one call for the three sets of a step: b200pt_trace_device_batch puts consecutive sets on the scene's two traversal
        # streams, so the tail of one persistent kernel overlaps the start of the next (same hit records as one call per set)
        scene.trace_device_batch(dev_sets, sync=False)

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
    for _ in range(args.steps):
        step_device()
    ev[1].record(stream)
    barrier()
    scene.counters()     # raises if ANY of the asynchronous launches above reported a traversal-stack overflow (sticky flag)
    ms_total = ev[0].elapsed_time(ev[1])
    # the same K steps once more with one launch per set on ONE stream and events around every kernel: the per-kernel durations
    # of the roofline (in the timed region above the sets of a step overlap, so a kernel has no duration of its own there)
    k = 2
    for _ in range(args.steps):
        for name, any_hit, n, rb, hb in dev:   # per-kernel events on the launching stream
            ev[k].record(stream)
            scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=False)
            ev[k + 1].record(stream)
            k += 2
    barrier()
    scene.counters()
    clocks = sampler.stop()
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
    ours_hits = {name: hb.download(np.float32, n * 4).reshape(n, 4).copy() for name, _, n, _, hb in dev} if rank == 0 else None

    # ---- e2e: the reference-facing call (host buffers in pinned memory, copies inside) -------------
    import ctypes as C
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    host = []
    for name, any_hit, (o, d, tm) in sets:
        n = len(o)
        bufs = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (o, d, tm)]
        outs = [torch.empty(n, dtype=torch.int32).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory(),
                torch.empty(2 * n, dtype=torch.float32).pin_memory()]
        host.append((any_hit, n, bufs, outs))
    batch = (api_ray_set(pkg) * len(host))()
    for i, (any_hit, n, bufs, outs) in enumerate(host):
        b = batch[i]
        b.o_xyz, b.d_xyz, b.tmax = (C.cast(x.data_ptr(), fp) for x in bufs)
        b.n, b.any_hit = n, any_hit
        b.prim_out = C.cast(outs[0].data_ptr(), ip)
        # any-hit callers want the occlusion flag only: t / barycentrics are not read back for that set
        b.t_out = None if any_hit else C.cast(outs[1].data_ptr(), fp)
        b.b1b2_out = None if any_hit else C.cast(outs[2].data_ptr(), fp)

    def step_host():
        # ONE call per step: the three ray sets share the copy / traverse / copy-back pipeline (b200pt_trace_batch)
        rc = lib.b200pt_trace_batch(scene.handle, batch, len(host))
        if rc:
            raise RuntimeError(lib.b200pt_last_error())

    for _ in range(max(1, args.warmup)):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()   # each call returns with its results on the host
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt.item())
    e2e_value = job_rays * args.steps / e2e_s / 1e6
    h2d = sum(n * 28 for _, n, _, _ in host)
    d2h = sum(n * (4 if any_hit else 16) for any_hit, n, _, _ in host)
    e2e_ids_equal = all(np.array_equal(outs[0].numpy(), ours_hits[name][:, 0].view(np.int32))
                        for (name, *_), (_, _, _, outs) in zip(dev, host)) if rank == 0 else None

    # ---- roofline of the dominant kernel --------------------------------------------------------
    pk, pk_src = peaks()
    peak = float(pk.get("hbm_gbs", 6650.0))
    dom = max(per_set_ms, key=per_set_ms.get)
    c = counts[dom]
    alg_bytes = 48 * c["rays"] + 80 * c["nodes"] + 48 * c["tris"]
    achieved = alg_bytes / (per_set_ms[dom] * 1e-3) / 1e9
    # k_trace is ISSUE bound: the tree is L2 resident, DRAM traffic is a few percent of the algorithmic
    # L2-level bytes.  `achieved` / `frac` stay the contract's algorithmic-bytes-over-HBM-peak figure; the counters that
    # name the real bound come from the committed ncu capture (profiles/traffic.json, stamped with its commit).
    roofline = {"bound": "issue", "kernel": "k_trace<closest> on the %s set" % dom, "achieved": round(achieved, 1),
                "peak": peak, "peak_source": pk_src, "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": None,
                "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": round(per_set_ms[dom], 4),
                "kernel_ms_source": "CUDA events around every launch in a second pass of the same %d steps, one launch per set on one stream, "
                                    "right after the timed region (whose steps overlap their three sets on two streams)" % args.steps,
                "nodes_per_ray": round(c["nodes"] / max(1, c["rays"]), 2), "tris_per_ray": round(c["tris"] / max(1, c["rays"]), 2)}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            tj = json.load(open(prof))
            roofline["traffic"] = tj.get("k_trace_dram_bytes_per_launch")
            roofline["ncu"] = {k2: tj[k2] for k2 in ("commit", "source", "issue_slots_busy_pct", "threads_per_inst", "warp_inst_per_ray", "alu_pipe_pct", "fma_pipe_pct", "note") if k2 in tj}
        except Exception:
            pass

    out = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (+u32 ids; f64 edge fallback)", "data": "synthetic",
        "config": c2_config(args, {n: cn for n, _, cn, _, _ in dev}),
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "call": "b200pt_trace_batch (pinned host buffers; outputs the caller passes as NULL are not copied back: the any-hit "
                        "set returns its occlusion flags only)", "ids_equal_device_resident": e2e_ids_equal},
        "gpu_launches": 3 * args.steps,
        "step": "b200pt_trace_device_batch: the three ray sets of a step in one call, consecutive sets on two traversal streams",
        "ms_per_step_one_stream": round(sum(per_set_ms.values()), 4),
        "clocks": clocks,
        "roofline": roofline,
        "notes": {"l2": "inputs larger than L2 (BVH8 nodes + triangles %.0f MB, rays %.0f MB)"
                        % ((build["n_bvh8_nodes"] * 80 + args.triangles * 48) / 1e6, total_rays * 32 / 1e6),
                  "parallelism": "C2: replicated scene, ray sets sharded by sample index, no data-path collective; "
                                 "render leg: sample-partitioned, one ncclReduce of the film"},
        "per_set_ms": {k2: round(v, 4) for k2, v in per_set_ms.items()},
        "per_set_mrays": {n: round(cn / per_set_ms[n] / 1e3, 1) for n, _, cn, _, _ in dev},
        "bvh_build": {k2: (round(v, 3) if isinstance(v, float) else v) for k2, v in build.items()},
        "setup_s": {"generate": round(t_gen, 2), "upload": round(t_upload, 2)},
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"], out["parity"] = cpu_baseline_and_parity(args, P, idx, sets, ours_hits)
    for _, _, _, rb, hb in dev:
        rb.free()
        hb.free()
    scene.close()
    if not args.no_render:
        out["render"] = render_leg_ours(pkg, args, local, rank, world, dist, torch)
        if world == 1:
            out["render_c1"] = render_c1_ours(pkg, local)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def api_ray_set(pkg):
    return pkg.api.RaySet


def cpu_baseline_and_parity(args, P, idx, sets, ours_hits):
    """The oracle's host BVH (oracle/oracle.c: Morton-split binary tree, the reference's traversal loop, slab test and
    watertight triangle test) on all host cores over the SAME three ray sets -- the checker of the hit ids at the full C2
    size, and the reported CPU baseline (the reference itself has no CPU path)."""
    from oracle import pyoracle
    cores = pyoracle.lib().oracle_get_threads()
    cap = args.cpu_rays if args.cpu_rays > 0 else None
    t0 = time.perf_counter()
    bvh = pyoracle.HostBVH(P, idx)
    t_build = time.perf_counter() - t0
    parity = {}
    n_rays = 0
    dt = 0.0
    for name, any_hit, (o, d, tm) in sets:
        n = len(o) if cap is None else min(len(o), cap)
        pick = np.arange(len(o)) if n == len(o) else np.linspace(0, len(o) - 1, n).astype(np.int64)
        oo, dd, tt = o[pick], d[pick], tm[pick]
        t0 = time.perf_counter()
        if any_hit:
            occ = bvh.any_hit(oo, dd, tt)
        else:
            prim, t, _ = bvh.closest_hit(oo, dd, tt)
        dt += time.perf_counter() - t0
        n_rays += n
        h = ours_hits[name][pick]
        if any_hit:
            parity[name] = {"rays": int(n), "mismatches": int((occ != (h[:, 0].view(np.int32) != 0)).sum())}
        else:
            mism, near = pyoracle.compare_closest(P, idx, oo, dd, tt, h[:, 0].view(np.int32), h[:, 1], prim, t)
            same = h[:, 0].view(np.int32) == prim
            t_bits = int((h[same & (prim >= 0), 1].view(np.uint32) != t[same & (prim >= 0)].view(np.uint32)).sum())
            parity[name] = {"rays": int(n), "mismatches": mism, "near_ties": near, "t_bit_differences_on_equal_ids": t_bits}
    bvh.close()
    parity["checker"] = "oracle host BVH over the full ray sets (ids bit-exact; near tie = the other primitive is accepted within 4 ulp of t)"
    base = {"value": round(n_rays / dt / 1e6, 4), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d rays (the three ray sets%s), host BVH traversal with the reference's loop and predicates on %d threads, "
                      "%.1f s; its BVH build %.1f s on one thread (not counted)" % (n_rays, "" if cap is None else ", subsampled", cores, dt, t_build)}
    return base, parity


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
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (+u32 ids; f64 edge fallback)", "data": "synthetic"}
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
        ref_render = ref_c1 = None
        if not args.no_render:
            # the render leg on a BOUNDED sample: the same 1920x1080 scene at --ref-render-spp samples per pixel (Msamples/s of a
            # path tracer does not depend on spp; the full 256 spp would take the reference about a minute)
            ref_render = render_reference(td, cornell_text(args.render_width, args.render_height), "cornell_%dx%d" % (args.render_width, args.render_height),
                                          args.ref_render_spp, args.render_width * args.render_height)
            ref_render["config"] = {"workload": RENDER_WORKLOAD % (args.render_width, args.render_height, args.render_spp),
                                    "spp_total": args.render_spp, "maxdepth": 5, "sampler": "stratified"}
            ref_render["sample"] = "%d of %d spp, one GPU (the reference is single-device)" % (args.ref_render_spp, args.render_spp)
            ref_c1 = render_reference(td, open(CORNELL).read(), "cornell_c1", 16, 1080 * 1080)
    if r.returncode != 0:
        base["unavailable"] = "ref_trace failed: " + (r.stdout + r.stderr)[-300:].replace("\n", " ")
        print(json.dumps(base))
        return
    info = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    build_line = [l for l in r.stdout.splitlines() if l.startswith("BVH building")]
    ms_per_step = sum(s["ms"] for s in info["sets"])
    rays = sum(s["rays"] for s in info["sets"])
    value = rays / (ms_per_step * 1e-3) / 1e6
    names = ("primary", "bounce", "shadow")
    sample = ("full workload; the reference has no CPU path, so this is its own CUDA build (sm_100) on GPU 0 with its BVH build on "
              "the %d host cores" % cores)
    base.update({
        "value": round(value, 2), "ms_per_step": round(ms_per_step, 3),
        "config": c2_config(args, {n: s["rays"] for n, s in zip(names, info["sets"])}),
        "e2e": {"value": round(value, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        # the contract's key for "what this arm timed"; the number is the reference's GPU build (also under reference_gpu)
        "cpu_baseline": {"value": round(value, 2), "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "reference_gpu": {"value": round(value, 2), "unit": UNIT, "what": "the reference's own CUDA build of HLBVH::intersect / fast_intersect on one B200"},
        "per_set_ms": {n: round(s["ms"], 3) for n, s in zip(names, info["sets"])},
        "reference_bvh_build_s": info["bvh_build_s"], "reference_load_s": info["load_s"],
        "reference_build_line": build_line[-1] if build_line else None, "wall_s": round(wall, 1),
        "render": ref_render, "render_c1": ref_c1,
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
