#!/usr/bin/env python3
"""Where does the host-buffer trace call (b200pt_trace, the e2e leg of bench.py) spend its time?  Per ray set of C2: raw
pinned H2D / D2H copy times of the same bytes, the device-resident trace time, and the wall time of the full call for the
current B200PT_TRACE_CHUNK.   python tools/e2e_probe.py [--triangles N]"""
import argparse, ctypes as C, importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--triangles", type=int, default=10_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    pkg = importlib.import_module("pbrt-minus_b200")
    lib = pkg.load_library()
    P, idx = pkg.workloads.triangle_soup(args.triangles)
    scene = pkg.HLBVH(P, idx, device=0, build=True)
    sets = bench.make_workload(pkg, args, scene, 0)
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    out = {"chunk": os.environ.get("B200PT_TRACE_CHUNK", "default")}
    for name, any_hit, (o, d, tm) in sets:
        n = len(o)
        bufs = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (o, d, tm)]
        outs = [torch.empty(n, dtype=torch.int32).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory(),
                torch.empty(2 * n, dtype=torch.float32).pin_memory()]
        dev_in = [torch.empty_like(b, device="cuda") for b in bufs]
        dev_out = [torch.empty_like(b, device="cuda") for b in outs]

        def timeit(fn):
            fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.reps):
                fn()
            torch.cuda.synchronize()
            return (time.perf_counter() - t0) / args.reps * 1e3

        def h2d():
            for a, b in zip(dev_in, bufs):
                a.copy_(b, non_blocking=True)

        def d2h():
            for a, b in zip(outs, dev_out):
                a.copy_(b, non_blocking=True)

        def call():
            rc = lib.b200pt_trace(scene.handle, C.cast(bufs[0].data_ptr(), fp), C.cast(bufs[1].data_ptr(), fp),
                                  C.cast(bufs[2].data_ptr(), fp), n, any_hit, C.cast(outs[0].data_ptr(), ip),
                                  C.cast(outs[1].data_ptr(), fp), C.cast(outs[2].data_ptr(), fp))
            assert rc == 0

        rays = scene.pack_rays(o, d, tm)
        rb, hb = scene.alloc(rays.nbytes), scene.alloc(n * 16)
        rb.upload(rays)

        def dev():
            scene.trace_device(rb, n, hb, any_hit=bool(any_hit), sync=True)

        out[name] = {"n": n, "h2d_ms": round(timeit(h2d), 3), "d2h_ms": round(timeit(d2h), 3), "device_trace_ms": round(timeit(dev), 3),
                     "call_ms": round(timeit(call), 3)}
    print(out)


if __name__ == "__main__":
    main()
