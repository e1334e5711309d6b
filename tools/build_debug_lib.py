#!/usr/bin/env python3
"""A debug variant of the library: csrc/render.cu compiled with -DB200PT_DEBUG_PIXELS(p)=<predicate on the pixel index>, linked
with the normal objects into pbrt-minus_b200/libb200pt_dbg.so (git-ignored; select it with B200PT_LIB).  The shade and emit
kernels then print the NEE record / emission of the chosen pixels (device printf), one line per sample.

  python tools/build_debug_lib.py 4081 5744        # pixel indices y * width + x
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "pbrt-minus_b200"))
import build as B  # noqa: E402

B.build()
pred = " || ".join("(p) == %d" % int(v) for v in sys.argv[1:]) or "false"
obj = os.path.join(B.OBJ, "render_dbg.o")
r = subprocess.run([B.NVCC] + [f for f in B.FLAGS if f not in ("-Xptxas", "-v")] + ["-DB200PT_DEBUG_PIXELS(p)=(%s)" % pred, "-c", os.path.join(B.CSRC, "render.cu"), "-o", obj],
                   stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
if r.returncode:
    sys.exit(r.stdout[-3000:])
objs = [os.path.join(B.OBJ, s[:-3] + ".o") for s in B._sources() if s != "render.cu"] + [obj]
out = os.path.join(ROOT, "pbrt-minus_b200", "libb200pt_dbg.so")
r = subprocess.run([B.NVCC, "-shared", "-o", out] + objs + ["-lcudart_static", "-ldl", "-lpthread", "-lrt"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
if r.returncode:
    sys.exit(r.stdout[-3000:])
print(out)
