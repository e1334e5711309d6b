"""In-tree build of libb200pt.so (hand-written CUDA for sm_100a) with nvcc.

`python pbrt-minus_b200/build.py` or `__graft_entry__.build()`.  nvcc cross-compiles without a GPU.
The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libb200pt.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "-Xptxas", "-v"]


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_mtime():
    inc = os.path.join(os.path.dirname(HERE), "include", "b200pt.h")
    paths = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))] + [inc]
    return max(os.path.getmtime(p) for p in paths)


def _compile(src):
    obj = os.path.join(OBJ, src[:-3] + ".o")
    spath = os.path.join(CSRC, src)
    if os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(spath), _headers_mtime()):
        return obj, ""
    r = subprocess.run([NVCC] + FLAGS + ["-c", spath, "-o", obj], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed on %s:\n%s" % (src, r.stdout))
    with open(obj[:-2] + ".ptxas.log", "w") as f:   # registers / spills per kernel (-Xptxas -v), one log per source
        f.write(r.stdout)
    return obj, r.stdout


def build(verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    with cf.ThreadPoolExecutor(max(1, os.cpu_count() or 4)) as ex:
        results = list(ex.map(_compile, _sources()))
    objs = [o for o, _ in results]
    log = "".join(l for _, l in results)
    if (not os.path.exists(LIB)) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        r = subprocess.run([NVCC, "-shared", "-o", LIB] + objs + ["-lcudart_static", "-ldl", "-lpthread", "-lrt"],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout)
    if verbose and log:
        sys.stdout.write(log)
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
