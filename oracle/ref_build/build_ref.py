#!/usr/bin/env python3
"""Build the UNMODIFIED reference (w3ntao/pbrt-minus) for sm_100 into oracle/_ref/.

TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is on the product path.

The reference is a CUDA-only program (no CPU render path), so "the reference itself" can only be
run on the GPU box.  This recipe compiles its sources where they lie under /root/reference (never
copied into the repo) with two headless shims first on the include path:

  shims/pbrt/gui/gl_helper.h   - no GLFW/OpenGL (no graphics stack on either machine)
  shims/ext/tinyexr/tinyexr.h  - the tinyexr submodule is missing from the archive

Outputs (all git-ignored, all travel to the GPU box with the gpurun snapshot):

  oracle/_ref/pbrt-minus-ref   the reference CLI; its Film::write_to_png additionally dumps the
                               linear float RGB film next to the PNG (<name>.f32, W*H*3 floats);
                               the dump is injected into a generated copy of base/film.cu at build
                               time (oracle/_ref/gen/film.cu), nothing else is touched.
  oracle/_ref/ref_trace        our ID-recording harness (ref_trace.cu) linked against the
                               reference's own objects: builds the reference HLBVH over a raw
                               triangle soup and traces ray files with the reference's traversal
                               order and the reference's Primitive::intersect, recording the
                               winning primitive index (the reference API returns none).
  oracle/_ref/ref_probe        host-only probe (ref_probe.cu): calls the reference's
                               __host__ __device__ functions on the CPU to emit golden vectors.

Usage: python oracle/ref_build/build_ref.py [-j N] [--only probe|trace|cli]
"""
import argparse
import concurrent.futures as cf
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("PBRT_MINUS_REF", "/root/reference")
OUT = os.path.join(REPO, "oracle", "_ref")
OBJ = os.path.join(OUT, "obj")
GEN = os.path.join(OUT, "gen")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
COMMON = ["-std=c++20", "--expt-relaxed-constexpr", "-arch=sm_100", "-O3", "-w",
          "-I" + os.path.join(HERE, "shims"), "-I" + os.path.join(REF, "src")]


def run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + "\n")
        raise SystemExit(1)
    return r.stdout


def newer(src, dst):
    return (not os.path.exists(dst)) or os.path.getmtime(src) > os.path.getmtime(dst)


def reference_sources():
    cm = open(os.path.join(REF, "CMakeLists.txt")).read()
    return re.findall(r"^\s*(src/pbrt/\S+\.cu)\s*$", cm, flags=re.M)


def gen_film_dump():
    """Generated copy of base/film.cu whose write_to_png also dumps the float film."""
    src = open(os.path.join(REF, "src/pbrt/base/film.cu")).read()
    marker = "    // Encode the image"
    assert src.count(marker) == 1, "reference film.cu changed; update the dump injection"
    dump = r'''
    {   // [oracle/ref_build] float dump of the linear output RGB (Film::get_pixel_rgb)
        std::vector<float> raw(size_t(width) * height * 3);
        for (int y = 0; y < height; y++) {
            for (int x = 0; x < width; x++) {
                const auto c = get_pixel_rgb(Point2i(x, y), splat_scale);
                raw[3 * (size_t(y) * width + x) + 0] = c.r;
                raw[3 * (size_t(y) * width + x) + 1] = c.g;
                raw[3 * (size_t(y) * width + x) + 2] = c.b;
            }
        }
        if (FILE *fp = fopen((filename + ".f32").c_str(), "wb")) {
            fwrite(raw.data(), sizeof(float), raw.size(), fp);
            fclose(fp);
        }
    }
'''
    os.makedirs(GEN, exist_ok=True)
    path = os.path.join(GEN, "film.cu")
    new = src.replace(marker, dump + marker)
    if not os.path.exists(path) or open(path).read() != new:
        open(path, "w").write(new)
    return path


def gen_light_debug():
    """Generated copy of light_samplers/power_light_sampler.cu that prints the light weights and the alias
    table (debug binary pbrt-minus-ref-debug only; the timed / golden binary stays unmodified)."""
    src = open(os.path.join(REF, "src/pbrt/light_samplers/power_light_sampler.cu")).read()
    marker = "    lights_power_distribution = allocator.create<Distribution1D>(lights_pmf, allocator);\n"
    assert src.count(marker) == 1, "reference power_light_sampler.cu changed"
    dump = r'''
    {   // [oracle/ref_build] debug dump
        cudaDeviceSynchronize();
        const AliasTable *t = lights_power_distribution->alias_table;
        for (int i = 0; i < light_num && i < 64; ++i) {
            printf("[lightdbg] %d weight %.9g pdf %.9g bin p %.9g first %d second %d\\n", i, (double)lights_pmf[i],
                   (double)t->pdfs[i], (double)t->bins[i].p, t->bins[i].first_idx, t->bins[i].second_idx);
        }
    }
'''
    os.makedirs(GEN, exist_ok=True)
    path = os.path.join(GEN, "power_light_sampler_debug.cu")
    new = src.replace(marker, marker + dump).replace("#include <pbrt/light_samplers/power_light_sampler.h>",
                                                     "#include <pbrt/light_samplers/power_light_sampler.h>\n#include <pbrt/distribution/alias_table.h>\n#include <cstdio>")
    new = "#define private public  // [oracle/ref_build] debug copy only: Distribution1D::alias_table is private\n" + new
    if not os.path.exists(path) or open(path).read() != new:
        open(path, "w").write(new)
    return path


def gen_bridge_scene_builder():
    """Generated copy of scene/scene_builder.cu whose wavefront branch of SceneBuilder::render calls libb200pt.so through
    pbrt-minus_b200/bridge/b200pt_bridge.h instead of WavefrontPathIntegrator::render + Film::write_to_png
    (scene_builder.cu:1036-1043).  The reference has no accessor API for the objects its parser builds, so this ONE translation
    unit is compiled with the access specifiers opened (every non-pbrt header is included before the macros)."""
    src = open(os.path.join(REF, "src/pbrt/scene/scene_builder.cu")).read()
    old = "        wavefront_path_integrator->render(film, preview);\n\n        film->write_to_png(output_filename);\n"
    assert src.count(old) == 1, "reference scene_builder.cu changed; update the bridge injection"
    new = "        b200pt_bridge_render(*this); // [b200pt bridge] the wavefront path integrator, through libb200pt.so\n"
    std = ["algorithm", "atomic", "chrono", "cmath", "condition_variable", "cstring", "filesystem", "fstream", "functional", "iomanip",
           "iostream", "iterator", "limits", "map", "memory", "mutex", "numbers", "numeric", "optional", "queue", "ranges", "set",
           "sstream", "stack", "string", "thread", "vector", "stdint.h", "cuda_fp16.h", "cuda/std/variant", "cuda/std/tuple",
           "cuda/std/optional", "ext/lodepng/lodepng.h", "pbrt/gpu/macro.h", "pbrt/gpu/gpu_memory_allocator.h", "pbrt/util/bit_arithmetic.h"]
    head = "// GENERATED by oracle/ref_build/build_ref.py from the reference's scene/scene_builder.cu -- do not edit\n"
    head += "".join("#include <%s>\n" % h for h in std)
    head += ("// the reference keeps the objects its parser builds behind private members and offers no accessors; this translation\n"
             "// unit has no kernels, so it is compiled by the host compiler alone with -fno-access-control (see build_ref.py)\n"
             "#include <cuda_runtime.h>\n#include <math.h>\nusing std::isnan;\nusing std::isinf;\nusing std::isfinite;\nusing std::min;\nusing std::max;\n"
             "template <typename T> T atomicAdd(T *address, T value) { T old = *address; *address += value; return old; } // device-only in the reference\n"
             "class SceneBuilder;\nvoid b200pt_bridge_render(const SceneBuilder &sb);\n")
    tail = ('\n#include "b200pt_bridge.h"\n'
            "void b200pt_bridge_render(const SceneBuilder &sb) { B200ptBridge::render(sb); }\n")
    os.makedirs(GEN, exist_ok=True)
    path = os.path.join(GEN, "scene_builder_b200.cu")
    text = head + src.replace(old, new) + tail
    if not os.path.exists(path) or open(path).read() != text:
        open(path, "w").write(text)
    return path


def build_probe():
    """Host-only probe of the reference's integer paths -> tests/golden/ref_probe.json."""
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "ref_probe")
    run([NVCC, "-std=c++20", "--expt-relaxed-constexpr", "-O2", "-w", "-Wno-deprecated-gpu-targets",
         "-I" + os.path.join(HERE, "shims"), "-I" + os.path.join(REF, "src"),
         os.path.join(HERE, "ref_probe.cu"), "-o", exe])
    out = run([exe])
    gold = os.path.join(REPO, "tests", "golden", "ref_probe.json")
    open(gold, "w").write(out)
    print("wrote", gold)
    return 0


def compile_one(job):
    src, obj, extra = job
    if newer(src, obj):
        run([NVCC] + COMMON + extra + ["-dc", src, "-o", obj])
    return obj


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-j", type=int, default=os.cpu_count() or 4)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    if not os.path.isdir(REF):
        print("reference tree not present at %s; keeping prebuilt oracle/_ref as is" % REF)
        return 0
    os.makedirs(OBJ, exist_ok=True)

    jobs = []
    for rel in reference_sources():
        src = os.path.join(REF, rel)
        if rel == "src/pbrt/base/film.cu":
            src = gen_film_dump()
        obj = os.path.join(OBJ, rel.replace("/", "_")[:-3] + ".o")
        jobs.append((src, obj, []))
    ext = [("src/ext/lodepng/lodepng.cpp", "ext_lodepng.o"), ("src/ext/rply/rply.cpp", "ext_rply.o")]
    for rel, o in ext:
        jobs.append((os.path.join(REF, rel), os.path.join(OBJ, o), ["-x", "cu"] if False else []))
    if args.only == "probe":
        return build_probe()
    for name in ("ref_trace",):
        p = os.path.join(HERE, name + ".cu")
        if os.path.exists(p):
            jobs.append((p, os.path.join(OBJ, "harness_" + name + ".o"), []))

    with cf.ThreadPoolExecutor(args.j) as ex:
        objs = list(ex.map(compile_one, jobs))

    main_o = [o for o in objs if o.endswith("src_pbrt_main.o")]
    harness = [o for o in objs if os.path.basename(o).startswith("harness_")]
    lib = [o for o in objs if o not in main_o and o not in harness]
    link = [NVCC, "-arch=sm_100"]
    tail = ["-lz", "-lpthread"]
    if args.only == "":
        build_probe()
    if args.only in ("", "cli"):
        run(link + main_o + lib + tail + ["-o", os.path.join(OUT, "pbrt-minus-ref")])
    if args.only in ("", "cli", "debug"):
        dbg_o = compile_one((gen_light_debug(), os.path.join(OBJ, "gen_power_light_sampler_debug.o"), []))
        lib_dbg = [o for o in lib if not o.endswith("src_pbrt_light_samplers_power_light_sampler.o")] + [dbg_o]
        run(link + main_o + lib_dbg + tail + ["-o", os.path.join(OUT, "pbrt-minus-ref-debug")])
    if args.only in ("", "cli", "bridge"):
        # pbrt-minus-b200: the reference's own front-end (lexer, parser, constructors) driving libb200pt.so through the bridge
        lib_dir = os.path.join(REPO, "pbrt-minus_b200")
        if os.path.exists(os.path.join(lib_dir, "libb200pt.so")):
            extra = ["-I" + os.path.join(REPO, "include"), "-I" + os.path.join(lib_dir, "bridge")]
            bridge_src = gen_bridge_scene_builder()
            bridge_o = os.path.join(OBJ, "gen_scene_builder_b200.o")
            hdr = os.path.join(lib_dir, "bridge", "b200pt_bridge.h")
            if newer(bridge_src, bridge_o) or newer(hdr, bridge_o) or newer(os.path.join(REPO, "include", "b200pt.h"), bridge_o):
                # host-only translation unit (no kernels, no __global__): g++ with access checks off, CUDA headers for the
                # __host__ __device__ annotations and cuda::std::variant; same ABI as the nvcc-compiled objects it links with
                run([os.environ.get("CXX", "g++"), "-std=c++20", "-O2", "-w", "-fno-access-control", "-fpermissive", "-x", "c++", "-fPIC",
                     "-I/usr/local/cuda/include", "-I" + os.path.join(HERE, "shims"), "-I" + os.path.join(REF, "src")] + extra +
                    ["-c", bridge_src, "-o", bridge_o])
            lib_b = [o for o in lib if not o.endswith("src_pbrt_scene_scene_builder.o")] + [bridge_o]
            run(link + main_o + lib_b + tail + ["-L" + lib_dir, "-lb200pt", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../../pbrt-minus_b200",
                                                 "-o", os.path.join(OUT, "pbrt-minus-b200")])
        else:
            print("pbrt-minus_b200/libb200pt.so not built yet: skipping oracle/_ref/pbrt-minus-b200")
    for h in harness:
        name = os.path.basename(h)[len("harness_"):-2]
        if args.only in ("", name.replace("ref_", "")):
            run(link + [h] + lib + tail + ["-o", os.path.join(OUT, name)])
    print("oracle/_ref built:", sorted(f for f in os.listdir(OUT) if not os.path.isdir(os.path.join(OUT, f))))
    return 0


if __name__ == "__main__":
    sys.exit(main())
