#!/usr/bin/env python3
"""Attribute the per-instruction samples of an `ncu --page source --print-source sass --csv` export to source lines.

ncu's CUDA-source view comes out empty for kernels built from many inlined headers, so the mapping is made here: the kernel's
SASS is disassembled with line info (`nvdisasm -g -c` of the cubin extracted from libb200pt.so) and matched to the ncu rows
instruction by instruction.

  python tools/ncu_source_hotspots.py SASS_CSV KERNEL_SUBSTRING MANGLED_NAME [--top N]
"""
import argparse
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def disassemble(mangled, src="render", lib=None):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib) if lib else os.path.join(ROOT, "pbrt-minus_b200", "libb200pt.so")], cwd=tmp, check=True, capture_output=True)
    cubin = os.path.join(tmp, src + ".sm_100a.cubin")
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
    start = None
    for i, l in enumerate(dis):
        if l.startswith(".text." + mangled + ":"):
            start = i
            break
    if start is None:
        sys.exit("function not found: " + mangled)
    out = []       # (opcode text, file, line, inlined-at chain)
    cur = ("?", 0)
    for l in dis[start + 1:]:
        if l.startswith("\t.section") or l.startswith(".text."):
            break
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)), m.group(3).strip())
            continue
        m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(.*?);', l)
        if m:
            out.append((m.group(1).strip(),) + tuple(cur))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("kernel")
    ap.add_argument("mangled")
    ap.add_argument("--src", default="render")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--lib", default=None, help="the build the capture was made with (default: the in-tree library)")
    ap.add_argument("--by", default="samples", choices=["samples", "inst"], help="rank lines by stall samples or by warp instructions executed")
    ap.add_argument("--window", type=int, default=0, help="also print the N hottest single instructions")
    args = ap.parse_args()
    csv.field_size_limit(10 ** 9)
    rows, take, hdr = [], False, None
    for r in csv.reader(open(args.csv)):
        if r and r[0] == "Kernel Name":
            if take:
                break
            take = args.kernel in r[1]
            continue
        if take:
            if r and r[0] == "Address":
                hdr = r
            elif hdr:
                rows.append(r)
    if not rows:
        sys.exit("kernel not in csv")
    ci = {n: hdr.index(n) for n in ("Source", "Warp Stall Sampling (All Samples)", "Instructions Executed", "Thread Instructions Executed")}
    dis = disassemble(args.mangled, args.src, args.lib)
    print("ncu rows %d, disassembled instructions %d" % (len(rows), len(dis)))
    n = min(len(rows), len(dis))
    by_line = collections.Counter()
    inst_by_line = collections.Counter()
    by_file = collections.Counter()
    inst_by_file = collections.Counter()
    total = 0
    total_inst = 0
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
    stall_tot = collections.Counter()
    stall_by_line = collections.defaultdict(collections.Counter)
    hot = []
    for k in range(n):
        r = rows[k]
        s = int(r[ci["Warp Stall Sampling (All Samples)"]] or 0)
        ie = int(r[ci["Instructions Executed"]] or 0)
        key = (dis[k][1], dis[k][2])
        by_line[key] += s
        inst_by_line[key] += ie
        by_file[dis[k][1]] += s
        inst_by_file[dis[k][1]] += ie
        total += s
        total_inst += ie
        for c in stall_cols:
            v = int(r[c] or 0)
            if v:
                stall_tot[hdr[c]] += v
                stall_by_line[key][hdr[c]] += v
        hot.append((s, k, r[ci["Source"]].strip(), key))
    print("samples %d, warp instructions executed %d" % (total, total_inst))
    print("stall reasons:", ", ".join("%s %.1f%%" % (k, 100.0 * v / max(1, sum(stall_tot.values()))) for k, v in stall_tot.most_common(8)))
    print("by file:", ", ".join("%s %.1f%%" % (k, 100.0 * v / max(1, total)) for k, v in by_file.most_common(12)))
    print("%-22s %8s %7s %9s  top stalls" % ("file:line", "samples", "%", "warp-inst"))
    ranked = by_line.most_common(args.top) if args.by == "samples" else [(k, by_line[k]) for k, _ in inst_by_line.most_common(args.top)]
    print("by file (warp instructions):", ", ".join("%s %.1f%%" % (k, 100.0 * v / max(1, total_inst)) for k, v in inst_by_file.most_common(12)))
    for key, s in ranked:
        st = ", ".join("%s %d" % (a.replace("stall_", ""), b) for a, b in stall_by_line[key].most_common(3))
        print("%-22s %8d %6.1f%% %9d  %s" % ("%s:%d" % key, s, 100.0 * s / max(1, total), inst_by_line[key], st))
    if args.window:
        print("hottest instructions:")
        for s, k, src, key in sorted(hot, reverse=True)[:args.window]:
            print("  %6d  #%d  %-60s %s:%d" % (s, k, src[:60], key[0], key[1]))


if __name__ == "__main__":
    main()
