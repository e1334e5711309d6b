#!/usr/bin/env python3
"""Parity bisection helper: render variants of a scene (one attribute block removed at a time, filter
swapped) with both renderers and print relMSE per variant.  GPU box only."""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    scene = sys.argv[1]
    spp = sys.argv[2] if len(sys.argv) > 2 else "64"
    text = open(scene).read()
    blocks = list(re.finditer(r"AttributeBegin.*?AttributeEnd\n", text, flags=re.S))
    variants = {"full": text, "mitchell": text.replace('PixelFilter "gaussian"', 'PixelFilter "mitchell"'),
                "depth5": re.sub(r'"integer maxdepth" \[\d+\]', '"integer maxdepth" [5]', text)}
    for i, b in enumerate(blocks):
        variants["minus_block%d" % i] = text[:b.start()] + text[b.end():]
    only = text
    for b in reversed(blocks[1:]):
        only = only[:b.start()] + only[b.end():]
    variants["only_block0"] = only
    if len(sys.argv) > 3 and sys.argv[3] == "lights":
        variants = {
            "side_rgb": text.replace('"spectrum L" [400 2 500 6 600 9 700 4]', '"rgb L" [0.5 0.7 0.9]'),
            "side_facing": text.replace('"point3 P" [-3.5 0.2 1  -3.5 2.2 1  -3.5 0.2 -1]', '"point3 P" [-3.5 0.2 1  -3.5 0.2 -1  -3.5 2.2 1]'),
            "panel_onesided": text.replace('"bool twosided" [true]', '"bool twosided" [false]').replace('[0 2 1 0 3 2]', '[0 1 2 0 2 3]'),
            "panel_flat": text.replace("  Rotate 20 0 0 1\n  Scale 1.5 1 1\n", ""),
            "panel_scale13": text.replace('"float scale" [12]', '"float scale" [1.3]'),
        }
    out = os.path.join(ROOT, "gpurun_out", "bisect")
    os.makedirs(out, exist_ok=True)
    for name, t in variants.items():
        p = os.path.join(out, name + ".pbrt")
        open(p, "w").write(t)
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "render_compare.py"), "--scene", p, "--spp", spp, "--tag", name,
                            "--out", out], capture_output=True, text=True)
        try:
            j = json.loads(r.stdout.strip().splitlines()[-1])
            print("%-16s relMSE %.3e  frac>1%% %.3f  mean_rel %.2e" % (name, j["compare"]["relMSE"], j["compare"]["frac_pixels_gt_1pct"],
                                                                      j["compare"]["mean_rel_diff_of_means"]))
        except Exception as e:
            print(name, "failed", e, r.stdout[-300:], r.stderr[-300:])


if __name__ == "__main__":
    main()
