#!/usr/bin/env python3
"""Extract the physical constant tables the spectral pipeline needs (CIE 1931 observer, CIE D65,
metal / glass refractive indices, the 95-sample tables of the RGB->spectrum fit) from the reference
headers where they lie under /root/reference, into pbrt-minus_b200/data/tables.npz.

These are published measurement data (CIE 015:2004; refractiveindex.info), not code; the numbers
must be the reference's so sensor responses and metal colours agree.  Re-run only when the
reference changes: python oracle/ref_build/extract_tables.py
"""
import os
import re
import sys

import numpy as np

REF = os.environ.get("PBRT_MINUS_REF", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(os.path.dirname(HERE)), "pbrt-minus_b200", "data", "tables.npz")


def arrays(path):
    src = open(path).read()
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"N\(([^()]*)\)", r"\1", src)   # the fit's D65 table is written N(x) = x / 10566.864...
    out = {}
    for m in re.finditer(r"(\w+)\s*(?:\[[^\]]*\])?\s*=\s*\{([^{}]*)\}", src):
        body = m.group(2)
        try:
            vals = [float(x.rstrip("fF")) for x in re.split(r"[\s,]+", body.strip()) if x]
        except ValueError:
            continue
        if len(vals) >= 8:
            out[m.group(1)] = np.array(vals, dtype=np.float64)
    return out


def main():
    d = os.path.join(REF, "src/pbrt/spectrum_util")
    cie = arrays(os.path.join(d, "spectrum_constants_cie.h"))
    metal = arrays(os.path.join(d, "spectrum_constants_metal.h"))
    glass = arrays(os.path.join(d, "spectrum_constants_glass.h"))
    fit = arrays(os.path.join(d, "rgb_to_spectrum_data.h"))
    enc = arrays(os.path.join(d, "color_encoding.h"))   # the 256-entry sRGB-to-linear table 8-bit images are decoded through
    out = {
        "cie_x": cie["CIE_X_VALUE_CPU"], "cie_y": cie["CIE_Y_VALUE_CPU"], "cie_z": cie["CIE_Z_VALUE_CPU"],
        "cie_lambda": cie["CIE_LAMBDA_CPU"], "illum_d65_interleaved": cie["CIE_Illum_D6500"],
        "fit_cie_x": fit["cie_x"], "fit_cie_y": fit["cie_y"], "fit_cie_z": fit["cie_z"], "fit_cie_d65": fit["cie_d65"] / 10566.864005283874576,
    }
    assert len(enc["SRGBToLinearLUT"]) == 256
    out["srgb_to_linear_lut"] = enc["SRGBToLinearLUT"]
    for k, v in list(metal.items()) + list(glass.items()):
        out["named_" + k] = v
    for k, v in out.items():
        print("%-28s %d" % (k, len(v)))
    assert len(out["cie_x"]) == 471 and len(out["fit_cie_x"]) == 95
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    sys.exit(main())
