"""Which object of tests/scenes/metals.pbrt separates our film from the reference's: renders variants of the scene (one metal
object at a time) with the reference CLI and with libb200pt, prints relMSE / differing pixels.  Needs oracle/_ref (GPU box)."""
import importlib, os, re, subprocess, sys, tempfile, pathlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T

text = open(os.path.join(ROOT, "tests", "scenes", "metals.pbrt")).read()
head, *blocks = re.split(r"(?m)^(?=# (?:tilted plate|sphere: copper|sphere: eta|small plate))", text)
names = ["plate_mix", "sphere_dirmix", "sphere_scale_k", "plate_rgb_eta"]
variants = {"none": head, "all": text}
if "--each" in sys.argv:
    for n, b in zip(names, blocks):
        variants[n] = head + b
variants["no_side_light"] = re.sub(r'AttributeBegin\n  AreaLightSource "diffuse" "rgb L" \[0.7 0.8 1\].*?AttributeEnd\n', "", head, flags=re.S)
# the same objects with plain spectra in place of the textures
variants["plate_plain_au"] = head + blocks[0].replace('"texture eta" ["inlay-eta"] "texture k" ["inlay-k"]', '"spectrum eta" ["metal-Au-eta"] "spectrum k" ["metal-Au-k"]')
variants["plate_const_tex"] = head + blocks[0].replace('"texture eta" ["inlay-eta"] "texture k" ["inlay-k"]', '"texture eta" ["au-eta"] "texture k" ["au-k"]')
exe = os.path.join(ROOT, "oracle", "_ref", "pbrt-minus-ref")
spp = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 64
for name, t in variants.items():
    tmp = pathlib.Path(tempfile.mkdtemp())
    import shutil
    shutil.copytree(os.path.join(ROOT, "tests", "scenes", "textures"), tmp / "textures")
    sp = tmp / "s.pbrt"; sp.write_text(t)
    png = str(tmp / "ref.png")
    r = subprocess.run([exe, str(sp), "--spp", str(spp), "--outfile", png], capture_output=True, text=True, cwd=str(tmp), timeout=900)
    if r.returncode != 0:
        print(name, "reference failed", (r.stdout + r.stderr)[-300:]); continue
    ref = np.fromfile(png + ".f32", dtype=np.float32).reshape(64, 96, 3)
    ours, _ = T.render(b200pt, t, spp, tmp)
    d = np.abs(ours - ref).max(axis=2) > 1e-4 * (1 + ref.max(axis=2))
    ys, xs = np.nonzero(d)
    box = (int(xs.min()), int(xs.max()), int(ys.min()), int(ys.max())) if d.any() else None
    print("%-16s relMSE %.3e  differing %4d / %d  box(x0,x1,y0,y1) %s  means %.5f %.5f" % (name, T.rel_mse(ours, ref), int(d.sum()), d.size, box, ours.mean(), ref.mean()), flush=True)
