#!/usr/bin/env python3
"""Generate golden films with the REFERENCE's own CUDA build (oracle/_ref/pbrt-minus-ref, float film
dump) on the GPU box: python tools/make_golden.py  -> gpurun_out/golden/*.f32 + manifest.json.
The outputs are then committed under tests/golden/ (this script is how they were made)."""
import json
import os
import re
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from render_compare import with_resolution  # noqa: E402

CASES = [
    # name, scene, (W, H) or None, spp, integrator, sampler override
    ("cornell_64_16spp", "cornell-box-specular.pbrt", (64, 64), 16, None, None),
    ("cornell_48_4096spp", "cornell-box-specular.pbrt", (48, 48), 4096, None, None),
    ("cornell_48_4096spp_megakernel", "cornell-box-specular.pbrt", (48, 48), 4096, "megakernelpath", None),
    ("cornell_64_9spp_independent", "cornell-box-specular.pbrt", (64, 64), 9, None, "independent"),
    ("materials_96x64_64spp", "materials.pbrt", None, 64, None, None),
    ("materials_96x64_1024spp", "materials.pbrt", None, 1024, None, None),
    # LightSource infinite / distant / spot + diffusetransmission
    ("lights_96x64_64spp", "lights.pbrt", None, 64, None, None),
    ("lights_96x64_1024spp", "lights.pbrt", None, 1024, None, None),
    # coated diffuse / coated conductor (LayeredBxDF)
    ("coated_96x64_64spp", "coated.pbrt", None, 64, None, None),
    ("coated_96x64_1024spp", "coated.pbrt", None, 1024, None, None),
    # ObjectBegin / ObjectInstance (the reference: per-instance BVH + TransformedPrimitive; here: flattened)
    ("instances_96x64_64spp", "instances.pbrt", None, 64, None, None),
    # Material "mix" (hash-resolved at hit time: statistical gate)
    ("mix_96x64_1024spp", "mix.pbrt", None, 1024, None, None),
    # per-vertex shading normals + uvs (incl. a mirrored instance and ReverseOrientation); one coated sheet -> statistical in part
    ("normals_96x64_64spp", "normals.pbrt", None, 64, None, None),
    ("normals_96x64_1024spp", "normals.pbrt", None, 1024, None, None),
    # homogeneous media, interface surfaces, a dielectric tank of coloured medium (volume_scatter + transmittance walks)
    ("media_96x64_64spp", "media.pbrt", None, 64, None, None),
    ("media_96x64_1024spp", "media.pbrt", None, 1024, None, None),
    # camera inside a haze, spot / distant / infinite lights, spheres (the reference aborts on a NaN pixel at 1024 spp)
    ("media2_96x64_64spp", "media2.pbrt", None, 64, None, None),
    # Loop subdivision surfaces (closed valence-4 / valence-3 meshes, an open sheet), limit normals
    ("loopsubdiv_96x64_64spp", "loopsubdiv.pbrt", None, 64, None, None),
    ("loopsubdiv_96x64_1024spp", "loopsubdiv.pbrt", None, 1024, None, None),
    # Shape "disk": annuli, partial sweeps, mirrored / squashed / quarter-turned
    ("disk_96x64_64spp", "disk.pbrt", None, 64, None, None),
    ("disk_96x64_1024spp", "disk.pbrt", None, 1024, None, None),
    # image / procedural textures (needs tests/scenes/textures/ beside the scene copy)
    ("textures_96x64_64spp", "textures.pbrt", None, 64, None, None),
    ("textures_96x64_1024spp", "textures.pbrt", None, 1024, None, None),
    # image infinite light (octahedral environment map)
    ("envmap_96x64_64spp", "envmap.pbrt", None, 64, None, None),
    ("envmap_96x64_1024spp", "envmap.pbrt", None, 1024, None, None),
    # C4 in miniature (workloads.procedural_scene, seed 7): many lights -> statistical gate only
    ("c4tiny_96x54_1024spp", "@c4tiny", None, 1024, None, None),
    # ---- round 2 ------------------------------------------------------------------------------------------------------
    # russian roulette (util/russian_roulette.h:12-29) fires only for 8 <= depth < maxdepth: maxdepth 12
    ("cornell_64_64spp_maxdepth12", "cornell-box-specular.pbrt", (64, 64), 64, None, None, [('Integrator "path"', 'Integrator "path" "integer maxdepth" [12]')]),
    ("materials_96x64_64spp_maxdepth12", "materials.pbrt", None, 64, None, None, [('"integer maxdepth" [7]', '"integer maxdepth" [12]')]),
    # thin-lens camera (cameras/perspective.cu:72-80)
    ("cornell_64_64spp_lens", "cornell-box-specular.pbrt", (64, 64), 64, None, None,
     [('Camera "perspective" "float fov" [37.5]', 'Camera "perspective" "float fov" [37.5] "float lensradius" [12] "float focaldistance" [900]')]),
    ("materials_96x64_64spp_lens", "materials.pbrt", None, 64, None, None,
     [('Camera "perspective" "float fov" [40]', 'Camera "perspective" "float fov" [40] "float lensradius" [0.05] "float focaldistance" [6]')]),
    # "bool regularize" (integrators/wavefront_path.cu:446-447, util/scattering.h:216-224)
    ("materials_96x64_64spp_regularize", "materials.pbrt", None, 64, None, None, [('"integer maxdepth" [7]', '"integer maxdepth" [7] "bool regularize" [true]')]),
    ("cornell_64_64spp_regularize", "cornell-box-specular.pbrt", (64, 64), 64, None, None, [('Integrator "path"', 'Integrator "path" "bool regularize" [true]')]),
    # C3's image at a size between 48x48 and full HD, converged: 480x270 at 4096 spp (about a minute of reference time)
    ("cornell_480x270_4096spp", "cornell-box-specular.pbrt", (480, 270), 4096, None, None),
    # emissive spheres (Sphere::sample / Sphere::pdf, shapes/sphere.cu:162-280)
    ("spherelight_96x64_64spp", "spherelight.pbrt", None, 64, None, None),
    ("spherelight_96x64_1024spp", "spherelight.pbrt", None, 1024, None, None),
    # planar / spherical / cylindrical texture mappings, float textures on conductor / dielectric roughness
    ("textures2_96x64_64spp", "textures2.pbrt", None, 64, None, None),
    ("textures2_96x64_1024spp", "textures2.pbrt", None, 1024, None, None),
    # occluders inside the margins of the reference's three-ray visibility test (base/integrator_base.cu:12-71)
    ("occluders_96x64_64spp", "occluders.pbrt", None, 64, None, None),
    ("occluders_96x64_1024spp", "occluders.pbrt", None, 1024, None, None),
    # spheres on medium transitions (a glass ball of dye, a diffuse-transmission ball of fog)
    ("mediaball_96x64_64spp", "mediaball.pbrt", None, 64, None, None),
    ("mediaball_96x64_1024spp", "mediaball.pbrt", None, 1024, None, None),
    # material-less quadrics: balls of smoke behind Material "interface" (a diffuse ball inside one of them)
    ("smokeball_96x64_64spp", "smokeball.pbrt", None, 64, None, None),
    ("smokeball_96x64_1024spp", "smokeball.pbrt", None, 1024, None, None),
    # spectrum textures (mix over a float image, scale, directionmix, constant) on conductor / coatedconductor eta and k
    ("metals_96x64_64spp", "metals.pbrt", None, 64, None, None),
    ("metals_96x64_1024spp", "metals.pbrt", None, 1024, None, None),
    # the same textures under a coating (LayeredBxDF walks are hash-seeded: a statistical golden, like coated.pbrt)
    ("metals_coated_96x64_1024spp", "metals.pbrt", None, 1024, None, None,
     [('Material "conductor" "texture eta" ["inlay-eta"] "texture k" ["inlay-k"] "float roughness" [0.12]',
       'Material "coatedconductor" "texture conductor.eta" ["inlay-eta"] "texture conductor.k" ["inlay-k"] "float interface.roughness" [0.02] "float conductor.roughness" [0.1] "float thickness" [0.02]')]),
]



ROUND2_START = [c[0] for c in CASES].index("cornell_64_64spp_maxdepth12")


def main():
    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    exe = os.path.join(ROOT, "oracle", "_ref", "pbrt-minus-ref")
    manifest = {}
    only = sys.argv[1:]
    import shutil
    shutil.copytree(os.path.join(ROOT, "tests", "scenes", "textures"), os.path.join(out, "textures"), dirs_exist_ok=True)
    for case in CASES:
        name, scene, res, spp, integ, sampler = case[:6]
        edits = case[6] if len(case) > 6 else []
        if only == ["round2"]:
            if CASES.index(case) < ROUND2_START:
                continue
        elif only and name not in only:
            continue
        if scene == "@c4tiny":
            sys.path.insert(0, ROOT)
            import importlib
            pkg = importlib.import_module("pbrt-minus_b200")
            path, _ = pkg.workloads.procedural_scene(out, **pkg.workloads.C4_TINY)
            text = open(path).read()
        else:
            text = open(os.path.join(ROOT, "tests", "scenes", scene)).read()
        if res:
            text = with_resolution(text, *res)
        if sampler:
            text = text.replace("WorldBegin", 'Sampler "%s"\nWorldBegin' % sampler, 1)
        for a, b in edits:
            assert a in text, (name, a)
            text = text.replace(a, b)
        sp = os.path.join(out, name + ".pbrt")
        open(sp, "w").write(text)
        png = os.path.join(out, name + ".png")
        cmd = [exe, sp, "--spp", str(spp), "--outfile", png] + (["--integrator", integ] if integ else [])
        r = subprocess.run(cmd, capture_output=True, text=True, cwd=out, timeout=3000)
        if r.returncode != 0 or not os.path.exists(png + ".f32"):
            print(name, "FAILED", (r.stdout + r.stderr)[-400:])
            continue
        m = re.search(r'xresolution"\s*\[\s*(\d+)', text)
        n = re.search(r'yresolution"\s*\[\s*(\d+)', text)
        W, H = int(m.group(1)), int(n.group(1))
        film = np.fromfile(png + ".f32", dtype=np.float32).reshape(H, W, 3)
        os.replace(png + ".f32", os.path.join(out, name + ".f32"))
        took = re.search(r"rendering took\s+([0-9.]+)", r.stdout)
        manifest[name] = {"scene": scene, "width": W, "height": H, "spp": spp, "integrator": integ or "path", "sampler": sampler or "stratified", "edits": [list(e) for e in edits],
                          "mean": float(film.mean()), "reference_render_s": float(took.group(1)) if took else None}
        print(name, manifest[name])
    json.dump(manifest, open(os.path.join(out, "manifest_new.json" if only else "manifest.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
