#!/bin/bash
# c4tiny through the bridge: the reference's alias table (B200PT_BRIDGE_DUMP) beside the library's (B200PT_DEBUG_LIGHTS), line by line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export B200PT_REQUIRE_GPU=1
python - <<'PY' 2>&1 | tail -30
import sys, os, re, numpy as np, tempfile, importlib, subprocess
sys.path.insert(0, "tests")
b200pt = importlib.import_module("pbrt-minus_b200")
import test_render_gpu as T
tmp = tempfile.mkdtemp()
m, ref = T.golden("c4tiny_96x54_1024spp")
path, info = b200pt.workloads.procedural_scene(tmp, **b200pt.workloads.C4_TINY)
env = dict(os.environ, B200PT_BRIDGE_DUMP="1", B200PT_DEBUG_LIGHTS="1")
out = os.path.join(tmp, "bridge.png")
r = subprocess.run([os.path.abspath("oracle/_ref/pbrt-minus-b200"), path, "--spp", "1024", "--outfile", out], capture_output=True, text=True, cwd=tmp, env=env, timeout=900)
print("rc", r.returncode)
refl = {int(a): (float(b), float(c), int(d), int(e)) for a, b, c, d, e in re.findall(r"\[reflight\] (\d+) pdf (\S+) bin p (\S+) first (-?\d+) second (-?\d+)", r.stderr)}
ours = {int(a): (float(w), float(b), float(c), int(d), int(e)) for a, w, b, c, d, e in re.findall(r"\[lightdbg\] (\d+) weight (\S+) pdf (\S+) bin p (\S+) first (-?\d+) second (-?\d+)", r.stderr)}
print(len(refl), len(ours), [l for l in r.stderr.splitlines() if "weights differ" in l])
npdf = sum(1 for i in refl if refl[i][0] != ours[i][1]); nbin = sum(1 for i in refl if refl[i][2:] != ours[i][3:]); nbp = sum(1 for i in refl if refl[i][1] != ours[i][2])
print("pdf differs", npdf, "bin (first, second) differs", nbin, "bin p differs", nbp)
shown = 0
for i in sorted(refl):
    if refl[i][0] != ours[i][1] or refl[i][2:] != ours[i][3:]:
        print(i, "ref", refl[i], "ours", ours[i]); shown += 1
        if shown >= 12: break
# the Python front-end's description of the same scene: illuminant table, light scales, spectra, weights
illum_ref = {int(a): b for a, b in re.findall(r"\[refillum\] (\d+) ([0-9a-f]{8})", r.stderr)}
sc = b200pt.pbrt_scene.load(path, spp=4)
ill = np.asarray(b200pt.spectral.illuminant_d65()[2], dtype=np.float32)
bad = [i for i in range(471) if "%08x" % int(ill[i:i+1].view(np.uint32)[0]) != illum_ref[i]]
print("illuminant entries differing", len(bad), bad[:8], [(float(ill[i]), illum_ref[i]) for i in bad[:3]])
recs = re.findall(r"\[reflightrec\] (\d+) tri (-?\d+) spectrum (-?\d+) scale ([0-9a-f]{8}) two_sided (\d)", r.stderr)
ps = sc.lights
nsc = sum(1 for (i, t, sp, scl, two) in recs if "%08x" % int(np.float32(ps[int(i)]["scale"]).view(np.uint32)) != scl)
ntri = sum(1 for (i, t, sp, scl, two) in recs if int(t) != ps[int(i)]["triangle"])
print("light scale bits differing", nsc, "triangle index differing", ntri, "of", len(recs))
r2 = subprocess.run([sys.executable, "-c", "import importlib,sys; sys.path.insert(0,%r); b=importlib.import_module('pbrt-minus_b200'); sc=b.pbrt_scene.load(%r,spp=4); i=b.WavefrontPathIntegrator(sc,device=0)" % (os.getcwd(), path)],
                    capture_output=True, text=True, env=env, timeout=900)
mine = {int(a): (float(w), float(b), float(c), int(d), int(e)) for a, w, b, c, d, e in re.findall(r"\[lightdbg\] (\d+) weight (\S+) pdf (\S+) bin p (\S+) first (-?\d+) second (-?\d+)", r2.stderr)}
print(len(mine), [l for l in r2.stderr.splitlines() if "weights differ" in l])
dw = [i for i in mine if mine[i][0] != ours[i][0]]
print("weights differing python vs bridge", len(dw), [(i, mine[i][0], ours[i][0]) for i in dw[:6]])
film = np.fromfile(out + ".f32", dtype=np.float32).reshape(m["height"], m["width"], 3)
print("bridge vs golden relMSE", T.rel_mse(film, ref), "identical", float((film == ref).all(axis=2).mean()))
PY
