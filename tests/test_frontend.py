"""CPU tests of the host side above the C-ABI: the .pbrt front-end (lexer, parameter dictionary,
transform stack, scene assembly), the spectral tables / RGB->spectrum fit, filter tables, the PLY
reader, ctypes struct layout vs include/b200pt.h, and the multi-GPU host logic (sample-index sharding +
film reduce) over gloo with world_size 2.  No kernels run here."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CORNELL = os.path.join(ROOT, "tests", "scenes", "cornell-box-specular.pbrt")


def test_lexer_and_parameters(b200pt):
    ps = b200pt.pbrt_scene
    text = '''# comment
    Shape "trianglemesh" "integer indices" [0 1 2] # trailing
       "point3 P" [ 0 0 0  1 0 0  0 1e0 -.5 ]
    Material "dielectric" "bool remaproughness" [ false ] "float eta" 1.5 "spectrum k" ["metal-Ag-k"]
    Translate 1 -2 3.5
    AttributeBegin
    '''
    d = ps.lex(text)
    assert [k for k, _ in d] == ["Shape", "Material", "Translate", "AttributeBegin"]
    p = ps.params_of(d[0][1][1:])
    assert p["indices"] == ("integer", [0.0, 1.0, 2.0]) and p["P"][1][-1] == -0.5
    m = ps.params_of(d[1][1][1:])
    assert ps.p_bool(m, "remaproughness", True) is False and ps.p_float(m, "eta", 0) == 1.5
    assert m["k"] == ("spectrum", ["metal-Ag-k"]) and d[2][1] == [1.0, -2.0, 3.5]
    with pytest.raises(ps.SceneError):
        ps.load(text='Camera "perspective"\nWorldBegin\nFooBar 1\n')          # unknown directive is fatal
    with pytest.raises(ps.SceneError):
        ps.load(text='Camera "perspective"\nWorldBegin\nMaterial "subsurface"\n')   # not a reference material (base/material.cu:31-87)


def test_transforms_follow_the_reference(b200pt):
    ps = b200pt.pbrt_scene
    t = ps.translate(1, 2, 3) * ps.rotate(30, [0, 0, 1]) * ps.scale(2, 1, 0.5)
    assert np.allclose(t.m.astype(np.float64) @ t.inv.astype(np.float64), np.eye(4), atol=1e-6)
    assert np.array_equal(t.inverse().m, t.inv)
    r = ps.rotate(90, [0, 0, 1]).m
    assert np.allclose(r[:3, :3] @ np.array([1, 0, 0]), [0, 1, 0], atol=1e-6)
    cam = ps.lookat([278, 273, -800], [278, 273, 0], [0, 1, 0])
    assert np.allclose(cam.inv[:3, 3], [278, 273, -800])                    # inv = world_from_camera
    P = ps.xf_points(ps.translate(1, 2, 3).m, [[1, 1, 1]])
    assert np.array_equal(P, np.array([[2, 3, 4]], np.float32))
    assert ps.swaps_handedness(ps.scale(1, -1, 1).m) and not ps.swaps_handedness(t.m)


def test_cornell_scene_assembly(b200pt):
    sc = b200pt.pbrt_scene.load(CORNELL, spp=16)
    assert sum(len(i) for i in sc.idx) == 22 and len(sc.spheres) == 1 and len(sc.lights) == 2
    assert sc.film == {"width": 1080, "height": 1080} and sc.filter[:4] == (2.0, 2.0, 64, 64)
    assert sc.spp == 16 and sc.max_depth == 5 and sc.sampler == "stratified" and sc.integrator == "path"
    assert [m["type"] for m in sc.materials] == [0, 0, 0, 0, 2, 1]           # default + grey/red/green + glass + silver
    assert np.array_equal(sc.camera["rfc"], np.eye(4, dtype=np.float32))     # camera-world: rotation only
    P = np.concatenate(sc.P)
    assert np.array_equal(P.min(0), np.float32([-278, -273, 800]))           # world - camera position, exact
    # AreaLightSource "rgb L" [1 1 1] "float scale" [20]: scale / photometric integral of the illuminant
    assert abs(sc.lights[0]["scale"] - 20.0 / 106.856895) < 1e-5
    ag = sc.spectra[sc.materials[5]["eta"]]
    assert ag["type"] == 4 and ag["lambda"][0] < 360 and ag["lambda"][-1] > 830 and len(ag["lambda"]) == len(ag["value"])
    assert sc.materials[5]["ur"] == 0.0025 and sc.materials[5]["remap"] == 1 and sc.materials[4]["remap"] == 0
    d = sc.to_desc()
    assert d.n_triangles == 22 and d.n_spheres == 1 and d.n_lights == 2 and d.n_materials == 6
    assert b200pt.pbrt_scene.load(CORNELL).spp == 4                          # CLI-less default (scene_builder.cu:249-251)


def test_rgb_to_spectrum_fit(b200pt):
    sp = b200pt.spectral
    lam = np.arange(360, 831, dtype=np.float64)
    x, y, z = (a.astype(np.float64) for a in sp.cie_xyz())
    ill = sp.illuminant_d65()[2].astype(np.float64)
    M = sp.srgb_color_space()["rgb_from_xyz"]
    for rgb in [(0.65, 0.05, 0.05), (0.12, 0.45, 0.15), (0.2, 0.3, 0.7), (0.9, 0.5, 0.3)]:
        c = sp.rgb_to_sigmoid(rgb).astype(np.float64)
        v = (c[0] * lam + c[1]) * lam + c[2]
        s = 0.5 + v / (2 * np.sqrt(1 + v * v))
        xyz = np.array([(s * ill * a).sum() for a in (x, y, z)]) / (ill * y).sum()
        assert np.allclose(M @ xyz, rgb, atol=0.02), (rgb, M @ xyz)          # reflectance under D65 round-trips
    g = sp.rgb_to_sigmoid((0.73, 0.73, 0.73))
    assert g[0] == 0 and g[1] == 0 and abs(g[2] - (0.73 - 0.5) / np.sqrt(0.73 * 0.27)) < 1e-6
    assert abs(sp.inner_product_dense(sp.illuminant_d65()[2], sp.cie_xyz()[1]) - 106.856895) < 1e-3
    w = sp.srgb_color_space()["w"]
    assert abs(w[0] - 0.3127) < 1e-3 and abs(w[1] - 0.3290) < 1e-3


def test_filter_tables(b200pt):
    ps = b200pt.pbrt_scene
    rx, ry, nx, ny, f = ps.filter_table("mitchell", {})
    assert (rx, ry, nx, ny) == (2.0, 2.0, 64, 64) and f.shape == (64, 64)
    assert np.allclose(f, f[::-1, ::-1]) and np.allclose(f, f.T) and f.min() < 0 < f.max()   # signed lobes
    rx, ry, nx, ny, g = ps.filter_table("gaussian", {})
    assert (nx, ny) == (48, 48) and g.min() >= 0
    assert ps.filter_table("box", {})[2:4] == (16, 16)
    with pytest.raises(ps.SceneError):
        ps.filter_table("sinc", {})


def test_ply_reader(b200pt, tmp_path):
    ps = b200pt.pbrt_scene
    P = np.float32([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]])
    a = tmp_path / "a.ply"
    a.write_text("ply\nformat ascii 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n"
                 "element face 1\nproperty list uchar int vertex_indices\nend_header\n" +
                 "".join("%g %g %g\n" % tuple(p) for p in P) + "4 0 1 2 3\n")
    Pa, ia, _, _ = ps.read_ply(str(a))
    assert np.array_equal(Pa, P) and ia.tolist() == [[0, 1, 2], [0, 2, 3]]
    b = tmp_path / "b.ply"
    hdr = ("ply\nformat binary_little_endian 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\n"
           "element face 2\nproperty list uchar int vertex_indices\nend_header\n").encode()
    faces = b"".join(bytes([3]) + np.int32(t).tobytes() for t in ([0, 1, 2], [0, 2, 3]))
    b.write_bytes(hdr + P.tobytes() + faces)
    Pb, ib, _, _ = ps.read_ply(str(b))
    assert np.array_equal(Pb, P) and ib.tolist() == [[0, 1, 2], [0, 2, 3]]


def test_ctypes_layout_matches_header(b200pt, tmp_path):
    """sizeof of every struct crossing the boundary, as gcc sees include/b200pt.h, equals the ctypes mirror"""
    names = ["b200pt_spectrum", "b200pt_material", "b200pt_light", "b200pt_sphere", "b200pt_camera", "b200pt_filter",
             "b200pt_film_desc", "b200pt_scene_desc", "b200pt_build_stats", "b200pt_render_opts", "b200pt_render_stats", "b200pt_medium", "b200pt_texture", "b200pt_image"]
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "b200pt.h"\nint main(void){' +
                   "".join('printf("%%zu\\n", sizeof(%s));' % n for n in names) + "return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    A = b200pt.api
    mirror = [A.Spectrum, A.Material, A.Light, A.Sphere, A.Camera, A.Filter, A.FilmDesc, A.SceneDesc, A.BuildStats, A.RenderOpts,
              A.RenderStats, A.Medium, A.Texture, A.Image]
    assert sizes == [ctypes.sizeof(m) for m in mirror]


def test_media_front_end(b200pt):
    """MakeNamedMedium / MediumInterface / Material "interface" (scene_builder.cu:166-171,631-676): per-triangle medium codes
    as include/b200pt.h defines them, camera medium = the exterior of the MediumInterface current at WorldBegin"""
    ps = b200pt.pbrt_scene
    sc = ps.load(os.path.join(ROOT, "tests", "scenes", "media.pbrt"))
    assert len(sc.media) == 2 and sc.camera_medium == -1
    # sigma spectra are handed over as the reference keeps them (medium/homogeneous_medium.cu:7-13): DenselySampledSpectrum(sigma,
    # scale), i.e. the spectrum times "scale" tabulated at 360..830 nm in fp32, so the record's own scale is 1
    assert sc.media[0]["g"] == pytest.approx(0.3) and sc.media[1]["scale"] == 1.0
    dye_a, dye_s = sc.spectra[sc.media[1]["sigma_a"]], sc.spectra[sc.media[1]["sigma_s"]]
    assert dye_a["type"] == 5 and len(dye_a["value"]) == 471 and dye_a["value"].dtype == np.float32
    # rgb (0.8 0.2 0.1) * 2: strong in the red, weak in the blue; (0.3 0.6 0.9) * 2 the other way round
    assert dye_a["value"][650 - 360] > 4 * dye_a["value"][450 - 360] > 0 and 1.2 < dye_a["value"][650 - 360] < 2.6
    assert dye_s["value"][450 - 360] > 2 * dye_s["value"][650 - 360] > 0
    codes = np.concatenate(sc.tri_medium)
    per_shape = [int(codes[i]) for i in (0, 2, 4, 6, 18, 30)]
    # light (GeometricPrimitive, no transition), floor, wall, fog box (interior fog, exterior vacuum), block (fog|fog is no
    # transition: SimplePrimitive), dye tank
    assert per_shape == [0xFFFFFFFF, 0, 0, 1, 0, 2]
    assert [m["type"] for m in sc.materials][3] == 7
    quad = 'Shape "trianglemesh" "integer indices" [0 1 2] "point3 P" [0 0 0 1 0 0 0 1 0]\n'
    head = 'Camera "perspective"\n'
    inside = ps.load(text='MakeNamedMedium "m" "string type" ["homogeneous"]\nMediumInterface "" "m"\n' + head + "WorldBegin\n" + quad)
    assert inside.camera_medium == 0 and inside.media[0]["scale"] == 1.0
    with pytest.raises(ps.SceneError):
        ps.load(text=head + 'WorldBegin\nMediumInterface "nope" ""\n' + quad)
    with pytest.raises(ps.SceneError):
        ps.load(text=head + 'WorldBegin\nMakeNamedMedium "m" "string type" ["uniformgrid"]\n')
    ball = ps.load(text=head + 'WorldBegin\nMakeNamedMedium "m" "string type" ["homogeneous"]\nMediumInterface "m" ""\nShape "sphere"\n')
    assert ball.spheres[0]["medium"] == 1 and ball.to_desc().spheres[0].medium == 1     # interior m (id 0 -> 1), exterior vacuum


def test_loop_subdivision_mesh_producer(b200pt):
    """pbrt-minus_b200/loop_subdiv.py against what shapes/loop_subdivide.cu guarantees: 4^levels faces, V - E + F preserved,
    the symmetric octahedron stays symmetric and converges to a sphere-like limit surface, a flat regular sheet stays flat with
    its boundary on the original outline, normals of a flat sheet are parallel"""
    import importlib
    ls = importlib.import_module("pbrt-minus_b200.loop_subdiv")
    P = [0, 0, 1, 1, 0, 0, 0, 1, 0, -1, 0, 0, 0, -1, 0, 0, 0, -1]
    I = [0, 1, 2, 0, 2, 3, 0, 3, 4, 0, 4, 1, 5, 2, 1, 5, 3, 2, 5, 4, 3, 5, 1, 4]
    for levels in (0, 1, 3):
        Pl, idx, N = ls.loop_subdivide(levels, I, P)
        assert idx.shape == (8 * 4 ** levels, 3) and Pl.dtype == np.float32 and idx.dtype == np.int32
        edges = {tuple(sorted((int(t[k]), int(t[(k + 1) % 3])))) for t in idx for k in range(3)}
        assert len(Pl) - len(edges) + len(idx) == 2                       # closed genus-0 surface
        assert idx.min() == 0 and idx.max() == len(Pl) - 1
    r = np.linalg.norm(Pl, axis=1)
    assert 0.40 < r.min() and r.max() <= 0.5 + 1e-6                        # the six original vertices end at 0.5
    assert np.allclose(sorted(map(tuple, np.round(np.abs(Pl), 5))), sorted(map(tuple, np.round(np.abs(Pl[:, [1, 2, 0]]), 5))))
    cosang = np.einsum("ij,ij->i", N, Pl) / (np.linalg.norm(N, axis=1) * r)
    assert (np.abs(cosang) > 0.93).all() and (np.sign(cosang) == np.sign(cosang[0])).all()
    # flat 4x4 regular grid
    g = [(x, y, 0) for y in range(4) for x in range(4)]
    gi = []
    for y in range(3):
        for x in range(3):
            a = y * 4 + x
            gi += [a, a + 1, a + 5, a, a + 5, a + 4]
    Pl, idx, N = ls.loop_subdivide(2, gi, np.array(g, np.float32).ravel())
    assert idx.shape == (18 * 16, 3) and np.abs(Pl[:, 2]).max() == 0
    assert Pl[:, :2].min() >= 0 and Pl[:, :2].max() <= 3
    assert np.abs(N[:, :2]).max() < 1e-5 and (N[:, 2] != 0).all() and (np.sign(N[:, 2]) == np.sign(N[0, 2])).all()
    ps = b200pt.pbrt_scene
    sc = ps.load(os.path.join(ROOT, "tests", "scenes", "loopsubdiv.pbrt"))
    assert sum(len(i) for i in sc.idx) == 4 + 8 * 64 + 4 * 64 + 18 * 16
    with pytest.raises(ps.SceneError):
        ps.load(text='Camera "perspective"\nWorldBegin\nShape "loopsubdiv" "integer indices" [0 1 2] "point3 P" [0 0 0 1 0 0 0 1 0 5 5 5]\n')


def test_texture_front_end(b200pt, tmp_path):
    """Texture directive -> texture records (children before parents), images through the sRGB table, materials naming textures"""
    ps = b200pt.pbrt_scene
    sc = ps.load(os.path.join(ROOT, "tests", "scenes", "textures.pbrt"))
    assert len(sc.images) == 2 and sc.images[0].shape == (16, 32, 3) and sc.images[0].dtype == np.float32
    assert sc.images[0].min() == 0.0 and sc.images[0].max() == 1.0
    assert abs(float(sc.images[0][5, 5, 0]) - 0.5775804) < 1e-6          # sRGB 200 -> linear (SRGBToLinearLUT[200])
    T = sc.textures
    for i, t in enumerate(T):
        kids = [t["a"], t["b"]] if t["type"] in (3, 4, 18) else [t["a"]] if t["type"] == 2 else []
        kids += [t["c"]] if t["type"] == 3 and t["c"] >= 0 else []
        assert all(0 <= k < i for k in kids)
    named = sc.named_spec_tex
    assert T[named["tiled"]]["map"] == [3.0, 2.5, 0.25, -0.1] and T[named["tiled"]]["f"][0] == pytest.approx(0.8)
    assert T[named["negative"]]["f"][1] == 1.0 and T[named["updown"]]["type"] == 4 and T[named["updown"]]["f"][:3] == [0.0, 1.0, 0.0]
    assert T[sc.named_float_tex["softmask"]]["type"] == 18
    mats = sc.materials
    assert [m["reflectance_tex1"] > 0 for m in mats] == [False, True, True, True, True, True, True]
    assert mats[5]["transmittance_tex1"] == named["dim"] + 1 and mats[6]["type"] == 1 and mats[6]["eta"] == -1
    d = sc.to_desc()
    assert d.n_textures == len(T) and d.n_images == 2 and bool(d.rgb2spec_fit)
    head = 'Camera "perspective"\nWorldBegin\n'
    for bad in ('Texture "t" "spectrum" "checkerboard"', 'Texture "t" "spectrum" "imagemap" "string filename" ["nope.png"]',
                'Texture "t" "spectrum" "imagemap" "string filename" ["x.exr"]', 'Texture "t" "spectrum" "wrinkled"',
                'Material "diffuse" "texture reflectance" ["undefined"]', 'Texture "t" "float" "mix"',
                'Texture "t" "spectrum" "imagemap" "string filename" ["textures/colours.png"] "string mapping" ["hexagonal"]'):
        with pytest.raises(ps.SceneError):
            ps.load(text=head + bad + "\n", scene_dir=os.path.join(ROOT, "tests", "scenes"))


def test_sample_ranges_tile(b200pt):
    for spp in (1, 4, 16, 1024, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            r = [b200pt.sample_range(spp, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp and all(r[k][1] == r[k + 1][0] for k in range(world - 1))
    with pytest.raises(ValueError):
        b200pt.sample_range(16, 2, 2)


_WORKER = r'''
import importlib, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
pkg = importlib.import_module("pbrt-minus_b200")
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
spp, npix = 64, 1000
def partial_film(b, e):            # stand-in for b200pt_render over [b, e): per-sample contributions are pure functions
    k = np.arange(b, e, dtype=np.float64)[:, None]
    p = np.arange(npix, dtype=np.float64)[None, :]
    w = 1.0 + 0.25 * np.sin(k + p)
    v = np.cos(0.1 * k * (1 + p % 7))
    return np.stack([(w * v).sum(0), (w * v * 0.5).sum(0), (w * v * 2).sum(0), w.sum(0)], axis=1)
b, e = pkg.sample_range(spp, rank, world)
film = torch.from_numpy(partial_film(b, e))
dist.reduce(film, dst=0, op=dist.ReduceOp.SUM)     # what b200pt_film_reduce does with ncclReduce on the GPUs
if rank == 0:
    full = partial_film(0, spp)
    assert np.allclose(film.numpy(), full, rtol=1e-12, atol=1e-9), "sharded film != full film"
    print("OK", world)
dist.destroy_process_group()
'''


def test_sample_sharding_and_film_reduce_gloo_world2(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(_WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29653", str(w), ROOT], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK 2" in r.stdout, r.stdout[-500:] + r.stderr[-1500:]
