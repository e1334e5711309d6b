"""Host-side mirror of the reference's interface for the hot path, over the C-ABI.

Names follow the reference: `HLBVH(primitives)` builds the accelerator (accelerator/hlbvh.h:76-78),
`intersect` / `fast_intersect` are its closest-hit / any-hit queries (hlbvh.h:96-100) -- batched
here, and returning primitive ids (the reference returns an interaction; the id is what parity is
defined on).  `WavefrontPathIntegrator.render(film)` mirrors integrators/wavefront_path.h:136-140.
"""
import ctypes as C

import numpy as np

from . import api
from .api import _check, _f32, _fp, _i32, _ip


class DeviceBuffer:
    """A raw device allocation owned by a scene (plain pointer + size; no torch types)."""

    def __init__(self, scene, nbytes):
        self.scene = scene
        self.nbytes = int(nbytes)
        p = C.c_void_p()
        _check(scene.lib.b200pt_device_alloc(scene.handle, self.nbytes, C.byref(p)))
        self.ptr = p.value

    def upload(self, array):
        a = np.ascontiguousarray(array)
        assert a.nbytes <= self.nbytes
        _check(self.scene.lib.b200pt_copy_to_device(self.scene.handle, self.ptr, a.ctypes.data, a.nbytes))

    def download(self, dtype, count):
        out = np.empty(count, dtype=dtype)
        assert out.nbytes <= self.nbytes
        _check(self.scene.lib.b200pt_copy_to_host(self.scene.handle, out.ctypes.data, self.ptr, out.nbytes))
        return out

    def free(self):
        if self.ptr:
            self.scene.lib.b200pt_device_free(self.scene.handle, self.ptr)
            self.ptr = None


class HLBVH:
    """Scene replica on one GPU + its accelerator.  `desc_extra` carries the render-side arrays."""

    def __init__(self, P, indices, device=0, build=True, desc=None, keep=None):
        self.lib = api.load_library()
        self._keep = [keep]
        if desc is None:
            P = _f32(P).reshape(-1, 3)
            indices = _i32(indices).reshape(-1, 3)
            desc = api.SceneDesc()
            desc.n_vertices = P.shape[0]
            desc.P = _fp(P)
            desc.n_triangles = indices.shape[0]
            desc.indices = _ip(indices)
            self._keep += [P, indices]
        self.n_triangles = int(desc.n_triangles)
        h = C.c_void_p()
        _check(self.lib.b200pt_scene_create(C.byref(desc), int(device), C.byref(h)))
        self.handle = h
        self.build_stats = None
        if build:
            self.build()

    def build(self):
        st = api.BuildStats()
        _check(self.lib.b200pt_bvh_build(self.handle, C.byref(st)))
        self.build_stats = st.as_dict()
        return self.build_stats

    def validate(self):
        """b200pt_bvh_validate: {bad_boxes, triangle_refs, nodes} -- a conservative tree has bad_boxes == 0"""
        out = (C.c_int64 * 3)()
        _check(self.lib.b200pt_bvh_validate(self.handle, out))
        return {"bad_boxes": out[0], "triangle_refs": out[1], "nodes": out[2]}

    def sorted_order(self):
        prim = np.empty(self.n_triangles, dtype=np.int32)
        keys = np.empty(self.n_triangles, dtype=np.uint64)
        _check(self.lib.b200pt_bvh_debug_sorted(self.handle, _ip(prim), keys.ctypes.data_as(C.POINTER(C.c_uint64))))
        return prim, keys

    def _trace(self, o, d, tmax, any_hit):
        o = _f32(o).reshape(-1, 3)
        d = _f32(d).reshape(-1, 3)
        n = o.shape[0]
        tmax = _f32(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
        prim = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float32)
        b = np.empty((n, 2), dtype=np.float32)
        _check(self.lib.b200pt_trace(self.handle, _fp(o), _fp(d), _fp(tmax), n, int(any_hit), _ip(prim), _fp(t), _fp(b)))
        return prim, t, b

    def intersect(self, o, d, tmax=np.inf):
        """closest hit: (prim ids [-1 = miss], t, (b1, b2))"""
        return self._trace(o, d, tmax, 0)

    def fast_intersect(self, o, d, tmax=np.inf):
        """any hit within tmax: bool array"""
        return self._trace(o, d, tmax, 1)[0].astype(bool)

    # ---- device-resident form (bench `value`: inputs already in HBM)
    def alloc(self, nbytes):
        return DeviceBuffer(self, nbytes)

    def pack_rays(self, o, d, tmax):
        o = _f32(o).reshape(-1, 3)
        d = _f32(d).reshape(-1, 3)
        n = o.shape[0]
        r = np.zeros((n, 8), dtype=np.float32)
        r[:, 0:3] = o
        r[:, 3] = np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,))
        r[:, 4:7] = d
        return r

    def trace_device(self, rays_buf, n, hits_buf, any_hit=False, sync=False):
        _check(self.lib.b200pt_trace_device(self.handle, rays_buf.ptr, int(n), int(any_hit), hits_buf.ptr, int(sync)))

    def trace_device_batch(self, sets, sync=False):
        """sets: [(rays_buf, n, hits_buf, any_hit)]; b200pt_trace_device_batch (two traversal streams, one call)"""
        arr = (api.DeviceRaySet * len(sets))()
        for i, (rb, n, hb, any_hit) in enumerate(sets):
            arr[i].rays_dev, arr[i].n, arr[i].any_hit, arr[i].hits_dev = rb.ptr, int(n), int(bool(any_hit)), hb.ptr
        _check(self.lib.b200pt_trace_device_batch(self.handle, arr, len(sets), int(sync)))

    def set_counting(self, enable):
        _check(self.lib.b200pt_trace_set_counting(self.handle, int(enable)))

    TUNE_KEYS = {"refill": 0, "tri_steps": 1, "ctas_per_sm": 2, "wf_refill": 3, "wf_tri_steps": 4, "shade_occ": 5, "stash": 6, "build_scratch_mb": 7}

    def tune(self, **kw):
        """scheduling knobs of the traversal kernels (b200pt_trace_tune): they change speed, never results"""
        for k, v in kw.items():
            _check(self.lib.b200pt_trace_tune(self.handle, self.TUNE_KEYS[k], int(v)))

    def counters(self):
        out = (C.c_int64 * 3)()
        _check(self.lib.b200pt_trace_counters(self.handle, out))
        return {"nodes": out[0], "tris": out[1], "rays": out[2]}

    def sync(self):
        _check(self.lib.b200pt_scene_sync(self.handle))

    @property
    def stream(self):
        return self.lib.b200pt_scene_stream(self.handle)

    def probe_triangle(self, o, d, tmax, tri):
        o = _f32(o).reshape(-1, 3)
        d = _f32(d).reshape(-1, 3)
        tri = _f32(tri).reshape(-1, 9)
        n = o.shape[0]
        tmax = _f32(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
        hit = np.empty(n, dtype=np.int32)
        bt = np.empty((n, 4), dtype=np.float32)
        _check(self.lib.b200pt_probe_triangle(self.handle, _fp(o), _fp(d), _fp(tmax), _fp(tri), n, _ip(hit), _fp(bt)))
        return hit, bt

    def probe_sampler(self, kind, spp, pixel, sample, pattern):
        pixel = _i32(pixel)
        sample = _i32(sample)
        pattern = _i32(pattern)
        per = int(sum(1 if p == 1 else 2 for p in pattern))
        out = np.empty((len(pixel), per), dtype=np.float32)
        _check(self.lib.b200pt_probe_sampler(self.handle, int(kind), int(spp), _ip(pixel), _ip(sample), len(pixel),
                                             _ip(pattern), len(pattern), _fp(out)))
        return out

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200pt_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Film:
    """RGBFilm (film/rgb_film.h:16-75) as a device accumulator: {sum X*w, sum Y*w, sum Z*w, sum w} per pixel."""

    def __init__(self, accel):
        self.accel = accel
        self.lib = accel.lib
        h = C.c_void_p()
        _check(self.lib.b200pt_film_create(accel.handle, C.byref(h)))
        self.handle = h
        w, hh = C.c_int32(), C.c_int32()
        _check(self.lib.b200pt_film_dims(self.handle, C.byref(w), C.byref(hh)))
        self.width, self.height = w.value, hh.value

    def clear(self):
        _check(self.lib.b200pt_film_clear(self.handle))

    def get_pixel_rgb(self):
        """[H, W, 3] linear output RGB (RGBFilm::get_pixel_rgb, film/rgb_film.cu:107-122)."""
        out = np.empty((self.height, self.width, 3), dtype=np.float32)
        _check(self.lib.b200pt_film_read(self.handle, _fp(out)))
        return out

    def raw(self):
        out = np.empty((self.height, self.width, 4), dtype=np.float32)
        _check(self.lib.b200pt_film_read_raw(self.handle, _fp(out)))
        return out

    @property
    def device_ptr(self):
        return self.lib.b200pt_film_device_ptr(self.handle)

    def reduce(self, nccl_comm, root=0, stream=None):
        _check(self.lib.b200pt_film_reduce(self.handle, nccl_comm, int(root), stream))

    def srgb8(self):
        """[H, W, 4] uint8: the reference's 8-bit sRGB pixels (Film::write_to_png, base/film.cu:57-99), encoded on the device;
        also returns the number of NaN pixels"""
        out = np.empty((self.height, self.width, 4), dtype=np.uint8)
        nan = C.c_int32(0)
        _check(self.lib.b200pt_film_read_srgb8(self.handle, out.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(nan)))
        return out, nan.value

    def srgb8_static(self):
        """srgb8() through b200pt_film_read_srgb8_static: no device allocation, films of at most 2,097,152 pixels"""
        out = np.empty((self.height, self.width, 4), dtype=np.uint8)
        nan = C.c_int32(0)
        _check(self.lib.b200pt_film_read_srgb8_static(self.handle, out.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(nan)))
        return out, nan.value

    def write_to_png(self, path):
        """Film::write_to_png (base/film.cu:57-99): sRGB-encode on the device, 8 bit; raises on NaN pixels as the reference aborts."""
        from . import png
        rgba, nan = self.srgb8()
        if nan:
            raise api.B200ptError(-4, "Film.write_to_png: %d pixels carry NaN components" % nan)
        png.write_png(path, rgba[..., :3])

    def close(self):
        if getattr(self, "handle", None):
            self.lib.b200pt_film_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def srgb8(rgb):
    """spectrum_util/color_encoding.h: linear -> sRGB 8 bit."""
    v = np.clip(rgb, 0.0, 1.0)
    e = np.where(v <= 0.0031308, 12.92 * v, 1.055 * np.power(v, 1.0 / 2.4) - 0.055)
    return np.clip(np.round(255.0 * e), 0, 255).astype(np.uint8)


class WavefrontPathIntegrator:
    """integrators/wavefront_path.h:136-140: WavefrontPathIntegrator(spp, sampler, params, base) + render(film).
    `scene` is a pbrt_scene.Scene; the accelerator is built here (scene_builder.cu:912)."""

    def __init__(self, scene, device=0, spp=None):
        self.scene = scene
        desc = scene.to_desc()
        self.accel = HLBVH(None, None, device=device, desc=desc, keep=scene)
        self.spp = spp if spp is not None else scene.spp
        self.last_stats = None

    def create_film(self):
        return Film(self.accel)

    def render(self, film, sample_begin=0, sample_end=None, pool_size=0):
        o = api.RenderOpts()
        o.spp_total = int(self.spp)
        o.sample_begin = int(sample_begin)
        o.sample_end = int(self.spp if sample_end is None else sample_end)
        o.max_depth = int(self.scene.max_depth)
        o.regularize = int(self.scene.regularize)
        o.sampler = 1 if self.scene.sampler == "independent" else 0
        o.pool_size = int(pool_size)
        st = api.RenderStats()
        _check(self.accel.lib.b200pt_render(self.accel.handle, film.handle, C.byref(o), C.byref(st)))
        self.last_stats = st.as_dict()
        return self.last_stats

    def close(self):
        self.accel.close()


def sample_range(spp, rank, world):
    """Sample-index sharding: rank r of G renders sample_idx in [r*spp/G, (r+1)*spp/G) of every
    pixel, with the sampler configured for the TOTAL spp so strata stay global.  Ranges tile [0, spp)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank %d of %d" % (rank, world))
    return (rank * spp) // world, ((rank + 1) * spp) // world


class NcclComm:
    """A raw ncclComm_t for b200pt_film_reduce, created next to a torch.distributed process group: rank 0
    makes the unique id, the group broadcasts its 128 bytes, every rank calls ncclCommInitRank.  NCCL is the
    copy torch loaded (libnccl.so.2), opened RTLD_GLOBAL so libb200pt.so resolves ncclReduce from it."""

    def __init__(self, rank, world, device):
        import glob
        import os
        import sys

        import torch
        import torch.distributed as dist
        cands = []
        for p in sys.path:
            cands += glob.glob(os.path.join(p, "nvidia", "nccl", "lib", "libnccl.so*"))
        cands += ["libnccl.so.2"]
        self.lib = None
        for c in cands:
            try:
                self.lib = C.CDLL(c, mode=C.RTLD_GLOBAL)
                break
            except OSError:
                continue
        if self.lib is None:
            raise api.B200ptError(-5, "libnccl.so.2 not found")
        uid = (C.c_char * 128)()
        if rank == 0:
            rc = self.lib.ncclGetUniqueId(C.byref(uid))
            if rc != 0:
                raise api.B200ptError(-5, "ncclGetUniqueId failed: %d" % rc)
        t = torch.tensor(list(bytes(uid)), dtype=torch.uint8, device="cuda:%d" % device)
        dist.broadcast(t, src=0)
        uid = (C.c_char * 128).from_buffer_copy(bytes(t.cpu().tolist()))
        torch.cuda.set_device(device)
        self.comm = C.c_void_p()

        class _Uid(C.Structure):
            _fields_ = [("internal", C.c_char * 128)]
        u = _Uid()
        C.memmove(C.byref(u), uid, 128)
        self.lib.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _Uid, C.c_int]
        rc = self.lib.ncclCommInitRank(C.byref(self.comm), int(world), u, int(rank))
        if rc != 0:
            raise api.B200ptError(-5, "ncclCommInitRank failed: %d" % rc)

    def close(self):
        if getattr(self, "comm", None) and self.comm.value:
            self.lib.ncclCommDestroy.argtypes = [C.c_void_p]
            self.lib.ncclCommDestroy(self.comm)
            self.comm = C.c_void_p()
