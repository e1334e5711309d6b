"""ctypes binding of libb200pt.so (include/b200pt.h) -- the test/bench harness' view of the C-ABI.

No torch types cross this boundary and there is NO CPU fallback: if the CUDA library is missing
or no device is present every call raises B200ptError (the product path must fail loudly).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200PT_LIB") or os.path.join(HERE, "libb200pt.so")   # B200PT_LIB: another build of the library (A/B runs)


class B200ptError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b200pt error %d: %s" % (code, msg))
        self.code = code


class Spectrum(C.Structure):
    _fields_ = [("type", C.c_int32), ("c", C.c_float * 4), ("n", C.c_int32),
                ("lambda_", C.POINTER(C.c_float)), ("value", C.POINTER(C.c_float))]


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("reflectance", C.c_int32), ("eta", C.c_int32), ("k", C.c_int32),
                ("u_roughness", C.c_float), ("v_roughness", C.c_float), ("remap_roughness", C.c_int32),
                ("transmittance", C.c_int32), ("scale", C.c_float),
                ("thickness", C.c_float), ("g", C.c_float), ("albedo", C.c_int32), ("max_depth", C.c_int32),
                ("n_samples", C.c_int32), ("c_u_roughness", C.c_float), ("c_v_roughness", C.c_float),
                ("c_eta", C.c_int32), ("c_k", C.c_int32), ("mix_a", C.c_int32), ("mix_b", C.c_int32), ("reflectance_tex1", C.c_int32), ("transmittance_tex1", C.c_int32),
                ("u_roughness_tex1", C.c_int32), ("v_roughness_tex1", C.c_int32), ("eta_tex1", C.c_int32), ("k_tex1", C.c_int32)]


class Light(C.Structure):
    _fields_ = [("triangle", C.c_int32), ("spectrum", C.c_int32), ("scale", C.c_float), ("two_sided", C.c_int32),
                ("type", C.c_int32), ("v", C.c_float * 3), ("light_from_render", C.c_float * 9),
                ("cos_falloff_start", C.c_float), ("cos_falloff_end", C.c_float),
                ("image", C.c_int32), ("render_from_light", C.c_float * 9)]


class Sphere(C.Structure):
    _fields_ = [("render_from_object", C.c_float * 16), ("object_from_render", C.c_float * 16),
                ("radius", C.c_float), ("z_min", C.c_float), ("z_max", C.c_float), ("phi_max", C.c_float),
                ("reverse_orientation", C.c_int32), ("transform_swaps_handedness", C.c_int32),
                ("material", C.c_int32), ("light", C.c_int32), ("kind", C.c_int32), ("medium", C.c_uint32)]


class Camera(C.Structure):
    _fields_ = [("camera_from_raster", C.c_float * 16), ("render_from_camera", C.c_float * 16),
                ("lens_radius", C.c_float), ("focal_distance", C.c_float), ("medium", C.c_int32)]


class Medium(C.Structure):
    _fields_ = [("sigma_a", C.c_int32), ("sigma_s", C.c_int32), ("scale", C.c_float), ("g", C.c_float)]


class Filter(C.Structure):
    _fields_ = [("radius_x", C.c_float), ("radius_y", C.c_float), ("nx", C.c_int32), ("ny", C.c_int32),
                ("f", C.POINTER(C.c_float))]


class FilmDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32),
                ("cie_x", C.POINTER(C.c_float)), ("cie_y", C.POINTER(C.c_float)), ("cie_z", C.POINTER(C.c_float)),
                ("imaging_ratio", C.c_float), ("output_rgb_from_sensor_rgb", C.c_float * 9),
                ("illuminant", C.POINTER(C.c_float))]


class Texture(C.Structure):
    _fields_ = [("type", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("c", C.c_int32), ("f", C.c_float * 4), ("map", C.c_float * 4),
                ("mapping", C.c_int32), ("texture_from_render", C.c_float * 12), ("v1", C.c_float * 3), ("v2", C.c_float * 3)]


class Image(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgb", C.POINTER(C.c_float))]


class Object(C.Structure):
    _fields_ = [("first_triangle", C.c_int64), ("n_triangles", C.c_int64), ("bounds", C.c_float * 6)]


class Instance(C.Structure):
    _fields_ = [("object", C.c_int32), ("render_from_instance", C.c_float * 16), ("instance_from_render", C.c_float * 16)]


class SceneDesc(C.Structure):
    _fields_ = [("n_vertices", C.c_int64), ("P", C.POINTER(C.c_float)), ("N", C.POINTER(C.c_float)),
                ("uv", C.POINTER(C.c_float)), ("n_triangles", C.c_int64), ("indices", C.POINTER(C.c_int32)),
                ("tri_material", C.POINTER(C.c_int32)), ("tri_light", C.POINTER(C.c_int32)),
                ("tri_flip", C.POINTER(C.c_uint8)),
                ("n_spheres", C.c_int32), ("spheres", C.POINTER(Sphere)),
                ("n_spectra", C.c_int32), ("spectra", C.POINTER(Spectrum)),
                ("n_materials", C.c_int32), ("materials", C.POINTER(Material)),
                ("n_lights", C.c_int32), ("lights", C.POINTER(Light)),
                ("n_media", C.c_int32), ("media", C.POINTER(Medium)), ("tri_medium", C.POINTER(C.c_uint32)),
                ("n_textures", C.c_int32), ("textures", C.POINTER(Texture)), ("n_images", C.c_int32), ("images", C.POINTER(Image)),
                ("rgb2spec_fit", C.POINTER(C.c_double)),
                ("camera", C.POINTER(Camera)), ("filter", C.POINTER(Filter)), ("film", C.POINTER(FilmDesc)),
                ("n_objects", C.c_int32), ("objects", C.POINTER(Object)), ("n_instances", C.c_int32), ("instances", C.POINTER(Instance))]


class BuildStats(C.Structure):
    _fields_ = [("n_triangles", C.c_int64), ("n_bvh2_nodes", C.c_int64), ("n_bvh8_nodes", C.c_int64),
                ("key_bits", C.c_int32), ("sort_passes", C.c_int32),
                ("ms_bounds_morton", C.c_float), ("ms_sort", C.c_float), ("ms_emit", C.c_float),
                ("ms_refit", C.c_float), ("ms_collapse", C.c_float), ("ms_total", C.c_float),
                ("device_bytes", C.c_int64), ("ms_wall", C.c_float), ("collapse_levels", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RenderOpts(C.Structure):
    _fields_ = [("spp_total", C.c_int32), ("sample_begin", C.c_int32), ("sample_end", C.c_int32),
                ("max_depth", C.c_int32), ("regularize", C.c_int32), ("sampler", C.c_int32),
                ("pool_size", C.c_int32), ("reserved", C.c_int32)]


class RenderStats(C.Structure):
    _fields_ = [("paths", C.c_int64), ("rays_closest", C.c_int64), ("rays_shadow", C.c_int64),
                ("passes", C.c_int64), ("kernel_launches", C.c_int64), ("ms_total", C.c_float),
                ("ms_trace", C.c_float), ("ms_shade", C.c_float), ("ms_generate", C.c_float), ("ms_logic", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class DeviceRaySet(C.Structure):
    _fields_ = [("rays_dev", C.c_void_p), ("n", C.c_int64), ("any_hit", C.c_int32), ("reserved", C.c_int32), ("hits_dev", C.c_void_p)]


class RaySet(C.Structure):
    """b200pt_ray_set (include/b200pt.h)"""
    _fields_ = [("o_xyz", C.POINTER(C.c_float)), ("d_xyz", C.POINTER(C.c_float)), ("tmax", C.POINTER(C.c_float)),
                ("n", C.c_int64), ("any_hit", C.c_int32), ("reserved", C.c_int32),
                ("prim_out", C.POINTER(C.c_int32)), ("t_out", C.POINTER(C.c_float)), ("b1b2_out", C.POINTER(C.c_float))]


# every symbol include/b200pt.h declares (tests check the library exports all of them)
SYMBOLS = [
    "b200pt_abi_version", "b200pt_last_error", "b200pt_device_count",
    "b200pt_scene_create", "b200pt_debug_dump_desc", "b200pt_scene_destroy", "b200pt_bvh_build", "b200pt_bvh_debug_sorted", "b200pt_bvh_validate",
    "b200pt_trace", "b200pt_trace_batch", "b200pt_trace_device", "b200pt_trace_device_batch", "b200pt_trace_set_counting", "b200pt_trace_counters", "b200pt_trace_tune",
    "b200pt_scene_stream", "b200pt_scene_sync",
    "b200pt_device_alloc", "b200pt_device_free", "b200pt_copy_to_device", "b200pt_copy_to_host",
    "b200pt_film_create", "b200pt_film_destroy", "b200pt_film_clear", "b200pt_render", "b200pt_film_reduce",
    "b200pt_film_device_ptr", "b200pt_film_dims", "b200pt_film_read", "b200pt_film_read_srgb8", "b200pt_film_read_srgb8_static", "b200pt_film_read_raw",
    "b200pt_probe_sampler", "b200pt_probe_triangle",
]

_lib = None


def load_library():
    """dlopen libb200pt.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200ptError(-3, "libb200pt.so is not built (run __graft_entry__.build()); there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    lib.b200pt_last_error.restype = C.c_char_p
    lib.b200pt_scene_create.argtypes = [C.POINTER(SceneDesc), i32, C.POINTER(vp)]
    lib.b200pt_debug_dump_desc.argtypes = [C.POINTER(SceneDesc), C.c_char_p]
    lib.b200pt_scene_destroy.argtypes = [vp]
    lib.b200pt_scene_destroy.restype = None
    lib.b200pt_bvh_build.argtypes = [vp, C.POINTER(BuildStats)]
    lib.b200pt_bvh_debug_sorted.argtypes = [vp, ip, C.POINTER(C.c_uint64)]
    lib.b200pt_bvh_validate.argtypes = [vp, C.POINTER(i64)]
    lib.b200pt_trace.argtypes = [vp, fp, fp, fp, i64, i32, ip, fp, fp]
    lib.b200pt_trace_batch.argtypes = [vp, C.POINTER(RaySet), i32]
    lib.b200pt_trace_device_batch.argtypes = [vp, C.POINTER(DeviceRaySet), i32, i32]
    lib.b200pt_trace_device.argtypes = [vp, vp, i64, i32, vp, i32]
    lib.b200pt_trace_set_counting.argtypes = [vp, i32]
    lib.b200pt_trace_counters.argtypes = [vp, C.POINTER(i64)]
    lib.b200pt_trace_tune.argtypes = [vp, i32, i32]
    lib.b200pt_scene_stream.argtypes = [vp]
    lib.b200pt_scene_stream.restype = vp
    lib.b200pt_scene_sync.argtypes = [vp]
    lib.b200pt_device_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    lib.b200pt_device_free.argtypes = [vp, vp]
    lib.b200pt_copy_to_device.argtypes = [vp, vp, vp, C.c_size_t]
    lib.b200pt_copy_to_host.argtypes = [vp, vp, vp, C.c_size_t]
    lib.b200pt_film_create.argtypes = [vp, C.POINTER(vp)]
    lib.b200pt_film_destroy.argtypes = [vp]
    lib.b200pt_film_destroy.restype = None
    lib.b200pt_film_clear.argtypes = [vp]
    lib.b200pt_render.argtypes = [vp, vp, C.POINTER(RenderOpts), C.POINTER(RenderStats)]
    lib.b200pt_film_reduce.argtypes = [vp, vp, i32, vp]
    lib.b200pt_film_device_ptr.argtypes = [vp]
    lib.b200pt_film_device_ptr.restype = vp
    lib.b200pt_film_dims.argtypes = [vp, ip, ip]
    lib.b200pt_film_read.argtypes = [vp, fp]
    lib.b200pt_film_read_raw.argtypes = [vp, fp]
    lib.b200pt_film_read_srgb8.argtypes = [vp, C.POINTER(C.c_uint8), ip]
    lib.b200pt_film_read_srgb8_static.argtypes = [vp, C.POINTER(C.c_uint8), ip]
    lib.b200pt_probe_sampler.argtypes = [vp, i32, i32, ip, ip, i32, ip, i32, fp]
    lib.b200pt_probe_triangle.argtypes = [vp, fp, fp, fp, fp, i64, ip, fp]
    _lib = lib
    return lib


def _check(rc):
    if rc != 0:
        raise B200ptError(rc, load_library().b200pt_last_error().decode("utf-8", "replace"))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32)) if a is not None else None
