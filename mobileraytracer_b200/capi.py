"""ctypes binding of libmobilert_b200.so (include/mobilert_b200.h). Fails loudly when the CUDA library is missing:
there is no CPU fallback anywhere in this package."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MRT_B200_LIB") or os.path.join(_HERE, "libmobilert_b200.so")   # MRT_B200_LIB: diagnostic builds only

MRT_OK, MRT_ERR_INVALID, MRT_ERR_CUDA, MRT_ERR_OOM, MRT_ERR_STOPPED, MRT_ERR_OVERFLOW = range(6)
SHADER_NOSHADOWS, SHADER_WHITTED, SHADER_PATHTRACER, SHADER_DEPTHMAP, SHADER_DIFFUSE = range(5)
ACC_NAIVE, ACC_REGULAR_GRID, ACC_BVH, ACC_B200 = 1, 2, 3, 4
SAMPLER_AUTO, SAMPLER_CONSTANT, SAMPLER_STATIC_HALTON, SAMPLER_STATIC_MT, SAMPLER_STATIC_PCG, SAMPLER_HALTON, SAMPLER_STRATIFIED, SAMPLER_STREAM = range(8)
FLAG_FP32_ACCUMULATE, FLAG_SORT_BY_MATERIAL, FLAG_REPLAY_SAMPLERS, FLAG_PROGRESSIVE = 1, 2, 4, 8
CAMERA_PERSPECTIVE, CAMERA_ORTHOGRAPHIC = 0, 1
STATIC_TRACE_MAX_RAYS = 65536   # MRT_STATIC_TRACE_MAX_RAYS
LIGHT_POINT, LIGHT_AREA = 0, 1

# numpy mirrors of the C records (the reference's own layouts, Triangle.hpp:18-27 etc.)
TRIANGLE = np.dtype([("AC", "<f4", 3), ("AB", "<f4", 3), ("A", "<f4", 3), ("nA", "<f4", 3), ("nB", "<f4", 3),
                     ("nC", "<f4", 3), ("tA", "<f4", 2), ("tB", "<f4", 2), ("tC", "<f4", 2), ("material", "<i4")])
SPHERE = np.dtype([("center", "<f4", 3), ("sq_radius", "<f4"), ("material", "<i4")])
PLANE = np.dtype([("normal", "<f4", 3), ("point", "<f4", 3), ("material", "<i4")])
MATERIAL = np.dtype([("Le", "<f4", 3), ("Kd", "<f4", 3), ("Ks", "<f4", 3), ("Kt", "<f4", 3), ("ior", "<f4")])
LIGHT = np.dtype([("kind", "<i4"), ("position", "<f4", 3), ("triangle", TRIANGLE), ("Le", "<f4", 3)])
assert TRIANGLE.itemsize == 100 and SPHERE.itemsize == 20 and PLANE.itemsize == 28 and MATERIAL.itemsize == 52
assert LIGHT.itemsize == 128


class SceneDesc(C.Structure):
    _fields_ = [("triangles", C.c_void_p), ("n_triangles", C.c_int64), ("spheres", C.c_void_p), ("n_spheres", C.c_int64),
                ("planes", C.c_void_p), ("n_planes", C.c_int64), ("materials", C.c_void_p), ("n_materials", C.c_int64),
                ("lights", C.c_void_p), ("n_lights", C.c_int64),
                ("textures", C.c_void_p), ("n_textures", C.c_int64), ("material_texture", C.c_void_p)]


class Texture(C.Structure):
    """mrt_texture: decoded image as stb_image returns it (rows top to bottom, `channels` bytes per texel)."""
    _fields_ = [("pixels", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32), ("channels", C.c_int32)]


class CameraDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("position", C.c_float * 3), ("look_at", C.c_float * 3), ("up", C.c_float * 3),
                ("a", C.c_float), ("b", C.c_float), ("right", C.c_float * 3)]


class RenderDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("samples_pixel", C.c_int32), ("samples_light", C.c_int32),
                ("shader", C.c_int32), ("accelerator", C.c_int32), ("pixel_sampler", C.c_int32), ("seed", C.c_uint32),
                ("flags", C.c_uint32), ("depth_max_point", C.c_float * 3), ("device", C.c_int32), ("light_sampler", C.c_int32),
                ("tile_begin", C.c_int32), ("tile_stride", C.c_int32)]


class SamplerTables(C.Structure):
    _fields_ = [("tile_order", C.c_void_p), ("shader_table", C.c_void_p), ("static_halton", C.c_void_p),
                ("static_mt", C.c_void_p), ("static_pcg", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("build_ms", C.c_double), ("upload_ms", C.c_double), ("frame_ms", C.c_double), ("extend_ms", C.c_double),
                ("shadow_ms", C.c_double), ("shade_ms", C.c_double), ("other_ms", C.c_double),
                ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_secondary", C.c_uint64),
                ("launches", C.c_uint64), ("extend_launches", C.c_uint64), ("shadow_launches", C.c_uint64),
                ("shade_launches", C.c_uint64), ("rays_extend_traced", C.c_uint64), ("rays_shadow_traced", C.c_uint64), ("bvh_nodes", C.c_uint64), ("bvh_bytes", C.c_uint64), ("scene_bytes", C.c_uint64),
                ("n_primitives", C.c_int64), ("pass_retries", C.c_uint64), ("queue_bytes", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# every symbol include/mobilert_b200.h declares
SYMBOLS = ["mrt_create", "mrt_destroy", "mrt_render_frame", "mrt_render_frame_device", "mrt_render_accumulate_device",
           "mrt_resolve_device", "mrt_set_camera", "mrt_set_pixel_stream", "mrt_set_stream", "mrt_stop", "mrt_set_replay_counter", "mrt_get_replay_counter", "mrt_get_sample", "mrt_get_total_rays", "mrt_get_stats",
           "mrt_last_error", "mrt_trace_rays", "mrt_trace_rays_static", "mrt_generate_rays", "mrt_shade_rays", "mrt_device_count", "mrt_version"]
# every symbol include/mobilert_b200_test.h declares (libmobilert_b200_test.so: the device functions one call at a time)
TEST_SYMBOLS = ["mrt_test_triangle_intersect", "mrt_test_sphere_intersect", "mrt_test_plane_intersect", "mrt_test_aabb_intersect",
                "mrt_test_incremental_avg", "mrt_test_halton", "mrt_test_build_grid", "mrt_test_last_error"]

_lib = None


class MrtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("mobilert_b200 error %d: %s" % (code, message))
        self.code = code


def lib():
    """Loads the CUDA library; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libmobilert_b200.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
        L.mrt_create.argtypes = [C.POINTER(SceneDesc), C.POINTER(CameraDesc), C.POINTER(RenderDesc), C.POINTER(SamplerTables), C.POINTER(vp)]
        L.mrt_destroy.argtypes = [vp]
        L.mrt_destroy.restype = None
        L.mrt_render_frame.argtypes = [vp, vp]
        L.mrt_render_frame_device.argtypes = [vp, vp]
        L.mrt_render_accumulate_device.argtypes = [vp, vp, i32, i32]
        L.mrt_resolve_device.argtypes = [vp, vp, vp]
        L.mrt_set_camera.argtypes = [vp, C.POINTER(CameraDesc)]
        L.mrt_set_stream.argtypes = [vp, vp]
        if hasattr(L, "mrt_set_pixel_stream"):   # absent from older diagnostic builds loaded through MRT_B200_LIB
            L.mrt_set_pixel_stream.argtypes = [vp, vp, i64]
        L.mrt_stop.argtypes = [vp]
        L.mrt_set_replay_counter.argtypes = [vp, C.c_uint32]
        L.mrt_get_replay_counter.argtypes = [vp, C.POINTER(C.c_uint32)]
        L.mrt_get_sample.argtypes = [vp, C.POINTER(i32)]
        L.mrt_get_total_rays.argtypes = [vp, C.POINTER(C.c_uint64)]
        L.mrt_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.mrt_last_error.argtypes = [vp]
        L.mrt_last_error.restype = C.c_char_p
        L.mrt_trace_rays.argtypes = [vp, i64] + [vp] * 10
        if hasattr(L, "mrt_trace_rays_static"):   # absent from older diagnostic builds loaded through MRT_B200_LIB
            L.mrt_trace_rays_static.argtypes = [vp, i64] + [vp] * 10
        L.mrt_generate_rays.argtypes = [vp, i64, vp, vp, vp]
        L.mrt_shade_rays.argtypes = [vp, i64, vp, vp, i32, vp]
        L.mrt_device_count.restype = C.c_int
        L.mrt_version.restype = C.c_char_p
        _lib = L
    return _lib


_test_lib = None
TEST_LIB_PATH = os.path.join(_HERE, "libmobilert_b200_test.so")


def test_lib():
    """libmobilert_b200_test.so (include/mobilert_b200_test.h): raises when it has not been built."""
    global _test_lib
    if _test_lib is None:
        if not os.path.exists(TEST_LIB_PATH):
            raise ImportError("libmobilert_b200_test.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(TEST_LIB_PATH)
        vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
        L.mrt_test_triangle_intersect.argtypes = [vp, vp, vp, f32, i32, vp, vp]
        L.mrt_test_sphere_intersect.argtypes = [vp, vp, vp, f32, vp, vp]
        L.mrt_test_plane_intersect.argtypes = [vp, vp, vp, f32, vp]
        L.mrt_test_aabb_intersect.argtypes = [vp, vp, vp, vp, vp]
        L.mrt_test_incremental_avg.argtypes = [vp, i32, i32, vp]
        L.mrt_test_halton.argtypes = [C.c_uint32, i64, C.c_uint32, vp]
        L.mrt_test_build_grid.argtypes = [C.POINTER(SceneDesc), i32, vp, vp, vp, i64, C.POINTER(i64)]
        L.mrt_test_last_error.restype = C.c_char_p
        _test_lib = L
    return _test_lib


def check_test(rc):
    if rc != MRT_OK:
        msg = test_lib().mrt_test_last_error()
        raise MrtError(rc, msg.decode() if msg else "?")
    return rc


def ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def check(rc, handle=None, allow=()):
    if rc != MRT_OK and rc not in allow:
        msg = lib().mrt_last_error(handle)
        raise MrtError(rc, msg.decode() if msg else "?")
    return rc
