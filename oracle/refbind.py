"""TEST INFRASTRUCTURE (oracle) — ctypes binding of oracle/_ref/libmobilert_ref.so.

That library is the UNMODIFIED reference engine (TiagoMSSantos/MobileRayTracer) compiled by
oracle/Makefile from the sources under /root/reference/app plus the shim in oracle/ref_shim/.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
The product package (mobileraytracer_b200) never does.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libmobilert_ref.so")

TRI_DTYPE = np.dtype([("AC", "<f4", 3), ("AB", "<f4", 3), ("A", "<f4", 3), ("nA", "<f4", 3), ("nB", "<f4", 3),
                      ("nC", "<f4", 3), ("tA", "<f4", 2), ("tB", "<f4", 2), ("tC", "<f4", 2), ("mat", "<i4")])
SPH_DTYPE = np.dtype([("center", "<f4", 3), ("sqRadius", "<f4"), ("mat", "<i4")])
PLA_DTYPE = np.dtype([("normal", "<f4", 3), ("point", "<f4", 3), ("mat", "<i4")])
assert TRI_DTYPE.itemsize == 100 and SPH_DTYPE.itemsize == 20 and PLA_DTYPE.itemsize == 28


class RefConfig(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("threads", C.c_int), ("shader", C.c_int),
                ("scene", C.c_int), ("spp", C.c_int), ("spl", C.c_int), ("accelerator", C.c_int),
                ("obj", C.c_char_p), ("mtl", C.c_char_p), ("cam", C.c_char_p), ("samplers", C.c_int),
                ("copies_x", C.c_int), ("copies_z", C.c_int)]


def available():
    return os.path.exists(REF_SO)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if "MRT_ORACLE_SEED" not in os.environ:
            os.environ["MRT_ORACLE_SEED"] = "1"
        L = C.CDLL(REF_SO)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.POINTER(RefConfig)]
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_render.restype = C.c_double
        L.ref_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.ref_raytrace.restype = C.c_double
        L.ref_raytrace.argtypes = [C.POINTER(RefConfig), C.c_void_p, C.c_int, C.c_int]
        L.ref_pixel_sampler_draws.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.ref_pixel_sampler_reset.argtypes = [C.c_void_p]
        L.ref_random_device_counter.restype = C.c_uint
        L.ref_random_device_counter.argtypes = [C.c_int, C.c_uint]
        L.ref_casted_rays.restype = C.c_uint64
        L.ref_casted_rays.argtypes = [C.c_void_p]
        for name in ("ref_times", "ref_counts", "ref_triangles", "ref_spheres", "ref_planes", "ref_materials",
                     "ref_lights", "ref_camera", "ref_depth_max_point"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_void_p]
        L.ref_light_triangle.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.ref_generate_rays.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_trace_rays.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 10
        L.ref_shade_rays.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_grid_cells.restype = C.c_longlong
        L.ref_grid_cells.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong]
        L.ref_triangle_intersect.restype = C.c_float
        L.ref_triangle_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p]
        L.ref_triangle_build.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_triangle_aabb.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_sphere_intersect.restype = C.c_float
        L.ref_sphere_intersect.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        L.ref_plane_intersect.restype = C.c_float
        L.ref_plane_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float]
        L.ref_plane_aabb.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_aabb_intersect.restype = C.c_int
        L.ref_aabb_intersect.argtypes = [C.c_void_p] * 4
        L.ref_aabb_area.restype = C.c_float
        L.ref_aabb_area.argtypes = [C.c_void_p] * 2
        L.ref_halton.restype = C.c_float
        L.ref_halton.argtypes = [C.c_uint, C.c_uint]
        L.ref_incremental_avg.restype = C.c_int32
        L.ref_incremental_avg.argtypes = [C.c_void_p, C.c_int32, C.c_int]
        L.ref_fresnel.restype = C.c_float
        L.ref_fresnel.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
        for t in ("tile", "shader", "static_halton", "static_mt", "static_pcg"):
            f = getattr(L, "mrt_ref_table_" + t)
            f.restype = C.POINTER(C.c_float)
            f.argtypes = [C.POINTER(C.c_uint)]
        L.mrt_ref_light_index_counter.restype = C.c_uint
        L.mrt_ref_light_index_counter.argtypes = []
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def f3(x):
    return np.ascontiguousarray(np.asarray(x, dtype=np.float32).reshape(-1))


def table(which):
    """Current content of one of the reference's shuffled tables:
    tile (Renderer.cpp:12), shader (Shader.cpp:20), static_halton, static_mt, static_pcg."""
    n = C.c_uint(0)
    ptr = getattr(lib(), "mrt_ref_table_" + which)(C.byref(n))
    return np.ctypeslib.as_array(ptr, shape=(n.value,)).copy()


def light_index_counter():
    """Value of the function-static counter inside Shader::getLightIndex (Shader.cpp:224): process-global, never reset."""
    return int(lib().mrt_ref_light_index_counter())


def random_device_counter(value=None):
    """Draws the seeded std::random_device stand-in has handed out; with a value, rewinds / sets it (returns the old count)."""
    return int(lib().ref_random_device_counter(0 if value is None else 1, 0 if value is None else int(value)))


class RefRenderer:
    """Scene -> Shader -> Renderer assembled like C_wrapper.cpp's work_thread and kept alive."""

    def __init__(self, width, height, shader, scene, spp=1, spl=1, accelerator=3, threads=1,
                 obj=None, mtl=None, cam=None, samplers=0, copies=(1, 1)):
        """samplers: 0 = the Linux front-end's (StaticHaltonSeq), 1 = Android's (StaticPCG) for pixel jitter, OBJ lights, RR;
        2..6 = pixel jitter from StaticMersenneTwister / HaltonSeq / Stratified / MersenneTwister / PCG (lights, RR as 0).
        copies: OBJ scenes only, the loaded triangles copied on a (nx, nz) grid (config 5's synthetic scene)."""
        self.cfg = RefConfig(width, height, threads, shader, scene, spp, spl, accelerator,
                             obj.encode() if obj else None, mtl.encode() if mtl else None,
                             cam.encode() if cam else None, samplers, int(copies[0]), int(copies[1]))
        self.width, self.height, self.spp, self.spl = width, height, spp, spl
        self.h = lib().ref_create(C.byref(self.cfg))
        if not self.h:
            raise RuntimeError("reference failed to assemble the scene")
        cnt = np.zeros(5, np.int32)
        lib().ref_counts(self.h, _p(cnt))
        self.n_tri, self.n_sph, self.n_pla, self.n_mat, self.n_lights = (int(v) for v in cnt)

    def close(self):
        if self.h:
            lib().ref_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def render(self, threads=1, repeats=1, bitmap=None):
        if bitmap is None:
            bitmap = np.zeros(self.width * self.height, np.int32)
        # For sizes that are not multiples of 16 the reference writes past the end of the bitmap (Renderer.cpp:149, the
        # pixel index is not clamped): render into a padded copy so that those writes land in our own slack.
        n = self.width * self.height
        work = np.zeros(n + 2 * self.width * (self.height // 16 + 2) + 64, np.int32)
        work[:n] = bitmap.reshape(-1)
        secs = lib().ref_render(self.h, _p(work), threads, repeats)
        if secs < 0:
            raise RuntimeError("the reference's renderFrame threw")
        bitmap.reshape(-1)[:] = work[:n]
        return bitmap.reshape(self.height, self.width), secs

    def casted_rays(self):
        return int(lib().ref_casted_rays(self.h))

    def pixel_sampler_draws(self, n):
        """The next n draws of the renderer's own pixel sampler object (advances it, like rendering would)."""
        out = np.zeros(n, np.float32)
        lib().ref_pixel_sampler_draws(self.h, n, _p(out))
        return out

    def pixel_sampler_reset(self):
        lib().ref_pixel_sampler_reset(self.h)

    def times(self):
        t = np.zeros(3, np.float64)
        lib().ref_times(self.h, _p(t))
        return {"load": t[0], "fill": t[1], "create": t[2]}

    def triangles(self):
        a = np.zeros(self.n_tri, TRI_DTYPE)
        lib().ref_triangles(self.h, _p(a))
        return a

    def spheres(self):
        a = np.zeros(self.n_sph, SPH_DTYPE)
        lib().ref_spheres(self.h, _p(a))
        return a

    def planes(self):
        a = np.zeros(self.n_pla, PLA_DTYPE)
        lib().ref_planes(self.h, _p(a))
        return a

    def materials(self):
        a = np.zeros((self.n_mat, 13), np.float32)
        lib().ref_materials(self.h, _p(a))
        return a

    def lights(self):
        a = np.zeros((self.n_lights, 16), np.float32)
        lib().ref_lights(self.h, _p(a))
        return a

    def light_triangles(self):
        a = np.zeros(self.n_lights, TRI_DTYPE)
        for i in range(self.n_lights):
            lib().ref_light_triangle(self.h, i, C.c_void_p(a.ctypes.data + 100 * i))
        return a

    def camera(self):
        a = np.zeros(15, np.float32)
        lib().ref_camera(self.h, _p(a))
        return {"kind": int(a[0]), "position": a[1:4].copy(), "direction": a[4:7].copy(), "right": a[7:10].copy(),
                "up": a[10:13].copy(), "a": a[13], "b": a[14]}

    def depth_max_point(self):
        a = np.zeros(3, np.float32)
        lib().ref_depth_max_point(self.h, _p(a))
        return a

    def generate_rays(self, uvdd):
        uvdd = np.ascontiguousarray(uvdd, np.float32).reshape(-1, 4)
        n = len(uvdd)
        o = np.zeros((n, 3), np.float32)
        d = np.zeros((n, 3), np.float32)
        lib().ref_generate_rays(self.h, n, _p(uvdd), _p(o), _p(d))
        return o, d

    def trace(self, o, d, origin_kind=None, origin_idx=None, shadow_dist=None):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        n = len(o)
        ok = np.ascontiguousarray(origin_kind, np.int32) if origin_kind is not None else None
        oi = np.ascontiguousarray(origin_idx, np.int32) if origin_idx is not None else None
        sd = np.ascontiguousarray(shadow_dist, np.float32) if shadow_dist is not None else None
        t = np.zeros(n, np.float32)
        kind = np.zeros(n, np.int32)
        idx = np.zeros(n, np.int32)
        mat = np.zeros(n, np.int32)
        nrm = np.zeros((n, 3), np.float32)
        lib().ref_trace_rays(self.h, n, _p(o), _p(d), _p(ok), _p(oi), _p(sd), _p(t), _p(kind), _p(idx), _p(mat), _p(nrm))
        return {"t": t, "kind": kind, "idx": idx, "mat": mat, "normal": nrm}

    def grid_cells(self, kind):
        """Cell lists of the Shader's RegularGrid<T> (handle created with accelerator=2): kind 0 planes, 1 spheres, 2 triangles.
        Returns (params[12] = bounds min, max, cellSizeInverted_, cellSize_; start[32^3 + 1]; items sorted inside each cell)."""
        params = np.zeros(12, np.float32)
        start = np.zeros(32 ** 3 + 1, np.uint32)
        n = lib().ref_grid_cells(self.h, kind, _p(params), _p(start), None, 0)
        if n < 0:
            raise RuntimeError("the reference handle was not created with the RegularGrid accelerator")
        items = np.zeros(max(int(n), 1), np.int32)
        lib().ref_grid_cells(self.h, kind, _p(params), _p(start), _p(items), int(n))
        return params, start, items[:int(n)]

    def shade(self, o, d, depth=1):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        rgb = np.zeros((len(o), 3), np.float32)
        lib().ref_shade_rays(self.h, len(o), _p(o), _p(d), depth, _p(rgb))
        return rgb


def raytrace(width, height, shader, scene, spp=1, spl=1, accelerator=3, threads=1, repeats=1,
             obj=None, mtl=None, cam=None):
    """The reference's own public entry RayTrace(config, false) (C_wrapper.h:15)."""
    cfg = RefConfig(width, height, threads, shader, scene, spp, spl, accelerator,
                    obj.encode() if obj else None, mtl.encode() if mtl else None, cam.encode() if cam else None)
    bitmap = np.zeros(width * height, np.int32)
    secs = lib().ref_raytrace(C.byref(cfg), _p(bitmap), repeats, 0)
    return bitmap.reshape(height, width), secs


# ---- primitive-level helpers ----
def triangle_build(A, B, Cc, mat=0):
    out = np.zeros(1, TRI_DTYPE)
    lib().ref_triangle_build(_p(f3(A)), _p(f3(B)), _p(f3(Cc)), mat, _p(out))
    return out


def triangle_intersect(tri, o, d, t_max=1e30, self_skip=False):
    out = np.zeros(8, np.float32)
    t = lib().ref_triangle_intersect(_p(tri), _p(f3(o)), _p(f3(d)), t_max, int(self_skip), _p(out))
    return float(t), out


def triangle_aabb(tri):
    out = np.zeros(6, np.float32)
    lib().ref_triangle_aabb(_p(tri), _p(out))
    return out


def sphere_intersect(center, radius, o, d, t_max=1e30):
    n = np.zeros(3, np.float32)
    t = lib().ref_sphere_intersect(_p(f3(center)), radius, _p(f3(o)), _p(f3(d)), t_max, _p(n))
    return float(t), n


def plane_intersect(point, normal, o, d, t_max=1e30):
    return float(lib().ref_plane_intersect(_p(f3(point)), _p(f3(normal)), _p(f3(o)), _p(f3(d)), t_max))


def plane_aabb(point, normal):
    out = np.zeros(6, np.float32)
    lib().ref_plane_aabb(_p(f3(point)), _p(f3(normal)), _p(out))
    return out


def aabb_intersect(mn, mx, o, d):
    return bool(lib().ref_aabb_intersect(_p(f3(mn)), _p(f3(mx)), _p(f3(o)), _p(f3(d))))


def aabb_area(mn, mx):
    return float(lib().ref_aabb_area(_p(f3(mn)), _p(f3(mx))))


def halton(index, base=2):
    return float(lib().ref_halton(index, base))


def incremental_avg(rgb, avg, n):
    return int(lib().ref_incremental_avg(_p(f3(rgb)), int(avg), int(n)))
