"""TEST INFRASTRUCTURE (oracle) — ctypes binding of oracle/libmrt_oracle.so, the plain C++ restatement of the reference's
hot path (oracle/mrt_oracle.cpp). Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(_HERE, "libmrt_oracle.so")

TRIANGLE = np.dtype([("AC", "<f4", 3), ("AB", "<f4", 3), ("A", "<f4", 3), ("nA", "<f4", 3), ("nB", "<f4", 3), ("nC", "<f4", 3),
                     ("tA", "<f4", 2), ("tB", "<f4", 2), ("tC", "<f4", 2), ("material", "<i4")])
SPHERE = np.dtype([("center", "<f4", 3), ("sq_radius", "<f4"), ("material", "<i4")])
PLANE = np.dtype([("normal", "<f4", 3), ("point", "<f4", 3), ("material", "<i4")])
MATERIAL = np.dtype([("Le", "<f4", 3), ("Kd", "<f4", 3), ("Ks", "<f4", 3), ("Kt", "<f4", 3), ("ior", "<f4")])
LIGHT = np.dtype([("kind", "<i4"), ("position", "<f4", 3), ("triangle", TRIANGLE), ("Le", "<f4", 3)])


class PortScene(C.Structure):
    _fields_ = [("tris", C.c_void_p), ("nTris", C.c_int64), ("spheres", C.c_void_p), ("nSpheres", C.c_int64), ("planes", C.c_void_p),
                ("nPlanes", C.c_int64), ("mats", C.c_void_p), ("nMats", C.c_int64), ("lights", C.c_void_p), ("nLights", C.c_int64)]


class PortCamera(C.Structure):
    _fields_ = [("kind", C.c_int32), ("position", C.c_float * 3), ("lookAt", C.c_float * 3), ("up", C.c_float * 3), ("a", C.c_float), ("b", C.c_float)]


class PortRender(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("spl", C.c_int32), ("shader", C.c_int32),
                ("depthMaxPoint", C.c_float * 3), ("tileOrder", C.c_void_p), ("shaderTable", C.c_void_p), ("staticHalton", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(PORT_SO):
            subprocess.check_call(["make", "-C", _HERE, "port"])
        L = C.CDLL(PORT_SO)
        vp = C.c_void_p
        L.port_create.restype = vp
        L.port_create.argtypes = [C.POINTER(PortScene), C.POINTER(PortCamera), C.POINTER(PortRender)]
        L.port_destroy.argtypes = [vp]
        L.port_render.argtypes = [vp, vp]
        L.port_casted_rays.restype = C.c_uint64
        L.port_casted_rays.argtypes = [vp]
        L.port_set_counters.argtypes = [vp, C.c_uint32, C.c_uint32]
        L.port_trace.argtypes = [vp, C.c_int64, vp, vp, vp, vp, vp]
        L.port_aabb_intersect.argtypes = [vp] * 4
        L.port_aabb_area.restype = C.c_float
        L.port_aabb_area.argtypes = [vp, vp]
        L.port_triangle_intersect.restype = C.c_float
        L.port_triangle_intersect.argtypes = [vp, vp, vp, C.c_float, C.c_int]
        L.port_triangle_aabb.argtypes = [vp, vp]
        L.port_sphere_intersect.restype = C.c_float
        L.port_sphere_intersect.argtypes = [vp, vp, vp, C.c_float]
        L.port_plane_intersect.restype = C.c_float
        L.port_plane_intersect.argtypes = [vp, vp, vp, C.c_float]
        L.port_plane_aabb.argtypes = [vp, vp]
        L.port_incremental_avg.restype = C.c_int32
        L.port_incremental_avg.argtypes = [vp, C.c_int32, C.c_int32]
        L.port_halton.restype = C.c_float
        L.port_halton.argtypes = [C.c_uint32, C.c_uint32]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and len(a) else None


def halton_base2(n):
    """fillArrayWithHaltonSeq before its shuffle (Utils.hpp:209-213): radical inverse base 2 of 0..n-1."""
    i = np.arange(n, dtype=np.uint32)
    bits = int(np.log2(n))
    rev = np.zeros(n, dtype=np.uint32)
    for b in range(bits):
        rev |= ((i >> b) & 1) << (bits - 1 - b)
    return (rev.astype(np.float64) / n).astype(np.float32)


def default_tables(seed=1):
    """Shuffled tables with numpy's generator (the reference's std::shuffle order is libstdc++ specific; parity runs take the
    compiled reference's own tables instead)."""
    rng = np.random.default_rng(seed)
    return {"tile_order": rng.permutation(halton_base2(256)), "shader_table": rng.permutation(halton_base2(1 << 20)),
            "static_halton": rng.permutation(halton_base2(1 << 20))}


class PortRenderer:
    def __init__(self, arrays, camera, width, height, shader, spp=1, spl=1, depth_max_point=(1, 1, 1), tables=None):
        """arrays: dict of numpy structured arrays (triangles, spheres, planes, materials, lights) in the C-ABI layouts;
        camera: dict(kind, position, look_at, up, a, b)."""
        self._keep = {k: np.ascontiguousarray(v) for k, v in arrays.items()}
        self.tables = {k: np.ascontiguousarray(v, np.float32) for k, v in (tables or default_tables()).items()}
        s = PortScene()
        for f, k in (("tris", "triangles"), ("spheres", "spheres"), ("planes", "planes"), ("mats", "materials"), ("lights", "lights")):
            a = self._keep[k]
            setattr(s, f, a.ctypes.data if len(a) else None)
            setattr(s, "n" + f[0].upper() + f[1:], len(a))
        c = PortCamera()
        c.kind = int(camera["kind"])
        for i in range(3):
            c.position[i], c.lookAt[i], c.up[i] = float(camera["position"][i]), float(camera["look_at"][i]), float(camera["up"][i])
        c.a, c.b = float(camera["a"]), float(camera["b"])
        r = PortRender()
        r.width, r.height, r.spp, r.spl, r.shader = width, height, spp, spl, shader
        for i in range(3):
            r.depthMaxPoint[i] = float(depth_max_point[i])
        r.tileOrder, r.shaderTable, r.staticHalton = (self.tables[k].ctypes.data for k in ("tile_order", "shader_table", "static_halton"))
        self.width, self.height = width, height
        self.h = lib().port_create(C.byref(s), C.byref(c), C.byref(r))

    def render(self, bitmap=None):
        if bitmap is None:
            bitmap = np.zeros(self.width * self.height, np.int32)
        lib().port_render(self.h, _p(bitmap))
        return bitmap.reshape(self.height, self.width)

    def casted_rays(self):
        return int(lib().port_casted_rays(self.h))

    def trace(self, o, d):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        n = len(o)
        t, kind, idx = np.zeros(n, np.float32), np.zeros(n, np.int32), np.zeros(n, np.int32)
        lib().port_trace(self.h, n, _p(o), _p(d), _p(t), _p(kind), _p(idx))
        return {"t": t, "kind": kind, "idx": idx}

    def close(self):
        if self.h:
            lib().port_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def render(width, height, shader, scene, spp=1, spl=1, tables=None, **kw):
    """Built-in scene `scene` (0..3) through the port; scene data come from the host-side input tables of the package."""
    from mobileraytracer_b200 import scene as S
    make_scene, make_cam, max_point = S.BUILTIN[scene]
    cam = make_cam(np.float32(width) / np.float32(height))
    camd = {"kind": cam.kind, "position": cam.position, "look_at": cam.look_at, "up": cam.up, "a": cam.a, "b": cam.b}
    r = PortRenderer(make_scene().arrays(), camd, width, height, shader, spp, spl, max_point, tables)
    bm = r.render()
    r.close()
    return bm
