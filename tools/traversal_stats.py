"""Diagnostic: average BVH8 node visits and triangle tests per ray of the wavefront kernels on the bench workload, from
the counting build (make -C mobileraytracer_b200/csrc stats). Run with MRT_B200_LIB=mobileraytracer_b200/libmobilert_b200_stats.so."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import mobileraytracer_b200 as mrt  # noqa: E402
from mobileraytracer_b200 import capi, loaders  # noqa: E402

h = bench.round_down16(bench.HEIGHT_ASKED)
obj, mtl, cam = bench.scene_files()
scene = loaders.load_obj(obj, mtl).arrays()
camera = loaders.load_camera(cam, np.float32(bench.WIDTH) / np.float32(h))
shader = int(sys.argv[1]) if len(sys.argv) > 1 else capi.SHADER_PATHTRACER
r = mrt.Renderer(scene, camera, bench.WIDTH, h, 1, 1, shader)
L = capi.lib()
out = (C.c_ulonglong * 5)()
L.mrt_debug_traversal_stats(out, 1)
r.renderFrame(np.zeros(bench.WIDTH * h, np.int32))
L.mrt_debug_traversal_stats(out, 1)
st = r.stats()
print("rays %d  nodes/ray %.2f  tris/ray %.2f   (extend %d, shadow %d rays; bvh nodes %d)" % (
    out[2], out[0] / out[2], out[1] / out[2], st["rays_extend_traced"], st["rays_shadow_traced"], st["bvh_nodes"]))
print("node visits that hit no child: %.1f %%; of all visits, %.1f %% hit none only because of the tbest bound (stale stack entries)" % (
    100.0 * out[3] / max(out[0], 1), 100.0 * out[4] / max(out[0], 1)))
