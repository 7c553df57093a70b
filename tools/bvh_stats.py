"""Diagnostic (stats library only): how full the BVH8 nodes of the bench scene are."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["MRT_B200_LIB"] = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mobileraytracer_b200", "libmobilert_b200_stats.so")
import bench  # noqa: E402
import mobileraytracer_b200 as mrt  # noqa: E402
from mobileraytracer_b200 import capi  # noqa: E402

obj, mtl, cam = bench.scene_files()
cfg = mrt.Config(width=256, height=144, shader=capi.SHADER_DEPTHMAP, sceneIndex=-1, samplesPixel=1, printStdOut=False,
                 objFilePath=obj, mtlFilePath=mtl, camFilePath=cam)
r = mrt.build_renderer(cfg)
out = (C.c_ulonglong * 18)()
capi.lib().mrt_debug_bvh_stats(r._h, out)
v = list(out)
nodes = sum(v[0:9])
print("nodes", nodes, "children/node", sum(k * v[k] for k in range(9)) / max(nodes, 1))
print("nodes by valid children 0..8:", v[0:9])
print("leaf slots by triangle count 0..7:", v[9:17], "inner children", v[17])
tris = sum(k * v[9 + k] for k in range(8))
print("triangles in leaves", tris, "per leaf slot", tris / max(sum(v[10:17]), 1))
