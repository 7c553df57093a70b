"""Short profiling target: a few samples per pixel of the bench workload (conference stand-in, PathTracer, 1920x1072)
through the C-ABI, so that `ncu` sees every kernel of the wavefront a handful of times. Not a benchmark."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import mobileraytracer_b200 as mrt  # noqa: E402
from mobileraytracer_b200 import capi, loaders  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--spp", type=int, default=2)
ap.add_argument("--shader", type=int, default=capi.SHADER_PATHTRACER)
ap.add_argument("--frames", type=int, default=1)
ap.add_argument("--flags", type=int, default=0, help="MRT_FLAG_* bits, e.g. 2 = sort by material")
a = ap.parse_args()
h = bench.round_down16(bench.HEIGHT_ASKED)
obj, mtl, cam = bench.scene_files()
scene = loaders.load_obj(obj, mtl).arrays()
camera = loaders.load_camera(cam, np.float32(bench.WIDTH) / np.float32(h))
r = mrt.Renderer(scene, camera, bench.WIDTH, h, a.spp, 1, a.shader, flags=a.flags)
bm = np.zeros(bench.WIDTH * h, np.int32)
for _ in range(a.frames):
    r.renderFrame(bm)
st = r.stats()
rays = st["rays_primary"] + st["rays_shadow"] + st["rays_secondary"]
print({k: st[k] for k in ("frame_ms", "extend_ms", "shadow_ms", "shade_ms", "other_ms", "launches", "build_ms", "bvh_nodes", "rays_extend_traced", "rays_shadow_traced")},
      "Mrays/s %.1f" % (rays / st["frame_ms"] / 1e3), "nonblack %.3f" % float((bm != -16777216).mean()))
r.close()
