"""BASELINE config 5 on one GPU: the conference stand-in replicated 30 times on a 6 x 5 grid (flattened, ~10 M triangles;
the reference has no instancing), GPU BVH build time, then a PathTracer frame at 3840x2160.
Reports build time, node count and Mrays/s. Not a benchmark line; a scale check of the builder and the queues."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import mobileraytracer_b200 as mrt  # noqa: E402
from mobileraytracer_b200 import capi, loaders  # noqa: E402


replicate = loaders.replicate_scene


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=2)
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--copies", type=int, default=30)
    a = ap.parse_args()
    obj, mtl, cam = bench.scene_files()
    base = loaders.load_obj(obj, mtl).arrays()
    nx = 6 if a.copies >= 30 else max(1, a.copies)
    nz = 5 if a.copies >= 30 else 1
    scene = replicate(base, nx, nz)
    w, h = bench.round_down16(a.width), bench.round_down16(a.height)
    camera = loaders.load_camera(cam, np.float32(w) / np.float32(h))
    t0 = time.time()
    r = mrt.Renderer(scene, camera, w, h, a.spp, 1, capi.SHADER_PATHTRACER)
    t1 = time.time()
    bm = np.zeros(w * h, np.int32)
    r.renderFrame(bm)
    st = r.stats()
    rays = st["rays_primary"] + st["rays_shadow"] + st["rays_secondary"]
    print({"triangles": int(len(scene["triangles"])), "lights": int(len(scene["lights"])), "width": w, "height": h, "spp": a.spp,
           "create_wall_s": round(t1 - t0, 3), "build_ms": st["build_ms"], "upload_ms": st["upload_ms"], "bvh_nodes": st["bvh_nodes"],
           "bvh_bytes": st["bvh_bytes"], "frame_ms": st["frame_ms"], "extend_ms": st["extend_ms"], "shadow_ms": st["shadow_ms"],
           "shade_ms": st["shade_ms"], "other_ms": st["other_ms"], "Mrays_s": rays / st["frame_ms"] / 1e3,
           "nonblack": float((bm != -16777216).mean())})
    r.close()
