"""BASELINE configs 1-3 through the RayTrace(Config) mirror on the GPU, next to the compiled reference (oracle/_ref) on the
host cores: ms/frame and Mrays/s (device events for the GPU frame; the reference's own "Rendering Time" for the CPU).
Parity for these configs is asserted in tests/; this tool only measures. Not the bench line (that is config 4)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import mobileraytracer_b200 as mrt  # noqa: E402
from mobileraytracer_b200 import capi  # noqa: E402


def gpu_case(name, cfg, frames=5):
    r = mrt.build_renderer(cfg)
    bm = np.zeros(cfg.width * cfg.height, np.int32)
    best = None
    for _ in range(frames):
        r.renderFrame(bm)
        st = r.stats()
        if best is None or st["frame_ms"] < best["frame_ms"]:
            best = st
    rays = best["rays_primary"] + best["rays_shadow"] + best["rays_secondary"]
    t0 = time.perf_counter()
    r.renderFrame(bm)                       # host bitmap in, host bitmap out: the e2e figure
    e2e_ms = (time.perf_counter() - t0) * 1e3
    out = {"config": name, "width": cfg.width, "height": cfg.height, "spp": cfg.samplesPixel, "rays": int(rays),
           "gpu_frame_ms": round(best["frame_ms"], 4), "gpu_mrays_s": round(rays / best["frame_ms"] / 1e3, 1),
           "gpu_e2e_ms_host_bitmap": round(e2e_ms, 3), "build_ms": round(best["build_ms"], 3)}
    r.close()
    return out


def cpu_case(out, shader, scene=None, spp=1, obj=None):
    try:
        from oracle import refbind
        threads = os.cpu_count() or 1
        kw = dict(obj=obj[0], mtl=obj[1], cam=obj[2]) if obj else {}
        r = refbind.RefRenderer(out["width"], out["height"], shader, -1 if obj else scene, spp=spp, spl=1, **kw)
        r.render(threads=threads)                      # warm-up frame
        before = r.casted_rays()
        _, secs = r.render(threads=threads)
        rays = r.casted_rays() - before
        r.close()
        out.update({"cpu_threads": threads, "cpu_frame_ms": round(secs * 1e3, 2), "cpu_mrays_s": round(rays / secs / 1e6, 2),
                    "speedup_frame": round(secs * 1e3 / out["gpu_frame_ms"], 1)})
    except Exception as ex:   # the reference library is optional infrastructure here
        out["cpu"] = "unavailable: %s" % ex
    return out


if __name__ == "__main__":
    rows = []
    c1 = mrt.Config(width=512, height=512, shader=capi.SHADER_WHITTED, sceneIndex=0, samplesPixel=1, samplesLight=1, printStdOut=False)
    rows.append(cpu_case(gpu_case("C1 Cornell, Whitted, 1 spp, 512x512", c1), 1, scene=0, spp=1))
    c2 = mrt.Config(width=1024, height=1024, shader=capi.SHADER_WHITTED, sceneIndex=3, samplesPixel=4, samplesLight=1, printStdOut=False)
    rows.append(cpu_case(gpu_case("C2 Spheres2, Whitted, 4 spp, 1024x1024", c2), 1, scene=3, spp=4))
    obj, mtl, cam = bench.scene_files()
    h = bench.round_down16(bench.HEIGHT_ASKED)
    for shader, nm in ((capi.SHADER_NOSHADOWS, "NoShadows"), (capi.SHADER_DEPTHMAP, "DepthMap")):
        c3 = mrt.Config(width=bench.WIDTH, height=h, shader=shader, sceneIndex=-1, samplesPixel=1, samplesLight=1, printStdOut=False,
                        objFilePath=obj, mtlFilePath=mtl, camFilePath=cam)
        rows.append(cpu_case(gpu_case("C3 conference stand-in, %s, 1 spp, 1920x1072" % nm, c3), shader, spp=1, obj=(obj, mtl, cam)))
    for r in rows:
        print(json.dumps(r))
