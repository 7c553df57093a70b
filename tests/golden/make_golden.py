"""Generates tests/golden/reference_frames.npz by running the UNMODIFIED reference (oracle/_ref, compiled from
/root/reference) in this container. The fixtures are the table-independent cases (1 spp, Constant(0.5) pixel sampler,
at most one light, so no shuffled-table draw changes a pixel): they stay valid on machines without /root/reference.
  frame_<scene>_<shader> : 64x64 ABGR8888 bitmap     rays_<scene>_<shader> : Ray objects constructed ("Casted rays")
  hit_<scene>_{t,kind,idx}: per-primary-ray ground truth of BVH<T>::trace over the three typed BVHs + lights
Run: python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import refbind as R  # noqa: E402
from tests import helpers as H  # noqa: E402

W = 64
out = {}
for scene in (0, 1, 2, 3):
    for shader in ((0, 1, 3, 4) if scene != 2 else (3, 4)):
        r = R.RefRenderer(W, W, shader, scene)
        bm, _ = r.render(threads=1)
        out["frame_%d_%d" % (scene, shader)] = bm.astype(np.int32)
        out["rays_%d_%d" % (scene, shader)] = np.int64(r.casted_rays())
        if shader == 3:
            o, d = r.generate_rays(H.primary_uvdd(W, W))
            tr = r.trace(o, d)
            out["hit_%d_t" % scene], out["hit_%d_kind" % scene], out["hit_%d_idx" % scene] = tr["t"], tr["kind"], tr["idx"]
            out["ray_%d_o" % scene], out["ray_%d_d" % scene] = o, d
        r.close()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_frames.npz"), **out)
print("wrote", len(out), "arrays")
