"""Diagnostic: noise calibration of the path-tracer comparison (reference vs reference with other tables, GPU vs reference)
and the normal mismatches on the conference stand-in."""
import os, sys
import numpy as np
sys.path.insert(0, ".")
from oracle import refbind as ref
from tests import helpers as H
import bench

w = h = 96; spp = 256
blk = lambda x: x.reshape(h // 4, 4, w // 4, 4, 3).mean(axis=(1, 3))
r = ref.RefRenderer(w, h, 2, 2, spp=spp); a, _ = r.render(threads=16); a = H.unpack(a).astype(float)
g = H.gpu_from_ref(ref, r, 2, 2, w, h, spp=spp); b = H.unpack(g.renderFrame()).astype(float); g.close()
r2 = ref.RefRenderer(w, h, 2, 2, spp=spp); c, _ = r2.render(threads=16); c = H.unpack(c).astype(float)
for name, x, y in (("ref-ref", a, c), ("gpu-ref", b, a), ("gpu-ref2", b, c)):
    d = np.abs(blk(x) - blk(y))
    print(name, "mean", x.mean(), y.mean(), "blockdiff mean %.2f p99 %.2f max %.2f" % (d.mean(), np.percentile(d, 99), d.max()), "signed", (blk(x) - blk(y)).mean(axis=(0, 1)))
sd = (blk(b) - blk(a)).mean(axis=2)
print("rows mean signed diff:", np.round(sd.mean(axis=1), 1))
obj, mtl, cam = bench.scene_files()
w2, h2 = 960, 528
r = ref.RefRenderer(w2, h2, 3, -1, obj=obj, mtl=mtl, cam=cam)
g = H.gpu_from_ref(ref, r, -1, 3, w2, h2, cam_file=cam)
o, d = r.generate_rays(H.primary_uvdd(w2, h2))
A, B = r.trace(o, d), g.traceRays(o, d)
same = (A["kind"] == B["kind"]) & (A["idx"] == B["idx"])
bad = same & ~(A["normal"] == B["normal"]).all(axis=1)
print("normal mismatches", int(bad.sum()), "nan ref", int(np.isnan(A["normal"][bad]).any(axis=1).sum()), "nan gpu", int(np.isnan(B["normal"][bad]).any(axis=1).sum()))
i = np.nonzero(bad)[0][:5]
print(A["normal"][i], B["normal"][i], A["kind"][i], A["idx"][i])
