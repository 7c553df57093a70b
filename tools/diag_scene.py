"""Diagnostic: prints per-ray disagreements between the CUDA path and the compiled reference for a built-in scene."""
import sys
import numpy as np
sys.path.insert(0, ".")
from oracle import refbind as ref
from tests import helpers as H

scene = int(sys.argv[1]) if len(sys.argv) > 1 else 1
w = h = 256
r = ref.RefRenderer(w, h, 3, scene)
g = H.gpu_from_ref(ref, r, scene, 3, w, h)
uvdd = H.primary_uvdd(w, h)
o, d = r.generate_rays(uvdd)
a = r.trace(o, d)
b = g.traceRays(o, d)
bad = np.nonzero(~((a["kind"] == b["kind"]) & (a["idx"] == b["idx"])))[0]
print("n_bad", len(bad), "of", len(o), "prims", r.n_tri, r.n_sph, r.n_pla)
for i in bad[:25]:
    print(i, divmod(int(i), w), "o", o[i], "d", d[i], "ref", a["kind"][i], a["idx"][i], a["t"][i], "gpu", b["kind"][i], b["idx"][i], b["t"][i])
import collections
print(collections.Counter(zip(a["kind"][bad].tolist(), a["idx"][bad].tolist(), b["kind"][bad].tolist(), b["idx"][bad].tolist())))
