"""TEST INFRASTRUCTURE (oracle) — picks the strongest checker available: the compiled reference (oracle/_ref) when its
.so is present, else the C++ restatement (oracle/libmrt_oracle.so). Only tests/, smoke() and bench.py's CPU legs use it."""
import numpy as np
from . import refbind


def unpack(bitmap):
    b = np.asarray(bitmap).astype(np.uint32).reshape(-1)
    return np.stack([b & 0xFF, (b >> 8) & 0xFF, (b >> 16) & 0xFF], axis=-1).astype(np.uint8)


def render(width, height, shader, scene, spp=1, spl=1, threads=1, **kw):
    if refbind.available():
        r = refbind.RefRenderer(width, height, shader, scene, spp=spp, spl=spl, **kw)
        bm, _ = r.render(threads=threads)
        r.close()
        return bm
    from . import portbind
    return portbind.render(width, height, shader, scene, spp=spp, spl=spl, **kw)
