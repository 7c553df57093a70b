"""Shared helpers of the parity tests: build a GPU renderer from the arrays the compiled reference exports."""
import numpy as np

from mobileraytracer_b200 import capi, scene as S
from mobileraytracer_b200.renderer import Renderer


def unpack(bitmap):
    b = np.asarray(bitmap).astype(np.uint32).reshape(-1)
    return np.stack([b & 0xFF, (b >> 8) & 0xFF, (b >> 16) & 0xFF], axis=-1).astype(np.int32)


def scene_arrays_from_ref(r):
    """Post-build primitive arrays of a RefRenderer -> the C-ABI's scene arrays (same byte layouts)."""
    tri = r.triangles().view(capi.TRIANGLE) if r.n_tri else np.zeros(0, capi.TRIANGLE)
    sph = r.spheres().view(capi.SPHERE) if r.n_sph else np.zeros(0, capi.SPHERE)
    pla = r.planes().view(capi.PLANE) if r.n_pla else np.zeros(0, capi.PLANE)
    mats = np.zeros(r.n_mat, capi.MATERIAL)
    m = r.materials()
    for i in range(r.n_mat):
        mats[i]["Le"], mats[i]["Kd"], mats[i]["Ks"], mats[i]["Kt"], mats[i]["ior"] = m[i, 0:3], m[i, 3:6], m[i, 6:9], m[i, 9:12], m[i, 12]
    lights = np.zeros(r.n_lights, capi.LIGHT)
    L = r.lights()
    LT = r.light_triangles()
    for i in range(r.n_lights):
        lights[i]["kind"] = int(L[i, 0])
        lights[i]["position"] = L[i, 1:4]
        lights[i]["triangle"] = LT[i:i + 1].view(capi.TRIANGLE)[0]
        lights[i]["Le"] = L[i, 10:13]
    return {"triangles": tri, "spheres": sph, "planes": pla, "materials": mats, "lights": lights}


def camera_from_config(scene_index, ratio, cam_file=None):
    if scene_index in S.BUILTIN:
        return S.BUILTIN[scene_index][1](np.float32(ratio))
    from mobileraytracer_b200 import loaders
    return loaders.load_camera(cam_file, np.float32(ratio))


def gpu_from_ref(refbind, r, scene_index, shader, width, height, spp=1, spl=1, flags=0, cam_file=None, seed=1, samplers=0, accelerator=None):
    """GPU renderer over exactly the reference's scene, camera and (current) sampler tables.
    samplers=1: the Android front-end's StaticPCG for pixel jitter, OBJ lights and russian roulette (RefRenderer(samplers=1))."""
    from mobileraytracer_b200 import capi
    arrays = scene_arrays_from_ref(r)
    tables = {"tile_order": refbind.table("tile"), "shader_table": refbind.table("shader"),
              "static_halton": refbind.table("static_halton")}
    kw = {} if accelerator is None else {"accelerator": accelerator}
    if samplers == 1:
        tables["static_pcg"] = refbind.table("static_pcg")
        kw.update(light_sampler=capi.SAMPLER_STATIC_PCG,
                  pixel_sampler=capi.SAMPLER_STATIC_PCG if spp > 1 else capi.SAMPLER_CONSTANT)
    cam = camera_from_config(scene_index, np.float32(width) / np.float32(height), cam_file)
    max_point = S.BUILTIN[scene_index][2] if scene_index in S.BUILTIN else (1, 1, 1)
    return Renderer(arrays, cam, width, height, spp, spl, shader, depth_max_point=max_point, tables=tables, flags=flags, seed=seed, **kw)


def primary_uvdd(width, height, step=1):
    ys, xs = np.mgrid[0:height:step, 0:width:step]
    inv_w, inv_h = np.float32(1.0) / np.float32(width), np.float32(1.0) / np.float32(height)
    u = xs.astype(np.float32) * inv_w
    v = ys.astype(np.float32) * inv_h
    z = np.zeros_like(u)
    return np.stack([u, v, z, z], axis=-1).reshape(-1, 4).astype(np.float32)


def compare_bitmaps(gpu, ref):
    d = np.abs(unpack(gpu) - unpack(ref))
    return {"exact": float((d == 0).all(axis=-1).mean()), "within1": float((d <= 1).mean()), "max": int(d.max()),
            "rmse": float(np.sqrt((d.astype(np.float64) ** 2).mean()))}


def resolve_reference(accum):
    """numpy statement of mrt_resolve_device (mrt_render.cu k_resolve) for checking the multi-GPU reduce + resolve."""
    a = np.asarray(accum, np.float32).reshape(-1, 4)
    out = np.zeros(len(a), np.uint32)
    ok = a[:, 3] > 0
    q = np.zeros((len(a), 3), np.uint32)
    q[ok] = np.minimum((a[ok, :3] / a[ok, 3:4]).astype(np.float32).astype(np.uint32), 255)
    out[ok] = 0xFF000000 | (q[ok, 2] << 16) | (q[ok, 1] << 8) | q[ok, 0]
    return out.view(np.int32)
