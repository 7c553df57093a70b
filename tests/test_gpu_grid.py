"""MRT_ACC_REGULAR_GRID on the GPU against the compiled reference built with Accelerator::ACC_REGULAR_GRID.
The reference's grid does not see what its BVH sees: a ray that starts outside a grid is walked from the clamped border
cell (RegularGrid.hpp:337-347), plane grids cover Plane::getAABB's finite flat boxes, and cell membership comes from
T::intersect(AABB). These tests therefore compare with the reference's GRID results, ray by ray and frame by frame, and
check that the grid mode is not simply the BVH mode under another name."""
import os

import numpy as np
import pytest

from tests import helpers as H

pytestmark = pytest.mark.gpu

@pytest.fixture(scope="module")
def mrt():
    import mobileraytracer_b200 as m
    if m.capi.lib().mrt_device_count() < 1:
        pytest.fail("no CUDA device: the product path has no CPU fallback")
    return m


def _rays(r, n, seed):
    """Rays from inside and from well outside the scene's bounds, in random directions."""
    rng = np.random.default_rng(seed)
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    if r.n_tri:
        t = r.triangles().view(H.capi.TRIANGLE)
        for p in (t["A"], t["A"] + t["AB"], t["A"] + t["AC"]):
            lo, hi = np.minimum(lo, p.min(axis=0)), np.maximum(hi, p.max(axis=0))
    if r.n_sph:
        s = r.spheres().view(H.capi.SPHERE)
        lo, hi = np.minimum(lo, (s["center"] - 1).min(axis=0)), np.maximum(hi, (s["center"] + 1).max(axis=0))
    if not np.isfinite(lo).all():
        lo, hi = np.full(3, -1.0), np.full(3, 1.0)
    c, e = (lo + hi) / 2, np.maximum((hi - lo) / 2, 0.5)
    scale = np.where(rng.uniform(size=(n, 1)) < 0.5, 0.95, 2.5)          # half of the origins lie outside the bounds
    o = (c + rng.uniform(-1, 1, (n, 3)) * e * scale).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    axis = rng.uniform(size=n) < 0.05                                     # some axis-parallel rays (tmax = RayLengthMax branches)
    d[axis] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, int(axis.sum()))] * rng.choice([-1.0, 1.0], (int(axis.sum()), 1)).astype(np.float32)
    return o, d, rng


def _compare_rays(r, g, o, d, rng, kind=None, idx=None):
    a, b = r.trace(o, d, kind, idx), g.traceRays(o, d, kind, idx)
    same = (a["kind"] == b["kind"]) & (a["idx"] == b["idx"])
    ties = ~same & (a["t"] == b["t"])
    assert ((~same) & ~ties).sum() == 0, (int((~same).sum()), int(ties.sum()))
    assert same.mean() >= 0.9999, same.mean()
    assert np.array_equal(a["t"], b["t"]), "t must be bit identical"
    surf = same & (a["kind"] != 0)
    assert np.array_equal(a["normal"][surf], b["normal"][surf]) and np.array_equal(a["mat"][surf], b["mat"][surf])
    dist = (rng.uniform(0.05, 3.0, len(o)) * np.maximum(np.linalg.norm(o, axis=1), 1.0)).astype(np.float32)
    sa, sb = r.trace(o, d, kind, idx, dist), g.traceRays(o, d, kind, idx, dist)
    assert np.array_equal(sa["kind"] != 0, sb["kind"] != 0), int(((sa["kind"] != 0) != (sb["kind"] != 0)).sum())
    return a


@pytest.mark.parametrize("scene", [0, 1, 2, 3])
@pytest.mark.parametrize("shader", [3, 2])
def test_grid_rays_match_the_reference_grid(mrt, ref, scene, shader):
    """Closest hits and occlusion flags, ray by ray. shader=3: per-thread walk inside the megakernel's device functions;
    shader=2: the handle is a PathTracer, so the answers come from the wavefront's grid extend / shadow kernels."""
    w = h = 64
    r = ref.RefRenderer(w, h, 3, scene, accelerator=2)
    g = H.gpu_from_ref(ref, r, scene, shader, w, h, accelerator=mrt.capi.ACC_REGULAR_GRID)
    uvdd = H.primary_uvdd(w, h)
    o, d = r.generate_rays(uvdd)
    rng = np.random.default_rng(5)
    _compare_rays(r, g, o, d, rng)
    o, d, rng = _rays(r, 60_000, 17 + scene)
    a = _compare_rays(r, g, o, d, rng)
    assert 0.02 < (a["kind"] != 0).mean() < 1.0
    g.close()
    r.close()


def test_grid_visibility_differs_from_the_bvh_and_we_follow_the_grid(mrt, ref):
    """Rays that start outside the bounds are where the two accelerators of the reference disagree; the GPU's grid mode must
    side with the reference's grid on every one of them (and its BVH mode with the reference's BVH)."""
    w = h = 64
    scene = 2
    rg = ref.RefRenderer(w, h, 3, scene, accelerator=2)
    rb = ref.RefRenderer(w, h, 3, scene, accelerator=3)
    gg = H.gpu_from_ref(ref, rg, scene, 3, w, h, accelerator=mrt.capi.ACC_REGULAR_GRID)
    gb = H.gpu_from_ref(ref, rb, scene, 3, w, h)
    o, d, _ = _rays(rg, 200_000, 3)
    ag, ab = rg.trace(o, d), rb.trace(o, d)
    differ = (ag["kind"] != ab["kind"]) | (ag["t"] != ab["t"])
    assert differ.sum() > 100, "the reference's grid and BVH should disagree on some of these rays"
    bg, bb = gg.traceRays(o, d), gb.traceRays(o, d)
    assert np.array_equal(bg["t"][differ], ag["t"][differ]) and np.array_equal(bg["kind"][differ], ag["kind"][differ])
    assert np.array_equal(bb["t"][differ], ab["t"][differ]) and np.array_equal(bb["kind"][differ], ab["kind"][differ])
    for x in (gg, gb, rg, rb):
        x.close()


@pytest.mark.parametrize("scene,shader", [(0, 1), (1, 1), (2, 3), (3, 1), (3, 0), (0, 3)])
def test_grid_frames_match_the_reference_grid(mrt, ref, scene, shader):
    w = h = 256
    r = ref.RefRenderer(w, h, shader, scene, accelerator=2)
    bm_ref, _ = r.render(threads=1)
    rays_ref = r.casted_rays()
    g = H.gpu_from_ref(ref, r, scene, shader, w, h, accelerator=mrt.capi.ACC_REGULAR_GRID)
    c = H.compare_bitmaps(g.renderFrame(), bm_ref)
    assert c["exact"] >= 0.999 and c["max"] <= 1, c
    assert g.getTotalCastedRays() == rays_ref, (g.getTotalCastedRays(), rays_ref)
    g.close()
    r.close()


@pytest.fixture(scope="module")
def small_room(tmp_path_factory):
    """The conference stand-in at a tenth of its detail: 16.7 k triangles, 36 materials (three specular), 8 area lights."""
    from tools import make_conference_standin as gen
    info = gen.write(str(tmp_path_factory.mktemp("grid_scene")), detail=0.1)
    return info["obj"], info["mtl"], info["cam"]


def test_grid_obj_scene_rays_and_frames(mrt, ref, small_room):
    """An OBJ scene through the triangle grid: incoherent rays (some leaving a triangle they must skip), a DepthMap frame,
    and a Whitted frame (shadow rays towards 8 area lights, reflections) with the reference's draw order replayed."""
    from mobileraytracer_b200 import capi
    obj, mtl, cam = small_room
    w = h = 128
    r = ref.RefRenderer(w, h, 3, -1, accelerator=2, obj=obj, mtl=mtl, cam=cam)
    g = H.gpu_from_ref(ref, r, -1, 3, w, h, cam_file=cam, accelerator=capi.ACC_REGULAR_GRID)
    o, d, rng = _rays(r, 100_000, 23)
    kind = np.where(rng.uniform(size=len(o)) < 0.1, 3, 0).astype(np.int32)
    idx = np.where(kind == 3, rng.integers(0, r.n_tri, len(o)), -1).astype(np.int32)
    a = _compare_rays(r, g, o, d, rng, kind, idx)
    assert (a["kind"] != 0).mean() > 0.4
    c = H.compare_bitmaps(g.renderFrame(), r.render(threads=1)[0])
    assert c["exact"] >= 0.999 and c["max"] <= 1, c
    g.close()
    r.close()
    r = ref.RefRenderer(w, h, 1, -1, accelerator=2, obj=obj, mtl=mtl, cam=cam)
    c0 = ref.light_index_counter()
    bm_ref, _ = r.render(threads=1)
    g = H.gpu_from_ref(ref, r, -1, 1, w, h, cam_file=cam, accelerator=capi.ACC_REGULAR_GRID, flags=capi.FLAG_REPLAY_SAMPLERS)
    g.setReplayCounter(c0)
    c = H.compare_bitmaps(g.renderFrame(), bm_ref)
    assert c["exact"] >= 0.999 and c["max"] <= 1 or c["exact"] >= 0.9999, c
    assert g.getTotalCastedRays() == r.casted_rays()
    g.close()
    r.close()


def test_grid_pathtracer_frame_is_statistically_the_reference_grid(mrt, ref):
    """PathTracer through the wavefront's grid kernels on scene 2 (area lights): mean image against the reference's grid
    frames at equal spp, bar from the spread between the reference's own frames."""
    w = h = 64
    spp, frames = 32, 6
    r = ref.RefRenderer(w, h, 2, 2, spp=spp, accelerator=2)
    refs = np.stack([H.unpack(r.render(threads=1)[0]).astype(np.float64) for _i in range(frames)])
    g = H.gpu_from_ref(ref, r, 2, 2, w, h, spp=spp, accelerator=mrt.capi.ACC_REGULAR_GRID)
    mine = H.unpack(g.renderFrame()).astype(np.float64)
    mean_ref = refs.mean(axis=0)
    spread = np.abs(refs - mean_ref).mean()
    assert np.abs(mine - mean_ref).mean() <= 1.5 * spread + 0.5, (np.abs(mine - mean_ref).mean(), spread)
    assert abs(mine.mean() - mean_ref.mean()) <= 0.02 * mean_ref.mean() + 0.5
    g.close()
    r.close()


@pytest.mark.parametrize("shader", [1, 2, 3])
def test_grid_mode_on_an_empty_scene_and_on_one_primitive(mrt, shader):
    """Degenerate inputs of the grid build: no primitives at all (bounds stay at +-RayLengthMax, every ray misses), and a single
    triangle (its grid has zero extent on one axis: the cell range collapses to cell 0 there, RegularGrid.hpp:243-259)."""
    from mobileraytracer_b200 import scene as S
    capi = mrt.capi
    w, h = 64, 48
    cam = S.Camera(capi.CAMERA_PERSPECTIVE, (0, 0, -3), (0, 0, 0), (0, 1, 0), 45.0, 45.0)
    g = mrt.Renderer(S.Scene(), cam, w, h, 2, 1, shader, accelerator=capi.ACC_REGULAR_GRID)
    bm = g.renderFrame()
    assert (np.asarray(bm).reshape(-1).view(np.uint32) == 0xFF000000).all()
    assert g.getTotalCastedRays() == 2 * w * h
    g.close()
    one = S.Scene()
    m = one.add_material(S.make_material(Kd=(0.5, 0.5, 0.5)))
    one.triangles_.append(S.make_triangle((-1, -1, 0), (1, -1, 0), (0, 1, 0), material=m))
    gg = mrt.Renderer(one, cam, w, h, 1, 1, shader, accelerator=capi.ACC_REGULAR_GRID)
    gb = mrt.Renderer(one, cam, w, h, 1, 1, shader)
    o = np.tile(np.array([[0.0, 0.0, -3.0]], np.float32), (4, 1))
    d = np.array([[0, 0, 1], [0.05, 0.02, 1], [1, 0, 0], [0, 0, -1]], np.float32)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    a, b = gg.traceRays(o, d), gb.traceRays(o, d)
    assert list(a["kind"]) == [3, 3, 0, 0] and np.array_equal(a["kind"], b["kind"]) and np.array_equal(a["t"], b["t"])
    gg.close()
    gb.close()

