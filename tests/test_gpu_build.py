"""The GPU BVH builder (csrc/mrt_bvh_build.cu, replaces BVH<T>::build, BVH.hpp:160-283): properties that do not need the
reference. The warp-aggregated bin atomics must build the very tree the one-atomic-per-primitive path builds (minimum, maximum
and count do not depend on the order), a rebuild must be reproducible, and the LBVH alternative must answer rays identically."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mrt():
    import mobileraytracer_b200 as m
    if m.capi.lib().mrt_device_count() < 1:
        pytest.fail("no CUDA device: the product path has no CPU fallback")
    return m


@pytest.fixture(scope="module")
def scene():
    import bench
    from mobileraytracer_b200 import loaders
    obj, mtl, cam = bench.scene_files()
    return loaders.load_obj(obj, mtl).arrays(), loaders.load_camera(cam, np.float32(1.0))


def _rays(n, seed, n_tri):
    rng = np.random.default_rng(seed)
    o = (rng.uniform(-1, 1, (n, 3)) * np.array([850.0, 380.0, 1350.0]) + np.array([0.0, 400.0, 0.0])).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    kind = np.where(rng.uniform(size=n) < 0.1, 3, 0).astype(np.int32)
    idx = np.where(kind == 3, rng.integers(0, n_tri, n), -1).astype(np.int32)
    dist = np.where(rng.uniform(size=n) < 0.4, rng.uniform(50.0, 2500.0, n), -1.0).astype(np.float32)
    return o, d, kind, idx, dist


def _build_and_trace(mrt, scene, rays, monkeypatch, **env):
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    arrays, cam = scene
    g = mrt.Renderer(arrays, cam, 256, 256, 1, 1, mrt.capi.SHADER_PATHTRACER)
    for k in env:
        monkeypatch.delenv(k)
    st = g.stats()
    out = g.traceRays(*rays)
    g.close()
    return st, out


def test_aggregated_atomics_build_the_same_tree(mrt, scene, monkeypatch):
    rays = _rays(100_000, 5, len(scene[0]["triangles"]))
    st_a, a = _build_and_trace(mrt, scene, rays, monkeypatch)                                # default: warp-aggregated everywhere
    st_b, b = _build_and_trace(mrt, scene, rays, monkeypatch, MRT_B200_BUILD_AGG="0")        # one atomic per primitive and bin word
    st_c, c = _build_and_trace(mrt, scene, rays, monkeypatch, MRT_B200_BUILD_AGG="64")       # mixed: aggregated for the first levels only
    st_d, d = _build_and_trace(mrt, scene, rays, monkeypatch)                                # and again: reproducible
    assert st_a["bvh_nodes"] == st_b["bvh_nodes"] == st_c["bvh_nodes"] == st_d["bvh_nodes"] > 1000
    assert st_a["bvh_bytes"] == st_b["bvh_bytes"]
    for other in (b, c, d):
        for key in ("t", "kind", "idx", "mat", "normal"):
            assert np.array_equal(a[key], other[key]), key
    assert (a["kind"] != 0).mean() > 0.5


def test_lbvh_tree_answers_rays_identically(mrt, scene, monkeypatch):
    """Another topology, the same hits: t bit-identical, a different primitive only at exactly the same distance."""
    rays = _rays(100_000, 6, len(scene[0]["triangles"]))
    st_a, a = _build_and_trace(mrt, scene, rays, monkeypatch)
    st_b, b = _build_and_trace(mrt, scene, rays, monkeypatch, MRT_B200_BVH="lbvh")
    assert st_a["bvh_nodes"] != st_b["bvh_nodes"]
    assert np.array_equal(a["t"], b["t"]) and np.array_equal(a["kind"], b["kind"])
    same = a["idx"] == b["idx"]
    assert same.mean() >= 0.9999
    assert np.array_equal(a["normal"][same], b["normal"][same])
