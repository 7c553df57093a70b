"""mrt_trace_rays_static (include/mobilert_b200.h): the allocation-free variant of mrt_trace_rays, whose queue scratch is a static
device array of MRT_STATIC_TRACE_MAX_RAYS rays. Checked against the existing entry point through the C-ABI, on two calls in
a row (the array keeps the previous call's contents), at the size limit and beyond it. Collected last on purpose (the file
name). Device buffers come from torch (plumbing); the library call itself allocates nothing."""
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
def room(tmp_path_factory):
    """The conference stand-in at a tenth of its detail (16.7 k triangles, 8 area lights), through the product's own loader."""
    from mobileraytracer_b200 import loaders
    from tools import make_conference_standin as gen
    info = gen.write(str(tmp_path_factory.mktemp("no_alloc_scene")), detail=0.1)
    return loaders.load_obj(info["obj"], info["mtl"]).arrays(), loaders.load_camera(info["cam"], np.float32(1.0))


def _rays(n, seed, n_tri, centre, extent):
    rng = np.random.default_rng(seed)
    o = (np.asarray(centre) + rng.uniform(-1, 1, (n, 3)) * np.asarray(extent)).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    kind = np.where((rng.uniform(size=n) < 0.1) & (n_tri > 0), 3, 0).astype(np.int32)
    idx = np.where(kind == 3, rng.integers(0, max(n_tri, 1), n), -1).astype(np.int32)
    dist = np.where(rng.uniform(size=n) < 0.4, rng.uniform(0.05, 2.0, n) * float(np.max(extent)), -1.0).astype(np.float32)   # >= 0: shadow ray
    return o, d, kind, idx, dist


def _static(mrt, g, o, d, kind, idx, dist):
    """mrt_trace_rays_static on device copies of the arrays; the outputs start as garbage."""
    import torch
    capi = mrt.capi
    dev = torch.device("cuda", 0)
    n = len(o)
    t_in = [torch.from_numpy(a).to(dev) for a in (o, d, kind, idx, dist)]
    out_t = torch.full((n,), -7.0, dtype=torch.float32, device=dev)
    out_kind, out_idx, out_mat = (torch.full((n,), -9, dtype=torch.int32, device=dev) for _i in range(3))
    out_n = torch.full((n, 3), 5.0, dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    capi.check(capi.lib().mrt_trace_rays_static(g._h, n, *(x.data_ptr() for x in t_in), out_t.data_ptr(), out_kind.data_ptr(),
                                                out_idx.data_ptr(), out_mat.data_ptr(), out_n.data_ptr()), g._h)
    return {"t": out_t.cpu().numpy(), "kind": out_kind.cpu().numpy(), "idx": out_idx.cpu().numpy(), "mat": out_mat.cpu().numpy(),
            "normal": out_n.cpu().numpy()}


def _same(a, b, exact_ids):
    """Bars of the per-ray parity tests (tests/test_gpu_conference.py): t bit-identical; a different primitive only at exactly
    the same distance. exact_ids: both sides ran the same kernels, so everything is bit-identical."""
    assert np.array_equal(a["t"], b["t"])
    assert np.array_equal(a["kind"] != 0, b["kind"] != 0)
    same = (a["kind"] == b["kind"]) & (a["idx"] == b["idx"])
    if exact_ids:
        assert same.all(), int((~same).sum())
    assert same.mean() >= 0.9999, same.mean()
    assert np.array_equal(a["mat"][same], b["mat"][same]) and np.array_equal(a["normal"][same], b["normal"][same])


@pytest.mark.parametrize("shader,accelerator", [(2, "b200"), (1, "b200"), (3, "grid")])
def test_static_scratch_entry_point_equals_mrt_trace_rays_twice_in_a_row(mrt, room, shader, accelerator):
    """shader 2: the handle's own mrt_trace_rays runs the same persistent kernels on cudaMalloc'ed queues -> bit-identical.
    shader 1: mrt_trace_rays answers through the per-thread traversal -> identical up to exact-t ties. grid: RegularGrid kernels."""
    capi = mrt.capi
    arrays, cam = room
    acc = capi.ACC_REGULAR_GRID if accelerator == "grid" else capi.ACC_B200
    g = mrt.Renderer(arrays, cam, 256, 256, 1, 1, shader, accelerator=acc)   # 256 x 256: the PathTracer handle's own queues hold 65536 rays
    n_tri = len(arrays["triangles"])
    # second call smaller than the first and with other rays: whatever the first call left in the array must not show
    for n, seed in ((50_000, 1), (12_345, 2), (capi.STATIC_TRACE_MAX_RAYS, 3)):
        o, d, kind, idx, dist = _rays(n, seed, n_tri, (0.0, 400.0, 0.0), (850.0, 380.0, 1350.0))
        want = g.traceRays(o, d, kind, idx, dist)
        got = _static(mrt, g, o, d, kind, idx, dist)
        assert (want["kind"] != 0).mean() > 0.3
        _same(want, got, exact_ids=(shader == 2 or accelerator == "grid"))
    g.close()


def test_static_scratch_with_planes_spheres_and_two_handles_in_turn(mrt):
    """Built-in scene 2 (planes, spheres, triangles, area lights) and scene 3 on two live handles used alternately: the shared
    array carries nothing from one handle's call into the other's."""
    from mobileraytracer_b200 import scene as S
    handles = []
    for index in (2, 3):
        make_scene, make_cam, max_point = S.BUILTIN[index]
        handles.append(mrt.Renderer(make_scene(), make_cam(np.float32(1.0)), 256, 256, 1, 1, 2, depth_max_point=max_point))
    for rnd in range(2):
        for k, g in enumerate(handles):
            o, d, kind, idx, dist = _rays(20_000 + 1000 * rnd, 10 * rnd + k, 0, (0.0, 0.0, 0.0), (1.5, 1.5, 1.5))
            _same(g.traceRays(o, d, None, None, dist), _static(mrt, g, o, d, kind, idx, dist), exact_ids=True)
    for g in handles:
        g.close()


def test_static_scratch_rejects_sizes_beyond_its_maximum(mrt, room):
    import torch
    capi = mrt.capi
    arrays, cam = room
    g = mrt.Renderer(arrays, cam, 64, 64, 1, 1, 1)
    n = capi.STATIC_TRACE_MAX_RAYS + 1
    buf = torch.zeros(3 * n, dtype=torch.float32, device="cuda:0")
    out = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    rc = capi.lib().mrt_trace_rays_static(g._h, n, buf.data_ptr(), buf.data_ptr(), None, None, None, buf.data_ptr(), out.data_ptr(),
                                          out.data_ptr(), out.data_ptr(), None)
    assert rc == capi.MRT_ERR_INVALID
    msg = capi.lib().mrt_last_error(g._h).decode()
    assert str(capi.STATIC_TRACE_MAX_RAYS) in msg and "static scratch" in msg, msg
    assert capi.lib().mrt_trace_rays_static(g._h, 0, buf.data_ptr(), buf.data_ptr(), None, None, None, buf.data_ptr(), out.data_ptr(),
                                            out.data_ptr(), out.data_ptr(), None) == capi.MRT_OK
    g.close()
