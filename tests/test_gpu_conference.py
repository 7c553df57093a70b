"""GPU parity tests on the conference stand-in (BASELINE configs 3 and 4) and the path tracer, through the C-ABI.
The compiled reference (oracle/_ref) is the checker where a ground truth exists (primary-ray primitive ids, DepthMap and
NoShadows frames, path-tracer statistics); full-size runs are checked through size-independent properties."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from tests import helpers as H

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def mrt():
    import mobileraytracer_b200 as m
    if m.capi.lib().mrt_device_count() < 1:
        pytest.fail("no CUDA device: the product path has no CPU fallback")
    return m


@pytest.fixture(scope="module")
def conf():
    import bench
    return bench.scene_files()


def test_known_answer_vectors_on_device(mrt):
    """The reference's own unit-test vectors (Unit_Testing/TestAABB.cpp, TestTriangle.cpp, TestPlane.cpp) through the
    device functions."""
    from mobileraytracer_b200.scene import make_triangle
    capi = mrt.capi
    L = capi.lib()
    T = capi.test_lib()
    ka = json.load(open(os.path.join(GOLD, "known_answers.json")))
    f3 = lambda x: np.ascontiguousarray(np.asarray(x, np.float32))
    for c in ka["aabb"]["ray"]:
        hit = C.c_int32(-1)
        capi.check_test(T.mrt_test_aabb_intersect(capi.ptr(f3(c["min"])), capi.ptr(f3(c["max"])), capi.ptr(f3(c["o"])), capi.ptr(f3(c["d"])), C.byref(hit)))
        assert bool(hit.value) == c["hit"], c["cite"]
    t = ka["triangle"]
    tri = np.asarray(make_triangle(t["A"], t["B"], t["C"], material=0)).reshape(1)
    for c in t["ray"]:
        o = f3(c["o"])
        d = (f3(c["to"]) - o).astype(np.float32)
        dist = C.c_float(0)
        capi.check_test(T.mrt_test_triangle_intersect(capi.ptr(tri), capi.ptr(o), capi.ptr(d), C.c_float(1e30), 1 if c.get("self") else 0, C.byref(dist), None))
        assert (dist.value < 1e30) == c["hit"], c["cite"]
    p = ka["plane"]
    pl = np.zeros(1, capi.PLANE)
    pl["normal"], pl["point"] = p["normal"], p["point"]
    for c in p["ray"]:
        dist = C.c_float(0)
        capi.check_test(T.mrt_test_plane_intersect(capi.ptr(pl), capi.ptr(f3(c["o"])), capi.ptr(f3(c["d"])), C.c_float(1e30), C.byref(dist)))
        assert (dist.value < 1e30) == c["hit"], c["cite"]


def test_golden_frames_through_the_front_end_path(mrt):
    """Committed frames of the unmodified reference (tests/golden/make_golden.py) vs RayTrace(Config) on the GPU."""
    g = np.load(os.path.join(GOLD, "reference_frames.npz"))
    for k in [k for k in g.files if k.startswith("frame_")]:
        _, scene, shader = k.split("_")
        cfg = mrt.Config(width=64, height=64, shader=int(shader), sceneIndex=int(scene), printStdOut=False)
        out = mrt.RayTrace(cfg)
        c = H.compare_bitmaps(cfg.bitmap, g[k])
        assert c["exact"] >= 0.998 and c["within1"] >= 0.999, (k, c)
        if int(scene) != 1:   # scene 1: the reference also counts nothing else; orthographic root-box NaN rays agree anyway
            assert out["casted_rays"] == int(g["rays_%s_%s" % (scene, shader)]), k


def test_conference_primary_rays_match_reference(mrt, ref, conf):
    """BASELINE config 3: primitive id per primary ray >= 99.99 % (mismatches only at exact-t ties), t bit identical,
    DepthMap frame equal; 960x528 so that the single-threaded reference walk stays in seconds."""
    obj, mtl, cam = conf
    w, h = 960, 528
    r = ref.RefRenderer(w, h, 3, -1, obj=obj, mtl=mtl, cam=cam)
    g = H.gpu_from_ref(ref, r, -1, 3, w, h, cam_file=cam)
    uvdd = H.primary_uvdd(w, h)
    o_ref, d_ref = r.generate_rays(uvdd)
    o, d = g.generateRays(uvdd)
    assert np.array_equal(o, o_ref) and np.array_equal(d, d_ref)
    a, b = r.trace(o_ref, d_ref), g.traceRays(o, d)
    same = (a["kind"] == b["kind"]) & (a["idx"] == b["idx"])
    ties = ~same & (a["t"] == b["t"])          # equal distance, different primitive: shared edges
    assert same.mean() >= 0.9999, (same.mean(), int(ties.sum()), int((~same).sum()))
    assert ((~same) & ~ties).sum() <= 2
    assert np.array_equal(a["t"][same], b["t"][same])
    surf = same & (a["kind"] != 4)              # light hits: the library reports the face normal, the reference the vertex normals
    assert np.array_equal(a["normal"][surf], b["normal"][surf])
    bm_ref, _ = r.render(threads=os.cpu_count() or 1)
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, bm_ref)
    assert c["exact"] >= 0.9999 and c["rmse"] <= 0.05, c
    assert g.getTotalCastedRays() == w * h
    g.close()
    r.close()


def test_conference_noshadows_frame_statistics(mrt, ref, conf):
    """NoShadows with 8 area lights draws a light and a point on it per pixel in execution order, so the frames agree in
    distribution: per-channel mean within 1 %, and exactly wherever no light sample matters (emitters, background)."""
    obj, mtl, cam = conf
    w, h = 480, 272
    r = ref.RefRenderer(w, h, 0, -1, obj=obj, mtl=mtl, cam=cam, spp=8)
    bm_ref, _ = r.render(threads=1)
    g = H.gpu_from_ref(ref, r, -1, 0, w, h, spp=8, cam_file=cam)
    bm = g.renderFrame()
    a, b = H.unpack(bm_ref).astype(float), H.unpack(bm).astype(float)
    assert np.all(np.abs(a.mean(0) - b.mean(0)) <= 0.01 * a.mean(0) + 0.3), (a.mean(0), b.mean(0))
    assert np.sqrt(((a - b) ** 2).mean()) < 25.0
    g.close()
    r.close()


def test_pathtracer_converges_to_the_reference(mrt, ref):
    """Scene 2 (Cornell box, two area lights, mirror + glass spheres): GPU wavefront path tracer vs the reference at 256 spp.
    Statistical bar: global mean luminance within 1 %, per-pixel difference consistent with Monte-Carlo noise."""
    w = h = 96
    spp = 256
    r = ref.RefRenderer(w, h, 2, 2, spp=spp)
    bm_ref, _ = r.render(threads=os.cpu_count() or 1)
    g = H.gpu_from_ref(ref, r, 2, 2, w, h, spp=spp)
    bm = g.renderFrame()
    a, b = H.unpack(bm_ref).astype(float), H.unpack(bm).astype(float)
    assert abs(a.mean() - b.mean()) <= 0.01 * a.mean(), (a.mean(), b.mean())
    # two independent 256-spp estimates of the same integrand: compare 4x4 block means (noise / 4) channel by channel
    blk = lambda x: x.reshape(h // 4, 4, w // 4, 4, 3).mean(axis=(1, 3))
    # noise calibration: a second reference render with re-shuffled tables differs from the first by Monte-Carlo noise only
    r2 = ref.RefRenderer(w, h, 2, 2, spp=spp)
    bm_ref2, _ = r2.render(threads=os.cpu_count() or 1)
    r2.close()
    noise = np.abs(blk(a) - blk(H.unpack(bm_ref2).astype(float)))
    diff = np.abs(blk(a) - blk(b))
    assert diff.mean() <= 1.2 * noise.mean() + 0.5, (diff.mean(), noise.mean())
    assert np.percentile(diff, 99) <= 1.3 * np.percentile(noise, 99) + 2.0, (np.percentile(diff, 99), np.percentile(noise, 99))
    assert np.abs((blk(a) - blk(b)).mean(axis=(0, 1))).max() < 1.5      # no systematic colour shift
    # ray mix: same order of magnitude of rays per sample as the reference
    ratio = g.getTotalCastedRays() / float(r.casted_rays())
    assert 0.97 < ratio < 1.03, ratio
    g.close()
    r.close()


def test_full_size_config4_properties(mrt, conf):
    """BASELINE config 4 at full size (1920x1072) through size-independent properties: determinism, linearity of the
    fp32 accumulators in the sample set (what the multi-GPU partition relies on), and resolve == mean."""
    import torch
    from mobileraytracer_b200 import loaders, capi
    obj, mtl, cam = conf
    w, h, spp = 1920, 1072, 4
    arrays = loaders.load_obj(obj, mtl).arrays()
    camera = loaders.load_camera(cam, np.float32(w) / np.float32(h))
    g = mrt.Renderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, flags=capi.FLAG_FP32_ACCUMULATE, seed=3)
    acc = [torch.zeros(h * w, 4, device="cuda") for _ in range(4)]
    g.renderAccumulateDevice(acc[0].data_ptr(), 0, 1)          # all samples
    g.renderAccumulateDevice(acc[1].data_ptr(), 0, 2)          # even samples (rank 0 of 2)
    g.renderAccumulateDevice(acc[2].data_ptr(), 1, 2)          # odd samples (rank 1 of 2)
    g.renderAccumulateDevice(acc[3].data_ptr(), 0, 1)          # again: determinism
    assert torch.equal(acc[0], acc[3])
    assert torch.equal(acc[0], acc[1] + acc[2])
    assert float(acc[0][:, 3].min()) == spp and float(acc[0][:, 3].max()) == spp
    bm = torch.zeros(h * w, dtype=torch.int32, device="cuda")
    g.resolveDevice(acc[0].data_ptr(), bm.data_ptr())
    expect = H.resolve_reference(acc[0].cpu().numpy())
    assert np.array_equal(bm.cpu().numpy(), expect)
    st = g.stats()
    assert st["rays_primary"] == w * h * spp and st["rays_extend_traced"] >= st["rays_primary"]
    g.close()


def test_stop_and_progress_polling(mrt):
    """Renderer::stopRender / getSample (Renderer.cpp:93-99,176-178) from another thread while renderFrame blocks."""
    import threading
    import time
    cfg = mrt.Config(width=1024, height=1024, shader=1, sceneIndex=3, samplesPixel=4000, printStdOut=False)
    r = mrt.build_renderer(cfg)
    bm = np.zeros(1024 * 1024, np.int32)
    th = threading.Thread(target=lambda: r.renderFrame(bm))
    th.start()
    mid, deadline = 0, time.time() + 20.0
    while mid == 0 and time.time() < deadline and th.is_alive():   # the first sample may wait for lazy CUDA module loading
        time.sleep(0.05)
        mid = r.getSample()
    r.stopRender()
    th.join(timeout=60)
    assert not th.is_alive()
    assert 0 < mid < 4000 and r.getSample() < 4000
    assert (bm != 0).any()
    r.close()


@pytest.mark.parametrize("shader,scene,spp", [(1, 3, 12000), (2, 0, 2048)])
def test_progressive_bitmap_builds_up_while_rendering(mrt, shader, scene, spp):
    """MRT_FLAG_PROGRESSIVE: the caller's HOST bitmap shows the frame while renderFrame is still running, as the
    reference's in-place pixel writes do for the Android / Qt UI threads (Renderer.cpp:149-157, JNI_layer.cpp:818-821);
    the finished frame is the same as without the flag. Megakernel (Whitted) and wavefront (PathTracer) paths."""
    import threading
    import time
    from mobileraytracer_b200 import capi
    w = h = 512
    plain = mrt.build_renderer(mrt.Config(width=w, height=h, shader=shader, sceneIndex=scene, samplesPixel=spp, printStdOut=False))
    expect = plain.renderFrame().copy()
    plain.close()
    r = mrt.build_renderer(mrt.Config(width=w, height=h, shader=shader, sceneIndex=scene, samplesPixel=spp, printStdOut=False,
                                      flags=capi.FLAG_PROGRESSIVE))
    bm = np.zeros(w * h, np.int32)
    th = threading.Thread(target=lambda: r.renderFrame(bm))
    th.start()
    partial_seen, deadline = 0, time.time() + 120.0
    while th.is_alive() and time.time() < deadline:
        s = r.getSample()
        filled = bool((bm != 0).mean() > 0.5)
        if filled and s < spp and th.is_alive():
            partial_seen += 1
        time.sleep(0.002)
    th.join(timeout=120)
    assert not th.is_alive()
    assert partial_seen > 0, "the host bitmap never showed a partial frame"
    assert np.array_equal(bm.reshape(h, w), expect)
    r.close()


@pytest.mark.parametrize("shader", [2, 0])
def test_conference_incoherent_rays_match_reference(mrt, ref, conf, shader):
    """Secondary-ray shaped work: 200 k rays from random points inside the room in random directions (closest hit, some
    leaving a triangle they must skip) and 200 k shadow rays with random lengths (any hit), ray by ray against the
    reference's BVH walk. shader=2: the handle is a PathTracer, so mrt_trace_rays answers through the persistent
    cooperative extend / shadow kernels of the wavefront (the bench's hot kernels); shader=0: per-thread traversal."""
    obj, mtl, cam = conf
    w, h = 64, 64
    r = ref.RefRenderer(w, h, 3, -1, obj=obj, mtl=mtl, cam=cam)
    g = H.gpu_from_ref(ref, r, -1, shader, w, h, cam_file=cam)
    rng = np.random.default_rng(11)
    n = 200_000
    o = (rng.uniform(-1, 1, (n, 3)) * np.array([850.0, 380.0, 1350.0]) + np.array([0.0, 400.0, 0.0])).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    kind = np.where(rng.uniform(size=n) < 0.1, 3, 0).astype(np.int32)
    idx = np.where(kind == 3, rng.integers(0, r.n_tri, n), -1).astype(np.int32)
    a, b = r.trace(o, d, kind, idx), g.traceRays(o, d, kind, idx)
    same = (a["kind"] == b["kind"]) & (a["idx"] == b["idx"])
    ties = ~same & (a["t"] == b["t"])
    assert same.mean() >= 0.9999, (same.mean(), int(ties.sum()), int((~same).sum()))
    assert ((~same) & ~ties).sum() == 0, "a different primitive is only acceptable at exactly the same distance"
    assert np.array_equal(a["t"][same], b["t"][same]), "t must be bit identical"
    surf = same & (a["kind"] == 3)
    assert np.array_equal(a["normal"][surf], b["normal"][surf]) and np.array_equal(a["mat"][surf], b["mat"][surf])
    assert (a["kind"] != 0).mean() > 0.9          # the room is closed: nearly every ray hits something
    # any hit: occlusion flags must agree exactly (which occluder is found first is free)
    dist = rng.uniform(50.0, 2500.0, n).astype(np.float32)
    sa, sb = r.trace(o, d, kind, idx, dist), g.traceRays(o, d, kind, idx, dist)
    occ_ref, occ_gpu = sa["kind"] != 0, sb["kind"] != 0
    assert 0.2 < occ_ref.mean() < 0.98
    assert np.array_equal(occ_ref, occ_gpu), int((occ_ref != occ_gpu).sum())
    g.close()
    r.close()


@pytest.mark.parametrize("shader", [1, 0])
def test_deterministic_shaders_with_area_lights_agree_statistically(mrt, ref, shader):
    """Scene 2 has two area lights, so Whitted / NoShadows draw light indices and light positions from the reference's
    global counters in its execution order (Shader.cpp:223-233, AreaLight.cpp:17-26): without a replay of that order
    (MRT_FLAG_REPLAY_SAMPLERS, tested below) the frames can only agree as Monte-Carlo estimates. Same bars as the path tracer:
    mean luminance within 1 %, block means within the noise measured between two reference renders, ray counts within 1 %."""
    w = h = 96
    spp = 128
    r = ref.RefRenderer(w, h, shader, 2, spp=spp)
    bm_ref, _ = r.render(threads=1)
    g = H.gpu_from_ref(ref, r, 2, shader, w, h, spp=spp)
    bm = g.renderFrame()
    a, b = H.unpack(bm_ref).astype(float), H.unpack(bm).astype(float)
    assert abs(a.mean() - b.mean()) <= 0.01 * a.mean(), (a.mean(), b.mean())
    blk = lambda x: x.reshape(h // 4, 4, w // 4, 4, 3).mean(axis=(1, 3))
    r2 = ref.RefRenderer(w, h, shader, 2, spp=spp)
    bm_ref2, _ = r2.render(threads=1)
    r2.close()
    noise = np.abs(blk(a) - blk(H.unpack(bm_ref2).astype(float)))
    diff = np.abs(blk(a) - blk(b))
    assert diff.mean() <= 1.2 * noise.mean() + 0.5, (diff.mean(), noise.mean())
    assert np.percentile(diff, 99) <= 1.3 * np.percentile(noise, 99) + 2.0, (np.percentile(diff, 99), np.percentile(noise, 99))
    ratio = g.getTotalCastedRays() / float(r.casted_rays())     # a shadow ray is only cast when the sampled light point is in front
    assert 0.99 < ratio < 1.01, ratio
    g.close()
    r.close()


@pytest.mark.parametrize("shader,spp,spl", [(1, 1, 1), (0, 1, 1), (1, 3, 2)])
def test_replay_samplers_reproduces_reference_draw_order(mrt, ref, shader, spp, spl):
    """MRT_FLAG_REPLAY_SAMPLERS: on scene 2 (two area lights) every getLightIndex / AreaLight::getPosition draw reads the
    table entry the reference's single-thread frame reads (Shader.cpp:223-233, AreaLight.cpp:17-26), so the frames agree
    like the single-light scenes do: bit for bit. Two frames in a row: the getLightIndex counter runs on across frames
    (function-static), the lights' own counters restart (Renderer.cpp:58-59)."""
    from mobileraytracer_b200 import capi
    w = h = 128
    r = ref.RefRenderer(w, h, shader, 2, spp=spp, spl=spl)
    c0 = ref.light_index_counter()
    bm_ref1, _ = r.render(threads=1)
    c1, rays1 = ref.light_index_counter(), r.casted_rays()
    bm_ref2, _ = r.render(threads=1)
    c2, rays2 = ref.light_index_counter(), r.casted_rays() - rays1
    assert c1 > c0
    g = H.gpu_from_ref(ref, r, 2, shader, w, h, spp=spp, spl=spl, flags=capi.FLAG_REPLAY_SAMPLERS)
    g.setReplayCounter(c0)
    bm1 = g.renderFrame().copy()
    assert g.getReplayCounter() == c1 & 0xFFFFFFFF       # as many index draws as the reference made
    assert g.getTotalCastedRays() == rays1               # shadow rays are cast only towards sampled points in front
    bm2 = g.renderFrame().copy()
    assert g.getReplayCounter() == c2 & 0xFFFFFFFF
    assert g.getTotalCastedRays() == rays2
    for bm, bm_ref in ((bm1, bm_ref1), (bm2, bm_ref2)):
        c = H.compare_bitmaps(bm, bm_ref)
        assert c["exact"] >= 0.999 and c["max"] <= 1 or c["exact"] >= 0.9999, c
    g.close()
    r.close()


def test_replay_samplers_rejected_for_path_tracer(mrt, ref):
    from mobileraytracer_b200 import capi
    r = ref.RefRenderer(32, 32, 2, 2, spp=1)
    with pytest.raises(Exception):
        H.gpu_from_ref(ref, r, 2, 2, 32, 32, flags=capi.FLAG_REPLAY_SAMPLERS)
    r.close()


@pytest.mark.parametrize("samplers", [0, 1])
def test_conference_whitted_with_eight_area_lights_replayed_exactly(mrt, ref, conf, samplers):
    """The conference stand-in has 8 area lights created by the front-end's sampler lambda (OBJLoader::fillScene):
    samplers=0 is the Linux front-end's StaticHaltonSeq (C_wrapper.cpp:109,145), samplers=1 the Android front-end's
    StaticPCG for pixel jitter and lights (JNI_layer.cpp:601,616-618). With MRT_FLAG_REPLAY_SAMPLERS the Whitted frame
    (2 spp, jittered) is the reference's single-thread frame bit for bit, and the casted-ray counts are equal."""
    from mobileraytracer_b200 import capi
    obj, mtl, cam = conf
    w, h = 256, 144
    r = ref.RefRenderer(w, h, 1, -1, obj=obj, mtl=mtl, cam=cam, spp=2, samplers=samplers)
    c0 = ref.light_index_counter()
    bm_ref, _ = r.render(threads=1)
    g = H.gpu_from_ref(ref, r, -1, 1, w, h, spp=2, cam_file=cam, flags=capi.FLAG_REPLAY_SAMPLERS, samplers=samplers)
    g.setReplayCounter(c0)
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, bm_ref)
    assert c["exact"] >= 0.9995 and c["within1"] >= 0.9999, c
    assert g.getReplayCounter() == ref.light_index_counter() & 0xFFFFFFFF
    assert abs(g.getTotalCastedRays() - r.casted_rays()) <= 4      # an exact-t tie on a shared edge may pick the other triangle
    g.close()
    r.close()
