"""GPU parity tests added in round 2 (VERDICT r1 "Next round" items 1, 4, 9 and the ADVICE findings), through the C-ABI:
the path tracer on the scene the bench numbers are quoted on, config 3 at its full size, config 5's replicated scene ray by
ray, every pixel sampler of the reference, queue overflow recovery, samples_light > 2, odd image sizes on the wavefront,
the tile partition and (with two devices) the NCCL-reduced frame against the single-rank frame."""
import os
import subprocess
import sys

import numpy as np
import pytest

from tests import helpers as H

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


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


def _box3(x, h, w):
    """3 x 3 box mean over the image plane of a (h * w, 3) array, edges replicated."""
    a = np.pad(x.reshape(h, w, 3), ((1, 1), (1, 1), (0, 0)), mode="edge")
    return sum(a[i:i + h, j:j + w] for i in range(3) for j in range(3)).reshape(-1, 3) / 9.0


def test_conference_pathtracer_image_matches_reference(mrt, ref, conf):
    """BASELINE config 4's own scene (8 area lights, 29 materials, three of them specular as in conference.mtl): GPU
    wavefront path tracer at 64 spp against 7 independent 64-spp frames of the reference (every RefRenderer re-shuffles the
    tables). SURVEY 8(d) bars: global mean luminance within 1 %; per-pixel difference to the mean of six reference frames
    within 3 sigma (+ 1 LSB of quantisation) on >= 99 % of the pixels, sigma from the scatter of those six frames (variance
    pooled over 3 x 3 pixels: six draws alone estimate a sigma to +-30 %). Path-traced pixels are heavy tailed, so the same
    statistic is taken for the seventh reference frame, which the six have not seen: the GPU frame must do as well as that
    control (minus one point), i.e. be indistinguishable from one more reference seed.
    Both sides use 64 samples per frame on purpose: the reference's 8-bit running average (incrementalAvg, Utils.cpp:66-90)
    truncates at every step, which biases a noisy pixel by about -n/4 LSB after n samples (n e_n = -sum k f_k with
    fractional parts f_k), and this library reproduces that arithmetic exactly, so frames are only comparable at equal n."""
    obj, mtl, cam = conf
    w, h, groups, spp = 240, 128, 7, 64
    frames, rays_ref, r = [], 0, None
    for i in range(groups):
        if r is not None:
            r.close()
        r = ref.RefRenderer(w, h, 2, -1, spp=spp, obj=obj, mtl=mtl, cam=cam)
        bm, _secs = r.render(threads=os.cpu_count() or 1)
        frames.append(H.unpack(bm).astype(np.float64))
        rays_ref += r.casted_rays()
    control, six = frames[0], np.stack(frames[1:])
    mean, var = six.mean(0), _box3(six.var(0, ddof=1), h, w)
    g = H.gpu_from_ref(ref, r, -1, 2, w, h, spp=spp, cam_file=cam)
    b = H.unpack(g.renderFrame()).astype(np.float64)
    assert abs(mean.mean() - b.mean()) <= 0.01 * mean.mean(), (mean.mean(), b.mean())
    tol = 3.0 * np.sqrt(var * (1.0 + 1.0 / (groups - 1))) + 1.0
    ok_gpu, ok_control = float((np.abs(b - mean) <= tol).mean()), float((np.abs(control - mean) <= tol).mean())
    assert ok_gpu >= min(0.99, ok_control - 0.01), (ok_gpu, ok_control)
    assert ok_gpu >= 0.97, (ok_gpu, ok_control)
    assert np.abs((b - mean).mean(axis=0)).max() < 1.0                            # no systematic colour shift
    # rank statistic: under "same distribution" the GPU value is the largest (smallest) of the seven values one time in seven;
    # a tie with m reference frames at the maximum counts 1 / (m + 1), which keeps the expectation at exactly 1/7
    mx, mn = six.max(0), six.min(0)
    top = float(np.where(b > mx, 1.0, np.where(b == mx, 1.0 / (1.0 + (six == mx).sum(0)), 0.0)).mean())
    bottom = float(np.where(b < mn, 1.0, np.where(b == mn, 1.0 / (1.0 + (six == mn).sum(0)), 0.0)).mean())
    assert abs(top - 1.0 / groups) < 0.025 and abs(bottom - 1.0 / groups) < 0.025, (top, bottom)
    ratio = g.getTotalCastedRays() / (rays_ref / groups)
    assert 0.98 < ratio < 1.02, ratio
    st = g.stats()
    assert st["rays_secondary"] > st["rays_primary"] * 0.5                        # specular children and diffuse bounces were traced
    g.close()
    r.close()


def test_conference_depthmap_full_size_frame(mrt, ref, conf):
    """BASELINE config 3 at its own size, 1920x1072: the DepthMap frame of the reference (all host threads; the shader draws
    nothing, so thread order cannot matter) against the GPU frame."""
    obj, mtl, cam = conf
    w, h = 1920, 1072
    r = ref.RefRenderer(w, h, 3, -1, obj=obj, mtl=mtl, cam=cam)
    bm_ref, _ = r.render(threads=os.cpu_count() or 1)
    g = H.gpu_from_ref(ref, r, -1, 3, w, h, cam_file=cam)
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, bm_ref)
    # the few pixels that differ are the stated epsilon edge cases (a ray lying exactly in a box plane of the reference's own
    # tree meets AABB::intersect's 0 * inf rule there): at most a handful in two million
    d = np.abs(H.unpack(bm) - H.unpack(bm_ref)).max(axis=1)
    assert c["exact"] >= 0.9999 and int((d > 1).sum()) <= 8, (c, int((d > 1).sum()))
    assert g.getTotalCastedRays() == w * h == r.casted_rays()
    g.close()
    r.close()


def test_config5_replicated_scene_rays_match_reference(mrt, ref, conf):
    """BASELINE config 5's synthetic scene at a size the reference's single-threaded BVH build finishes in seconds: the room
    copied 3 x 2 times (1.97 M triangles, 0.12 GB of BVH8: the scene no longer sits in a corner of the L2 next to the
    queues). Closest-hit and any-hit rays from random points of the WHOLE grid, ray by ray against the reference's walk,
    through the wavefront's persistent extend / shadow kernels (PathTracer handle)."""
    obj, mtl, cam = conf
    w = h = 64
    r = ref.RefRenderer(w, h, 3, -1, obj=obj, mtl=mtl, cam=cam, copies=(3, 2))
    assert r.n_tri > 1_900_000
    g = H.gpu_from_ref(ref, r, -1, 2, w, h, cam_file=cam)
    rng = np.random.default_rng(5)
    n = 150_000
    cell = rng.integers(0, 6, n)
    off = np.stack([(cell // 2 - 1) * 4200.0, np.zeros(n), (cell % 2 - 1) * 5200.0], axis=1)
    o = (rng.uniform(-1, 1, (n, 3)) * np.array([850.0, 380.0, 1350.0]) + np.array([0.0, 400.0, 0.0]) + off).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    a, b = r.trace(o, d), g.traceRays(o, d)
    same = (a["kind"] == b["kind"]) & (a["idx"] == b["idx"])
    ties = ~same & (a["t"] == b["t"])
    assert same.mean() >= 0.9999, (same.mean(), int(ties.sum()), int((~same).sum()))
    assert ((~same) & ~ties).sum() == 0
    assert np.array_equal(a["t"][same], b["t"][same])
    assert (a["kind"] != 0).mean() > 0.9
    dist = rng.uniform(50.0, 6000.0, n).astype(np.float32)
    sa, sb = r.trace(o, d, None, None, dist), g.traceRays(o, d, None, None, dist)
    assert np.array_equal(sa["kind"] != 0, sb["kind"] != 0)
    # and one primary-ray frame from the config-5 camera position
    bm_ref, _ = r.render(threads=os.cpu_count() or 1)
    g2 = H.gpu_from_ref(ref, r, -1, 3, w, h, cam_file=cam)
    c = H.compare_bitmaps(g2.renderFrame(), bm_ref)
    assert c["exact"] >= 0.999, c
    g.close(); g2.close()
    r.close()


@pytest.mark.parametrize("samplers,kind", [(2, "STATIC_MT"), (3, "HALTON"), (4, "STRATIFIED")])
def test_every_pixel_sampler_reproduces_the_reference_frame(mrt, ref, samplers, kind):
    """north_star (4) / SURVEY a4: pixel jitter from StaticMersenneTwister (StaticMersenneTwister.cpp:6-16), HaltonSeq
    (HaltonSeq.cpp:10-19, 1.0 after its domain) and Stratified (Stratified.cpp:10-19), index for index: scene 3, Whitted,
    4 spp, frame equal to the reference's single-thread frame."""
    from mobileraytracer_b200 import capi, scene as S
    w = h = 128
    spp = 4
    r = ref.RefRenderer(w, h, 1, 3, spp=spp, samplers=samplers)
    arrays = H.scene_arrays_from_ref(r)
    tables = {"tile_order": ref.table("tile"), "shader_table": ref.table("shader"), "static_halton": ref.table("static_halton")}
    if kind == "STATIC_MT":
        tables["static_mt"] = ref.table("static_mt")
    cam = S.BUILTIN[3][1](np.float32(1.0))
    g = mrt.Renderer(arrays, cam, w, h, spp, 1, capi.SHADER_WHITTED, depth_max_point=S.BUILTIN[3][2], tables=tables,
                     pixel_sampler=getattr(capi, "SAMPLER_" + kind))
    bm_ref, _ = r.render(threads=1)
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, bm_ref)
    assert c["within1"] >= 0.999 and c["exact"] >= 0.995, c
    assert g.getTotalCastedRays() == r.casted_rays()
    # and it is not another sampler's jitter in disguise
    g2 = mrt.Renderer(arrays, cam, w, h, spp, 1, capi.SHADER_WHITTED, depth_max_point=S.BUILTIN[3][2], tables=tables,
                      pixel_sampler=capi.SAMPLER_STATIC_HALTON)
    assert not np.array_equal(g2.renderFrame(), bm)
    g.close(); g2.close()
    r.close()


@pytest.mark.parametrize("samplers", [5, 6])
def test_stateful_generators_as_pixel_samplers_match_reference_frames(mrt, ref, samplers):
    """MersenneTwister (MersenneTwister.cpp:5-11) and PCG (PCG.cpp:6-13) keep a thread_local generator that is never reset, so
    their sequence can only come from the sampler object itself: a drop-in host takes the frame's draws from it and hands them
    over (MRT_SAMPLER_STREAM, mrt_set_pixel_stream). Checked end to end: one thread takes the draws, a second thread renders the
    reference frame with a generator that starts from the same seed (the oracle's seeded std::random_device is rewound in
    between), and the GPU frame fed with the draws equals that frame."""
    import threading
    from mobileraytracer_b200 import capi, scene as S
    w = h = 128
    spp = 4
    out = {}
    start = ref.random_device_counter()

    def reference_draws():
        ref.random_device_counter(start)
        r = ref.RefRenderer(w, h, 1, 3, spp=spp, samplers=samplers)
        out["draws"] = r.pixel_sampler_draws(2 * spp * w * h)
        out["arrays"] = H.scene_arrays_from_ref(r)
        r.close()

    def reference_frame():
        ref.random_device_counter(start)
        r = ref.RefRenderer(w, h, 1, 3, spp=spp, samplers=samplers)
        out["tables"] = {"tile_order": ref.table("tile").copy(), "shader_table": ref.table("shader").copy(), "static_halton": ref.table("static_halton").copy()}
        bm, _ = r.render(threads=1)
        out["bm"], out["rays"] = bm.copy(), r.casted_rays()
        r.close()
    # every thread constructs its own thread_local generator at its first draw
    for fn in (reference_draws, reference_frame):
        t = threading.Thread(target=fn)
        t.start()
        t.join()
    assert 0.0 <= out["draws"].min() and out["draws"].max() <= 1.0 and abs(out["draws"].mean() - 0.5) < 0.01
    cam = S.BUILTIN[3][1](np.float32(1.0))
    g = mrt.Renderer(out["arrays"], cam, w, h, spp, 1, capi.SHADER_WHITTED, depth_max_point=S.BUILTIN[3][2], tables=out["tables"],
                     pixel_sampler=capi.SAMPLER_STREAM)
    with pytest.raises(Exception):
        g.renderFrame()                       # no stream yet
    g.setPixelStream(out["draws"])
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, out["bm"])
    assert c["within1"] >= 0.999 and c["exact"] >= 0.995, c
    assert g.getTotalCastedRays() == out["rays"]
    g.close()


def _write_box_scene(tmp_path, kd, ks, tf):
    """A closed 2 x 2 x 2 box with an emissive ceiling patch; every wall has the given Kd / Ks / Tf (-> Kt)."""
    obj, mtl, cam = (str(tmp_path / ("box" + e)) for e in (".obj", ".mtl", ".cam"))
    with open(mtl, "w") as fh:
        fh.write("newmtl wall\n  Ns 10\n  Ni 1.5\n  d 1.0\n  Tf %g %g %g\n  illum 2\n  Ka 0 0 0\n  Kd %g %g %g\n  Ks %g %g %g\n  Ke 0 0 0\n\n" % (*tf, *kd, *ks))
        fh.write("newmtl light\n  Ke 4 4 4\n")
    v = [(-1, 0, -1), (1, 0, -1), (1, 0, 1), (-1, 0, 1), (-1, 2, -1), (1, 2, -1), (1, 2, 1), (-1, 2, 1),
         (-0.4, 1.98, -0.4), (0.4, 1.98, -0.4), (0.4, 1.98, 0.4), (-0.4, 1.98, 0.4)]
    quads = [(0, 1, 2, 3), (7, 6, 5, 4), (0, 4, 5, 1), (1, 5, 6, 2), (2, 6, 7, 3), (3, 7, 4, 0)]
    with open(obj, "w") as fh:
        fh.write("mtllib box.mtl\n")
        for p in v:
            fh.write("v %g %g %g\n" % p)
        fh.write("usemtl wall\n")
        for a, b, c, d in quads:
            fh.write("f %d %d %d\nf %d %d %d\n" % (a + 1, b + 1, c + 1, a + 1, c + 1, d + 1))
        fh.write("usemtl light\nf 9 10 11\nf 9 11 12\n")
    with open(cam, "w") as fh:
        fh.write("t perspective\np 0.0 1.0 -0.9\nl 0.0 1.0 1.0\nu 0.0 1.0 0.0\nf 60 60\n")
    return obj, mtl, cam


def test_branching_materials_overflow_the_queues_and_are_retried(mrt, ref, tmp_path):
    """a material with Kd, Ks and Kt > 0 triples the rays per depth level, which no fixed queue factor
    holds; the reference renders such scenes (its own teapot.mtl / CornellBox-Water.mtl have them). The pass that
    overflowed is repeated with fewer samples / larger queues and the frame still matches the reference statistically."""
    obj, mtl, cam = _write_box_scene(tmp_path, (0.5, 0.5, 0.5), (0.3, 0.3, 0.3), (0.8, 0.8, 0.8))
    w = h = 64
    r = ref.RefRenderer(w, h, 2, -1, spp=64, obj=obj, mtl=mtl, cam=cam)
    m = r.materials()
    wall = [i for i in range(r.n_mat) if m[i, 3] > 0][0]
    assert m[wall, 6] > 0, "the reference's loader did not give the wall a specular term"
    branching = 2 + (1 if m[wall, 9] > 0 else 0)
    bm_ref, _ = r.render(threads=os.cpu_count() or 1)
    g = H.gpu_from_ref(ref, r, -1, 2, w, h, spp=64, cam_file=cam)
    bm = g.renderFrame()                       # must not raise MRT_ERR_OVERFLOW
    st = g.stats()
    assert st["pass_retries"] > 0, (branching, st)
    a, b = H.unpack(bm_ref).astype(float), H.unpack(bm).astype(float)
    assert abs(a.mean() - b.mean()) <= 0.03 * a.mean() + 1.0, (a.mean(), b.mean())
    ratio = g.getTotalCastedRays() / float(r.casted_rays())
    assert 0.95 < ratio < 1.05, ratio
    # a second frame on the same handle (the queues keep their new size) gives the same picture
    bm2 = g.renderFrame()
    assert np.abs(H.unpack(bm2).astype(float).mean() - b.mean()) < 1.0
    g.close()
    r.close()


def test_four_light_samples_per_hit(mrt, ref):
    """samples_light > 2 used to overflow the shadow queue at level 0. Scene 2, spl = 4."""
    w = h = 96
    spp, spl = 64, 4
    r = ref.RefRenderer(w, h, 2, 2, spp=spp, spl=spl)
    bm_ref, _ = r.render(threads=os.cpu_count() or 1)
    g = H.gpu_from_ref(ref, r, 2, 2, w, h, spp=spp, spl=spl)
    bm = g.renderFrame()
    a, b = H.unpack(bm_ref).astype(float), H.unpack(bm).astype(float)
    assert abs(a.mean() - b.mean()) <= 0.015 * a.mean(), (a.mean(), b.mean())
    ratio = g.getTotalCastedRays() / float(r.casted_rays())
    assert 0.98 < ratio < 1.02, ratio
    assert g.stats()["pass_retries"] == 0
    g.close()
    r.close()


def test_pathtracer_on_an_image_size_that_is_not_a_multiple_of_16(mrt, ref):
    """for such sizes the reference's tiles wrap and overlap, several path slots of one sample hit the same
    pixel and the running average has to see them one after the other (Renderer.cpp:125-157). Scene 2 at 100 x 90:
    8-bit mode twice (deterministic), against the fp32-accumulate mode and against the reference's mean."""
    from mobileraytracer_b200 import capi
    w, h, spp = 100, 90, 32
    r = ref.RefRenderer(w, h, 2, 2, spp=spp)
    bm_ref, _ = r.render(threads=1)
    g = H.gpu_from_ref(ref, r, 2, 2, w, h, spp=spp)
    canvas = np.full(w * h, 0x00123456, np.int32)
    bm1 = g.renderFrame(canvas.copy()).copy()
    bm2 = g.renderFrame(canvas.copy()).copy()
    assert np.array_equal(bm1, bm2)
    untouched_ref = H.unpack(bm_ref).sum(axis=1) == 0                     # pixels no tile reaches stay as the caller left them
    assert np.array_equal(bm1.reshape(-1)[untouched_ref] == 0x00123456, np.ones(int(untouched_ref.sum()), bool)) or untouched_ref.sum() == 0
    a, b = H.unpack(bm_ref).astype(float), H.unpack(bm1).astype(float)
    covered = ~untouched_ref
    assert abs(a[covered].mean() - b[covered].mean()) <= 0.02 * a[covered].mean(), (a[covered].mean(), b[covered].mean())
    g.close()
    r.close()


def test_tile_partition_composes_the_full_frame(mrt, conf):
    """handles that render tiles tb, tb + 4, ... each write only their own pixels, and the sum of the
    four fp32 accumulators is the accumulator of the un-partitioned frame, bit for bit (the jitter draws of a pixel-sample do
    not depend on the partition). Also x sample halves: 4 tile groups x 2 sample groups = the 8-GPU layout."""
    import torch
    from mobileraytracer_b200 import loaders, capi
    obj, mtl, cam = conf
    w, h, spp = 480, 272, 4
    arrays = loaders.load_obj(obj, mtl).arrays()
    camera = loaders.load_camera(cam, np.float32(w) / np.float32(h))
    full = mrt.Renderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, flags=capi.FLAG_FP32_ACCUMULATE, seed=3)
    acc_full = torch.zeros(h * w, 4, device="cuda")
    full.renderAccumulateDevice(acc_full.data_ptr(), 0, 1)
    total = torch.zeros_like(acc_full)
    rays = 0
    for tb in range(4):
        part = mrt.Renderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, flags=capi.FLAG_FP32_ACCUMULATE, seed=3,
                            tile_begin=tb, tile_stride=4)
        for sb in range(2):
            acc = torch.zeros_like(acc_full)
            part.renderAccumulateDevice(acc.data_ptr(), sb, 2)
            touched = acc[:, 3] > 0
            assert 0.2 < float(touched.float().mean()) < 0.3          # a quarter of the tiles
            assert float(acc[touched, 3].min()) == spp // 2
            total += acc
            rays += part.getTotalCastedRays()
        assert part.stats()["queue_bytes"] < 0.3 * full.stats()["queue_bytes"]
        part.close()
    assert torch.equal(total, acc_full)
    assert rays == full.getTotalCastedRays()
    full.close()


def test_two_rank_nccl_frame_equals_the_single_rank_frame(mrt, conf):
    """ADVICE r1 (high) / VERDICT weak #8: DistributedRenderer over two NCCL ranks gives the frame one rank gives in
    fp32-accumulate mode (the reduce, the resolve and the copy are ordered on one explicit stream)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    out = os.path.join(ROOT, "gpurun_out", "two_rank_frame.npy")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    env = dict(os.environ, MRT_TWO_RANK_OUT=out)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(ROOT, "tests", "two_rank_frame.py")]
    subprocess.run(cmd, check=True, env=env, timeout=600, cwd=ROOT)
    got = np.load(out)
    from mobileraytracer_b200 import loaders, capi
    obj, mtl, cam = conf
    w, h, spp = 480, 272, 8
    arrays = loaders.load_obj(obj, mtl).arrays()
    camera = loaders.load_camera(cam, np.float32(w) / np.float32(h))
    one = mrt.Renderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, flags=capi.FLAG_FP32_ACCUMULATE, seed=1)
    want = one.renderFrame().reshape(-1)
    one.close()
    assert np.array_equal(got, want), float((got != want).mean())


def test_cpp_adapter_renders_the_golden_frames(mrt, tmp_path):
    """the C++ drop-in adapter (include/mobilert_b200_renderer.hpp), compiled against the reference's own
    headers and fed the reference's own Scene / Camera objects (Scenes.cpp) by tools/adapter_demo/adapter_main.cpp the way
    C_wrapper.cpp's work_thread would, renders on the GPU through the C-ABI; the frames equal the ones the unmodified reference
    rendered (tests/golden/reference_frames.npz). The executable is built where /root/reference exists (oracle/Makefile,
    target adapter) and travels prebuilt, like oracle/_ref/libmobilert_ref.so."""
    exe = os.path.join(ROOT, "oracle", "_ref", "b200_adapter_demo")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/b200_adapter_demo not built (needs /root/reference at build time)")
    g = np.load(os.path.join(ROOT, "tests", "golden", "reference_frames.npz"))
    frames = [k for k in g.files if k.startswith("frame_")]
    assert len(frames) >= 12
    for k in frames:
        _, scene, shader = k.split("_")
        out = str(tmp_path / (k + ".raw"))
        r = subprocess.run([exe, scene, shader, "64", "64", "1", "1", out], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, (k, r.stdout, r.stderr)
        bm = np.fromfile(out, np.int32)
        c = H.compare_bitmaps(bm, g[k])
        assert c["exact"] >= 0.998 and c["within1"] >= 0.999, (k, c)
        if int(scene) != 1:
            assert int(r.stdout.split()[1]) == int(g["rays_%s_%s" % (scene, shader)]), (k, r.stdout)


def test_cpp_adapter_on_an_obj_scene(mrt, ref, conf, tmp_path):
    """The same executable on the conference stand-in through the reference's OBJLoader + CameraFactory (DepthMap, 480x272):
    equal to the reference's own frame of that scene."""
    exe = os.path.join(ROOT, "oracle", "_ref", "b200_adapter_demo")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/b200_adapter_demo not built (needs /root/reference at build time)")
    obj, mtl, cam = conf
    w, h = 480, 272
    out = str(tmp_path / "conf.raw")
    r = subprocess.run([exe, obj, mtl, cam, "3", str(w), str(h), "1", "1", out], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout, r.stderr)
    rr = ref.RefRenderer(w, h, 3, -1, obj=obj, mtl=mtl, cam=cam)
    bm_ref, _ = rr.render(threads=os.cpu_count() or 1)
    rr.close()
    c = H.compare_bitmaps(np.fromfile(out, np.int32), bm_ref)
    assert c["exact"] >= 0.9999, c


@pytest.mark.parametrize("key", ["material", "primitive"])
def test_sort_by_material_renders_the_same_frame(mrt, conf, key, monkeypatch):
    """north_star (3) / VERDICT r1 missing #3: MRT_FLAG_SORT_BY_MATERIAL bins the hits of every level before shading. The draws
    of a vertex depend on the path, not on the order it is shaded in, so the accumulators are the same bit for bit."""
    import torch
    from mobileraytracer_b200 import loaders, capi
    obj, mtl, cam = conf
    w, h, spp = 480, 272, 4
    arrays = loaders.load_obj(obj, mtl).arrays()
    camera = loaders.load_camera(cam, np.float32(w) / np.float32(h))
    acc = []
    for flags in (capi.FLAG_FP32_ACCUMULATE, capi.FLAG_FP32_ACCUMULATE | capi.FLAG_SORT_BY_MATERIAL):
        monkeypatch.setenv("MRT_B200_SORT_KEY", key)
        g = mrt.Renderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, flags=flags, seed=3)
        a = torch.zeros(h * w, 4, device="cuda")
        g.renderAccumulateDevice(a.data_ptr(), 0, 1)
        acc.append((a, g.getTotalCastedRays()))
        g.close()
    assert acc[0][1] == acc[1][1]
    assert torch.equal(acc[0][0], acc[1][0])


def test_cpp_render_thread_state_machine(mrt, tmp_path):
    """SURVEY 8(f) rank 3 / VERDICT r1 missing #6: MobileRT::B200RenderThread (include/mobilert_b200_renderer.hpp) is the render
    thread + State machine the Android front-end wraps around renderFrame (JNI_layer.cpp:406-462,744-913). The adapter
    executable renders scene 3 (Whitted, 512 x 512, many samples) on a worker thread with MRT_FLAG_PROGRESSIVE while its main
    thread polls getSample() and the bitmap like the UI does: partial frames are visible, the finished frame equals the
    blocking call's frame, and stopRender(true) ends a frame early and leaves the machine IDLE."""
    exe = os.path.join(ROOT, "oracle", "_ref", "b200_adapter_demo")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/b200_adapter_demo not built (needs /root/reference at build time)")
    spp = 3000
    args = ["3", "1", "512", "512", str(spp), "1"]
    sync_out, async_out, stop_out = (str(tmp_path / n) for n in ("sync.raw", "async.raw", "stop.raw"))
    r = subprocess.run([exe] + args + [sync_out], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout, r.stderr)
    r = subprocess.run([exe] + args + [async_out], capture_output=True, text=True, timeout=600, env=dict(os.environ, MRT_DEMO_ASYNC="1"))
    assert r.returncode == 0, (r.stdout, r.stderr)
    f = dict(zip(r.stdout.split()[1::2], r.stdout.split()[2::2]))     # "async polls N busy N partial N stopped N final_sample N state N secs X"
    assert int(f["busy"]) > 0 and int(f["partial"]) > 0, r.stdout
    assert int(f["final_sample"]) == spp and int(f["state"]) == 0, r.stdout
    assert np.array_equal(np.fromfile(async_out, np.int32), np.fromfile(sync_out, np.int32))
    r = subprocess.run([exe] + args + [stop_out], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, MRT_DEMO_ASYNC="1", MRT_DEMO_STOP_AT="200"))
    assert r.returncode == 0, (r.stdout, r.stderr)
    f = dict(zip(r.stdout.split()[1::2], r.stdout.split()[2::2]))
    assert int(f["stopped"]) == 1 and 200 <= int(f["final_sample"]) < spp and int(f["state"]) == 0, r.stdout
    assert (np.fromfile(stop_out, np.int32) != 0).mean() > 0.5


def test_queues_shrink_instead_of_failing_on_a_full_device(mrt, conf):
    """mrt_create sizes the wavefront queues for as few passes as fit 40 GB. On a device that something else has filled it
    must come up with smaller passes (more of them) rather than fail, and render the same accumulators."""
    import torch
    from mobileraytracer_b200 import loaders, capi
    obj, mtl, cam = conf
    w, h, spp = 1920, 1072, 64
    arrays = loaders.load_obj(obj, mtl).arrays()
    camera = loaders.load_camera(cam, np.float32(w) / np.float32(h))
    roomy = mrt.Renderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, flags=capi.FLAG_FP32_ACCUMULATE, seed=5)
    a = torch.zeros(h * w, 4, device="cuda")
    roomy.renderAccumulateDevice(a.data_ptr(), 0, 16)        # 4 of the 64 samples
    full_bytes = roomy.stats()["queue_bytes"]
    roomy.close()
    torch.cuda.synchronize()
    free, _total = torch.cuda.mem_get_info()
    if free < (60 << 30):
        pytest.skip("the device is shared: not enough free memory to stage the experiment")
    filler = torch.empty(free - (14 << 30), dtype=torch.uint8, device="cuda")   # leave about 14 GB
    try:
        tight = mrt.Renderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, flags=capi.FLAG_FP32_ACCUMULATE, seed=5)
        b = torch.zeros(h * w, 4, device="cuda")
        tight.renderAccumulateDevice(b.data_ptr(), 0, 16)
        assert 0 < tight.stats()["queue_bytes"] < full_bytes
        tight.close()
    finally:
        del filler
        torch.cuda.empty_cache()
    assert full_bytes > (20 << 30)
    assert torch.equal(a, b)
