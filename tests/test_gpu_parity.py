"""GPU parity tests proper: the CUDA path, called through the C-ABI, against the compiled reference (oracle/_ref) on the
same inputs. Bars: primitive id per primary ray >= 99.99 %, bitmap >= 99.9 % of channels within 1 LSB and
RMSE <= 1.0 for the deterministic shaders. Integer helpers are bit exact."""
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


# ---- primitive level -------------------------------------------------------------------------------------------------
def test_device_primitives_match_reference(mrt, ref):
    import ctypes as C
    capi = mrt.capi
    L = capi.lib()
    T = capi.test_lib()
    rng = np.random.default_rng(7)
    n_hit = 0
    for i in range(200):
        A, B, Cc = (rng.uniform(-2, 2, 3).astype(np.float32) for _ in range(3))
        tri = ref.triangle_build(A, B, Cc, 0)
        # aim at a point inside (or slightly outside) the triangle
        w = rng.uniform(-0.1, 1.1, 2)
        target = A + w[0] * (B - A) * 0.5 + w[1] * (Cc - A) * 0.5
        o = rng.uniform(-5, 5, 3).astype(np.float32)
        d = (target - o).astype(np.float32)
        d /= np.linalg.norm(d)
        t_ref, out_ref = ref.triangle_intersect(tri, o, d)
        t = C.c_float(0)
        out = np.zeros(8, np.float32)
        capi.check_test(T.mrt_test_triangle_intersect(capi.ptr(tri), capi.ptr(o), capi.ptr(d.astype(np.float32)), C.c_float(1e30), 0, C.byref(t), capi.ptr(out)))
        assert np.float32(t.value) == np.float32(t_ref)
        if t_ref < 1e29:
            n_hit += 1
            assert np.array_equal(out[:3], out_ref[:3]) and np.array_equal(out[5:], out_ref[5:])
    assert n_hit > 50
    for i in range(200):
        c = rng.uniform(-2, 2, 3).astype(np.float32)
        rad = np.float32(rng.uniform(0.1, 1.5))
        o = rng.uniform(-5, 5, 3).astype(np.float32)
        d = (c + rng.uniform(-1.2, 1.2, 3) * rad - o).astype(np.float32)
        d /= np.linalg.norm(d)
        t_ref, n_ref = ref.sphere_intersect(c, float(rad), o, d)
        s = np.zeros(1, capi.SPHERE)
        s["center"], s["sq_radius"] = c, rad * rad
        t = C.c_float(0)
        nrm = np.zeros(3, np.float32)
        capi.check_test(T.mrt_test_sphere_intersect(capi.ptr(s), capi.ptr(o), capi.ptr(d), C.c_float(1e30), C.byref(t), capi.ptr(nrm)))
        assert np.float32(t.value) == np.float32(t_ref)
        if t_ref < 1e29:
            assert np.array_equal(nrm, n_ref)
    axes = np.eye(3, dtype=np.float32)
    for i in range(100):
        nrm = axes[i % 3] * (1 if i % 2 else -1)
        p = rng.uniform(-2, 2, 3).astype(np.float32)
        o = rng.uniform(-5, 5, 3).astype(np.float32)
        d = rng.normal(size=3).astype(np.float32)
        d /= np.linalg.norm(d)
        t_ref = ref.plane_intersect(p, nrm, o, d)
        pl = np.zeros(1, capi.PLANE)
        pl["normal"], pl["point"] = nrm, p
        t = C.c_float(0)
        capi.check_test(T.mrt_test_plane_intersect(capi.ptr(pl), capi.ptr(o), capi.ptr(d), C.c_float(1e30), C.byref(t)))
        assert np.float32(t.value) == np.float32(t_ref)
        mn, mx = np.minimum(p, o * 0.3), np.maximum(p, o * 0.3) + np.float32(0.1)
        hit = C.c_int32(0)
        capi.check_test(T.mrt_test_aabb_intersect(capi.ptr(mn), capi.ptr(mx), capi.ptr(o), capi.ptr(d), C.byref(hit)))
        assert bool(hit.value) == ref.aabb_intersect(mn, mx, o, d)
    for i in range(300):
        rgb = rng.uniform(-0.2, 3.0, 3).astype(np.float32)
        avg = int(rng.integers(0, 2 ** 31 - 1)) | -16777216
        avg = int(np.int32(np.uint32(avg & 0xFFFFFFFF)))
        n = int(rng.integers(1, 200))
        out = C.c_int32(0)
        capi.check_test(T.mrt_test_incremental_avg(capi.ptr(rgb), avg, n, C.byref(out)))
        assert out.value == ref.incremental_avg(rgb, avg, n)
    h = np.zeros(4096, np.float32)
    capi.check_test(T.mrt_test_halton(0, 4096, 2, capi.ptr(h)))
    assert np.array_equal(h, np.array([ref.halton(i, 2) for i in range(4096)], np.float32))


# ---- per-ray: camera rays, primitive ids, t ----------------------------------------------------------------------------
@pytest.mark.parametrize("scene", [0, 1, 2, 3])
def test_primary_rays_and_hits_match_reference(mrt, ref, scene):
    w = h = 256
    r = ref.RefRenderer(w, h, 3, scene)
    g = H.gpu_from_ref(ref, r, scene, 3, w, h)
    uvdd = H.primary_uvdd(w, h)
    o_ref, d_ref = r.generate_rays(uvdd)
    o, d = g.generateRays(uvdd)
    assert np.array_equal(o, o_ref) and np.array_equal(d, d_ref), "camera rays must be bit identical"
    a = r.trace(o_ref, d_ref)
    b = g.traceRays(o, d)
    same = (a["kind"] == b["kind"]) & (a["idx"] == b["idx"])
    assert same.mean() >= 0.9999, "primitive id agreement %.6f" % same.mean()
    assert np.array_equal(a["t"][same], b["t"][same]), "t must be bit identical on agreeing rays"
    assert np.array_equal(a["normal"][same], b["normal"][same])
    assert np.array_equal(a["mat"][same], b["mat"][same])
    g.close()
    r.close()


@pytest.mark.parametrize("scene,shader", [(0, 1), (0, 0), (0, 3), (0, 4), (1, 1), (3, 1), (3, 0), (2, 3), (2, 4)])
def test_frame_matches_reference_1spp(mrt, ref, scene, shader):
    """BASELINE config 1 (scene 0, Whitted, 512x512, 1 spp) and its deterministic neighbours: bitmap equality."""
    w = h = 512
    r = ref.RefRenderer(w, h, shader, scene)
    bm_ref, _ = r.render(threads=1)
    rays_ref = r.casted_rays()
    g = H.gpu_from_ref(ref, r, scene, shader, w, h)
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, bm_ref)
    assert c["within1"] >= 0.999 and c["rmse"] <= 1.0, c
    assert c["exact"] >= 0.999, c
    if scene != 2:
        assert g.getTotalCastedRays() == rays_ref, (g.getTotalCastedRays(), rays_ref)
    g.close()
    r.close()


def test_builtin_scene_data_path_matches_reference(mrt, ref):
    """The package's own scene builders (scene.py) + RayTrace entry, i.e. the drop-in path a front-end would use."""
    for scene, shader in ((0, 1), (3, 1), (1, 4)):
        bm_ref, _ = ref.raytrace(256, 256, shader, scene)
        cfg = mrt.Config(width=256, height=256, shader=shader, sceneIndex=scene, printStdOut=False)
        mrt.RayTrace(cfg)
        c = H.compare_bitmaps(cfg.bitmap, bm_ref)
        assert c["exact"] >= 0.999, (scene, shader, c)


def test_config2_whitted_4spp_halton_jitter(mrt, ref):
    """BASELINE config 2: scene 3, Whitted, 4 spp, 1024x1024, pixel jitter from the reference's shuffled Halton table,
    accumulated with the reference's 8-bit running average."""
    w = h = 1024
    r = ref.RefRenderer(w, h, 1, 3, spp=4)
    bm_ref, _ = r.render(threads=1)
    g = H.gpu_from_ref(ref, r, 3, 1, w, h, spp=4)
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, bm_ref)
    assert c["within1"] >= 0.999 and c["rmse"] <= 1.0 and c["exact"] >= 0.995, c
    assert g.getTotalCastedRays() == r.casted_rays()
    g.close()
    r.close()


def test_odd_sizes_follow_the_reference_tile_walk(mrt, ref):
    """Sizes that are not multiples of 16: the reference's tiles wrap and overlap (Renderer.cpp:125-133). Pixels written by
    exactly one tile must match, pixels no tile reaches must keep the caller's content; pixels several tiles write are
    order dependent in the reference itself (thread races) and are left out."""
    for (w, h) in ((200, 120), (264, 264), (250, 100)):
        r = ref.RefRenderer(w, h, 1, 0)
        bm_ref = np.full(w * h, 0x11223344, np.int32)
        bm_ref, _ = r.render(threads=1, bitmap=bm_ref)
        g = H.gpu_from_ref(ref, r, 0, 1, w, h)
        bm = g.renderFrame(np.full(w * h, 0x11223344, np.int32)).reshape(-1)
        bm_ref = bm_ref.reshape(-1)
        bsx, bsy = w // 16, h // 16
        domain = (w // bsx) * (h // bsy)
        count = np.zeros(w * h + bsx * bsy + w * bsy, np.int32)
        for tv in ref.table("tile"):
            rb = int(np.floor(np.float32(tv) * np.float32(domain) + np.float32(0.5)))  # roundf: halves away from zero
            pixel = rb * bsx % (w * h)
            sy = ((pixel // w) * bsy) % h
            sx = pixel % w
            for y in range(sy, sy + bsy):
                count[y * w + sx:y * w + sx + bsx] += 1
        count = count[:w * h]
        once, never = count == 1, count == 0
        assert once.sum() > 0.5 * w * h
        assert (bm[never] == 0x11223344).all() and (bm_ref[never] == 0x11223344).all()
        c = H.compare_bitmaps(bm[once], bm_ref[once])
        assert c["exact"] >= 0.999, (w, h, c)
        g.close()
        r.close()


def _write_textured_scene(d):
    """A small OBJ scene with a textured material (map_Kd + vt, some coordinates outside [0, 1)), an untextured one and
    a .cam file; returns the three paths."""
    from PIL import Image
    rng = np.random.default_rng(5)
    Image.fromarray(rng.integers(0, 256, (32, 64, 3), dtype=np.uint8)).save(os.path.join(d, "tex.png"))
    Image.fromarray(rng.integers(0, 256, (16, 16, 4), dtype=np.uint8)).save(os.path.join(d, "tex4.png"))
    with open(os.path.join(d, "s.mtl"), "w") as fh:
        fh.write("newmtl tex\nKd 0.5 0.5 0.5\nmap_Kd tex.png\n\nnewmtl tex4\nKd 0.2 0.9 0.2\nmap_Kd tex4.png\n\nnewmtl plain\nKd 0.8 0.3 0.1\n")
    v, vt, faces = [], [], []
    def quad(x0, y0, x1, y1, z, t0, t1, mat):
        b = len(v) + 1
        v.extend([(x0, y0, z), (x1, y0, z), (x1, y1, z), (x0, y1, z)])
        tb = len(vt) + 1
        vt.extend([(t0[0], t0[1]), (t1[0], t0[1]), (t1[0], t1[1]), (t0[0], t1[1])])
        faces.append((mat, [(b, tb), (b + 1, tb + 1), (b + 2, tb + 2), (b + 3, tb + 3)]))
    quad(-2.0, -2.0, 0.0, 0.0, 0.0, (0.05, 0.05), (0.95, 0.9), "tex")
    quad(0.0, -2.0, 2.0, 0.0, 0.3, (0.1, 0.2), (2.6, 1.7), "tex")          # wraps: fract() per vertex, as the reference does
    quad(-2.0, 0.0, 0.0, 2.0, -0.2, (0.0, 0.0), (0.99, 0.99), "tex4")
    quad(0.0, 0.0, 2.0, 2.0, 0.1, (0.3, 0.3), (0.6, 0.6), "plain")
    with open(os.path.join(d, "s.obj"), "w") as fh:
        fh.write("mtllib s.mtl\n")
        fh.write("".join("v %.4f %.4f %.4f\n" % p for p in v))
        fh.write("".join("vt %.4f %.4f\n" % p for p in vt))
        for mat, idx in faces:
            fh.write("usemtl %s\nf %s\n" % (mat, " ".join("%d/%d" % p for p in idx)))
    with open(os.path.join(d, "s.cam"), "w") as fh:
        # off-axis on purpose: a ray lying exactly in the seam plane between two quads meets the 0 * inf case of the reference's
        # inner-node box test (AABB.cpp:34-54), which depends on the reference's own tree; only the root-box case is reproduced
        fh.write("t perspective\np 0.13 0.07 -5.0\nl 0.02 0.01 0.0\nu 0.0 1.0 0.0\nf 45 45\n")
    return os.path.join(d, "s.obj"), os.path.join(d, "s.mtl"), os.path.join(d, "s.cam")


def test_textured_obj_scene_matches_reference(mrt, ref, tmp_path):
    """Texture::loadColor + the per-hit Kd override of Shader::rayTrace (Shader.cpp:113-121): a textured OBJ scene loaded by
    this package's loader (PIL) and rendered on the GPU against the reference's own loader (stb_image) and renderer, with the
    DiffuseMaterial shader (returns Kd, no sampling)."""
    obj, mtl, cam = _write_textured_scene(str(tmp_path))
    w = h = 256
    r = ref.RefRenderer(w, h, 4, -1, obj=obj, mtl=mtl, cam=cam)
    bm_ref, _ = r.render(threads=1)
    cfg = mrt.Config(width=w, height=h, shader=4, sceneIndex=-1, samplesPixel=1, samplesLight=1, printStdOut=False,
                     objFilePath=obj, mtlFilePath=mtl, camFilePath=cam)
    mrt.RayTrace(cfg)
    c = H.compare_bitmaps(cfg.bitmap.reshape(h, w), bm_ref)
    colours = len(np.unique(cfg.bitmap))
    assert colours > 200, "the textures must be visible (%d distinct colours)" % colours
    assert c["exact"] >= 0.999 and c["max"] <= 1, c
    r.close()


@pytest.mark.parametrize("scene,shader", [(0, 1), (1, 1), (3, 1), (2, 3)])
def test_naive_accelerator_semantics_match_reference(mrt, ref, scene, shader):
    """Config::accelerator = ACC_NAIVE (Naive.hpp:80-91): planes are unbounded and every primitive is tested in scene order.
    The frame must equal the reference's Naive frame (which differs from its BVH frame where plane boxes cut planes off)."""
    w = h = 256
    r = ref.RefRenderer(w, h, shader, scene, accelerator=1)
    bm_ref, _ = r.render(threads=1)
    arrays = H.scene_arrays_from_ref(r)
    cam = H.camera_from_config(scene, np.float32(w) / np.float32(h), None)
    from mobileraytracer_b200 import scene as S
    g = mrt.Renderer(arrays, cam, w, h, 1, 1, shader, accelerator=mrt.capi.ACC_NAIVE, depth_max_point=S.BUILTIN[scene][2],
                     tables={"tile_order": ref.table("tile"), "shader_table": ref.table("shader"), "static_halton": ref.table("static_halton")})
    c = H.compare_bitmaps(g.renderFrame(), bm_ref)
    assert c["exact"] >= 0.999 and c["max"] <= 1, c
    g.close()
    r.close()


@pytest.mark.parametrize("shader", [0, 1, 2, 3, 4])
def test_empty_scene_renders_black(mrt, shader):
    """No primitives, no lights, no materials: every ray misses, rgb stays 0 (Shader.cpp:122) and the pixel is 0xFF000000;
    one primary ray per pixel is counted and nothing else."""
    from mobileraytracer_b200 import scene as S
    w, h = 64, 48
    cam = S.Camera(mrt.capi.CAMERA_PERSPECTIVE, (0, 0, -3), (0, 0, 0), (0, 1, 0), 45.0, 45.0)
    g = mrt.Renderer(S.Scene(), cam, w, h, 2, 1, shader)
    bm = g.renderFrame()
    assert (np.asarray(bm).reshape(-1).view(np.uint32) == 0xFF000000).all()
    assert g.getTotalCastedRays() == 2 * w * h
    t = g.traceRays(np.zeros((5, 3), np.float32), np.tile(np.array([[0, 0, 1]], np.float32), (5, 1)))
    assert (t["kind"] == 0).all() and (t["t"] == np.float32(1e30)).all()
    g.close()


def test_android_front_end_pixel_sampler_static_pcg(mrt, ref):
    """The Android front-end jitters pixels with StaticPCG (JNI_layer.cpp:616-618): scene 3, Whitted, 4 spp, the
    reference's PCG table handed over through mrt_sampler_tables::static_pcg -> same frame."""
    w = h = 256
    r = ref.RefRenderer(w, h, 1, 3, spp=4, samplers=1)
    bm_ref, _ = r.render(threads=1)
    g = H.gpu_from_ref(ref, r, 3, 1, w, h, spp=4, samplers=1)
    bm = g.renderFrame()
    c = H.compare_bitmaps(bm, bm_ref)
    assert c["within1"] >= 0.999 and c["exact"] >= 0.995, c
    assert g.getTotalCastedRays() == r.casted_rays()
    # and it is not the Halton jitter in disguise
    g2 = H.gpu_from_ref(ref, r, 3, 1, w, h, spp=4, samplers=0)
    assert not np.array_equal(g2.renderFrame(), bm)
    g.close(); g2.close()
    r.close()


def test_own_pcg_table_is_seeded_and_reproducible(mrt):
    """Without caller tables the library fills its own StaticPCG table (pcg32 XSH-RR, Utils.hpp:246-252): a jittered frame
    must be reproducible from the seed, change with it, and differ from the un-jittered frame."""
    import mobileraytracer_b200 as m
    from mobileraytracer_b200 import capi, scene as S
    make_scene, make_cam, mp = S.BUILTIN[3]
    frames = []
    for seed, sampler in ((5, capi.SAMPLER_STATIC_PCG), (5, capi.SAMPLER_STATIC_PCG), (6, capi.SAMPLER_STATIC_PCG), (5, capi.SAMPLER_CONSTANT)):
        r = m.Renderer(make_scene(), make_cam(np.float32(1.0)), 128, 128, 4, 1, capi.SHADER_WHITTED, depth_max_point=mp,
                       pixel_sampler=sampler, seed=seed)
        frames.append(r.renderFrame().copy())
        r.close()
    assert np.array_equal(frames[0], frames[1])
    assert not np.array_equal(frames[0], frames[2]) and not np.array_equal(frames[0], frames[3])
