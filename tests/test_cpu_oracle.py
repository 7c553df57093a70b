"""CPU tests (-m "not gpu"): the oracle itself. The plain C++ restatement (oracle/mrt_oracle.cpp) is checked against
 * the reference's own known-answer unit tests (tests/golden/known_answers.json, transcribed from Unit_Testing/*.cpp),
 * golden frames produced by the unmodified compiled reference (tests/golden/reference_frames.npz, make_golden.py),
 * the compiled reference itself, bit for bit, wherever oracle/_ref is present (this container)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from oracle import portbind as P

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def f3(x):
    return np.ascontiguousarray(np.asarray(x, np.float32))


def make_tri(A, B, Cc):
    from mobileraytracer_b200.scene import make_triangle
    return np.asarray(make_triangle(A, B, Cc, material=0)).reshape(1)


@pytest.fixture(scope="module")
def ka():
    return json.load(open(os.path.join(GOLD, "known_answers.json")))


def test_port_known_answers_aabb(ka):
    L = P.lib()
    for c in ka["aabb"]["surface_area"]:
        assert L.port_aabb_area(P._p(f3(c["min"])), P._p(f3(c["max"]))) == np.float32(c["area"]), c["cite"]
    for c in ka["aabb"]["ray"]:
        hit = L.port_aabb_intersect(P._p(f3(c["min"])), P._p(f3(c["max"])), P._p(f3(c["o"])), P._p(f3(c["d"])))
        assert bool(hit) == c["hit"], c["cite"]


def test_port_known_answers_triangle(ka):
    L = P.lib()
    t = ka["triangle"]
    tri = make_tri(t["A"], t["B"], t["C"])
    box = np.zeros(6, np.float32)
    L.port_triangle_aabb(P._p(tri), P._p(box))
    assert np.array_equal(box[:3], f3(t["aabb"]["min"])) and np.array_equal(box[3:], f3(t["aabb"]["max"]))
    for c in t["ray"]:
        o = f3(c["o"])
        d = (f3(c["to"]) - o).astype(np.float32)
        dist = L.port_triangle_intersect(P._p(tri), P._p(o), P._p(d), C.c_float(1e30), 1 if c.get("self") else 0)
        assert (dist < 1e30) == c["hit"], c["cite"]


def test_port_known_answers_plane(ka):
    L = P.lib()
    p = ka["plane"]
    pl = np.zeros(1, P.PLANE)
    pl["normal"], pl["point"] = p["normal"], p["point"]
    box = np.zeros(6, np.float32)
    L.port_plane_aabb(P._p(pl), P._p(box))
    np.testing.assert_allclose(box[:3], p["aabb"]["min"], rtol=4e-7 * 4)      # ASSERT_FLOAT_EQ = 4 ulp
    np.testing.assert_allclose(box[3:], p["aabb"]["max"], rtol=4e-7 * 4)
    for c in p["ray"]:
        dist = L.port_plane_intersect(P._p(pl), P._p(f3(c["o"])), P._p(f3(c["d"])), C.c_float(1e30))
        assert (dist < 1e30) == c["hit"], c["cite"]


def test_port_matches_golden_reference_frames():
    g = np.load(os.path.join(GOLD, "reference_frames.npz"))
    frames = [k for k in g.files if k.startswith("frame_")]
    assert len(frames) == 14
    for k in frames:
        _, scene, shader = k.split("_")
        bm = P.render(64, 64, int(shader), int(scene))
        assert np.array_equal(bm, g[k]), k


def test_port_matches_golden_primary_hits():
    from mobileraytracer_b200 import scene as S
    g = np.load(os.path.join(GOLD, "reference_frames.npz"))
    for scene in (0, 1, 2, 3):
        make_scene, make_cam, mp = S.BUILTIN[scene]
        cam = make_cam(np.float32(1.0))
        camd = {"kind": cam.kind, "position": cam.position, "look_at": cam.look_at, "up": cam.up, "a": cam.a, "b": cam.b}
        r = P.PortRenderer(make_scene().arrays(), camd, 64, 64, 3)
        tr = r.trace(g["ray_%d_o" % scene], g["ray_%d_d" % scene])
        assert np.array_equal(tr["kind"], g["hit_%d_kind" % scene]) and np.array_equal(tr["t"], g["hit_%d_t" % scene]), scene
        r.close()


def test_port_helpers_match_closed_forms():
    L = P.lib()
    # haltonSequence base 2 = bit reversal (Utils.cpp:43-53)
    h = np.array([L.port_halton(i, 2) for i in range(1024)], np.float32)
    assert np.array_equal(h, P.halton_base2(1024))
    assert L.port_halton(5, 3) == np.float32(np.float32(np.float32(1) / np.float32(3)) * 2 + np.float32(np.float32(np.float32(1) / np.float32(3)) / np.float32(3)) * 1)
    # incrementalAvg (Utils.cpp:66-90): running mean of the truncated 8-bit samples
    rgb = f3([0.5, 0.25, 1.5])
    first = L.port_incremental_avg(P._p(rgb), 0x11223344, 1)
    assert np.uint32(first & 0xFFFFFFFF) == 0xFF000000 | (255 << 16) | (63 << 8) | 127
    second = L.port_incremental_avg(P._p(f3([0.0, 0.0, 0.0])), first, 2)
    assert np.uint32(second & 0xFFFFFFFF) == 0xFF000000 | (127 << 16) | (31 << 8) | 63   # the first sample was clamped to 255 when stored


def test_port_equals_compiled_reference_bit_for_bit(ref):
    """Pins the restatement: every deterministic shader on the built-in scenes, odd size included, plus ray counts."""
    for scene, shader, w, h in ((0, 1, 96, 96), (0, 0, 96, 96), (1, 1, 96, 96), (3, 1, 128, 96), (3, 0, 96, 96), (2, 3, 96, 96), (0, 1, 200, 120)):
        r = ref.RefRenderer(w, h, shader, scene)
        before = r.casted_rays()
        bm_ref, _ = r.render(threads=1)
        rays = r.casted_rays() - before
        tabs = {"tile_order": ref.table("tile"), "shader_table": ref.table("shader"), "static_halton": ref.table("static_halton")}
        from mobileraytracer_b200 import scene as S
        make_scene, make_cam, mp = S.BUILTIN[scene]
        cam = make_cam(np.float32(w) / np.float32(h))
        camd = {"kind": cam.kind, "position": cam.position, "look_at": cam.look_at, "up": cam.up, "a": cam.a, "b": cam.b}
        p = P.PortRenderer(make_scene().arrays(), camd, w, h, shader, 1, 1, mp, tabs)
        bm = p.render()
        assert np.array_equal(bm, bm_ref), (scene, shader, w, h)
        assert p.casted_rays() == rays, (scene, shader)
        p.close()
        r.close()


def test_port_4spp_jitter_equals_compiled_reference(ref):
    """BASELINE config 2 at reduced size: scene 3, Whitted, 4 spp with the shuffled-Halton pixel jitter and the 8-bit
    running average."""
    r = ref.RefRenderer(128, 128, 1, 3, spp=4)
    bm_ref, _ = r.render(threads=1)
    tabs = {"tile_order": ref.table("tile"), "shader_table": ref.table("shader"), "static_halton": ref.table("static_halton")}
    bm = P.render(128, 128, 1, 3, spp=4, tables=tabs)
    assert np.array_equal(bm, bm_ref)
    r.close()


def test_port_pathtracer_agrees_statistically_with_reference(ref):
    """The path tracer consumes global counters in execution order, so only the distribution can be compared."""
    r = ref.RefRenderer(64, 64, 2, 2, spp=32)
    bm_ref, _ = r.render(threads=1)
    tabs = {"tile_order": ref.table("tile"), "shader_table": ref.table("shader"), "static_halton": ref.table("static_halton")}
    bm = P.render(64, 64, 2, 2, spp=32, tables=tabs)
    a = (bm_ref.astype(np.uint32) & 255).astype(float)
    b = (bm.astype(np.uint32) & 255).astype(float)
    assert abs(a.mean() - b.mean()) < 0.01 * a.mean()
    assert np.abs(a - b).mean() < 12.0
    r.close()
