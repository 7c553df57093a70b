"""CPU tests (-m "not gpu") of the host side: the C-ABI library loads and exports every symbol the header declares (no
compute calls: there is no GPU here and no CPU fallback), argument validation, the scene / OBJ / camera ingest against
the reference's own loader, and the multi-GPU partition + reduce + resolve logic over gloo with world_size 2."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from mobileraytracer_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "mobilert_b200.h")).read()
    declared = set(re.findall(r"\b(mrt_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    L = capi.lib()
    for sym in sorted(declared):
        assert getattr(L, sym) is not None, sym
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    assert b"sm_100a" in L.mrt_version()
    # the second library: the device functions one call at a time (include/mobilert_b200_test.h); none of them in the product
    hdr_t = open(os.path.join(ROOT, "include", "mobilert_b200_test.h")).read()
    declared_t = set(re.findall(r"\b(mrt_test_[a-z0-9_]+)\s*\(", hdr_t))
    T = capi.test_lib()
    for sym in sorted(declared_t):
        assert getattr(T, sym) is not None, sym
        assert not hasattr(L, sym), sym + " must not ship in the product library"
    assert declared_t == set(capi.TEST_SYMBOLS), declared_t ^ set(capi.TEST_SYMBOLS)


def test_no_cpu_fallback_and_argument_checks():
    from mobileraytracer_b200 import capi
    L = capi.lib()
    h = C.c_void_p()
    assert L.mrt_create(None, None, None, None, C.byref(h)) == capi.MRT_ERR_INVALID
    assert L.mrt_stop(None) == capi.MRT_ERR_INVALID
    if L.mrt_device_count() == 0:
        import mobileraytracer_b200 as mrt
        with pytest.raises(capi.MrtError) as e:
            mrt.RayTrace(mrt.Config(width=32, height=32, printStdOut=False))
        assert e.value.code == capi.MRT_ERR_CUDA and "no CPU fallback" in str(e.value)


def test_record_layouts_match_the_reference():
    from mobileraytracer_b200 import capi
    assert capi.TRIANGLE.itemsize == 100 and capi.SPHERE.itemsize == 20 and capi.PLANE.itemsize == 28   # static_asserts of ref_api.cpp
    assert C.sizeof(capi.Stats) == 7 * 8 + 15 * 8   # mrt_stats (include/mobilert_b200.h)


def test_round_down_to_multiple_of():
    from mobileraytracer_b200 import round_down_to_multiple_of as rd
    assert rd(1080, 16) == 1072 and rd(1920, 16) == 1920 and rd(513, 16) == 513 and rd(530, 16) == 528   # Utils.cpp:26-31 (rest > 1)


def test_builtin_scene_arrays_equal_the_reference(ref):
    from mobileraytracer_b200 import scene as S
    for idx in (0, 1, 2, 3):
        r = ref.RefRenderer(32, 32, 3, idx)
        mine = S.BUILTIN[idx][0]().arrays()
        def rows(a, n):
            return sorted(bytes(x) for x in np.ascontiguousarray(a).view(np.uint8).reshape(n, -1))
        if r.n_tri:
            assert rows(mine["triangles"], r.n_tri) == rows(r.triangles(), r.n_tri), idx
        if r.n_sph:
            assert rows(mine["spheres"], r.n_sph) == rows(r.spheres(), r.n_sph), idx
        if r.n_pla:
            assert rows(mine["planes"], r.n_pla) == rows(r.planes(), r.n_pla), idx
        m = r.materials()
        for i in range(r.n_mat):
            mm = mine["materials"][i]
            assert np.array_equal(np.concatenate([mm["Le"], mm["Kd"], mm["Ks"], mm["Kt"], [mm["ior"]]]).astype(np.float32), m[i, :13]), (idx, i)
        r.close()


def test_obj_loader_matches_the_reference_loader(ref):
    from mobileraytracer_b200 import loaders
    base = "/root/reference/app/src/androidTest/resources/CornellBox/CornellBox-Water"
    if not os.path.exists(base + ".obj"):
        pytest.skip("reference fixtures not present on this machine")
    r = ref.RefRenderer(32, 32, 0, -1, obj=base + ".obj", mtl=base + ".mtl", cam=base + ".cam")
    arr = loaders.load_obj(base + ".obj", base + ".mtl").arrays()
    assert (len(arr["triangles"]), len(arr["lights"]), len(arr["materials"])) == (r.n_tri, r.n_lights, r.n_mat) == (7086, 2, 6)
    def geom(a):   # everything but the material index (de-duplication order depends on the reference's loader threads)
        return sorted(bytes(x[:96]) for x in np.ascontiguousarray(a).view(np.uint8).reshape(len(a), -1))
    assert geom(arr["triangles"]) == geom(r.triangles())
    ref_mats = {tuple(np.round(x[:13], 6)) for x in r.materials()}
    my_mats = {tuple(np.round(np.concatenate([m["Le"], m["Kd"], m["Ks"], m["Kt"], [m["ior"]]]), 6)) for m in arr["materials"]}
    assert ref_mats == my_mats
    # triangle -> material VALUE agrees
    rm, rt = r.materials(), r.triangles().view(np.dtype(arr["triangles"].dtype))
    kd_ref = {bytes(np.ascontiguousarray(t).view(np.uint8)[:36]): tuple(rm[t["material"], 3:6]) for t in rt}
    for t in arr["triangles"][::97]:
        assert kd_ref[bytes(np.ascontiguousarray(t).view(np.uint8)[:36])] == tuple(arr["materials"][t["material"]]["Kd"])
    cam = loaders.load_camera(base + ".cam", np.float32(1.0))
    c = r.camera()
    assert np.array_equal(cam.position, c["position"]) and np.float32(np.float32(cam.a) * np.float32(np.pi) / np.float32(180)) == c["a"]
    r.close()


def test_conference_standin_is_deterministic_and_loadable(tmp_path):
    from tools import make_conference_standin as gen
    from mobileraytracer_b200 import loaders
    a = gen.write(str(tmp_path / "a"), detail=0.25)
    b = gen.write(str(tmp_path / "b"), detail=0.25)
    assert open(a["obj"], "rb").read() == open(b["obj"], "rb").read()
    arr = loaders.load_obj(a["obj"], a["mtl"]).arrays()
    assert len(arr["lights"]) == 8 and len(arr["triangles"]) == a["triangles"] - 8 and len(arr["materials"]) <= 36
    assert np.isfinite(arr["triangles"]["nA"]).all()


def test_sample_partition_plan():
    from mobileraytracer_b200.multigpu import plan_samples
    for total in (1, 7, 64, 512):
        for world in (1, 2, 3, 8):
            got = sorted(s for r in range(world) for s in plan_samples(total, r, world))
            assert got == list(range(total))
            sizes = [len(plan_samples(total, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


_GLOO_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["MRT_ROOT"])
from mobileraytracer_b200.multigpu import plan_samples
from tests.helpers import resolve_reference
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
W = H = 16
spp = 9
rng = np.random.default_rng(123)                      # every rank draws the same per-sample images
samples = rng.uniform(0, 1.3, (spp, W * H, 3)).astype(np.float32)
acc = np.zeros((W * H, 4), np.float32)
for s in plan_samples(spp, rank, world):               # what mrt_render_accumulate_device adds per sample
    acc[:, :3] += np.floor(samples[s] * np.float32(255.0))
    acc[:, 3] += 1
t = torch.from_numpy(acc)
dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
if rank == 0:
    full = np.zeros((W * H, 4), np.float32)
    for s in range(spp):
        full[:, :3] += np.floor(samples[s] * np.float32(255.0)); full[:, 3] += 1
    assert np.array_equal(t.numpy(), full), "reduced accumulators differ from the single-rank sum"
    bm = resolve_reference(t.numpy())
    mean = np.minimum((full[:, :3] / spp).astype(np.uint32), 255)
    assert np.array_equal(bm.view(np.uint32) & 0xFF, mean[:, 0]) and np.array_equal((bm.view(np.uint32) >> 16) & 0xFF, mean[:, 2])
    print("GLOO_OK")
dist.destroy_process_group()
'''


def test_partition_reduce_resolve_over_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    env = dict(os.environ, MRT_ROOT=ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29611", str(script)], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "GLOO_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_cpp_adapter_compiles_against_reference(tmp_path):
    """include/mobilert_b200_renderer.hpp (the B200Renderer adapter of INTEGRATION.md) against the reference's own headers.
    Needs /root/reference (this container only); the GPU box has no reference tree."""
    import glob
    import subprocess
    ref = "/root/reference/app"
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present")
    tp = ref + "/third_party"
    src = tmp_path / "adapter_check.cpp"
    src.write_text('#include "mobilert_b200_renderer.hpp"\nint main() { return 0; }\n')
    cmd = ["/usr/bin/g++", "-std=c++17", "-fsyntax-only", "-fno-access-control", "-w", "-I", os.path.join(ROOT, "include"), "-I", ref,
           "-I", ref + "/System_dependent/Native", "-isystem", tp, "-isystem", tp + "/glm", "-isystem", tp + "/stb",
           "-isystem", tp + "/pcg-cpp/include"]
    for d in sorted(glob.glob(tp + "/boost/libs/*/include")):
        cmd += ["-isystem", d]
    r = subprocess.run(cmd + [str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]


def test_tile_and_sample_partition_covers_every_tile_sample_once():
    """plan_partition (multigpu.py): tile_groups x sample_groups ranks; every (tile, sample) pair belongs to exactly one rank."""
    from mobileraytracer_b200.multigpu import plan_partition
    for world, tg in ((1, 1), (2, 1), (2, 2), (4, 4), (8, 4), (8, 1), (8, 8)):
        seen = np.zeros((256, 16), np.int32)
        for rank in range(world):
            tb, ts, sb, ss = plan_partition(rank, world, tg)
            seen[tb::ts, sb::ss] += 1
        assert (seen == 1).all(), (world, tg)
    with pytest.raises(ValueError):
        plan_partition(0, 8, 3)


def _host_grid(capi, arrays, kind):
    """mrt_test_build_grid (host code of the product's MRT_ACC_REGULAR_GRID build; no device involved)."""
    T = capi.test_lib()
    sd = capi.SceneDesc()
    for name in ("triangles", "spheres", "planes", "materials", "lights"):
        a = arrays[name]
        setattr(sd, name, a.ctypes.data if len(a) else None)
        setattr(sd, "n_" + name, len(a))
    params, start, n = np.zeros(12, np.float32), np.zeros(32 ** 3 + 1, np.uint32), C.c_int64(0)
    capi.check_test(T.mrt_test_build_grid(C.byref(sd), kind, capi.ptr(params), capi.ptr(start), None, 0, C.byref(n)))
    items = np.zeros(max(n.value, 1), np.int32)
    capi.check_test(T.mrt_test_build_grid(C.byref(sd), kind, capi.ptr(params), capi.ptr(start), capi.ptr(items), n.value, C.byref(n)))
    return params, start, items[:n.value]


def test_regular_grid_cell_lists_equal_the_reference(ref, tmp_path):
    """The host-side grid build behind MRT_ACC_REGULAR_GRID (csrc/mrt_grid.h) against RegularGrid<T> of the compiled reference
    (RegularGrid.hpp:112-283 with Plane / Sphere / Triangle::intersect(AABB)): bounds, cell sizes and every cell's primitive
    set, bit for bit, for planes, spheres and triangles of the built-in scenes and a 16.7 k-triangle OBJ scene."""
    from mobileraytracer_b200 import capi
    from tests.helpers import scene_arrays_from_ref
    from tools import make_conference_standin as gen
    info = gen.write(str(tmp_path), detail=0.1)
    cases = [(s, {}) for s in (0, 1, 2, 3)] + [(-1, dict(obj=info["obj"], mtl=info["mtl"], cam=info["cam"]))]
    checked = 0
    for scene, kw in cases:
        r = ref.RefRenderer(32, 32, 3, scene, accelerator=2, **kw)
        arrays = scene_arrays_from_ref(r)
        for kind, count in ((0, r.n_pla), (1, r.n_sph), (2, r.n_tri)):
            if count == 0:
                continue
            mp, ms, mi = _host_grid(capi, arrays, kind)
            rp, rs, ri = r.grid_cells(kind)
            assert np.array_equal(mp.view(np.uint32), rp.view(np.uint32)), (scene, kind, mp, rp)
            assert np.array_equal(ms, rs) and np.array_equal(mi, ri), (scene, kind, len(mi), len(ri))
            assert len(mi) >= count
            checked += 1
        r.close()
    assert checked >= 10
