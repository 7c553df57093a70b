"""bench.py — the reference's headline metric on the configuration BASELINE.json quotes it on.

  metric   : Mrays/s (rays as the reference counts them: every Ray constructed = primary + shadow + secondary,
             app/MobileRT/Ray.cpp:11,26; printed as "Total Millions rays per second", C_wrapper.cpp:246-256) and ms/frame
  workload : BASELINE config 4 — conference scene (procedural stand-in, the .obj is missing from the reference snapshot),
             PathTracer, 1 sample per light, 1920x1072 (1080 rounded down to x16 as both front-ends do), 64 samples per
             pixel PER GPU; rank g renders samples g, g+N, ... of 64*N, one NCCL reduce, rank 0 quantises ("weak").
  step     : one frame.

  python bench.py [--gpus N --steps K --warmup W]           this library (sm_100a CUDA through the C-ABI)
  python bench.py --impl reference [...]                    the reference's own CPU renderer (oracle/_ref) on the host cores

Prints ONE JSON line (rank 0)."""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH, HEIGHT_ASKED, SPP, SPL = 1920, 1080, 64, 1
SHADER_PATHTRACER = 2
BYTES_PER_CLOSEST_RAY = 48     # 32 B ray record read + 16 B hit record written
BYTES_PER_SHADOW_RAY = 36      # 32 B read + 4 B visibility


def round_down16(v):
    # MobileRT::roundDownToMultipleOf (Utils.cpp:26-31) as the front-ends call it (Qt main.cpp:36-44)
    rest = v % 16
    return v - rest if rest > 1 else v


def scene_files():
    """Writes the conference stand-in once per box (deterministic bytes) and returns its paths."""
    from tools import make_conference_standin as gen
    out = os.environ.get("MRT_SCENE_DIR", os.path.join(tempfile.gettempdir(), "mrt_b200_scenes"))
    marker = os.path.join(out, "conference_standin.v2.ok")
    if not os.path.exists(marker):
        info = gen.write(out)
        with open(marker, "w") as fh:
            fh.write(str(info["triangles"]))
    base = os.path.join(out, "conference_standin")
    return base + ".obj", base + ".mtl", base + ".cam"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=10).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.splitlines()[0].split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=15)
        sm = [float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def base_line(args, height):
    return {"metric": "Mrays/s (primary+shadow+secondary rays, reference definition) on the conference path tracer", "unit": "Mrays/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic (procedural conference stand-in, seed 7, 328120 triangles; conference.obj is absent from the reference snapshot)",
            "config": {"workload": "BASELINE config 4: conference (stand-in), PathTracer, 1 spl, %dx%d, %d spp per GPU (sample-partitioned, NCCL reduce)" % (WIDTH, height, SPP),
                       "width": WIDTH, "height": height, "spp_per_gpu": SPP, "spp_total": SPP * args.gpus, "samples_light": SPL,
                       "shader": "PathTracer", "accelerator": "BVH", "parallelism": "sample-partition x%d" % args.gpus}}


# ---- the reference's own CPU renderer (oracle/_ref): the --impl reference arm and the cpu_baseline leg --------------
class ReferenceCpu:
    """The unmodified reference engine (compiled from /root/reference by oracle/Makefile into oracle/_ref), driven like
    C_wrapper.cpp's work_thread: OBJ scene, BVH accelerator, PathTracer, all host threads."""

    def __init__(self, height, spp):
        from oracle import refbind
        if not refbind.available():
            raise RuntimeError("oracle/_ref/libmobilert_ref.so missing (built by __graft_entry__.build() where /root/reference exists)")
        obj, mtl, cam = scene_files()
        self.threads = os.cpu_count() or 1
        self.spp = spp
        t0 = time.perf_counter()
        self.r = refbind.RefRenderer(WIDTH, height, SHADER_PATHTRACER, -1, spp=spp, spl=SPL, obj=obj, mtl=mtl, cam=cam)
        self.create_s = time.perf_counter() - t0
        self.bitmap = np.zeros(WIDTH * height, np.int32)
        self.rays_before = self.r.casted_rays()

    def frame(self):
        """One frame of `spp` samples; returns (seconds, rays cast by it)."""
        _, secs = self.r.render(threads=self.threads, bitmap=self.bitmap)
        now = self.r.casted_rays()
        rays, self.rays_before = now - self.rays_before, now
        return secs, rays


def cpu_baseline_sample(height):
    """Bounded sample of the same workload on the host cores: full resolution, 1 spp warm-up, then as many samples per
    pixel as fit in about 10 s (1..4)."""
    ref = ReferenceCpu(height, 1)
    s1, _ = ref.frame()
    spp = int(max(1, min(4, round(10.0 / max(s1, 1e-3)))))
    ref.r.close()
    ref = ReferenceCpu(height, spp)
    secs, rays = ref.frame()
    out = {"value": rays / secs / 1e6, "unit": "Mrays/s", "cores": ref.threads, "kind": "reference",
           "sample": "same scene, PathTracer, %dx%d, %d spp (of 64), 1 frame, %d threads, %.2f s; reference BVH+scene create %.2f s" % (
               WIDTH, height, spp, ref.threads, secs, ref.create_s),
           "ms_per_frame_at_64spp_extrapolated": secs * 1e3 * SPP / spp}
    ref.r.close()
    return out


def run_reference(args, height):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    spp = 1
    ref = ReferenceCpu(height, spp)
    for _ in range(args.warmup):
        ref.frame()
    tot_s, tot_rays = 0.0, 0
    for _ in range(args.steps):
        s, r = ref.frame()
        tot_s += s
        tot_rays += r
    line = base_line(args, height)
    v = tot_rays / tot_s / 1e6
    sample = "each step = one %dx%d PathTracer frame at %d spp (bounded sample of the 64-spp frame), %d host threads" % (WIDTH, height, spp, ref.threads)
    line.update({"impl": "reference", "value": v, "ms_per_step": tot_s * 1e3 / args.steps,
                 "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": ref.threads, "kind": "reference", "sample": sample},
                 "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    line["config"]["reference_sample"] = sample
    print(json.dumps(line))


# ---- this library -----------------------------------------------------------------------------------------------
def run_b200(args, height):
    import torch
    import torch.distributed as dist
    import mobileraytracer_b200 as mrt
    from mobileraytracer_b200 import capi, loaders
    from mobileraytracer_b200.multigpu import DistributedRenderer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if capi.lib().mrt_device_count() < 1:
        raise RuntimeError("no CUDA device: this library has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus

    obj, mtl, cam_path = scene_files() if rank == 0 else (None, None, None)
    if world > 1:
        dist.barrier()
        obj, mtl, cam_path = scene_files()
    scene = loaders.load_obj(obj, mtl)
    camera = loaders.load_camera(cam_path, np.float32(WIDTH) / np.float32(height))
    arrays = scene.arrays()
    n_tri, n_lights = len(arrays["triangles"]), len(arrays["lights"])
    stream = torch.cuda.current_stream(dev)

    # value: device-resident frame (scene, queues, accumulator and bitmap all in HBM)
    dr = DistributedRenderer(arrays, camera, WIDTH, height, SPP * world, SPL, capi.SHADER_PATHTRACER, seed=1)
    stats0 = dr.local.stats()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(args.warmup):
        dr.renderFrameDevice()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    if clocks:
        clocks.start()
    ext_ms = sh_ms = shade_ms = other_ms = 0.0
    ext_rays = sh_rays = launches = 0

    def step_value():
        nonlocal ext_ms, sh_ms, shade_ms, other_ms, ext_rays, sh_rays, launches
        dr.renderFrameDevice()
        st = dr.local.stats()     # host-side read of figures the library took with CUDA events on this stream
        ext_ms += st["extend_ms"]; sh_ms += st["shadow_ms"]; shade_ms += st["shade_ms"]; other_ms += st["other_ms"]
        ext_rays += st["rays_extend_traced"]; sh_rays += st["rays_shadow_traced"]
        launches += st["launches"] + (1 if rank == 0 else 0)

    total_ms = timed(step_value, args.steps)
    clk = clocks.summary() if clocks else None
    rays_frame = dr.totalCastedRays()
    st = dr.local.stats()
    value = rays_frame * args.steps / (total_ms * 1e-3) / 1e6

    # e2e: the reference-facing call with HOST buffers. N=1: mrt_set_camera + mrt_render_frame (exact integer running
    # average) into the caller's pinned bitmap; N>1: DistributedRenderer.renderFrame (reduce + resolve + D2H on rank 0).
    host_bitmap = torch.zeros(WIDTH * height, dtype=torch.int32).pin_memory()
    if world == 1:
        r1 = mrt.Renderer(arrays, camera, WIDTH, height, SPP, SPL, capi.SHADER_PATHTRACER, seed=1)
        r1.setStream(stream.cuda_stream)
        hb = host_bitmap.numpy()

        def step_e2e():
            r1.setCamera(camera)
            r1.renderFrame(hb)
        rays_e2e_frame = None
    else:
        def step_e2e():
            dr.renderFrame(host_bitmap)
    e2e_warm = max(1, min(args.warmup, 2))
    for _ in range(e2e_warm):
        step_e2e()
    e2e_ms = timed(step_e2e, args.steps)
    rays_e2e_frame = r1.getTotalCastedRays() if world == 1 else dr.totalCastedRays()
    e2e_value = rays_e2e_frame * args.steps / (e2e_ms * 1e-3) / 1e6

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
        else:
            peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
        achieved = BYTES_PER_CLOSEST_RAY * ext_rays / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None
        traffic = None
        tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("k_wf_extend_dram_bytes_per_launch")
        n_ext_launch = st["extend_launches"] * args.steps
        line = base_line(args, height)
        line.update({
            "value": value, "ms_per_step": total_ms / args.steps,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_frame": e2e_ms / args.steps, "h2d_bytes_per_step": 52,
                    "d2h_bytes_per_step": WIDTH * height * 4,
                    "path": "mrt_set_camera + mrt_render_frame (pinned host ABGR8888 bitmap)" if world == 1 else
                            "DistributedRenderer.renderFrame: mrt_render_accumulate_device per rank + NCCL reduce + mrt_resolve_device + D2H on rank 0"},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"kernel": "k_wf_extend (closest-hit BVH8 traversal)", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak if achieved else None, "traffic": traffic, "peak_source": peak_src,
                         "bytes_per_ray": BYTES_PER_CLOSEST_RAY, "rays_per_launch": ext_rays / max(n_ext_launch, 1),
                         "launch_ms_avg": ext_ms / max(n_ext_launch, 1),
                         "note": "compulsory queue bytes only; the 25 MB scene is L2-resident, so the kernel is L2-gather/latency bound, see profiles/"},
            "kernel_ms_per_frame": {"extend": ext_ms / args.steps, "shadow": sh_ms / args.steps, "shade": shade_ms / args.steps, "other": other_ms / args.steps},
            "rays_per_frame": {"total_all_gpus": rays_frame, "primary": st["rays_primary"], "shadow": st["rays_shadow"], "secondary": st["rays_secondary"],
                               "extend_traced": st["rays_extend_traced"], "shadow_traced": st["rays_shadow_traced"]},
            "kernel_grays_per_s": {"extend_closest_hit": ext_rays / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None,
                                   "shadow_any_hit": sh_rays / (sh_ms * 1e-3) / 1e9 if sh_ms > 0 else None,
                                   "note": "rays traversed by the kernel / summed CUDA-event time of its launches, all ranks' average"},
            "grays_per_s_primary_plus_shadow": (st["rays_primary"] + st["rays_shadow"]) * world * args.steps / (total_ms * 1e-3) / 1e9,
            "scene": {"triangles": n_tri, "area_lights": n_lights, "bvh_nodes": stats0["bvh_nodes"], "bvh_bytes": stats0["bvh_bytes"],
                      "build_ms": stats0["build_ms"], "upload_ms": stats0["upload_ms"]},
        })
        line["config"]["l2"] = ("not flushed between steps: every step streams its ray / hit / shadow queues (several GB per frame, far above the "
                                "126 MB L2) while the %.0f MB scene stays L2-resident by design" % (stats0["scene_bytes"] / 1e6))
        if args.gpus == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline_sample(height)
            except Exception as ex:   # the checker is optional infrastructure; the GPU numbers stand without it
                line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference", "sample": "unavailable: %s" % ex}
        print(json.dumps(line))
    barrier()
    dr.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    height = round_down16(HEIGHT_ASKED)
    if args.impl == "reference":
        run_reference(args, height)
    else:
        run_b200(args, height)


if __name__ == "__main__":
    main()
