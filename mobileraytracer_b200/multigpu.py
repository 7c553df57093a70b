"""Sample- and tile-partitioned rendering across the GPUs of one box: one process per GPU, scene replicated, one NCCL reduce.

The reference has no multi-device path; its only parallel axis is the tile loop of
Renderer::renderScene (app/MobileRT/Renderer.cpp:117-135), which works because every (pixel, sample) is independent.
The same independence is used here. With G = tile_groups x sample_groups ranks, rank g = tg + tile_groups * sg renders

    tiles   tg, tg + tile_groups, ...      of the 256 tiles of the reference's shuffled tile walk (Renderer.cpp:189-197)
    samples sg, sg + sample_groups, ...    of every pixel of those tiles

into an fp32 RGBA accumulator (sum of the per-sample trunc(255 c) values, .w = sample count). The accumulators are summed
onto rank 0 by ONE torch.distributed reduce (NCCL over NVLink / NVSwitch) and rank 0 quantises once (mrt_resolve_device:
mean, truncate, clamp = what incrementalAvg converges to without its per-step truncation, app/MobileRT/Utils/Utils.cpp:66-90).
BASELINE config 4 uses the sample partition alone (tile_groups = 1), config 5 tiles x samples.

Everything a frame does on the device (zeroing, the library's kernels, the reduce, the resolve, the copy to the host) is
enqueued on ONE explicit stream owned by this object, so the steps are ordered by the stream and nothing else: torch's
default stream has the handle 0, which mrt_set_stream reads as "the handle's own stream", and that left the resolve
unordered against the NCCL kernel in the first version.

torch is plumbing here (device buffers, the process group); every ray is traced by libmobilert_b200.so."""
import numpy as np

from . import capi
from .renderer import Renderer


def plan_samples(samples_total, rank, world):
    """Sample indices rank `rank` of `world` renders: rank, rank + world, ... < samples_total."""
    return list(range(rank, samples_total, world))


def plan_partition(rank, world, tile_groups=1):
    """(tile_begin, tile_stride, sample_begin, sample_stride) of rank `rank`: tile_groups x (world / tile_groups) ranks."""
    if tile_groups < 1 or world % tile_groups != 0:
        raise ValueError("tile_groups must divide the world size")
    sample_groups = world // tile_groups
    return rank % tile_groups, tile_groups, rank // tile_groups, sample_groups


class DistributedRenderer:
    """Renderer surface over a process group. Every rank constructs it with the same scene / camera / tables."""

    def __init__(self, scene, camera, width, height, samples_total, samplesLight=1, shader=capi.SHADER_PATHTRACER,
                 depth_max_point=(1, 1, 1), seed=1, tables=None, group=None, device=None, tile_groups=1):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.width, self.height, self.samples_total = width, height, samples_total
        self.tile_begin, self.tile_stride, self.sample_begin, self.sample_stride = plan_partition(self.rank, self.world, tile_groups)
        self.local = Renderer(scene, camera, width, height, samples_total, samplesLight, shader, depth_max_point=depth_max_point,
                              seed=seed, flags=capi.FLAG_FP32_ACCUMULATE, tables=tables, device=self.device.index,
                              tile_begin=self.tile_begin, tile_stride=self.tile_stride)
        self.stream = torch.cuda.Stream(self.device)
        self.local.setStream(self.stream.cuda_stream)
        with torch.cuda.stream(self.stream):
            self.accum = torch.zeros(height * width, 4, dtype=torch.float32, device=self.device)
            self.bitmap = torch.zeros(height * width, dtype=torch.int32, device=self.device)
        self.stream.synchronize()
        self.events = None

    def _mark(self, name):
        if self.events is not None:
            e = self.torch.cuda.Event(enable_timing=True)
            e.record(self.stream)
            self.events.append((name, e))

    def renderFrameDevice(self):
        """All ranks: render the local share, reduce onto rank 0, resolve there. Returns the device bitmap on rank 0."""
        torch = self.torch
        with torch.cuda.stream(self.stream):
            self._mark("begin")
            self.accum.zero_()
            self._mark("zero")
            self.local.renderAccumulateDevice(self.accum.data_ptr(), self.sample_begin, self.sample_stride)   # same stream, blocking
            self._mark("render")
            if self.world > 1:
                self.dist.reduce(self.accum, dst=0, op=self.dist.ReduceOp.SUM, group=self.group)   # the stream waits for NCCL
            self._mark("reduce")
            if self.rank == 0:
                self.local.resolveDevice(self.accum.data_ptr(), self.bitmap.data_ptr())
                self._mark("resolve")
                return self.bitmap
        return None

    def renderFrame(self, host_bitmap=None):
        """Renderer::renderFrame for the whole group: rank 0 receives the ABGR8888 frame in a (pinned) host tensor."""
        dev = self.renderFrameDevice()
        if self.rank != 0:
            return None
        if host_bitmap is None:
            host_bitmap = self.torch.empty(self.height * self.width, dtype=self.torch.int32, pin_memory=True)
        with self.torch.cuda.stream(self.stream):
            host_bitmap.copy_(dev, non_blocking=True)
            self._mark("d2h")
        self.stream.synchronize()
        return host_bitmap

    def timePhases(self, on=True):
        """Start (or stop) recording one CUDA event per frame phase on the frame stream; phaseTimes() reads them."""
        self.events = [] if on else None

    def phaseTimes(self):
        """ms per phase summed over the frames since timePhases(): zero, render, reduce, resolve, d2h."""
        out = {}
        self.stream.synchronize()
        ev = self.events or []
        for (_, e0), (name, e1) in zip(ev, ev[1:]):
            if name != "begin":
                out[name] = out.get(name, 0.0) + e0.elapsed_time(e1)
        self.events = [] if self.events is not None else None
        return out

    def totalCastedRays(self):
        """Rays cast by all ranks in the last frame (Renderer::getTotalCastedRays summed over the group)."""
        t = self.torch.tensor([self.local.getTotalCastedRays()], dtype=self.torch.int64, device=self.device)
        if self.world > 1:
            self.dist.all_reduce(t, group=self.group)
        return int(t.item())

    def close(self):
        self.local.close()
