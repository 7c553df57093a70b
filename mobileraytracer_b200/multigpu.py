"""Sample-partitioned rendering across the GPUs of one box: one process per GPU, scene replicated, one NCCL reduce.

The reference has no multi-device path; its only parallel axis is the tile loop of
Renderer::renderScene (app/MobileRT/Renderer.cpp:117-135), which works because every (pixel, sample) is independent.
The same independence is used here: rank g renders samples g, g + G, g + 2G, ... of every pixel into an fp32 RGBA
accumulator (sum of the per-sample trunc(255 c) values, .w = sample count), the accumulators are summed onto rank 0
by ONE torch.distributed reduce (NCCL over NVLink / NVSwitch) and rank 0 quantises once
(mrt_resolve_device: mean, truncate, clamp = what incrementalAvg converges to without its per-step truncation,
app/MobileRT/Utils/Utils.cpp:66-90).

torch is plumbing here (device buffers, the process group); every ray is traced by libmobilert_b200.so."""
import numpy as np

from . import capi
from .renderer import Renderer


def plan_samples(samples_total, rank, world):
    """Sample indices rank `rank` of `world` renders: rank, rank + world, ... < samples_total."""
    return list(range(rank, samples_total, world))


class DistributedRenderer:
    """Renderer surface over a process group. Every rank constructs it with the same scene / camera / tables."""

    def __init__(self, scene, camera, width, height, samples_total, samplesLight=1, shader=capi.SHADER_PATHTRACER,
                 depth_max_point=(1, 1, 1), seed=1, tables=None, group=None, device=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.width, self.height, self.samples_total = width, height, samples_total
        self.local = Renderer(scene, camera, width, height, samples_total, samplesLight, shader, depth_max_point=depth_max_point,
                              seed=seed, flags=capi.FLAG_FP32_ACCUMULATE, tables=tables, device=self.device.index)
        self.local.setStream(torch.cuda.current_stream(self.device).cuda_stream)
        self.accum = torch.zeros(height * width, 4, dtype=torch.float32, device=self.device)
        self.bitmap = torch.zeros(height * width, dtype=torch.int32, device=self.device)

    def renderFrameDevice(self):
        """All ranks: render the local share, reduce onto rank 0, resolve there. Returns the device bitmap on rank 0."""
        self.accum.zero_()
        self.local.renderAccumulateDevice(self.accum.data_ptr(), self.rank, self.world)
        if self.world > 1:
            self.dist.reduce(self.accum, dst=0, op=self.dist.ReduceOp.SUM, group=self.group)
        if self.rank == 0:
            self.local.resolveDevice(self.accum.data_ptr(), self.bitmap.data_ptr())
            return self.bitmap
        return None

    def renderFrame(self, host_bitmap=None):
        """Renderer::renderFrame for the whole group: rank 0 receives the ABGR8888 frame in a (pinned) host tensor."""
        dev = self.renderFrameDevice()
        if self.rank != 0:
            return None
        if host_bitmap is None:
            host_bitmap = self.torch.empty(self.height * self.width, dtype=self.torch.int32, pin_memory=True)
        host_bitmap.copy_(dev, non_blocking=True)
        self.torch.cuda.current_stream(self.device).synchronize()
        return host_bitmap

    def totalCastedRays(self):
        """Rays cast by all ranks in the last frame (Renderer::getTotalCastedRays summed over the group)."""
        t = self.torch.tensor([self.local.getTotalCastedRays()], dtype=self.torch.int64, device=self.device)
        if self.world > 1:
            self.dist.all_reduce(t, group=self.group)
        return int(t.item())

    def close(self):
        self.local.close()
