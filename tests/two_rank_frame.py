"""Helper of tests/test_gpu_round2.py::test_two_rank_nccl_frame_equals_the_single_rank_frame: launched with torchrun on two
GPUs; rank 0 saves the NCCL-reduced frame of DistributedRenderer (conference stand-in, PathTracer, 480x272, 8 spp) to
$MRT_TWO_RANK_OUT."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import bench
    from mobileraytracer_b200 import capi, loaders
    from mobileraytracer_b200.multigpu import DistributedRenderer
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if dist.get_rank() == 0:
        bench.scene_files()
    dist.barrier()
    obj, mtl, cam = bench.scene_files()
    w, h, spp = 480, 272, 8
    arrays = loaders.load_obj(obj, mtl).arrays()
    camera = loaders.load_camera(cam, np.float32(w) / np.float32(h))
    dr = DistributedRenderer(arrays, camera, w, h, spp, 1, capi.SHADER_PATHTRACER, seed=1)
    frame = None
    for _ in range(3):          # several frames back to back: an unordered resolve would show up as a flickering result
        frame = dr.renderFrame()
        if frame is not None:
            frame = frame.numpy().copy()
    if dist.get_rank() == 0:
        np.save(os.environ["MRT_TWO_RANK_OUT"], frame)
    dist.barrier()
    dr.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
