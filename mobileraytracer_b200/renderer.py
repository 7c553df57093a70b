"""Host-side mirror of the reference interface for the hot path, over the C-ABI of libmobilert_b200.so.

  Config      <- MobileRT::Config (app/MobileRT/Config.hpp:12-83)
  Renderer    <- MobileRT::Renderer public surface (app/MobileRT/Renderer.hpp:41-63): renderFrame, stopRender,
                 getSample, getTotalCastedRays
  RayTrace    <- extern "C" RayTrace(Config&, bool async) / stopRender() (app/System_dependent/Native/C_wrapper.h:12-20)

Every ray is traced by the CUDA library; nothing here computes pixels on the host."""
import ctypes as C
import threading
import time
from dataclasses import dataclass, field

import numpy as np

from . import capi
from . import scene as S


@dataclass
class Config:
    width: int = 512
    height: int = 512
    threads: int = 1            # kept for signature compatibility; the GPU ignores it
    shader: int = capi.SHADER_WHITTED
    sceneIndex: int = 0
    samplesPixel: int = 1
    samplesLight: int = 1
    repeats: int = 1
    accelerator: int = capi.ACC_B200
    printStdOut: bool = True
    objFilePath: str = ""
    mtlFilePath: str = ""
    camFilePath: str = ""
    bitmap: np.ndarray = field(default=None, repr=False)
    # additions of this implementation
    seed: int = 1
    flags: int = 0
    device: int = -1


class Renderer:
    """Owns one mrt_handle: scene + camera + samplers on the device, like Renderer owns shader_/camera_/samplerPixel_."""

    def __init__(self, scene, camera, width, height, samplesPixel, samplesLight=1, shader=capi.SHADER_WHITTED,
                 accelerator=capi.ACC_B200, depth_max_point=(1, 1, 1), pixel_sampler=capi.SAMPLER_AUTO, seed=1, flags=0,
                 tables=None, device=-1, light_sampler=capi.SAMPLER_AUTO, tile_begin=0, tile_stride=1):
        L = capi.lib()
        arr = scene.arrays() if isinstance(scene, S.Scene) else scene
        self._keep = arr
        sd = capi.SceneDesc()
        for name in ("triangles", "spheres", "planes", "materials", "lights"):
            a = arr[name]
            setattr(sd, name, a.ctypes.data if len(a) else None)
            setattr(sd, "n_" + name, len(a))
        texs = arr.get("textures") or []
        if texs:
            self._tex_pixels = [np.ascontiguousarray(t, np.uint8) for t in texs]
            self._tex_recs = (capi.Texture * len(texs))()
            for rec, px in zip(self._tex_recs, self._tex_pixels):
                hh, ww = px.shape[0], px.shape[1]
                rec.pixels, rec.width, rec.height, rec.channels = px.ctypes.data, ww, hh, (px.shape[2] if px.ndim == 3 else 1)
            self._mat_tex = np.ascontiguousarray(arr["material_texture"], np.int32)
            assert len(self._mat_tex) == len(arr["materials"])
            sd.textures, sd.n_textures, sd.material_texture = C.addressof(self._tex_recs), len(texs), self._mat_tex.ctypes.data
        rd = capi.RenderDesc()
        rd.width, rd.height, rd.samples_pixel, rd.samples_light = width, height, samplesPixel, samplesLight
        rd.shader, rd.accelerator, rd.pixel_sampler, rd.seed, rd.flags, rd.device = shader, accelerator, pixel_sampler, seed, flags, device
        rd.light_sampler = light_sampler
        rd.tile_begin, rd.tile_stride = tile_begin, tile_stride
        for k in range(3):
            rd.depth_max_point[k] = float(depth_max_point[k])
        tb = None
        if tables is not None:
            tb = capi.SamplerTables()
            self._tables = {k: np.ascontiguousarray(v, np.float32) for k, v in tables.items() if v is not None}
            for k, v in self._tables.items():
                setattr(tb, k, v.ctypes.data)
        cd = camera.desc()
        h = C.c_void_p()
        capi.check(L.mrt_create(C.byref(sd), C.byref(cd), C.byref(rd), C.byref(tb) if tb is not None else None, C.byref(h)))
        self._h = h
        self.width, self.height, self.samplesPixel, self.samplesLight, self.shader = width, height, samplesPixel, samplesLight, shader
        self.camera_ = camera

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().mrt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- MobileRT::Renderer surface
    def renderFrame(self, bitmap=None, numThreads=1):
        """Renderer::renderFrame(int32* bitmap, int numThreads): blocking; fills the caller's HOST bitmap (ABGR8888)."""
        if bitmap is None:
            bitmap = np.zeros(self.width * self.height, np.int32)
        assert bitmap.dtype == np.int32 and bitmap.size == self.width * self.height and bitmap.flags.c_contiguous
        capi.check(capi.lib().mrt_render_frame(self._h, capi.ptr(bitmap)), self._h, allow=(capi.MRT_ERR_STOPPED,))
        return bitmap.reshape(self.height, self.width)

    def stopRender(self):
        capi.lib().mrt_stop(self._h)

    def setReplayCounter(self, light_index_counter):
        """FLAG_REPLAY_SAMPLERS: value of Shader::getLightIndex's process-global counter at the next frame's first draw."""
        capi.check(capi.lib().mrt_set_replay_counter(self._h, C.c_uint32(int(light_index_counter) & 0xFFFFFFFF)), self._h)

    def getReplayCounter(self):
        v = C.c_uint32(0)
        capi.check(capi.lib().mrt_get_replay_counter(self._h, C.byref(v)), self._h)
        return v.value

    def getSample(self):
        v = C.c_int32(0)
        capi.lib().mrt_get_sample(self._h, C.byref(v))
        return v.value

    def getTotalCastedRays(self):
        v = C.c_uint64(0)
        capi.lib().mrt_get_total_rays(self._h, C.byref(v))
        return v.value

    # ---- device-resident entry points (headless replacement of the display path; multi-GPU partitions)
    def renderFrameDevice(self, device_ptr):
        capi.check(capi.lib().mrt_render_frame_device(self._h, C.c_void_p(device_ptr)), self._h, allow=(capi.MRT_ERR_STOPPED,))

    def renderAccumulateDevice(self, device_ptr, sample_begin, sample_stride):
        capi.check(capi.lib().mrt_render_accumulate_device(self._h, C.c_void_p(device_ptr), sample_begin, sample_stride), self._h,
                   allow=(capi.MRT_ERR_STOPPED,))

    def resolveDevice(self, accum_ptr, bitmap_ptr):
        capi.check(capi.lib().mrt_resolve_device(self._h, C.c_void_p(accum_ptr), C.c_void_p(bitmap_ptr)), self._h)

    def setCamera(self, camera):
        self.camera_ = camera
        cd = camera.desc()
        capi.check(capi.lib().mrt_set_camera(self._h, C.byref(cd)), self._h)

    def setStream(self, cuda_stream):
        """Run this renderer's kernels on the caller's CUDA stream (an integer cudaStream_t, e.g. torch's `cuda_stream`).
        torch's default stream has the handle 0, which mrt_set_stream reads as "back to the handle's own stream": it is
        passed as cudaStreamLegacy (0x1), the same stream under the name the C-ABI can tell apart. None restores the own stream."""
        if cuda_stream is None:
            handle = None
        else:
            handle = C.c_void_p(int(cuda_stream) if int(cuda_stream) != 0 else 1)
        capi.check(capi.lib().mrt_set_stream(self._h, handle), self._h)

    def setPixelStream(self, draws):
        """SAMPLER_STREAM: the pixel sampler's draws of the next frame(s), taken by the host from the sampler object itself
        (MersenneTwister / PCG keep hidden generator state), in the order Renderer::renderScene makes them."""
        d = np.ascontiguousarray(draws, np.float32)
        capi.check(capi.lib().mrt_set_pixel_stream(self._h, capi.ptr(d), len(d)), self._h)

    def stats(self):
        st = capi.Stats()
        capi.lib().mrt_get_stats(self._h, C.byref(st))
        return st.as_dict()

    # ---- per-ray access for parity tests
    def generateRays(self, uvdd):
        uvdd = np.ascontiguousarray(uvdd, np.float32).reshape(-1, 4)
        o = np.zeros((len(uvdd), 3), np.float32)
        d = np.zeros((len(uvdd), 3), np.float32)
        capi.check(capi.lib().mrt_generate_rays(self._h, len(uvdd), capi.ptr(uvdd), capi.ptr(o), capi.ptr(d)), self._h)
        return o, d

    def traceRays(self, o, d, origin_kind=None, origin_idx=None, shadow_dist=None):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        n = len(o)
        ok = np.ascontiguousarray(origin_kind, np.int32) if origin_kind is not None else None
        oi = np.ascontiguousarray(origin_idx, np.int32) if origin_idx is not None else None
        sd = np.ascontiguousarray(shadow_dist, np.float32) if shadow_dist is not None else None
        t = np.zeros(n, np.float32)
        kind, idx, mat = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.int32)
        nrm = np.zeros((n, 3), np.float32)
        capi.check(capi.lib().mrt_trace_rays(self._h, n, capi.ptr(o), capi.ptr(d), capi.ptr(ok), capi.ptr(oi), capi.ptr(sd),
                                             capi.ptr(t), capi.ptr(kind), capi.ptr(idx), capi.ptr(mat), capi.ptr(nrm)), self._h)
        return {"t": t, "kind": kind, "idx": idx, "mat": mat, "normal": nrm}

    def shadeRays(self, o, d, depth=1):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        rgb = np.zeros((len(o), 3), np.float32)
        capi.check(capi.lib().mrt_shade_rays(self._h, len(o), capi.ptr(o), capi.ptr(d), depth, capi.ptr(rgb)), self._h)
        return rgb


def round_down_to_multiple_of(value, multiple):
    """MobileRT::roundDownToMultipleOf (app/MobileRT/Utils/Utils.cpp:26-31), including its `rest > 1` quirk."""
    rest = value % multiple
    return value - rest if rest > 1 else value


_renderer = None
_renderer_lock = threading.Lock()


def build_renderer(config, tables=None):
    """What work_thread does before the first frame (C_wrapper.cpp:76-203)."""
    ratio = np.float32(config.width) / np.float32(config.height)
    if config.sceneIndex in S.BUILTIN:
        make_scene, make_cam, max_point = S.BUILTIN[config.sceneIndex]
        scene, camera = make_scene(), make_cam(ratio)
    else:
        from . import loaders
        scene = loaders.load_obj(config.objFilePath, config.mtlFilePath)
        camera = loaders.load_camera(config.camFilePath, ratio)
        max_point = (1, 1, 1)
    return Renderer(scene, camera, config.width, config.height, config.samplesPixel, config.samplesLight, config.shader,
                    capi.ACC_B200 if config.accelerator in (capi.ACC_BVH, capi.ACC_B200) else config.accelerator,
                    depth_max_point=max_point, seed=config.seed, flags=config.flags, tables=tables, device=config.device)


def RayTrace(config, async_=False, tables=None):
    """extern "C" void RayTrace(MobileRT::Config&, bool async): renders config.repeats frames into config.bitmap and
    reports the reference's own timing lines (C_wrapper.cpp:246-256). Returns a dict of those figures (sync mode)."""
    def work():
        global _renderer
        if config.bitmap is None:
            config.bitmap = np.zeros(config.width * config.height, np.int32)
        t0 = time.perf_counter()
        r = build_renderer(config, tables)
        with _renderer_lock:
            _renderer = r
        t1 = time.perf_counter()
        repeats = config.repeats
        while True:
            r.renderFrame(config.bitmap, config.threads)
            repeats -= 1
            if repeats <= 0:
                break
        t2 = time.perf_counter()
        rays = r.getTotalCastedRays() * max(config.repeats, 1)
        out = {"creating_s": t1 - t0, "rendering_s": t2 - t1, "casted_rays": rays, "mrays_per_s": rays / (t2 - t1) / 1e6,
               "stats": r.stats()}
        if config.printStdOut:
            print("Creating Time in secs = %g" % out["creating_s"])
            print("Rendering Time in secs = %g" % out["rendering_s"])
            print("Casted rays = %d" % rays)
            print("Total Millions rays per second = %g" % out["mrays_per_s"])
        with _renderer_lock:
            _renderer = None
        r.close()
        return out
    if async_:
        th = threading.Thread(target=work, daemon=True)
        th.start()
        return th
    return work()


def stopRender():
    """extern "C" void stopRender() (C_wrapper.cpp:271-275)."""
    with _renderer_lock:
        if _renderer is not None:
            _renderer.stopRender()
