// Wavefront path tracer (PathTracer::shade, Components/Shaders/PathTracer.cpp:22-142): generate / extend / shade /
// shadow / finalize kernels over compacted ray queues (mrt_wavefront.cu).
#pragma once
#include "mrt_render.h"

namespace mrt {

struct Wavefront;   // opaque: queues, chain records, counters

struct WavefrontTimes { double extend_ms = 0, shadow_ms = 0, shade_ms = 0, other_ms = 0; };

cudaError_t wavefront_create(int width, int height, Wavefront **out);
void wavefront_destroy(Wavefront *w);
// One sample of every pixel. avg_n = the sample number used by the integer running average (bitmap target).
cudaError_t wavefront_render_sample(Wavefront *w, const SceneDev &s, const CameraRec &cam, const RenderParams &rp, const int2 *d_tile_origin,
                                    int sample, int avg_n, FrameTargets tg, DeviceCounters *d_counters, bool sort_by_material,
                                    WavefrontTimes *times, uint64_t *launches, cudaStream_t st);

// After the stream has been synchronised: adds the device time of every launch since the last call, per kernel class.
cudaError_t wavefront_collect_times(Wavefront *w, WavefrontTimes *times);

}  // namespace mrt
