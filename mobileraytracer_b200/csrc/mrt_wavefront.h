// Wavefront path tracer (PathTracer::shade, Components/Shaders/PathTracer.cpp:22-142): generate / extend / shade /
// shadow / finalize kernels over compacted ray queues (mrt_wavefront.cu).
#pragma once
#include "mrt_render.h"

namespace mrt {

struct Wavefront;   // opaque: queues, chain records, counters

struct WavefrontTimes { double extend_ms = 0, shadow_ms = 0, shade_ms = 0, other_ms = 0; uint64_t extend_launches = 0; };

constexpr int kWfMaxBatch = 64;
constexpr int kWfCounts = 32;   // Queues::counts entries (see mrt_wavefront.cu)
// batch: samples of every pixel traced together in one pass (1..kWfMaxBatch); the queues are sized for it and for
// samples_light shadow rays per hit
// sort_mode: 0 off; 1 = MRT_FLAG_SORT_BY_MATERIAL (the shade kernel visits the hits of a level binned by material); 2 = binned by
// primitive id instead (diagnostic: MRT_B200_SORT_KEY=primitive)
cudaError_t wavefront_create(size_t paths_per_sample, int batch, int samples_light, int sort_mode, uint32_t n_primitives, Wavefront **out);
int wavefront_batch(const Wavefront *w);
uint64_t wavefront_capacity(const Wavefront *w);   // rays one pass can hold
size_t wavefront_bytes(const Wavefront *w);        // device memory of the queues
cudaError_t wavefront_grow(Wavefront *w, bool *grown);   // doubles the queue capacities (after a single-sample pass overflowed)
void wavefront_destroy(Wavefront *w);
// n_samples (<= batch) samples of every pixel in one pass over the queues: generate, then per depth level extend / shade /
// shadow / fold, then the child chains fold into their parents. samples[i] = sample index (selects the jitter draws).
// Touches no frame target: the result stays in the per-path accumulators until wavefront_resolve_batch, so a pass whose
// queues overflowed (d_counters->overflow) can be repeated with fewer samples.
// d_tile_base[k] = valid pixels in tiles 0..k-1 (device, 257 entries), roots_per_sample = d_tile_base[256]: camera path slots are deterministic.
cudaError_t wavefront_trace_batch(Wavefront *w, const SceneDev &s, const CameraRec &cam, const RenderParams &rp, const int2 *d_tile_origin,
                                  const uint32_t *d_tile_base, uint32_t roots_per_sample, int n_samples, const int *samples, DeviceCounters *d_counters,
                                  uint64_t *launches, cudaStream_t st);
// Camera paths of the last traced pass -> pixels, samples applied in the order given; avg_n[i] = the sample number used by
// the integer running average (bitmap target). tiles_disjoint: no two path slots of a sample share a pixel.
cudaError_t wavefront_resolve_batch(Wavefront *w, const uint32_t *d_tile_base, int n_tiles, uint32_t roots_per_sample, bool tiles_disjoint, int n_samples,
                                    const int *avg_n, FrameTargets tg, uint64_t *launches, cudaStream_t st);

// Per-ray results through the extend / shadow kernels themselves (test entry; same outputs as launch_trace_rays). n <= queue capacity.
cudaError_t wavefront_trace_rays(Wavefront *w, const SceneDev &s, int64_t n, const float *d_o, const float *d_d, const int32_t *d_okind,
                                 const int32_t *d_oidx, const float *d_sdist, float *d_t, int32_t *d_kind, int32_t *d_idx, int32_t *d_mat,
                                 float *d_normal, DeviceCounters *dc, cudaStream_t st);

// The same, allocation-free: the queues are a file-scope device array of wavefront_static_capacity() rays (mrt_wavefront.cu,
// g_static_scratch), shared by every call on a device. No Wavefront object is needed. n beyond the capacity: cudaErrorInvalidValue.
#ifndef MRT_STATIC_TRACE_MAX_RAYS
#define MRT_STATIC_TRACE_MAX_RAYS 65536
#endif
uint32_t wavefront_static_capacity();
cudaError_t wavefront_trace_rays_static(const SceneDev &s, int64_t n, const float *d_o, const float *d_d, const int32_t *d_okind,
                                        const int32_t *d_oidx, const float *d_sdist, float *d_t, int32_t *d_kind, int32_t *d_idx, int32_t *d_mat,
                                        float *d_normal, DeviceCounters *dc, cudaStream_t st);

// After the stream has been synchronised: adds the device time of every launch since the last call, per kernel class.
cudaError_t wavefront_collect_times(Wavefront *w, WavefrontTimes *times);

}  // namespace mrt
