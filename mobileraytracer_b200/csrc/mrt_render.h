// Host launchers of the render / traversal kernels (mrt_render.cu).
#pragma once
#include "mrt_types.h"

namespace mrt {

struct FrameTargets {
    int32_t *bitmap;      // device ABGR8888, updated in place with the reference's integer running average, or null
    float4 *accum;        // device fp32 RGBA accumulator (adds trunc(255*c) per channel, .w += 1), or null
};

struct DeviceCounters {   // device memory, zeroed per frame
    unsigned long long rays_primary, rays_shadow, rays_secondary;
    unsigned long long extend_traced, shadow_traced;   // wavefront: queue sizes the traversal kernels consumed
    int overflow;
    int pad;
};

// Exact replay of the reference's draw order (MRT_FLAG_REPLAY_SAMPLERS): arrays indexed by the pixel-sample's position
// in the reference's single-thread order within one sample (tile, then row-major inside the tile).
struct ReplayArgs {
    int32_t mode;            // ReplayMode (mrt_shade.cuh)
    int32_t n_lights;
    uint32_t *index_draws;   // [P] mode 1 writes the count, modes 2/3 read the counter value at the pixel-sample's first draw
    uint32_t *light_draws;   // [P * n_lights] mode 2 writes the counts, mode 3 reads the counter values
};

// One sample of every pixel through the fused megakernel (deterministic shaders: NoShadows, Whitted, DepthMap, Diffuse).
cudaError_t launch_render_sample(const SceneDev &s, const CameraRec &cam, const RenderParams &rp, const int2 *d_tile_origin,
                                 int sample, int sample_number_for_avg, FrameTargets tg, DeviceCounters *d_counters, cudaStream_t st,
                                 ReplayArgs replay = ReplayArgs {0, 0, nullptr, nullptr});
cudaError_t launch_resolve(const float4 *accum, int32_t *bitmap, int n_pixels, cudaStream_t st);

cudaError_t launch_trace_rays(const SceneDev &s, int64_t n, const float *o, const float *d, const int32_t *okind, const int32_t *oidx,
                              const float *shadow_dist, float *out_t, int32_t *out_kind, int32_t *out_idx, int32_t *out_mat,
                              float *out_normal, DeviceCounters *d_counters, cudaStream_t st);
cudaError_t launch_generate_rays(const CameraRec &cam, int64_t n, const float *uvdd, float *o, float *d, cudaStream_t st);
cudaError_t launch_shade_rays(const SceneDev &s, const RenderParams &rp, int64_t n, const float *o, const float *d, int depth,
                              float *rgb, DeviceCounters *d_counters, cudaStream_t st);

}  // namespace mrt
