// Render kernels. Built with -fmad=false (see mrt_math.cuh).
#include "mrt_render.h"
#include "mrt_shade.cuh"

namespace mrt {

// Shadow and secondary rays of the warp -> the frame's counters. The camera rays are not counted here: every path slot of a
// rendered sample holds exactly one, so the frame driver adds samples x path slots on the host (mrt_api.cu); one same-address
// atomicAdd per warp (64 k per 1080p launch, about 0.7 ns each at the L2) was a third of the DepthMap / NoShadows frame time.
MRT_D void flush_counters(const RayCounters &rc, DeviceCounters *dc) {
    const uint32_t sh = __reduce_add_sync(0xffffffffu, rc.shadow);
    const uint32_t se = __reduce_add_sync(0xffffffffu, rc.secondary);
    if ((threadIdx.x & 31) == 0) {
        if (sh) { atomicAdd(&dc->rays_shadow, static_cast<unsigned long long>(sh)); }
        if (se) { atomicAdd(&dc->rays_secondary, static_cast<unsigned long long>(se)); }
    }
}

// Renderer::renderScene's pixel body (Renderer.cpp:130-157) for one sample of every pixel: jitter, camera ray,
// Shader::rayTrace, incrementalAvg. A warp covers an 8x4 pixel patch of one tile so that its rays stay coherent.
template <int SHADER>
__global__ void __launch_bounds__(128) k_render_sample(SceneDev s, CameraRec cam, RenderParams rp, const int2 *__restrict__ tile_origin,
                                                       int sample, int avg_n, FrameTargets tg, DeviceCounters *dc, ReplayArgs rep) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int patch = blockIdx.x * 4 + warp;
    RayCounters rc = {0u, 0u, 0u};
    int overflow = 0;
    uint32_t q = 0;
    int x, y, pix;
    if (tile_pixel(rp, tile_origin, patch, lane, sample, x, y, pix, q)) {
        const float invImgWidth = 1.0f / static_cast<float>(rp.width), invImgHeight = 1.0f / static_cast<float>(rp.height);
        const float pixelWidth = 0.5f / static_cast<float>(rp.width), pixelHeight = 0.5f / static_cast<float>(rp.height);
        const float u = static_cast<float>(x) * invImgWidth, v = static_cast<float>(y) * invImgHeight;
        const float r1 = pixel_draw(s, rp, 2u * q), r2 = pixel_draw(s, rp, 2u * q + 1u);
        const float deviationU = (r1 - 0.5f) * 2.0f * pixelWidth;
        const float deviationV = (r2 - 0.5f) * 2.0f * pixelHeight;
        v3 o, d;
        generate_ray(cam, u, v, deviationU, deviationV, o, d);
        rc.primary = 1u;
        DrawState ds = {q, 0u, rp.seed};
        bool counting = false;
        if (rep.mode != REPLAY_OFF) {
            // position of this pixel-sample in the reference's order within the sample (q counts across samples)
            const uint32_t qi = q - static_cast<uint32_t>(sample) * static_cast<uint32_t>(kNumberOfTiles * rp.bsx * rp.bsy);
            ds.mode = rep.mode;
            if (rep.mode >= REPLAY_COUNT_LIGHTS) { ds.a_next = rep.index_draws[qi]; }
            if (rep.mode == REPLAY_EXACT) {
#pragma unroll
                for (int k = 0; k < kReplayMaxLights; ++k) { if (k < rep.n_lights) { ds.l_next[k] = rep.light_draws[qi * rep.n_lights + k]; } }
            }
            if (rep.mode != REPLAY_EXACT) {
                RayCounters scratch = {0u, 0u, 0u};
                (void) ray_trace<SHADER>(s, rp, o, d, 1, ds, scratch, &overflow);
                if (rep.mode == REPLAY_COUNT_INDEX) { rep.index_draws[qi] = ds.a_next; }
                else {
#pragma unroll
                    for (int k = 0; k < kReplayMaxLights; ++k) { if (k < rep.n_lights) { rep.light_draws[qi * rep.n_lights + k] = ds.l_next[k]; } }
                }
                counting = true; rc.primary = 0u;
            }
        }
        const v3 rgb = counting ? mk3(0.0f, 0.0f, 0.0f) : ray_trace<SHADER>(s, rp, o, d, 1, ds, rc, &overflow);
        // (the counting passes are launched without targets)
        if (tg.bitmap != nullptr) { tg.bitmap[pix] = incremental_avg(rgb, tg.bitmap[pix], avg_n); }
        if (tg.accum != nullptr) {
            float4 a = tg.accum[pix];
            a.x += static_cast<float>(f2u_x86(rgb.x * 255.0f)); a.y += static_cast<float>(f2u_x86(rgb.y * 255.0f));
            a.z += static_cast<float>(f2u_x86(rgb.z * 255.0f)); a.w += 1.0f;
            tg.accum[pix] = a;
        }
    }
    flush_counters(rc, dc);
    if (overflow) { dc->overflow = 1; }
}

// fp32 accumulator -> ABGR8888: the mean of the per-sample integer values, truncated, clamped (Utils.cpp:78-86)
__global__ void k_resolve(const float4 *__restrict__ accum, int32_t *__restrict__ bitmap, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    const float4 a = accum[i];
    if (!(a.w > 0.0f)) { return; }   // pixel never rendered (F8): leave the caller's content
    const uint32_t r = min(static_cast<uint32_t>(a.x / a.w), 255u), g = min(static_cast<uint32_t>(a.y / a.w), 255u), b = min(static_cast<uint32_t>(a.z / a.w), 255u);
    bitmap[i] = static_cast<int32_t>(0xFF000000u | b << 16 | g << 8 | r);
}

__global__ void __launch_bounds__(128) k_trace_rays(SceneDev s, int64_t n, const float *__restrict__ o, const float *__restrict__ d,
                                                    const int32_t *__restrict__ okind, const int32_t *__restrict__ oidx,
                                                    const float *__restrict__ sdist, float *out_t, int32_t *out_kind, int32_t *out_idx,
                                                    int32_t *out_mat, float *out_normal, DeviceCounters *dc) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    const v3 ro = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    const int sk = okind ? okind[i] : HIT_NONE, si = oidx ? oidx[i] : -1;
    int overflow = 0;
    if (sdist != nullptr && sdist[i] >= 0.0f) {
        const bool occluded = trace_shadow(s, ro, rd, sdist[i], sk, si, &overflow);
        out_t[i] = occluded ? 0.0f : sdist[i];
        out_kind[i] = occluded ? 1 : 0; out_idx[i] = -1; out_mat[i] = -1;
    } else {
        HitRec h;
        trace_closest(s, ro, rd, sk, si, h, &overflow);
        out_t[i] = h.t; out_kind[i] = h.kind;
        int idx = h.idx, mat = -1;
        v3 nrm = mk3(0.0f, 1.0f, 0.0f);
        if (h.kind != HIT_NONE) {
            Surface sf;
            resolve_surface(s, ro, rd, h, sf);
            nrm = sf.normal;
            if (h.kind == HIT_PLANE) { idx = s.planes[h.idx].id; mat = s.planes[h.idx].material; }
            else if (h.kind == HIT_SPHERE) { idx = s.spheres[h.idx].id; mat = s.spheres[h.idx].material; }
            else if (h.kind == HIT_TRIANGLE) { mat = __float_as_int(reinterpret_cast<const float4 *>(s.tri_shade + h.idx)[3].w); }
        }
        out_idx[i] = idx; out_mat[i] = mat;
        if (out_normal) { out_normal[3 * i] = nrm.x; out_normal[3 * i + 1] = nrm.y; out_normal[3 * i + 2] = nrm.z; }
    }
    if (overflow) { dc->overflow = 1; }
}

__global__ void k_generate_rays(CameraRec cam, int64_t n, const float *__restrict__ uvdd, float *o, float *d) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    v3 ro, rd;
    generate_ray(cam, uvdd[4 * i], uvdd[4 * i + 1], uvdd[4 * i + 2], uvdd[4 * i + 3], ro, rd);
    o[3 * i] = ro.x; o[3 * i + 1] = ro.y; o[3 * i + 2] = ro.z;
    d[3 * i] = rd.x; d[3 * i + 1] = rd.y; d[3 * i + 2] = rd.z;
}

template <int SHADER>
__global__ void __launch_bounds__(128) k_shade_rays(SceneDev s, RenderParams rp, int64_t n, const float *__restrict__ o,
                                                    const float *__restrict__ d, int depth, float *rgb, DeviceCounters *dc) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    RayCounters rc = {0u, 0u, 0u};
    int overflow = 0;
    DrawState ds = {static_cast<uint32_t>(i), 0u, rp.seed};
    const v3 c = ray_trace<SHADER>(s, rp, mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), depth, ds, rc, &overflow);
    rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    if (overflow) { dc->overflow = 1; }
}

// ---- launchers ------------------------------------------------------------------------------------------------
cudaError_t launch_render_sample(const SceneDev &s, const CameraRec &cam, const RenderParams &rp, const int2 *d_tile_origin,
                                 int sample, int avg_n, FrameTargets tg, DeviceCounters *dc, cudaStream_t st, ReplayArgs rep) {
    const int patches = rp.n_tiles * ((rp.bsx + 7) / 8) * ((rp.bsy + 3) / 4);
    if (patches == 0) { return cudaSuccess; }
    const int blocks = (patches + 3) / 4;
    switch (rp.shader) {
        case 1: k_render_sample<1><<<blocks, 128, 0, st>>>(s, cam, rp, d_tile_origin, sample, avg_n, tg, dc, rep); break;
        case 3: k_render_sample<3><<<blocks, 128, 0, st>>>(s, cam, rp, d_tile_origin, sample, avg_n, tg, dc, rep); break;
        case 4: k_render_sample<4><<<blocks, 128, 0, st>>>(s, cam, rp, d_tile_origin, sample, avg_n, tg, dc, rep); break;
        default: k_render_sample<0><<<blocks, 128, 0, st>>>(s, cam, rp, d_tile_origin, sample, avg_n, tg, dc, rep); break;
    }
    return cudaGetLastError();
}
cudaError_t launch_resolve(const float4 *accum, int32_t *bitmap, int n, cudaStream_t st) {
    k_resolve<<<(n + 255) / 256, 256, 0, st>>>(accum, bitmap, n);
    return cudaGetLastError();
}
cudaError_t launch_trace_rays(const SceneDev &s, int64_t n, const float *o, const float *d, const int32_t *okind, const int32_t *oidx,
                              const float *sdist, float *out_t, int32_t *out_kind, int32_t *out_idx, int32_t *out_mat, float *out_normal,
                              DeviceCounters *dc, cudaStream_t st) {
    if (n == 0) { return cudaSuccess; }
    k_trace_rays<<<static_cast<unsigned>((n + 127) / 128), 128, 0, st>>>(s, n, o, d, okind, oidx, sdist, out_t, out_kind, out_idx, out_mat, out_normal, dc);
    return cudaGetLastError();
}
cudaError_t launch_generate_rays(const CameraRec &cam, int64_t n, const float *uvdd, float *o, float *d, cudaStream_t st) {
    if (n == 0) { return cudaSuccess; }
    k_generate_rays<<<static_cast<unsigned>((n + 127) / 128), 128, 0, st>>>(cam, n, uvdd, o, d);
    return cudaGetLastError();
}
cudaError_t launch_shade_rays(const SceneDev &s, const RenderParams &rp, int64_t n, const float *o, const float *d, int depth,
                              float *rgb, DeviceCounters *dc, cudaStream_t st) {
    if (n == 0) { return cudaSuccess; }
    const unsigned blocks = static_cast<unsigned>((n + 127) / 128);
    switch (rp.shader) {
        case 1: k_shade_rays<1><<<blocks, 128, 0, st>>>(s, rp, n, o, d, depth, rgb, dc); break;
        case 3: k_shade_rays<3><<<blocks, 128, 0, st>>>(s, rp, n, o, d, depth, rgb, dc); break;
        case 4: k_shade_rays<4><<<blocks, 128, 0, st>>>(s, rp, n, o, d, depth, rgb, dc); break;
        default: k_shade_rays<0><<<blocks, 128, 0, st>>>(s, rp, n, o, d, depth, rgb, dc); break;
    }
    return cudaGetLastError();
}
}  // namespace mrt
