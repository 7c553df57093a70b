// libmobilert_b200_test.so (include/mobilert_b200_test.h): the device functions of the hot path, one call at a time, for the
// reference's own known-answer vectors (Unit_Testing/TestTriangle.cpp, TestAABB.cpp, TestPlane.cpp). Test infrastructure of
// the product's device code: it compiles the same headers (mrt_math.cuh) as libmobilert_b200.so but is not linked into it,
// so a front-end that links the product library gets none of these symbols. Built with -fmad=false like the product.
#include "../../include/mobilert_b200_test.h"
#include "mrt_grid.h"
#include "mrt_math.cuh"
#include <cerrno>
#include <cstring>
#include <string>
#include <vector>

namespace mrt {

// ---- one-thread test kernels ----------------------------------------------------------------------------------
__global__ void k_test_triangle(const float *t, const float *o, const float *d, float tmax, int self_skip, float *out) {
    // t = the reference's 100-byte record: AC, AB, A, nA, nB, nC, tA, tB, tC, mat
    const v3 AC = mk3(t[0], t[1], t[2]), AB = mk3(t[3], t[4], t[5]), A = mk3(t[6], t[7], t[8]);
    const v3 ro = mk3(o[0], o[1], o[2]), rd = mk3(d[0], d[1], d[2]);
    float tt, u, v;
    out[0] = tmax;
    out[1] = 0.0f; out[2] = 1.0f; out[3] = 0.0f; out[4] = -1.0f; out[5] = -1.0f; out[6] = 0.0f; out[7] = 0.0f; out[8] = 0.0f;
    if (self_skip) { return; }
    if (triangle_intersect(A, AB, AC, ro, rd, tt, u, v) && tt < tmax) {
        const float w = 1.0f - u - v;
        const v3 nA = mk3(t[9], t[10], t[11]), nB = mk3(t[12], t[13], t[14]), nC = mk3(t[15], t[16], t[17]);
        const v3 nrm = normalize((nA * w + nB * u) + nC * v);
        const v3 p = ro + rd * tt;
        out[0] = tt; out[1] = nrm.x; out[2] = nrm.y; out[3] = nrm.z;
        out[4] = (t[18] * w + t[20] * u) + t[22] * v; out[5] = (t[19] * w + t[21] * u) + t[23] * v;
        out[6] = p.x; out[7] = p.y; out[8] = p.z;
    }
}
__global__ void k_test_sphere(const float *s4, const float *o, const float *d, float tmax, float *out) {
    const v3 ro = mk3(o[0], o[1], o[2]), rd = mk3(d[0], d[1], d[2]), c = mk3(s4[0], s4[1], s4[2]);
    float t;
    out[0] = tmax; out[1] = 0.0f; out[2] = 1.0f; out[3] = 0.0f;
    if (sphere_intersect(c, s4[3], ro, rd, t) && t < tmax) {
        const v3 n = normalize((ro + rd * t) - c);
        out[0] = t; out[1] = n.x; out[2] = n.y; out[3] = n.z;
    }
}
__global__ void k_test_plane(const float *p6, const float *o, const float *d, float tmax, float *out) {
    float t;
    out[0] = tmax;
    if (plane_intersect(mk3(p6[0], p6[1], p6[2]), mk3(p6[3], p6[4], p6[5]), mk3(o[0], o[1], o[2]), mk3(d[0], d[1], d[2]), t) && t < tmax) { out[0] = t; }
}
__global__ void k_test_aabb(const float *b, const float *o, const float *d, int32_t *out) {
    out[0] = aabb_intersect_exact(mk3(b[0], b[1], b[2]), mk3(b[3], b[4], b[5]), mk3(o[0], o[1], o[2]), mk3(d[0], d[1], d[2])) ? 1 : 0;
}
__global__ void k_test_incavg(const float *rgb, int32_t avg, int32_t n, int32_t *out) { out[0] = incremental_avg(mk3(rgb[0], rgb[1], rgb[2]), avg, n); }
__global__ void k_test_halton(uint32_t first, int64_t n, uint32_t base, float *out) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) { out[i] = halton_sequence(first + static_cast<uint32_t>(i), base); }
}


static cudaError_t launch_test_triangle(const float *tri25, const float *o, const float *d, float tmax, int self_skip, float *out9, cudaStream_t st) {
    k_test_triangle<<<1, 1, 0, st>>>(tri25, o, d, tmax, self_skip, out9); return cudaGetLastError();
}
static cudaError_t launch_test_sphere(const float *sph4, const float *o, const float *d, float tmax, float *out4, cudaStream_t st) {
    k_test_sphere<<<1, 1, 0, st>>>(sph4, o, d, tmax, out4); return cudaGetLastError();
}
static cudaError_t launch_test_plane(const float *pl6, const float *o, const float *d, float tmax, float *out1, cudaStream_t st) {
    k_test_plane<<<1, 1, 0, st>>>(pl6, o, d, tmax, out1); return cudaGetLastError();
}
static cudaError_t launch_test_aabb(const float *box6, const float *o, const float *d, int32_t *out, cudaStream_t st) {
    k_test_aabb<<<1, 1, 0, st>>>(box6, o, d, out); return cudaGetLastError();
}
static cudaError_t launch_test_incavg(const float *rgb, int32_t avg, int32_t n, int32_t *out, cudaStream_t st) {
    k_test_incavg<<<1, 1, 0, st>>>(rgb, avg, n, out); return cudaGetLastError();
}
static cudaError_t launch_test_halton(uint32_t first, int64_t n, uint32_t base, float *out, cudaStream_t st) {
    if (n == 0) { return cudaSuccess; }
    k_test_halton<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(first, n, base, out); return cudaGetLastError();
}


}  // namespace mrt

using namespace mrt;

namespace {
thread_local std::string g_error;
int fail(int code, const std::string &msg) { g_error = msg; errno = 0; return code; }
struct TmpDev {
    std::vector<void *> ptrs;
    ~TmpDev() { for (void *p : ptrs) { cudaFree(p); } }
    template <typename T> cudaError_t get(T **out, size_t count) {
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, (count ? count : 1) * sizeof(T));
        if (e == cudaSuccess) { ptrs.push_back(p); *out = static_cast<T *>(p); }
        return e;
    }
    template <typename T> cudaError_t put(T **out, const T *src, size_t count, cudaStream_t st) {
        cudaError_t e = get(out, count);
        if (e == cudaSuccess && count) { e = cudaMemcpyAsync(*out, src, count * sizeof(T), cudaMemcpyHostToDevice, st); }
        return e;
    }
};
#define TEST_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { return fail(MRT_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); } } while (0)
}  // namespace

extern "C" {

const char *mrt_test_last_error(void) { return g_error.c_str(); }

int mrt_test_triangle_intersect(const mrt_triangle *tri, const float *o, const float *d, float t_max, int32_t self_skip,
                                float *out_t, float *out8) {
    if (!tri || !o || !d || !out_t) { return fail(MRT_ERR_INVALID, "null argument"); }
    TmpDev tmp;
    float *d_tri, *d_o, *d_d, *d_out;
    TEST_CUDA(tmp.put(&d_tri, reinterpret_cast<const float *>(tri), 25, nullptr));
    TEST_CUDA(tmp.put(&d_o, o, 3, nullptr)); TEST_CUDA(tmp.put(&d_d, d, 3, nullptr)); TEST_CUDA(tmp.get(&d_out, 9));
    TEST_CUDA(launch_test_triangle(d_tri, d_o, d_d, t_max, self_skip, d_out, nullptr));
    float out[9];
    TEST_CUDA(cudaMemcpy(out, d_out, sizeof(out), cudaMemcpyDeviceToHost));
    *out_t = out[0];
    if (out8) { std::memcpy(out8, out + 1, 8 * sizeof(float)); }
    errno = 0;
    return MRT_OK;
}
int mrt_test_sphere_intersect(const mrt_sphere *s, const float *o, const float *d, float t_max, float *out_t, float *out_normal) {
    if (!s || !o || !d || !out_t) { return fail(MRT_ERR_INVALID, "null argument"); }
    TmpDev tmp;
    float *d_s, *d_o, *d_d, *d_out;
    const float s4[4] = {s->center[0], s->center[1], s->center[2], s->sq_radius};
    TEST_CUDA(tmp.put(&d_s, s4, 4, nullptr)); TEST_CUDA(tmp.put(&d_o, o, 3, nullptr)); TEST_CUDA(tmp.put(&d_d, d, 3, nullptr)); TEST_CUDA(tmp.get(&d_out, 4));
    TEST_CUDA(launch_test_sphere(d_s, d_o, d_d, t_max, d_out, nullptr));
    float out[4];
    TEST_CUDA(cudaMemcpy(out, d_out, sizeof(out), cudaMemcpyDeviceToHost));
    *out_t = out[0];
    if (out_normal) { std::memcpy(out_normal, out + 1, 3 * sizeof(float)); }
    errno = 0;
    return MRT_OK;
}
int mrt_test_plane_intersect(const mrt_plane *p, const float *o, const float *d, float t_max, float *out_t) {
    if (!p || !o || !d || !out_t) { return fail(MRT_ERR_INVALID, "null argument"); }
    TmpDev tmp;
    float *d_p, *d_o, *d_d, *d_out;
    const float p6[6] = {p->normal[0], p->normal[1], p->normal[2], p->point[0], p->point[1], p->point[2]};
    TEST_CUDA(tmp.put(&d_p, p6, 6, nullptr)); TEST_CUDA(tmp.put(&d_o, o, 3, nullptr)); TEST_CUDA(tmp.put(&d_d, d, 3, nullptr)); TEST_CUDA(tmp.get(&d_out, 1));
    TEST_CUDA(launch_test_plane(d_p, d_o, d_d, t_max, d_out, nullptr));
    TEST_CUDA(cudaMemcpy(out_t, d_out, sizeof(float), cudaMemcpyDeviceToHost));
    errno = 0;
    return MRT_OK;
}
int mrt_test_aabb_intersect(const float *box_min, const float *box_max, const float *o, const float *d, int32_t *out_hit) {
    if (!box_min || !box_max || !o || !d || !out_hit) { return fail(MRT_ERR_INVALID, "null argument"); }
    TmpDev tmp;
    float *d_b, *d_o, *d_d; int32_t *d_out;
    const float b6[6] = {box_min[0], box_min[1], box_min[2], box_max[0], box_max[1], box_max[2]};
    TEST_CUDA(tmp.put(&d_b, b6, 6, nullptr)); TEST_CUDA(tmp.put(&d_o, o, 3, nullptr)); TEST_CUDA(tmp.put(&d_d, d, 3, nullptr)); TEST_CUDA(tmp.get(&d_out, 1));
    TEST_CUDA(launch_test_aabb(d_b, d_o, d_d, d_out, nullptr));
    TEST_CUDA(cudaMemcpy(out_hit, d_out, sizeof(int32_t), cudaMemcpyDeviceToHost));
    errno = 0;
    return MRT_OK;
}
int mrt_test_incremental_avg(const float *rgb, int32_t avg, int32_t num_sample, int32_t *out) {
    if (!rgb || !out || num_sample <= 0) { return fail(MRT_ERR_INVALID, "bad argument"); }
    TmpDev tmp;
    float *d_rgb; int32_t *d_out;
    TEST_CUDA(tmp.put(&d_rgb, rgb, 3, nullptr)); TEST_CUDA(tmp.get(&d_out, 1));
    TEST_CUDA(launch_test_incavg(d_rgb, avg, num_sample, d_out, nullptr));
    TEST_CUDA(cudaMemcpy(out, d_out, sizeof(int32_t), cudaMemcpyDeviceToHost));
    errno = 0;
    return MRT_OK;
}
int mrt_test_halton(uint32_t first, int64_t n, uint32_t base, float *out) {
    if (!out || n < 0 || base < 2) { return fail(MRT_ERR_INVALID, "bad argument"); }
    TmpDev tmp;
    float *d_out;
    TEST_CUDA(tmp.get(&d_out, static_cast<size_t>(n)));
    TEST_CUDA(launch_test_halton(first, n, base, d_out, nullptr));
    TEST_CUDA(cudaMemcpy(out, d_out, static_cast<size_t>(n) * sizeof(float), cudaMemcpyDeviceToHost));
    errno = 0;
    return MRT_OK;
}


// Host code only (no device needed): the cell lists libmobilert_b200.so builds for MRT_ACC_REGULAR_GRID (mrt_grid.h)
int mrt_test_build_grid(const mrt_scene_desc *scene, int32_t kind, float *out_params12, uint32_t *out_start, int32_t *out_items, int64_t items_cap,
                        int64_t *out_n_items) {
    if (!scene || kind < 0 || kind > 2 || !out_params12 || !out_start || !out_n_items) { return fail(MRT_ERR_INVALID, "null argument or kind outside 0..2"); }
    std::vector<HostBox> boxes;
    GridHost g;
    if (kind == 0) {
        for (int64_t i = 0; i < scene->n_planes; ++i) { boxes.push_back(hb_plane(scene->planes[i].normal, scene->planes[i].point)); }
        build_ref_grid(boxes, [&](int64_t i, const HostBox &cell) { return grid_plane_overlaps(scene->planes[i].normal, scene->planes[i].point, cell); }, g);
    } else if (kind == 1) {
        for (int64_t i = 0; i < scene->n_spheres; ++i) { boxes.push_back(hb_sphere(scene->spheres[i].center, scene->spheres[i].sq_radius)); }
        build_ref_grid(boxes, [&](int64_t i, const HostBox &cell) { return grid_sphere_overlaps(scene->spheres[i].center, scene->spheres[i].sq_radius, cell); }, g);
    } else {
        for (int64_t i = 0; i < scene->n_triangles; ++i) { boxes.push_back(hb_triangle(scene->triangles[i].A, scene->triangles[i].AB, scene->triangles[i].AC)); }
        build_ref_grid(boxes, [&](int64_t i, const HostBox &cell) { return grid_triangle_overlaps(scene->triangles[i].A, scene->triangles[i].AB, scene->triangles[i].AC, cell); }, g);
    }
    for (int k = 0; k < 3; ++k) { out_params12[k] = g.bmin[k]; out_params12[3 + k] = g.bmax[k]; out_params12[6 + k] = g.cellInv[k]; out_params12[9 + k] = g.cellSize[k]; }
    std::memcpy(out_start, g.start.data(), g.start.size() * sizeof(uint32_t));
    *out_n_items = static_cast<int64_t>(g.items.size());
    if (out_items) { std::memcpy(out_items, g.items.data(), static_cast<size_t>(std::min<int64_t>(items_cap, *out_n_items)) * sizeof(int32_t)); }
    errno = 0;
    return MRT_OK;
}

}  // extern "C"
