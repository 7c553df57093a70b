// C-ABI of libmobilert_b200.so (include/mobilert_b200.h): handle, scene upload, frame driver.
// Host logic only; every ray is traced by the CUDA kernels of mrt_render.cu / mrt_wavefront.cu. No CPU fallback.
#include "../../include/mobilert_b200.h"
#include "mrt_build.h"
#include "mrt_grid.h"
#include "mrt_refbvh.h"
#include "mrt_render.h"
#include "mrt_wavefront.h"
#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <random>
#include <string>
#include <vector>

using namespace mrt;

namespace {

thread_local std::string g_create_error;

struct DevMem {
    void *p = nullptr;
    size_t bytes = 0;
    ~DevMem() { release(); }
    void release() { if (p) { cudaFree(p); p = nullptr; bytes = 0; } }
    cudaError_t alloc(size_t n) { release(); bytes = n; return cudaMalloc(&p, n ? n : 16); }
    cudaError_t upload(const void *src, size_t n, cudaStream_t st) {
        cudaError_t e = alloc(n);
        if (e != cudaSuccess || n == 0) { return e; }
        return cudaMemcpyAsync(p, src, n, cudaMemcpyHostToDevice, st);
    }
    template <typename T> T *as() const { return static_cast<T *>(p); }
};

}  // namespace

struct mrt_handle {
    bool naive = false;              // MRT_ACC_NAIVE semantics for planes / spheres / the triangle root box
    bool grid = false;               // MRT_ACC_REGULAR_GRID: the reference's three 32^3 grids instead of the BVHs
    uint32_t roots_per_sample = 0;   // valid pixels of one sample (wavefront path slots)
    bool tiles_disjoint = true;      // no two path slots of a sample share a pixel (false for image sizes that are not multiples of 16)
    uint32_t replay_index_counter = 0;   // MRT_FLAG_REPLAY_SAMPLERS: value of Shader::getLightIndex's counter at the next frame's first draw
    int device = 0;
    cudaStream_t stream = nullptr;       // stream in use (own_stream unless mrt_set_stream installed the caller's)
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::vector<cudaEvent_t> sample_events;
    SceneDev scene {};
    CameraRec camera {};
    RenderParams rp {};
    uint32_t flags = 0;
    Bvh8 bvh {};
    DevMem d_grid_start[3], d_grid_items[3], d_grid_tris;
    DevMem d_tris9, d_tri_shade, d_sphere_nodes, d_spheres, d_plane_nodes, d_planes, d_materials, d_lights;
    DevMem d_tile, d_shader_table, d_static_halton, d_static_mt, d_static_pcg, d_inv_tile, d_tile_base, d_textures, d_texels, d_bitmap, d_accum, d_counters, d_counters_saved, d_pixel_stream;
    std::vector<int> plane_slot_of_id, sphere_slot_of_id;
    int32_t *h_pinned_bitmap = nullptr;
    // MRT_FLAG_PROGRESSIVE: the caller's host bitmap is refreshed while the frame renders (set for the duration of a frame)
    int32_t *progress_host = nullptr;
    bool progress_pinned = false;
    size_t progress_bytes = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_progress = nullptr, ev_copied = nullptr;
    size_t pinned_bytes = 0;
    std::atomic<int> stop {0};
    std::atomic<int> sample {0};
    std::atomic<unsigned long long> total_rays {0};
    mrt_stats stats {};
    std::string error;
    std::mutex mutex;
    Wavefront *wavefront = nullptr;
};

namespace {

int fail(mrt_handle *h, int code, const std::string &msg) {
    if (h) { h->error = msg; } else { g_create_error = msg; }
    errno = 0;
    return code;
}
int cuda_fail(mrt_handle *h, cudaError_t e, const char *what) {
    const int code = (e == cudaErrorMemoryAllocation) ? MRT_ERR_OOM : MRT_ERR_CUDA;
    return fail(h, code, std::string(what) + ": " + cudaGetErrorString(e));
}
#define API_CUDA(h, x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { return cuda_fail((h), e_, #x); } } while (0)

// Camera ctor (MobileRT/Camera.cpp:14-19) + Perspective ctor (Perspective.cpp:8-14, degToRad Camera.cpp:28-31)
// + Orthographic ctor (Orthographic.cpp:7-13)
CameraRec make_camera(const mrt_camera_desc &c) {
    CameraRec r {};
    if (c.kind == MRT_CAMERA_PERSPECTIVE_STATE || c.kind == MRT_CAMERA_ORTHOGRAPHIC_STATE) {   // a constructed camera's members, taken as they are
        r.kind = c.kind == MRT_CAMERA_PERSPECTIVE_STATE ? MRT_CAMERA_PERSPECTIVE : MRT_CAMERA_ORTHOGRAPHIC;
        r.position = mk3(c.position[0], c.position[1], c.position[2]);
        r.direction = mk3(c.look_at[0], c.look_at[1], c.look_at[2]);
        r.up = mk3(c.up[0], c.up[1], c.up[2]);
        r.right = mk3(c.right[0], c.right[1], c.right[2]);
        r.a = c.a; r.b = c.b;
        return r;
    }
    r.kind = c.kind;
    const v3 pos = mk3(c.position[0], c.position[1], c.position[2]);
    const v3 look = mk3(c.look_at[0], c.look_at[1], c.look_at[2]);
    const v3 up = mk3(c.up[0], c.up[1], c.up[2]);
    r.position = pos;
    r.direction = normalize(look - pos);
    r.right = cross(up, r.direction);
    r.up = cross(r.direction, r.right);
    if (c.kind == MRT_CAMERA_PERSPECTIVE) {
        const float pi = 3.14159265358979323846264338327950288f;
        r.a = (c.a * pi) / 180.0f;
        r.b = (c.b * pi) / 180.0f;
    } else {
        r.a = c.a / 2.0f;
        r.b = c.b / 2.0f;
    }
    return r;
}

// fillArrayWithHaltonSeq (MobileRT/Utils/Utils.hpp:209-218) with a caller-visible seed instead of std::random_device
void make_halton_table(std::vector<float> &t, size_t n, uint32_t seed) {
    t.resize(n);
    for (size_t i = 0; i < n; ++i) { t[i] = halton_sequence(static_cast<uint32_t>(i), 2u); }
    std::mt19937 gen(seed);
    std::shuffle(t.begin(), t.end(), gen);
}
// fillArrayWithMersenneTwister (Utils.hpp:230-235)
void make_mt_table(std::vector<float> &t, size_t n, uint32_t seed) {
    t.resize(n);
    std::uniform_real_distribution<float> dist(0.0f, 1.0f);
    std::mt19937 gen(seed);
    for (size_t i = 0; i < n; ++i) { t[i] = dist(gen); }
}

// fillArrayWithPCG (Utils.hpp:246-252): pcg32 (XSH-RR 64/32, pcg_random.hpp) through uniform_real_distribution<float>,
// i.e. one 32-bit draw / 2^32 (libstdc++ generate_canonical<float, 24>), kept below 1; seeded by the caller here
void make_pcg_table(std::vector<float> &t, size_t n, uint32_t seed) {
    t.resize(n);
    const uint64_t mult = 6364136223846793005ull, inc = 1442695040888963407ull;
    uint64_t state = 0ull;
    state = state * mult + inc; state += seed; state = state * mult + inc;
    for (size_t i = 0; i < n; ++i) {
        const uint64_t old = state;
        state = old * mult + inc;
        const uint32_t xs = static_cast<uint32_t>(((old >> 18) ^ old) >> 27), rot = static_cast<uint32_t>(old >> 59);
        const uint32_t r = (xs >> rot) | (xs << ((32u - rot) & 31u));
        float f = static_cast<float>(r) / 4294967296.0f;
        if (f >= 1.0f) { f = std::nextafter(1.0f, 0.0f); }
        t[i] = f;
    }
}

int upload_scene(mrt_handle *h, const mrt_scene_desc &sc) {
    cudaStream_t st = h->stream;
    const int64_t nt = sc.n_triangles, nl = sc.n_lights;
    int64_t n_area = 0;
    for (int64_t i = 0; i < nl; ++i) { if (sc.lights[i].kind == MRT_LIGHT_AREA) { ++n_area; } }
    if (nt + nl > 0x7ffffff0ll) { return fail(h, MRT_ERR_INVALID, "too many primitives"); }
    // BVH8 primitive ids: [0, nt) scene triangles, nt + i = light i (point lights get a degenerate, unhittable triangle)
    const int64_t nprims = nt + (n_area > 0 ? nl : 0);
    std::vector<float> tris9(static_cast<size_t>(nprims) * 9);
    std::vector<TriShade> shade(static_cast<size_t>(nt > 0 ? nt : 1));
    for (int64_t i = 0; i < nt; ++i) {
        const mrt_triangle &t = sc.triangles[i];
        float *o = &tris9[static_cast<size_t>(i) * 9];
        for (int k = 0; k < 3; ++k) { o[k] = t.A[k]; o[3 + k] = t.AB[k]; o[6 + k] = t.AC[k]; }
        TriShade &s = shade[static_cast<size_t>(i)];
        s.s0 = make_float4(t.nA[0], t.nA[1], t.nA[2], t.nB[0]);
        s.s1 = make_float4(t.nB[1], t.nB[2], t.nC[0], t.nC[1]);
        s.s2 = make_float4(t.nC[2], t.tA[0], t.tA[1], t.tB[0]);
        float matbits;
        std::memcpy(&matbits, &t.material, 4);
        s.s3 = make_float4(t.tB[1], t.tC[0], t.tC[1], matbits);
        if (t.material < 0 || t.material >= sc.n_materials) { return fail(h, MRT_ERR_INVALID, "triangle material index out of range"); }
    }
    // root box of the reference's BVH<Triangle>: union of Triangle::getAABB (Triangle.cpp:116-123) over the scene triangles
    float rmin[3] = {0, 0, 0}, rmax[3] = {0, 0, 0};
    for (int64_t i = 0; i < nt; ++i) {
        const mrt_triangle &t = sc.triangles[i];
        for (int k = 0; k < 3; ++k) {
            const float a = t.A[k], b = t.A[k] + t.AB[k], c = t.A[k] + t.AC[k];
            const float lo = std::min(a, std::min(b, c)), hi = std::max(a, std::max(b, c));
            rmin[k] = i == 0 ? lo : std::min(rmin[k], lo); rmax[k] = i == 0 ? hi : std::max(rmax[k], hi);
        }
    }
    h->scene.tri_bmin = mk3(rmin[0], rmin[1], rmin[2]); h->scene.tri_bmax = mk3(rmax[0], rmax[1], rmax[2]);
    std::vector<LightRec> lights(static_cast<size_t>(nl > 0 ? nl : 1));
    for (int64_t i = 0; i < nl; ++i) {
        const mrt_light &l = sc.lights[i];
        LightRec &r = lights[static_cast<size_t>(i)];
        std::memset(&r, 0, sizeof(r));
        r.kind = l.kind;
        for (int k = 0; k < 3; ++k) { r.Le[k] = l.Le[k]; }
        if (l.kind == MRT_LIGHT_AREA) {
            for (int k = 0; k < 3; ++k) { r.A[k] = l.triangle.A[k]; r.AB[k] = l.triangle.AB[k]; r.AC[k] = l.triangle.AC[k]; }
        } else {
            for (int k = 0; k < 3; ++k) { r.A[k] = l.position[k]; }
        }
        if (n_area > 0) {
            float *o = &tris9[static_cast<size_t>(nt + i) * 9];
            for (int k = 0; k < 3; ++k) { o[k] = r.A[k]; o[3 + k] = l.kind == MRT_LIGHT_AREA ? r.AB[k] : 0.0f; o[6 + k] = l.kind == MRT_LIGHT_AREA ? r.AC[k] : 0.0f; }
        }
    }
    std::vector<MaterialRec> mats(static_cast<size_t>(sc.n_materials > 0 ? sc.n_materials : 1));
    for (int64_t i = 0; i < sc.n_materials; ++i) {
        MaterialRec &m = mats[static_cast<size_t>(i)];
        std::memset(&m, 0, sizeof(m));
        for (int k = 0; k < 3; ++k) { m.Le[k] = sc.materials[i].Le[k]; m.Kd[k] = sc.materials[i].Kd[k]; m.Ks[k] = sc.materials[i].Ks[k]; m.Kt[k] = sc.materials[i].Kt[k]; }
        m.ior = sc.materials[i].ior;
        m.texture = -1;
        if (sc.material_texture != nullptr && sc.n_textures > 0) {
            const int32_t tx = sc.material_texture[i];
            if (tx >= sc.n_textures) { return fail(h, MRT_ERR_INVALID, "material texture index out of range"); }
            m.texture = tx < 0 ? -1 : tx;
        }
    }
    // textures: one byte buffer + a record per texture
    std::vector<TextureRec> texrecs(static_cast<size_t>(sc.n_textures > 0 ? sc.n_textures : 1));
    std::vector<uint8_t> texels;
    for (int64_t i = 0; i < sc.n_textures; ++i) {
        const mrt_texture &tx = sc.textures[i];
        if (!tx.pixels || tx.width <= 0 || tx.height <= 0 || tx.channels < 1 || tx.channels > 4) { return fail(h, MRT_ERR_INVALID, "bad texture"); }
        const size_t bytes = static_cast<size_t>(tx.width) * tx.height * tx.channels;
        if (texels.size() + bytes > 0xfffffff0ull) { return fail(h, MRT_ERR_INVALID, "textures too large"); }
        texrecs[static_cast<size_t>(i)] = TextureRec{static_cast<uint32_t>(texels.size()), tx.width, tx.height, tx.channels, static_cast<uint32_t>(bytes)};
        texels.insert(texels.end(), tx.pixels, tx.pixels + bytes);
    }
    texels.resize(texels.size() + 16, 0);   // loadColor reads index + 2 without a bound (Texture.cpp:42-46): keep that read inside the buffer
    // planes and spheres: reference-layout BVHs built on the host
    std::vector<HostBox> boxes;
    std::vector<RefNode> pnodes, snodes;
    std::vector<int> porder, sorder;
    for (int64_t i = 0; i < sc.n_planes; ++i) { boxes.push_back(hb_plane(sc.planes[i].normal, sc.planes[i].point)); }
    // MRT_ACC_NAIVE (Naive<T>::intersect, Accelerators/Naive.hpp:80-91): every plane and sphere is tested in the caller's order
    // and no box decides what a ray can see; expressed as one leaf holding all primitives behind an unbounded box
    auto naive_leaf = [](size_t count, std::vector<RefNode> &nodes, std::vector<int> &order) {
        nodes.clear(); order.clear();
        if (count == 0) { return; }
        nodes.push_back(RefNode{-3.0e38f, -3.0e38f, -3.0e38f, 3.0e38f, 3.0e38f, 3.0e38f, 0, static_cast<int32_t>(count)});
        for (size_t i = 0; i < count; ++i) { order.push_back(static_cast<int>(i)); }
    };
    // MRT_ACC_REGULAR_GRID: caller order as well (the grids' cell lists hold caller indices), no binary tree at all
    const bool flat = h->naive || h->grid;
    GridHost grids[3];
    const auto grid_t0 = std::chrono::steady_clock::now();
    if (flat) { naive_leaf(static_cast<size_t>(sc.n_planes), pnodes, porder); } else { build_ref_bvh(boxes, pnodes, porder); }
    if (h->grid) { build_ref_grid(boxes, [&](int64_t i, const HostBox &cell) { return grid_plane_overlaps(sc.planes[i].normal, sc.planes[i].point, cell); }, grids[0]); }
    boxes.clear();
    for (int64_t i = 0; i < sc.n_spheres; ++i) { boxes.push_back(hb_sphere(sc.spheres[i].center, sc.spheres[i].sq_radius)); }
    if (flat) { naive_leaf(static_cast<size_t>(sc.n_spheres), snodes, sorder); } else { build_ref_bvh(boxes, snodes, sorder); }
    if (h->grid) { build_ref_grid(boxes, [&](int64_t i, const HostBox &cell) { return grid_sphere_overlaps(sc.spheres[i].center, sc.spheres[i].sq_radius, cell); }, grids[1]); }
    std::vector<TriGeom> grid_tris;
    if (h->grid) {
        boxes.clear();
        boxes.reserve(static_cast<size_t>(nt));
        grid_tris.resize(static_cast<size_t>(nt > 0 ? nt : 1));
        for (int64_t i = 0; i < nt; ++i) {
            const mrt_triangle &t = sc.triangles[i];
            boxes.push_back(hb_triangle(t.A, t.AB, t.AC));
            float idbits;
            const int32_t id = static_cast<int32_t>(i);
            std::memcpy(&idbits, &id, 4);
            grid_tris[static_cast<size_t>(i)] = TriGeom{make_float4(t.A[0], t.A[1], t.A[2], t.AB[0]), make_float4(t.AB[1], t.AB[2], t.AC[0], t.AC[1]), make_float4(t.AC[2], idbits, 0.0f, 0.0f)};
        }
        build_ref_grid(boxes, [&](int64_t i, const HostBox &cell) { return grid_triangle_overlaps(sc.triangles[i].A, sc.triangles[i].AB, sc.triangles[i].AC, cell); }, grids[2]);
        boxes.clear();
    }
    const float grid_build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - grid_t0).count();
    std::vector<PlaneRec> planes(porder.size() ? porder.size() : 1);
    h->plane_slot_of_id.assign(porder.size(), 0);
    for (size_t sidx = 0; sidx < porder.size(); ++sidx) {
        const mrt_plane &p = sc.planes[porder[sidx]];
        planes[sidx] = PlaneRec{p.normal[0], p.normal[1], p.normal[2], p.point[0], p.point[1], p.point[2], p.material, porder[sidx]};
        h->plane_slot_of_id[static_cast<size_t>(porder[sidx])] = static_cast<int>(sidx);
        if (p.material < 0 || p.material >= sc.n_materials) { return fail(h, MRT_ERR_INVALID, "plane material index out of range"); }
    }
    std::vector<SphereRec> spheres(sorder.size() ? sorder.size() : 1);
    h->sphere_slot_of_id.assign(sorder.size(), 0);
    for (size_t sidx = 0; sidx < sorder.size(); ++sidx) {
        const mrt_sphere &p = sc.spheres[sorder[sidx]];
        spheres[sidx] = SphereRec{p.center[0], p.center[1], p.center[2], p.sq_radius, p.material, sorder[sidx], 0, 0};
        h->sphere_slot_of_id[static_cast<size_t>(sorder[sidx])] = static_cast<int>(sidx);
        if (p.material < 0 || p.material >= sc.n_materials) { return fail(h, MRT_ERR_INVALID, "sphere material index out of range"); }
    }
    cudaEvent_t e0, e1;
    API_CUDA(h, cudaEventCreate(&e0)); API_CUDA(h, cudaEventCreate(&e1));
    API_CUDA(h, cudaEventRecord(e0, st));
    if (!h->grid) { API_CUDA(h, h->d_tris9.upload(tris9.data(), tris9.size() * sizeof(float), st)); }
    API_CUDA(h, h->d_tri_shade.upload(shade.data(), shade.size() * sizeof(TriShade), st));
    API_CUDA(h, h->d_lights.upload(lights.data(), lights.size() * sizeof(LightRec), st));
    API_CUDA(h, h->d_materials.upload(mats.data(), mats.size() * sizeof(MaterialRec), st));
    API_CUDA(h, h->d_textures.upload(texrecs.data(), texrecs.size() * sizeof(TextureRec), st));
    API_CUDA(h, h->d_texels.upload(texels.data(), texels.size(), st));
    API_CUDA(h, h->d_plane_nodes.upload(pnodes.data(), pnodes.size() * sizeof(RefNode), st));
    API_CUDA(h, h->d_planes.upload(planes.data(), planes.size() * sizeof(PlaneRec), st));
    API_CUDA(h, h->d_sphere_nodes.upload(snodes.data(), snodes.size() * sizeof(RefNode), st));
    API_CUDA(h, h->d_spheres.upload(spheres.data(), spheres.size() * sizeof(SphereRec), st));
    size_t grid_bytes = 0;
    if (h->grid) {
        API_CUDA(h, h->d_grid_tris.upload(grid_tris.data(), grid_tris.size() * sizeof(TriGeom), st));
        grid_bytes += grid_tris.size() * sizeof(TriGeom);
        for (int k = 0; k < 3; ++k) {
            API_CUDA(h, h->d_grid_start[k].upload(grids[k].start.data(), grids[k].start.size() * sizeof(uint32_t), st));
            API_CUDA(h, h->d_grid_items[k].upload(grids[k].items.data(), grids[k].items.size() * sizeof(int32_t), st));
            grid_bytes += grids[k].start.size() * sizeof(uint32_t) + grids[k].items.size() * sizeof(int32_t);
            GridDev &g = h->scene.grids[k];
            g.bmin = mk3(grids[k].bmin[0], grids[k].bmin[1], grids[k].bmin[2]);
            g.cellInv = mk3(grids[k].cellInv[0], grids[k].cellInv[1], grids[k].cellInv[2]);
            g.cellSize = mk3(grids[k].cellSize[0], grids[k].cellSize[1], grids[k].cellSize[2]);
            g.start = h->d_grid_start[k].as<uint32_t>(); g.items = h->d_grid_items[k].as<int32_t>();
        }
    }
    API_CUDA(h, cudaEventRecord(e1, st));
    API_CUDA(h, cudaEventSynchronize(e1));
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    h->stats.upload_ms = ms;
    h->stats.n_primitives = nt + sc.n_spheres + sc.n_planes;
    SceneDev &s = h->scene;
    const size_t records_bytes = shade.size() * sizeof(TriShade) + mats.size() * sizeof(MaterialRec) + lights.size() * sizeof(LightRec);
    if (h->grid) {
        // no BVH8: the grids are the acceleration structure (built on the host above; build_ms is that wall time)
        h->stats.build_ms = grid_build_ms;
        h->stats.bvh_nodes = 0;
        h->stats.bvh_bytes = grid_bytes;
        h->stats.scene_bytes = grid_bytes + records_bytes;
        s.grid = 1;
        s.nodes = nullptr; s.tri_geom = h->d_grid_tris.as<TriGeom>();
        s.n_bvh_prims = static_cast<int32_t>(nt);
    } else {
        // acceleration structure over triangles + light triangles, on the GPU
        // binned-SAH hierarchy by default; MRT_B200_BVH=lbvh selects the Morton-order hierarchy (build-speed / quality experiments)
        const char *bvh_env = std::getenv("MRT_B200_BVH");
        const char *leaf_env = std::getenv("MRT_B200_LEAF");   // triangles per leaf slot, 1..7 (tuning experiments)
        const uint32_t leaf = leaf_env ? static_cast<uint32_t>(std::atoi(leaf_env)) & 7u : 4u;
        const char *fat_env = std::getenv("MRT_B200_FATLEAF");   // 1: never split a subtree that fits one leaf slot (tuning experiments)
        const uint32_t build_flags = ((bvh_env && std::strcmp(bvh_env, "lbvh") == 0) ? 0u : 1u) | (leaf << 8) | ((fat_env && std::atoi(fat_env)) ? 2u : 0u);
        API_CUDA(h, build_bvh8(h->d_tris9.as<float>(), static_cast<int>(nprims), build_flags, st, &h->bvh));
        h->stats.build_ms = h->bvh.build_ms;
        h->stats.bvh_nodes = h->bvh.n_nodes;
        h->stats.bvh_bytes = h->bvh.bytes();
        h->stats.scene_bytes = h->bvh.bytes() + records_bytes;
        h->d_tris9.release();   // geometry now lives in the BVH8 leaf array
        s.grid = 0;
        s.nodes = h->bvh.nodes; s.tri_geom = h->bvh.tri_geom;
        s.n_bvh_prims = static_cast<int32_t>(nprims);
    }
    {
        const char *tp = std::getenv("MRT_B200_POSTPONE"), *tr = std::getenv("MRT_B200_REFILL");   // tuning experiments only
        s.tune_postpone = tp ? std::atoi(tp) : 12;
        s.tune_refill = tr ? std::atoi(tr) : 26;
        s.prmt_one = 0x3F800000u;
        s.naive = (h->naive && !h->grid) ? 1 : 0;
        const char *tw = std::getenv("MRT_B200_WAIT");
        s.tune_wait = tw ? std::atoi(tw) : 20;
        const char *twa = std::getenv("MRT_B200_WAIT_ANY");
        s.tune_wait_any = twa ? std::atoi(twa) : (tw ? std::atoi(tw) : 24);
    }
    s.tri_shade = h->d_tri_shade.as<TriShade>();
    s.n_triangles = static_cast<int32_t>(nt);
    s.sphere_nodes = h->d_sphere_nodes.as<RefNode>(); s.spheres = h->d_spheres.as<SphereRec>(); s.n_spheres = static_cast<int32_t>(sc.n_spheres);
    s.plane_nodes = h->d_plane_nodes.as<RefNode>(); s.planes = h->d_planes.as<PlaneRec>(); s.n_planes = static_cast<int32_t>(sc.n_planes);
    s.materials = h->d_materials.as<MaterialRec>(); s.n_materials = static_cast<int32_t>(sc.n_materials);
    s.textures = h->d_textures.as<TextureRec>(); s.texels = h->d_texels.as<uint8_t>(); s.n_textures = static_cast<int32_t>(sc.n_textures);
    s.lights = h->d_lights.as<LightRec>(); s.n_lights = static_cast<int32_t>(nl);
    return MRT_OK;
}

// true when the 2^20 values are exactly the set {k / 2^20 : 0 <= k < 2^20}, each once: the radical inverses in base 2 of
// 0 .. 2^20 - 1 (Utils.cpp:43-53) in any order, i.e. any shuffle of fillArrayWithHaltonSeq's table
bool is_permutation_of_dyadic_grid(const float *t) {
    std::vector<bool> seen(kArraySize, false);
    for (uint32_t i = 0; i < kArraySize; ++i) {
        const float v = t[i] * static_cast<float>(kArraySize);
        if (!(v >= 0.0f) || v >= static_cast<float>(kArraySize)) { return false; }
        const uint32_t k = static_cast<uint32_t>(v);
        if (static_cast<float>(k) != v || seen[k]) { return false; }
        seen[k] = true;
    }
    return true;
}

int upload_tables(mrt_handle *h, const mrt_render_desc &rd, const mrt_sampler_tables *tb) {
    cudaStream_t st = h->stream;
    std::vector<float> tile, shader, sh, smt, spcg;
    const float *p_tile = tb ? tb->tile_order : nullptr, *p_shader = tb ? tb->shader_table : nullptr;
    const float *p_sh = tb ? tb->static_halton : nullptr, *p_smt = tb ? tb->static_mt : nullptr;
    if (!p_tile) { make_halton_table(tile, kNumberOfTiles, rd.seed * 4u + 0u); p_tile = tile.data(); }
    if (!p_shader) { make_halton_table(shader, kArraySize, rd.seed * 4u + 1u); p_shader = shader.data(); }
    if (!p_sh) { make_halton_table(sh, kArraySize, rd.seed * 4u + 2u); p_sh = sh.data(); }
    const bool need_mt = rd.pixel_sampler == MRT_SAMPLER_STATIC_MT;
    if (need_mt && !p_smt) { make_mt_table(smt, kArraySize, rd.seed * 4u + 3u); p_smt = smt.data(); }
    const float *p_pcg = tb ? tb->static_pcg : nullptr;
    const bool need_pcg = rd.pixel_sampler == MRT_SAMPLER_STATIC_PCG || rd.light_sampler == MRT_SAMPLER_STATIC_PCG;
    if (need_pcg && !p_pcg) { make_pcg_table(spcg, kArraySize, rd.seed * 4u + 5u); p_pcg = spcg.data(); }
    // Renderer::renderScene's tile -> first pixel mapping (Renderer.cpp:125-133), integer for integer
    std::vector<int2> origin(kNumberOfTiles, make_int2(0, 0));
    std::vector<uint32_t> tile_base(kNumberOfTiles + 1, 0u);   // valid pixels before each tile this handle renders (the wavefront's deterministic path slots)
    const int n_sel = h->rp.n_tiles;
    {
        const int w = rd.width, hgt = rd.height, bsx = h->rp.bsx, bsy = h->rp.bsy;
        if (bsx > 0 && bsy > 0) {
            const int domainSize = (w / bsx) * (hgt / bsy), resolution = w * hgt;
            for (int k = 0; k < kNumberOfTiles; ++k) {
                if (!(p_tile[k] >= 0.0f && p_tile[k] < 1.0f)) { return fail(h, MRT_ERR_INVALID, "tile_order values must lie in [0, 1)"); }
                const int roundBlock = static_cast<int>(::roundf(p_tile[k] * static_cast<float>(domainSize)));
                const int pixel = static_cast<int>((static_cast<long long>(roundBlock) * bsx) % resolution);
                const int startY = ((pixel / w) * bsy) % hgt;
                origin[static_cast<size_t>(k)] = make_int2(pixel % w, startY);   // startX = (pixel + y * w) % w
            }
            for (int j = 0; j < n_sel; ++j) {
                // a pixel index past the end of the image is dropped (tile_pixel, mrt_shade.cuh); those are the last ones of the tile
                const size_t k = static_cast<size_t>(h->rp.tile_begin + j * h->rp.tile_stride);
                uint32_t valid = 0;
                for (int ly = 0; ly < bsy; ++ly) {
                    const long long first = static_cast<long long>(origin[k].y + ly) * w + origin[k].x;
                    const long long room = static_cast<long long>(resolution) - first;
                    valid += static_cast<uint32_t>(room <= 0 ? 0 : (room < bsx ? room : bsx));
                }
                tile_base[static_cast<size_t>(j) + 1] = tile_base[static_cast<size_t>(j)] + valid;
            }
        }
    }
    API_CUDA(h, h->d_tile.upload(p_tile, kNumberOfTiles * sizeof(float), st));
    API_CUDA(h, h->d_shader_table.upload(p_shader, kArraySize * sizeof(float), st));
    API_CUDA(h, h->d_static_halton.upload(p_sh, kArraySize * sizeof(float), st));
    if (need_mt) { API_CUDA(h, h->d_static_mt.upload(p_smt, kArraySize * sizeof(float), st)); }
    if (need_pcg) { API_CUDA(h, h->d_static_pcg.upload(p_pcg, kArraySize * sizeof(float), st)); }
    API_CUDA(h, h->d_inv_tile.upload(origin.data(), origin.size() * sizeof(int2), st));
    API_CUDA(h, h->d_tile_base.upload(tile_base.data(), tile_base.size() * sizeof(uint32_t), st));
    h->roots_per_sample = tile_base[static_cast<size_t>(n_sel)];
    {   // the wavefront's parallel resolve needs every pixel to belong to at most one path slot of a sample
        bool disjoint = rd.width % 16 == 0 && rd.height % 16 == 0;
        std::vector<long long> firsts;
        for (int k = 0; k < kNumberOfTiles; ++k) { firsts.push_back(static_cast<long long>(origin[static_cast<size_t>(k)].y) * rd.width + origin[static_cast<size_t>(k)].x); }
        std::sort(firsts.begin(), firsts.end());
        if (std::adjacent_find(firsts.begin(), firsts.end()) != firsts.end()) { disjoint = false; }   // a caller's tile table may repeat a tile
        h->tiles_disjoint = disjoint;
    }
    API_CUDA(h, cudaStreamSynchronize(st));
    SceneDev &s = h->scene;
    s.tile_order = h->d_tile.as<float>(); s.shader_table = h->d_shader_table.as<float>(); s.static_halton = h->d_static_halton.as<float>();
    int ps = rd.pixel_sampler;
    if (ps == MRT_SAMPLER_AUTO) { ps = rd.samples_pixel > 1 ? MRT_SAMPLER_STATIC_HALTON : MRT_SAMPLER_CONSTANT; }  // C_wrapper.cpp:144-148
    h->rp.pixel_sampler = ps;
    {   // Sampler::Sampler (MobileRT/Sampler.cpp:14-18): both factors divide by width / sqrt(NumberOfTiles)
        const uint32_t w = static_cast<uint32_t>(rd.width), hgt = static_cast<uint32_t>(rd.height), per = w / 16u;
        h->rp.sampler_domain = per > 0u ? (w / per) * (hgt / per) : 0u;
        h->rp.pixel_stream_n = 1u;
    }
    s.pixel_table = ps == MRT_SAMPLER_STATIC_HALTON ? s.static_halton
                  : ps == MRT_SAMPLER_STATIC_MT ? h->d_static_mt.as<float>()
                  : ps == MRT_SAMPLER_STATIC_PCG ? h->d_static_pcg.as<float>() : nullptr;
    // from here on `static_halton` is the table of the area lights' and the russian roulette's samplers
    s.shader_table_is_perm = is_permutation_of_dyadic_grid(p_shader) ? 1 : 0;
    s.static_halton_is_perm = is_permutation_of_dyadic_grid(p_sh) ? 1 : 0;
    if (std::getenv("MRT_B200_TABLE_GATHER")) { s.shader_table_is_perm = 0; s.static_halton_is_perm = 0; }   // A/B: always read the tables
    if (rd.light_sampler == MRT_SAMPLER_STATIC_PCG) { s.static_halton = h->d_static_pcg.as<float>(); s.static_halton_is_perm = 0; }
    return MRT_OK;
}

}  // namespace

extern "C" {

int mrt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
    errno = 0;
    return n;
}
const char *mrt_version(void) { return "mobilert_b200 0.1 (sm_100a)"; }

const char *mrt_last_error(mrt_handle *h) { return h ? h->error.c_str() : g_create_error.c_str(); }

void mrt_destroy(mrt_handle *h) {
    if (!h) { return; }
    cudaSetDevice(h->device);
    if (h->stream) { cudaStreamSynchronize(h->stream); }
    if (h->wavefront) { wavefront_destroy(h->wavefront); }
    free_bvh8(&h->bvh);
    for (cudaEvent_t e : h->sample_events) { cudaEventDestroy(e); }
    if (h->ev0) { cudaEventDestroy(h->ev0); }
    if (h->ev1) { cudaEventDestroy(h->ev1); }
    if (h->copy_stream) { cudaStreamSynchronize(h->copy_stream); cudaStreamDestroy(h->copy_stream); }
    if (h->ev_progress) { cudaEventDestroy(h->ev_progress); }
    if (h->ev_copied) { cudaEventDestroy(h->ev_copied); }
    if (h->h_pinned_bitmap) { cudaFreeHost(h->h_pinned_bitmap); }
    cudaStream_t st = h->own_stream;
    delete h;   // DevMem destructors free the device arrays
    if (st) { cudaStreamDestroy(st); }
    errno = 0;
}

int mrt_create(const mrt_scene_desc *scene, const mrt_camera_desc *camera, const mrt_render_desc *render,
               const mrt_sampler_tables *tables, mrt_handle **out) {
    if (!scene || !camera || !render || !out) { return fail(nullptr, MRT_ERR_INVALID, "null argument"); }
    *out = nullptr;
    if (render->width <= 0 || render->height <= 0 || render->samples_pixel < 0 || render->samples_light < 0) { return fail(nullptr, MRT_ERR_INVALID, "bad render size / sample counts"); }
    if (render->shader < 0 || render->shader > 4) { return fail(nullptr, MRT_ERR_INVALID, "unknown shader"); }
    if (render->accelerator != MRT_ACC_B200 && render->accelerator != MRT_ACC_BVH && render->accelerator != MRT_ACC_NAIVE && render->accelerator != MRT_ACC_REGULAR_GRID) {
        return fail(nullptr, MRT_ERR_INVALID, "accelerator must be MRT_ACC_B200 / MRT_ACC_BVH (BVH semantics), MRT_ACC_NAIVE or MRT_ACC_REGULAR_GRID (the reference's visibility rules for those)");
    }
    if (scene->n_triangles < 0 || scene->n_spheres < 0 || scene->n_planes < 0 || scene->n_materials < 0 || scene->n_lights < 0) { return fail(nullptr, MRT_ERR_INVALID, "negative count"); }
    if ((scene->n_triangles && !scene->triangles) || (scene->n_spheres && !scene->spheres) || (scene->n_planes && !scene->planes) ||
        (scene->n_materials && !scene->materials) || (scene->n_lights && !scene->lights)) { return fail(nullptr, MRT_ERR_INVALID, "null array with non-zero count"); }
    if (render->pixel_sampler < MRT_SAMPLER_AUTO || render->pixel_sampler > MRT_SAMPLER_STREAM) { return fail(nullptr, MRT_ERR_INVALID, "unknown pixel_sampler"); }
    if (render->light_sampler != MRT_SAMPLER_AUTO && render->light_sampler != MRT_SAMPLER_STATIC_HALTON && render->light_sampler != MRT_SAMPLER_STATIC_PCG) {
        return fail(nullptr, MRT_ERR_INVALID, "light_sampler must be MRT_SAMPLER_AUTO, _STATIC_HALTON or _STATIC_PCG");
    }
    if (render->tile_stride > 1 && (render->tile_begin < 0 || render->tile_begin >= render->tile_stride || render->tile_stride > kNumberOfTiles)) {
        return fail(nullptr, MRT_ERR_INVALID, "tile partition: need 0 <= tile_begin < tile_stride <= 256");
    }
    if (render->tile_stride > 1 && (render->flags & MRT_FLAG_REPLAY_SAMPLERS)) { return fail(nullptr, MRT_ERR_INVALID, "MRT_FLAG_REPLAY_SAMPLERS needs every tile of the frame on this handle"); }
    if (render->flags & MRT_FLAG_REPLAY_SAMPLERS) {
        if (render->shader == MRT_SHADER_PATHTRACER) { return fail(nullptr, MRT_ERR_INVALID, "MRT_FLAG_REPLAY_SAMPLERS: the PathTracer's ray tree depends on the draws themselves; only NoShadows / Whitted can be replayed"); }
        if (scene->n_lights > kReplayMaxLights) { return fail(nullptr, MRT_ERR_INVALID, "MRT_FLAG_REPLAY_SAMPLERS supports up to 8 lights"); }
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(nullptr, MRT_ERR_CUDA, "no CUDA device: this library has no CPU fallback"); }
    mrt_handle *h = new (std::nothrow) mrt_handle;
    if (!h) { return fail(nullptr, MRT_ERR_OOM, "host allocation failed"); }
    int rc = MRT_OK;
    do {
        if (render->device >= 0) { h->device = render->device; } else { cudaGetDevice(&h->device); }
        cudaError_t e = cudaSetDevice(h->device);
        if (e == cudaSuccess) { e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking); h->stream = h->own_stream; }
        if (e == cudaSuccess) { e = cudaEventCreate(&h->ev0); }
        if (e == cudaSuccess) { e = cudaEventCreate(&h->ev1); }
        if (e != cudaSuccess) { rc = cuda_fail(h, e, "device setup"); break; }
        h->flags = render->flags;
        h->naive = render->accelerator == MRT_ACC_NAIVE;
        h->grid = render->accelerator == MRT_ACC_REGULAR_GRID;
        h->rp.width = render->width; h->rp.height = render->height; h->rp.spp = render->samples_pixel; h->rp.spl = render->samples_light;
        h->rp.shader = render->shader; h->rp.seed = render->seed;
        h->rp.bsx = render->width / 16; h->rp.bsy = render->height / 16;   // Renderer.cpp:33-34
        h->rp.tile_stride = render->tile_stride > 1 ? render->tile_stride : 1;
        h->rp.tile_begin = render->tile_stride > 1 ? render->tile_begin : 0;
        h->rp.n_tiles = (kNumberOfTiles - h->rp.tile_begin + h->rp.tile_stride - 1) / h->rp.tile_stride;
        h->rp.depth_max_point = mk3(render->depth_max_point[0], render->depth_max_point[1], render->depth_max_point[2]);
        h->camera = make_camera(*camera);
        if ((rc = upload_scene(h, *scene)) != MRT_OK) { break; }
        if ((rc = upload_tables(h, *render, tables)) != MRT_OK) { break; }
        const size_t npix = static_cast<size_t>(render->width) * render->height;
        if ((e = h->d_bitmap.alloc(npix * sizeof(int32_t))) != cudaSuccess) { rc = cuda_fail(h, e, "bitmap"); break; }
        if ((e = h->d_accum.alloc(npix * sizeof(float4))) != cudaSuccess) { rc = cuda_fail(h, e, "accumulator"); break; }
        if ((e = h->d_counters.alloc(sizeof(DeviceCounters))) != cudaSuccess) { rc = cuda_fail(h, e, "counters"); break; }
        if ((e = h->d_counters_saved.alloc(sizeof(DeviceCounters))) != cudaSuccess) { rc = cuda_fail(h, e, "counters"); break; }
        if ((e = cudaMemsetAsync(h->d_bitmap.p, 0, npix * sizeof(int32_t), h->stream)) != cudaSuccess) { rc = cuda_fail(h, e, "memset"); break; }
        if ((e = cudaMallocHost(reinterpret_cast<void **>(&h->h_pinned_bitmap), npix * sizeof(int32_t))) != cudaSuccess) { rc = cuda_fail(h, e, "pinned bitmap"); break; }
        h->pinned_bytes = npix * sizeof(int32_t);
        if (render->shader == MRT_SHADER_PATHTRACER) {
            const char *sk = std::getenv("MRT_B200_SORT_KEY");   // diagnostic: "primitive" bins by primitive id instead of material
            const int sort_mode = (render->flags & MRT_FLAG_SORT_BY_MATERIAL) ? ((sk && std::strcmp(sk, "primitive") == 0) ? 2 : 1) : 0;
            const char *be = std::getenv("MRT_B200_BATCH");   // samples traced together per wavefront pass
            // default: as few equally large passes as keep the queues (about 390 + 100 * samples_light bytes per pixel and sample) below
            // 40 GB. Every pass ends with the persistent kernels of each level draining, so fewer, larger passes are faster: the 64-spp
            // frame of config 4 takes 182.8 ms in four passes of 16, 181.1 in two of 32 (33 GB), 180.3 in one of 64 (65 GB).
            const double per_sample = (390.0 + 100.0 * std::max(1, render->samples_light)) * static_cast<double>(std::max<uint32_t>(h->roots_per_sample, 1u));
            const int cap = static_cast<int>(std::max(1.0, std::min(static_cast<double>(kWfMaxBatch), 40.0e9 / std::max(per_sample, 1.0))));
            const int spp = std::max(1, render->samples_pixel);
            const int passes = (spp + cap - 1) / cap;
            const int balanced = (spp + passes - 1) / passes;
            int batch = std::max(1, std::min(be ? std::atoi(be) : balanced, spp));
            // a device that other handles or processes have filled: smaller passes instead of no renderer
            while ((e = wavefront_create(h->roots_per_sample, batch, render->samples_light, sort_mode, static_cast<uint32_t>(h->scene.n_bvh_prims), &h->wavefront)) == cudaErrorMemoryAllocation && batch > 1) {
                cudaGetLastError();
                wavefront_destroy(h->wavefront); h->wavefront = nullptr;   // frees what the failed attempt had allocated
                batch = (batch + 1) / 2;
            }
            if (e != cudaSuccess) { rc = cuda_fail(h, e, "wavefront queues"); break; }
        }
        if ((e = cudaStreamSynchronize(h->stream)) != cudaSuccess) { rc = cuda_fail(h, e, "create sync"); break; }
    } while (false);
    if (rc != MRT_OK) { g_create_error = h->error; mrt_destroy(h); errno = 0; return rc; }
    *out = h;
    errno = 0;
    return MRT_OK;
}

int mrt_set_camera(mrt_handle *h, const mrt_camera_desc *camera) {
    if (!h || !camera) { return fail(h, MRT_ERR_INVALID, "null argument"); }
    std::lock_guard<std::mutex> lock(h->mutex);
    h->camera = make_camera(*camera);
    errno = 0;
    return MRT_OK;
}

int mrt_set_pixel_stream(mrt_handle *h, const float *draws, int64_t n) {
    if (!h || !draws || n < 1 || n > 0x7fffffffll) { return fail(h, MRT_ERR_INVALID, "bad argument"); }
    if (h->rp.pixel_sampler != MRT_SAMPLER_STREAM) { return fail(h, MRT_ERR_INVALID, "the handle was not created with MRT_SAMPLER_STREAM"); }
    std::lock_guard<std::mutex> lock(h->mutex);
    API_CUDA(h, cudaSetDevice(h->device));
    API_CUDA(h, cudaStreamSynchronize(h->stream));
    API_CUDA(h, h->d_pixel_stream.upload(draws, static_cast<size_t>(n) * sizeof(float), h->stream));
    API_CUDA(h, cudaStreamSynchronize(h->stream));
    h->scene.pixel_table = h->d_pixel_stream.as<float>();
    h->rp.pixel_stream_n = static_cast<uint32_t>(n);
    errno = 0;
    return MRT_OK;
}

int mrt_set_stream(mrt_handle *h, void *cuda_stream) {
    if (!h) { return MRT_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(h->mutex);
    cudaStreamSynchronize(h->stream);
    // NULL restores the handle's own stream; the legacy default stream is named by cudaStreamLegacy ((cudaStream_t)0x1)
    h->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : h->own_stream;
    errno = 0;
    return MRT_OK;
}

int mrt_set_replay_counter(mrt_handle *h, uint32_t light_index_counter) {
    if (!h) { return MRT_ERR_INVALID; }
    h->replay_index_counter = light_index_counter; errno = 0; return MRT_OK;
}
int mrt_get_replay_counter(mrt_handle *h, uint32_t *light_index_counter) {
    if (!h || !light_index_counter) { return MRT_ERR_INVALID; }
    *light_index_counter = h->replay_index_counter; errno = 0; return MRT_OK;
}
int mrt_stop(mrt_handle *h) { if (!h) { return MRT_ERR_INVALID; } h->stop.store(1); errno = 0; return MRT_OK; }
int mrt_get_sample(mrt_handle *h, int32_t *sample) { if (!h || !sample) { return MRT_ERR_INVALID; } *sample = h->sample.load(); errno = 0; return MRT_OK; }
int mrt_get_total_rays(mrt_handle *h, uint64_t *rays) { if (!h || !rays) { return MRT_ERR_INVALID; } *rays = h->total_rays.load(); errno = 0; return MRT_OK; }
int mrt_get_stats(mrt_handle *h, mrt_stats *stats) {
    if (!h || !stats) { return MRT_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(h->mutex);
    *stats = h->stats;
    errno = 0;
    return MRT_OK;
}

}  // extern "C"

namespace {

// Frame driver shared by the three render entry points. Exactly one of bitmap / accum is the target:
//   bitmap: the reference's in-place integer running average, samples applied in order (Renderer.cpp:117-157)
//   accum : fp32 sums of the selected samples (multi-GPU partitions), resolved later
// The reference's worker threads write every finished pixel straight into the caller's bitmap, which the UI thread shows
// while the frame renders (DrawView / MainRenderer on Android, mainwindow.cpp:45-66 on Qt). Here the bitmap lives on the
// device; with MRT_FLAG_PROGRESSIVE a second stream copies it to the caller's buffer after every sample (or wavefront
// pass), skipping a refresh while the previous one is still in flight.
void CUDART_CB progress_to_caller(void *p) {
    mrt_handle *h = static_cast<mrt_handle *>(p);
    if (h->progress_host) { std::memcpy(h->progress_host, h->h_pinned_bitmap, h->progress_bytes); }
}
cudaError_t progress_refresh(mrt_handle *h, const int32_t *d_bitmap) {
    if (!h->progress_host || !d_bitmap) { return cudaSuccess; }
    if (cudaEventQuery(h->ev_copied) != cudaSuccess) { cudaGetLastError(); return cudaSuccess; }
    cudaError_t e = cudaEventRecord(h->ev_progress, h->stream);
    if (e == cudaSuccess) { e = cudaStreamWaitEvent(h->copy_stream, h->ev_progress, 0); }
    if (e == cudaSuccess) { e = cudaMemcpyAsync(h->progress_pinned ? h->progress_host : h->h_pinned_bitmap, d_bitmap, h->progress_bytes, cudaMemcpyDeviceToHost, h->copy_stream); }
    if (e == cudaSuccess && !h->progress_pinned) { e = cudaLaunchHostFunc(h->copy_stream, progress_to_caller, h); }
    if (e == cudaSuccess) { e = cudaEventRecord(h->ev_copied, h->copy_stream); }
    return e;
}

int render_samples(mrt_handle *h, int32_t *d_bitmap, float4 *d_accum, int sample_begin, int sample_stride) {
    cudaStream_t st = h->stream;
    API_CUDA(h, cudaSetDevice(h->device));
    if (h->rp.pixel_sampler == MRT_SAMPLER_STREAM && h->d_pixel_stream.p == nullptr) { return fail(h, MRT_ERR_INVALID, "MRT_SAMPLER_STREAM: call mrt_set_pixel_stream before rendering"); }
    h->stop.store(0);
    h->sample.store(0);
    API_CUDA(h, cudaMemsetAsync(h->d_counters.p, 0, sizeof(DeviceCounters), st));
    API_CUDA(h, cudaEventRecord(h->ev0, st));
    FrameTargets tg {d_bitmap, d_accum};
    WavefrontTimes wt {};
    uint64_t launches = 0;
    uint64_t samples_rendered = 0;   // megakernel path: samples launched (one camera ray per path slot each, Ray.cpp:11)
    const int spp = h->rp.spp;
    h->stats.pass_retries = 0;
    bool stopped = false;
    if (h->rp.shader == MRT_SHADER_PATHTRACER) {
        // Passes of up to `batch` samples. A pass leaves its result in the per-path accumulators and only then is applied
        // to the target, so a pass whose queues overflowed (a hit may emit up to three continuation rays and
        // samples_light shadow rays; the reference accepts any material set and samplesLight) is repeated with half the
        // samples, and a single sample with doubled queues, instead of failing the frame.
        std::vector<int> todo_s, todo_n;
        for (int s = sample_begin, k = 0; s < spp; s += sample_stride, ++k) { todo_s.push_back(s); todo_n.push_back(d_bitmap ? s + 1 : k + 1); }
        int batch = wavefront_batch(h->wavefront);
        DeviceCounters *dc = h->d_counters.as<DeviceCounters>();
        for (size_t pos = 0; pos < todo_s.size();) {
            if (h->stop.load()) { stopped = true; break; }
            const int nb = static_cast<int>(std::min<size_t>(static_cast<size_t>(batch), todo_s.size() - pos));
            API_CUDA(h, cudaMemcpyAsync(h->d_counters_saved.p, dc, sizeof(DeviceCounters), cudaMemcpyDeviceToDevice, st));
            API_CUDA(h, wavefront_trace_batch(h->wavefront, h->scene, h->camera, h->rp, h->d_inv_tile.as<int2>(), h->d_tile_base.as<uint32_t>(), h->roots_per_sample, nb,
                                              todo_s.data() + pos, dc, &launches, st));
            int overflow = 0;
            API_CUDA(h, cudaMemcpyAsync(&overflow, &dc->overflow, sizeof(int), cudaMemcpyDeviceToHost, st));
            API_CUDA(h, cudaStreamSynchronize(st));
            if (overflow & 1) {   // bit 0: a ray / shadow / chain queue was full (bit 1: traversal stack, which no retry cures)
                API_CUDA(h, cudaMemcpyAsync(dc, h->d_counters_saved.p, sizeof(DeviceCounters), cudaMemcpyDeviceToDevice, st));
                ++h->stats.pass_retries;
                if (nb > 1) { batch = std::max(1, nb / 2); continue; }
                bool grown = false;
                API_CUDA(h, wavefront_grow(h->wavefront, &grown));
                if (!grown) { return fail(h, MRT_ERR_OVERFLOW, "ray queues overflowed at the largest size this library allocates"); }
                continue;
            }
            if (overflow) { break; }   // reported below
            API_CUDA(h, wavefront_resolve_batch(h->wavefront, h->d_tile_base.as<uint32_t>(), h->rp.n_tiles, h->roots_per_sample, h->tiles_disjoint, nb, todo_n.data() + pos, tg, &launches, st));
            pos += static_cast<size_t>(nb);
            h->sample.store(todo_s[pos - 1] + 1);
            API_CUDA(h, progress_refresh(h, d_bitmap));
        }
    } else {
        const size_t depth = 4;   // samples in flight, so that mrt_stop() takes effect within a few samples
        int k = 0;
        // Exact replay (MRT_FLAG_REPLAY_SAMPLERS). The ray tree of NoShadows / Whitted does not depend on the draws, so
        // two counting passes per sample find how many draws of each counter precede every pixel-sample in the
        // reference's single-thread order: Shader::getLightIndex's counter (function-static, Shader.cpp:224: it runs on
        // across frames and renderers) and each AreaLight's own sampler (reset per frame, Renderer.cpp:58-59).
        const bool replay = (h->flags & MRT_FLAG_REPLAY_SAMPLERS) && h->scene.n_lights > 0 &&
                            (h->rp.shader == MRT_SHADER_NOSHADOWS || h->rp.shader == MRT_SHADER_WHITTED);
        const int nl = h->scene.n_lights;
        const size_t slots = static_cast<size_t>(kNumberOfTiles) * h->rp.bsx * h->rp.bsy;
        DevMem d_index, d_light;
        std::vector<uint32_t> hv;
        uint32_t light_counter[kReplayMaxLights] = {};
        if (replay) {
            if (sample_begin != 0 || sample_stride != 1) { return fail(h, MRT_ERR_INVALID, "MRT_FLAG_REPLAY_SAMPLERS needs every sample of the frame on this handle"); }
            API_CUDA(h, d_index.alloc(slots * sizeof(uint32_t)));
            API_CUDA(h, d_light.alloc(slots * nl * sizeof(uint32_t)));
            hv.resize(slots * static_cast<size_t>(nl));
        }
        for (int s = sample_begin; s < spp; s += sample_stride, ++k) {
            if (static_cast<size_t>(k) >= depth) {
                API_CUDA(h, cudaEventSynchronize(h->sample_events[static_cast<size_t>(k) % depth]));
                h->sample.store(s - static_cast<int>(depth - 1) * sample_stride);
            }
            if (h->stop.load()) { stopped = true; break; }
            ReplayArgs rep {REPLAY_OFF, nl, d_index.as<uint32_t>(), d_light.as<uint32_t>()};
            if (replay) {
                const FrameTargets none {nullptr, nullptr};
                API_CUDA(h, cudaMemsetAsync(d_index.p, 0, slots * sizeof(uint32_t), st));
                rep.mode = REPLAY_COUNT_INDEX;
                API_CUDA(h, launch_render_sample(h->scene, h->camera, h->rp, h->d_inv_tile.as<int2>(), s, 1, none, h->d_counters.as<DeviceCounters>(), st, rep));
                API_CUDA(h, cudaMemcpyAsync(hv.data(), d_index.p, slots * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                API_CUDA(h, cudaStreamSynchronize(st));
                for (size_t i = 0; i < slots; ++i) { const uint32_t c = hv[i]; hv[i] = h->replay_index_counter; h->replay_index_counter += c; }
                API_CUDA(h, cudaMemcpyAsync(d_index.p, hv.data(), slots * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
                API_CUDA(h, cudaMemsetAsync(d_light.p, 0, slots * nl * sizeof(uint32_t), st));
                rep.mode = REPLAY_COUNT_LIGHTS;
                API_CUDA(h, launch_render_sample(h->scene, h->camera, h->rp, h->d_inv_tile.as<int2>(), s, 1, none, h->d_counters.as<DeviceCounters>(), st, rep));
                API_CUDA(h, cudaMemcpyAsync(hv.data(), d_light.p, slots * nl * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                API_CUDA(h, cudaStreamSynchronize(st));
                for (size_t i = 0; i < slots; ++i) {
                    for (int l = 0; l < nl; ++l) { uint32_t &v = hv[i * nl + l]; const uint32_t c = v; v = light_counter[l]; light_counter[l] += c; }
                }
                API_CUDA(h, cudaMemcpyAsync(d_light.p, hv.data(), slots * nl * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
                rep.mode = REPLAY_EXACT;
                launches += 2;
            }
            API_CUDA(h, launch_render_sample(h->scene, h->camera, h->rp, h->d_inv_tile.as<int2>(), s, d_bitmap ? s + 1 : k + 1, tg,
                                             h->d_counters.as<DeviceCounters>(), st, rep));
            ++launches;
            ++samples_rendered;
            API_CUDA(h, progress_refresh(h, d_bitmap));
            while (h->sample_events.size() < depth) { cudaEvent_t e; API_CUDA(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->sample_events.push_back(e); }
            API_CUDA(h, cudaEventRecord(h->sample_events[static_cast<size_t>(k) % depth], st));
        }
    }
    API_CUDA(h, cudaEventRecord(h->ev1, st));
    DeviceCounters hc {};
    API_CUDA(h, cudaMemcpyAsync(&hc, h->d_counters.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaStreamSynchronize(st));
    hc.rays_primary += samples_rendered * h->roots_per_sample;
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, h->ev0, h->ev1);
    if (h->wavefront) { API_CUDA(h, wavefront_collect_times(h->wavefront, &wt)); }
    {
        std::lock_guard<std::mutex> lock(h->mutex);
        h->stats.frame_ms = ms;
        h->stats.extend_ms = wt.extend_ms; h->stats.shadow_ms = wt.shadow_ms; h->stats.shade_ms = wt.shade_ms; h->stats.other_ms = wt.other_ms;
        h->stats.rays_primary = hc.rays_primary; h->stats.rays_shadow = hc.rays_shadow; h->stats.rays_secondary = hc.rays_secondary;
        h->stats.launches = launches;
        h->stats.rays_extend_traced = hc.extend_traced; h->stats.rays_shadow_traced = hc.shadow_traced;
        const uint64_t per_class = h->rp.shader == MRT_SHADER_PATHTRACER ? wt.extend_launches : 0u;   // 6 levels x (extend, shade, shadow) per pass
        h->stats.extend_launches = per_class; h->stats.shade_launches = per_class; h->stats.shadow_launches = per_class;
        h->stats.queue_bytes = h->wavefront ? wavefront_bytes(h->wavefront) : 0u;
    }
    h->total_rays.store(hc.rays_primary + hc.rays_shadow + hc.rays_secondary);
    if (!stopped) { h->sample.store(spp); }
    if (hc.overflow) { return fail(h, MRT_ERR_OVERFLOW, "traversal stack or ray queue overflow"); }
    if (stopped) { return fail(h, MRT_ERR_STOPPED, "stopped"); }
    errno = 0;
    return MRT_OK;
}

int render_to_device_bitmap(mrt_handle *h, int32_t *d_bitmap) {
    if (h->flags & MRT_FLAG_FP32_ACCUMULATE) {
        const size_t npix = static_cast<size_t>(h->rp.width) * h->rp.height;
        API_CUDA(h, cudaMemsetAsync(h->d_accum.p, 0, npix * sizeof(float4), h->stream));
        const int rc = render_samples(h, nullptr, h->d_accum.as<float4>(), 0, 1);
        if (rc != MRT_OK && rc != MRT_ERR_STOPPED) { return rc; }
        API_CUDA(h, launch_resolve(h->d_accum.as<float4>(), d_bitmap, static_cast<int>(npix), h->stream));
        API_CUDA(h, cudaStreamSynchronize(h->stream));
        return rc;
    }
    return render_samples(h, d_bitmap, nullptr, 0, 1);
}

}  // namespace

extern "C" {

int mrt_render_frame_device(mrt_handle *h, void *device_bitmap) {
    if (!h || !device_bitmap) { return fail(h, MRT_ERR_INVALID, "null argument"); }
    return render_to_device_bitmap(h, static_cast<int32_t *>(device_bitmap));
}

int mrt_render_frame(mrt_handle *h, int32_t *host_bitmap) {
    if (!h || !host_bitmap) { return fail(h, MRT_ERR_INVALID, "null argument"); }
    API_CUDA(h, cudaSetDevice(h->device));
    const size_t bytes = static_cast<size_t>(h->rp.width) * h->rp.height * sizeof(int32_t);
    // a caller that hands in page-locked memory (cudaHostAlloc / cudaHostRegister) gets the DMA straight into it
    cudaPointerAttributes attr {};
    const bool pinned = cudaPointerGetAttributes(&attr, host_bitmap) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    // The frame is a read-modify-write of the caller's bitmap (Renderer.cpp:149-157). When the 16x16 tile grid covers the
    // image exactly every pixel is overwritten at sample 1 (incrementalAvg weights the old value by n-1 = 0), so the old
    // content is never needed; otherwise the pixels no tile reaches must keep the caller's content: upload it first.
    const bool covers_all = h->rp.width % 16 == 0 && h->rp.height % 16 == 0;
    if (!covers_all) {
        if (pinned) { API_CUDA(h, cudaMemcpyAsync(h->d_bitmap.p, host_bitmap, bytes, cudaMemcpyHostToDevice, h->stream)); }
        else { std::memcpy(h->h_pinned_bitmap, host_bitmap, bytes); API_CUDA(h, cudaMemcpyAsync(h->d_bitmap.p, h->h_pinned_bitmap, bytes, cudaMemcpyHostToDevice, h->stream)); }
    }
    if (h->flags & MRT_FLAG_PROGRESSIVE) {
        if (!h->copy_stream) {
            API_CUDA(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
            API_CUDA(h, cudaEventCreateWithFlags(&h->ev_progress, cudaEventDisableTiming));
            API_CUDA(h, cudaEventCreateWithFlags(&h->ev_copied, cudaEventDisableTiming));
        }
        if (!covers_all) { API_CUDA(h, cudaStreamSynchronize(h->stream)); }   // the staging buffer is free again
        h->progress_pinned = pinned; h->progress_bytes = bytes; h->progress_host = host_bitmap;
    }
    const int rc = render_to_device_bitmap(h, h->d_bitmap.as<int32_t>());
    if (h->progress_host) {
        cudaStreamSynchronize(h->copy_stream);   // no refresh may land after the final copy below
        h->progress_host = nullptr;
    }
    if (rc != MRT_OK && rc != MRT_ERR_STOPPED) { return rc; }
    if (pinned) {
        API_CUDA(h, cudaMemcpyAsync(host_bitmap, h->d_bitmap.p, bytes, cudaMemcpyDeviceToHost, h->stream));
        API_CUDA(h, cudaStreamSynchronize(h->stream));
    } else {
        API_CUDA(h, cudaMemcpyAsync(h->h_pinned_bitmap, h->d_bitmap.p, bytes, cudaMemcpyDeviceToHost, h->stream));
        API_CUDA(h, cudaStreamSynchronize(h->stream));
        std::memcpy(host_bitmap, h->h_pinned_bitmap, bytes);
    }
    errno = 0;
    return rc;
}

int mrt_render_accumulate_device(mrt_handle *h, void *device_accum, int32_t sample_begin, int32_t sample_stride) {
    if (!h || !device_accum || sample_begin < 0 || sample_stride <= 0) { return fail(h, MRT_ERR_INVALID, "bad argument"); }
    return render_samples(h, nullptr, static_cast<float4 *>(device_accum), sample_begin, sample_stride);
}

int mrt_resolve_device(mrt_handle *h, const void *device_accum, void *device_bitmap) {
    if (!h || !device_accum || !device_bitmap) { return fail(h, MRT_ERR_INVALID, "null argument"); }
    API_CUDA(h, cudaSetDevice(h->device));
    API_CUDA(h, launch_resolve(static_cast<const float4 *>(device_accum), static_cast<int32_t *>(device_bitmap), h->rp.width * h->rp.height, h->stream));
    API_CUDA(h, cudaStreamSynchronize(h->stream));
    errno = 0;
    return MRT_OK;
}

}  // extern "C"

// ---- per-ray and primitive-level entry points (parity tests) -----------------------------------------------------
namespace {
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
}  // namespace

extern "C" {

int mrt_trace_rays(mrt_handle *h, int64_t n, const float *origins, const float *directions, const int32_t *origin_kind,
                   const int32_t *origin_idx, const float *shadow_dist, float *out_t, int32_t *out_kind, int32_t *out_idx,
                   int32_t *out_material, float *out_normal) {
    if (!h || n < 0 || !origins || !directions || !out_t || !out_kind || !out_idx || !out_material) { return fail(h, MRT_ERR_INVALID, "bad argument"); }
    API_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    TmpDev tmp;
    const size_t N = static_cast<size_t>(n);
    float *d_o, *d_d, *d_sd = nullptr, *d_t, *d_n = nullptr;
    int32_t *d_ok = nullptr, *d_oi = nullptr, *d_kind, *d_idx, *d_mat;
    API_CUDA(h, tmp.put(&d_o, origins, 3 * N, st));
    API_CUDA(h, tmp.put(&d_d, directions, 3 * N, st));
    std::vector<int32_t> slots;
    if (origin_kind && origin_idx) {
        // caller indices -> device slots for planes (spheres never self-skip)
        slots.assign(origin_idx, origin_idx + N);
        for (size_t i = 0; i < N; ++i) {
            if (origin_kind[i] == HIT_PLANE && slots[i] >= 0 && static_cast<size_t>(slots[i]) < h->plane_slot_of_id.size()) { slots[i] = h->plane_slot_of_id[static_cast<size_t>(slots[i])]; }
        }
        API_CUDA(h, tmp.put(&d_ok, origin_kind, N, st));
        API_CUDA(h, tmp.put(&d_oi, slots.data(), N, st));
    }
    if (shadow_dist) { API_CUDA(h, tmp.put(&d_sd, shadow_dist, N, st)); }
    API_CUDA(h, tmp.get(&d_t, N)); API_CUDA(h, tmp.get(&d_kind, N)); API_CUDA(h, tmp.get(&d_idx, N)); API_CUDA(h, tmp.get(&d_mat, N));
    if (out_normal) { API_CUDA(h, tmp.get(&d_n, 3 * N)); API_CUDA(h, cudaMemsetAsync(d_n, 0, 3 * N * sizeof(float), st)); }
    API_CUDA(h, cudaMemsetAsync(h->d_counters.p, 0, sizeof(DeviceCounters), st));
    // a PathTracer handle answers through its own persistent extend / shadow kernels, any other through the per-thread traversal
    if (h->wavefront && wavefront_capacity(h->wavefront) >= static_cast<uint64_t>(n)) {
        API_CUDA(h, wavefront_trace_rays(h->wavefront, h->scene, n, d_o, d_d, d_ok, d_oi, d_sd, d_t, d_kind, d_idx, d_mat, d_n, h->d_counters.as<DeviceCounters>(), st));
    } else {
        API_CUDA(h, launch_trace_rays(h->scene, n, d_o, d_d, d_ok, d_oi, d_sd, d_t, d_kind, d_idx, d_mat, d_n, h->d_counters.as<DeviceCounters>(), st));
    }
    DeviceCounters hc {};
    API_CUDA(h, cudaMemcpyAsync(out_t, d_t, N * 4, cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaMemcpyAsync(out_kind, d_kind, N * 4, cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaMemcpyAsync(out_idx, d_idx, N * 4, cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaMemcpyAsync(out_material, d_mat, N * 4, cudaMemcpyDeviceToHost, st));
    if (out_normal) { API_CUDA(h, cudaMemcpyAsync(out_normal, d_n, 3 * N * 4, cudaMemcpyDeviceToHost, st)); }
    API_CUDA(h, cudaMemcpyAsync(&hc, h->d_counters.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaStreamSynchronize(st));
    if (hc.overflow) { return fail(h, MRT_ERR_OVERFLOW, "traversal stack overflow"); }
    errno = 0;
    return MRT_OK;
}

// No counterpart in the reference: mrt_trace_rays for callers that may not
// allocate. DEVICE arrays in and out, the kernels' queue scratch in a static device array (mrt_wavefront.cu); no cudaMalloc,
// cudaFree or any allocating routine on this path.
int mrt_trace_rays_static(mrt_handle *h, int64_t n, const float *d_origins, const float *d_directions, const int32_t *d_origin_kind,
                          const int32_t *d_origin_idx, const float *d_shadow_dist, float *d_out_t, int32_t *d_out_kind, int32_t *d_out_idx,
                          int32_t *d_out_material, float *d_out_normal) {
    if (!h || n < 0 || !d_origins || !d_directions || !d_out_t || !d_out_kind || !d_out_idx || !d_out_material) { return fail(h, MRT_ERR_INVALID, "bad argument"); }
    if ((d_origin_kind == nullptr) != (d_origin_idx == nullptr)) { return fail(h, MRT_ERR_INVALID, "origin_kind and origin_idx go together"); }
    if (static_cast<uint64_t>(n) > wavefront_static_capacity()) {
        return fail(h, MRT_ERR_INVALID, "mrt_trace_rays_static: n = " + std::to_string(n) + " exceeds the static scratch capacity of " +
                                        std::to_string(wavefront_static_capacity()) + " rays (MRT_STATIC_TRACE_MAX_RAYS); split the batch or use mrt_trace_rays");
    }
    API_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    const size_t N = static_cast<size_t>(n);
    if (d_out_normal) { API_CUDA(h, cudaMemsetAsync(d_out_normal, 0, 3 * N * sizeof(float), st)); }
    API_CUDA(h, cudaMemsetAsync(h->d_counters.p, 0, sizeof(DeviceCounters), st));
    API_CUDA(h, wavefront_trace_rays_static(h->scene, n, d_origins, d_directions, d_origin_kind, d_origin_idx, d_shadow_dist, d_out_t, d_out_kind,
                                            d_out_idx, d_out_material, d_out_normal, h->d_counters.as<DeviceCounters>(), st));
    DeviceCounters hc {};
    API_CUDA(h, cudaMemcpyAsync(&hc, h->d_counters.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaStreamSynchronize(st));
    if (hc.overflow) { return fail(h, MRT_ERR_OVERFLOW, "traversal stack overflow"); }
    errno = 0;
    return MRT_OK;
}

int mrt_generate_rays(mrt_handle *h, int64_t n, const float *uvdd, float *out_origins, float *out_directions) {
    if (!h || n < 0 || !uvdd || !out_origins || !out_directions) { return fail(h, MRT_ERR_INVALID, "bad argument"); }
    API_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    TmpDev tmp;
    const size_t N = static_cast<size_t>(n);
    float *d_in, *d_o, *d_d;
    API_CUDA(h, tmp.put(&d_in, uvdd, 4 * N, st));
    API_CUDA(h, tmp.get(&d_o, 3 * N)); API_CUDA(h, tmp.get(&d_d, 3 * N));
    API_CUDA(h, launch_generate_rays(h->camera, n, d_in, d_o, d_d, st));
    API_CUDA(h, cudaMemcpyAsync(out_origins, d_o, 3 * N * 4, cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaMemcpyAsync(out_directions, d_d, 3 * N * 4, cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaStreamSynchronize(st));
    errno = 0;
    return MRT_OK;
}

int mrt_shade_rays(mrt_handle *h, int64_t n, const float *origins, const float *directions, int32_t depth, float *out_rgb) {
    if (!h || n < 0 || !origins || !directions || !out_rgb) { return fail(h, MRT_ERR_INVALID, "bad argument"); }
    if (h->rp.shader == MRT_SHADER_PATHTRACER) { return fail(h, MRT_ERR_INVALID, "mrt_shade_rays serves the deterministic shaders only"); }
    API_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    TmpDev tmp;
    const size_t N = static_cast<size_t>(n);
    float *d_o, *d_d, *d_rgb;
    API_CUDA(h, tmp.put(&d_o, origins, 3 * N, st));
    API_CUDA(h, tmp.put(&d_d, directions, 3 * N, st));
    API_CUDA(h, tmp.get(&d_rgb, 3 * N));
    API_CUDA(h, cudaMemsetAsync(h->d_counters.p, 0, sizeof(DeviceCounters), st));
    API_CUDA(h, launch_shade_rays(h->scene, h->rp, n, d_o, d_d, depth, d_rgb, h->d_counters.as<DeviceCounters>(), st));
    API_CUDA(h, cudaMemcpyAsync(out_rgb, d_rgb, 3 * N * 4, cudaMemcpyDeviceToHost, st));
    API_CUDA(h, cudaStreamSynchronize(st));
    errno = 0;
    return MRT_OK;
}

}  // extern "C"

#ifdef MRT_TRAVERSAL_STATS
// diagnostic library only: out[0..8] = nodes with k valid children, out[9..16] = leaf slots holding 0..7 triangles, out[17] = inner children
extern "C" int mrt_debug_bvh_stats(mrt_handle *h, unsigned long long out[18]) {
    std::vector<mrt::Node8> nodes(h->bvh.n_nodes);
    cudaMemcpy(nodes.data(), h->bvh.nodes, nodes.size() * sizeof(mrt::Node8), cudaMemcpyDeviceToHost);
    for (int i = 0; i < 18; ++i) { out[i] = 0; }
    for (const mrt::Node8 &n : nodes) {
        const uint32_t *w = reinterpret_cast<const uint32_t *>(&n);
        const uint32_t imask = w[3] >> 24;
        const uint8_t *meta = reinterpret_cast<const uint8_t *>(w + 6);
        int valid = 0;
        for (int k = 0; k < 8; ++k) {
            if (imask & (1u << k)) { ++valid; ++out[17]; }
            else if (meta[k] != 0) { ++valid; ++out[9 + (meta[k] >> 5)]; }
        }
        ++out[valid];
    }
    return 0;
}
#endif
