// Host-side builder of the reference's RegularGrid<T> (MobileRT/Accelerators/RegularGrid.hpp:112-297), selected with
// MRT_ACC_REGULAR_GRID. The reference builds one 32 x 32 x 32 grid per primitive type (Shader.cpp:56-62) over that
// type's own bounds (Scene::getBounds, Scene.hpp:52-62) and files a primitive under every cell of its box's cell range
// that passes T::intersect(AABB). Those overlap tests are generous in type-specific ways (a triangle's edges count as
// unbounded rays, the cell diagonal as an unbounded ray; a plane only looks at two cell corners), the start cell of a
// ray is clamped into the grid, and the walk never leaves through a face it did not enter - so which primitives a ray
// gets to test is a property of these lists, not of geometry. The lists are therefore rebuilt here decision by decision
// and the device walk (mrt_traverse.cuh, trace_grid) consumes them as they are.
// Order inside a cell: ascending caller index (the reference's order depends on its thread timing, RegularGrid.hpp:191-203).
#pragma once
#include "mrt_refbvh.h"
#include <cfloat>
#include <climits>
#include <cstdint>
#include <thread>

namespace mrt {

constexpr int kGridSize = 32, kGridShift = 5, kGridCells = kGridSize * kGridSize * kGridSize;

struct GridHost {
    float bmin[3] = {0, 0, 0}, bmax[3] = {0, 0, 0};
    float cellInv[3] = {0, 0, 0}, cellSize[3] = {0, 0, 0};
    std::vector<uint32_t> start;   // kGridCells + 1 offsets into items
    std::vector<int32_t> items;    // caller indices, cell after cell
};

// static_cast<int32_t>(float) as the reference's x86-64 build performs it (cvttss2si: out of range and NaN give INT_MIN)
inline int32_t grid_to_int(float v) {
    if (!(v > -2147483904.0f && v < 2147483648.0f)) { return INT_MIN; }
    return static_cast<int32_t>(v);
}

// Scene::getBounds (Scene.hpp:52-62) with glm::min / glm::max's operand order (getBoxBounds, Scene.cpp:36-39)
inline void grid_bounds(const std::vector<HostBox> &boxes, float bmin[3], float bmax[3]) {
    for (int k = 0; k < 3; ++k) { bmin[k] = kRayLengthMax; bmax[k] = -kRayLengthMax; }
    for (const HostBox &b : boxes) {
        for (int k = 0; k < 3; ++k) {
            bmin[k] = (bmin[k] < b.mn[k]) ? bmin[k] : b.mn[k];
            bmax[k] = (b.mx[k] < bmax[k]) ? bmax[k] : b.mx[k];
        }
    }
    for (int k = 0; k < 3; ++k) { bmin[k] = bmin[k] - kEpsilon; bmax[k] = bmax[k] + kEpsilon; }
}

// Triangle::getAABB (Triangle.cpp:116-123)
inline HostBox hb_triangle(const float A[3], const float AB[3], const float AC[3]) {
    HostBox b;
    for (int k = 0; k < 3; ++k) {
        const float pb = A[k] + AB[k], pc = A[k] + AC[k];
        b.mn[k] = std::min(A[k], std::min(pb, pc));
        b.mx[k] = std::max(A[k], std::max(pb, pc));
    }
    return b;
}

// ---- T::intersect(const AABB &) ------------------------------------------------------------------------------

// Plane::intersect(AABB) (Plane.cpp:146-155) on Plane::distance (Plane.cpp:117-138): the signed distances of the box's
// max and min corners have different signs (or one is zero)
inline bool grid_plane_overlaps(const float n[3], const float p[3], const HostBox &cell) {
    const float d = n[0] * -p[0] + n[1] * -p[1] + n[2] * -p[2];
    const float den = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    const float distP = (n[0] * cell.mx[0] + n[1] * cell.mx[1] + n[2] * cell.mx[2] + d) / den;
    const float distN = (n[0] * cell.mn[0] + n[1] * cell.mn[1] + n[2] * cell.mn[2] + d) / den;
    return (distP <= 0 && distN >= 0) || (distP >= 0 && distN <= 0);
}

// Sphere::intersect(AABB) (Sphere.cpp:102-123): squared distance from the centre to the box against the squared radius
inline bool grid_sphere_overlaps(const float c[3], float sqRadius, const HostBox &cell) {
    float dmin = 0.0f;
    for (int k = 0; k < 3; ++k) {
        if (c[k] < cell.mn[k]) { dmin = dmin + (c[k] - cell.mn[k]) * (c[k] - cell.mn[k]); }
        else if (c[k] > cell.mx[k]) { dmin = dmin + (c[k] - cell.mx[k]) * (c[k] - cell.mx[k]); }
    }
    return dmin <= sqRadius;
}

// the slab walk Triangle::intersect(AABB) runs per edge (Triangle.cpp:143-201): the edge is a ray from `orig` along `vec`
// with no upper bound on t, tNear starts at the smallest normal float
inline bool grid_edge_hits_box(const float orig[3], const float vec[3], const HostBox &box) {
    float tNear = FLT_MIN, tFar = FLT_MAX;
    for (int k = 0; k < 3; ++k) {
        if (std::fabs(vec[k]) < FLT_EPSILON) {
            if (orig[k] < box.mn[k] || (orig[k] + vec[k]) > box.mx[k]) { return false; }
        } else {
            float t1 = (box.mn[k] - orig[k]) / vec[k];
            float t2 = (box.mx[k] - orig[k]) / vec[k];
            if (t1 > t2) { std::swap(t1, t2); }
            tNear = std::max(t1, tNear);
            tFar = std::min(t2, tFar);
            if (tNear > tFar || tFar < 0) { return false; }   // Triangle::isNearFarInvalid (Triangle.cpp:132-134)
        }
    }
    return true;
}

inline void grid_cross(const float a[3], const float b[3], float r[3]) {
    r[0] = a[1] * b[2] - b[1] * a[2];
    r[1] = a[2] * b[0] - b[2] * a[0];
    r[2] = a[0] * b[1] - b[0] * a[1];
}
inline float grid_dot(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// Triangle::intersect(Intersection) (Triangle.cpp:63-99) for a ray of length RayLengthMax: only "is there a hit"
inline bool grid_ray_hits_triangle(const float A[3], const float AB[3], const float AC[3], const float o[3], const float d[3]) {
    float pv[3], toCam[3], up[3];
    grid_cross(d, AC, pv);
    const float det = grid_dot(AB, pv);
    if (std::fabs(det) < kEpsilon) { return false; }
    const float inv = 1.0f / det;
    for (int k = 0; k < 3; ++k) { toCam[k] = o[k] - A[k]; }
    const float u = inv * grid_dot(toCam, pv);
    if (u < 0.0f || u > 1.0f) { return false; }
    grid_cross(toCam, AB, up);
    const float v = inv * grid_dot(d, up);
    if (v < 0.0f || (u + v) > 1.0f) { return false; }
    const float t = inv * grid_dot(AC, up);
    return !(t < kEpsilon || t >= kRayLengthMax);
}

// Triangle::intersect(AABB) (Triangle.cpp:142-229): any of the three edge rays meets the box, or the ray from the box's
// min corner along its diagonal meets the triangle, or that diagonal is parallel to the triangle's plane
inline bool grid_triangle_overlaps(const float A[3], const float AB[3], const float AC[3], const HostBox &cell) {
    if (grid_edge_hits_box(A, AB, cell) || grid_edge_hits_box(A, AC, cell)) { return true; }
    float B[3], BC[3], diag[3];
    for (int k = 0; k < 3; ++k) {
        B[k] = A[k] + AB[k];
        const float c = A[k] + AC[k];
        BC[k] = c - B[k];
        diag[k] = cell.mx[k] - cell.mn[k];
    }
    if (grid_edge_hits_box(B, BC, cell)) { return true; }
    if (grid_ray_hits_triangle(A, AB, AC, cell.mn, diag)) { return true; }
    float pv[3];
    grid_cross(diag, AC, pv);
    return std::fabs(grid_dot(AB, pv)) < kEpsilon;
}

// ---- RegularGrid<T>::RegularGrid + addPrimitivesThreadWork (RegularGrid.hpp:112-149, 212-283) ---------------
// overlaps(i, cell) is T::intersect(AABB) of primitive i.
template <typename OverlapFn>
void build_ref_grid(const std::vector<HostBox> &boxes, OverlapFn overlaps, GridHost &g) {
    const int64_t n = static_cast<int64_t>(boxes.size());
    grid_bounds(boxes, g.bmin, g.bmax);
    float size[3], dcell[3], dreci[3];
    for (int k = 0; k < 3; ++k) {
        size[k] = g.bmax[k] - g.bmin[k];
        g.cellInv[k] = static_cast<float>(kGridSize) / size[k];
        g.cellSize[k] = size[k] * (1.0f / static_cast<float>(kGridSize));
        dcell[k] = size[k] / static_cast<float>(kGridSize);
        dreci[k] = dcell[k] > 0 ? 1.0f / dcell[k] : 1.0f;
    }
    g.start.assign(static_cast<size_t>(kGridCells) + 1, 0u);
    g.items.clear();
    if (n == 0) { return; }
    unsigned hw = std::thread::hardware_concurrency();
    const int64_t n_workers = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(hw ? hw : 1, 32), n / 4096 + 1));
    std::vector<std::vector<uint64_t>> found(static_cast<size_t>(n_workers));   // (cell << 32) | primitive, primitives ascending per worker
    auto work = [&](int64_t wk) {
        const int64_t i0 = n * wk / n_workers, i1 = n * (wk + 1) / n_workers;
        std::vector<uint64_t> &out = found[static_cast<size_t>(wk)];
        for (int64_t i = i0; i < i1; ++i) {
            const HostBox &b = boxes[static_cast<size_t>(i)];
            int32_t lo[3], hi[3];
            for (int k = 0; k < 3; ++k) {
                int32_t c1 = grid_to_int((b.mn[k] - g.bmin[k]) * dreci[k]);
                const int32_t raw2 = grid_to_int((b.mx[k] - g.bmin[k]) * dreci[k]);
                int32_t c2 = raw2 == INT_MAX ? INT_MAX : raw2 + 1;
                c1 = std::max(0, c1);
                c2 = std::min(c2, kGridSize - 1);
                c2 = std::fabs(size[k]) < FLT_EPSILON ? 0 : c2;
                c1 = std::min(c1, c2);
                lo[k] = c1; hi[k] = c2;
            }
            for (int32_t x = lo[0]; x <= hi[0]; ++x) {
                for (int32_t y = lo[1]; y <= hi[1]; ++y) {
                    for (int32_t z = lo[2]; z <= hi[2]; ++z) {
                        HostBox cell;
                        cell.mn[0] = g.bmin[0] + static_cast<float>(x) * dcell[0];
                        cell.mn[1] = g.bmin[1] + static_cast<float>(y) * dcell[1];
                        cell.mn[2] = g.bmin[2] + static_cast<float>(z) * dcell[2];
                        for (int k = 0; k < 3; ++k) { cell.mx[k] = cell.mn[k] + dcell[k]; }
                        if (overlaps(i, cell)) {
                            // negative coordinates cannot occur for the lists the walk reads (it clamps into 0..31), so they are dropped
                            if (x < 0 || y < 0 || z < 0) { continue; }
                            const uint32_t idx = static_cast<uint32_t>(x) + (static_cast<uint32_t>(y) << kGridShift) + (static_cast<uint32_t>(z) << (2 * kGridShift));
                            out.push_back((static_cast<uint64_t>(idx) << 32) | static_cast<uint64_t>(i));
                        }
                    }
                }
            }
        }
    };
    if (n_workers == 1) { work(0); }
    else {
        std::vector<std::thread> threads;
        for (int64_t wk = 1; wk < n_workers; ++wk) { threads.emplace_back(work, wk); }
        work(0);
        for (std::thread &t : threads) { t.join(); }
    }
    // counting sort by cell; workers hold ascending primitive ranges, so every cell's list comes out ascending
    for (const auto &v : found) { for (uint64_t e : v) { ++g.start[static_cast<size_t>(e >> 32) + 1]; } }
    for (size_t c = 0; c < static_cast<size_t>(kGridCells); ++c) { g.start[c + 1] += g.start[c]; }
    g.items.assign(g.start[kGridCells], 0);
    std::vector<uint32_t> fill(g.start.begin(), g.start.end() - 1);
    for (const auto &v : found) { for (uint64_t e : v) { g.items[fill[static_cast<size_t>(e >> 32)]++] = static_cast<int32_t>(e & 0xffffffffull); } }
}

}  // namespace mrt
