// Host-side builder of the reference's binary BVH layout for the (few) planes and spheres of a scene.
// Which rays may see a plane at all depends on the box the reference gives it (Plane::getAABB, Plane.cpp:103-109:
// a finite, zero-thickness box) and on how BVH<T>::build groups those boxes into leaves (BVH.hpp:160-283), so this
// path follows the same decisions step by step: longest axis, ten std::partition buckets, full SAH sweep
// (BVH.hpp:396-439), leaves of at most four primitives, children stored adjacently.
// The big triangle set never goes through here; it is built on the GPU (mrt_bvh_build.cu).
#pragma once
#include "mrt_types.h"
#include <algorithm>
#include <cmath>
#include <vector>

namespace mrt {

struct HostBox { float mn[3], mx[3]; };

inline HostBox hb_union(const HostBox &a, const HostBox &b) {
    HostBox r;
    for (int k = 0; k < 3; ++k) { r.mn[k] = std::min(a.mn[k], b.mn[k]); r.mx[k] = std::max(a.mx[k], b.mx[k]); }
    return r;
}
// AABB::getSurfaceArea (AABB.cpp:61-71)
inline float hb_area(const HostBox &b) {
    const float l0 = b.mx[0] - b.mn[0], l1 = b.mx[1] - b.mn[1], l2 = b.mx[2] - b.mn[2];
    const float bottomTop = 2 * l0 * l2, sideXY = 2 * l0 * l1, sideZY = 2 * l2 * l1;
    return bottomTop + sideXY + sideZY;
}
// Sphere::getAABB (Sphere.cpp:88-94)
inline HostBox hb_sphere(const float c[3], float sqRadius) {
    const float r = std::sqrt(sqRadius);
    HostBox b;
    for (int k = 0; k < 3; ++k) { b.mn[k] = c[k] - r; b.mx[k] = c[k] + r; }
    return b;
}
// Plane::getRightVector + getAABB (Plane.cpp:79-109). A normal that is not axis aligned yields a NaN box, as in the reference.
inline HostBox hb_plane(const float n[3], const float p[3]) {
    float r[3] = {0, 0, 0};
    if (n[0] >= 1) { r[1] = 1; r[2] = 1; } else if (n[1] >= 1) { r[0] = 1; r[2] = 1; } else if (n[2] >= 1) { r[0] = 1; r[1] = 1; }
    else if (n[0] <= -1) { r[1] = 1; r[2] = 1; } else if (n[1] <= -1) { r[0] = 1; r[2] = 1; } else if (n[2] <= -1) { r[0] = 1; r[1] = 1; }
    const float inv = 1.0f / std::sqrt((r[0] * r[0] + r[1] * r[1]) + r[2] * r[2]);
    HostBox b;
    for (int k = 0; k < 3; ++k) { const float rk = r[k] * inv; b.mn[k] = p[k] + rk * -100.0f; b.mx[k] = p[k] + rk * 100.0f; }
    return b;
}

// Returns the nodes and `order` (order[i] = input index of the primitive stored at position i).
inline void build_ref_bvh(const std::vector<HostBox> &prims, std::vector<RefNode> &nodes, std::vector<int> &order) {
    const int n = static_cast<int>(prims.size());
    nodes.clear(); order.clear();
    if (n == 0) { nodes.push_back(RefNode{0, 0, 0, 0, 0, 0, 0, 0}); return; }
    struct BN { HostBox box; float c[3]; int old; };
    std::vector<BN> bn(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) {
        bn[i].box = prims[i]; bn[i].old = i;
        for (int k = 0; k < 3; ++k) { bn[i].c[k] = prims[i].mn[k] + (prims[i].mx[k] - prims[i].mn[k]) / 2.0f; }  // AABB::getCentroid
    }
    nodes.assign(static_cast<size_t>(2 * n - 1), RefNode{0, 0, 0, 0, 0, 0, 0, 0});
    struct Task { int node, begin, end; };
    std::vector<Task> stack;
    int cur = 0, begin = 0, end = n, maxNode = 0;
    bool more = true;
    while (more) {
        HostBox sb = bn[begin].box;
        for (int i = begin + 1; i < end; ++i) { sb = hb_union(sb, bn[i].box); }
        const float md[3] = {sb.mx[0] - sb.mn[0], sb.mx[1] - sb.mn[1], sb.mx[2] - sb.mn[2]};
        const int axis = (md[0] >= md[1] && md[0] >= md[2]) ? 0 : ((md[1] >= md[0] && md[1] >= md[2]) ? 1 : 2);
        const float stepAxis = md[axis] / 10.0f, startBox = sb.mn[axis];
        auto it = std::partition(bn.begin() + begin, bn.begin() + end, [&](const BN &b) { return b.c[axis] < startBox + stepAxis; });
        for (int bucket = 2; bucket < 10; ++bucket) {
            const float limit = startBox + stepAxis * static_cast<float>(bucket);
            it = std::partition(it, bn.begin() + end, [&](const BN &b) { return b.c[axis] < limit; });
        }
        HostBox nb = bn[begin].box;
        for (int i = begin + 1; i < end; ++i) { nb = hb_union(bn[i].box, nb); }
        RefNode &node = nodes[static_cast<size_t>(cur)];
        node.mnx = nb.mn[0]; node.mny = nb.mn[1]; node.mnz = nb.mn[2]; node.mxx = nb.mx[0]; node.mxy = nb.mx[1]; node.mxz = nb.mx[2];
        const int size = end - begin;
        if (size <= 4) {
            node.indexOffset = begin; node.numPrimitives = size;
            if (stack.empty()) { more = false; } else { const Task t = stack.back(); stack.pop_back(); cur = t.node; begin = t.begin; end = t.end; }
        } else {
            // getSplitIndexSah over the bucket-ordered boxes
            const int numBoxes = size - 1;
            std::vector<float> leftArea(static_cast<size_t>(numBoxes)), rightArea(static_cast<size_t>(numBoxes));
            HostBox lb = bn[begin].box; leftArea[0] = hb_area(lb);
            for (int i = 1; i < numBoxes; ++i) { lb = hb_union(lb, bn[begin + i].box); leftArea[i] = hb_area(lb); }
            HostBox rb = bn[begin + numBoxes].box; rightArea[numBoxes - 1] = hb_area(rb);
            for (int i = numBoxes - 2; i >= 0; --i) { rb = hb_union(rb, bn[begin + i + 1].box); rightArea[i] = hb_area(rb); }
            int split = 1;
            float minSah = leftArea[0] + static_cast<float>(numBoxes) * rightArea[0];
            for (int i = 1; i < numBoxes; ++i) {
                const float sah = static_cast<float>(i + 1) * leftArea[i] + static_cast<float>(size - (i + 1)) * rightArea[i];
                if (sah < minSah) { split = i + 1; minSah = sah; }
            }
            const int left = maxNode + 1, right = left + 1;
            node.indexOffset = left; node.numPrimitives = 0;
            maxNode = std::max(right, maxNode);
            stack.push_back(Task{right, begin + split, end});
            cur = left; end = begin + split;
        }
    }
    nodes.resize(static_cast<size_t>(maxNode + 1));
    order.resize(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) { order[i] = bn[i].old; }
}

}  // namespace mrt
