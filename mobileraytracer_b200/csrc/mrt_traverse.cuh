// Closest-hit and any-hit traversal (replaces BVH<T>::intersect, MobileRT/Accelerators/BVH.hpp:326-384, as driven by
// Shader::rayTrace / Shader::shadowTrace, MobileRT/Shader.cpp:86-158).
//  * triangles and area-light triangles: compressed 8-wide BVH, octant-ordered descent, t-max culling, per-thread stack
//  * planes / spheres: the reference's own binary BVH (built on the host from the same algorithm) walked exactly like
//    BVH.hpp does, because Plane::getAABB's finite flat box (Plane.cpp:103-109) decides which rays may see a plane.
// Parity rule: at equal t the reference keeps the FIRST hit of its left-first DFS, i.e. the lowest primitive index
// of its own array; the triangle test below therefore breaks exact-t ties towards the lower primitive id.
#pragma once
#include "mrt_types.h"

namespace mrt {

constexpr int kStack8 = 40;     // node-group stack entries per thread (local memory, L1 resident)
constexpr int kStackRef = 64;

struct HitRec {
    float t;
    int32_t kind;   // HitKind
    int32_t idx;    // triangle: caller index; light: light index; plane / sphere: device slot (see PlaneRec::id)
    float u, v;     // barycentrics (triangles / lights)
};

// ---- reference-layout binary BVH over planes or spheres ---------------------------------------------------
// PRIM: 0 planes, 1 spheres.  Returns true as soon as a shadow ray found an occluder.
template <int PRIM>
MRT_D bool trace_ref_bvh(const SceneDev &s, v3 o, v3 d, bool shadow, int skipKind, int skipIdx, HitRec &h, int *overflow) {
    const RefNode *nodes = PRIM == 0 ? s.plane_nodes : s.sphere_nodes;
    const int nprims = PRIM == 0 ? s.n_planes : s.n_spheres;
    if (nprims == 0) { return false; }
    int stack[kStackRef];
    int sp = 1;
    int boxIndex = 0;
    stack[0] = 0;
    do {
        const RefNode node = nodes[boxIndex];
        if (aabb_intersect_exact(mk3(node.mnx, node.mny, node.mnz), mk3(node.mxx, node.mxy, node.mxz), o, d)) {
            if (node.numPrimitives > 0) {
                for (int i = 0; i < node.numPrimitives; ++i) {
                    float t;
                    bool ok;
                    int id, kind;
                    if (PRIM == 0) {
                        const PlaneRec p = s.planes[node.indexOffset + i];
                        id = node.indexOffset + i; kind = HIT_PLANE;   // device slot; PlaneRec::id maps back to the caller's index
                        ok = !(skipKind == HIT_PLANE && skipIdx == id) &&
                             plane_intersect(mk3(p.nx, p.ny, p.nz), mk3(p.px, p.py, p.pz), o, d, t);
                    } else {
                        const SphereRec sp_ = s.spheres[node.indexOffset + i];
                        id = node.indexOffset + i; kind = HIT_SPHERE;
                        ok = sphere_intersect(mk3(sp_.cx, sp_.cy, sp_.cz), sp_.sqRadius, o, d, t);
                    }
                    if (ok && t < h.t) {
                        h.t = t; h.kind = kind; h.idx = id;
                        if (shadow) { return true; }
                    }
                }
                --sp; boxIndex = stack[sp];
            } else {
                const int left = node.indexOffset, right = node.indexOffset + 1;
                const RefNode l = nodes[left], r = nodes[right];
                const bool tl = aabb_intersect_exact(mk3(l.mnx, l.mny, l.mnz), mk3(l.mxx, l.mxy, l.mxz), o, d);
                const bool tr = aabb_intersect_exact(mk3(r.mnx, r.mny, r.mnz), mk3(r.mxx, r.mxy, r.mxz), o, d);
                if (!tl && !tr) {
                    --sp; boxIndex = stack[sp];
                } else {
                    boxIndex = tl ? left : right;
                    if (tl && tr) {
                        if (sp >= kStackRef) { *overflow = 1; return false; }
                        stack[sp++] = right;
                    }
                }
            }
        } else {
            --sp; boxIndex = stack[sp];
        }
    } while (sp > 0);
    return false;
}

// ---- compressed 8-wide BVH ----------------------------------------------------------------------------------
MRT_D float u8f(uint32_t word, int byte) { return static_cast<float>((word >> (8 * byte)) & 0xffu); }

// ANY: stop at the first hit with t < h.t (shadow rays; area-light triangles are not occluders, Shader.cpp:132-158).
// skipPrim: primitive id the ray leaves (Triangle.cpp:64), or -1.  bestPrim: in/out, -1 when the current best hit
// is not a BVH8 primitive.
template <bool ANY>
MRT_D bool trace_bvh8(const SceneDev &s, v3 o, v3 d, int skipPrim, float &tbest, int &bestPrim, float &bu, float &bv, int *overflow) {
    if (s.n_bvh_prims == 0) { return false; }
    // robust reciprocal: a zero component becomes a huge finite slope so that slabs stay well defined
    const float big = 1.0e20f;
    const float idx = fabsf(d.x) > 1.0e-20f ? 1.0f / d.x : (d.x < 0.0f ? -big : big);
    const float idy = fabsf(d.y) > 1.0e-20f ? 1.0f / d.y : (d.y < 0.0f ? -big : big);
    const float idz = fabsf(d.z) > 1.0e-20f ? 1.0f / d.z : (d.z < 0.0f ? -big : big);
    const uint32_t oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    const uint32_t octinv = 7u - oct;

    uint2 stack[kStack8];
    int sp = 0;
    uint2 ngroup = make_uint2(0u, 0x80000000u);
    const float4 *nodes4 = reinterpret_cast<const float4 *>(s.nodes);
    const float4 *tris4 = reinterpret_cast<const float4 *>(s.tri_geom);

    while (true) {
        uint2 tgroup = make_uint2(0u, 0u);
        if (ngroup.y & 0xff000000u) {
            const uint32_t bit = 31u - __clz(ngroup.y);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y & 0xff000000u) {
                if (sp >= kStack8) { *overflow = 1; return false; }
                stack[sp++] = ngroup;
            }
            const uint32_t slot = (bit - 24u) ^ octinv;
            const uint32_t rel = __popc(ngroup.y & ((1u << slot) - 1u) & 0xffu);
            const float4 *np = nodes4 + 5ull * (ngroup.x + rel);
            const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            // n0 = px py pz (ex ey ez imask) | n1 = child_base tri_base meta[0..3] meta[4..7]
            // n2 = qlox[0..7] qloy[0..7] | n3 = qloz[0..7] qhix[0..7] | n4 = qhiy[0..7] qhiz[0..7]
            const uint32_t e = __float_as_uint(n0.w);
            const float sx = __uint_as_float((e & 0xffu) << 23), sy = __uint_as_float(((e >> 8) & 0xffu) << 23), sz = __uint_as_float(((e >> 16) & 0xffu) << 23);
            const uint32_t imask = e >> 24;
            const float ax = sx * idx, ay = sy * idy, az = sz * idz;
            const float ox = (n0.x - o.x) * idx, oy = (n0.y - o.y) * idy, oz = (n0.z - o.z) * idz;
            // conservative slabs: the offsets are widened by their own rounding error (relative 1e-6), the final
            // comparison by the rounding of the fused multiply-add itself
            const float sxo = 1.0e-6f * fabsf(ox), syo = 1.0e-6f * fabsf(oy), szo = 1.0e-6f * fabsf(oz);
            const float ox0 = ox - sxo, ox1 = ox + sxo, oy0 = oy - syo, oy1 = oy + syo, oz0 = oz - szo, oz1 = oz + szo;
            const uint32_t meta[2] = {__float_as_uint(n1.z), __float_as_uint(n1.w)};
            const uint32_t qlx[2] = {__float_as_uint(n2.x), __float_as_uint(n2.y)}, qly[2] = {__float_as_uint(n2.z), __float_as_uint(n2.w)};
            const uint32_t qlz[2] = {__float_as_uint(n3.x), __float_as_uint(n3.y)}, qhx[2] = {__float_as_uint(n3.z), __float_as_uint(n3.w)};
            const uint32_t qhy[2] = {__float_as_uint(n4.x), __float_as_uint(n4.y)}, qhz[2] = {__float_as_uint(n4.z), __float_as_uint(n4.w)};
            uint32_t hitmask = 0u;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int w = c >> 2, b = c & 3;
                const uint32_t m = (meta[w] >> (8 * b)) & 0xffu;
                if (m == 0u) { continue; }
                const float lx = u8f(qlx[w], b), hx = u8f(qhx[w], b);
                const float ly = u8f(qly[w], b), hy = u8f(qhy[w], b);
                const float lz = u8f(qlz[w], b), hz = u8f(qhz[w], b);
                const float t0x = __fmaf_rn(oct & 1u ? hx : lx, ax, ox0), t1x = __fmaf_rn(oct & 1u ? lx : hx, ax, ox1);
                const float t0y = __fmaf_rn(oct & 2u ? hy : ly, ay, oy0), t1y = __fmaf_rn(oct & 2u ? ly : hy, ay, oy1);
                const float t0z = __fmaf_rn(oct & 4u ? hz : lz, az, oz0), t1z = __fmaf_rn(oct & 4u ? lz : hz, az, oz1);
                const float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
                const float tmax = fminf(fminf(t1x, t1y), fminf(t1z, tbest));
                if (tmin <= tmax * 1.000001f) {
                    const uint32_t inner = (m & 0x18u) == 0x18u ? 1u : 0u;   // low 5 bits >= 24
                    const uint32_t bitIndex = inner ? (24u + (static_cast<uint32_t>(c) ^ octinv)) : (m & 0x1fu);
                    hitmask |= (inner ? 1u : (m >> 5)) << bitIndex;
                }
            }
            ngroup.x = __float_as_uint(n1.x);
            ngroup.y = (hitmask & 0xff000000u) | imask;
            tgroup.x = __float_as_uint(n1.y);
            tgroup.y = hitmask & 0x00ffffffu;
        } else {
            tgroup = ngroup;
            ngroup = make_uint2(0u, 0u);
        }
        while (tgroup.y) {
            const uint32_t i = __ffs(tgroup.y) - 1u;
            tgroup.y &= tgroup.y - 1u;
            const float4 *tp = tris4 + 3ull * (tgroup.x + i);
            const float4 g0 = __ldg(tp), g1 = __ldg(tp + 1), g2 = __ldg(tp + 2);
            const int prim = static_cast<int>(__float_as_uint(g2.y));
            if (prim == skipPrim) { continue; }
            if (ANY && prim >= s.n_triangles) { continue; }
            float t, u, v;
            if (triangle_intersect(mk3(g0.x, g0.y, g0.z), mk3(g0.w, g1.x, g1.y), mk3(g1.z, g1.w, g2.x), o, d, t, u, v)) {
                if (t < tbest || (!ANY && t == tbest && bestPrim >= 0 && prim < bestPrim)) {
                    tbest = t; bestPrim = prim; bu = u; bv = v;
                    if (ANY) { return true; }
                }
            }
        }
        if (!(ngroup.y & 0xff000000u)) {
            if (sp == 0) { break; }
            ngroup = stack[--sp];
        }
    }
    return false;
}

// The reference tests the root box of BVH<Triangle> with AABB::intersect before anything else (BVH.hpp:341-342); that
// test is topology independent, and its NaN rule (an axis-parallel ray whose origin lies exactly on a box plane
// misses, AABB.cpp:34-54) is visible in images of the orthographic scenes, so it is reproduced exactly here.
MRT_D bool triangles_root_hit(const SceneDev &s, v3 o, v3 d) {
    return s.n_triangles > 0 && aabb_intersect_exact(s.tri_bmin, s.tri_bmax, o, d);
}

// Shader::traceLights (Shader.cpp:166-171) as a plain loop: only used for the rare rays that fail the root test above.
MRT_D void trace_lights_linear(const SceneDev &s, v3 o, v3 d, HitRec &h) {
    for (int i = 0; i < s.n_lights; ++i) {
        const LightRec &l = s.lights[i];
        if (l.kind == 0) { continue; }
        float t, u, v;
        if (triangle_intersect(mk3(l.A[0], l.A[1], l.A[2]), mk3(l.AB[0], l.AB[1], l.AB[2]), mk3(l.AC[0], l.AC[1], l.AC[2]), o, d, t, u, v) && t < h.t) {
            h.t = t; h.kind = HIT_LIGHT; h.idx = i; h.u = u; h.v = v;
        }
    }
}

// Shader::rayTrace's geometric part (Shader.cpp:86-111): planes, then spheres, then triangles share one running t;
// area lights are tested last with the same strict "<" rule, which the tie rule of trace_bvh8 reproduces because
// light primitive ids come after all scene triangles.
MRT_D void trace_closest(const SceneDev &s, v3 o, v3 d, int skipKind, int skipIdx, HitRec &h, int *overflow) {
    h.t = kRayLengthMax; h.kind = HIT_NONE; h.idx = -1; h.u = 0.0f; h.v = 0.0f;
    trace_ref_bvh<0>(s, o, d, false, skipKind, skipIdx, h, overflow);
    trace_ref_bvh<1>(s, o, d, false, skipKind, skipIdx, h, overflow);
    if (s.n_bvh_prims == 0) { return; }
    if (!triangles_root_hit(s, o, d)) {
        if (s.n_bvh_prims > s.n_triangles) { trace_lights_linear(s, o, d, h); }
        return;
    }
    int bestPrim = -1;
    trace_bvh8<false>(s, o, d, skipKind == HIT_TRIANGLE ? skipIdx : -1, h.t, bestPrim, h.u, h.v, overflow);
    if (bestPrim >= 0) {
        if (bestPrim < s.n_triangles) { h.kind = HIT_TRIANGLE; h.idx = bestPrim; }
        else { h.kind = HIT_LIGHT; h.idx = bestPrim - s.n_triangles; }
    }
}

// Shader::shadowTrace (Shader.cpp:132-158): true when something lies within `distance`.
MRT_D bool trace_shadow(const SceneDev &s, v3 o, v3 d, float distance, int skipKind, int skipIdx, int *overflow) {
    HitRec h;
    h.t = distance; h.kind = HIT_NONE; h.idx = -1; h.u = 0.0f; h.v = 0.0f;
    if (trace_ref_bvh<0>(s, o, d, true, skipKind, skipIdx, h, overflow)) { return true; }
    if (trace_ref_bvh<1>(s, o, d, true, skipKind, skipIdx, h, overflow)) { return true; }
    if (!triangles_root_hit(s, o, d)) { return false; }
    int bestPrim = -1;
    return trace_bvh8<true>(s, o, d, skipKind == HIT_TRIANGLE ? skipIdx : -1, h.t, bestPrim, h.u, h.v, overflow);
}

}  // namespace mrt
