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
constexpr int kPend = 6;        // parked triangle groups per thread

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
// Traversal state of one ray, advanced in steps so that the same code serves the per-thread callers (megakernel, tests)
// and the persistent wavefront kernels that refill idle lanes between steps (mrt_wavefront.cu).
// Instruction budget (ncu, profiles/): the node test is issue bound, split between the ALU pipe (PRMT, FMNMX, FSETP)
// and the FMA pipe (FADD, FFMA); the 8-bit plane coordinates are turned into floats with PRMT + a magic-number FADD
// instead of I2F (I2F.U8 runs on the quarter-rate XU pipe and was the top stall), and everything is branch free per child.
struct Trav8 {
    v3 o, d;
    float idx, idy, idz;       // approximate reciprocals (slab tests are widened below, so 1 ulp does not matter)
    uint32_t oct, octinv4;     // direction octant; (7 - oct) replicated into four bytes
    int skipPrim;
    float tbest; int bestPrim; float bu, bv;
    uint2 ngroup, tgroup;      // x = base index, y = pending bits (node group: inner hits in 31..24, imask in 7..0)
    int sp, np;
    uint2 pend[kPend];         // parked leaf groups (x = triangle base, y = pending triangle bits)
#ifdef MRT_TRAVERSAL_STATS
    uint32_t n_nodes, n_tris;  // diagnostic build only (libmobilert_b200_stats.so): node visits / triangle tests of this ray
#endif
    uint2 stack[kStack8];
};

#ifdef MRT_TRAVERSAL_STATS
static __device__ unsigned long long g_stat_nodes, g_stat_tris, g_stat_rays;   // one copy per translation unit
#define MRT_STAT(x) x
#else
#define MRT_STAT(x)
#endif

MRT_D float rcp_robust(float d) {
    // a zero component becomes a huge finite slope so that slabs stay well defined
    const float big = 1.0e20f;
    return fabsf(d) > 1.0e-20f ? __frcp_rn(d) : (d < 0.0f ? -big : big);
}

MRT_D void trav8_init(Trav8 &t, v3 o, v3 d, int skipPrim, float tbest) {
    t.o = o; t.d = d;
    t.idx = rcp_robust(d.x); t.idy = rcp_robust(d.y); t.idz = rcp_robust(d.z);
    t.oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    t.octinv4 = (7u - t.oct) * 0x01010101u;
    t.skipPrim = skipPrim;
    t.tbest = tbest; t.bestPrim = -1; t.bu = 0.0f; t.bv = 0.0f;
    t.ngroup = make_uint2(0u, 0x80000000u);   // the root
    t.tgroup = make_uint2(0u, 0u);
    t.sp = 0; t.np = 0;
    MRT_STAT(t.n_nodes = 0u; t.n_tris = 0u;)
}

// byte b of w -> float, via 0x4B0000bb = 2^23 + bb
template <int B> MRT_D float u8_to_float(uint32_t w) {
    return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7440u | B)) - 8388608.0f;
}

// Four children (bytes 0..3 of the quantised plane words) against the ray; returns their contribution to the hit mask.
template <int B>
MRT_D uint32_t child_hit(uint32_t nx, uint32_t ny, uint32_t nz, uint32_t fx, uint32_t fy, uint32_t fz, float ax, float ay, float az,
                         float ox0, float oy0, float oz0, float ox1, float oy1, float oz1, float tbest, uint32_t bits4, uint32_t index4) {
    const float t0x = __fmaf_rn(u8_to_float<B>(nx), ax, ox0), t0y = __fmaf_rn(u8_to_float<B>(ny), ay, oy0), t0z = __fmaf_rn(u8_to_float<B>(nz), az, oz0);
    const float t1x = __fmaf_rn(u8_to_float<B>(fx), ax, ox1), t1y = __fmaf_rn(u8_to_float<B>(fy), ay, oy1), t1z = __fmaf_rn(u8_to_float<B>(fz), az, oz1);
    const float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
    const float tmax = fminf(fminf(t1x, t1y), fminf(t1z, tbest));
    const uint32_t count = (bits4 >> (8 * B)) & 0xffu, index = (index4 >> (8 * B)) & 0x1fu;   // count: 1 for an inner child, 1..7 triangles for a leaf
    return tmin <= tmax * 1.000001f ? ((1u << count) - 1u) << index : 0u;
}

// Visits the next pending inner node of t.ngroup: afterwards t.ngroup holds that node's hit children and
// `tnew` the hit leaf triangles (the previous group, if it still has pending nodes, goes to the stack).
MRT_D bool trav8_node_step(const SceneDev &s, Trav8 &t, uint2 &tnew, int *overflow) {
    MRT_STAT(++t.n_nodes;)
    const uint32_t bit = 31u - __clz(t.ngroup.y);
    t.ngroup.y &= ~(1u << bit);
    if (t.ngroup.y & 0xff000000u) {
        if (t.sp >= kStack8) { *overflow = 1; return false; }
        t.stack[t.sp++] = t.ngroup;
    }
    const uint32_t slot = (bit - 24u) ^ (t.octinv4 & 0xffu);
    const uint32_t rel = __popc(t.ngroup.y & ((1u << slot) - 1u) & 0xffu);
    const float4 *np = reinterpret_cast<const float4 *>(s.nodes) + 5ull * (t.ngroup.x + rel);
    const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
    // n0 = px py pz (ex ey ez imask) | n1 = child_base tri_base meta[0..3] meta[4..7]
    // n2 = qlox[0..7] qloy[0..7] | n3 = qloz[0..7] qhix[0..7] | n4 = qhiy[0..7] qhiz[0..7]
    const uint32_t e = __float_as_uint(n0.w);
    const float ax = __uint_as_float((e & 0xffu) << 23) * t.idx, ay = __uint_as_float(((e >> 8) & 0xffu) << 23) * t.idy,
                az = __uint_as_float(((e >> 16) & 0xffu) << 23) * t.idz;
    const float ox = (n0.x - t.o.x) * t.idx, oy = (n0.y - t.o.y) * t.idy, oz = (n0.z - t.o.z) * t.idz;
    // conservative slabs: the offsets are widened by their own rounding error (relative 1e-6), the final comparison
    // by the rounding of the fused multiply-adds
    const float sxo = 1.0e-6f * fabsf(ox), syo = 1.0e-6f * fabsf(oy), szo = 1.0e-6f * fabsf(oz);
    const float ox0 = ox - sxo, ox1 = ox + sxo, oy0 = oy - syo, oy1 = oy + syo, oz0 = oz - szo, oz1 = oz + szo;
    // near / far plane words per axis, chosen once per node by the ray octant
    const bool fx_ = t.oct & 1u, fy_ = t.oct & 2u, fz_ = t.oct & 4u;
    const uint32_t qlx0 = __float_as_uint(n2.x), qlx1 = __float_as_uint(n2.y), qly0 = __float_as_uint(n2.z), qly1 = __float_as_uint(n2.w);
    const uint32_t qlz0 = __float_as_uint(n3.x), qlz1 = __float_as_uint(n3.y), qhx0 = __float_as_uint(n3.z), qhx1 = __float_as_uint(n3.w);
    const uint32_t qhy0 = __float_as_uint(n4.x), qhy1 = __float_as_uint(n4.y), qhz0 = __float_as_uint(n4.z), qhz1 = __float_as_uint(n4.w);
    const uint32_t nx0 = fx_ ? qhx0 : qlx0, nx1 = fx_ ? qhx1 : qlx1, fx0 = fx_ ? qlx0 : qhx0, fx1 = fx_ ? qlx1 : qhx1;
    const uint32_t ny0 = fy_ ? qhy0 : qly0, ny1 = fy_ ? qhy1 : qly1, fy0 = fy_ ? qly0 : qhy0, fy1 = fy_ ? qly1 : qhy1;
    const uint32_t nz0 = fz_ ? qhz0 : qlz0, nz1 = fz_ ? qhz1 : qlz1, fz0 = fz_ ? qlz0 : qhz0, fz1 = fz_ ? qlz1 : qhz1;
    const uint32_t imask = e >> 24;
    uint32_t hitmask = 0u;
#pragma unroll
    for (int w = 0; w < 2; ++w) {
        const uint32_t meta4 = __float_as_uint(w == 0 ? n1.z : n1.w);
        // per byte: inner children carry (1 << 5) | (24 + slot); their bit index is XORed with the inverse octant so that the
        // highest pending bit is always the nearest child along the ray. Leaves carry (triangle count << 5) | triangle offset.
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
        const uint32_t index4 = (meta4 ^ (t.octinv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t nx = w == 0 ? nx0 : nx1, ny = w == 0 ? ny0 : ny1, nz = w == 0 ? nz0 : nz1;
        const uint32_t fx = w == 0 ? fx0 : fx1, fy = w == 0 ? fy0 : fy1, fz = w == 0 ? fz0 : fz1;
        hitmask |= child_hit<0>(nx, ny, nz, fx, fy, fz, ax, ay, az, ox0, oy0, oz0, ox1, oy1, oz1, t.tbest, bits4, index4);
        hitmask |= child_hit<1>(nx, ny, nz, fx, fy, fz, ax, ay, az, ox0, oy0, oz0, ox1, oy1, oz1, t.tbest, bits4, index4);
        hitmask |= child_hit<2>(nx, ny, nz, fx, fy, fz, ax, ay, az, ox0, oy0, oz0, ox1, oy1, oz1, t.tbest, bits4, index4);
        hitmask |= child_hit<3>(nx, ny, nz, fx, fy, fz, ax, ay, az, ox0, oy0, oz0, ox1, oy1, oz1, t.tbest, bits4, index4);
    }
    t.ngroup.x = __float_as_uint(n1.x);
    t.ngroup.y = (hitmask & 0xff000000u) | imask;
    tnew.x = __float_as_uint(n1.y);
    tnew.y = hitmask & 0x00ffffffu;
    return true;
}

// Tests the next pending triangle of t.tgroup. Returns true when an ANY-hit ray found its occluder.
// Triangle::intersect (Triangle.cpp:63-99) with a sound prefilter: the sign / magnitude tests below only reject what the
// reference's own comparisons reject for certain; survivors run the reference's exact arithmetic (one IEEE division).
template <bool ANY>
MRT_D bool trav8_triangle_step(const SceneDev &s, Trav8 &t) {
    const uint32_t i = __ffs(t.tgroup.y) - 1u;
    t.tgroup.y &= t.tgroup.y - 1u;
    const float4 *tp = reinterpret_cast<const float4 *>(s.tri_geom) + 3ull * (t.tgroup.x + i);
    const float4 g0 = __ldg(tp), g1 = __ldg(tp + 1), g2 = __ldg(tp + 2);
    const int prim = static_cast<int>(__float_as_uint(g2.y));
    if (prim == t.skipPrim) { return false; }
    if (ANY && prim >= s.n_triangles) { return false; }      // area lights do not occlude (Shader.cpp:132-158)
    MRT_STAT(++t.n_tris;)
    const v3 A = mk3(g0.x, g0.y, g0.z), AB = mk3(g0.w, g1.x, g1.y), AC = mk3(g1.z, g1.w, g2.x);
    const v3 perpendicularVector = cross(t.d, AC);
    const float det = dot(AB, perpendicularVector);
    const float adet = fabsf(det);
    if (adet < kEpsilon) { return false; }
    const uint32_t sgn = __float_as_uint(det) & 0x80000000u;
    const float hi = adet * 1.00001f;
    const v3 vectorToCamera = t.o - A;
    const float du = dot(vectorToCamera, perpendicularVector);
    const float su = __uint_as_float(__float_as_uint(du) ^ sgn);
    if (su < -1.0e-25f || su > hi) { return false; }
    const v3 upPerpendicularVector = cross(vectorToCamera, AB);
    const float dv = dot(t.d, upPerpendicularVector);
    const float sv = __uint_as_float(__float_as_uint(dv) ^ sgn);
    if (sv < -1.0e-25f || su + sv > hi) { return false; }
    const float dt = dot(AC, upPerpendicularVector);
    const float st = __uint_as_float(__float_as_uint(dt) ^ sgn);
    if (st < -1.0e-25f || st > hi * fminf(t.tbest, 1.0e25f)) { return false; }
    // exact path
    const float inv = 1.0f / det;
    const float u = inv * du;
    if (u < 0.0f || u > 1.0f) { return false; }
    const float v = inv * dv;
    if (v < 0.0f || (u + v) > 1.0f) { return false; }
    const float tt = inv * dt;
    if (tt < kEpsilon) { return false; }
    if (tt < t.tbest || (!ANY && tt == t.tbest && t.bestPrim >= 0 && prim < t.bestPrim)) {
        t.tbest = tt; t.bestPrim = prim; t.bu = u; t.bv = v;
        return ANY;
    }
    return false;
}

// One scheduling step for this lane: a node visit when inner children are pending, then - only when the warp agrees -
// the triangle phase. A triangle test costs about a fifth of a node test, but after any single node step only a few
// lanes have hit a leaf (ncu, profiles/: 2-4 of 32 threads active in the triangle code when it ran after every step).
// So hit leaf groups are parked in t.pend and the warp enters the triangle phase together, when at least
// s.tune_postpone lanes have parked work, when nobody has node work left, or when this lane's parking space is full.
// Deferring a test only delays the shrinking of tbest; the result does not depend on the order (exact-t ties go to the
// lower primitive id). Returns false when the ray is finished (ANY: occluded iff bestPrim >= 0).
constexpr int kPostponeLanes = 12;
template <bool ANY>
MRT_D bool trav8_step(const SceneDev &s, Trav8 &t, int *overflow) {
    if (!(t.ngroup.y & 0xff000000u) && t.sp > 0) { t.ngroup = t.stack[--t.sp]; }
    bool node_work = (t.ngroup.y & 0xff000000u) != 0u;
    if (node_work) {
        uint2 tnew;
        if (!trav8_node_step(s, t, tnew, overflow)) { return false; }
        if (tnew.y) { t.pend[t.np++] = tnew; }
        node_work = (t.ngroup.y & 0xff000000u) != 0u || t.sp > 0;
    }
    const uint32_t mask = __activemask();
    const uint32_t ready = __ballot_sync(mask, t.np > 0);
    const uint32_t busy = __ballot_sync(mask, node_work);
    const uint32_t full = __ballot_sync(mask, t.np >= kPend);
    if (__popc(ready) >= s.tune_postpone || busy == 0u || full != 0u) {
        if (t.np > 0) {
            t.tgroup = t.pend[--t.np];
            while (t.tgroup.y) {
                if (trav8_triangle_step<ANY>(s, t)) { return false; }
            }
        }
    }
    return node_work || t.np > 0;
}

// ANY: stop at the first hit with t < tbest (shadow rays). skipPrim: primitive id the ray leaves (Triangle.cpp:64), or -1.
// bestPrim: out, -1 when nothing nearer than the incoming tbest was found.
template <bool ANY>
MRT_D bool trace_bvh8(const SceneDev &s, v3 o, v3 d, int skipPrim, float &tbest, int &bestPrim, float &bu, float &bv, int *overflow) {
    if (s.n_bvh_prims == 0) { return false; }
    Trav8 t;
    trav8_init(t, o, d, skipPrim, tbest);
    while (trav8_step<ANY>(s, t, overflow)) { }
    MRT_STAT(atomicAdd(&g_stat_nodes, t.n_nodes); atomicAdd(&g_stat_tris, t.n_tris); atomicAdd(&g_stat_rays, 1ull);)
    if (t.bestPrim >= 0) { tbest = t.tbest; bestPrim = t.bestPrim; bu = t.bu; bv = t.bv; }
    return ANY && t.bestPrim >= 0;
}

// The reference tests the root box of BVH<Triangle> with AABB::intersect before anything else (BVH.hpp:341-342); that
// test is topology independent, and its NaN rule (an axis-parallel ray whose origin lies exactly on a box plane
// misses, AABB.cpp:34-54) is visible in images of the orthographic scenes, so it is reproduced exactly here.
MRT_D bool triangles_root_hit(const SceneDev &s, v3 o, v3 d) {
    if (s.n_triangles <= 0) { return false; }
    // without a zero direction component the test has no NaN case and only rejects rays that cannot hit anything inside
    if (d.x != 0.0f && d.y != 0.0f && d.z != 0.0f) { return true; }
    return aabb_intersect_exact(s.tri_bmin, s.tri_bmax, o, d);
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
