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
#ifdef MRT_TRAVERSAL_STATS
    uint32_t n_nodes, n_tris;  // diagnostic build only (libmobilert_b200_stats.so): node visits / triangle tests of this ray
#endif
};
// The indexed arrays live outside Trav8 so that its scalar fields stay in registers (with the arrays inside, the whole
// struct sat in local memory: 9 LDL per node step in the first ncu SASS view).
struct TravMem {
    uint2 stack[kStack8 + 1];  // node groups still to visit; entry kStack8 only ever receives the unconditional store of a full stack
    uint2 pend[kPend];         // parked leaf groups (x = triangle base, y = pending triangle bits); per-thread callers only
};
// (a shared-memory short stack for the first entries was measured slower in every size, profiles/r1_ab_short_stack_s43.txt)
MRT_D void stack_put(TravMem &tm, int i, uint2 v) { tm.stack[i] = v; }
MRT_D uint2 stack_get(const TravMem &tm, int i) { return tm.stack[i]; }

#ifdef MRT_TRAVERSAL_STATS
static __device__ unsigned long long g_stat_nodes, g_stat_tris, g_stat_rays;   // one copy per translation unit
static __device__ unsigned long long g_stat_fruitless, g_stat_stale;           // node visits that hit no child / that hit none only because of tbest
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

// byte B of w -> the float 1 + b * 2^-15 (0x3F80bb00): one PRMT, no integer-to-float conversion and no bias subtraction;
// the "1 +" is folded into the slab offset once per node and axis (see trav8_node_step). `one` is 0x3F800000 held in a
// register the compiler cannot see through (SceneDev::prmt_one, a kernel parameter set by the host): PRMT takes one immediate, and with a known constant there nvcc keeps the
// 48 selectors of a node step in registers instead (40 extra MOV / IMAD per step in the first version, ncu SASS view).
template <int B> MRT_D float u8_to_unit(uint32_t w, uint32_t one) {
    return __uint_as_float(__byte_perm(w, one, 0x7604u | (B << 4)));
}
// One child (byte B of the quantised plane words) against the ray.
template <int B>
MRT_D bool child_hit(uint32_t nx, uint32_t ny, uint32_t nz, uint32_t fx, uint32_t fy, uint32_t fz, float ax, float ay, float az,
                     float ox0, float oy0, float oz0, float ox1, float oy1, float oz1, float tbest, uint32_t one) {
    const float t0x = __fmaf_rn(u8_to_unit<B>(nx, one), ax, ox0), t0y = __fmaf_rn(u8_to_unit<B>(ny, one), ay, oy0), t0z = __fmaf_rn(u8_to_unit<B>(nz, one), az, oz0);
    const float t1x = __fmaf_rn(u8_to_unit<B>(fx, one), ax, ox1), t1y = __fmaf_rn(u8_to_unit<B>(fy, one), ay, oy1), t1z = __fmaf_rn(u8_to_unit<B>(fz, one), az, oz1);
    const float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
    const float tmax = fminf(fminf(t1x, t1y), fminf(t1z, tbest));
    return tmin <= tmax;
}

#ifndef MRT_NO_SIGN_PACK
#define MRT_SIGN_PACK   // measured on the bench frame: extend -1.3 %, shadow -0.9 % (profiles/r2_ab_sign_pack.txt)
#endif
#ifdef MRT_SIGN_PACK
// The same test as a number: tmax - tmin, whose sign bit is clear exactly when the child is hit (x - x is +0; all operands are
// finite, see rcp_robust). The caller shifts the eight sign bits into one word with one funnel shift each, which replaces
// FSETP + a predicated add per child (both on the ALU pipe that bounds the traversal kernels) by FADD (FMA pipe) + SHF.
template <int B>
MRT_D uint32_t child_gap(uint32_t nx, uint32_t ny, uint32_t nz, uint32_t fx, uint32_t fy, uint32_t fz, float ax, float ay, float az,
                         float ox0, float oy0, float oz0, float ox1, float oy1, float oz1, float tbest, uint32_t one) {
    const float t0x = __fmaf_rn(u8_to_unit<B>(nx, one), ax, ox0), t0y = __fmaf_rn(u8_to_unit<B>(ny, one), ay, oy0), t0z = __fmaf_rn(u8_to_unit<B>(nz, one), az, oz0);
    const float t1x = __fmaf_rn(u8_to_unit<B>(fx, one), ax, ox1), t1y = __fmaf_rn(u8_to_unit<B>(fy, one), ay, oy1), t1z = __fmaf_rn(u8_to_unit<B>(fz, one), az, oz1);
    const float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
    const float tmax = fminf(fminf(t1x, t1y), fminf(t1z, tbest));
    return __float_as_uint(tmax - tmin);
}
#endif

// Visits the next pending inner node of t.ngroup: afterwards t.ngroup holds that node's hit children and
// `tnew` the hit leaf triangles (the previous group, if it still has pending nodes, goes to the stack).
//
// Slab arithmetic per axis (s = 2^e the node's quantisation step, q the 8-bit plane, r = 1/d):
//   t = (p + q s - o) r = m A + O   with m = 1 + q 2^-15 (exact, from PRMT), A = 2^15 s r, O = (p - o) r - A.
// The node stores e + 15, so A is one shift, one mask and one multiply. Rounding: r, (p - o) r and O are each off by at
// most 2^-24 relative to |(p - o) r| + |A|, the fused multiply-add by 2^-24 |t|; the offsets are therefore widened by
// 1e-6 (|(p - o) r| + |A|) (about 2^-20, four times the bound) on the near side downwards and on the far side upwards.
// The test is conservative: it can only accept more children than exact arithmetic would, never fewer.
// Core: leaves the hit leaf slots in `lh` (bit k = slot k) with the node's meta words m0 / m1 and triangle base.
// `perm`: optional 8 x 256 byte table perm[o * 256 + m] = the mask m with every bit k moved to k ^ o (shared memory of the
// wavefront kernels: one LDS instead of three predicated swaps on the ALU pipe, which bounds those kernels).
MRT_D bool trav8_node_core(const SceneDev &s, Trav8 &t, TravMem &tm_, uint32_t &tbase, uint32_t &lh, uint32_t &m0, uint32_t &m1, int *overflow,
                           const uint8_t *perm = nullptr) {
    MRT_STAT(++t.n_nodes;)
    const uint32_t bit = 31u - __clz(t.ngroup.y);
    t.ngroup.y &= ~(1u << bit);
    {   // the rest of the group goes to the stack: an unconditional store plus a conditional bump of sp (no divergent branch)
        const bool push = (t.ngroup.y & 0xff000000u) != 0u;
        if (push && t.sp >= kStack8) { *overflow = 1; return false; }
        stack_put(tm_, t.sp < kStack8 ? t.sp : kStack8, t.ngroup);   // sp == kStack8 without a push: a scratch slot, never popped
        t.sp += push ? 1 : 0;
    }
    const uint32_t one = s.prmt_one;
    const uint32_t octinv = t.octinv4 & 0xffu;
    const uint32_t slot = (bit - 24u) ^ octinv;
    const uint32_t rel = __popc(t.ngroup.y & ((1u << slot) - 1u) & 0xffu);
    const float4 *np = reinterpret_cast<const float4 *>(s.nodes) + 5ull * (t.ngroup.x + rel);
    const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
    // n0 = px py pz (ex+15 ey+15 ez+15 imask) | n1 = child_base tri_base meta[0..3] meta[4..7]
    // n2 = qlox[0..7] qloy[0..7] | n3 = qloz[0..7] qhix[0..7] | n4 = qhiy[0..7] qhiz[0..7]
    const uint32_t e = __float_as_uint(n0.w);
    const float ax = __uint_as_float((e << 23) & 0x7f800000u) * t.idx, ay = __uint_as_float((e << 15) & 0x7f800000u) * t.idy,
                az = __uint_as_float((e << 7) & 0x7f800000u) * t.idz;
    const float px = (n0.x - t.o.x) * t.idx, py = (n0.y - t.o.y) * t.idy, pz = (n0.z - t.o.z) * t.idz;
    const float ox = px - ax, oy = py - ay, oz = pz - az;
    const float wx = 1.0e-6f * (fabsf(px) + fabsf(ax)), wy = 1.0e-6f * (fabsf(py) + fabsf(ay)), wz = 1.0e-6f * (fabsf(pz) + fabsf(az));
    const float ox0 = ox - wx, ox1 = ox + wx, oy0 = oy - wy, oy1 = oy + wy, oz0 = oz - wz, oz1 = oz + wz;
    // near / far plane words per axis, chosen once per node by the ray octant
    const bool fx_ = t.oct & 1u, fy_ = t.oct & 2u, fz_ = t.oct & 4u;
    const uint32_t qlx0 = __float_as_uint(n2.x), qlx1 = __float_as_uint(n2.y), qly0 = __float_as_uint(n2.z), qly1 = __float_as_uint(n2.w);
    const uint32_t qlz0 = __float_as_uint(n3.x), qlz1 = __float_as_uint(n3.y), qhx0 = __float_as_uint(n3.z), qhx1 = __float_as_uint(n3.w);
    const uint32_t qhy0 = __float_as_uint(n4.x), qhy1 = __float_as_uint(n4.y), qhz0 = __float_as_uint(n4.z), qhz1 = __float_as_uint(n4.w);
    const uint32_t nx0 = fx_ ? qhx0 : qlx0, nx1 = fx_ ? qhx1 : qlx1, fx0 = fx_ ? qlx0 : qhx0, fx1 = fx_ ? qlx1 : qhx1;
    const uint32_t ny0 = fy_ ? qhy0 : qly0, ny1 = fy_ ? qhy1 : qly1, fy0 = fy_ ? qly0 : qhy0, fy1 = fy_ ? qly1 : qhy1;
    const uint32_t nz0 = fz_ ? qhz0 : qlz0, nz1 = fz_ ? qhz1 : qlz1, fz0 = fz_ ? qlz0 : qhz0, fz1 = fz_ ? qlz1 : qhz1;
    const uint32_t imask = e >> 24;
#ifdef MRT_SIGN_PACK
    uint32_t miss = 0u;   // child 7 first, so that child k ends up in bit k
#define MRT_CHILD(B, NX, NY, NZ, FX, FY, FZ) \
    miss = __funnelshift_l(child_gap<B>(NX, NY, NZ, FX, FY, FZ, ax, ay, az, ox0, oy0, oz0, ox1, oy1, oz1, t.tbest, one), miss, 1);
    MRT_CHILD(3, nx1, ny1, nz1, fx1, fy1, fz1) MRT_CHILD(2, nx1, ny1, nz1, fx1, fy1, fz1)
    MRT_CHILD(1, nx1, ny1, nz1, fx1, fy1, fz1) MRT_CHILD(0, nx1, ny1, nz1, fx1, fy1, fz1)
    MRT_CHILD(3, nx0, ny0, nz0, fx0, fy0, fz0) MRT_CHILD(2, nx0, ny0, nz0, fx0, fy0, fz0)
    MRT_CHILD(1, nx0, ny0, nz0, fx0, fy0, fz0) MRT_CHILD(0, nx0, ny0, nz0, fx0, fy0, fz0)
#undef MRT_CHILD
    const uint32_t hits = ~miss & 0xffu;   // bit k: child slot k is hit
#else
    uint32_t hits = 0u;   // bit k: child slot k is hit
#define MRT_CHILD(B, NX, NY, NZ, FX, FY, FZ, K) \
    if (child_hit<B>(NX, NY, NZ, FX, FY, FZ, ax, ay, az, ox0, oy0, oz0, ox1, oy1, oz1, t.tbest, one)) { hits |= 1u << (K); }
    MRT_CHILD(0, nx0, ny0, nz0, fx0, fy0, fz0, 0) MRT_CHILD(1, nx0, ny0, nz0, fx0, fy0, fz0, 1)
    MRT_CHILD(2, nx0, ny0, nz0, fx0, fy0, fz0, 2) MRT_CHILD(3, nx0, ny0, nz0, fx0, fy0, fz0, 3)
    MRT_CHILD(0, nx1, ny1, nz1, fx1, fy1, fz1, 4) MRT_CHILD(1, nx1, ny1, nz1, fx1, fy1, fz1, 5)
    MRT_CHILD(2, nx1, ny1, nz1, fx1, fy1, fz1, 6) MRT_CHILD(3, nx1, ny1, nz1, fx1, fy1, fz1, 7)
#undef MRT_CHILD
#endif
#ifdef MRT_TRAVERSAL_STATS
    if (hits == 0u) {
        atomicAdd(&g_stat_fruitless, 1ull);
        uint32_t far = 0u;   // the same eight tests without the tbest bound
#define MRT_CHILD(B, NX, NY, NZ, FX, FY, FZ, K) \
        if (child_hit<B>(NX, NY, NZ, FX, FY, FZ, ax, ay, az, ox0, oy0, oz0, ox1, oy1, oz1, 1.0e30f, one)) { far |= 1u << (K); }
        MRT_CHILD(0, nx0, ny0, nz0, fx0, fy0, fz0, 0) MRT_CHILD(1, nx0, ny0, nz0, fx0, fy0, fz0, 1)
        MRT_CHILD(2, nx0, ny0, nz0, fx0, fy0, fz0, 2) MRT_CHILD(3, nx0, ny0, nz0, fx0, fy0, fz0, 3)
        MRT_CHILD(0, nx1, ny1, nz1, fx1, fy1, fz1, 4) MRT_CHILD(1, nx1, ny1, nz1, fx1, fy1, fz1, 5)
        MRT_CHILD(2, nx1, ny1, nz1, fx1, fy1, fz1, 6) MRT_CHILD(3, nx1, ny1, nz1, fx1, fy1, fz1, 7)
#undef MRT_CHILD
        if (far != 0u) { atomicAdd(&g_stat_stale, 1ull); }
    }
#endif
    // inner children: bit (slot ^ octinv), so that the highest pending bit is always the nearest child along the ray
    // (slot s of a node holds the child lying towards +axis k for every set bit k of s, see k_collapse8)
    uint32_t in = hits & imask;
    if (perm != nullptr) {
        in = perm[(octinv << 8) | in];
    } else {
        if (octinv & 1u) { in = ((in & 0x55u) << 1) | ((in & 0xaau) >> 1); }
        if (octinv & 2u) { in = ((in & 0x33u) << 2) | ((in & 0xccu) >> 2); }
        if (octinv & 4u) { in = ((in & 0x0fu) << 4) | ((in & 0xf0u) >> 4); }
    }
    t.ngroup.x = __float_as_uint(n1.x);
    t.ngroup.y = (in << 24) | imask;
    // leaf children: meta byte = (triangle count 1..7 << 5) | offset 0..23 into the node's triangle run (0 for an empty slot)
    lh = hits & ~imask;
    m0 = __float_as_uint(n1.z); m1 = __float_as_uint(n1.w);
    tbase = __float_as_uint(n1.y);
    return true;
}

// meta byte of leaf slot k
MRT_D uint32_t leaf_meta(uint32_t m0, uint32_t m1, uint32_t k) { return __byte_perm(m0, m1, k) & 0xffu; }

// Per-thread form: `tnew` = (triangle base, bit mask of the hit leaf triangles within the node's run of <= 24).
MRT_D bool trav8_node_step(const SceneDev &s, Trav8 &t, TravMem &tm_, uint2 &tnew, int *overflow) {
    uint32_t tbase, lh, m0, m1;
    if (!trav8_node_core(s, t, tm_, tbase, lh, m0, m1, overflow)) { return false; }
    uint32_t tm = 0u;
    while (lh) {
        const uint32_t k = __ffs(lh) - 1u;
        lh &= lh - 1u;
        const uint32_t mb = leaf_meta(m0, m1, k);
        tm |= ((1u << (mb >> 5)) - 1u) << (mb & 31u);
    }
    tnew.x = tbase;
    tnew.y = tm;
    return true;
}

// Triangle::intersect (Triangle.cpp:63-99) on a packed triangle, without the "t < current length" rule (the callers apply
// it so that they can break exact-t ties). A sound prefilter comes first: the sign / magnitude tests only reject what the
// reference's own comparisons reject for certain (tbest is only used there); survivors run the reference's exact
// arithmetic (one IEEE division). True when 0 <= u, v, u + v <= 1 and Epsilon <= tt.
MRT_D bool tri_hit(const float4 g0, const float4 g1, const float4 g2, v3 o, v3 d, float tbest, float &tt, float &u, float &v) {
    const v3 A = mk3(g0.x, g0.y, g0.z), AB = mk3(g0.w, g1.x, g1.y), AC = mk3(g1.z, g1.w, g2.x);
    const v3 perpendicularVector = cross(d, AC);
    const float det = dot(AB, perpendicularVector);
    const float adet = fabsf(det);
    if (adet < kEpsilon) { return false; }
    const uint32_t sgn = __float_as_uint(det) & 0x80000000u;
    const float hi = adet * 1.00001f;
    const v3 vectorToCamera = o - A;
    const float du = dot(vectorToCamera, perpendicularVector);
    const float su = __uint_as_float(__float_as_uint(du) ^ sgn);
    if (su < -1.0e-25f || su > hi) { return false; }
    const v3 upPerpendicularVector = cross(vectorToCamera, AB);
    const float dv = dot(d, upPerpendicularVector);
    const float sv = __uint_as_float(__float_as_uint(dv) ^ sgn);
    if (sv < -1.0e-25f || su + sv > hi) { return false; }
    const float dt = dot(AC, upPerpendicularVector);
    const float st = __uint_as_float(__float_as_uint(dt) ^ sgn);
    if (st < -1.0e-25f || st > hi * fminf(tbest, 1.0e25f)) { return false; }
    // exact path
    const float inv = 1.0f / det;
    u = inv * du;
    if (u < 0.0f || u > 1.0f) { return false; }
    v = inv * dv;
    if (v < 0.0f || (u + v) > 1.0f) { return false; }
    tt = inv * dt;
    return !(tt < kEpsilon);
}

// Tests the next pending triangle of t.tgroup. Returns true when an ANY-hit ray found its occluder.
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
    float tt, u, v;
    if (!tri_hit(g0, g1, g2, t.o, t.d, t.tbest, tt, u, v)) { return false; }
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
MRT_D bool trav8_step(const SceneDev &s, Trav8 &t, TravMem &tm_, int *overflow) {
    if (!(t.ngroup.y & 0xff000000u) && t.sp > 0) { t.ngroup = stack_get(tm_, --t.sp); }
    bool node_work = (t.ngroup.y & 0xff000000u) != 0u;
    if (node_work) {
        uint2 tnew;
        if (!trav8_node_step(s, t, tm_, tnew, overflow)) { return false; }
        if (tnew.y) { tm_.pend[t.np++] = tnew; }
        node_work = (t.ngroup.y & 0xff000000u) != 0u || t.sp > 0;
    }
    const uint32_t mask = __activemask();
    const uint32_t ready = __ballot_sync(mask, t.np > 0);
    const uint32_t busy = __ballot_sync(mask, node_work);
    const uint32_t full = __ballot_sync(mask, t.np >= kPend);
    if (__popc(ready) >= s.tune_postpone || busy == 0u || full != 0u) {
        if (t.np > 0) {
            t.tgroup = tm_.pend[--t.np];
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
    TravMem tm_;
    trav8_init(t, o, d, skipPrim, tbest);
    while (trav8_step<ANY>(s, t, tm_, overflow)) { }
    MRT_STAT(atomicAdd(&g_stat_nodes, t.n_nodes); atomicAdd(&g_stat_tris, t.n_tris); atomicAdd(&g_stat_rays, 1ull);)
    if (t.bestPrim >= 0) { tbest = t.tbest; bestPrim = t.bestPrim; bu = t.bu; bv = t.bv; }
    return ANY && t.bestPrim >= 0;
}

// The reference tests the root box of BVH<Triangle> with AABB::intersect before anything else (BVH.hpp:341-342); that
// test is topology independent, and its NaN rule (an axis-parallel ray whose origin lies exactly on a box plane
// misses, AABB.cpp:34-54) is visible in images of the orthographic scenes, so it is reproduced exactly here.
MRT_D bool triangles_root_hit(const SceneDev &s, v3 o, v3 d) {
    if (s.n_triangles <= 0) { return false; }
    if (s.naive) { return true; }                            // Naive<Triangle> has no box at all (Naive.hpp:80-91)
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

// ---- the reference's regular grids (MRT_ACC_REGULAR_GRID) ---------------------------------------------------------
// RegularGrid<T>::intersect (Accelerators/RegularGrid.hpp:333-513) on the cell lists of mrt_grid.h, one thread per ray:
// the 3-D DDA starts in the cell of the ray origin CLAMPED into the grid (a ray that starts outside is walked from the
// nearest border cell with the boundary distances of that cell, RegularGrid.hpp:337-347), tests every primitive of a
// cell against the running nearest hit, and once any cell has produced a hit it stops as soon as the hit lies before the
// next cell boundary (the reference's second loop, :461-512; re-testing the first hit's cell there changes nothing). A
// shadow ray returns at its first hit (:436-438). PRIM: 0 planes, 1 spheres, 2 triangles. Performance is not a goal
// here; the point is that a front-end configured for this accelerator sees the reference's visibility, quirks included.
MRT_D int32_t grid_to_int(float v) {   // static_cast<int32_t> of the reference's x86-64 build: out of range and NaN give INT_MIN
    return (v > -2147483904.0f && v < 2147483648.0f) ? __float2int_rz(v) : static_cast<int32_t>(0x80000000u);
}

template <int PRIM>
MRT_D bool trace_grid(const SceneDev &s, v3 o, v3 d, bool shadow, int skipKind, int skipIdx, HitRec &h) {
    const int nprims = PRIM == 0 ? s.n_planes : (PRIM == 1 ? s.n_spheres : s.n_triangles);
    if (nprims == 0) { return false; }
    constexpr int kSize = 32, kShift = 5;
    const GridDev &g = s.grids[PRIM];
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    const float bmin[3] = {g.bmin.x, g.bmin.y, g.bmin.z}, csize[3] = {g.cellSize.x, g.cellSize.y, g.cellSize.z}, cinv[3] = {g.cellInv.x, g.cellInv.y, g.cellInv.z};
    int32_t cell[3], step[3], out[3];
    float tmax[3], tdelta[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        int32_t c = grid_to_int((oo[k] - bmin[k]) * cinv[k]);
        c = min(c, kSize - 1);
        c = max(c, 0);
        float cb;
        if (dd[k] > 0.0f) { step[k] = 1; out[k] = kSize; cb = bmin[k] + (static_cast<float>(c) + 1.0f) * csize[k]; }
        else { step[k] = -1; out[k] = -1; cb = bmin[k] + static_cast<float>(c) * csize[k]; }
        if (fabsf(dd[k]) > 1.1920929e-7f) {
            const float r = 1.0f / dd[k];
            tmax[k] = (cb - oo[k]) * r;
            tdelta[k] = csize[k] * static_cast<float>(step[k]) * r;
        } else { tmax[k] = kRayLengthMax; tdelta[k] = 0.0f; }
        cell[k] = c;
    }
    bool found = false;
    while (true) {
        const uint32_t idx = static_cast<uint32_t>(cell[0]) + (static_cast<uint32_t>(cell[1]) << kShift) + (static_cast<uint32_t>(cell[2]) << (2 * kShift));
        const uint32_t first = __ldg(g.start + idx), last = __ldg(g.start + idx + 1);
        const float before = h.t;
        for (uint32_t i = first; i < last; ++i) {
            const int id = __ldg(g.items + i);
            float t, u = 0.0f, v = 0.0f;
            bool ok;
            if (PRIM == 0) {
                const PlaneRec p = s.planes[id];
                ok = !(skipKind == HIT_PLANE && skipIdx == id) && plane_intersect(mk3(p.nx, p.ny, p.nz), mk3(p.px, p.py, p.pz), o, d, t);
            } else if (PRIM == 1) {
                const SphereRec sp_ = s.spheres[id];
                ok = sphere_intersect(mk3(sp_.cx, sp_.cy, sp_.cz), sp_.sqRadius, o, d, t);
            } else {
                const float4 *tp = reinterpret_cast<const float4 *>(s.tri_geom) + 3ull * static_cast<uint32_t>(id);
                const float4 g0 = __ldg(tp), g1 = __ldg(tp + 1), g2 = __ldg(tp + 2);
                ok = !(skipKind == HIT_TRIANGLE && skipIdx == id) && tri_hit(g0, g1, g2, o, d, h.t, t, u, v);
            }
            if (ok && t < h.t) {
                h.t = t; h.kind = PRIM == 0 ? HIT_PLANE : (PRIM == 1 ? HIT_SPHERE : HIT_TRIANGLE); h.idx = id; h.u = u; h.v = v;
                if (shadow) { return true; }
            }
        }
        found = found || h.t < before;
        if (tmax[0] < tmax[1]) {
            if (tmax[0] < tmax[2]) {
                if (found && h.t < tmax[0]) { break; }
                cell[0] += step[0]; if (cell[0] == out[0]) { break; } tmax[0] = tmax[0] + tdelta[0];
            } else {
                if (found && h.t < tmax[2]) { break; }
                cell[2] += step[2]; if (cell[2] == out[2]) { break; } tmax[2] = tmax[2] + tdelta[2];
            }
        } else {
            if (tmax[1] < tmax[2]) {
                if (found && h.t < tmax[1]) { break; }
                cell[1] += step[1]; if (cell[1] == out[1]) { break; } tmax[1] = tmax[1] + tdelta[1];
            } else {
                if (found && h.t < tmax[2]) { break; }
                cell[2] += step[2]; if (cell[2] == out[2]) { break; } tmax[2] = tmax[2] + tdelta[2];
            }
        }
    }
    return false;
}

// Shader::rayTrace / shadowTrace with Accelerator::ACC_REGULAR_GRID (Shader.cpp:97-102,142-147): the three grids share one
// running t, then every light is tested (Shader.cpp:166-171)
MRT_D void trace_closest_grid(const SceneDev &s, v3 o, v3 d, int skipKind, int skipIdx, HitRec &h) {
    trace_grid<0>(s, o, d, false, skipKind, skipIdx, h);
    trace_grid<1>(s, o, d, false, skipKind, skipIdx, h);
    trace_grid<2>(s, o, d, false, skipKind, skipIdx, h);
    if (s.n_lights > 0) { trace_lights_linear(s, o, d, h); }
}
MRT_D bool trace_shadow_grid(const SceneDev &s, v3 o, v3 d, float distance, int skipKind, int skipIdx) {
    HitRec h;
    h.t = distance; h.kind = HIT_NONE; h.idx = -1; h.u = 0.0f; h.v = 0.0f;
    return trace_grid<0>(s, o, d, true, skipKind, skipIdx, h) || trace_grid<1>(s, o, d, true, skipKind, skipIdx, h) ||
           trace_grid<2>(s, o, d, true, skipKind, skipIdx, h);
}

// Shader::rayTrace's geometric part (Shader.cpp:86-111): planes, then spheres, then triangles share one running t;
// area lights are tested last with the same strict "<" rule, which the tie rule of trace_bvh8 reproduces because
// light primitive ids come after all scene triangles.
MRT_D void trace_closest(const SceneDev &s, v3 o, v3 d, int skipKind, int skipIdx, HitRec &h, int *overflow) {
    h.t = kRayLengthMax; h.kind = HIT_NONE; h.idx = -1; h.u = 0.0f; h.v = 0.0f;
    if (s.grid) { trace_closest_grid(s, o, d, skipKind, skipIdx, h); return; }
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
    if (s.grid) { return trace_shadow_grid(s, o, d, distance, skipKind, skipIdx); }
    HitRec h;
    h.t = distance; h.kind = HIT_NONE; h.idx = -1; h.u = 0.0f; h.v = 0.0f;
    if (trace_ref_bvh<0>(s, o, d, true, skipKind, skipIdx, h, overflow)) { return true; }
    if (trace_ref_bvh<1>(s, o, d, true, skipKind, skipIdx, h, overflow)) { return true; }
    if (!triangles_root_hit(s, o, d)) { return false; }
    int bestPrim = -1;
    return trace_bvh8<true>(s, o, d, skipKind == HIT_TRIANGLE ? skipIdx : -1, h.t, bestPrim, h.u, h.v, overflow);
}

}  // namespace mrt
