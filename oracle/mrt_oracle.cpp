// TEST INFRASTRUCTURE (oracle) — NOT product code; never linked into or called by libmobilert_b200.so.
//
// Plain C++ restatement of the reference's hot path (TiagoMSSantos/MobileRayTracer), one function per reference
// function, each citing the file:line it follows (paths relative to /root/reference/app). It exists so that the
// checker travels to machines without /root/reference; where the compiled reference (oracle/_ref) is present the tests
// pin this file against it bit for bit (tests/test_oracle_port.py), and against the reference's own known-answer unit
// tests (Unit_Testing/TestAABB.cpp, TestTriangle.cpp, TestPlane.cpp; tests/golden/).
// Parity status: PINNED (bitmaps of the deterministic shaders equal the compiled reference's on the built-in scenes).
// Built with -ffp-contract=off: the reference is compiled for baseline x86-64 without FMA (CMakeLists.txt:121-123).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

constexpr float Epsilon = 1.0e-06F, EpsilonLarge = 1.0e-05F, RayLengthMax = 1.0e+30F;   // Utils/Constants.hpp:22-40
constexpr int RayDepthMin = 1, RayDepthMax = 6, NumberOfTiles = 256, StackSize = 512;     // Constants.hpp:45-62
constexpr uint32_t ArrayMask = 0xFFFFFu;                                                  // Constants.hpp:79

struct V3 { float x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
// glm 1.0.1 func_geometric.inl: dot = (x*x' + y*y') + z*z'; cross; normalize = v * (1 / sqrt(dot)); reflect; refract
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 cross(V3 x, V3 y) { return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
inline float length(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return a * (1.0F / std::sqrt(dot(a, a))); }
inline V3 reflect(V3 I, V3 N) { return I - N * dot(N, I) * 2.0F; }
inline V3 refract(V3 I, V3 N, float eta) {
    const float d = dot(N, I), k = 1.0F - eta * eta * (1.0F - d * d);
    if (k >= 0.0F) { return eta * I - (eta * d + std::sqrt(k)) * N; }
    return {0, 0, 0};
}
inline bool hasPositiveValue(V3 a) { return a.x > 0 || a.y > 0 || a.z > 0; }   // Utils.hpp:278-281

// byte layouts = the reference's (Triangle.hpp:18-27, Sphere.hpp, Plane.hpp)
struct Triangle { V3 AC, AB, A, nA, nB, nC; float tA[2], tB[2], tC[2]; int32_t material; };
struct Sphere { V3 center; float sqRadius; int32_t material; };
struct Plane { V3 normal, point; int32_t material; };
struct Material { V3 Le, Kd, Ks, Kt; float ior; };
struct Light { int32_t kind; V3 position; Triangle triangle; V3 Le; };   // kind 0 point, 1 area
static_assert(sizeof(Triangle) == 100 && sizeof(Sphere) == 20 && sizeof(Plane) == 28 && sizeof(Material) == 52 && sizeof(Light) == 128, "layout");

struct AABB { V3 mn, mx; };
// AABB::intersect (Accelerators/AABB.cpp:34-54); std::min/max argument order matters when a NaN shows up
inline bool aabbIntersect(const AABB &b, V3 o, V3 d) {
    const float ix = 1.0F / d.x;
    const float t1x = (b.mn.x - o.x) * ix, t2x = (b.mx.x - o.x) * ix;
    float tMin = std::min(t1x, t2x), tMax = std::max(t1x, t2x);
    const float iy = 1.0F / d.y;
    const float t1y = (b.mn.y - o.y) * iy, t2y = (b.mx.y - o.y) * iy;
    tMin = std::max(tMin, std::min(t1y, t2y)); tMax = std::min(tMax, std::max(t1y, t2y));
    const float iz = 1.0F / d.z;
    const float t1z = (b.mn.z - o.z) * iz, t2z = (b.mx.z - o.z) * iz;
    tMin = std::max(tMin, std::min(t1z, t2z)); tMax = std::min(tMax, std::max(t1z, t2z));
    return tMax >= std::max(tMin, 0.0F);
}
inline float aabbArea(const AABB &b) {   // AABB::getSurfaceArea (AABB.cpp:61-71)
    const float l0 = b.mx.x - b.mn.x, l1 = b.mx.y - b.mn.y, l2 = b.mx.z - b.mn.z;
    return 2 * l0 * l2 + 2 * l0 * l1 + 2 * l2 * l1;
}
inline AABB surroundingBox(const AABB &a, const AABB &b) {   // AABB.cpp:113-119
    return {{std::min(a.mn.x, b.mn.x), std::min(a.mn.y, b.mn.y), std::min(a.mn.z, b.mn.z)},
            {std::max(a.mx.x, b.mx.x), std::max(a.mx.y, b.mx.y), std::max(a.mx.z, b.mx.z)}};
}
inline V3 centroid(const AABB &b) { return b.mn + (b.mx - b.mn) / 2.0F; }   // AABB::getCentroid (AABB.cpp:81-85)

enum Kind { NONE = 0, PLANE = 1, SPHERE = 2, TRIANGLE = 3, LIGHT = 4 };
struct Ray { V3 o, d; int depth; int skipKind, skipIdx; bool shadow; };
struct Hit { float length = RayLengthMax; V3 point {}, normal {}; int kind = NONE, idx = -1, material = -1; bool isLight = false; V3 Le {}; };

// ---- primitives -------------------------------------------------------------------------------------------
// Triangle::intersect (Shapes/Triangle.cpp:63-109)
inline void intersectTriangle(const Triangle &t, int idx, const Ray &r, Hit &h, bool asLight = false) {
    if (!asLight && r.skipKind == TRIANGLE && r.skipIdx == idx) { return; }
    const V3 p = cross(r.d, t.AC);
    const float det = dot(t.AB, p);
    if (std::fabs(det) < Epsilon) { return; }
    const float inv = 1.0F / det;
    const V3 toCam = r.o - t.A;
    const float u = inv * dot(toCam, p);
    if (u < 0.0F || u > 1.0F) { return; }
    const V3 q = cross(toCam, t.AB);
    const float v = inv * dot(r.d, q);
    if (v < 0.0F || (u + v) > 1.0F) { return; }
    const float dist = inv * dot(t.AC, q);
    if (dist < Epsilon || dist >= h.length) { return; }
    const float w = 1.0F - u - v;
    h.length = dist; h.point = r.o + r.d * dist;
    h.normal = normalize((t.nA * w + t.nB * u) + t.nC * v);
    h.kind = asLight ? LIGHT : TRIANGLE; h.idx = idx; h.material = asLight ? -1 : t.material; h.isLight = asLight;
}
inline AABB aabbOf(const Triangle &t) {   // Triangle::getAABB (Triangle.cpp:116-123)
    const V3 B = t.A + t.AB, C = t.A + t.AC;
    return {{std::min(t.A.x, std::min(B.x, C.x)), std::min(t.A.y, std::min(B.y, C.y)), std::min(t.A.z, std::min(B.z, C.z))},
            {std::max(t.A.x, std::max(B.x, C.x)), std::max(t.A.y, std::max(B.y, C.y)), std::max(t.A.z, std::max(B.z, C.z))}};
}
// Sphere::intersect (Shapes/Sphere.cpp:42-81): nearest root only, no self skip
inline void intersectSphere(const Sphere &s, int idx, const Ray &r, Hit &h) {
    const V3 oc = s.center - r.o;
    const float proj = dot(oc, r.d), mag = length(oc);
    const float a = dot(r.d, r.d), b = 2.0F * -proj, c = mag * mag - s.sqRadius;
    const float disc = b * b - 4.0F * a * c;
    if (disc < 0.0F) { return; }
    const float root = std::sqrt(disc);
    const float d1 = -b + root, d2 = -b - root;
    const float dist = std::min(d1, d2) / (2.0F * a);
    if (dist < EpsilonLarge || dist >= h.length) { return; }
    h.length = dist; h.point = r.o + r.d * dist; h.normal = normalize(h.point - s.center);
    h.kind = SPHERE; h.idx = idx; h.material = s.material; h.isLight = false;
}
inline AABB aabbOf(const Sphere &s) {   // Sphere::getAABB (Sphere.cpp:88-94)
    const float r = std::sqrt(s.sqRadius);
    return {{s.center.x - r, s.center.y - r, s.center.z - r}, {s.center.x + r, s.center.y + r, s.center.z + r}};
}
// Plane::intersect (Shapes/Plane.cpp:38-72)
inline void intersectPlane(const Plane &p, int idx, const Ray &r, Hit &h) {
    if (r.skipKind == PLANE && r.skipIdx == idx) { return; }
    const float proj = dot(p.normal, r.d);
    if (std::fabs(proj) < Epsilon) { return; }
    const float dist = dot(p.normal, p.point - r.o) / proj;
    if (dist < Epsilon || dist >= h.length) { return; }
    h.length = dist; h.point = r.o + r.d * dist; h.normal = p.normal;
    h.kind = PLANE; h.idx = idx; h.material = p.material; h.isLight = false;
}
inline AABB aabbOf(const Plane &p) {   // Plane::getRightVector + getAABB (Plane.cpp:79-109): finite, flat box
    V3 r {0, 0, 0};
    const V3 n = p.normal;
    if (n.x >= 1) { r = {0, 1, 1}; } else if (n.y >= 1) { r = {1, 0, 1}; } else if (n.z >= 1) { r = {1, 1, 0}; }
    else if (n.x <= -1) { r = {0, 1, 1}; } else if (n.y <= -1) { r = {1, 0, 1}; } else if (n.z <= -1) { r = {1, 1, 0}; }
    r = normalize(r);
    return {p.point + r * -100.0F, p.point + r * 100.0F};
}
inline void intersectPrim(const Triangle &t, int i, const Ray &r, Hit &h) { intersectTriangle(t, i, r, h); }
inline void intersectPrim(const Sphere &s, int i, const Ray &r, Hit &h) { intersectSphere(s, i, r, h); }
inline void intersectPrim(const Plane &p, int i, const Ray &r, Hit &h) { intersectPlane(p, i, r, h); }

// ---- BVH<T> (Accelerators/BVH.hpp) ------------------------------------------------------------------------------
template <typename T> struct BVH {
    struct Node { AABB box; int indexOffset = 0, numPrimitives = 0; };   // BVH.hpp:56-60
    std::vector<Node> boxes;
    std::vector<T> prims;
    std::vector<int> original;   // original[i] = caller index of prims[i]

    // BVH::build (BVH.hpp:160-283) with getSplitIndexSah (:396-439)
    void build(const std::vector<T> &in) {
        prims.clear(); boxes.clear(); original.clear();
        const int n = static_cast<int>(in.size());
        if (n == 0) { return; }
        struct BN { AABB box; V3 c; int old; };
        std::vector<BN> bn(static_cast<size_t>(n));
        for (int i = 0; i < n; ++i) { bn[i].box = aabbOf(in[i]); bn[i].c = centroid(bn[i].box); bn[i].old = i; }
        boxes.assign(static_cast<size_t>(2 * n - 1), Node {});
        struct Task { int node, begin, end; };
        std::vector<Task> stack;
        int cur = 0, begin = 0, end = n, maxNode = 0;
        while (true) {
            AABB sb = bn[begin].box;
            for (int i = begin + 1; i < end; ++i) { sb = surroundingBox(sb, bn[i].box); }
            const float md[3] = {sb.mx.x - sb.mn.x, sb.mx.y - sb.mn.y, sb.mx.z - sb.mn.z};
            const int axis = (md[0] >= md[1] && md[0] >= md[2]) ? 0 : ((md[1] >= md[0] && md[1] >= md[2]) ? 1 : 2);
            auto comp = [axis](const BN &b) { return axis == 0 ? b.c.x : (axis == 1 ? b.c.y : b.c.z); };
            const float step = md[axis] / 10.0F, start = axis == 0 ? sb.mn.x : (axis == 1 ? sb.mn.y : sb.mn.z);
            auto it = std::partition(bn.begin() + begin, bn.begin() + end, [&](const BN &b) { return comp(b) < start + step; });
            for (int bucket = 2; bucket < 10; ++bucket) {
                const float limit = start + step * static_cast<float>(bucket);
                it = std::partition(it, bn.begin() + end, [&](const BN &b) { return comp(b) < limit; });
            }
            AABB nb = bn[begin].box;
            for (int i = begin + 1; i < end; ++i) { nb = surroundingBox(bn[i].box, nb); }
            Node &node = boxes[static_cast<size_t>(cur)];
            node.box = nb;
            const int size = end - begin;
            if (size <= 4) {
                node.indexOffset = begin; node.numPrimitives = size;
                if (stack.empty()) { break; }
                const Task t = stack.back(); stack.pop_back();
                cur = t.node; begin = t.begin; end = t.end;
            } else {
                const int nb1 = size - 1;
                std::vector<float> la(static_cast<size_t>(nb1)), ra(static_cast<size_t>(nb1));
                AABB lb = bn[begin].box; la[0] = aabbArea(lb);
                for (int i = 1; i < nb1; ++i) { lb = surroundingBox(lb, bn[begin + i].box); la[i] = aabbArea(lb); }
                AABB rb = bn[begin + nb1].box; ra[nb1 - 1] = aabbArea(rb);
                for (int i = nb1 - 2; i >= 0; --i) { rb = surroundingBox(rb, bn[begin + i + 1].box); ra[i] = aabbArea(rb); }
                int split = 1;
                float minSah = la[0] + static_cast<float>(nb1) * ra[0];
                for (int i = 1; i < nb1; ++i) {
                    const float sah = static_cast<float>(i + 1) * la[i] + static_cast<float>(size - (i + 1)) * ra[i];
                    if (sah < minSah) { split = i + 1; minSah = sah; }
                }
                const int left = maxNode + 1, right = left + 1;
                node.indexOffset = left; node.numPrimitives = 0;
                maxNode = std::max(right, maxNode);
                stack.push_back(Task {right, begin + split, end});
                cur = left; end = begin + split;
            }
        }
        boxes.resize(static_cast<size_t>(maxNode + 1));
        prims.resize(static_cast<size_t>(n)); original.resize(static_cast<size_t>(n));
        for (int i = 0; i < n; ++i) { prims[i] = in[bn[i].old]; original[i] = bn[i].old; }
    }

    // BVH::intersect (BVH.hpp:326-384). Primitive indices reported in the Hit are positions in `prims` (post-build order),
    // which is also what the reference's pointer identity (Ray::primitive_) refers to.
    void intersect(const Ray &r, Hit &h) const {
        if (prims.empty()) { return; }
        int stackIdx[StackSize];
        int sp = 1, box = 0;
        stackIdx[0] = 0;
        do {
            const Node &node = boxes[static_cast<size_t>(box)];
            if (aabbIntersect(node.box, r.o, r.d)) {
                if (node.numPrimitives > 0) {
                    for (int i = 0; i < node.numPrimitives; ++i) {
                        const float last = h.length;
                        intersectPrim(prims[static_cast<size_t>(node.indexOffset + i)], node.indexOffset + i, r, h);
                        if (r.shadow && h.length < last) { return; }
                    }
                    box = stackIdx[--sp];
                } else {
                    const int left = node.indexOffset, right = left + 1;
                    const bool tl = aabbIntersect(boxes[static_cast<size_t>(left)].box, r.o, r.d);
                    const bool tr = aabbIntersect(boxes[static_cast<size_t>(right)].box, r.o, r.d);
                    if (!tl && !tr) { box = stackIdx[--sp]; }
                    else { box = tl ? left : right; if (tl && tr) { stackIdx[sp++] = right; } }
                }
            } else {
                box = stackIdx[--sp];
            }
        } while (sp > 0);
    }
};

// ---- Shader base (MobileRT/Shader.cpp) ---------------------------------------------------------------------------
struct Counter { uint32_t n = 0; uint32_t next() { return n++; } };

struct Engine {
    BVH<Plane> bvhPlanes; BVH<Sphere> bvhSpheres; BVH<Triangle> bvhTriangles;
    std::vector<Material> materials;
    std::vector<Light> lights;
    std::vector<Counter> lightSamplers;        // one StaticHaltonSeq per AreaLight (OBJLoader.cpp:383, Scenes.cpp:152-153)
    const float *tileOrder = nullptr, *shaderTable = nullptr, *staticHalton = nullptr;
    Counter hemiCounter, lightIndexCounter;    // function-static counters of Shader.cpp:189,224 (never reset)
    Counter pixelCounter, rrCounter;           // samplerPixel_ (Renderer) and samplerRussianRoulette_ (PathTracer)
    int shader = 1, spl = 1, spp = 1, width = 0, height = 0;
    bool pixelJitter = false;                  // C_wrapper.cpp:144-148: StaticHaltonSeq when spp > 1, else Constant(0.5)
    V3 depthMaxPoint {1, 1, 1};
    // camera (Camera.cpp:14-19, Perspective.cpp:8-14 / Orthographic.cpp:7-13)
    int camKind = 0; V3 camPos {}, camDir {}, camRight {}, camUp {}; float camA = 0, camB = 0;
    std::atomic<uint64_t> rays {0};            // Ray::counter (Ray.cpp:11,26)

    Ray makeRay(V3 d, V3 o, int depth, bool shadow, int skipKind, int skipIdx) { rays.fetch_add(1, std::memory_order_relaxed); return {o, d, depth, skipKind, skipIdx, shadow}; }

    // Shader::traceLights (Shader.cpp:166-171) + AreaLight::intersect (AreaLight.cpp:32-41)
    void traceLights(const Ray &r, Hit &h) const {
        for (size_t i = 0; i < lights.size(); ++i) {
            if (lights[i].kind != 1) { continue; }
            const float last = h.length;
            intersectTriangle(lights[i].triangle, static_cast<int>(i), r, h, true);
            if (h.length < last) { h.Le = lights[i].Le; }
        }
    }
    // geometric part of Shader::rayTrace (Shader.cpp:86-111)
    void traceClosest(const Ray &r, Hit &h) const {
        bvhPlanes.intersect(r, h); bvhSpheres.intersect(r, h); bvhTriangles.intersect(r, h);
        traceLights(r, h);
    }
    // Shader::shadowTrace (Shader.cpp:132-158)
    bool shadowTrace(float distance, const Ray &r) const {
        Hit h; h.length = distance;
        bvhPlanes.intersect(r, h); if (h.length < distance) { return true; }
        bvhSpheres.intersect(r, h); if (h.length < distance) { return true; }
        bvhTriangles.intersect(r, h);
        return h.length < distance;
    }
    // Shader::getLightIndex (Shader.cpp:223-233)
    uint32_t getLightIndex() {
        const float rnd = shaderTable[lightIndexCounter.next() & ArrayMask];
        return static_cast<uint32_t>(std::floor(rnd * static_cast<uint32_t>(lights.size()) * 0.99999F));
    }
    // PointLight::getPosition (PointLight.cpp:16-18) / AreaLight::getPosition (AreaLight.cpp:17-26)
    V3 lightPosition(uint32_t i) {
        const Light &l = lights[i];
        if (l.kind == 0) { return l.position; }
        float r = staticHalton[lightSamplers[i].next() & ArrayMask];
        float s = staticHalton[lightSamplers[i].next() & ArrayMask];
        if (r + s >= 1.0F) { r = 1.0F - r; s = 1.0F - s; }
        return (l.triangle.A + r * l.triangle.AB) + s * l.triangle.AC;
    }
    // Shader::getCosineSampleHemisphere (Shader.cpp:188-216)
    V3 cosineSampleHemisphere(V3 normal) {
        const float r1 = shaderTable[hemiCounter.next() & ArrayMask], r2 = shaderTable[hemiCounter.next() & ArrayMask];
        const float phi = 6.28318530717958647692528676655900576F * r1, cosTheta = std::sqrt(r2);
        V3 u = std::fabs(normal.x) > 0.1F ? V3 {0, 1, 0} : V3 {1, 0, 0};
        u = normalize(cross(u, normal));
        const V3 v = cross(normal, u);
        const V3 dir = (u * (std::cos(phi) * cosTheta) + v * (std::sin(phi) * cosTheta)) + normal * std::sqrt(1.0F - r2);
        return normalize(dir);
    }
    struct Surf { V3 Le, Kd, Ks, Kt; float ior; };
    Surf surface(const Hit &h) const {
        if (h.isLight) { return {h.Le, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}, 1.0F}; }
        const Material &m = materials[static_cast<size_t>(h.material)];
        return {m.Le, m.Kd, m.Ks, m.Kt, m.ior};
    }
    static void skipOf(const Hit &h, int &kind, int &idx) {   // Ray::primitive_: triangles and planes only (Sphere.cpp:77 stores nullptr)
        kind = (h.kind == TRIANGLE || h.kind == PLANE) ? h.kind : NONE; idx = kind == NONE ? -1 : h.idx;
    }

    // direct light of Whitted (Whitted.cpp:37-64) / NoShadows (NoShadows.cpp:25-43) / PathTracer (PathTracer.cpp:55-87)
    V3 directLight(const Hit &h, const Surf &sf, int depth, bool shadows) {
        V3 rgb {0, 0, 0};
        if (hasPositiveValue(sf.Kd) && !lights.empty()) {
            int sk, si; skipOf(h, sk, si);
            for (int i = 0; i < spl; ++i) {
                const uint32_t chosen = getLightIndex();
                const V3 lp = lightPosition(chosen);
                V3 toLight = lp - h.point;
                const float dist = length(toLight);
                toLight = normalize(toLight);
                const float cosNl = dot(h.normal, toLight);
                if (cosNl > 0.0F) {
                    bool lit = true;
                    if (shadows) { const Ray sr = makeRay(toLight, h.point, depth + 1, true, sk, si); lit = !shadowTrace(dist, sr); }
                    if (lit) { rgb = rgb + lights[chosen].Le * cosNl; }
                }
            }
            rgb = rgb * sf.Kd;
            rgb = rgb / static_cast<float>(spl);
        }
        return rgb;
    }

    // Shader::rayTrace (Shader.cpp:86-123) + the shade() of the selected shader. Returns shade()'s bool.
    bool rayTrace(V3 &rgb, const Ray &ray) {
        Hit h;
        traceClosest(ray, h);
        if (!(h.length < RayLengthMax)) { return false; }
        const Surf sf = surface(h);
        int sk, si; skipOf(h, sk, si);
        switch (shader) {
            case 3: {   // DepthMap::shade (DepthMap.cpp:12-17)
                const float maxDist = length(depthMaxPoint - ray.o) * 1.1F;
                const float depth = std::max((maxDist - h.length) / maxDist, 0.0F);
                rgb = {depth, depth, depth};
                return false;
            }
            case 4:     // DiffuseMaterial::shade (DiffuseMaterial.cpp:11-27)
                if (hasPositiveValue(sf.Kd)) { rgb = sf.Kd; } else if (hasPositiveValue(sf.Ks)) { rgb = sf.Ks; }
                else if (hasPositiveValue(sf.Kt)) { rgb = sf.Kt; } else if (hasPositiveValue(sf.Le)) { rgb = sf.Le; }
                return false;
            case 0:     // NoShadows::shade (NoShadows.cpp:12-46)
                if (hasPositiveValue(sf.Le)) { rgb = sf.Le; return true; }
                rgb = directLight(h, sf, ray.depth, false);
                rgb = rgb + sf.Kd * 0.1F;
                return false;
            case 1: {   // Whitted::shade (Whitted.cpp:13-93)
                if (ray.depth > RayDepthMax) { return false; }
                if (hasPositiveValue(sf.Le)) { rgb = sf.Le; return true; }
                rgb = directLight(h, sf, ray.depth, true);
                if (hasPositiveValue(sf.Ks)) {
                    V3 li {0, 0, 0};
                    rayTrace(li, makeRay(reflect(ray.d, h.normal), h.point, ray.depth + 1, false, sk, si));
                    rgb = rgb + sf.Ks * li;
                }
                if (hasPositiveValue(sf.Kt)) {
                    V3 li {0, 0, 0};
                    rayTrace(li, makeRay(refract(ray.d, h.normal, 1.0F / sf.ior), h.point, ray.depth + 1, false, sk, si));
                    rgb = rgb + sf.Kt * li;
                }
                rgb = rgb + sf.Kd * 0.1F;
                return false;
            }
            default: {  // PathTracer::shade (PathTracer.cpp:22-142)
                if (ray.depth > RayDepthMax) { return false; }
                if (hasPositiveValue(sf.Le)) { rgb = sf.Le; return true; }
                V3 Ld {0, 0, 0}, LiD {0, 0, 0}, LiS {0, 0, 0}, LiT {0, 0, 0};
                bool intersectedLight = false;
                if (hasPositiveValue(sf.Kd)) {
                    Ld = directLight(h, sf, ray.depth, true);
                    if (ray.depth <= RayDepthMin || staticHalton[rrCounter.next() & ArrayMask] > 0.5F) {
                        const V3 nd = cosineSampleHemisphere(h.normal);
                        V3 li {0, 0, 0};
                        intersectedLight = rayTrace(li, makeRay(nd, h.point, ray.depth + 1, false, sk, si));
                        LiD = LiD + sf.Kd * li;
                        if (ray.depth > RayDepthMin) { LiD = LiD / (0.5F * 0.5F); }
                        if (hasPositiveValue(Ld) && intersectedLight) { LiD = {0, 0, 0}; }
                    }
                }
                if (hasPositiveValue(sf.Ks)) {
                    V3 li {0, 0, 0};
                    rayTrace(li, makeRay(reflect(ray.d, h.normal), h.point, ray.depth + 1, false, sk, si));
                    LiS = LiS + sf.Ks * li;
                }
                if (hasPositiveValue(sf.Kt)) {
                    V3 li {0, 0, 0};
                    rayTrace(li, makeRay(refract(ray.d, h.normal, 1.0F / sf.ior), h.point, ray.depth + 1, false, sk, si));
                    LiT = LiT + sf.Kt * li;
                }
                rgb = rgb + Ld; rgb = rgb + LiD; rgb = rgb + LiS; rgb = rgb + LiT;
                return intersectedLight;
            }
        }
    }

    // Perspective::generateRay + fastArcTan (Perspective.cpp:16-46) / Orthographic::generateRay (Orthographic.cpp:15-23)
    static float fastArcTan(float v) {
        const float a = std::fabs(v);
        return 0.785398163397448309615660845819875721F * v - (v * (a - 1.0F)) * (0.2447F + (0.0663F * a));
    }
    Ray generateRay(float u, float v, float du, float dv) {
        if (camKind == 0) {
            const float rf = fastArcTan(camA * (u - 0.5F)) + du;
            const V3 right = camRight * rf;
            const float uf = fastArcTan(camB * (0.5F - v)) + dv;
            const V3 up = camUp * uf;
            const V3 dest = ((camPos + camDir) + right) + up;
            return makeRay(normalize(dest - camPos), camPos, 1, false, NONE, -1);
        }
        const V3 right = camRight * ((u - 0.5F) * camA) + camRight * du;
        const V3 up = camUp * ((0.5F - v) * camB) + camUp * dv;
        return makeRay(camDir, (camPos + right) + up, 1, false, NONE, -1);
    }
};

// MobileRT::incrementalAvg (Utils/Utils.cpp:66-90); float -> uint32 as x86-64 does it (cvttss2si to 64 bit, truncated)
inline uint32_t f2u(float f) {
    if (!(f == f) || f >= 9.2233720e18F || f <= -9.2233720e18F) { return 0u; }
    return static_cast<uint32_t>(static_cast<uint64_t>(static_cast<int64_t>(f)));
}
inline int32_t incrementalAvg(V3 s, int32_t avg, int32_t num) {
    const uint32_t a = static_cast<uint32_t>(avg), n = static_cast<uint32_t>(num);
    const uint32_t r = ((n - 1u) * (a & 0xFFu) + f2u(s.x * 255.0F)) / n;
    const uint32_t g = ((n - 1u) * ((a >> 8u) & 0xFFu) + f2u(s.y * 255.0F)) / n;
    const uint32_t b = ((n - 1u) * ((a >> 16u) & 0xFFu) + f2u(s.z * 255.0F)) / n;
    return static_cast<int32_t>(0xFF000000u | std::min(b, 255u) << 16u | std::min(g, 255u) << 8u | std::min(r, 255u));
}
// MobileRT::haltonSequence (Utils/Utils.cpp:43-53)
inline float haltonSequence(uint32_t index, uint32_t base) {
    float f = 1.0F, result = 0.0F;
    const float b = static_cast<float>(base);
    while (index > 0u) { f /= b; result += f * static_cast<float>(index % base); index = index / base; }
    return result;
}

// Renderer::renderFrame / renderScene / getTile, single-threaded (Renderer.cpp:53-88,107-170,189-197)
void renderFrame(Engine &e, int32_t *bitmap) {
    const int w = e.width, hgt = e.height, bsx = w / 16, bsy = hgt / 16;
    if (bsx == 0 || bsy == 0) { return; }
    const int domainSize = (w / bsx) * (hgt / bsy), resolution = w * hgt;
    const float invW = 1.0F / static_cast<float>(w), invH = 1.0F / static_cast<float>(hgt);
    const float pixelW = 0.5F / static_cast<float>(w), pixelH = 0.5F / static_cast<float>(hgt);
    e.pixelCounter.n = 0; e.rrCounter.n = 0;
    for (Counter &c : e.lightSamplers) { c.n = 0; }          // Shader::resetSampling (Shader.cpp:176-180)
    for (int sample = 0; sample < e.spp; ++sample) {
        for (int k = 0; k < NumberOfTiles; ++k) {
            const float tile = e.tileOrder[k];
            const int roundBlock = static_cast<int>(::roundf(tile * static_cast<float>(domainSize)));
            const int pixel = roundBlock * bsx % resolution;
            const int startY = ((pixel / w) * bsy) % hgt, endY = startY + bsy;
            for (int y = startY; y < endY; ++y) {
                const float v = static_cast<float>(y) * invH;
                const int yWidth = y * w, startX = (pixel + yWidth) % w, endX = startX + bsx;
                for (int x = startX; x < endX; ++x) {
                    const float u = static_cast<float>(x) * invW;
                    float r1 = 0.5F, r2 = 0.5F;
                    if (e.pixelJitter) { r1 = e.staticHalton[e.pixelCounter.next() & ArrayMask]; r2 = e.staticHalton[e.pixelCounter.next() & ArrayMask]; }
                    const float du = (r1 - 0.5F) * 2.0F * pixelW, dv = (r2 - 0.5F) * 2.0F * pixelH;
                    const Ray ray = e.generateRay(u, v, du, dv);
                    V3 rgb {0, 0, 0};
                    e.rayTrace(rgb, ray);
                    const int idx = yWidth + x;
                    if (idx < resolution) { bitmap[idx] = incrementalAvg(rgb, bitmap[idx], sample + 1); }
                }
            }
        }
    }
}

}  // namespace

// ---- flat C API (ctypes: oracle/portbind.py) ------------------------------------------------------------------------
extern "C" {

struct PortScene { const Triangle *tris; int64_t nTris; const Sphere *spheres; int64_t nSpheres; const Plane *planes; int64_t nPlanes;
                   const Material *mats; int64_t nMats; const Light *lights; int64_t nLights; };
struct PortCamera { int32_t kind; float position[3], lookAt[3], up[3]; float a, b; };
struct PortRender { int32_t width, height, spp, spl, shader; float depthMaxPoint[3]; const float *tileOrder, *shaderTable, *staticHalton; };

void *port_create(const PortScene *s, const PortCamera *c, const PortRender *r) {
    Engine *e = new Engine;
    e->bvhTriangles.build(std::vector<Triangle>(s->tris, s->tris + s->nTris));
    e->bvhSpheres.build(std::vector<Sphere>(s->spheres, s->spheres + s->nSpheres));
    e->bvhPlanes.build(std::vector<Plane>(s->planes, s->planes + s->nPlanes));
    e->materials.assign(s->mats, s->mats + s->nMats);
    e->lights.assign(s->lights, s->lights + s->nLights);
    e->lightSamplers.assign(static_cast<size_t>(s->nLights), Counter {});
    e->width = r->width; e->height = r->height; e->spp = r->spp; e->spl = r->spl; e->shader = r->shader;
    e->pixelJitter = r->spp > 1;
    e->depthMaxPoint = {r->depthMaxPoint[0], r->depthMaxPoint[1], r->depthMaxPoint[2]};
    e->tileOrder = r->tileOrder; e->shaderTable = r->shaderTable; e->staticHalton = r->staticHalton;
    const V3 pos {c->position[0], c->position[1], c->position[2]}, look {c->lookAt[0], c->lookAt[1], c->lookAt[2]}, up {c->up[0], c->up[1], c->up[2]};
    e->camKind = c->kind; e->camPos = pos;
    e->camDir = normalize(look - pos); e->camRight = cross(up, e->camDir); e->camUp = cross(e->camDir, e->camRight);   // Camera.cpp:14-19
    if (c->kind == 0) { const float pi = 3.14159265358979323846264338327950288F; e->camA = (c->a * pi) / 180.0F; e->camB = (c->b * pi) / 180.0F; }
    else { e->camA = c->a / 2.0F; e->camB = c->b / 2.0F; }
    return e;
}
void port_destroy(void *p) { delete static_cast<Engine *>(p); }
void port_render(void *p, int32_t *bitmap) { renderFrame(*static_cast<Engine *>(p), bitmap); }
uint64_t port_casted_rays(void *p) { return static_cast<Engine *>(p)->rays.load(); }
void port_set_counters(void *p, uint32_t hemi, uint32_t lightIndex) { auto *e = static_cast<Engine *>(p); e->hemiCounter.n = hemi; e->lightIndexCounter.n = lightIndex; }

// per-ray ground truth: kind/idx in the CALLER's array order, like the GPU library reports them
void port_trace(void *p, int64_t n, const float *o, const float *d, float *outT, int32_t *outKind, int32_t *outIdx) {
    Engine &e = *static_cast<Engine *>(p);
    for (int64_t i = 0; i < n; ++i) {
        const Ray r {{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, {d[3 * i], d[3 * i + 1], d[3 * i + 2]}, 1, NONE, -1, false};
        Hit h;
        e.traceClosest(r, h);
        outT[i] = h.length; outKind[i] = h.kind;
        int idx = h.idx;
        if (h.kind == TRIANGLE) { idx = e.bvhTriangles.original[static_cast<size_t>(idx)]; }
        else if (h.kind == SPHERE) { idx = e.bvhSpheres.original[static_cast<size_t>(idx)]; }
        else if (h.kind == PLANE) { idx = e.bvhPlanes.original[static_cast<size_t>(idx)]; }
        outIdx[i] = idx;
    }
}

// primitive-level known-answer entry points (Unit_Testing/TestAABB.cpp, TestTriangle.cpp, TestPlane.cpp)
int port_aabb_intersect(const float *mn, const float *mx, const float *o, const float *d) {
    return aabbIntersect({{mn[0], mn[1], mn[2]}, {mx[0], mx[1], mx[2]}}, {o[0], o[1], o[2]}, {d[0], d[1], d[2]}) ? 1 : 0;
}
float port_aabb_area(const float *mn, const float *mx) { return aabbArea({{mn[0], mn[1], mn[2]}, {mx[0], mx[1], mx[2]}}); }
float port_triangle_intersect(const Triangle *t, const float *o, const float *d, float tMax, int selfSkip) {
    const Ray r {{o[0], o[1], o[2]}, {d[0], d[1], d[2]}, 1, selfSkip ? TRIANGLE : NONE, selfSkip ? 0 : -1, false};
    Hit h; h.length = tMax;
    intersectTriangle(*t, 0, r, h);
    return h.length;
}
void port_triangle_aabb(const Triangle *t, float *out6) { const AABB b = aabbOf(*t); std::memcpy(out6, &b, 24); }
float port_sphere_intersect(const Sphere *s, const float *o, const float *d, float tMax) {
    const Ray r {{o[0], o[1], o[2]}, {d[0], d[1], d[2]}, 1, NONE, -1, false};
    Hit h; h.length = tMax;
    intersectSphere(*s, 0, r, h);
    return h.length;
}
float port_plane_intersect(const Plane *pl, const float *o, const float *d, float tMax) {
    const Ray r {{o[0], o[1], o[2]}, {d[0], d[1], d[2]}, 1, NONE, -1, false};
    Hit h; h.length = tMax;
    intersectPlane(*pl, 0, r, h);
    return h.length;
}
void port_plane_aabb(const Plane *pl, float *out6) { const AABB b = aabbOf(*pl); std::memcpy(out6, &b, 24); }
int32_t port_incremental_avg(const float *rgb, int32_t avg, int32_t n) { return incrementalAvg({rgb[0], rgb[1], rgb[2]}, avg, n); }
float port_halton(uint32_t index, uint32_t base) { return haltonSequence(index, base); }

}  // extern "C"
