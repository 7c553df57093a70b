// Vector math and primitive intersection tests in the reference's exact fp32 operation order.
// The reference is compiled for baseline x86-64 (no FMA contraction, app/CMakeLists.txt:121-123), so this
// translation unit is built with -fmad=false: every * and + below rounds separately, as glm 1.0.1 does on
// the CPU. Where fused arithmetic is wanted (conservative BVH8 box tests) __fmaf_rn is written explicitly.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define MRT_HD __host__ __device__ __forceinline__
#define MRT_D __device__ __forceinline__

namespace mrt {

// MobileRT/Utils/Constants.hpp:22-79
constexpr float kEpsilon = 1.0e-06f;
constexpr float kEpsilonLarge = 1.0e-05f;
constexpr float kRayLengthMax = 1.0e+30f;
constexpr int kRayDepthMin = 1;
constexpr int kRayDepthMax = 6;
constexpr int kNumberOfTiles = 256;
constexpr uint32_t kArrayMask = 0xFFFFFu;
constexpr uint32_t kArraySize = kArrayMask + 1u;

struct v3 { float x, y, z; };

MRT_HD v3 mk3(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
MRT_HD v3 operator+(v3 a, v3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
MRT_HD v3 operator-(v3 a, v3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
MRT_HD v3 operator*(v3 a, v3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
MRT_HD v3 operator*(v3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
MRT_HD v3 operator*(float s, v3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
MRT_HD v3 operator/(v3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
MRT_HD v3 operator-(v3 a) { return mk3(-a.x, -a.y, -a.z); }
// glm::dot(vec3): tmp = a*b; tmp.x + tmp.y + tmp.z  (glm/detail/func_geometric.inl compute_dot<vec<3>>)
MRT_HD float dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// glm::cross (func_geometric.inl compute_cross)
MRT_HD v3 cross(v3 x, v3 y) { return mk3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
MRT_HD float length(v3 a) { return sqrtf(dot(a, a)); }
// glm::normalize = v * inversesqrt(dot(v,v)); inversesqrt(x) = 1 / sqrt(x)
MRT_HD v3 normalize(v3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
// glm::reflect: I - N * dot(N, I) * 2
MRT_HD v3 reflect(v3 I, v3 N) { return I - N * dot(N, I) * 2.0f; }
// glm::refract (func_geometric.inl:113-122)
MRT_HD v3 refract(v3 I, v3 N, float eta) {
    const float dotValue = dot(N, I);
    const float k = 1.0f - eta * eta * (1.0f - dotValue * dotValue);
    if (k >= 0.0f) { return eta * I - (eta * dotValue + sqrtf(k)) * N; }
    return mk3(0.0f, 0.0f, 0.0f);
}
// MobileRT::hasPositiveValue (Utils.hpp:278-281)
MRT_HD bool has_positive(v3 a) { return a.x > 0.0f || a.y > 0.0f || a.z > 0.0f; }
// std::min / std::max argument-order semantics (they differ from fminf/fmaxf when a NaN is involved)
MRT_HD float std_min(float a, float b) { return (b < a) ? b : a; }
MRT_HD float std_max(float a, float b) { return (a < b) ? b : a; }

// ---- AABB::intersect (MobileRT/Accelerators/AABB.cpp:34-54), exact --------------------------------------
MRT_HD bool aabb_intersect_exact(v3 bmin, v3 bmax, v3 o, v3 d) {
    const float invDirX = 1.0f / d.x;
    const float t1X = (bmin.x - o.x) * invDirX;
    const float t2X = (bmax.x - o.x) * invDirX;
    float tMin = std_min(t1X, t2X);
    float tMax = std_max(t1X, t2X);
    {
        const float invDir = 1.0f / d.y;
        const float t1 = (bmin.y - o.y) * invDir;
        const float t2 = (bmax.y - o.y) * invDir;
        tMin = std_max(tMin, std_min(t1, t2));
        tMax = std_min(tMax, std_max(t1, t2));
    }
    {
        const float invDir = 1.0f / d.z;
        const float t1 = (bmin.z - o.z) * invDir;
        const float t2 = (bmax.z - o.z) * invDir;
        tMin = std_max(tMin, std_min(t1, t2));
        tMax = std_min(tMax, std_max(t1, t2));
    }
    return tMax >= std_max(tMin, 0.0f);
}

// ---- Triangle::intersect (MobileRT/Shapes/Triangle.cpp:63-99): Moller-Trumbore, exact order ---------------
// Returns true when Epsilon <= t (caller applies the "< current length" rule so that it can break ties).
MRT_HD bool triangle_intersect(v3 A, v3 AB, v3 AC, v3 o, v3 d, float &t, float &u, float &v) {
    const v3 perpendicularVector = cross(d, AC);
    const float normalizedProjection = dot(AB, perpendicularVector);
    if (fabsf(normalizedProjection) < kEpsilon) { return false; }
    const float normalizedProjectionInv = 1.0f / normalizedProjection;
    const v3 vectorToCamera = o - A;
    u = normalizedProjectionInv * dot(vectorToCamera, perpendicularVector);
    if (u < 0.0f || u > 1.0f) { return false; }
    const v3 upPerpendicularVector = cross(vectorToCamera, AB);
    v = normalizedProjectionInv * dot(d, upPerpendicularVector);
    if (v < 0.0f || (u + v) > 1.0f) { return false; }
    t = normalizedProjectionInv * dot(AC, upPerpendicularVector);
    return !(t < kEpsilon);
}

// ---- Sphere::intersect (MobileRT/Shapes/Sphere.cpp:42-66): nearest root only -------------------------------
MRT_HD bool sphere_intersect(v3 center, float sqRadius, v3 o, v3 d, float &t) {
    const v3 originToCenter = center - o;
    const float projectionOnDirection = dot(originToCenter, d);
    const float originToCenterMagnitude = length(originToCenter);
    const float a = dot(d, d);
    const float b = 2.0f * -projectionOnDirection;
    const float c = originToCenterMagnitude * originToCenterMagnitude - sqRadius;
    const float discriminant = b * b - 4.0f * a * c;
    if (discriminant < 0.0f) { return false; }
    const float rootDiscriminant = sqrtf(discriminant);
    const float distanceToIntersection1 = -b + rootDiscriminant;
    const float distanceToIntersection2 = -b - rootDiscriminant;
    t = std_min(distanceToIntersection1, distanceToIntersection2) / (2.0f * a);
    return !(t < kEpsilonLarge);
}

// ---- Plane::intersect (MobileRT/Shapes/Plane.cpp:38-62) -----------------------------------------------------
MRT_HD bool plane_intersect(v3 normal, v3 point, v3 o, v3 d, float &t) {
    const float normalizedProjection = dot(normal, d);
    if (fabsf(normalizedProjection) < kEpsilon) { return false; }
    const v3 vecToPlane = point - o;
    const float scalarProjectionVecToPlaneOnNormal = dot(normal, vecToPlane);
    t = scalarProjectionVecToPlaneOnNormal / normalizedProjection;
    return !(t < kEpsilon);
}

// ---- MobileRT::haltonSequence (MobileRT/Utils/Utils.cpp:43-53) ----------------------------------------------
MRT_HD float halton_sequence(uint32_t index, uint32_t base) {
    float fraction = 1.0f;
    float nextValue = 0.0f;
    const float baseInFloat = static_cast<float>(base);
    while (index > 0u) {
        fraction /= baseInFloat;
        nextValue += fraction * static_cast<float>(index % base);
        index = index / base;  // floor(index / base) on unsigned operands is the integer quotient
    }
    return nextValue;
}

// ---- MobileRT::incrementalAvg (MobileRT/Utils/Utils.cpp:66-90) ---------------------------------------------
// static_cast<uint32>(float) is undefined for negative / huge values in C++; x86-64 cvttss2si (64-bit, then
// truncated to 32 bits) is what the reference binary does, reproduced here with a 64-bit conversion.
MRT_HD uint32_t f2u_x86(float f) {
#ifdef __CUDA_ARCH__
    if (!(f == f) || f >= 9.2233720e18f || f <= -9.2233720e18f) { return 0u; }
    return static_cast<uint32_t>(static_cast<unsigned long long>(static_cast<long long>(f)));
#else
    if (!(f == f) || f >= 9.2233720e18f || f <= -9.2233720e18f) { return 0u; }
    return static_cast<uint32_t>(static_cast<unsigned long long>(static_cast<long long>(f)));
#endif
}
MRT_HD int32_t incremental_avg(v3 sample, int32_t avg, int32_t numSample) {
    const uint32_t avgUnsigned = static_cast<uint32_t>(avg);
    const uint32_t n = static_cast<uint32_t>(numSample);
    const uint32_t lastRed = avgUnsigned & 0xFFu;
    const uint32_t lastGreen = (avgUnsigned >> 8u) & 0xFFu;
    const uint32_t lastBlue = (avgUnsigned >> 16u) & 0xFFu;
    const uint32_t samplerRed = f2u_x86(sample.x * 255.0f);
    const uint32_t samplerGreen = f2u_x86(sample.y * 255.0f);
    const uint32_t samplerBlue = f2u_x86(sample.z * 255.0f);
    const uint32_t currentRed = ((n - 1u) * lastRed + samplerRed) / n;
    const uint32_t currentGreen = ((n - 1u) * lastGreen + samplerGreen) / n;
    const uint32_t currentBlue = ((n - 1u) * lastBlue + samplerBlue) / n;
    const uint32_t retR = currentRed < 255u ? currentRed : 255u;
    const uint32_t retG = currentGreen < 255u ? currentGreen : 255u;
    const uint32_t retB = currentBlue < 255u ? currentBlue : 255u;
    return static_cast<int32_t>(0xFF000000u | retB << 16u | retG << 8u | retR);
}

// ---- Perspective::fastArcTan (Components/Cameras/Perspective.cpp:37-43) ------------------------------------
MRT_HD float fast_arctan(float value) {
    const float absValue = fabsf(value);
    return 0.785398163397448309615660845819875721f * value - (value * (absValue - 1.0f)) * (0.2447f + (0.0663f * absValue));
}

}  // namespace mrt
