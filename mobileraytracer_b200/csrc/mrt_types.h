// Device-side data layout (HBM) shared by the builder, the traversal kernels and the host API.
#pragma once
#include "mrt_math.cuh"

namespace mrt {

// Compressed 8-wide BVH node, 80 bytes, 16-byte aligned (5 x float4 loads).
// Child boxes are quantised to 8 bits per plane relative to (p, 2^e): lo = p + qlo * 2^e (rounded outward);
// ex/ey/ez hold the IEEE exponent byte of 2^e plus 15.
// meta[i]: empty 0 | inner: (1 << 5) | (24 + i) | leaf: (triangle count 1..7 << 5) | triangle offset 0..23.
// Inner children are stored contiguously from child_base in slot order (imask bit i = slot i is inner);
// the triangles of all leaf slots are stored contiguously from tri_base (<= 24 per node).
struct __align__(16) Node8 {
    float px, py, pz;
    uint8_t ex, ey, ez, imask;
    uint32_t child_base;
    uint32_t tri_base;
    uint8_t meta[8];
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(Node8) == 80, "Node8 must be 80 bytes");

// Triangle geometry in BVH leaf order, 48 bytes = 3 x float4:
//   g0 = (A.x, A.y, A.z, AB.x)  g1 = (AB.y, AB.z, AC.x, AC.y)  g2 = (AC.z, as_float(prim id), 0, 0)
// prim id = index into the caller's triangle array, or n_triangles + light index for an area-light triangle.
struct __align__(16) TriGeom { float4 g0, g1, g2; };
static_assert(sizeof(TriGeom) == 48, "TriGeom must be 48 bytes");

// Triangle shading record in caller order, 64 bytes = 4 x float4:
//   s0 = (nA.xyz, nB.x) s1 = (nB.yz, nC.xy) s2 = (nC.z, tA.x, tA.y, tB.x) s3 = (tB.y, tC.x, tC.y, as_float(material))
struct __align__(16) TriShade { float4 s0, s1, s2, s3; };

// Binary node in the reference's own BVH layout (BVH.hpp:56-60), used for the (few) planes and spheres whose
// visibility depends on the reference's box quirks (Plane::getAABB, Plane.cpp:103-109).
struct RefNode { float mnx, mny, mnz, mxx, mxy, mxz; int32_t indexOffset; int32_t numPrimitives; };
static_assert(sizeof(RefNode) == 32, "RefNode must be 32 bytes");

struct SphereRec { float cx, cy, cz, sqRadius; int32_t material; int32_t id; int32_t pad0, pad1; };
struct PlaneRec { float nx, ny, nz, px, py, pz; int32_t material; int32_t id; };

struct __align__(16) MaterialRec { float Le[3], Kd[3], Ks[3], Kt[3]; float ior; int32_t texture; float pad[2]; };   // texture: index into SceneDev::textures or -1
// Texture (MobileRT/Texture.hpp): texels of all textures live in one byte buffer
struct TextureRec { uint32_t offset; int32_t width, height, channels; uint32_t bytes; };
static_assert(sizeof(MaterialRec) == 64, "MaterialRec must be 64 bytes");

struct __align__(16) LightRec { int32_t kind; float A[3], AB[3], AC[3], Le[3]; float pad[3]; };
static_assert(sizeof(LightRec) == 64, "LightRec must be 64 bytes");

struct CameraRec {
    int32_t kind;
    v3 position, direction, right, up;
    float a, b;  // perspective: hFov, vFov in radians; orthographic: sizeH/2, sizeV/2
};

// One RegularGrid<T> of the reference (mrt_grid.h): cell c = x + (y << 5) + (z << 10) holds items[start[c] .. start[c + 1])
struct GridDev { v3 bmin, cellInv, cellSize; const uint32_t *start; const int32_t *items; };

enum HitKind : int32_t { HIT_NONE = 0, HIT_PLANE = 1, HIT_SPHERE = 2, HIT_TRIANGLE = 3, HIT_LIGHT = 4 };

// Everything a traversal / shading kernel needs, passed by value as a kernel parameter.
struct SceneDev {
    const Node8 *nodes;        // BVH8 over triangles + area-light triangles
    const TriGeom *tri_geom;   // leaf order
    const TriShade *tri_shade; // caller order (scene triangles only)
    int32_t n_triangles;       // scene triangles (prim ids >= n_triangles are lights)
    int32_t n_bvh_prims;       // scene triangles + area-light triangles
    v3 tri_bmin, tri_bmax;     // exact box of the scene triangles = the root box of the reference's BVH<Triangle>
    const RefNode *sphere_nodes; const SphereRec *spheres; int32_t n_spheres;
    const RefNode *plane_nodes;  const PlaneRec *planes;   int32_t n_planes;
    const MaterialRec *materials; int32_t n_materials;
    const TextureRec *textures; const uint8_t *texels; int32_t n_textures;
    const LightRec *lights; int32_t n_lights;
    const float *tile_order;     // 256
    const float *shader_table;   // 1 << 20
    const float *static_halton;  // 1 << 20
    int32_t tune_postpone, tune_refill, tune_wait, tune_wait_any;   // traversal scheduling thresholds in lanes (see mrt_traverse.cuh / mrt_wavefront.cu)
    int32_t naive;               // 1: Naive accelerator semantics (no triangle root-box test; planes / spheres behind an unbounded box)
    uint32_t prmt_one;           // 0x3F800000, deliberately a run-time value (mrt_traverse.cuh, u8_to_unit)
    const float *pixel_table;    // table the pixel sampler reads (static_halton or static_mt), or null for Constant
    // 1: the table is a permutation of {k / 2^20} (what fillArrayWithHaltonSeq + std::shuffle produce, Utils.hpp:209-218), so the
    // value of a uniformly chosen slot can be computed instead of gathered (wavefront path tracer, mrt_wavefront.cu pt_draw)
    int32_t shader_table_is_perm, static_halton_is_perm;
    // MRT_ACC_REGULAR_GRID: 1 = every ray walks the reference's three grids (0 planes, 1 spheres, 2 triangles) instead of the
    // BVHs; planes / spheres / tri_geom are then stored in the caller's order and there is no BVH8 (nodes == nullptr)
    int32_t grid;
    GridDev grids[3];
};

struct RenderParams {
    int32_t width, height, spp, spl, shader;
    int32_t bsx, bsy;           // tile size (Renderer.cpp:33-34)
    int32_t pixel_sampler;      // MRT_SAMPLER_* resolved
    v3 depth_max_point;
    uint32_t seed;
    int32_t tile_begin, tile_stride;   // this handle renders tiles tile_begin, tile_begin + tile_stride, ... of the shuffled tile walk (multi-GPU tile partition)
    int32_t n_tiles;                   // how many tiles that is (kNumberOfTiles without a partition)
    uint32_t sampler_domain;           // Sampler::domainSize_ (Sampler.cpp:14-18) of the HaltonSeq / Stratified pixel samplers
    uint32_t pixel_stream_n;           // MRT_SAMPLER_STREAM: draws in the caller's stream (>= 1)
};

// MRT_FLAG_REPLAY_SAMPLERS (mrt_shade.cuh DrawState, mrt_render.h ReplayArgs)
constexpr int kReplayMaxLights = 8;
enum ReplayMode : int32_t {
    REPLAY_OFF = 0,           // hashed slots
    REPLAY_COUNT_INDEX = 1,   // count this pixel-sample's Shader::getLightIndex draws, shade nothing
    REPLAY_COUNT_LIGHTS = 2,  // getLightIndex draws exact; count the draws each light's own sampler receives
    REPLAY_EXACT = 3          // every draw at the reference's counter value
};

}  // namespace mrt
