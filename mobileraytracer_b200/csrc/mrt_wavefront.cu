// Wavefront path tracer: PathTracer::shade (Components/Shaders/PathTracer.cpp:22-142) driven by Shader::rayTrace
// (MobileRT/Shader.cpp:86-123), re-expressed as generate / extend / shade / shadow / fold / finalize / resolve kernels
// over compacted ray queues in HBM. Built with -fmad=false (see mrt_math.cuh).
//
// The reference's recursion has one non-local rule: at a vertex with direct light (Ld > 0) the WHOLE indirect diffuse
// term is dropped when the diffuse chain below it ends on an emitter (PathTracer.cpp:99-110, the bool returned up the
// recursion). A "chain" is therefore the unit of bookkeeping: a maximal sequence of diffuse bounces; specular /
// transmission rays start child chains that are folded into the parent at the depth of the spawning vertex.
// With j* = the first depth of the chain whose direct light was > 0, the recursion's result is
//   not ended on an emitter : sum of the contributions of all depths
//   ended on an emitter     : sum of the contributions of the depths <= j* (+ the emitter term when there is no j*)
// Levels are processed in depth order, so a 32-byte accumulator per chain is enough: `pre` collects what counts in both
// cases (depths <= j*, everything while j* is still unknown, the emitter term), `post` what only counts when the chain
// does not end on an emitter; total = pre + (ended ? 0 : post). The first version kept one sum per depth in a 128-byte
// record per chain: 4.2 GB of records per 16-sample pass, written by generate, updated by atomics from the shadow
// kernel (5x its algorithmic DRAM bytes, profiles/r1_traffic_s40.csv) and read back by resolve.
#include "mrt_wavefront.h"
#include "mrt_shade.cuh"
#include <cstdlib>
#include <mutex>
#include <vector>

namespace mrt {

constexpr int kLevels = kRayDepthMax;          // depths 1..6 are traced; a depth-7 ray contributes nothing (PathTracer.cpp:24-26)
constexpr int kWfThreads = 128;
#ifndef MRT_TRAV_MIN_BLOCKS
#define MRT_TRAV_MIN_BLOCKS 8     // CTAs per SM of the persistent traversal kernels (64 registers per thread)
#endif
#ifndef MRT_SHADE_MIN_BLOCKS
#define MRT_SHADE_MIN_BLOCKS 6    // CTAs per SM of the shade kernel: 768 threads, 80 registers per thread (1024 threads = 64 registers spilled: +37 %; 512: +24 %; 256: 2.2x, profiles/r2_ab_shade.txt)
#endif

// One sector per chain. info: bits 0..2 = j* (0: no vertex had direct light yet), bit 3 = ended on an emitter,
// bits 4..6 = depth of the parent's vertex that spawned this chain. parent: parent chain, or -1 - pixel for a camera path.
struct __align__(16) PathAcc {
    float pre[3];
    uint32_t info;
    float post[3];
    int32_t parent;
};
static_assert(sizeof(PathAcc) == 32, "PathAcc must be one 32-byte sector");
constexpr uint32_t kAccFirstMask = 7u, kAccEnded = 8u;

struct BatchInts { int32_t v[kWfMaxBatch]; };   // per sample of a pass: sample index, or sample number of the running average

struct Queues {
    // ray queue (SoA of float4): r0 = o.xyz d.x | r1 = d.y d.z skip depth | r2 = T.xyz chain | r3 = q branch - -
    // level 0 (camera rays) only has r0 and r1 = d.y d.z q pixel: T = 1, chain = slot, depth = 1, nothing to skip
    float4 *r0[2], *r1[2], *r2[2], *r3[2];
    float4 *hit;                 // t u v (kind << 28 | idx)
    float4 *s0, *s1, *s2;        // shadow queue: o.xyz d.x | d.y d.z dist skip | c.xyz (chain | depth-1 << 28 | Ld > 0 << 31)
    uint8_t *vis;                // per shadow ray: 1 = reached the light (written by the shadow kernel, consumed by k_wf_fold)
    PathAcc *acc;                // per chain; the root chain of camera path `slot` is chain `slot`
    uint32_t *counts;            // [0..7] ray count per level, [8] shadow count, [9] chain pool size, [10..17] chain pool size after each level,
                                 // [20],[21] work counters
    uint32_t cap_rays, cap_shadow, cap_chains;
    // MRT_FLAG_SORT_BY_MATERIAL: the shade kernel visits the hits of a level in the order perm[0 .. n), binned by sort key
    uint32_t *perm;              // null when the flag is off
    uint32_t *sort_bins;         // kSortBins + 1 counters / offsets
    int32_t sort_key;            // 0: material index of the hit primitive; 1: primitive id (hit locality)
    uint32_t sort_shift;         // sort_key 1: bin = primitive id >> sort_shift
};

struct WfMark { cudaEvent_t ev; int cls; };

struct Wavefront {
    Queues q {};
    size_t paths_per_sample = 0;      // camera paths of one sample (the pixels of the tiles this handle renders)
    int sm_count = 148;
    int batch = 1;
    int spl = 1;
    int factor = 2;                   // ray-queue capacity in camera paths of a full pass (wf_alloc_queues)
    int sort_mode = 0;                // 0: off; 1: bin the hits by material before shading; 2: by primitive id (A/B of hit locality)
    uint32_t sort_shift = 0;
    size_t bytes = 0;
    std::vector<void *> allocs;
    std::vector<WfMark> marks;        // one event after every launch; the class says what ran before it
    std::vector<cudaEvent_t> pool;
    size_t used = 0;
    // fold(level) next to extend(level + 1): see wavefront_trace_batch
    int overlap = 0;
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
};

// A fresh accumulator, written as two full 16-byte stores: the whole sector is overwritten, so the L2 never has to
// read-modify-write it (a record with pad bytes left unwritten cost 66 MB of DRAM reads per generate launch, ncu).
MRT_D void acc_init(PathAcc *a, v3 pre, uint32_t info, int32_t parent) {
    float4 *p = reinterpret_cast<float4 *>(a);
    p[0] = make_float4(pre.x, pre.y, pre.z, __uint_as_float(info));
    p[1] = make_float4(0.0f, 0.0f, 0.0f, __int_as_float(parent));
}

MRT_D uint32_t pack_skip(int kind, int idx) { return kind == HIT_NONE ? 0u : (static_cast<uint32_t>(kind) << 28) | (static_cast<uint32_t>(idx) & 0x0fffffffu); }
MRT_D void unpack_skip(uint32_t s, int &kind, int &idx) { kind = static_cast<int>(s >> 28); idx = kind == HIT_NONE ? -1 : static_cast<int>(s & 0x0fffffffu); }

// ---- generate: Renderer::renderScene's jitter + Camera::generateRay (Renderer.cpp:130-141) ---------------------------
__global__ void __launch_bounds__(kWfThreads) k_wf_generate(Queues Q, SceneDev s, CameraRec cam, RenderParams rp,
                                                            const int2 *__restrict__ tile_origin, const uint32_t *__restrict__ tile_base,
                                                            uint32_t roots_per_sample, BatchInts samples) {
    const int sample = samples.v[blockIdx.y];                       // blockIdx.y: position of the sample in the pass
    const uint32_t slot_base = blockIdx.y * roots_per_sample;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int patch = blockIdx.x * (kWfThreads / 32) + warp;
    uint32_t q = 0;
    int x, y, pix;
    int tile, rank;
    if (tile_pixel(rp, tile_origin, patch, lane, sample, x, y, pix, q, &tile, &rank)) {
        const float invImgWidth = 1.0f / static_cast<float>(rp.width), invImgHeight = 1.0f / static_cast<float>(rp.height);
        const float pixelWidth = 0.5f / static_cast<float>(rp.width), pixelHeight = 0.5f / static_cast<float>(rp.height);
        const float u = static_cast<float>(x) * invImgWidth, v = static_cast<float>(y) * invImgHeight;
        const float r1 = pixel_draw(s, rp, 2u * q), r2 = pixel_draw(s, rp, 2u * q + 1u);
        v3 o, d;
        generate_ray(cam, u, v, (r1 - 0.5f) * 2.0f * pixelWidth, (r2 - 0.5f) * 2.0f * pixelHeight, o, d);
        // deterministic slot: no atomics (64 k same-address atomics per sample made this kernel latency bound, profiles/)
        const uint32_t slot = slot_base + __ldg(tile_base + tile) + static_cast<uint32_t>(rank);
        // 32 bytes per camera ray; its accumulator (chain `slot`) is initialised by the shade kernel of level 0
        Q.r0[0][slot] = make_float4(o.x, o.y, o.z, d.x);
        Q.r1[0][slot] = make_float4(d.y, d.z, __uint_as_float(q), __int_as_float(pix));
    }
    // (the camera rays are counted by k_wf_begin_levels: one same-address atomicAdd per warp here, 1 M per launch at about 0.7 ns
    // each, was three quarters of this kernel's time)
}

// the chain pool starts right after the root chains: camera paths of batch sample i are chains [i, i + 1) * roots_per_sample
__global__ void k_wf_begin_levels(Queues Q, uint32_t n_samples, uint32_t roots_per_sample, DeviceCounters *dc) {
    Q.counts[0] = n_samples * roots_per_sample; Q.counts[9] = n_samples * roots_per_sample;
    dc->rays_primary += static_cast<unsigned long long>(n_samples) * roots_per_sample;   // every path slot holds one camera ray (Ray.cpp:11)
}

// ---- extend / shadow: persistent traversal kernels --------------------------------------------------------------
// One CTA of 4 warps per scheduling slot, resident for the whole launch. A lane that has finished its ray fetches the
// next one from the level's work counter (one warp-aggregated atomic per refill), so the lanes of a warp stay busy although
// their rays need very different numbers of steps; refills happen when fewer than s.tune_refill lanes still hold a ray.
//
// Node tests and triangle tests are scheduled separately, because after any single node step only a few lanes have hit
// a leaf (ncu, profiles/r1_s25_extend_*: 7 of 32 threads active in per-lane triangle code, which then cost a third of
// the kernel for 3 % of its arithmetic):
//   node phase     : every lane with pending inner nodes visits one BVH8 node; a lane that hit leaf children appends ONE
//                    node item (owner lane, triangle base, hit leaf slots, the node's meta words) to a per-warp list in
//                    shared memory, at a position taken from the ballot (no atomics, no per-leaf loop);
//   triangle phase : whenever 32 node items are waiting, every lane expands one item into the bit mask of its triangles, a
//                    warp scan lays the triangles out in a list, and the lanes test one TRIANGLE each per round. The
//                    ray of the owner is read from shared memory; a hit is committed with one 64-bit atomicMin on the
//                    owner's key (t bits << 32 | primitive id), which implements "nearest t, then lowest primitive id"
//                    (the reference keeps the first hit of its left-first walk at equal t, mrt_traverse.cuh)
//                    independently of the order the triangles are processed in.
// A lane whose node work is finished may still own items in the list; it retires once the list has been emptied, which
// is forced when s.tune_wait (closest hit) / s.tune_wait_any (any hit) lanes are waiting or nobody has node work left.
constexpr int kCounterExtend = 20, kCounterShadow = 21;   // Queues::counts slots used as work counters
#ifndef MRT_QUEUE_AHEAD
#define MRT_QUEUE_AHEAD 2048
#endif
constexpr uint32_t kQueueAhead = MRT_QUEUE_AHEAD;          // rays between a refill and the queue entries it prefetches
constexpr uint32_t kNodeCap = 64;                         // node items per warp: < 32 left after a pass + at most one per lane and step
constexpr uint32_t kTriCap = 256;                         // triangle entries of one (sub-)pass

struct __align__(16) WarpShared {
    unsigned long long key[32];   // per lane: best hit so far, (float bits of t) << 32 | primitive id
    unsigned long long key0[32];  // the key the ray started with: tbest of the plane / sphere pass (or the light distance), id 0
    float2 uv[32];                // barycentrics of that hit
    // ray state that only the triangle phase and the retire step need lives here instead of in registers (the node step
    // is register bound: at 64 registers it spilled its own operands, ncu SASS view of the first cooperative version)
    float ox[32], oy[32], oz[32], dx[32], dy[32], dz[32];
    int32_t skip[32];             // primitive id the ray leaves (Triangle.cpp:64), or -1
    uint32_t cur[32];             // queue index of the lane's ray
    uint32_t pre[32];             // packed plane / sphere hit found before the triangles (kind << 28 | idx)
    uint4 nitems[kNodeCap];       // one per node step that hit leaves: triangle base | hit leaf slots + owner lane << 8 | meta words
    uint16_t tris[kTriCap];       // the triangles of one pass: lane holding the node item << 5 | offset in the node's triangle run
};

MRT_D uint32_t pack_hit(const HitRec &h) { return (static_cast<uint32_t>(h.kind) << 28) | (static_cast<uint32_t>(h.idx) & 0x0fffffffu); }

// The warp tests the triangles of the node items nitems[first, first + n), n <= 32. Must be called by all 32 lanes.
// Every lane expands one node item (hit leaf slots + meta bytes) into the bit mask of its triangles within the node's
// run of <= 24; a warp scan gives every item its range in the triangle list, then the lanes take one TRIANGLE each per
// round. The expansion used to happen per leaf at append time, in a loop that ran with 3.8 of 32 lanes active and cost
// 6 % of the kernel's instructions, and the per-leaf triangle loop kept 9-15 lanes busy (profiles/r1_extend_s31).
template <bool ANY>
MRT_D void coop_triangle_pass(const SceneDev &s, WarpShared &w, uint32_t first, uint32_t n, int lane) {
    const uint32_t full = 0xffffffffu;
    const bool valid = static_cast<uint32_t>(lane) < n;
    const uint4 it = valid ? w.nitems[first + lane] : make_uint4(0u, 0u, 0u, 0u);
    const uint32_t owner = (it.y >> 8) & 31u;
    uint32_t tm = 0u;
    for (uint32_t lh = it.y & 0xffu; lh != 0u; lh &= lh - 1u) {
        const uint32_t mb = leaf_meta(it.z, it.w, __ffs(lh) - 1u);
        tm |= ((1u << (mb >> 5)) - 1u) << (mb & 31u);
    }
    const uint32_t cnt_all = __popc(tm);
    const uint32_t total = __reduce_add_sync(full, cnt_all);
    const uint32_t step = total <= kTriCap ? 32u : 8u;      // 8 node items hold at most 192 triangles
    for (uint32_t lo = 0; lo < n; lo += step) {
        const uint32_t cnt = (static_cast<uint32_t>(lane) - lo < step) ? cnt_all : 0u;
        uint32_t incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(full, incl, d);
            if (lane >= d) { incl += v; }
        }
        const uint32_t tot = __shfl_sync(full, incl, 31);
        {
            uint32_t pos = incl - cnt;
            for (uint32_t m = cnt ? tm : 0u; m != 0u; m &= m - 1u) {
                w.tris[pos++] = static_cast<uint16_t>((static_cast<uint32_t>(lane) << 5) | (__ffs(m) - 1u));
            }
        }
        __syncwarp();
        for (uint32_t base = 0; base < tot; base += 32u) {
            const uint32_t idx = base + static_cast<uint32_t>(lane);
            const bool have = idx < tot;
            const uint32_t e = have ? w.tris[idx] : 0u;
            const uint32_t tb = __shfl_sync(full, it.x, e >> 5);
            const int ow = static_cast<int>(__shfl_sync(full, owner, e >> 5));
            unsigned long long mykey = ~0ull;
            float bu = 0.0f, bv = 0.0f;
            if (have) {
                const float4 *tp = reinterpret_cast<const float4 *>(s.tri_geom) + 3ull * (tb + (e & 31u));
                const float4 g0 = __ldg(tp), g1 = __ldg(tp + 1), g2 = __ldg(tp + 2);
                const uint32_t prim = __float_as_uint(g2.y);
                // self-hit skip (Triangle.cpp:64); area lights do not occlude (Shader.cpp:132-158)
                if (!(static_cast<int>(prim) == w.skip[ow] || (ANY && static_cast<int>(prim) >= s.n_triangles))) {
                    MRT_STAT(atomicAdd(&g_stat_tris, 1ull);)
                    const v3 o = mk3(w.ox[ow], w.oy[ow], w.oz[ow]);
                    const v3 d = mk3(w.dx[ow], w.dy[ow], w.dz[ow]);
                    const float tbest = __uint_as_float(reinterpret_cast<const uint2 *>(&w.key[ow])->y);
                    float tt, u, v;
                    if (tri_hit(g0, g1, g2, o, d, tbest, tt, u, v)) {
                        mykey = (static_cast<unsigned long long>(__float_as_uint(tt)) << 32) | prim; bu = u; bv = v;
                    }
                }
            }
            bool won = false;
            if (mykey != ~0ull) { won = atomicMin(&w.key[ow], mykey) > mykey; }
            if (!ANY) {
                __syncwarp();
                if (won && w.key[ow] == mykey) { w.uv[ow] = make_float2(bu, bv); }   // the one final winner of this round
            }
            __syncwarp();
        }
    }
}

// ANY = false: closest hit for every ray of the level (Shader::rayTrace's geometric part, Shader.cpp:86-111) -> Q.hit
// ANY = true : any hit within the light distance for every shadow ray (Shader::shadowTrace, Shader.cpp:132-158)
template <bool ANY>
MRT_D void wf_traverse(const SceneDev &s, const Queues &Q, int level, DeviceCounters *dc) {
    __shared__ WarpShared wsh[kWfThreads / 32];
#ifndef MRT_NO_PERM_TABLE
    __shared__ uint8_t perm[8 * 256];   // perm[o * 256 + m]: bit k of m moved to k ^ o (ray-octant order of a node's children)
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += kWfThreads) {
        const uint32_t o = i >> 8, m = i & 255u;
        uint32_t r = 0u;
        for (uint32_t k = 0; k < 8u; ++k) { if (m & (1u << k)) { r |= 1u << (k ^ o); } }
        perm[i] = static_cast<uint8_t>(r);
    }
    __syncthreads();
    const uint8_t *permp = perm;
#else
    const uint8_t *permp = nullptr;
#endif
    WarpShared &w = wsh[threadIdx.x >> 5];
    const uint32_t full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const uint32_t lane_bit = 1u << lane;
    const int b = level & 1;
    const uint32_t n = ANY ? min(Q.counts[8], Q.cap_shadow) : min(Q.counts[level], Q.cap_rays);
    uint32_t *work = &Q.counts[ANY ? kCounterShadow : kCounterExtend];
    int overflow = 0;
    bool exhausted = false;
    bool node_work = false;          // this lane has pending inner nodes
    uint32_t hasmask = 0u;           // lanes holding a ray (warp-uniform)
    uint32_t outmask = 0u;           // lanes that may own items in the list (warp-uniform)
    uint32_t busy = 0u;              // ballot of node_work (warp-uniform)
    uint32_t count = 0;              // node items in the warp's list (warp-uniform)
    Trav8 t;
    TravMem tmem;
    t.o = mk3(0.0f, 0.0f, 0.0f);
    t.ngroup = make_uint2(0u, 0u); t.sp = 0; t.tbest = 0.0f;
    MRT_STAT(t.n_nodes = 0u; t.n_tris = 0u;)

    // empties the warp's item list (top chunk first); afterwards nobody owns pending items
#define MRT_FLUSH_ITEMS()                                                                         \
    do {                                                                                          \
        while (count > 0u) {                                                                      \
            const uint32_t m_ = count < 32u ? count : 32u;                                        \
            coop_triangle_pass<ANY>(s, w, count - m_, m_, lane);                               \
            count -= m_;                                                                          \
        }                                                                                         \
        outmask = 0u;                                                                             \
    } while (0)
    // after triangle passes: the lane's own best hit bounds its further node tests; an occluded shadow ray stops
#define MRT_REFRESH_BEST()                                                                        \
    do {                                                                                          \
        if (hasmask & lane_bit) {                                                                 \
            const unsigned long long k_ = w.key[lane];                                            \
            if (ANY) { if (k_ != w.key0[lane]) { t.ngroup.y = 0u; t.sp = 0; node_work = false; } }        \
            else { t.tbest = __uint_as_float(static_cast<uint32_t>(k_ >> 32)); }                  \
        }                                                                                         \
        busy = __ballot_sync(full, node_work);                                                    \
    } while (0)

    while (true) {
        // retire: node work done and no item of the lane left in the list. A finished lane idles until the next refill
        // anyway, so the (divergent) retire block runs together with the refill instead of once per finished lane
        // (it ran in 45 % of the iterations with 4.7 lanes active, profiles/r1_extend_s31).
        const uint32_t done = hasmask & ~busy & ~outmask;
        const bool refill_now = !exhausted && __popc(~hasmask | done) > 32 - s.tune_refill;
        if (done && (refill_now || exhausted)) {
            if (done & lane_bit) {
                const unsigned long long k = w.key[lane], key0 = w.key0[lane];
                const uint32_t cur = w.cur[lane];
                if (ANY) {
                    Q.vis[cur] = k == key0 ? 1 : 0;      // one byte per shadow ray; k_wf_fold adds the light (PathTracer.cpp:72-80)
                } else {
                    HitRec h;
                    const uint32_t pre = w.pre[lane];
                    h.t = __uint_as_float(static_cast<uint32_t>(key0 >> 32)); h.kind = static_cast<int>(pre >> 28);
                    h.idx = h.kind == HIT_NONE ? -1 : static_cast<int>(pre & 0x0fffffffu); h.u = 0.0f; h.v = 0.0f;
                    if (k != key0) {
                        const int prim = static_cast<int>(static_cast<uint32_t>(k));
                        const float2 uv = w.uv[lane];
                        h.t = __uint_as_float(static_cast<uint32_t>(k >> 32)); h.u = uv.x; h.v = uv.y;
                        if (prim < s.n_triangles) { h.kind = HIT_TRIANGLE; h.idx = prim; }
                        else { h.kind = HIT_LIGHT; h.idx = prim - s.n_triangles; }
                    }
                    Q.hit[cur] = make_float4(h.t, h.u, h.v, __uint_as_float(pack_hit(h)));
                }
                MRT_STAT(atomicAdd(&g_stat_nodes, t.n_nodes); atomicAdd(&g_stat_rays, 1ull); t.n_nodes = 0u;)
            }
            hasmask &= ~done;
        }
        // refill the idle lanes (one warp-aggregated atomicAdd on the level's work counter per refill; handing every warp chunks
        // of 64-128 rays per atomicAdd instead was measured slower twice, extend +1.7 %, shadow +2.8 % even when a refill is
        // completed from the next chunk: warps that run side by side then no longer work on neighbouring queue entries,
        // profiles/r2_ab_sign_pack_and_tuning.txt)
        if (refill_now) {
            const uint32_t idle = ~hasmask;
            uint32_t base = 0;
            if (lane == 0) { base = atomicAdd(work, static_cast<uint32_t>(__popc(idle))); }
            base = __shfl_sync(full, base, 0);
            if (base + __popc(idle) >= n) { exhausted = true; }
            bool got = false;
            if (idle & lane_bit) {
                const uint32_t cur = base + __popc(idle & (lane_bit - 1u));
                if (cur < n) {
                    HitRec h;
                    const float4 a = ANY ? Q.s0[cur] : Q.r0[b][cur], c = ANY ? Q.s1[cur] : Q.r1[b][cur];
#ifndef MRT_NO_QUEUE_PREFETCH
                    // the ray that somebody will fetch kQueueAhead refills from now: start its trip from HBM to L2 (the refill
                    // is a divergent block the whole warp waits for; every queue entry is prefetched exactly once)
                    if (cur + kQueueAhead < n) {
                        asm volatile("prefetch.global.L2 [%0];" :: "l"((ANY ? Q.s0 : Q.r0[b]) + cur + kQueueAhead));
                        asm volatile("prefetch.global.L2 [%0];" :: "l"((ANY ? Q.s1 : Q.r1[b]) + cur + kQueueAhead));
                    }
#endif
                    const v3 o = mk3(a.x, a.y, a.z), d = mk3(a.w, c.x, c.y);
                    int sk, si;
                    unpack_skip(ANY ? __float_as_uint(c.w) : (level == 0 ? 0u : __float_as_uint(c.z)), sk, si);
                    h.t = ANY ? c.z : kRayLengthMax; h.kind = HIT_NONE; h.idx = -1; h.u = 0.0f; h.v = 0.0f;
                    bool occluded = false;
                    if (ANY) {
                        occluded = trace_ref_bvh<0>(s, o, d, true, sk, si, h, &overflow) || trace_ref_bvh<1>(s, o, d, true, sk, si, h, &overflow);
                    } else {
                        trace_ref_bvh<0>(s, o, d, false, sk, si, h, &overflow);
                        trace_ref_bvh<1>(s, o, d, false, sk, si, h, &overflow);
                    }
                    if (ANY && occluded) { Q.vis[cur] = 0; }
                    if (!occluded) {
                        if (s.n_bvh_prims > 0 && triangles_root_hit(s, o, d)) {
                            trav8_init(t, o, d, sk == HIT_TRIANGLE ? si : -1, h.t);
                            const unsigned long long key0 = static_cast<unsigned long long>(__float_as_uint(h.t)) << 32;
                            w.key[lane] = key0; w.key0[lane] = key0;
                            w.ox[lane] = o.x; w.oy[lane] = o.y; w.oz[lane] = o.z; w.dx[lane] = d.x; w.dy[lane] = d.y; w.dz[lane] = d.z;
                            w.skip[lane] = sk == HIT_TRIANGLE ? si : -1;
                            w.cur[lane] = cur; w.pre[lane] = pack_hit(h);
                            got = true; node_work = true;
                        } else if (ANY) {
                            Q.vis[cur] = 1;
                        } else {
                            if (s.n_bvh_prims > s.n_triangles) { trace_lights_linear(s, o, d, h); }
                            Q.hit[cur] = make_float4(h.t, h.u, h.v, __uint_as_float(pack_hit(h)));
                        }
                    }
                }
            }
            hasmask |= __ballot_sync(full, got);
            busy = __ballot_sync(full, node_work);
        }
        if (busy == 0u) {
            if (count > 0u) { MRT_FLUSH_ITEMS(); MRT_REFRESH_BEST(); continue; }
            if (hasmask == 0u && exhausted) { break; }
            continue;
        }
        // node phase
        uint32_t tbase = 0u, lh = 0u, m0 = 0u, m1 = 0u;
        {
            // pop speculatively (the load does not wait inside a divergent branch)
            const uint2 top = stack_get(tmem, t.sp > 0 ? t.sp - 1 : 0);
            if (node_work) {
                if (!(t.ngroup.y & 0xff000000u)) { t.ngroup = top; --t.sp; }
                if (!trav8_node_core(s, t, tmem, tbase, lh, m0, m1, &overflow, permp)) { t.ngroup.y = 0u; t.sp = 0; lh = 0u; }
                node_work = (t.ngroup.y & 0xff000000u) != 0u || t.sp > 0;
            }
        }
        // append one node item per lane that hit leaves: positions from the ballot, no atomics, no loop
        {
            const uint32_t emit = __ballot_sync(full, lh != 0u);
            if (emit != 0u) {
                if (lh != 0u) {
                    w.nitems[count + __popc(emit & (lane_bit - 1u))] = make_uint4(tbase, lh | (static_cast<uint32_t>(lane) << 8), m0, m1);
                }
                count += __popc(emit);
                outmask |= emit;
                __syncwarp();
            }
        }
        // triangle phase: full chunks only
        if (count >= 32u) {
            do {
                coop_triangle_pass<ANY>(s, w, count - 32u, 32u, lane);
                count -= 32u;
            } while (count >= 32u);
            if (count == 0u) { outmask = 0u; }
            MRT_REFRESH_BEST();
        } else {
            busy = __ballot_sync(full, node_work);
        }
        // lanes that only wait for their items: force the triangle phase when enough of them pile up
        if (__popc(hasmask & ~busy & outmask) >= (ANY ? s.tune_wait_any : s.tune_wait)) { MRT_FLUSH_ITEMS(); MRT_REFRESH_BEST(); }
    }
#undef MRT_FLUSH_ITEMS
#undef MRT_REFRESH_BEST
    if (overflow) { atomicOr(&dc->overflow, 2); }   // traversal stack
}

__global__ void __launch_bounds__(kWfThreads, MRT_TRAV_MIN_BLOCKS) k_wf_extend(SceneDev s, Queues Q, int level, DeviceCounters *dc) { wf_traverse<false>(s, Q, level, dc); }
__global__ void __launch_bounds__(kWfThreads, MRT_TRAV_MIN_BLOCKS) k_wf_shadow(SceneDev s, Queues Q, DeviceCounters *dc) { wf_traverse<true>(s, Q, 0, dc); }

// MRT_ACC_REGULAR_GRID: the same queue records in and out, one thread per ray walking the reference's grids (trace_grid,
// mrt_traverse.cuh). Not tuned: this mode exists for its visibility rules, not for speed.
__global__ void __launch_bounds__(256) k_wf_extend_grid(SceneDev s, Queues Q, int level) {
    const int b = level & 1;
    const uint32_t n = min(Q.counts[level], Q.cap_rays);
    for (uint32_t cur = blockIdx.x * blockDim.x + threadIdx.x; cur < n; cur += gridDim.x * blockDim.x) {
        const float4 a = Q.r0[b][cur], c = Q.r1[b][cur];
        int sk, si;
        unpack_skip(level == 0 ? 0u : __float_as_uint(c.z), sk, si);
        HitRec h;
        h.t = kRayLengthMax; h.kind = HIT_NONE; h.idx = -1; h.u = 0.0f; h.v = 0.0f;
        trace_closest_grid(s, mk3(a.x, a.y, a.z), mk3(a.w, c.x, c.y), sk, si, h);
        Q.hit[cur] = make_float4(h.t, h.u, h.v, __uint_as_float(pack_hit(h)));
    }
}
__global__ void __launch_bounds__(256) k_wf_shadow_grid(SceneDev s, Queues Q) {
    const uint32_t n = min(Q.counts[8], Q.cap_shadow);
    for (uint32_t cur = blockIdx.x * blockDim.x + threadIdx.x; cur < n; cur += gridDim.x * blockDim.x) {
        const float4 a = Q.s0[cur], c = Q.s1[cur];
        int sk, si;
        unpack_skip(__float_as_uint(c.w), sk, si);
        Q.vis[cur] = trace_shadow_grid(s, mk3(a.x, a.y, a.z), mk3(a.w, c.x, c.y), c.z, sk, si) ? 0 : 1;
    }
}

// Shader::getCosineSampleHemisphere (Shader.cpp:188-216)
MRT_D v3 cosine_sample_hemisphere(v3 normal, float uniformRandom1, float uniformRandom2) {
    const float phi = 6.28318530717958647692528676655900576f * uniformRandom1;
    const float r2 = uniformRandom2;
    const float cosTheta = sqrtf(r2);
    v3 u = fabsf(normal.x) > 0.1f ? mk3(0.0f, 1.0f, 0.0f) : mk3(1.0f, 0.0f, 0.0f);
    u = normalize(cross(u, normal));
    const v3 v = cross(normal, u);
    v3 direction = (u * (cosf(phi) * cosTheta) + v * (sinf(phi) * cosTheta)) + normal * sqrtf(1.0f - r2);
    return normalize(direction);
}

// ---- optional: bin the hits of a level by material (north_star 3: "ray queues ... sorted by material") -------------
// A counting sort of hit indices: histogram per CTA chunk in shared memory, one global atomicAdd per bin and CTA to reserve
// the bin's range, then the scatter. The shade kernel then walks perm[] instead of 0 .. n-1. Every material runs the same
// shading code here (materials are data, Material.hpp:18-43), so what binning can buy is locality of the records the
// shade kernel gathers (materials; with sort_key 1 the 64-byte TriShade records), paid for with three more kernels per
// level and gathered instead of streamed hit / ray records. Measured: profiles/r2_ab_material_sort.txt.
constexpr uint32_t kSortBins = 1024, kSortChunk = 8192;

MRT_D uint32_t sort_bin(const SceneDev &s, const Queues &Q, uint32_t i) {
    const uint32_t packed = __float_as_uint(Q.hit[i].w);
    const uint32_t kind = packed >> 28, idx = packed & 0x0fffffffu;
    if (kind != HIT_TRIANGLE) { return kSortBins - 1u - min(kind, 7u); }          // misses, lights, planes, spheres: bins of their own at the end
    if (Q.sort_key == 1) { return min(idx >> Q.sort_shift, kSortBins - 9u); }
    const uint32_t m = static_cast<uint32_t>(__float_as_int(__ldg(reinterpret_cast<const float4 *>(s.tri_shade + idx) + 3).w));
    return min(m, kSortBins - 9u);
}

__global__ void __launch_bounds__(256) k_wf_sort_count(SceneDev s, Queues Q, int level) {
    __shared__ uint32_t local[kSortBins];
    const uint32_t n = min(Q.counts[level], Q.cap_rays);
    for (uint32_t k = threadIdx.x; k < kSortBins; k += 256u) { local[k] = 0u; }
    __syncthreads();
    const uint32_t c0 = blockIdx.x * kSortChunk, c1 = min(c0 + kSortChunk, n);
    for (uint32_t i = c0 + threadIdx.x; i < c1; i += 256u) { atomicAdd(&local[sort_bin(s, Q, i)], 1u); }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < kSortBins; k += 256u) { if (local[k]) { atomicAdd(&Q.sort_bins[k], local[k]); } }
}

__global__ void __launch_bounds__(kSortBins) k_wf_sort_scan(Queues Q) {   // exclusive scan of the bin counts (one CTA)
    __shared__ uint32_t tmp[kSortBins];
    const uint32_t k = threadIdx.x, v = Q.sort_bins[k];
    tmp[k] = v;
    __syncthreads();
    for (uint32_t d = 1; d < kSortBins; d <<= 1) {
        const uint32_t add = k >= d ? tmp[k - d] : 0u;
        __syncthreads();
        tmp[k] += add;
        __syncthreads();
    }
    Q.sort_bins[k] = tmp[k] - v;
}

__global__ void __launch_bounds__(256) k_wf_sort_scatter(SceneDev s, Queues Q, int level) {
    __shared__ uint32_t local[kSortBins], base[kSortBins];
    const uint32_t n = min(Q.counts[level], Q.cap_rays);
    for (uint32_t k = threadIdx.x; k < kSortBins; k += 256u) { local[k] = 0u; }
    __syncthreads();
    const uint32_t c0 = blockIdx.x * kSortChunk, c1 = min(c0 + kSortChunk, n);
    for (uint32_t i = c0 + threadIdx.x; i < c1; i += 256u) { atomicAdd(&local[sort_bin(s, Q, i)], 1u); }
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < kSortBins; k += 256u) { base[k] = local[k] ? atomicAdd(&Q.sort_bins[k], local[k]) : 0u; local[k] = 0u; }
    __syncthreads();
    for (uint32_t i = c0 + threadIdx.x; i < c1; i += 256u) {
        const uint32_t k = sort_bin(s, Q, i);
        Q.perm[base[k] + atomicAdd(&local[k], 1u)] = i;
    }
}

// ---- shade: PathTracer::shade for every hit of the level; emits shadow rays and the next level's rays --------------
// Where a draw of the reference's tables comes from (Shader.cpp:188-233, Sampler.hpp:58-63, PathTracer.cpp:91): the
// reference hands out table[counter++] in execution order, which no parallel schedule can replay for a path tracer
// (SURVEY H2), so every (pixel-sample q, vertex `branch`, ordinal, stream) reads its own decorrelated slot of the SAME
// tables. The ordinals are fixed per vertex, independent of control flow.
// The tables are shuffles of the radical inverses of 0 .. 2^20 - 1, i.e. permutations of {k / 2^20}; the value found in a
// uniformly chosen slot is therefore k / 2^20 for a uniformly chosen k, and is computed instead of gathered when the host
// has verified that property (SceneDev::*_is_perm). Six 4-byte gathers per hit cost six 32-byte sectors of L2 -> L1
// traffic, 40 % of what the shade kernel moved while it ran at 73 % of the measured L2 gather peak (profiles/).
MRT_D float pt_draw(const float *__restrict__ table, int is_perm, uint32_t q, uint32_t e, uint32_t stream, uint32_t seed) {
    const uint32_t i = hash_u32(q, e, stream ^ seed) & kArrayMask;
    // k / 2^20 without an integer-to-float conversion (I2F runs on the quarter-rate XU pipe): 1.mantissa - 1
    return is_perm ? __uint_as_float(0x3F800000u | (i << 3)) - 1.0f : __ldg(table + i);
}

#ifndef MRT_SHADE_THREADS
#define MRT_SHADE_THREADS 128     // 128 x 6 CTAs per SM: 4.49 ms per 16-sample pass; 256 x 3: 4.89 (more warps wait at the two barriers); 512 x 1: 6.78
#endif
constexpr int kShadeThreads = MRT_SHADE_THREADS;
constexpr int kShadeMatCap = 64, kShadeLightCap = 256;   // records staged in shared memory (64 bytes each); larger scenes read them from global memory

// One light sample of a vertex (PathTracer.cpp:60-82): Shader::getLightIndex (Shader.cpp:223-233), then
// PointLight::getPosition (PointLight.cpp:16-18) or AreaLight::getPosition (AreaLight.cpp:17-26). True when the light point
// lies in front of the surface, i.e. when the reference casts the shadow ray.
struct LightSample { v3 dir; float dist; v3 ld; };
MRT_D bool pt_light_sample(const SceneDev &s, const RenderParams &rp, const float4 *lightp, const Surface &sf, uint32_t q, uint32_t eb, uint32_t k,
                           LightSample &ls) {
    const float rl = pt_draw(s.shader_table, s.shader_table_is_perm, q, eb + k, DS_LIGHT_INDEX, rp.seed);
    const uint32_t chosen = static_cast<uint32_t>(floorf(rl * static_cast<float>(static_cast<uint32_t>(s.n_lights)) * 0.99999f));
    // LightRec as four float4: kind A.xyz | AB.xyz AC.x | AC.yz Le.xy | Le.z - - -
    const float4 l0 = lightp[4u * chosen], l1 = lightp[4u * chosen + 1u], l2 = lightp[4u * chosen + 2u], l3 = lightp[4u * chosen + 3u];
    v3 lightPosition = mk3(l0.y, l0.z, l0.w);
    if (__float_as_int(l0.x) != 0) {
        float r = pt_draw(s.static_halton, s.static_halton_is_perm, q, eb + 2u * k, DS_LIGHT_POS, rp.seed);
        float g = pt_draw(s.static_halton, s.static_halton_is_perm, q, eb + 2u * k + 1u, DS_LIGHT_POS, rp.seed);
        if (r + g >= 1.0f) { r = 1.0f - r; g = 1.0f - g; }
        lightPosition = (lightPosition + r * mk3(l1.x, l1.y, l1.z)) + g * mk3(l1.w, l2.x, l2.y);
    }
    v3 vectorToLight = lightPosition - sf.point;
    ls.dist = length(vectorToLight);
    ls.dir = normalize(vectorToLight);
    const float cosNormalLight = dot(sf.normal, ls.dir);
    // Ld's share of this light sample: (Le * cos) * kD / samplesLight
    ls.ld = ((mk3(l2.z, l2.w, l3.x) * cosNormalLight) * sf.Kd) / static_cast<float>(rp.spl);
    return cosNormalLight > 0.0f;
}

// Queue slots are handed out per CTA: every thread first decides what its hit emits (phase 1), a block-wide prefix sum turns
// the counts into offsets and ONE thread per queue bumps the global counter for all of the CTA's hits, then the records are written
// (phase 2). (Laying the CTA's records out by direction octant instead of hit order was measured too: traversal -0.5 %, this
// kernel +12 %, profiles/r2_ab_material_sort_and_octant_bins.txt.) With one warp-aggregated atomicAdd per push site the kernel spent 48 % of its stall samples waiting for the two
// same-address atomics of the shadow and ray queues (2 M per 33 M-hit launch, about 0.7 ns each at the L2;
// profiles/r2_shade_before_ncu_summary.txt). Consecutive hits now also get consecutive slots.
template <bool L0>
__global__ void __launch_bounds__(kShadeThreads, MRT_SHADE_MIN_BLOCKS) k_wf_shade(SceneDev s, RenderParams rp, Queues Q, int level, DeviceCounters *dc) {
    // materials and lights: a few KB that every hit needs; out of shared memory instead of 2 x 64 bytes of L2 -> L1 traffic per hit
    __shared__ float4 sm_mat[kShadeMatCap * 4];
    __shared__ float4 sm_light[kShadeLightCap * 4];
    __shared__ unsigned long long sm_warp[kShadeThreads / 32];   // per warp: shadow rays | rays << 24 | chains << 44 of this iteration
    __shared__ uint32_t sm_base[3];
    const bool mat_sh = s.n_materials <= kShadeMatCap, light_sh = s.n_lights <= kShadeLightCap;
    if (mat_sh) { for (int k = threadIdx.x; k < s.n_materials * 4; k += kShadeThreads) { sm_mat[k] = __ldg(reinterpret_cast<const float4 *>(s.materials) + k); } }
    if (light_sh) { for (int k = threadIdx.x; k < s.n_lights * 4; k += kShadeThreads) { sm_light[k] = __ldg(reinterpret_cast<const float4 *>(s.lights) + k); } }
    __syncthreads();
    const float4 *matp = mat_sh ? sm_mat : reinterpret_cast<const float4 *>(s.materials);
    const float4 *lightp = light_sh ? sm_light : reinterpret_cast<const float4 *>(s.lights);
    const int b = level & 1, nb = b ^ 1;
    const uint32_t n = min(Q.counts[level], Q.cap_rays);
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    RayCounters rc = {0u, 0u, 0u};
    int overflow = 0;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += stride) {   // the same trip count for every thread of the CTA
        const uint32_t j = base + threadIdx.x;
        const uint32_t i = (Q.perm != nullptr && j < n) ? Q.perm[j] : j;   // MRT_FLAG_SORT_BY_MATERIAL: hits binned by material
        // ---- phase 1: the surface, and how many shadow rays / rays / chains this hit emits
        uint32_t ns = 0u, nr = 0u, nc = 0u;
        bool shading = false, first_lit = false, go = false, spec = false, trans = false;
        Surface sf;
        LightSample ls0;
        v3 d = mk3(0.0f, 0.0f, 0.0f), T = mk3(1.0f, 1.0f, 1.0f);
        uint32_t q = 0u, branch = 1u, chain = i, eb = 0u;
        int depth = 1;
        float r_h0 = 0.0f, r_h1 = 0.0f;
        if (j < n) {
            // the records of this thread's next iteration start their trip from HBM to L2 now
            if (Q.perm == nullptr && i + stride < n) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(Q.hit + i + stride));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(Q.r0[b] + i + stride));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(Q.r1[b] + i + stride));
                if (!L0) {
                    asm volatile("prefetch.global.L2 [%0];" :: "l"(Q.r2[b] + i + stride));
                    asm volatile("prefetch.global.L2 [%0];" :: "l"(Q.r3[b] + i + stride));
                }
            }
            // hit and ray records are independent loads: issue all of them before the first use
            const float4 hq = Q.hit[i];
            const float4 a = Q.r0[b][i], c1 = Q.r1[b][i];
            float4 c2 = make_float4(1.0f, 1.0f, 1.0f, 0.0f), c3 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (!L0) { c2 = Q.r2[b][i]; c3 = Q.r3[b][i]; }
            q = __float_as_uint(L0 ? c1.z : c3.x); branch = L0 ? 1u : __float_as_uint(c3.y);
            depth = L0 ? 1 : __float_as_int(c1.w);
            chain = L0 ? i : __float_as_uint(c2.w);
            eb = branch << 6;
            // the vertex's russian-roulette and hemisphere draws (table gathers, when they are gathers, go out before the hit is looked at)
            const float r_rr = pt_draw(s.static_halton, s.static_halton_is_perm, q, eb, DS_RR, rp.seed);
            r_h0 = pt_draw(s.shader_table, s.shader_table_is_perm, q, eb, DS_HEMI, rp.seed);
            r_h1 = pt_draw(s.shader_table, s.shader_table_is_perm, q, eb + 1u, DS_HEMI, rp.seed);
            const uint32_t packed = __float_as_uint(hq.w);
            HitRec h;
            h.t = hq.x; h.u = hq.y; h.v = hq.z; h.kind = static_cast<int>(packed >> 28); h.idx = static_cast<int>(packed & 0x0fffffffu);
            const int32_t root_parent = L0 ? -1 - __float_as_int(c1.w) : 0;   // camera path: -1 - pixel
            PathAcc *ch = Q.acc + chain;
            if (h.kind == HIT_NONE) {                                   // Shader.cpp:122: no intersection, rgb stays 0
                if (L0) { acc_init(ch, mk3(0.0f, 0.0f, 0.0f), 0u, root_parent); }
            } else {
                const v3 o = mk3(a.x, a.y, a.z);
                d = mk3(a.w, c1.x, c1.y); T = mk3(c2.x, c2.y, c2.z);
                resolve_surface(s, o, d, h, sf, matp, lightp);
                if (has_positive(sf.Le)) {                              // PathTracer.cpp:28-33: stop on a light source
                    const v3 e = T * sf.Le;
                    if (L0) { acc_init(ch, e, kAccEnded, root_parent); }
                    else {
                        // the emitter term only counts while no vertex of the chain had direct light (PathTracer.cpp:107-110)
                        const uint32_t info = ch->info;
                        if ((info & kAccFirstMask) == 0u) { ch->pre[0] += e.x; ch->pre[1] += e.y; ch->pre[2] += e.z; }
                        ch->info = info | kAccEnded;
                    }
                } else {
                    if (L0) { acc_init(ch, mk3(0.0f, 0.0f, 0.0f), 0u, root_parent); }
                    shading = true;
                    const bool diffuse = has_positive(sf.Kd);
                    if (diffuse && s.n_lights > 0) {                    // PathTracer.cpp:55-87: direct light, one shadow ray per light sample
                        first_lit = pt_light_sample(s, rp, lightp, sf, q, eb, 0u, ls0);
                        ns = first_lit ? 1u : 0u;
                        for (int k = 1; k < rp.spl; ++k) {
                            LightSample ls;
                            if (pt_light_sample(s, rp, lightp, sf, q, eb, static_cast<uint32_t>(k), ls)) { ++ns; }
                        }
                        rc.shadow += ns;
                    }
                    const bool traceable = depth + 1 <= kRayDepthMax;   // deeper rays are constructed (counted) but add nothing
                    go = diffuse && (depth <= kRayDepthMin || r_rr > 0.5f);   // PathTracer.cpp:89-91: russian roulette
                    spec = has_positive(sf.Ks); trans = has_positive(sf.Kt);   // PathTracer.cpp:113-134
                    rc.secondary += (go ? 1u : 0u) + (spec ? 1u : 0u) + (trans ? 1u : 0u);
                    if (traceable) { nc = (spec ? 1u : 0u) + (trans ? 1u : 0u); nr = nc + (go ? 1u : 0u); }
                    else { go = false; spec = false; trans = false; }
                }
            }
        }
        // ---- slots: prefix sums over the CTA, one atomicAdd per queue
        unsigned long long mine = static_cast<unsigned long long>(ns) | (static_cast<unsigned long long>(nr) << 24) | (static_cast<unsigned long long>(nc) << 44);
        unsigned long long incl = mine;
#pragma unroll
        for (int k = 1; k < 32; k <<= 1) {
            const unsigned long long v = __shfl_up_sync(0xffffffffu, incl, k);
            if (lane >= static_cast<uint32_t>(k)) { incl += v; }
        }
        if (lane == 31u) { sm_warp[warp] = incl; }
        __syncthreads();
        unsigned long long before = 0ull, total = 0ull;
#pragma unroll
        for (uint32_t wq = 0; wq < kShadeThreads / 32; ++wq) {
            const unsigned long long v = sm_warp[wq];
            total += v;
            if (wq < warp) { before += v; }
        }
        if (threadIdx.x < 3u) {
            const uint32_t cnt = threadIdx.x == 0u ? static_cast<uint32_t>(total & 0xffffffu)
                               : threadIdx.x == 1u ? static_cast<uint32_t>((total >> 24) & 0xfffffu) : static_cast<uint32_t>(total >> 44);
            uint32_t *ctr = &Q.counts[threadIdx.x == 0u ? 8 : (threadIdx.x == 1u ? level + 1 : 9)];
            sm_base[threadIdx.x] = cnt ? atomicAdd(ctr, cnt) : 0u;
        }
        __syncthreads();
        const unsigned long long excl = before + incl - mine;
        uint32_t sslot = sm_base[0] + static_cast<uint32_t>(excl & 0xffffffu);
        uint32_t rslot = sm_base[1] + static_cast<uint32_t>((excl >> 24) & 0xfffffu);
        uint32_t cslot = sm_base[2] + static_cast<uint32_t>(excl >> 44);
        if (!shading) { continue; }
        // ---- phase 2: write the records
        const uint32_t skip = pack_skip(sf.skipKind, sf.skipIdx);
        if (ns != 0u) {
            for (int k = 0; k < rp.spl; ++k) {
                LightSample ls = ls0;
                const bool lit = k == 0 ? first_lit : pt_light_sample(s, rp, lightp, sf, q, eb, static_cast<uint32_t>(k), ls);
                if (!lit) { continue; }
                const v3 cc = T * ls.ld;
                if (sslot < Q.cap_shadow) {
                    Q.s0[sslot] = make_float4(sf.point.x, sf.point.y, sf.point.z, ls.dir.x);
                    Q.s1[sslot] = make_float4(ls.dir.y, ls.dir.z, ls.dist, __uint_as_float(skip));
                    const uint32_t tag = chain | (static_cast<uint32_t>(depth - 1) << 28) | (has_positive(ls.ld) ? 0x80000000u : 0u);
                    Q.s2[sslot] = make_float4(cc.x, cc.y, cc.z, __uint_as_float(tag));
                } else { overflow = 1; }
                ++sslot;
            }
        }
        if (go) {                                                   // PathTracer.cpp:89-111: indirect diffuse
            const v3 nd = cosine_sample_hemisphere(sf.normal, r_h0, r_h1);
            v3 Tn = T * sf.Kd;
            if (depth > kRayDepthMin) { Tn = Tn / (0.5f * 0.5f); }
            if (rslot < Q.cap_rays) {
                Q.r0[nb][rslot] = make_float4(sf.point.x, sf.point.y, sf.point.z, nd.x);
                Q.r1[nb][rslot] = make_float4(nd.y, nd.z, __uint_as_float(skip), __int_as_float(depth + 1));
                Q.r2[nb][rslot] = make_float4(Tn.x, Tn.y, Tn.z, __uint_as_float(chain));
                Q.r3[nb][rslot] = make_float4(__uint_as_float(q), __uint_as_float(branch * 4u + 1u), 0.0f, 0.0f);
            } else { overflow = 1; }
            ++rslot;
        }
#pragma unroll
        for (int kind = 0; kind < 2; ++kind) {                      // PathTracer.cpp:113-134: specular reflection, then transmission
            if (!(kind == 0 ? spec : trans)) { continue; }
            const v3 K = kind == 0 ? sf.Ks : sf.Kt;
            const v3 nd = kind == 0 ? reflect(d, sf.normal) : refract(d, sf.normal, 1.0f / sf.ior);
            // the accumulator is initialised whenever the pool slot exists, so that finalize never reads a stale record
            if (cslot < Q.cap_chains) { acc_init(Q.acc + cslot, mk3(0.0f, 0.0f, 0.0f), static_cast<uint32_t>(depth) << 4, static_cast<int32_t>(chain)); }
            if (cslot < Q.cap_chains && rslot < Q.cap_rays) {
                const v3 Tn = T * K;
                Q.r0[nb][rslot] = make_float4(sf.point.x, sf.point.y, sf.point.z, nd.x);
                Q.r1[nb][rslot] = make_float4(nd.y, nd.z, __uint_as_float(skip), __int_as_float(depth + 1));
                Q.r2[nb][rslot] = make_float4(Tn.x, Tn.y, Tn.z, __uint_as_float(cslot));
                Q.r3[nb][rslot] = make_float4(__uint_as_float(q), __uint_as_float(branch * 4u + 2u + static_cast<uint32_t>(kind)), 0.0f, 0.0f);
            } else { overflow = 1; }
            ++cslot; ++rslot;
        }
    }
    const uint32_t sh = __reduce_add_sync(0xffffffffu, rc.shadow), se = __reduce_add_sync(0xffffffffu, rc.secondary);
    if ((threadIdx.x & 31) == 0) {
        if (sh) { atomicAdd(&dc->rays_shadow, static_cast<unsigned long long>(sh)); }
        if (se) { atomicAdd(&dc->rays_secondary, static_cast<unsigned long long>(se)); }
    }
    if (overflow) { atomicOr(&dc->overflow, 1); }   // a queue was full: the frame driver repeats the pass with fewer samples
}

// bookkeeping after shade(level) and after shadow(level)
__global__ void k_wf_after_shade(Queues Q, int level, DeviceCounters *dc) {
    dc->extend_traced += min(Q.counts[level], Q.cap_rays);
    Q.counts[kCounterExtend] = 0u;            // work counter of the persistent extend kernel
    Q.counts[10 + level] = min(Q.counts[9], Q.cap_chains);   // chain pool size after this level
}

// ---- fold: the light of every unoccluded shadow ray goes into its chain (PathTracer.cpp:72-82) ---------------------
// A streaming pass over the shadow queue in queue order (the shade kernel filled it in roughly chain order, so the
// accumulator sectors are touched nearly sequentially). The shadow kernel used to do this itself with three float atomics
// and an atomicOr per unoccluded ray at the moment the ray retired, i.e. in random order. All rays of one chain in one
// level belong to the same depth, so "no j* yet, or j* == this depth" decides pre / post without ordering between them.
__global__ void k_wf_fold(Queues Q) {
    const uint32_t n = min(Q.counts[8], Q.cap_shadow);
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        if (Q.vis[j] == 0) { continue; }
        const float4 cc = Q.s2[j];
        const uint32_t tag = __float_as_uint(cc.w);
        PathAcc *a = Q.acc + (tag & 0x0fffffffu);
        const uint32_t depth = ((tag >> 28) & 7u) + 1u;
        const uint32_t first = a->info & kAccFirstMask;
        float *dst = (first == 0u || first == depth) ? a->pre : a->post;
        if (cc.x != 0.0f) { atomicAdd(dst + 0, cc.x); }
        if (cc.y != 0.0f) { atomicAdd(dst + 1, cc.y); }
        if (cc.z != 0.0f) { atomicAdd(dst + 2, cc.z); }
        if ((tag & 0x80000000u) && first == 0u) { atomicOr(&a->info, depth); }
    }
}
__global__ void k_wf_after_fold(Queues Q, DeviceCounters *dc) {
    dc->shadow_traced += min(Q.counts[8], Q.cap_shadow);
    Q.counts[kCounterShadow] = 0u;
    Q.counts[8] = 0u;                         // shadow queue consumed
}

MRT_D v3 acc_total(const float4 p0, const float4 p1) {
    v3 t = mk3(p0.x, p0.y, p0.z);
    if (!(__float_as_uint(p0.w) & kAccEnded)) { t = t + mk3(p1.x, p1.y, p1.z); }
    return t;
}

// child chains created while shading `level` (pool range [counts[10+level-1], counts[10+level])) fold into their parents:
// a child spawned at depth pd counts like the parent's own light of that depth (see the header comment)
__global__ void k_wf_finalize_level(Queues Q, int level) {
    const uint32_t begin = min(level == 0 ? Q.counts[0] : Q.counts[10 + level - 1], Q.cap_chains), end = min(Q.counts[10 + level], Q.cap_chains);
    for (uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x; i < end; i += gridDim.x * blockDim.x) {
        const float4 *cp = reinterpret_cast<const float4 *>(Q.acc + i);
        const float4 p0 = cp[0], p1 = cp[1];
        const v3 t = acc_total(p0, p1);
        const int32_t parent = __float_as_int(p1.w);
        const uint32_t pd = (__float_as_uint(p0.w) >> 4) & 7u;
        if (parent < 0 || static_cast<uint32_t>(parent) >= Q.cap_chains || pd == 0u) { continue; }
        PathAcc *p = Q.acc + parent;
        const uint32_t first = p->info & kAccFirstMask;
        float *dst = (first == 0u || pd <= first) ? p->pre : p->post;
        if (t.x != 0.0f) { atomicAdd(dst + 0, t.x); }
        if (t.y != 0.0f) { atomicAdd(dst + 1, t.y); }
        if (t.z != 0.0f) { atomicAdd(dst + 2, t.z); }
    }
}

// camera paths -> pixels (Renderer.cpp:149-157). One thread owns one path slot of the tile walk and applies the batch's
// samples in order, so the 8-bit running average stays exact while the pixel is read and written once per pass.
// Requires that no two slots of a sample map to the same pixel (tiles_disjoint, mrt_api.cu).
__global__ void k_wf_resolve(Queues Q, uint32_t roots_per_sample, int n_samples, BatchInts avg_n, FrameTargets tg) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= roots_per_sample) { return; }
    const int32_t pix = -1 - Q.acc[r].parent;
    int32_t bm = tg.bitmap != nullptr ? tg.bitmap[pix] : 0;
    float4 ac = tg.accum != nullptr ? tg.accum[pix] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll 4
    for (int b = 0; b < n_samples; ++b) {
        const float4 *cp = reinterpret_cast<const float4 *>(Q.acc + static_cast<size_t>(b) * roots_per_sample + r);
        const v3 rgb = acc_total(cp[0], cp[1]);
        bm = incremental_avg(rgb, bm, avg_n.v[b]);
        ac.x += static_cast<float>(f2u_x86(rgb.x * 255.0f)); ac.y += static_cast<float>(f2u_x86(rgb.y * 255.0f));
        ac.z += static_cast<float>(f2u_x86(rgb.z * 255.0f)); ac.w += 1.0f;
    }
    if (tg.bitmap != nullptr) { tg.bitmap[pix] = bm; }
    if (tg.accum != nullptr) { tg.accum[pix] = ac; }
}

// Image sizes that are not multiples of 16 make the reference's tiles wrap and overlap (Renderer.cpp:125-133), so several
// path slots of one sample can hit the same pixel; the reference applies them one after the other in tile order. One CTA
// walks the samples and tiles in that order (slots of one tile never share a pixel). Slow, and only used for such sizes;
// both front-ends round the size to a multiple of 16 first.
__global__ void __launch_bounds__(1024) k_wf_resolve_serial(Queues Q, const uint32_t *__restrict__ tile_base, int n_tiles, uint32_t roots_per_sample,
                                                            int n_samples, BatchInts avg_n, FrameTargets tg) {
    for (int b = 0; b < n_samples; ++b) {
        for (int k = 0; k < n_tiles; ++k) {
            const uint32_t lo = tile_base[k], hi = tile_base[k + 1];
            for (uint32_t r = lo + threadIdx.x; r < hi; r += blockDim.x) {
                const float4 *cp = reinterpret_cast<const float4 *>(Q.acc + static_cast<size_t>(b) * roots_per_sample + r);
                const float4 p0 = cp[0], p1 = cp[1];
                const v3 rgb = acc_total(p0, p1);
                const int32_t pix = -1 - __float_as_int(p1.w);
                if (tg.bitmap != nullptr) { tg.bitmap[pix] = incremental_avg(rgb, tg.bitmap[pix], avg_n.v[b]); }
                if (tg.accum != nullptr) {
                    float4 ac = tg.accum[pix];
                    ac.x += static_cast<float>(f2u_x86(rgb.x * 255.0f)); ac.y += static_cast<float>(f2u_x86(rgb.y * 255.0f));
                    ac.z += static_cast<float>(f2u_x86(rgb.z * 255.0f)); ac.w += 1.0f;
                    tg.accum[pix] = ac;
                }
            }
            __syncthreads();
        }
    }
}

// ---- host side -----------------------------------------------------------------------------------------------------
namespace {
template <typename T> cudaError_t wf_alloc(Wavefront *w, T **p, size_t count) {
    void *v = nullptr;
    cudaError_t e = cudaMalloc(&v, (count ? count : 1) * sizeof(T));
    if (e == cudaSuccess) { w->allocs.push_back(v); *p = static_cast<T *>(v); }
    return e;
}
enum WfClass { WF_OTHER = 0, WF_EXTEND = 1, WF_SHADE = 2, WF_SHADOW = 3 };
}  // namespace

#define WF_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { return e_; } } while (0)

int wavefront_batch(const Wavefront *w) { return w->batch; }
uint64_t wavefront_capacity(const Wavefront *w) { return w->q.cap_rays < w->q.cap_shadow ? w->q.cap_rays : w->q.cap_shadow; }
size_t wavefront_bytes(const Wavefront *w) { return w->bytes; }

// Queue capacities for `paths` camera paths per pass. A hit can emit three continuation rays (diffuse, specular,
// transmission) and samples_light shadow rays, so no fixed factor is safe for every material set: the frame driver
// (mrt_api.cu) retries a pass that overflowed with half the samples and, for a single sample, with grown queues.
static cudaError_t wf_alloc_queues(Wavefront *w, size_t paths, int factor) {
    Queues &q = w->q;
    const size_t lim = 0x0ffffff0u;
    const size_t spl = static_cast<size_t>(w->spl < 1 ? 1 : w->spl);
    q.cap_rays = static_cast<uint32_t>(std::min<size_t>(static_cast<size_t>(factor) * paths + 1024, lim));
    q.cap_shadow = static_cast<uint32_t>(std::min<size_t>(static_cast<size_t>(q.cap_rays) * spl, lim));
    q.cap_chains = static_cast<uint32_t>(std::min<size_t>((static_cast<size_t>(factor) + 1) * paths + 1024, lim));
    w->bytes = 0;
    for (int b = 0; b < 2; ++b) {
        WF_TRY(wf_alloc(w, &q.r0[b], q.cap_rays)); WF_TRY(wf_alloc(w, &q.r1[b], q.cap_rays));
        WF_TRY(wf_alloc(w, &q.r2[b], q.cap_rays)); WF_TRY(wf_alloc(w, &q.r3[b], q.cap_rays));
    }
    WF_TRY(wf_alloc(w, &q.hit, q.cap_rays));
    WF_TRY(wf_alloc(w, &q.s0, q.cap_shadow)); WF_TRY(wf_alloc(w, &q.s1, q.cap_shadow)); WF_TRY(wf_alloc(w, &q.s2, q.cap_shadow));
    WF_TRY(wf_alloc(w, &q.vis, static_cast<size_t>(q.cap_shadow) + 64));
    WF_TRY(wf_alloc(w, &q.acc, q.cap_chains));
    WF_TRY(wf_alloc(w, &q.counts, kWfCounts));
    q.perm = nullptr; q.sort_bins = nullptr;
    if (w->sort_mode != 0) {
        WF_TRY(wf_alloc(w, &q.perm, q.cap_rays));
        WF_TRY(wf_alloc(w, &q.sort_bins, kSortBins + 1));
        q.sort_key = w->sort_mode == 2 ? 1 : 0;
        q.sort_shift = w->sort_shift;
    }
    w->bytes = static_cast<size_t>(q.cap_rays) * (8 * 16 + 16) + static_cast<size_t>(q.cap_shadow) * (3 * 16 + 1) + static_cast<size_t>(q.cap_chains) * sizeof(PathAcc);
    w->factor = factor;
    return cudaSuccess;
}
static void wf_free_queues(Wavefront *w) {
    for (void *p : w->allocs) { cudaFree(p); }
    w->allocs.clear();
}

cudaError_t wavefront_create(size_t paths_per_sample, int batch, int spl, int sort_mode, uint32_t n_primitives, Wavefront **out) {
    Wavefront *w = new Wavefront;
    *out = w;
    w->paths_per_sample = paths_per_sample; w->spl = spl;
    w->sort_mode = sort_mode;
    while ((n_primitives >> w->sort_shift) > kSortBins - 9u) { ++w->sort_shift; }
    w->batch = batch < 1 ? 1 : (batch > kWfMaxBatch ? kWfMaxBatch : batch);
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&w->sm_count, cudaDevAttrMultiProcessorCount, dev);
    const char *ov = std::getenv("MRT_B200_OVERLAP");
    // on by default while nodes and triangles fit the 126 MB L2 (80 B per 10 primitives + 48 B per primitive): with the 550 MB scene of
    // config 5 the extend kernel needs its eighth CTA per SM to hide HBM latency and the frame gets 1 % slower instead (197.8 vs 195.9 ms)
    w->overlap = ov ? std::atoi(ov) : (n_primitives <= 2000000u ? 1 : 0);   // MRT_B200_OVERLAP=0 / 1: A/B
    if (w->overlap) {
        WF_TRY(cudaStreamCreateWithFlags(&w->side, cudaStreamNonBlocking));
        WF_TRY(cudaEventCreateWithFlags(&w->ev_fork, cudaEventDisableTiming));
        WF_TRY(cudaEventCreateWithFlags(&w->ev_join, cudaEventDisableTiming));
    }
    return wf_alloc_queues(w, paths_per_sample * static_cast<size_t>(w->batch), 2);
}

// doubles the queues (after a single-sample pass overflowed); false when the limit is reached
cudaError_t wavefront_grow(Wavefront *w, bool *grown) {
    *grown = false;
    const size_t paths = w->paths_per_sample * static_cast<size_t>(w->batch);
    const int factor = w->factor * 2;
    const double need = static_cast<double>(w->bytes) * 2.0;
    if (w->q.cap_rays >= 0x0ffffff0u || need > 64.0e9) { return cudaSuccess; }
    wf_free_queues(w);
    cudaError_t e = wf_alloc_queues(w, paths, factor);
    if (e == cudaSuccess) { *grown = true; }
    return e;
}

void wavefront_destroy(Wavefront *w) {
    if (!w) { return; }
    for (cudaEvent_t e : w->pool) { cudaEventDestroy(e); }
    if (w->side) { cudaStreamSynchronize(w->side); cudaStreamDestroy(w->side); }
    if (w->ev_fork) { cudaEventDestroy(w->ev_fork); }
    if (w->ev_join) { cudaEventDestroy(w->ev_join); }
    wf_free_queues(w);
    delete w;
}

static cudaError_t wf_mark(Wavefront *hst, int cls, cudaStream_t st) {
    if (hst->used == hst->pool.size()) {
        if (hst->pool.size() >= 16384) { return cudaSuccess; }   // stop timing very long frames, keep rendering
        cudaEvent_t e;
        WF_TRY(cudaEventCreate(&e));
        hst->pool.push_back(e);
    }
    cudaEvent_t e = hst->pool[hst->used++];
    hst->marks.push_back(WfMark{e, cls});
    return cudaEventRecord(e, st);
}

// to be called after the stream has been synchronised: adds the device time of every launch since the last call
cudaError_t wavefront_collect_times(Wavefront *w, WavefrontTimes *t) {
    for (size_t i = 1; i < w->marks.size(); ++i) {
        if (w->marks[i].cls < 0) { continue; }   // a "begin" mark: nothing ran between it and its predecessor
        float ms = 0.0f;
        WF_TRY(cudaEventElapsedTime(&ms, w->marks[i - 1].ev, w->marks[i].ev));
        switch (w->marks[i].cls) {
            case WF_EXTEND: t->extend_ms += ms; t->extend_launches += 1; break;
            case WF_SHADE: t->shade_ms += ms; break;
            case WF_SHADOW: t->shadow_ms += ms; break;
            default: t->other_ms += ms; break;
        }
    }
    w->marks.clear();
    w->used = 0;
    return cudaSuccess;
}

cudaError_t wavefront_trace_batch(Wavefront *w, const SceneDev &s, const CameraRec &cam, const RenderParams &rp, const int2 *d_tile_origin,
                                  const uint32_t *d_tile_base, uint32_t roots_per_sample, int n_samples, const int *samples, DeviceCounters *dc,
                                  uint64_t *launches, cudaStream_t st) {
    Queues &Q = w->q;
    if (n_samples < 1 || n_samples > w->batch) { return cudaErrorInvalidValue; }
    const int grid = w->sm_count * MRT_TRAV_MIN_BLOCKS;
    WF_TRY(wf_mark(w, -1, st));
    WF_TRY(cudaMemsetAsync(Q.counts, 0, kWfCounts * sizeof(uint32_t), st));
    const int patches = rp.n_tiles * ((rp.bsx + 7) / 8) * ((rp.bsy + 3) / 4);
    if (patches == 0) { return cudaSuccess; }
    if (static_cast<unsigned long long>(roots_per_sample) * static_cast<unsigned long long>(n_samples) > Q.cap_rays) { return cudaErrorInvalidValue; }
    BatchInts sl {};
    for (int i = 0; i < n_samples; ++i) { sl.v[i] = samples[i]; }
    const dim3 ggrid(static_cast<unsigned>((patches + kWfThreads / 32 - 1) / (kWfThreads / 32)), static_cast<unsigned>(n_samples));
    k_wf_generate<<<ggrid, kWfThreads, 0, st>>>(Q, s, cam, rp, d_tile_origin, d_tile_base, roots_per_sample, sl);
    k_wf_begin_levels<<<1, 1, 0, st>>>(Q, static_cast<uint32_t>(n_samples), roots_per_sample, dc);
    WF_TRY(wf_mark(w, WF_OTHER, st));
    *launches += 2;
    const int sgrid = w->sm_count * MRT_SHADE_MIN_BLOCKS;
    // overlap: k_wf_fold(level) streams the shadow queue at half the HBM rate and nothing needs its result before
    // k_wf_shade(level + 1), while k_wf_extend(level + 1) is bound by instruction issue. So the extend kernel of the next level is
    // launched with one CTA per SM fewer right after shadow(level), and fold(level) runs next to it on a second stream in the
    // registers and thread slots that leaves free (256 threads x 16 registers per CTA, one CTA per SM). The extend kernel gets 4 %
    // slower, the fold kernels disappear from the critical path: 180.9 -> 178.0 ms per 64-spp frame, 23.25 -> 22.92 ms per 8-spp
    // frame (profiles/r2_ab_fold_overlap.txt). The per-kernel times the library reports are event intervals: the fold interval now
    // lies inside the extend interval of the next level.
    const bool overlap = w->overlap != 0 && !s.grid;
    const int xgrid = (overlap && w->overlap == 1) ? w->sm_count * (MRT_TRAV_MIN_BLOCKS - 1) : grid;   // overlap == 2 (A/B): full grid, fold only fills the tails (slower)
    static const int fold_mult = std::getenv("MRT_B200_FOLD_GRID") ? std::atoi(std::getenv("MRT_B200_FOLD_GRID")) : 1;   // CTAs per SM of the overlapped fold (A/B: 1 / 2 / 4 = 177.98 / 178.30 / 178.33 ms)
    bool extend_launched = false, joined = true;
    for (int level = 0; level < kLevels; ++level) {
        if (!extend_launched) {
            if (s.grid) { k_wf_extend_grid<<<grid, 256, 0, st>>>(s, Q, level); } else { k_wf_extend<<<grid, kWfThreads, 0, st>>>(s, Q, level, dc); }
            WF_TRY(wf_mark(w, WF_EXTEND, st));
        }
        extend_launched = false;
        if (!joined) { WF_TRY(cudaStreamWaitEvent(st, w->ev_join, 0)); joined = true; }   // fold(level - 1) has updated the accumulators
        if (Q.perm != nullptr) {
            WF_TRY(cudaMemsetAsync(Q.sort_bins, 0, (kSortBins + 1) * sizeof(uint32_t), st));
            const unsigned sblocks = static_cast<unsigned>((static_cast<size_t>(n_samples) * roots_per_sample * static_cast<size_t>(w->factor) + kSortChunk - 1) / kSortChunk);
            k_wf_sort_count<<<sblocks, 256, 0, st>>>(s, Q, level);
            k_wf_sort_scan<<<1, kSortBins, 0, st>>>(Q);
            k_wf_sort_scatter<<<sblocks, 256, 0, st>>>(s, Q, level);
            *launches += 3;
        }
        if (level == 0) { k_wf_shade<true><<<sgrid, kShadeThreads, 0, st>>>(s, rp, Q, level, dc); }
        else { k_wf_shade<false><<<sgrid, kShadeThreads, 0, st>>>(s, rp, Q, level, dc); }
        WF_TRY(wf_mark(w, WF_SHADE, st));
        k_wf_after_shade<<<1, 1, 0, st>>>(Q, level, dc);
        if (s.grid) { k_wf_shadow_grid<<<grid, 256, 0, st>>>(s, Q); } else { k_wf_shadow<<<grid, kWfThreads, 0, st>>>(s, Q, dc); }
        WF_TRY(wf_mark(w, WF_SHADOW, st));
        if (overlap && level + 1 < kLevels) {
            WF_TRY(cudaEventRecord(w->ev_fork, st));
            k_wf_extend<<<xgrid, kWfThreads, 0, st>>>(s, Q, level + 1, dc);                 // first, so that its CTAs are placed first
            WF_TRY(wf_mark(w, WF_EXTEND, st));
            extend_launched = true;
            WF_TRY(cudaStreamWaitEvent(w->side, w->ev_fork, 0));
            WF_TRY(wf_mark(w, -1, w->side));
            k_wf_fold<<<w->sm_count * fold_mult, 256, 0, w->side>>>(Q);
            k_wf_after_fold<<<1, 1, 0, w->side>>>(Q, dc);
            WF_TRY(wf_mark(w, WF_OTHER, w->side));
            WF_TRY(cudaEventRecord(w->ev_join, w->side));
            WF_TRY(wf_mark(w, -1, st));   // the next mark on the frame stream starts a new interval
            joined = false;
        } else {
            k_wf_fold<<<w->sm_count * 8, 256, 0, st>>>(Q);
            k_wf_after_fold<<<1, 1, 0, st>>>(Q, dc);
            WF_TRY(wf_mark(w, WF_OTHER, st));
        }
        *launches += 6;
    }
    for (int level = kLevels - 2; level >= 0; --level) {   // no chain is created while shading the last level
        k_wf_finalize_level<<<w->sm_count * 2, 256, 0, st>>>(Q, level);
        *launches += 1;
    }
    WF_TRY(wf_mark(w, WF_OTHER, st));
    return cudaGetLastError();
}

cudaError_t wavefront_resolve_batch(Wavefront *w, const uint32_t *d_tile_base, int n_tiles, uint32_t roots_per_sample, bool tiles_disjoint, int n_samples,
                                    const int *avg_n, FrameTargets tg, uint64_t *launches, cudaStream_t st) {
    if (n_samples < 1 || n_samples > w->batch || roots_per_sample == 0) { return n_samples < 1 || n_samples > w->batch ? cudaErrorInvalidValue : cudaSuccess; }
    BatchInts an {};
    for (int i = 0; i < n_samples; ++i) { an.v[i] = avg_n[i]; }
    WF_TRY(wf_mark(w, -1, st));
    if (tiles_disjoint) { k_wf_resolve<<<(roots_per_sample + 255) / 256, 256, 0, st>>>(w->q, roots_per_sample, n_samples, an, tg); }
    else { k_wf_resolve_serial<<<1, 1024, 0, st>>>(w->q, d_tile_base, n_tiles, roots_per_sample, n_samples, an, tg); }
    *launches += 1;
    WF_TRY(wf_mark(w, WF_OTHER, st));
    return cudaGetLastError();
}

}  // namespace mrt

#ifdef MRT_TRAVERSAL_STATS
// diagnostic library only: node visits / triangle tests / rays of the wavefront kernels since the last reset
extern "C" int mrt_debug_traversal_stats(unsigned long long out[5], int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(&out[0], mrt::g_stat_nodes, 8); cudaMemcpyFromSymbol(&out[1], mrt::g_stat_tris, 8); cudaMemcpyFromSymbol(&out[2], mrt::g_stat_rays, 8);
    cudaMemcpyFromSymbol(&out[3], mrt::g_stat_fruitless, 8); cudaMemcpyFromSymbol(&out[4], mrt::g_stat_stale, 8);
    if (reset) {
        const unsigned long long z = 0;
        cudaMemcpyToSymbol(mrt::g_stat_nodes, &z, 8); cudaMemcpyToSymbol(mrt::g_stat_tris, &z, 8); cudaMemcpyToSymbol(mrt::g_stat_rays, &z, 8);
        cudaMemcpyToSymbol(mrt::g_stat_fruitless, &z, 8); cudaMemcpyToSymbol(mrt::g_stat_stale, &z, 8);
    }
    return 0;
}
#endif

// ---- per-ray ground truth through the wavefront kernels (mrt_trace_rays on a PathTracer handle) ------------------------
// The parity tests compare these results with the reference ray by ray, so that the persistent extend / shadow kernels
// themselves (not only the per-thread traversal of the megakernel) are pinned bit for bit. The closest-hit rays go in as
// level 2 of the ray queue (same buffer as level 0, but with the skip primitive where a camera ray carries its q).
namespace mrt {

constexpr int kTestLevel = 2;

__global__ void k_wf_test_fill(Queues Q, int64_t n, const float *__restrict__ o, const float *__restrict__ d, const int32_t *__restrict__ okind,
                               const int32_t *__restrict__ oidx, const float *__restrict__ sdist) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    const int sk = okind ? okind[i] : HIT_NONE, si = oidx ? oidx[i] : -1;
    const uint32_t skip = pack_skip(sk, si);
    if (sdist != nullptr && sdist[i] >= 0.0f) {
        const uint32_t slot = atomicAdd(&Q.counts[8], 1u);
        Q.s0[slot] = make_float4(o[3 * i], o[3 * i + 1], o[3 * i + 2], d[3 * i]);
        Q.s1[slot] = make_float4(d[3 * i + 1], d[3 * i + 2], sdist[i], __uint_as_float(skip));
        Q.s2[slot] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(static_cast<uint32_t>(i)));   // where the answer goes
    } else {
        const uint32_t slot = atomicAdd(&Q.counts[kTestLevel], 1u);
        Q.r0[0][slot] = make_float4(o[3 * i], o[3 * i + 1], o[3 * i + 2], d[3 * i]);
        Q.r1[0][slot] = make_float4(d[3 * i + 1], d[3 * i + 2], __uint_as_float(skip), __int_as_float(1));
        Q.r3[0][slot] = make_float4(__uint_as_float(static_cast<uint32_t>(i)), 0.0f, 0.0f, 0.0f);
    }
}

__global__ void k_wf_test_gather(SceneDev s, Queues Q, float *out_t, int32_t *out_kind, int32_t *out_idx, int32_t *out_mat, float *out_normal) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < Q.counts[kTestLevel]) {
        const float4 a = Q.r0[0][i], c1 = Q.r1[0][i], hq = Q.hit[i];
        const uint32_t dst = __float_as_uint(Q.r3[0][i].x), packed = __float_as_uint(hq.w);
        const v3 ro = mk3(a.x, a.y, a.z), rd = mk3(a.w, c1.x, c1.y);
        HitRec h;
        h.t = hq.x; h.u = hq.y; h.v = hq.z; h.kind = static_cast<int>(packed >> 28); h.idx = h.kind == HIT_NONE ? -1 : static_cast<int>(packed & 0x0fffffffu);
        int idx = h.idx, mat = -1;
        v3 nrm = mk3(0.0f, 1.0f, 0.0f);
        if (h.kind != HIT_NONE) {
            Surface sf;
            resolve_surface(s, ro, rd, h, sf);
            nrm = sf.normal;
            if (h.kind == HIT_PLANE) { idx = s.planes[h.idx].id; mat = s.planes[h.idx].material; }
            else if (h.kind == HIT_SPHERE) { idx = s.spheres[h.idx].id; mat = s.spheres[h.idx].material; }
            else if (h.kind == HIT_TRIANGLE) { mat = __float_as_int(reinterpret_cast<const float4 *>(s.tri_shade + h.idx)[3].w); }
        }
        out_t[dst] = h.t; out_kind[dst] = h.kind; out_idx[dst] = idx; out_mat[dst] = mat;
        if (out_normal) { out_normal[3 * dst] = nrm.x; out_normal[3 * dst + 1] = nrm.y; out_normal[3 * dst + 2] = nrm.z; }
    }
    if (i < Q.counts[8]) {   // the shadow rays of this test: the visibility byte the shadow kernel wrote
        const uint32_t dst = __float_as_uint(Q.s2[i].w);
        const bool occluded = Q.vis[i] == 0;
        out_t[dst] = occluded ? 0.0f : Q.s1[i].z;
        out_kind[dst] = occluded ? 1 : 0; out_idx[dst] = -1; out_mat[dst] = -1;
    }
}

cudaError_t wavefront_trace_rays(Wavefront *w, const SceneDev &s, int64_t n, const float *d_o, const float *d_d, const int32_t *d_okind,
                                 const int32_t *d_oidx, const float *d_sdist, float *d_t, int32_t *d_kind, int32_t *d_idx, int32_t *d_mat,
                                 float *d_normal, DeviceCounters *dc, cudaStream_t st) {
    Queues &Q = w->q;
    if (n < 0 || static_cast<uint64_t>(n) > Q.cap_rays || static_cast<uint64_t>(n) > Q.cap_shadow) { return cudaErrorInvalidValue; }
    if (n == 0) { return cudaSuccess; }
    WF_TRY(cudaMemsetAsync(Q.counts, 0, kWfCounts * sizeof(uint32_t), st));
    const int blocks = static_cast<int>((n + 255) / 256), grid = w->sm_count * MRT_TRAV_MIN_BLOCKS;
    k_wf_test_fill<<<blocks, 256, 0, st>>>(Q, n, d_o, d_d, d_okind, d_oidx, d_sdist);
    if (s.grid) {
        k_wf_extend_grid<<<grid, 256, 0, st>>>(s, Q, kTestLevel);
        k_wf_shadow_grid<<<grid, 256, 0, st>>>(s, Q);
    } else {
        k_wf_extend<<<grid, kWfThreads, 0, st>>>(s, Q, kTestLevel, dc);
        k_wf_shadow<<<grid, kWfThreads, 0, st>>>(s, Q, dc);
    }
    k_wf_test_gather<<<blocks, 256, 0, st>>>(s, Q, d_t, d_kind, d_idx, d_mat, d_normal);
    return cudaGetLastError();
}

// ---- the same per-ray path without any allocation (mrt_trace_rays_static, include/mobilert_b200.h) -------------------
// The queues the extend / shadow kernels work on are scratch memory; here they live in a file-scope device array instead of
// cudaMalloc'ed buffers, so that a caller that may not allocate (stream capture, a tight serving loop) can still trace rays.
// Every GPU that loads the library holds the array until the process exits. All calls share it and it keeps the previous
// call's contents: the counters at its start are cleared per call, every other record that is read was written by the same
// call first. Calls that overlap on one GPU race on it.
//   [0, 256)                     kWfCounts counters (Queues::counts)
//   7 x kStaticRays x 16 bytes   r0 r1 r3 hit s0 s1 s2
//   kStaticRays bytes            vis
constexpr uint32_t kStaticRays = MRT_STATIC_TRACE_MAX_RAYS;
constexpr size_t kStaticScratchBytes = 256 + static_cast<size_t>(kStaticRays) * (7 * sizeof(float4) + 1);
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mutex;
unsigned char *g_static_ptr[kMaxDevices];   // device address of g_static_scratch on each device (the symbol itself is not one)
int g_static_sms[kMaxDevices];
}  // namespace

uint32_t wavefront_static_capacity() { return kStaticRays; }

cudaError_t wavefront_trace_rays_static(const SceneDev &s, int64_t n, const float *d_o, const float *d_d, const int32_t *d_okind,
                                        const int32_t *d_oidx, const float *d_sdist, float *d_t, int32_t *d_kind, int32_t *d_idx, int32_t *d_mat,
                                        float *d_normal, DeviceCounters *dc, cudaStream_t st) {
    if (n < 0 || static_cast<uint64_t>(n) > kStaticRays) { return cudaErrorInvalidValue; }
    if (n == 0) { return cudaSuccess; }
    int dev = -1;
    WF_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { return cudaErrorInvalidDevice; }
    unsigned char *base = nullptr;
    int sms = 0;
    {
        std::lock_guard<std::mutex> lock(g_static_mutex);
        if (g_static_ptr[dev] == nullptr) {   // once per device, with that device current
            void *p = nullptr;
            WF_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
            WF_TRY(cudaDeviceGetAttribute(&g_static_sms[dev], cudaDevAttrMultiProcessorCount, dev));
            g_static_ptr[dev] = static_cast<unsigned char *>(p);
        }
        base = g_static_ptr[dev];
        sms = g_static_sms[dev];
    }
    Queues Q {};
    Q.counts = reinterpret_cast<uint32_t *>(base);
    float4 *rec = reinterpret_cast<float4 *>(base + 256);
    Q.r0[0] = rec; Q.r1[0] = rec + kStaticRays; Q.r3[0] = rec + 2ull * kStaticRays; Q.hit = rec + 3ull * kStaticRays;
    Q.s0 = rec + 4ull * kStaticRays; Q.s1 = rec + 5ull * kStaticRays; Q.s2 = rec + 6ull * kStaticRays;
    Q.vis = reinterpret_cast<uint8_t *>(rec + 7ull * kStaticRays);
    Q.cap_rays = kStaticRays; Q.cap_shadow = kStaticRays; Q.cap_chains = 0;
    WF_TRY(cudaMemsetAsync(Q.counts, 0, kWfCounts * sizeof(uint32_t), st));
    const int blocks = static_cast<int>((n + 255) / 256), grid = sms * MRT_TRAV_MIN_BLOCKS;
    k_wf_test_fill<<<blocks, 256, 0, st>>>(Q, n, d_o, d_d, d_okind, d_oidx, d_sdist);
    if (s.grid) {
        k_wf_extend_grid<<<grid, 256, 0, st>>>(s, Q, kTestLevel);
        k_wf_shadow_grid<<<grid, 256, 0, st>>>(s, Q);
    } else {
        k_wf_extend<<<grid, kWfThreads, 0, st>>>(s, Q, kTestLevel, dc);
        k_wf_shadow<<<grid, kWfThreads, 0, st>>>(s, Q, dc);
    }
    k_wf_test_gather<<<blocks, 256, 0, st>>>(s, Q, d_t, d_kind, d_idx, d_mat, d_normal);
    return cudaGetLastError();
}

}  // namespace mrt
