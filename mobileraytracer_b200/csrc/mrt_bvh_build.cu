// GPU builder for the compressed 8-wide BVH (replaces BVH<T>::build, MobileRT/Accelerators/BVH.hpp:160-283,
// getSplitIndexSah :396-439). The topology does not have to match the reference's binary tree; only hit
// results must. Chain of kernels, all on one stream, no host-side geometry work:
//   k_prim_bounds -> k_morton -> radix sort (hist/scan/scatter x 8) -> k_hierarchy (Karras 2012) -> k_refit
//   -> [k_top_* : binned-SAH rebuild of the upper tree over LBVH clusters] -> k_collapse8 (level-synchronous)
#include "mrt_build.h"
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace mrt {

#define CUDA_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { return e_; } } while (0)

// order-preserving float <-> int mapping so that atomicMin/atomicMax on ints order floats
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i ^ ((i >> 31) & 0x7fffffff); }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7fffffff)); }

// ---------------------------------------------------------------------------------------------------------
// 1. primitive boxes (Triangle::getAABB, Triangle.cpp:116-123) and centroid bounds
// ---------------------------------------------------------------------------------------------------------
__global__ void k_prim_bounds(const float *__restrict__ tris9, int n, float *__restrict__ plo, float *__restrict__ phi,
                              int *__restrict__ cbounds /* 6 ordered ints: min xyz, max xyz */) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float c[3] = {0, 0, 0};
    const bool valid = i < n;
    if (valid) {
        const float *t = tris9 + 9ll * i;
        for (int k = 0; k < 3; ++k) {
            const float a = t[k], b = a + t[3 + k], cc = a + t[6 + k];
            const float lo = fminf(a, fminf(b, cc)), hi = fmaxf(a, fmaxf(b, cc));
            plo[3ll * i + k] = lo; phi[3ll * i + k] = hi;
            c[k] = 0.5f * lo + 0.5f * hi;
        }
    }
    // warp reduce then one atomic per warp
    for (int k = 0; k < 3; ++k) {
        float mn = valid ? c[k] : 3.0e38f, mx = valid ? c[k] : -3.0e38f;
        for (int o = 16; o > 0; o >>= 1) {
            mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if ((threadIdx.x & 31) == 0) { atomicMin(&cbounds[k], f2ord(mn)); atomicMax(&cbounds[3 + k], f2ord(mx)); }
    }
}

__device__ __forceinline__ unsigned long long expand21(unsigned long long v) {
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

// 2. 63-bit Morton codes of the box centroids
__global__ void k_morton(const float *__restrict__ plo, const float *__restrict__ phi, int n, const int *__restrict__ cbounds,
                         unsigned long long *__restrict__ keys, uint32_t *__restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    unsigned long long code = 0;
    for (int k = 0; k < 3; ++k) {
        const float mn = ord2f(cbounds[k]), mx = ord2f(cbounds[3 + k]);
        const float c = 0.5f * plo[3ll * i + k] + 0.5f * phi[3ll * i + k];
        const float ext = mx - mn;
        float f = ext > 0.0f ? (c - mn) / ext : 0.0f;
        f = fminf(fmaxf(f, 0.0f), 1.0f);
        const unsigned long long q = static_cast<unsigned long long>(fminf(f * 2097152.0f, 2097151.0f));
        code |= expand21(q) << (2 - k);
    }
    keys[i] = code;
    vals[i] = static_cast<uint32_t>(i);
}

// ---------------------------------------------------------------------------------------------------------
// 3. LSD radix sort of (u64 key, u32 value), 8 bits per pass, stable. Block = 256 threads = 8 warps, each warp
//    owns a contiguous segment of the block's tile so that ranks can be produced with warp match + popc.
// ---------------------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_SEG = 512;                  // keys per warp
constexpr int RS_TILE = RS_WARPS * RS_SEG;   // keys per block

__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const unsigned long long *__restrict__ keys, int n, int shift,
                                                        uint32_t *__restrict__ hist, int nblocks) {
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const long long base = static_cast<long long>(blockIdx.x) * RS_TILE;
    for (int j = threadIdx.x; j < RS_TILE; j += RS_THREADS) {
        const long long i = base + j;
        if (i < n) { atomicAdd(&sh[(keys[i] >> shift) & 0xff], 1u); }
    }
    __syncthreads();
    hist[static_cast<long long>(threadIdx.x) * nblocks + blockIdx.x] = sh[threadIdx.x];
}

// exclusive scan over the digit-major histogram (256 * nblocks entries) by one block
__global__ void __launch_bounds__(1024) k_rs_scan(uint32_t *__restrict__ hist, long long total) {
    __shared__ uint32_t warpsum[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) { carry = 0; }
    __syncthreads();
    for (long long base = 0; base < total; base += 1024) {
        const long long i = base + threadIdx.x;
        const uint32_t v = i < total ? hist[i] : 0u;
        uint32_t x = v;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((threadIdx.x & 31) >= o) { x += y; } }
        if ((threadIdx.x & 31) == 31) { warpsum[threadIdx.x >> 5] = x; }
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warpsum[threadIdx.x];
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, w, o); if (threadIdx.x >= o) { w += y; } }
            warpsum[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t woff = (threadIdx.x >> 5) ? warpsum[(threadIdx.x >> 5) - 1] : 0u;
        const uint32_t incl = x + woff + carry;
        if (i < total) { hist[i] = incl - v; }
        __syncthreads();
        if (threadIdx.x == 1023) { carry = incl; }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ vals,
                                                           unsigned long long *__restrict__ okeys, uint32_t *__restrict__ ovals,
                                                           int n, int shift, const uint32_t *__restrict__ hist, int nblocks) {
    __shared__ uint32_t wcnt[RS_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int j = threadIdx.x; j < RS_WARPS * 256; j += RS_THREADS) { (&wcnt[0][0])[j] = 0; }
    __syncthreads();
    const long long wbase = static_cast<long long>(blockIdx.x) * RS_TILE + static_cast<long long>(warp) * RS_SEG;
    // phase 1: per-warp digit counts
    for (int j = lane; j < RS_SEG; j += 32) {
        const long long i = wbase + j;
        if (i < n) { atomicAdd(&wcnt[warp][(keys[i] >> shift) & 0xff], 1u); }
    }
    __syncthreads();
    // phase 2: per digit, exclusive scan over warps + global base of this block
    {
        const int d = threadIdx.x;
        uint32_t run = hist[static_cast<long long>(d) * nblocks + blockIdx.x];
        for (int w = 0; w < RS_WARPS; ++w) { const uint32_t c = wcnt[w][d]; wcnt[w][d] = run; run += c; }
    }
    __syncthreads();
    // phase 3: stable scatter, 32 keys at a time per warp
    for (int j0 = 0; j0 < RS_SEG; j0 += 32) {
        const long long i = wbase + j0 + lane;
        const bool valid = i < n;
        const unsigned long long key = valid ? keys[i] : 0ull;
        const uint32_t digit = valid ? static_cast<uint32_t>((key >> shift) & 0xff) : 0x100u + lane;
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        if (valid) {
            const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
            const uint32_t base = wcnt[warp][digit];
            const uint32_t dst = base + rank;
            okeys[dst] = key;
            ovals[dst] = vals[i];
            __syncwarp(peers);
            if (rank == 0) { wcnt[warp][digit] = base + __popc(peers); }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------------
// 4. LBVH hierarchy (Karras 2012). Internal nodes 0..n-2, leaves n-1..2n-2 (leaf j = node n-1+j).
// ---------------------------------------------------------------------------------------------------------
struct BinTree {
    int *left, *right, *parent;   // [2n-1] (left/right only for internal nodes)
    int *first, *count;           // [2n-1] sorted-primitive range covered by the node
    float *lo, *hi;               // [2n-1][3]
    int *flags;                   // [n-1] refit arrival counters
};

__device__ __forceinline__ int lbvh_delta(const unsigned long long *keys, int n, int i, int j) {
    if (j < 0 || j >= n) { return -1; }
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) { return 64 + __clz(i ^ j); }
    return __clzll(static_cast<long long>(a ^ b));
}

__global__ void k_hierarchy(const unsigned long long *__restrict__ keys, int n, BinTree t) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) { return; }
    const int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) { lmax <<= 1; }
    int l = 0;
    for (int s = lmax >> 1; s >= 1; s >>= 1) {
        if (lbvh_delta(keys, n, i, i + (l + s) * d) > dmin) { l += s; }
    }
    const int j = i + l * d;
    const int dnode = lbvh_delta(keys, n, i, j);
    int s = 0;
    for (int tt = (l + 1) >> 1; ; tt = (tt + 1) >> 1) {
        if (lbvh_delta(keys, n, i, i + (s + tt) * d) > dnode) { s += tt; }
        if (tt == 1) { break; }
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = (lo == gamma) ? (n - 1 + gamma) : gamma;
    const int right = (hi == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
    t.left[i] = left; t.right[i] = right;
    t.parent[left] = i; t.parent[right] = i;
    t.first[i] = lo; t.count[i] = hi - lo + 1;
    if (i == 0) { t.parent[0] = -1; }
}

// 5. bottom-up refit: second arrival at an internal node merges the children boxes
__global__ void k_refit(const uint32_t *__restrict__ sorted, const float *__restrict__ plo, const float *__restrict__ phi, int n, BinTree t) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) { return; }
    const int leaf = n - 1 + j;
    const uint32_t p = sorted[j];
    for (int k = 0; k < 3; ++k) { t.lo[3ll * leaf + k] = plo[3ll * p + k]; t.hi[3ll * leaf + k] = phi[3ll * p + k]; }
    t.first[leaf] = j; t.count[leaf] = 1;
    if (n == 1) { t.parent[leaf] = -1; return; }
    int node = t.parent[leaf];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&t.flags[node], 1) == 0) { return; }
        __threadfence();
        const int l = t.left[node], r = t.right[node];
        for (int k = 0; k < 3; ++k) {
            const float a = __ldcg(&t.lo[3ll * l + k]), b = __ldcg(&t.lo[3ll * r + k]);
            const float c = __ldcg(&t.hi[3ll * l + k]), d = __ldcg(&t.hi[3ll * r + k]);
            t.lo[3ll * node + k] = fminf(a, b);
            t.hi[3ll * node + k] = fmaxf(c, d);
        }
        node = t.parent[node];
    }
}

// ---------------------------------------------------------------------------------------------------------
// 5b. binned-SAH hierarchy (default). Top-down, level-synchronous, no physical partitioning: every primitive carries the
//     id of the node it currently belongs to and a 64-bit root-to-leaf path key; per level
//       k_sah_bin    each primitive adds its box to one of 16 bins per axis of its node (atomics on ordered ints)
//       k_sah_split  each open node sweeps its 3 x 16 bins, picks the cheapest SAH plane, creates two children
//                    (boxes and counts come straight from the bins)
//       k_sah_assign each primitive moves to its child and widens that child's centroid bounds
//     until every node holds one primitive. Sorting the primitives by path key then makes every subtree contiguous
//     (left before right), which is all the 8-wide collapse needs. Replaces the reference's single-threaded
//     10-bucket + full-sweep SAH build (BVH.hpp:160-283, 396-439); the topology differs, the hit results do not.
// ---------------------------------------------------------------------------------------------------------
#ifndef MRT_SAH_BINS
#define MRT_SAH_BINS 16
#endif
constexpr int SAH_BINS = MRT_SAH_BINS;
constexpr int SAH_BIN_WORDS = 7;                      // ordered min xyz, ordered max xyz, count
constexpr int SAH_NODE_WORDS = 3 * SAH_BINS * SAH_BIN_WORDS;
constexpr int SAH_RANK_SPLIT_DEPTH = 36;              // from this depth on, halve by arrival order so that depth stays < 64

struct SahState {
    int *node_of;                 // [n] current node of each primitive
    unsigned long long *path;     // [n] path key: bit (63 - depth) = went right at that depth
    int *cb;                      // [2n-1][6] ordered centroid bounds of each node
    int *split;                   // [2n-1] axis << 8 | bin (axis 3 = split by arrival rank)
    int *depth;                   // [2n-1]
    int *bins;                    // [capacity][SAH_NODE_WORDS]
    int *rank;                    // [2n-1] arrival counters for rank splits
    int *ctl;                     // [0] level begin, [1] level end, [2] node counter, [3] open nodes in the level
};

__device__ __forceinline__ int sah_bin_of(float c, float cmin, float cmax) {
    const float ext = cmax - cmin;
    if (!(ext > 0.0f)) { return 0; }
    const int b = static_cast<int>((c - cmin) * (static_cast<float>(SAH_BINS) / ext));
    return min(max(b, 0), SAH_BINS - 1);
}

__global__ void k_iota(uint32_t *v, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { v[i] = static_cast<uint32_t>(i); }
}

__global__ void k_sah_init(int n, const float *__restrict__ plo, const float *__restrict__ phi, BinTree t, SahState st) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) { return; }
    st.node_of[i] = 0;
    st.path[i] = 0ull;
    for (int k = 0; k < 3; ++k) {
        const float lo = plo[3ll * i + k], hi = phi[3ll * i + k];
        const float c = 0.5f * lo + 0.5f * hi;
        atomicMin(&st.cb[k], f2ord(c)); atomicMax(&st.cb[3 + k], f2ord(c));
        atomicMin(reinterpret_cast<int *>(&t.lo[k]), f2ord(lo)); atomicMax(reinterpret_cast<int *>(&t.hi[k]), f2ord(hi));
    }
}
// root box was accumulated as ordered ints: turn it back into floats
__global__ void k_sah_fix_root(int n, BinTree t, SahState st) {
    for (int k = 0; k < 3; ++k) { t.lo[k] = ord2f(reinterpret_cast<int *>(t.lo)[k]); t.hi[k] = ord2f(reinterpret_cast<int *>(t.hi)[k]); }
    t.count[0] = n; t.parent[0] = -1; st.depth[0] = 0;
    st.ctl[0] = 0; st.ctl[1] = 1; st.ctl[2] = 1; st.ctl[3] = n > 1 ? 1 : 0;
}

__global__ void k_sah_clear_bins(SahState st) {
    const long long words = static_cast<long long>(st.ctl[1] - st.ctl[0]) * SAH_NODE_WORDS;
    for (long long j = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; j < words; j += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int w = static_cast<int>(j % SAH_BIN_WORDS);
        st.bins[j] = w < 3 ? 0x7fffffff : (w < 6 ? static_cast<int>(0x80000000) : 0);
    }
}

// AGG: lanes of a warp that add to the same bin combine their boxes first (match.any + redux) and one of them issues the seven
// atomics. The primitives arrive in scene order, so neighbours mostly share node and bin while the level has few nodes - which
// is exactly when every atomic of the level lands on a handful of addresses: the first ten levels of the 9.8 M-triangle
// scene took 700 of the build's 827 ms that way (MRT_B200_BUILD_TRACE). Min, max and count do not depend on the order.
template <bool AGG>
__global__ void k_sah_bin(int n, const float *__restrict__ plo, const float *__restrict__ phi, BinTree t, SahState st) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int nd = i < n ? st.node_of[i] : -1;
    const bool active = nd >= st.ctl[0] && t.count[nd] > 1;
    const unsigned act = AGG ? __ballot_sync(0xffffffffu, active) : 0u;
    if (!active) { return; }
    float lo[3], hi[3];
    for (int k = 0; k < 3; ++k) { lo[k] = plo[3ll * i + k]; hi[k] = phi[3ll * i + k]; }
    int *nb = st.bins + static_cast<long long>(nd - st.ctl[0]) * SAH_NODE_WORDS;
    int olo[3], ohi[3];
    for (int k = 0; k < 3; ++k) { olo[k] = f2ord(lo[k]); ohi[k] = f2ord(hi[k]); }
    for (int k = 0; k < 3; ++k) {
        const float c = 0.5f * lo[k] + 0.5f * hi[k];
        const int b = sah_bin_of(c, ord2f(st.cb[6ll * nd + k]), ord2f(st.cb[6ll * nd + 3 + k]));
        int *w = nb + (k * SAH_BINS + b) * SAH_BIN_WORDS;
        if (AGG) {
            const unsigned peers = __match_any_sync(act, (nd - st.ctl[0]) * SAH_BINS + b);
            const int m0 = __reduce_min_sync(peers, olo[0]), m1 = __reduce_min_sync(peers, olo[1]), m2 = __reduce_min_sync(peers, olo[2]);
            const int x0 = __reduce_max_sync(peers, ohi[0]), x1 = __reduce_max_sync(peers, ohi[1]), x2 = __reduce_max_sync(peers, ohi[2]);
            if ((threadIdx.x & 31) == __ffs(peers) - 1) {
                atomicMin(&w[0], m0); atomicMin(&w[1], m1); atomicMin(&w[2], m2);
                atomicMax(&w[3], x0); atomicMax(&w[4], x1); atomicMax(&w[5], x2);
                atomicAdd(&w[6], __popc(peers));
            }
        } else {
            atomicMin(&w[0], olo[0]); atomicMin(&w[1], olo[1]); atomicMin(&w[2], olo[2]);
            atomicMax(&w[3], ohi[0]); atomicMax(&w[4], ohi[1]); atomicMax(&w[5], ohi[2]);
            atomicAdd(&w[6], 1);
        }
    }
}

struct SahBox { float lo[3], hi[3]; };
__device__ __forceinline__ void sah_box_reset(SahBox &b) { for (int k = 0; k < 3; ++k) { b.lo[k] = 3.0e38f; b.hi[k] = -3.0e38f; } }
__device__ __forceinline__ void sah_box_add(SahBox &b, const int *w) {
    for (int k = 0; k < 3; ++k) { b.lo[k] = fminf(b.lo[k], ord2f(w[k])); b.hi[k] = fmaxf(b.hi[k], ord2f(w[3 + k])); }
}
__device__ __forceinline__ float sah_box_area(const SahBox &b) {
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// one thread per open node of the level
__global__ void k_sah_split(BinTree t, SahState st) {
    const int begin = st.ctl[0], end = st.ctl[1];
    for (int nd = begin + blockIdx.x * blockDim.x + threadIdx.x; nd < end; nd += gridDim.x * blockDim.x) {
        const int cnt = t.count[nd];
        if (cnt <= 1) { continue; }
        const int *nb = st.bins + static_cast<long long>(nd - begin) * SAH_NODE_WORDS;
        float best = 3.0e38f; int bestAxis = 3, bestBin = 0, bestLeft = 0;
        if (st.depth[nd] < SAH_RANK_SPLIT_DEPTH) {
            for (int k = 0; k < 3; ++k) {
                float rightArea[SAH_BINS]; int rightCnt[SAH_BINS];
                SahBox rb; sah_box_reset(rb);
                int rc = 0;
                for (int b = SAH_BINS - 1; b >= 1; --b) {
                    const int *w = nb + (k * SAH_BINS + b) * SAH_BIN_WORDS;
                    if (w[6] > 0) { sah_box_add(rb, w); rc += w[6]; }
                    rightArea[b] = rc > 0 ? sah_box_area(rb) : 0.0f; rightCnt[b] = rc;
                }
                SahBox lb; sah_box_reset(lb);
                int lc = 0;
                for (int b = 0; b < SAH_BINS - 1; ++b) {       // plane between bin b and b + 1
                    const int *w = nb + (k * SAH_BINS + b) * SAH_BIN_WORDS;
                    if (w[6] > 0) { sah_box_add(lb, w); lc += w[6]; }
                    const int rcn = rightCnt[b + 1];
                    if (lc == 0 || rcn == 0) { continue; }
                    const float cost = sah_box_area(lb) * static_cast<float>(lc) + rightArea[b + 1] * static_cast<float>(rcn);
                    if (cost < best) { best = cost; bestAxis = k; bestBin = b; bestLeft = lc; }
                }
            }
        }
        const int left = atomicAdd(&st.ctl[2], 2), right = left + 1;
        t.left[nd] = left; t.right[nd] = right;
        t.parent[left] = nd; t.parent[right] = nd;
        st.depth[left] = st.depth[nd] + 1; st.depth[right] = st.depth[nd] + 1;
        st.split[nd] = (bestAxis << 8) | bestBin;
        if (bestAxis == 3) { bestLeft = cnt / 2; }
        t.count[left] = bestLeft; t.count[right] = cnt - bestLeft;
        st.rank[nd] = 0;
        SahBox lb, rb; sah_box_reset(lb); sah_box_reset(rb);
        if (bestAxis < 3) {
            for (int b = 0; b < SAH_BINS; ++b) {
                const int *w = nb + (bestAxis * SAH_BINS + b) * SAH_BIN_WORDS;
                if (w[6] > 0) { sah_box_add(b <= bestBin ? lb : rb, w); }
            }
            for (int k = 0; k < 3; ++k) { t.lo[3ll * left + k] = lb.lo[k]; t.hi[3ll * left + k] = lb.hi[k]; t.lo[3ll * right + k] = rb.lo[k]; t.hi[3ll * right + k] = rb.hi[k]; }
        } else {
            // rank split: the boxes are accumulated by k_sah_assign (as ordered ints, fixed up by k_sah_next_level)
            for (int k = 0; k < 3; ++k) {
                reinterpret_cast<int *>(t.lo)[3ll * left + k] = 0x7fffffff; reinterpret_cast<int *>(t.hi)[3ll * left + k] = static_cast<int>(0x80000000);
                reinterpret_cast<int *>(t.lo)[3ll * right + k] = 0x7fffffff; reinterpret_cast<int *>(t.hi)[3ll * right + k] = static_cast<int>(0x80000000);
            }
        }
        for (int c = left; c <= right; ++c) {
            for (int k = 0; k < 3; ++k) { st.cb[6ll * c + k] = 0x7fffffff; st.cb[6ll * c + 3 + k] = static_cast<int>(0x80000000); }
        }
    }
}

// one thread per primitive: moves it to the child its node's split sends it to
template <bool AGG>
__global__ void k_sah_assign(int n, const float *__restrict__ plo, const float *__restrict__ phi, BinTree t, SahState st) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int nd = i < n ? st.node_of[i] : -1;
    const bool active = nd >= st.ctl[0] && t.count[nd] > 1;
    const unsigned act = AGG ? __ballot_sync(0xffffffffu, active) : 0u;
    if (!active) { return; }
    const int sp = st.split[nd], axis = sp >> 8, bin = sp & 0xff;
    float c[3], lo[3], hi[3];
    for (int k = 0; k < 3; ++k) { lo[k] = plo[3ll * i + k]; hi[k] = phi[3ll * i + k]; c[k] = 0.5f * lo[k] + 0.5f * hi[k]; }
    bool goRight;
    if (axis < 3) {
        goRight = sah_bin_of(c[axis], ord2f(st.cb[6ll * nd + axis]), ord2f(st.cb[6ll * nd + 3 + axis])) > bin;
    } else {
        goRight = atomicAdd(&st.rank[nd], 1) >= t.count[t.left[nd]];
    }
    const int child = goRight ? t.right[nd] : t.left[nd];
    st.node_of[i] = child;
    if (goRight) { st.path[i] |= 1ull << (63 - st.depth[nd]); }
    if (AGG) {   // centroid bounds of the child: one lane per child and warp (see k_sah_bin)
        const unsigned peers = __match_any_sync(act, child);
        const bool leader = (threadIdx.x & 31) == __ffs(peers) - 1;
        for (int k = 0; k < 3; ++k) {
            const int oc = f2ord(c[k]);
            const int mn = __reduce_min_sync(peers, oc), mx = __reduce_max_sync(peers, oc);
            if (leader) { atomicMin(&st.cb[6ll * child + k], mn); atomicMax(&st.cb[6ll * child + 3 + k], mx); }
        }
    } else {
        for (int k = 0; k < 3; ++k) { atomicMin(&st.cb[6ll * child + k], f2ord(c[k])); atomicMax(&st.cb[6ll * child + 3 + k], f2ord(c[k])); }
    }
    if (axis == 3) {
        for (int k = 0; k < 3; ++k) {
            atomicMin(reinterpret_cast<int *>(&t.lo[3ll * child + k]), f2ord(lo[k])); atomicMax(reinterpret_cast<int *>(&t.hi[3ll * child + k]), f2ord(hi[k]));
        }
    }
}

// advances the level window to the children just created; fixes up the boxes of rank-split children
__global__ void k_sah_next_level(BinTree t, SahState st) {
    __shared__ int open;
    const int begin = st.ctl[1], end = st.ctl[2];
    if (threadIdx.x == 0) { open = 0; }
    __syncthreads();
    int mine = 0;
    for (int c = begin + threadIdx.x; c < end; c += blockDim.x) {
        const int p = t.parent[c];
        if ((st.split[p] >> 8) == 3) {
            for (int k = 0; k < 3; ++k) { t.lo[3ll * c + k] = ord2f(reinterpret_cast<int *>(t.lo)[3ll * c + k]); t.hi[3ll * c + k] = ord2f(reinterpret_cast<int *>(t.hi)[3ll * c + k]); }
        }
        if (t.count[c] > 1) { ++mine; }
    }
    if (mine) { atomicAdd(&open, mine); }
    __syncthreads();
    if (threadIdx.x == 0) { st.ctl[0] = begin; st.ctl[1] = end; st.ctl[3] = open; }
}

// after the path-key sort: leaf node of sorted position j
__global__ void k_sah_leaves(int n, const uint32_t *__restrict__ sorted, BinTree t, SahState st) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) { return; }
    const int leaf = st.node_of[sorted[j]];
    t.first[leaf] = j;
}
// first primitive of every inner node = first of its leftmost leaf (subtrees are contiguous, left before right)
__global__ void k_sah_first(int nodes, BinTree t) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nodes || t.count[b] <= 1) { return; }
    int c = b;
    while (t.count[c] > 1) { c = t.left[c]; }
    t.first[b] = t.first[c];
}

// ---------------------------------------------------------------------------------------------------------
// 6. collapse to 8-wide + quantise + pack triangles. One thread per wide node, level by level.
// ---------------------------------------------------------------------------------------------------------
struct WorkItem { int bin; int wide; };

struct CollapseArgs {
    BinTree t; int n;
    const uint32_t *sorted; const float *tris9;
    Node8 *nodes; TriGeom *tri_geom;
    uint32_t *node_counter, *tri_counter;
    const WorkItem *in; int n_in; WorkItem *out; uint32_t *n_out;
    uint32_t max_nodes;
    int max_leaf;                 // primitives per leaf slot, 1..7
    int fat_leaves;               // 1: a subtree that fits one leaf slot is never opened further
    int *error;
};

__device__ __forceinline__ float box_area(const float *lo, const float *hi) {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_collapse8(CollapseArgs a) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= a.n_in) { return; }
    const WorkItem item = a.in[w];
    const BinTree &t = a.t;
    int kids[8]; int nk = 0;
    // a binary subtree with <= max_leaf primitives can live in one leaf slot (leaves have count 1 in both hierarchies)
    const int ML = a.max_leaf;
    if (t.count[item.bin] <= ML) { kids[nk++] = item.bin; }
    else { kids[nk++] = t.left[item.bin]; kids[nk++] = t.right[item.bin]; }
    while (nk < 8) {
        // open the child with the largest surface area; subtrees that would need their own wide node go first
        // (lexicographic: adding a large constant to the area instead would swallow the area in fp32 rounding and
        // always open the first child, i.e. walk down the left spine)
        int best = -1; float bestA = -1.0f; bool bestBig = false;
        for (int k = 0; k < nk; ++k) {
            const int b = kids[k];
            if (t.count[b] <= 1) { continue; }   // a single primitive cannot be opened
            const float A = box_area(&t.lo[3ll * b], &t.hi[3ll * b]);
            const bool big = t.count[b] > ML;
            if (!big && a.fat_leaves) { continue; }
            if (best < 0 || (big && !bestBig) || (big == bestBig && A > bestA)) { bestA = A; best = k; bestBig = big; }
        }
        if (best < 0) { break; }
        const int b = kids[best];
        kids[best] = t.left[b];
        kids[nk++] = t.right[b];
    }
    // node frame
    float plo[3], phi[3];
    for (int k = 0; k < 3; ++k) { plo[k] = 3.0e38f; phi[k] = -3.0e38f; }
    for (int c = 0; c < nk; ++c) for (int k = 0; k < 3; ++k) {
        plo[k] = fminf(plo[k], t.lo[3ll * kids[c] + k]); phi[k] = fmaxf(phi[k], t.hi[3ll * kids[c] + k]);
    }
    int e8[3]; float scale[3];
    for (int k = 0; k < 3; ++k) {
        const float ext = phi[k] - plo[k];
        int e = 1;
        if (ext > 0.0f) {
            int ex; frexpf(ext / 255.0f, &ex);      // ext/255 = m * 2^ex, m in [0.5,1)  =>  2^ex >= ext/255
            e = ex + 127;
            if (e < 1) { e = 1; }
            while (e < 239 && __int_as_float(e << 23) * 255.0f < ext) { ++e; }
            if (__int_as_float(e << 23) * 255.0f < ext) { *a.error = 1; }   // extent beyond 2^119: outside the node format
        }
        e8[k] = e; scale[k] = __int_as_float(e << 23);
    }
    // greedy slot assignment: slot s prefers children lying in direction (bit k of s set => +axis k)
    float cen[8][3];
    float pc[3];
    for (int k = 0; k < 3; ++k) { pc[k] = 0.5f * plo[k] + 0.5f * phi[k]; }
    for (int c = 0; c < nk; ++c) for (int k = 0; k < 3; ++k) {
        cen[c][k] = (0.5f * t.lo[3ll * kids[c] + k] + 0.5f * t.hi[3ll * kids[c] + k]) - pc[k];
    }
    int slot_of[8]; int child_in_slot[8];
    for (int s = 0; s < 8; ++s) { child_in_slot[s] = -1; }
    for (int c = 0; c < nk; ++c) { slot_of[c] = -1; }
    for (int round = 0; round < nk; ++round) {
        float bestCost = -3.0e38f; int bc = -1, bs = -1;
        for (int c = 0; c < nk; ++c) {
            if (slot_of[c] >= 0) { continue; }
            for (int s = 0; s < 8; ++s) {
                if (child_in_slot[s] >= 0) { continue; }
                const float cost = ((s & 1) ? cen[c][0] : -cen[c][0]) + ((s & 2) ? cen[c][1] : -cen[c][1]) + ((s & 4) ? cen[c][2] : -cen[c][2]);
                if (cost > bestCost) { bestCost = cost; bc = c; bs = s; }
            }
        }
        slot_of[bc] = bs; child_in_slot[bs] = bc;
    }
    // count inner children / leaf triangles
    // the node's triangle mask has 24 bits: while the leaf slots hold more than that, the fullest one becomes an inner child
    bool force_inner[8] = {false, false, false, false, false, false, false, false};
    while (true) {
        int total = 0, big = -1, bigc = 1;
        for (int c = 0; c < nk; ++c) {
            const int cc = t.count[kids[c]];
            if (cc <= ML && !force_inner[c]) { total += cc; if (cc > bigc) { bigc = cc; big = c; } }
        }
        if (total <= 24 || big < 0) { break; }
        force_inner[big] = true;
    }
    uint32_t imask = 0; int ninner = 0, ntris = 0;
    for (int s = 0; s < 8; ++s) {
        const int c = child_in_slot[s];
        if (c < 0) { continue; }
        const int b = kids[c];
        if (t.count[b] > ML || force_inner[c]) { imask |= 1u << s; ++ninner; } else { ntris += t.count[b]; }
    }
    const uint32_t child_base = ninner ? atomicAdd(a.node_counter, static_cast<uint32_t>(ninner)) : 0u;
    const uint32_t tri_base = ntris ? atomicAdd(a.tri_counter, static_cast<uint32_t>(ntris)) : 0u;
    if (ninner && child_base + ninner > a.max_nodes) { *a.error = 1; return; }
    Node8 node;
    node.px = plo[0]; node.py = plo[1]; node.pz = plo[2];
    // stored biased by 15: the traversal's slab scale is 2^15 * 2^e (mrt_traverse.cuh, trav8_node_step)
    node.ex = static_cast<uint8_t>(e8[0] + 15); node.ey = static_cast<uint8_t>(e8[1] + 15); node.ez = static_cast<uint8_t>(e8[2] + 15);
    node.imask = static_cast<uint8_t>(imask);
    node.child_base = child_base; node.tri_base = tri_base;
    int inner_rank = 0, tri_off = 0;
    uint32_t out_base = 0;
    if (ninner) { out_base = atomicAdd(a.n_out, static_cast<uint32_t>(ninner)); }
    for (int s = 0; s < 8; ++s) {
        const int c = child_in_slot[s];
        uint8_t q[6] = {0, 0, 0, 0, 0, 0};
        if (c < 0) {
            node.meta[s] = 0;
        } else {
            const int b = kids[c];
            for (int k = 0; k < 3; ++k) {
                const float clo = t.lo[3ll * b + k], chi = t.hi[3ll * b + k];
                int ql = static_cast<int>(floorf((clo - plo[k]) / scale[k]));
                int qh = static_cast<int>(ceilf((chi - plo[k]) / scale[k]));
                ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
                while (ql > 0 && plo[k] + static_cast<float>(ql) * scale[k] > clo) { --ql; }
                while (qh < 255 && plo[k] + static_cast<float>(qh) * scale[k] < chi) { ++qh; }
                q[k] = static_cast<uint8_t>(ql); q[3 + k] = static_cast<uint8_t>(qh);
            }
            if (imask & (1u << s)) {
                node.meta[s] = static_cast<uint8_t>((1u << 5) | (24u + s));
                a.out[out_base + inner_rank] = WorkItem{b, static_cast<int>(child_base) + inner_rank};
                ++inner_rank;
            } else {
                const int cnt = t.count[b];
                node.meta[s] = static_cast<uint8_t>((static_cast<uint32_t>(cnt) << 5) | static_cast<uint32_t>(tri_off));   // binary count 1..7
                const int first = t.first[b];
                for (int j = 0; j < cnt; ++j) {
                    const uint32_t prim = a.sorted[first + j];
                    const float *tr = a.tris9 + 9ll * prim;
                    TriGeom g;
                    g.g0 = make_float4(tr[0], tr[1], tr[2], tr[3]);
                    g.g1 = make_float4(tr[4], tr[5], tr[6], tr[7]);
                    g.g2 = make_float4(tr[8], __uint_as_float(prim), 0.0f, 0.0f);
                    a.tri_geom[tri_base + tri_off + j] = g;
                }
                tri_off += cnt;
            }
        }
        node.qlox[s] = q[0]; node.qloy[s] = q[1]; node.qloz[s] = q[2];
        node.qhix[s] = q[3]; node.qhiy[s] = q[4]; node.qhiz[s] = q[5];
    }
    a.nodes[item.wide] = node;
}

// ---------------------------------------------------------------------------------------------------------
// host driver
// ---------------------------------------------------------------------------------------------------------
namespace {
// Every temporary of a build comes out of ONE device allocation: a build used to make about 35 cudaMalloc / cudaFree calls
// (13 GB of bins among them at 9.8 M triangles), and those calls - each a trip through the driver that can stall behind
// whatever else the process or the box is doing - made the measured build time vary between 120 and 700 ms for the same 100 ms
// of kernels. Buffers are carved on 256-byte boundaries; the arena is freed when the build returns.
struct Arena {
    char *base = nullptr;
    size_t cap = 0, off = 0;
    ~Arena() { if (base) { cudaFree(base); } }
    cudaError_t reserve(size_t bytes) { cap = bytes; off = 0; return cudaMalloc(reinterpret_cast<void **>(&base), bytes); }
    void *take(size_t bytes) {
        const size_t o = (off + 255) & ~static_cast<size_t>(255);
        if (o + bytes > cap) { return nullptr; }
        off = o + bytes;
        return base + o;
    }
};
struct DevBuf {
    Arena *arena = nullptr;
    void *p = nullptr;
    cudaError_t alloc(size_t bytes) { p = arena->take(bytes ? bytes : 16); return p ? cudaSuccess : cudaErrorMemoryAllocation; }
    template <typename T> T *as() { return static_cast<T *>(p); }
};
}  // namespace

void free_bvh8(Bvh8 *b) {
    if (b->nodes) { cudaFree(b->nodes); }
    if (b->tri_geom) { cudaFree(b->tri_geom); }
    *b = Bvh8{};
}

cudaError_t build_bvh8(const float *d_tris9, int n, uint32_t flags, cudaStream_t stream, Bvh8 *out) {
    *out = Bvh8{};
    cudaEvent_t ev0, ev1;
    CUDA_TRY(cudaEventCreate(&ev0));
    CUDA_TRY(cudaEventCreate(&ev1));
    CUDA_TRY(cudaEventRecord(ev0, stream));
    const bool trace = std::getenv("MRT_B200_BUILD_TRACE") != nullptr;   // diagnostic: wall time per phase / level on stderr
    auto tprev = std::chrono::steady_clock::now();
    auto lap = [&](const char *what, int a, int b) {
        if (!trace) { return; }
        cudaStreamSynchronize(stream);
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "build %-10s %3d %9d  %8.3f ms\n", what, a, b, std::chrono::duration<double, std::milli>(now - tprev).count());
        tprev = now;
    };
    const uint32_t max_nodes = static_cast<uint32_t>(n > 16 ? n : 16);
    CUDA_TRY(cudaMalloc(&out->tri_geom, static_cast<size_t>(n > 0 ? n : 1) * sizeof(TriGeom)));
    out->n_prims = static_cast<uint32_t>(n);
    if (n == 0) {
        // empty root: no child is ever hit
        CUDA_TRY(cudaMalloc(&out->nodes, sizeof(Node8)));
        CUDA_TRY(cudaMemsetAsync(out->nodes, 0, sizeof(Node8), stream));
        out->n_nodes = 1;
        CUDA_TRY(cudaStreamSynchronize(stream));
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
        return cudaSuccess;
    }
    lap("out-alloc", 0, n);
    const int T = 256;
    const int nb = (n + T - 1) / T;
    const bool sah = (flags & 1u) != 0u && n > 1;
    const int rs_blocks = (n + RS_TILE - 1) / RS_TILE;
    const int nn = 2 * n - 1;
    Arena arena;
    {
        const size_t N = static_cast<size_t>(n), NN = static_cast<size_t>(nn), M = max_nodes;
        size_t total = 24 * N + 24 + 16 * N + 8 * N + 1024ull * rs_blocks + 20 * NN + 24 * NN + 4 * N   // boxes, keys, values, histograms, binary tree
                     + 32 + 2 * sizeof(WorkItem) * M + sizeof(Node8) * M                                  // collapse: counters, two work queues, nodes before the trim
                     + 64 * 256;                                                                          // alignment slack of at most 64 buffers
        if (sah) { total += 4 * N + 24 * NN + 12 * NN + 16 + 4ull * SAH_NODE_WORDS * N; }                 // node_of, centroid bounds, split / depth / rank, control, bins
        CUDA_TRY(arena.reserve(total));
    }
    DevBuf plo, phi, cb, keys0, keys1, vals0, vals1, hist, misc;
    DevBuf tl, tr, tp, tf, tc, tlo, thi, tflags, q0, q1, wide;
    DevBuf s_node, s_cb, s_split, s_depth, s_bins, s_rank, s_ctl;
    for (DevBuf *b : {&plo, &phi, &cb, &keys0, &keys1, &vals0, &vals1, &hist, &misc, &tl, &tr, &tp, &tf, &tc, &tlo, &thi, &tflags, &q0, &q1, &wide,
                      &s_node, &s_cb, &s_split, &s_depth, &s_bins, &s_rank, &s_ctl}) { b->arena = &arena; }
    CUDA_TRY(wide.alloc(sizeof(Node8) * static_cast<size_t>(max_nodes)));
    Node8 *wide_nodes = wide.as<Node8>();
    CUDA_TRY(plo.alloc(sizeof(float) * 3ull * n)); CUDA_TRY(phi.alloc(sizeof(float) * 3ull * n));
    CUDA_TRY(cb.alloc(sizeof(int) * 6));
    CUDA_TRY(keys0.alloc(8ull * n)); CUDA_TRY(keys1.alloc(8ull * n));
    CUDA_TRY(vals0.alloc(4ull * n)); CUDA_TRY(vals1.alloc(4ull * n));
    CUDA_TRY(hist.alloc(sizeof(uint32_t) * 256ull * rs_blocks));
    {
        const int init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, static_cast<int>(0x80000000), static_cast<int>(0x80000000), static_cast<int>(0x80000000)};
        CUDA_TRY(cudaMemcpyAsync(cb.p, init, sizeof(init), cudaMemcpyHostToDevice, stream));
    }
    unsigned long long *ka = keys0.as<unsigned long long>(), *kb = keys1.as<unsigned long long>();
    uint32_t *va = vals0.as<uint32_t>(), *vb = vals1.as<uint32_t>();
    BinTree t;
    CUDA_TRY(tl.alloc(4ull * nn)); CUDA_TRY(tr.alloc(4ull * nn)); CUDA_TRY(tp.alloc(4ull * nn));
    CUDA_TRY(tf.alloc(4ull * nn)); CUDA_TRY(tc.alloc(4ull * nn));
    CUDA_TRY(tlo.alloc(12ull * nn)); CUDA_TRY(thi.alloc(12ull * nn)); CUDA_TRY(tflags.alloc(4ull * n));
    t.left = tl.as<int>(); t.right = tr.as<int>(); t.parent = tp.as<int>(); t.first = tf.as<int>(); t.count = tc.as<int>();
    t.lo = tlo.as<float>(); t.hi = thi.as<float>(); t.flags = tflags.as<int>();
    lap("alloc", 0, n);
    if (sah) {
        SahState st;
        CUDA_TRY(s_node.alloc(4ull * n)); CUDA_TRY(s_cb.alloc(24ull * nn)); CUDA_TRY(s_split.alloc(4ull * nn)); CUDA_TRY(s_depth.alloc(4ull * nn));
        CUDA_TRY(s_rank.alloc(4ull * nn)); CUDA_TRY(s_ctl.alloc(16));
        CUDA_TRY(s_bins.alloc(4ull * SAH_NODE_WORDS * static_cast<size_t>(n)));
        st.node_of = s_node.as<int>(); st.path = ka; st.cb = s_cb.as<int>(); st.split = s_split.as<int>(); st.depth = s_depth.as<int>();
        st.bins = s_bins.as<int>(); st.rank = s_rank.as<int>(); st.ctl = s_ctl.as<int>();
        {
            const int init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, static_cast<int>(0x80000000), static_cast<int>(0x80000000), static_cast<int>(0x80000000)};
            CUDA_TRY(cudaMemcpyAsync(st.cb, init, sizeof(init), cudaMemcpyHostToDevice, stream));
            CUDA_TRY(cudaMemcpyAsync(t.lo, init, 12, cudaMemcpyHostToDevice, stream));
            CUDA_TRY(cudaMemcpyAsync(t.hi, init + 3, 12, cudaMemcpyHostToDevice, stream));
        }
        k_prim_bounds<<<nb, T, 0, stream>>>(d_tris9, n, plo.as<float>(), phi.as<float>(), cb.as<int>());
        k_sah_init<<<nb, T, 0, stream>>>(n, plo.as<float>(), phi.as<float>(), t, st);
        k_sah_fix_root<<<1, 1, 0, stream>>>(n, t, st);
        int ctl[4] = {0, 1, 1, 1};
        lap("alloc", 1, n);
        lap("init", 0, n);
        const char *agg_env = std::getenv("MRT_B200_BUILD_AGG");
        const int agg_window = agg_env ? std::atoi(agg_env) : 0x7fffffff;   // measured: never slower, 6.6x faster build at 9.8 M triangles
        for (int level = 0; level < 64; ++level) {
            const int window = ctl[1] - ctl[0];
            const int gridw = (window + 127) / 128;
            k_sah_clear_bins<<<std::max(1, std::min(4096, window)), 256, 0, stream>>>(st);
            // warp-aggregated atomics (MRT_B200_BUILD_AGG = largest level window that uses them; 0 turns them off: A/B only)
            const bool agg = window <= agg_window;
            if (agg) { k_sah_bin<true><<<nb, T, 0, stream>>>(n, plo.as<float>(), phi.as<float>(), t, st); }
            else { k_sah_bin<false><<<nb, T, 0, stream>>>(n, plo.as<float>(), phi.as<float>(), t, st); }
            k_sah_split<<<gridw, 128, 0, stream>>>(t, st);
            if (agg) { k_sah_assign<true><<<nb, T, 0, stream>>>(n, plo.as<float>(), phi.as<float>(), t, st); }
            else { k_sah_assign<false><<<nb, T, 0, stream>>>(n, plo.as<float>(), phi.as<float>(), t, st); }
            k_sah_next_level<<<1, 1024, 0, stream>>>(t, st);
            CUDA_TRY(cudaMemcpyAsync(ctl, st.ctl, sizeof(ctl), cudaMemcpyDeviceToHost, stream));
            CUDA_TRY(cudaStreamSynchronize(stream));
            lap("level", level, window);
            if (ctl[3] == 0) { break; }
        }
        if (ctl[3] != 0) { return cudaErrorUnknown; }   // deeper than 64 levels: cannot happen with the rank-split rule
        // path-key sort -> subtrees contiguous, left before right
        k_iota<<<nb, T, 0, stream>>>(va, n);
        for (int pass = 0; pass < 8; ++pass) {
            const int shift = pass * 8;
            k_rs_hist<<<rs_blocks, RS_THREADS, 0, stream>>>(ka, n, shift, hist.as<uint32_t>(), rs_blocks);
            k_rs_scan<<<1, 1024, 0, stream>>>(hist.as<uint32_t>(), 256ll * rs_blocks);
            k_rs_scatter<<<rs_blocks, RS_THREADS, 0, stream>>>(ka, va, kb, vb, n, shift, hist.as<uint32_t>(), rs_blocks);
            unsigned long long *tk = ka; ka = kb; kb = tk;
            uint32_t *tv = va; va = vb; vb = tv;
        }
        lap("sort", 0, n);
        k_sah_leaves<<<nb, T, 0, stream>>>(n, va, t, st);
        k_sah_first<<<(ctl[2] + T - 1) / T, T, 0, stream>>>(ctl[2], t);
        lap("leaves", 0, ctl[2]);
    } else {
        k_prim_bounds<<<nb, T, 0, stream>>>(d_tris9, n, plo.as<float>(), phi.as<float>(), cb.as<int>());
        k_morton<<<nb, T, 0, stream>>>(plo.as<float>(), phi.as<float>(), n, cb.as<int>(), ka, va);
        for (int pass = 0; pass < 8; ++pass) {
            const int shift = pass * 8;
            k_rs_hist<<<rs_blocks, RS_THREADS, 0, stream>>>(ka, n, shift, hist.as<uint32_t>(), rs_blocks);
            k_rs_scan<<<1, 1024, 0, stream>>>(hist.as<uint32_t>(), 256ll * rs_blocks);
            k_rs_scatter<<<rs_blocks, RS_THREADS, 0, stream>>>(ka, va, kb, vb, n, shift, hist.as<uint32_t>(), rs_blocks);
            unsigned long long *tk = ka; ka = kb; kb = tk;
            uint32_t *tv = va; va = vb; vb = tv;
        }
        // ka/va now hold the sorted keys / primitive ids
        CUDA_TRY(cudaMemsetAsync(t.flags, 0, 4ull * n, stream));
        if (n > 1) { k_hierarchy<<<(n - 1 + T - 1) / T, T, 0, stream>>>(ka, n, t); }
        k_refit<<<nb, T, 0, stream>>>(va, plo.as<float>(), phi.as<float>(), n, t);
    }
    // collapse, level by level
    CUDA_TRY(misc.alloc(sizeof(uint32_t) * 8));
    uint32_t *d_node_counter = misc.as<uint32_t>(), *d_tri_counter = d_node_counter + 1, *d_nout = d_node_counter + 2;
    int *d_error = reinterpret_cast<int *>(d_node_counter + 3);
    {
        const uint32_t init[4] = {1u, 0u, 0u, 0u};
        CUDA_TRY(cudaMemcpyAsync(misc.p, init, sizeof(init), cudaMemcpyHostToDevice, stream));
    }
    CUDA_TRY(q0.alloc(sizeof(WorkItem) * static_cast<size_t>(max_nodes)));
    CUDA_TRY(q1.alloc(sizeof(WorkItem) * static_cast<size_t>(max_nodes)));
    {
        const WorkItem root{n == 1 ? 0 : 0, 0};  // node 0 is the root (internal 0, or the only leaf when n == 1)
        CUDA_TRY(cudaMemcpyAsync(q0.p, &root, sizeof(root), cudaMemcpyHostToDevice, stream));
    }
    CollapseArgs a;
    a.t = t; a.n = n; a.sorted = va; a.tris9 = d_tris9; a.nodes = wide_nodes; a.tri_geom = out->tri_geom;
    a.node_counter = d_node_counter; a.tri_counter = d_tri_counter; a.n_out = d_nout; a.max_nodes = max_nodes; a.error = d_error;
    a.fat_leaves = (flags & 2u) ? 1 : 0;
    a.max_leaf = static_cast<int>((flags >> 8) & 7u) ? static_cast<int>((flags >> 8) & 7u) : 3;
    WorkItem *qi = q0.as<WorkItem>(), *qo = q1.as<WorkItem>();
    uint32_t n_in = 1, levels = 0;
    while (n_in > 0) {
        a.in = qi; a.n_in = static_cast<int>(n_in); a.out = qo;
        CUDA_TRY(cudaMemsetAsync(d_nout, 0, sizeof(uint32_t), stream));
        k_collapse8<<<(n_in + 127) / 128, 128, 0, stream>>>(a);
        uint32_t h[4];
        CUDA_TRY(cudaMemcpyAsync(h, misc.p, sizeof(h), cudaMemcpyDeviceToHost, stream));
        CUDA_TRY(cudaStreamSynchronize(stream));
        if (h[3] != 0u) { return cudaErrorMemoryAllocation; }
        n_in = h[2];
        out->n_nodes = h[0];
        WorkItem *tq = qi; qi = qo; qo = tq;
        ++levels;
        if (levels > 256) { return cudaErrorUnknown; }
    }
    out->levels = levels;
    // the handle keeps exactly the nodes that exist (the collapse wrote into an n-node array: 787 MB for 78 MB at 9.8 M triangles)
    CUDA_TRY(cudaMalloc(&out->nodes, static_cast<size_t>(out->n_nodes) * sizeof(Node8)));
    CUDA_TRY(cudaMemcpyAsync(out->nodes, wide_nodes, static_cast<size_t>(out->n_nodes) * sizeof(Node8), cudaMemcpyDeviceToDevice, stream));
    lap("collapse", static_cast<int>(levels), static_cast<int>(out->n_nodes));
    CUDA_TRY(cudaEventRecord(ev1, stream));
    CUDA_TRY(cudaEventSynchronize(ev1));
    CUDA_TRY(cudaEventElapsedTime(&out->build_ms, ev0, ev1));
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    if (trace) { std::fprintf(stderr, "build total (CUDA events, stream time incl. host gaps) %.3f ms\n", out->build_ms); }
    CUDA_TRY(cudaGetLastError());
    return cudaSuccess;
}

}  // namespace mrt
