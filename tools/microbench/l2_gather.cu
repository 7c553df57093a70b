// Own roofline microbenchmark: what the B200 delivers for the access pattern of the BVH8
// traversal kernels - every lane gathers one record (80-byte node = 5 x 16 B, or 48-byte triangle = 3 x 16 B) from a random
// place in an L2-resident buffer, with the launch shape of k_wf_extend (148 x 8 CTAs x 128 threads, persistent).
// Not part of the product library: built into libmrt_microbench.so by __graft_entry__.build(), called by bench.py only.
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

__device__ __forceinline__ uint32_t mix(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

// Each lane follows its own pseudo-random sequence of record indices; the index of step k+1 does not depend on the data
// of step k (a ray's next node does, but the warps of a traversal kernel hide that latency with 32 resident warps per
// SM, which is what this loop's independent steps stand for). Every 16-byte piece of a record is loaded and consumed.
template <int PIECES>
__global__ void __launch_bounds__(128, 8) k_gather(const float4 *__restrict__ buf, uint32_t n_records, int steps, float *sink) {
    uint32_t s = mix(blockIdx.x * 128u + threadIdx.x + 1u);
    float acc = 0.0f;
    for (int k = 0; k < steps; ++k) {
        s = s * 1664525u + 1013904223u;
        const uint32_t r = static_cast<uint32_t>((static_cast<unsigned long long>(mix(s)) * n_records) >> 32);
        const float4 *p = buf + static_cast<size_t>(r) * PIECES;
#pragma unroll
        for (int j = 0; j < PIECES; ++j) { const float4 v = __ldg(p + j); acc += v.x + v.y + v.z + v.w; }
    }
    if (acc == 123.456f) { sink[0] = acc; }
}

}  // namespace

// record_bytes: 80 or 48. buffer_bytes: size of the gathered buffer (20 MB stays in the 126 MB L2; 2 GB does not).
// Returns GB/s of record bytes gathered (best of `repeats` launches), or a negative CUDA error code.
extern "C" double mrt_microbench_gather(int record_bytes, long long buffer_bytes, int steps, int repeats) {
    const int pieces = record_bytes / 16;
    if (pieces != 5 && pieces != 3) { return -1.0; }
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint32_t n_records = static_cast<uint32_t>(buffer_bytes / record_bytes);
    float4 *buf = nullptr; float *sink = nullptr;
    if (cudaMalloc(&buf, static_cast<size_t>(n_records) * record_bytes) != cudaSuccess) { return -2.0; }
    cudaMalloc(&sink, 16);
    cudaMemset(buf, 0, static_cast<size_t>(n_records) * record_bytes);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = sms * 8;
    double best = 0.0;
    for (int r = 0; r < repeats + 1; ++r) {   // the first launch warms the L2
        cudaEventRecord(e0);
        if (pieces == 5) { k_gather<5><<<grid, 128>>>(buf, n_records, steps, sink); } else { k_gather<3><<<grid, 128>>>(buf, n_records, steps, sink); }
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -3.0; break; }
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double gbs = static_cast<double>(grid) * 128.0 * steps * record_bytes / (ms * 1e-3) / 1e9;
        if (r > 0 && gbs > best) { best = gbs; }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(buf); cudaFree(sink);
    return best;
}
