// Micro-benchmark: how fast can 192-byte rows be gathered from HBM on this GPU, as a function of
// how local the gather order is?  (Design input for the vote kernel; not part of the library.)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench tools/gather_bench.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>

template <int ROWS_IN_FLIGHT>
__global__ void __launch_bounds__(256) gather_kernel(const uint8_t *__restrict__ payload, const uint32_t *__restrict__ order,
                                                     uint32_t n_rows, uint32_t stride, uint32_t *__restrict__ sink)
{
    const int lane = threadIdx.x & 31, gl = lane & 7, grp = lane >> 3;
    const uint32_t groups = (gridDim.x * blockDim.x) >> 3;
    uint32_t g = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 4 + grp;
    uint32_t acc = 0;
    const uint32_t per = (n_rows + groups - 1) / groups;
    const uint32_t lo = g * per, hi = min(n_rows, lo + per);
    for (uint32_t r0 = lo; r0 < hi; r0 += ROWS_IN_FLIGHT) {
        uint2 s[ROWS_IN_FLIGHT];
        uint4 q[ROWS_IN_FLIGHT];
#pragma unroll
        for (int k = 0; k < ROWS_IN_FLIGHT; k++) {
            uint32_t r = order[min(r0 + k, hi - 1)];
            const uint8_t *src = payload + (uint64_t)r * stride;
            s[k] = __ldg(reinterpret_cast<const uint2 *>(src + gl * 8));
            q[k] = __ldg(reinterpret_cast<const uint4 *>(src + 64 + gl * 16));
        }
#pragma unroll
        for (int k = 0; k < ROWS_IN_FLIGHT; k++) acc ^= s[k].x ^ s[k].y ^ q[k].x ^ q[k].y ^ q[k].z ^ q[k].w;
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

int main(int argc, char **argv)
{
    const uint32_t n = argc > 1 ? atoi(argv[1]) : 20000000, stride = 192;
    uint8_t *payload; uint32_t *order, *sink;
    cudaMalloc(&payload, (size_t)n * stride); cudaMalloc(&order, (size_t)n * 4); cudaMalloc(&sink, 4);
    cudaMemset(payload, 1, (size_t)n * stride);
    std::vector<uint32_t> h(n);
    std::mt19937 rng(1);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const uint32_t windows[] = {1, 16, 256, 4096, 65536, 1u << 20, n};
    for (uint32_t w : windows) {
        for (uint32_t i = 0; i < n; i++) h[i] = i;
        if (w > 1) for (uint32_t a = 0; a < n; a += w) std::shuffle(h.begin() + a, h.begin() + std::min<uint64_t>(n, (uint64_t)a + w), rng);
        cudaMemcpy(order, h.data(), (size_t)n * 4, cudaMemcpyHostToDevice);
        for (int blocks_per_sm : {2, 4, 8}) {
            float best = 1e9;
            for (int rep = 0; rep < 4; rep++) {
                cudaEventRecord(e0);
                gather_kernel<4><<<148 * blocks_per_sm, 256>>>(payload, order, n, stride, sink);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); best = std::min(best, ms);
            }
            printf("window %9u  blocks/SM %d  rows-in-flight 4: %.3f ms  %.0f GB/s\n", w, blocks_per_sm, best, (double)n * stride / best / 1e6);
            best = 1e9;
            for (int rep = 0; rep < 4; rep++) {
                cudaEventRecord(e0);
                gather_kernel<8><<<148 * blocks_per_sm, 256>>>(payload, order, n, stride, sink);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); best = std::min(best, ms);
            }
            printf("window %9u  blocks/SM %d  rows-in-flight 8: %.3f ms  %.0f GB/s\n", w, blocks_per_sm, best, (double)n * stride / best / 1e6);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
