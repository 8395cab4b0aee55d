// Stable LSD radix sort of (u32 key, u32 value) pairs, 8 bits per pass (sm_100a).
//
// One pass = per-tile digit histogram -> exclusive scan of the digit-major
// count matrix (scan.cuh) -> stable scatter.  Stability is what keeps the
// reads of a family in arrival order after the sort, which the reference's
// "first read is the template" and first-appearance tie rules need
// (consensus_helper.py:561).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ccgpu {

constexpr int RADIX_THREADS = 256;
constexpr int RADIX_ITEMS = 16;
constexpr int RADIX_TILE = RADIX_THREADS * RADIX_ITEMS;   // 4096 keys per CTA
constexpr int RADIX_WARPS = RADIX_THREADS / 32;
constexpr int RADIX_BINS = 256;

inline uint64_t radix_tiles(uint64_t n) { return (n + RADIX_TILE - 1) / RADIX_TILE; }

// counts[d * ntiles + tile] = number of keys of `tile` whose digit is d
__global__ void __launch_bounds__(RADIX_THREADS)
radix_hist_kernel(const uint32_t *__restrict__ keys, uint64_t n, int shift, uint32_t ntiles,
                  uint32_t *__restrict__ counts)
{
    __shared__ uint32_t h[RADIX_BINS];
    const int tid = threadIdx.x;
    h[tid] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * RADIX_TILE;
#pragma unroll 4
    for (int i = 0; i < RADIX_ITEMS; i++) {
        uint64_t idx = base + (uint64_t)i * RADIX_THREADS + tid;
        if (idx < n) atomicAdd(&h[(keys[idx] >> shift) & 0xFF], 1u);
    }
    __syncthreads();
    counts[(uint64_t)tid * ntiles + blockIdx.x] = h[tid];
}

// vals_in == nullptr means "value = position" (first pass over the hashed keys)
__global__ void __launch_bounds__(RADIX_THREADS)
radix_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                     uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n,
                     int shift, uint32_t ntiles, const uint32_t *__restrict__ offsets)
{
    __shared__ uint32_t wcnt[RADIX_WARPS][RADIX_BINS];
    __shared__ uint32_t goff[RADIX_BINS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < RADIX_WARPS * RADIX_BINS; k += RADIX_THREADS) (&wcnt[0][0])[k] = 0;
    __syncthreads();
    // each warp owns RADIX_ITEMS*32 consecutive keys; order inside a warp is (item, lane)
    const uint64_t wbase = (uint64_t)blockIdx.x * RADIX_TILE + (uint64_t)warp * (RADIX_ITEMS * 32);
    uint32_t key[RADIX_ITEMS], val[RADIX_ITEMS];
    uint16_t rank[RADIX_ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < RADIX_ITEMS; i++) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        bool valid = idx < n;
        key[i] = valid ? keys_in[idx] : 0xFFFFFFFFu;
        val[i] = valid ? (vals_in ? vals_in[idx] : (uint32_t)idx) : 0u;
    }
#pragma unroll
    for (int i = 0; i < RADIX_ITEMS; i++) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        bool valid = idx < n;
        uint32_t d = (key[i] >> shift) & 0xFF;
        // invalid lanes vote in a bin of their own so they never disturb a real digit
        uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : 0x100u);
        uint32_t before = __popc(peers & lt);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = wcnt[warp][d];
            wcnt[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = (uint16_t)(old + before);
        __syncwarp();
    }
    __syncthreads();
    {   // digit `tid`: exclusive prefix over the warps, and the tile's global offset
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RADIX_WARPS; w++) {
            uint32_t t = wcnt[w][tid];
            wcnt[w][tid] = run;
            run += t;
        }
        goff[tid] = offsets[(uint64_t)tid * ntiles + blockIdx.x];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < RADIX_ITEMS; i++) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        if (idx < n) {
            uint32_t d = (key[i] >> shift) & 0xFF;
            uint32_t dst = goff[d] + wcnt[warp][d] + rank[i];
            keys_out[dst] = key[i];
            vals_out[dst] = val[i];
        }
    }
}

}  // namespace ccgpu
