// Onesweep LSD radix sort of (u32 key, u32 value) pairs, 8 bits per pass (sm_100a).
//
// One kernel per pass: a tile ranks its keys (per-warp match_any, stable), publishes its digit
// counts, obtains its global offsets by decoupled look-back over the tiles before it, reorders
// the tile through shared memory and writes runs of equal digit contiguously.  The digit totals
// of all passes come from one histogram kernel that reads the keys once.  Stability keeps the
// reads of a family in arrival order (consensus_helper.py:561).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ccgpu {

constexpr int OS_THREADS = 256;
#ifndef OS_ITEMS_CFG
#define OS_ITEMS_CFG 16
#endif
#ifndef OS_BLOCKS_CFG
#define OS_BLOCKS_CFG 3
#endif
constexpr int OS_ITEMS = OS_ITEMS_CFG;
constexpr int OS_TILE = OS_THREADS * OS_ITEMS;      // keys per CTA
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_BINS = 256;
constexpr int OS_MAX_PASSES = 4;

inline uint64_t os_tiles(uint64_t n) { return (n + OS_TILE - 1) / OS_TILE; }

// ghist[pass][digit] += occurrences, for every pass at once
__global__ void __launch_bounds__(256)
os_histogram_kernel(const uint32_t *__restrict__ keys, uint64_t n, int passes, uint32_t *__restrict__ ghist)
{
    __shared__ uint32_t h[OS_MAX_PASSES][OS_BINS];
    for (int k = threadIdx.x; k < OS_MAX_PASSES * OS_BINS; k += blockDim.x) (&h[0][0])[k] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t k = keys[i];
        for (int p = 0; p < passes; p++) atomicAdd(&h[p][(k >> (8 * p)) & 0xFF], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < passes * OS_BINS; k += blockDim.x) {
        uint32_t v = (&h[0][0])[k];
        if (v) atomicAdd(&ghist[k], v);
    }
}

// ghist[pass][digit] -> exclusive prefix over the digits of each pass (one warp per pass)
__global__ void os_scan_hist_kernel(uint32_t *__restrict__ ghist, int passes)
{
    const int lane = threadIdx.x & 31, p = threadIdx.x >> 5;
    if (p >= passes) return;
    uint32_t v[8], sum = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) { v[k] = ghist[p * OS_BINS + lane * 8 + k]; sum += v[k]; }
    uint32_t incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    uint32_t run = incl - sum;
#pragma unroll
    for (int k = 0; k < 8; k++) { ghist[p * OS_BINS + lane * 8 + k] = run; run += v[k]; }
}

__device__ __forceinline__ uint32_t os_ld(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void os_st(uint32_t *p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// status[tile][digit]: top 2 bits = 0 empty, 1 tile count, 2 inclusive prefix; low 30 bits = value
// vals_in == nullptr means "value = position" (first pass over the hashed keys)
__global__ void __launch_bounds__(OS_THREADS, OS_BLOCKS_CFG)
os_pass_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
               uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n, int shift,
               const uint32_t *__restrict__ digit_base, uint32_t *__restrict__ status, uint32_t *__restrict__ ticket)
{
    __shared__ uint32_t wcnt[OS_WARPS][OS_BINS];
    __shared__ uint32_t bin_start[OS_BINS];     // first sorted position of the digit inside the tile
    __shared__ uint32_t goff[OS_BINS];          // global position of the digit's first key of this tile
    __shared__ uint32_t s_key[OS_TILE];
    __shared__ uint32_t s_val[OS_TILE];
    __shared__ uint32_t s_tile;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int k = tid; k < OS_WARPS * OS_BINS; k += OS_THREADS) (&wcnt[0][0])[k] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t tbase = (uint64_t)tile * OS_TILE;
    const uint64_t wbase = tbase + (uint64_t)warp * (OS_ITEMS * 32);
    uint32_t key[OS_ITEMS], val[OS_ITEMS];
    uint16_t rank[OS_ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        bool valid = idx < n;
        key[i] = valid ? keys_in[idx] : 0xFFFFFFFFu;
        val[i] = valid ? (vals_in ? vals_in[idx] : (uint32_t)idx) : 0u;
    }
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        bool valid = idx < n;
        uint32_t d = (key[i] >> shift) & 0xFF;
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
    // digit `tid`: prefix over the warps, tile total, look-back for the tile's global offset
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < OS_WARPS; w++) {
        uint32_t t = wcnt[w][tid];
        wcnt[w][tid] = total;
        total += t;
    }
    uint32_t *mine = status + (uint64_t)tile * OS_BINS + tid;
    uint32_t excl = 0;
    if (tile == 0) {
        os_st(mine, (2u << 30) | total);
    } else {
        os_st(mine, (1u << 30) | total);
        for (int64_t p = (int64_t)tile - 1;; --p) {
            uint32_t s;
            do { s = os_ld(status + (uint64_t)p * OS_BINS + tid); } while ((s >> 30) == 0);
            excl += s & 0x3FFFFFFFu;
            if ((s >> 30) == 2) break;
        }
        os_st(mine, (2u << 30) | (excl + total));
    }
    goff[tid] = digit_base[tid] + excl;
    // exclusive prefix of the tile totals over the digits -> where each digit starts inside the tile
    uint32_t incl = total;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    __shared__ uint32_t wsum[OS_WARPS];
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < OS_WARPS; w++) if (w < warp) woff += wsum[w];
    bin_start[tid] = woff + incl - total;
    __syncthreads();
    // reorder the tile in shared memory
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        uint64_t idx = wbase + (uint64_t)i * 32 + lane;
        if (idx < n) {
            uint32_t d = (key[i] >> shift) & 0xFF;
            uint32_t pos = bin_start[d] + wcnt[warp][d] + rank[i];
            s_key[pos] = key[i];
            s_val[pos] = val[i];
        }
    }
    __syncthreads();
    const uint32_t count = (uint32_t)min((uint64_t)OS_TILE, n - tbase);
    for (uint32_t e = tid; e < count; e += OS_THREADS) {
        uint32_t k = s_key[e];
        uint32_t d = (k >> shift) & 0xFF;
        uint32_t dst = goff[d] + (e - bin_start[d]);
        keys_out[dst] = k;
        vals_out[dst] = s_val[e];
    }
}

}  // namespace ccgpu
