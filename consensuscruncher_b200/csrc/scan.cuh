// Single-pass prefix sums with decoupled look-back (sm_100a).
//
// Three flavours share one kernel body:
//   exclusive scan of u32 values                       -> out[i]
//   compaction of the indices whose flag has bit 0 set -> list[rank] = i
//   exclusive scan of the popcounts of u32 words       -> out[i]  (ranks inside a bitmap)
// Tiles take their number from an atomic ticket so that a tile only ever waits
// on tiles that have already started.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ccgpu {

constexpr int SCAN_THREADS = 256;
#ifndef SCAN_ITEMS_CFG
#define SCAN_ITEMS_CFG 32
#endif
constexpr int SCAN_ITEMS = SCAN_ITEMS_CFG;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

struct ScanScratch {
    unsigned long long *status;   // one word per tile: flag << 32 | value
    uint32_t *ticket;             // tile ticket counter
    uint64_t max_tiles;
};

__device__ __forceinline__ unsigned long long ld_status(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_status(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

enum { SCAN_MODE_EXCL = 0, SCAN_MODE_COMPACT = 1, SCAN_MODE_POPC = 2 };   // POPC: exclusive scan of popcounts

// In: u32 values (MODE_EXCL), u8 flags (MODE_COMPACT) or u32 bitmap words (MODE_POPC).
// MODE_COMPACT can also write the inverse map: row_out[i] = *row_base + rank of i if flagged, else
// row_alt[i] (or -1 without row_alt) - the "which output row is family i" arrays, for free.
template <int MODE, typename InT>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_kernel(const InT *in, uint32_t *out, uint32_t *carry,
            uint64_t n, unsigned long long *status, uint32_t *ticket, uint32_t *total,
            int32_t *row_out = nullptr, const int32_t *row_alt = nullptr, const uint32_t *row_base = nullptr)
{
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    __shared__ uint32_t s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t base = (uint64_t)tile * SCAN_TILE + (uint64_t)tid * SCAN_ITEMS;

    InT v[SCAN_ITEMS];
    uint32_t w[SCAN_ITEMS];
    uint32_t sum = 0;
    // a thread's 32 consecutive items are one 32-byte (flags) or 128-byte (words) piece: 16-byte loads
    constexpr int PER_VEC = 16 / (int)sizeof(InT);
    static_assert(SCAN_ITEMS % PER_VEC == 0, "items per thread must fill whole 16-byte loads");
    if (base + SCAN_ITEMS <= n) {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS / PER_VEC; k++) {
            const uint4 q = *reinterpret_cast<const uint4 *>(in + base + k * PER_VEC);
            memcpy(&v[k * PER_VEC], &q, 16);
        }
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < n) ? in[base + i] : (InT)0;
    }
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        w[i] = (MODE == SCAN_MODE_EXCL) ? (uint32_t)v[i] : (MODE == SCAN_MODE_COMPACT ? ((uint32_t)v[i] & 1u) : (uint32_t)__popc((uint32_t)v[i]));
        sum += w[i];
    }
    // block-wide exclusive prefix of the per-thread sums
    uint32_t incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, block_total = 0;
#pragma unroll
    for (int k = 0; k < SCAN_THREADS / 32; k++) {
        uint32_t t = s_warp[k];
        if (k < warp) warp_off += t;
        block_total += t;
    }
    if (warp == 0) {
        // look-back by a whole warp: 32 predecessors per round, nearest first
        uint32_t excl = 0;
        if (tile != 0) {
            if (lane == 0) st_status(&status[tile], (1ull << 32) | block_total);
            for (int64_t p = (int64_t)tile - 1;; p -= 32) {
                const int64_t idx = p - lane;
                unsigned long long s = idx >= 0 ? ld_status(&status[idx]) : (2ull << 32);   // before tile 0: prefix 0
                while (__any_sync(0xffffffffu, (s >> 32) == 0))
                    if ((s >> 32) == 0) s = ld_status(&status[idx]);
                const uint32_t done = __ballot_sync(0xffffffffu, (s >> 32) == 2);
                const int first = done ? __ffs(done) - 1 : 31;          // nearest tile that already knows its prefix
                uint32_t v = lane <= first ? (uint32_t)s : 0u;
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
                excl += v;
                if (done) break;
            }
        }
        if (lane == 0) {
            st_status(&status[tile], (2ull << 32) | (uint32_t)(excl + block_total));
            s_prefix = excl;
            if (total != nullptr && tile == gridDim.x - 1)
                *total = excl + block_total;  // the last tile publishes the grand total
        }
    }
    __syncthreads();
    uint32_t run = s_prefix + warp_off + (incl - sum);
    const uint32_t first_row = (MODE == SCAN_MODE_COMPACT && row_out != nullptr && row_base != nullptr) ? *row_base : 0u;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        if (base + i < n) {
            if (MODE != SCAN_MODE_COMPACT) {
                if (base + SCAN_ITEMS > n) out[base + i] = run;       // whole pieces are stored below, 16 bytes at a time
            } else if (w[i]) {
                out[run] = (uint32_t)(base + i);
            }
        }
        const uint32_t step = w[i];
        if (MODE == SCAN_MODE_COMPACT) w[i] = w[i] ? first_row + run + 1u : 0u;     // row + 1, 0 = none (for row_out below)
        else w[i] = run;                                                            // the exclusive prefix itself
        run += step;
    }
    if (MODE != SCAN_MODE_COMPACT && base + SCAN_ITEMS <= n) {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i += 4)
            *reinterpret_cast<uint4 *>(out + base + i) = make_uint4(w[i], w[i + 1], w[i + 2], w[i + 3]);
    }
    if (MODE == SCAN_MODE_COMPACT && row_out != nullptr) {
        // a thread owns 32 consecutive entries = one 128-byte line: written as eight 16-byte stores
        static_assert(SCAN_ITEMS % 4 == 0, "row_out is written four entries at a time");
        if (base + SCAN_ITEMS <= n) {
#pragma unroll
            for (int i = 0; i < SCAN_ITEMS; i += 4) {
                int4 a = row_alt != nullptr ? *reinterpret_cast<const int4 *>(row_alt + base + i) : make_int4(-1, -1, -1, -1);
                int4 v;
                v.x = w[i] ? (int32_t)(w[i] - 1u) : a.x;
                v.y = w[i + 1] ? (int32_t)(w[i + 1] - 1u) : a.y;
                v.z = w[i + 2] ? (int32_t)(w[i + 2] - 1u) : a.z;
                v.w = w[i + 3] ? (int32_t)(w[i + 3] - 1u) : a.w;
                *reinterpret_cast<int4 *>(row_out + base + i) = v;
            }
        } else {
            for (int i = 0; i < SCAN_ITEMS && base + i < n; i++)
                row_out[base + i] = w[i] ? (int32_t)(w[i] - 1u) : (row_alt != nullptr ? row_alt[base + i] : -1);
        }
    }
}

inline uint64_t scan_tiles(uint64_t n) { return (n + SCAN_TILE - 1) / SCAN_TILE; }

// ---- chained scan of several sums at once -------------------------------------------------------
// A status word carries a state (2 bits: 0 nothing yet, 1 this tile's own sums, 2 sums of every tile up to
// and including this one) and two 31-bit sums; a tile owns W consecutive words (status[tile * W + w]), so
// one kernel can carry 2 * W running sums.  Each word is looked back on its own: predecessors publish
// their words one after the other.  Called by the first warp of the CTA; tot[] / excl[] hold 2 * W sums.
__device__ __forceinline__ unsigned long long pair_pack(uint32_t state, uint32_t a, uint32_t b)
{
    return ((unsigned long long)state << 62) | ((unsigned long long)a << 31) | b;
}

template <int W>
__device__ __forceinline__ void lookback_pairs(unsigned long long *status, uint32_t tile, const uint32_t *tot, uint32_t *excl,
                                               int lane)
{
#pragma unroll
    for (int w = 0; w < W; w++) {
        uint32_t ea = 0, eb = 0;
        if (tile != 0) {
            if (lane == 0) st_status(&status[(uint64_t)tile * W + w], pair_pack(1, tot[2 * w], tot[2 * w + 1]));
            for (int64_t p = (int64_t)tile - 1;; p -= 32) {
                const int64_t idx = p - lane;
                unsigned long long s = idx >= 0 ? ld_status(&status[idx * W + w]) : pair_pack(2, 0, 0);   // before tile 0: nothing
                while (__any_sync(0xffffffffu, (s >> 62) == 0))
                    if ((s >> 62) == 0) s = ld_status(&status[idx * W + w]);
                const uint32_t done = __ballot_sync(0xffffffffu, (s >> 62) == 2);
                const int first = done ? __ffs(done) - 1 : 31;            // nearest tile that already knows its prefix
                uint32_t va = lane <= first ? (uint32_t)(s >> 31) & 0x7FFFFFFFu : 0u;
                uint32_t vb = lane <= first ? (uint32_t)s & 0x7FFFFFFFu : 0u;
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) {
                    va += __shfl_xor_sync(0xffffffffu, va, d);
                    vb += __shfl_xor_sync(0xffffffffu, vb, d);
                }
                ea += va; eb += vb;
                if (done) break;
            }
        }
        if (lane == 0) st_status(&status[(uint64_t)tile * W + w], pair_pack(2, ea + tot[2 * w], eb + tot[2 * w + 1]));
        excl[2 * w] = ea; excl[2 * w + 1] = eb;
    }
}

}  // namespace ccgpu
