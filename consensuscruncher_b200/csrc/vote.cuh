// SSCS vote for families of 2..127 reads: sub-warp groups, rows staged through registers.
//
// consensus_maker (SSCS_maker.py:103-155) + read_mode / consensus_flag
// (consensus_helper.py:483-535) for the bulk of the families.  A group of G lanes owns one
// family; a lane owns 16 consecutive positions of it (8 sequence bytes + 16 quality bytes
// per read, one 8-byte and one 16-byte load), so a warp votes 32/G families at once and
// keeps 32/G times more independent loads in flight than one family per warp.
//
// Per read and 4 positions the work is byte-lane SIMD in 32-bit registers:
//   pass mask (Q >= 30), base & mask, one-hot counts per base, count of passing reads,
//   and a sticky check that every passing base is exactly one of A,C,G,T.
// Qualities are not accumulated: a passing quality is >= 30, so the capped sum
// min(60, sum) is 60 as soon as two passing reads support the winner.  Only where the
// winner has exactly one supporter is that read's own quality needed; those positions
// are resolved by a second walk over the family's rows (L1/L2 hits).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.cuh"

namespace ccgpu {

constexpr uint32_t VOTE_SMALL_MAX = 127;     // byte-lane compares below need counts < 128
constexpr int VOTE_ROWS_PER_BATCH = 4;       // rows a lane keeps in flight

__device__ __forceinline__ uint32_t pass7_of(uint32_t q)
{
    return (((q & 0x7F7F7F7Fu) + 0x62626262u) | q) & 0x80808080u;    // bit 7 of each byte: q >= 30
}

__device__ __forceinline__ uint32_t bytes_of(uint32_t bit7) { return (bit7 >> 7) * 0xFFu; }

// bit 7 of each byte: a >= b, for bytes < 128
__device__ __forceinline__ uint32_t ge7(uint32_t a, uint32_t b) { return ((a | 0x80808080u) - b) & 0x80808080u; }

// bit 7 of each byte: byte != 0, for bytes < 128
__device__ __forceinline__ uint32_t nz7(uint32_t x) { return (x + 0x7F7F7F7Fu) & 0x80808080u; }

__device__ __forceinline__ uint32_t sel(uint32_t a, uint32_t b, uint32_t mask) { return (a & mask) | (b & ~mask); }

struct Quad {            // accumulators of 4 positions (one byte lane each)
    uint32_t cA, cC, cG, cT, pc;
};

__device__ __forceinline__ void quad_add(Quad &a, uint32_t nib, uint32_t q, uint32_t &err)
{
    uint32_t p1 = pass7_of(q) >> 7;
    uint32_t b = nib & (p1 * 0xFFu);
    err |= ((b | 0x80808080u) - 0x01010101u) & b;        // more than one bit set in a passing base
    a.cA += b & 0x01010101u;
    a.cC += (b >> 1) & 0x01010101u;
    a.cG += (b >> 2) & 0x01010101u;
    a.cT += (b >> 3) & 0x01010101u;
    a.pc += p1;
}

struct QuadOut {
    uint32_t code;       // winner or N per byte
    uint32_t qual;       // consensus quality per byte (valid unless `single` is set there)
    uint32_t win;        // winner's one-hot code per byte
    uint32_t single;     // 0xFF where the winner has exactly one supporter
};

// Winner, cutoff decision and quality of 4 positions.  thr[p] = smallest count c with
// (double)c / (double)p >= cutoff, built on the host with the same IEEE division Python does.
__device__ __forceinline__ QuadOut quad_finish(const Quad &a, const uint8_t *thr, bool unanimous_ok, uint32_t &err)
{
    QuadOut o;
    uint32_t m1 = bytes_of(ge7(a.cA, a.cC));              // ties keep the earlier base: A,C,G,T order
    uint32_t m2 = bytes_of(ge7(a.cG, a.cT));
    uint32_t bAC = sel(a.cA, a.cC, m1), wAC = sel(0x01010101u, 0x02020202u, m1);
    uint32_t bGT = sel(a.cG, a.cT, m2), wGT = sel(0x04040404u, 0x08080808u, m2);
    uint32_t m3 = bytes_of(ge7(bAC, bGT));
    uint32_t best = sel(bAC, bGT, m3);
    o.win = sel(wAC, wGT, m3);
    uint32_t sum = a.cA + a.cC + a.cG + a.cT;
    err |= sum ^ a.pc;                                     // a passing base that is none of A,C,G,T
    uint32_t pass = a.pc;
    uint32_t split = nz7(best ^ pass);                     // not unanimous
    uint32_t ok = bytes_of(nz7(best)) & ~bytes_of(split);  // unanimous and at least one passing read
    if (!unanimous_ok) ok = 0;
    while (split) {                                        // rare: look the threshold up per position
        int k = (__ffs(split) - 1) >> 3;
        uint32_t c = (best >> (8 * k)) & 0xFFu, p = (pass >> (8 * k)) & 0xFFu;
        if (c >= thr[p]) ok |= 0xFFu << (8 * k);
        split &= ~(0x80u << (8 * k));
    }
    o.code = sel(o.win, 0x0F0F0F0Fu, ok);
    uint32_t two = bytes_of(ge7(best, 0x02020202u));
    o.single = bytes_of(nz7(best)) & ~two;
    o.qual = two & 0x3C3C3C3Cu;                            // 60 where two or more passing reads agree
    return o;
}

// second walk: quality of the only read supporting the winner
__device__ __forceinline__ uint32_t quad_pick(uint32_t nib, uint32_t q, uint32_t win, uint32_t single)
{
    uint32_t b = nib & bytes_of(pass7_of(q));
    uint32_t same = ~bytes_of(nz7(b ^ win)) & single;      // win is never 0, so `same` implies passing
    return q & same;
}

__device__ __forceinline__ uint32_t cap60(uint32_t q)
{
    uint32_t lo = q & 0x7F7F7F7Fu;
    uint32_t big = bytes_of(ge7(lo, 0x3C3C3C3Cu) | (q & 0x80808080u));
    return sel(0x3C3C3C3Cu, lo, big);
}

struct LaneRow {
    uint2 s;    // 16 bases
    uint4 q;    // 16 qualities
};

__device__ __forceinline__ LaneRow load_lane_row(const uint8_t *__restrict__ payload, uint32_t stride, uint32_t half,
                                                 uint32_t read, uint32_t pos0)
{
    const uint8_t *src = payload + (uint64_t)read * stride;
    LaneRow r;
    r.s = __ldg(reinterpret_cast<const uint2 *>(src + (pos0 >> 1)));
    r.q = __ldg(reinterpret_cast<const uint4 *>(src + half + pos0));
    return r;
}

// quads of a lane: (even positions of word 0, odd of word 0, even of word 1, odd of word 1)
__device__ __forceinline__ void split_row(const LaneRow &r, uint32_t nib[4], uint32_t q[4])
{
    nib[0] = (r.s.x >> 4) & 0x0F0F0F0Fu; nib[1] = r.s.x & 0x0F0F0F0Fu;
    nib[2] = (r.s.y >> 4) & 0x0F0F0F0Fu; nib[3] = r.s.y & 0x0F0F0F0Fu;
    q[0] = __byte_perm(r.q.x, r.q.y, 0x6420); q[1] = __byte_perm(r.q.x, r.q.y, 0x7531);
    q[2] = __byte_perm(r.q.z, r.q.w, 0x6420); q[3] = __byte_perm(r.q.z, r.q.w, 0x7531);
}

// mode of one field over a family held by a lane group (slow path: fields differ inside the family)
template <typename Get>
__device__ uint32_t group_mode(Get get, uint32_t n, bool is_flag, int gl, int G, uint32_t gmask)
{
    uint32_t best = 0, best_cnt = 0, counted = 0;
    int best_rank = 9;
    for (uint32_t i = 0; counted < n && i < n; i++) {
        uint32_t vi = get(i), cnt = 0;
        bool seen_before = false;
        for (uint32_t j0 = 0; j0 < n; j0 += G) {
            uint32_t j = j0 + gl;
            bool m = j < n && get(j) == vi;
            uint32_t b = __ballot_sync(gmask, m) & gmask;
            if (cnt == 0 && b && j0 + (uint32_t)(__ffs(b) - 1 - (__ffs(gmask) - 1)) < i) { seen_before = true; break; }
            cnt += __popc(b);
        }
        if (seen_before) continue;
        counted += cnt;
        int rank = is_flag ? flag_rank(vi) : 4;
        if (cnt > best_cnt || (cnt == best_cnt && rank < best_rank)) { best = vi; best_cnt = cnt; best_rank = rank; }
    }
    return best;
}

// vstart / vsize: segment of `perm` of every SSCS row (filled by vote_segments_kernel)
template <int G>
__global__ void __launch_bounds__(256)
vote_small_kernel(const uint4 *__restrict__ fields, const uint8_t *__restrict__ payload, uint32_t stride,
                  uint32_t lpad, const uint32_t *__restrict__ perm, const uint32_t *__restrict__ vstart,
                  const uint32_t *__restrict__ vsize, const uint32_t *__restrict__ sscs_list, uint32_t n_rows,
                  const uint8_t *__restrict__ thr_table, int unanimous_ok, uint8_t *__restrict__ out_rows,
                  uint8_t *__restrict__ fam_mapq, uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen,
                  uint8_t *__restrict__ fam_rg, uint32_t *__restrict__ err_flag)
{
    __shared__ uint8_t thr[128];
    if (threadIdx.x < 128) thr[threadIdx.x] = thr_table[threadIdx.x];
    __syncthreads();
    constexpr int GROUPS = 32 / G;
    const int lane = threadIdx.x & 31, gl = lane % G, grp = lane / G;
    const uint32_t gmask = (G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u)) << (grp * G);
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t row = warp * GROUPS + grp;
    const uint32_t half = lpad >> 1;
    uint32_t start = 0, n = 0;
    if (row < n_rows) {
        start = __ldg(&vstart[row]);
        n = __ldg(&vsize[row]);
        if (n > VOTE_SMALL_MAX) n = 0;                     // large families go to vote_big_kernel
    }
    uint8_t *out = out_rows + (uint64_t)row * stride;
    uint32_t err = 0;

    for (uint32_t base = 0; base < lpad; base += G * 16) { // one pass unless lpad > 16 * G
        const uint32_t pos0 = base + gl * 16;
        const bool active = n != 0 && pos0 < lpad;
        Quad acc[4];
#pragma unroll
        for (int g = 0; g < 4; g++) acc[g] = Quad{0, 0, 0, 0, 0};
        if (active) {
            for (uint32_t j0 = 0; j0 < n; j0 += VOTE_ROWS_PER_BATCH) {
                uint32_t idx[VOTE_ROWS_PER_BATCH];
                LaneRow r[VOTE_ROWS_PER_BATCH];
#pragma unroll
                for (int k = 0; k < VOTE_ROWS_PER_BATCH; k++)
                    idx[k] = __ldg(&perm[start + min(j0 + k, n - 1)]);
#pragma unroll
                for (int k = 0; k < VOTE_ROWS_PER_BATCH; k++) r[k] = load_lane_row(payload, stride, half, idx[k], pos0);
#pragma unroll
                for (int k = 0; k < VOTE_ROWS_PER_BATCH; k++) {
                    if (j0 + k < n) {
                        uint32_t nib[4], q[4];
                        split_row(r[k], nib, q);
#pragma unroll
                        for (int g = 0; g < 4; g++) quad_add(acc[g], nib[g], q[g], err);
                    }
                }
            }
            QuadOut o[4];
            uint32_t any_single = 0;
#pragma unroll
            for (int g = 0; g < 4; g++) {
                o[g] = quad_finish(acc[g], thr, unanimous_ok != 0, err);
                any_single |= o[g].single;
            }
            if (any_single) {                              // lane-local: re-walk the rows for the lone supporters
                uint32_t pick[4] = {0, 0, 0, 0};
                for (uint32_t j = 0; j < n; j++) {
                    LaneRow r = load_lane_row(payload, stride, half, __ldg(&perm[start + j]), pos0);
                    uint32_t nib[4], q[4];
                    split_row(r, nib, q);
#pragma unroll
                    for (int g = 0; g < 4; g++) pick[g] |= quad_pick(nib[g], q[g], o[g].win, o[g].single);
                }
#pragma unroll
                for (int g = 0; g < 4; g++) o[g].qual |= cap60(pick[g]) & o[g].single;
            }
            uint2 so;
            uint4 qo;
            so.x = (o[0].code << 4) | o[1].code;
            so.y = (o[2].code << 4) | o[3].code;
            qo.x = __byte_perm(o[0].qual, o[1].qual, 0x5140); qo.y = __byte_perm(o[0].qual, o[1].qual, 0x7362);
            qo.z = __byte_perm(o[2].qual, o[3].qual, 0x5140); qo.w = __byte_perm(o[2].qual, o[3].qual, 0x7362);
            *reinterpret_cast<uint2 *>(out + (pos0 >> 1)) = so;
            *reinterpret_cast<uint4 *>(out + half + pos0) = qo;
        }
    }
    if (n != 0)
        for (uint32_t x = half + lpad + gl * 4; x < stride; x += G * 4) *reinterpret_cast<uint32_t *>(out + x) = 0;
    if (err) atomicOr(err_flag, 1u);

    // ---- mode fields (create_aligned_segment, consensus_helper.py:569-583) -------------------
    // every lane of the warp takes part in the ballots; groups without a family carry n = 0
    auto meta_of = [&](uint32_t j) { return __ldg(&fields[2 * (uint64_t)__ldg(&perm[start + j]) + 1]); };
    uint32_t nmax = n;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) nmax = max(nmax, __shfl_xor_sync(0xffffffffu, nmax, d));
    if (nmax == 0) return;
    uint4 m0 = n ? meta_of(0) : make_uint4(0, 0, 0, 0);
    bool uniform = true, has_rg = true;
    for (uint32_t j0 = 0; j0 < nmax; j0 += G) {
        uint32_t j = j0 + gl;
        uint4 m = j < n ? meta_of(j) : m0;
        uint32_t same = __ballot_sync(0xffffffffu, m.z == m0.z && m.w == m0.w);
        uint32_t rg_ok = __ballot_sync(0xffffffffu, j >= n || (m.w >> 24) != 0xFFu);
        uniform = uniform && (same & gmask) == gmask;
        has_rg = has_rg && (rg_ok & gmask) == gmask;
    }
    uint32_t tlen = m0.z, flag = m0.w & 0xFFFF, mapq = (m0.w >> 16) & 0xFF, rg = m0.w >> 24;
    if (n != 0 && !uniform) {                              // group-uniform branch; ballots use the group mask
        tlen = group_mode([&](uint32_t j) { return meta_of(j).z; }, n, false, gl, G, gmask);
        flag = group_mode([&](uint32_t j) { return meta_of(j).w & 0xFFFFu; }, n, true, gl, G, gmask);
        mapq = group_mode([&](uint32_t j) { return (meta_of(j).w >> 16) & 0xFFu; }, n, false, gl, G, gmask);
        rg = has_rg ? group_mode([&](uint32_t j) { return meta_of(j).w >> 24; }, n, false, gl, G, gmask) : 0xFFu;
    }
    if (n != 0 && gl == 0) {
        const uint32_t c = __ldg(&sscs_list[row]);
        fam_tlen[c] = (int32_t)tlen; fam_flag[c] = (uint16_t)flag;
        fam_mapq[c] = (uint8_t)mapq; fam_rg[c] = (uint8_t)rg;
    }
}

// segment of every SSCS row, and the list of rows too large for vote_small_kernel
__global__ void __launch_bounds__(256)
vote_segments_kernel(const uint32_t *__restrict__ sscs_list, const uint32_t *__restrict__ fam_start,
                     const uint32_t *__restrict__ fam_size, uint32_t n_rows, uint32_t *__restrict__ vstart,
                     uint32_t *__restrict__ vsize, uint32_t *__restrict__ big_list, uint32_t *__restrict__ n_big)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    uint32_t c = sscs_list[r];
    uint32_t n = fam_size[c];
    vstart[r] = fam_start[c];
    vsize[r] = n;
    if (n > VOTE_SMALL_MAX) big_list[atomicAdd(n_big, 1u)] = r;
}

// one warp per large family, looping over however many the device-side counter says
__global__ void __launch_bounds__(256)
vote_big_kernel(const uint4 *__restrict__ fields, const uint8_t *__restrict__ payload, uint32_t stride, uint32_t lpad,
                const uint32_t *__restrict__ perm, const uint32_t *__restrict__ vstart,
                const uint32_t *__restrict__ vsize, const uint32_t *__restrict__ sscs_list,
                const uint32_t *__restrict__ big_list, const uint32_t *__restrict__ n_big, double cutoff,
                uint8_t *__restrict__ out_rows, uint8_t *__restrict__ fam_mapq, uint16_t *__restrict__ fam_flag,
                int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg, uint32_t *__restrict__ err_flag)
{
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t total = *n_big;
    for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < total; i += warps) {
        const uint32_t row = big_list[i];
        vote_family_warp(fields, payload, stride, lpad, perm, vstart[row], vsize[row], cutoff,
                         out_rows + (uint64_t)row * stride, sscs_list[row], fam_mapq, fam_flag, fam_tlen, fam_rg,
                         err_flag);
    }
}

}  // namespace ccgpu
