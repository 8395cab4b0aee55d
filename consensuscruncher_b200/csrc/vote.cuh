// SSCS vote for families of 2..127 reads: sub-warp lane groups, rows staged through a lane-private
// cp.async ring in shared memory.
//
// consensus_maker (SSCS_maker.py:103-155) + read_mode / consensus_flag
// (consensus_helper.py:483-535) for the bulk of the families.  A group of G lanes owns one
// family; a lane owns 16 consecutive positions of it (8 sequence bytes + 16 quality bytes
// per read, one 8-byte and one 16-byte asynchronous copy), so a warp votes 32/G families at once
// and keeps 32/G times more independent copies in flight than one family per warp.
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

#ifndef VOTE_DEPTH_CFG
#define VOTE_DEPTH_CFG 2
#endif
#ifndef VOTE_BLOCKS_CFG
#define VOTE_BLOCKS_CFG 3
#endif

namespace ccgpu {

constexpr uint32_t VOTE_SMALL_MAX = 127;     // byte-lane compares below need counts < 128
#ifndef VOTE_RB_CFG
#define VOTE_RB_CFG 4
#endif
constexpr int VOTE_ROWS_PER_BATCH = VOTE_RB_CFG;   // rows per staged batch
static_assert(12 % VOTE_ROWS_PER_BATCH == 0, "nibble counters are flushed every 12 reads");

__device__ __forceinline__ void cp_async_16_sh(uint32_t smem, const void *gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem), "l"(gmem) : "memory");
}

__device__ __forceinline__ void cp_async_8_sh(uint32_t smem, const void *gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem), "l"(gmem) : "memory");
}

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }

template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ uint32_t pass7_of(uint32_t q)
{
    return (((q & 0x7F7F7F7Fu) + 0x62626262u) | q) & 0x80808080u;    // bit 7 of each byte: q >= 30
}

// 0xFF in every byte whose bit 7 is set (PRMT sign-replicate mode), 0x00 elsewhere
__device__ __forceinline__ uint32_t bytes_of(uint32_t bit7)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, %1, 0xBA98;" : "=r"(r) : "r"(bit7));   // selector bit 3 = replicate the byte's sign
    return r;
}

// bit 7 of each byte: byte != 0, for bytes < 128
__device__ __forceinline__ uint32_t nz7(uint32_t x) { return (x + 0x7F7F7F7Fu) & 0x80808080u; }

// bit 7 of each byte: a >= b, resp. byte != 0, for bytes < 128; bits 0-6 are garbage, which is enough
// when the result only feeds bytes_of().  (a + 0x80 == a | 0x80 for such bytes; written as a sum so
// that the compare is one three-input add.)
__device__ __forceinline__ uint32_t ge7u(uint32_t a, uint32_t b) { return a + 0x80808080u - b; }
__device__ __forceinline__ uint32_t nz7u(uint32_t x) { return x + 0x7F7F7F7Fu; }

__device__ __forceinline__ uint32_t sel(uint32_t a, uint32_t b, uint32_t mask) { return (a & mask) | (b & ~mask); }

// Experiment kept behind a switch (profiles/r2_vote_sass.md): the ALU pipe (LOP3 / PRMT / IADD3 / SHF) is 78 %
// busy while the FMA pipe idles at 25 %, so VOTE_FMA_CFG >= 1 writes the additions that feed the byte-lane
// compares - and >= 2 the constant right shifts too - as integer multiply-adds with a run-time 1 / -1 / power
// of two (same value mod 2^32, issued on the FMA pipe).  Measured slower (1.031 / 1.054 ms against 1.013):
// a + C - b is ONE three-input add but TWO multiply-adds, and the kernel is bound by issue slots, not by a pipe.
#ifndef VOTE_FMA_CFG
#define VOTE_FMA_CFG 0
#endif

// x + y as an integer multiply-add: with `one` a run-time 1 the compiler cannot fold it back
// into an add, so it issues on the FMA pipe instead of the (saturated) ALU pipe.
__device__ __forceinline__ uint32_t fadd(uint32_t x, uint32_t y, uint32_t one) { return x * one + y; }

// ge7u / nz7u on the FMA pipe (`one` is the run-time 1 of fadd)
__device__ __forceinline__ uint32_t ge7f(uint32_t a, uint32_t b, uint32_t one)
{
#if VOTE_FMA_CFG >= 1
    return fadd(b * (0u - one) + a, 0x80808080u, one);
#else
    return ge7u(a, b);
#endif
}
template <uint32_t B>
__device__ __forceinline__ uint32_t ge7c(uint32_t a, uint32_t one)      // b a compile-time constant: one multiply-add
{
#if VOTE_FMA_CFG >= 1
    return fadd(a, 0x80808080u - B, one);
#else
    return ge7u(a, B);
#endif
}
__device__ __forceinline__ uint32_t nz7f(uint32_t x, uint32_t one)
{
#if VOTE_FMA_CFG >= 1
    return fadd(x, 0x7F7F7F7Fu, one);
#else
    return nz7u(x);
#endif
}
// x >> S for a constant S: the high word of x * 2^(32-S)
template <int S>
__device__ __forceinline__ uint32_t shr(uint32_t x, uint32_t one)
{
#if VOTE_FMA_CFG >= 2
    return __umulhi(x, one << (32 - S));
#else
    return x >> S;
#endif
}

__device__ __forceinline__ uint32_t prmt(uint32_t lo, uint32_t hi, uint32_t selector)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(selector));
    return r;
}

// A quad is 4 consecutive positions = 2 sequence bytes = one 16-bit group of 4 base codes.  That
// group is used directly as a PRMT selector: the payload keeps the even position of a byte in the
// LOW nibble, so selector nibble i, byte lane i and the i-th quality byte are all position i.
struct Quad {
    uint32_t ac;        // per byte lane: count of A in the low nibble, of C in the high nibble
    uint32_t gt;        // per byte lane: count of G plus T in the low nibble, of T in the high nibble
    uint32_t pc;        // number of passing reads
    uint32_t qany;      // OR of the passing qualities: the supporter's quality when one read passes
    uint32_t tA, tC, tG, tT;   // counts flushed out of the nibbles (families of more than 15 reads)
};

// One read, 4 positions.  The base codes select their increments from two byte tables:
//   A (1) -> ac += 0x01, C (2) -> ac += 0x10, G (4) -> gt += 0x01, T (8) -> gt += 0x11
// (T reaches the table through PRMT's sign-replicate mode: entry 0 of the G/T table has its
// top bit set, so selector 8 yields 0xFF, masked to 0x11).  Every other code, N and IUPAC
// included, selects 0: it is counted in pc but in no base, which quad_finish reports.
//
// The pass mask assumes qualities below 128 (SAM allows 0..93): q + 98 then cannot carry into the
// next byte.  The caller ORs the qualities it feeds in; a family that shows a byte >= 128 is voted
// again, exactly, by vote_big_kernel (see `hi` in vote_small_kernel).
__device__ __forceinline__ void quad_add(Quad &a, uint32_t sel16, uint32_t q, uint32_t one)
{
    uint32_t m = bytes_of(fadd(q, 0x62626262u, one));                       // 0xFF where q >= 30
    a.ac = fadd(prmt(0x00100100u, 0u, sel16) & m, a.ac, one);
    a.gt = fadd(prmt(0x00000080u, 0x00000001u, sel16) & m & 0x11111111u, a.gt, one);
    a.pc = m * 0xFEFEFEFFu + a.pc;         // 0xFF bytes -> 0x01 bytes: 0xFEFEFEFF is the inverse of 255 mod 2^32
    a.qany |= q & m;
}

// every 12 reads the nibble counters are emptied into byte counters
__device__ __forceinline__ void quad_flush(Quad &a)
{
    uint32_t t = (a.gt >> 4) & 0x0F0F0F0Fu;
    a.tA += a.ac & 0x0F0F0F0Fu;
    a.tC += (a.ac >> 4) & 0x0F0F0F0Fu;
    a.tT += t;
    a.tG += (a.gt & 0x0F0F0F0Fu) - t;
    a.ac = 0; a.gt = 0;
}

// 4 base codes of a 16-bit group, one per byte lane (position order)
__device__ __forceinline__ uint32_t expand_sel(uint32_t sel16)
{
    return __byte_perm(sel16 & 0x0F0Fu, (sel16 >> 4) & 0x0F0Fu, 0x5140);
}
__device__ __forceinline__ uint32_t expand_sel(uint32_t sel16, uint32_t one)
{
    return __byte_perm(sel16 & 0x0F0Fu, shr<4>(sel16, one) & 0x0F0Fu, 0x5140);
}

__device__ __forceinline__ uint32_t cap60(uint32_t q)
{
    uint32_t lo = q & 0x7F7F7F7Fu;
    uint32_t big = bytes_of(ge7u(lo, 0x3C3C3C3Cu) | q);
    return sel(0x3C3C3C3Cu, lo, big);
}

// min(60, q) for qualities below 128 (vote_small_kernel: families with larger bytes are voted again, exactly)
__device__ __forceinline__ uint32_t cap60_lo(uint32_t q)
{
    return sel(0x3C3C3C3Cu, q, bytes_of(ge7u(q, 0x3C3C3C3Cu)));
}
__device__ __forceinline__ uint32_t cap60_lo(uint32_t q, uint32_t one)
{
    return sel(0x3C3C3C3Cu, q, bytes_of(ge7c<0x3C3C3C3Cu>(q, one)));
}

struct QuadOut {
    uint32_t code;       // winner or N per byte
    uint32_t qual;       // consensus quality per byte (valid unless `single` is set there)
    uint32_t win;        // winner's one-hot code per byte
    uint32_t single;     // 0xFF where the winner has exactly one supporter
};

// Winner, cutoff decision and quality of 4 positions.  thr[p] = smallest count c with
// (double)c / (double)p >= cutoff, built on the host with the same IEEE division Python does.
struct ThrMagic {       // thr[p] == (p * k + r) >> 8 for every p the kernel can meet (found on the host), or use_table
    uint32_t k, r2;      // r2 = r | r << 16
    int use_table;
};

template <bool TOTALS>
__device__ __forceinline__ QuadOut quad_finish(const Quad &a, const uint8_t *thr, bool unanimous_ok, ThrMagic tm,
                                               uint32_t &err, uint32_t one)
{
    QuadOut o;
    const uint32_t tnow = shr<4>(a.gt, one) & 0x0F0F0F0Fu;
    uint32_t cA = a.ac & 0x0F0F0F0Fu, cC = shr<4>(a.ac, one) & 0x0F0F0F0Fu, cT = tnow;
    uint32_t cG = fadd(tnow, a.gt & 0x0F0F0F0Fu, 0u - one);
    if (TOTALS) { cA += a.tA; cC += a.tC; cT += a.tT; cG += a.tG; }
    uint32_t m1 = bytes_of(ge7f(cA, cC, one));             // ties keep the earlier base: A,C,G,T order
    uint32_t m2 = bytes_of(ge7f(cG, cT, one));
    uint32_t bAC = sel(cA, cC, m1), wAC = sel(0x01010101u, 0x02020202u, m1);
    uint32_t bGT = sel(cG, cT, m2), wGT = sel(0x04040404u, 0x08080808u, m2);
    uint32_t m3 = bytes_of(ge7f(bAC, bGT, one));
    uint32_t best = sel(bAC, bGT, m3);
    o.win = sel(wAC, wGT, m3);
    uint32_t sum = fadd(cA, fadd(cC, a.gt & 0x0F0F0F0Fu, one), one);   // the G+T nibble already holds cG + cT
    if (TOTALS) sum += a.tG + a.tT;
    err |= sum ^ a.pc;                                     // a passing base that is none of A,C,G,T (N, IUPAC, '=')
    uint32_t pass = a.pc;
    uint32_t ok;
    if (!tm.use_table) {
        // threshold of every position at once: two 16-bit lanes per multiply, no table, no divergence
        uint32_t te = shr<8>((pass & 0x00FF00FFu) * tm.k + tm.r2, one) & 0x00FF00FFu;
        uint32_t to = ((shr<8>(pass, one) & 0x00FF00FFu) * tm.k + tm.r2) & 0xFF00FF00u;
        ok = bytes_of(ge7f(best, te | to, one)) & bytes_of(nz7f(best, one));
    } else {
        uint32_t split = nz7(best ^ pass);                 // not unanimous
        ok = bytes_of(nz7f(best, one)) & ~bytes_of(split);  // unanimous and at least one passing read
        if (!unanimous_ok) ok = 0;
        while (split) {                                    // look the threshold up per position
            int k = (__ffs(split) - 1) >> 3;
            uint32_t c = (best >> (8 * k)) & 0xFFu, p = (pass >> (8 * k)) & 0xFFu;
            if (c >= thr[p]) ok |= 0xFFu << (8 * k);
            split &= ~(0x80u << (8 * k));
        }
    }
    o.code = sel(o.win, 0x0F0F0F0Fu, ok);
    uint32_t two = bytes_of(ge7c<0x02020202u>(best, one));
    uint32_t lone = bytes_of(nz7f(best, one)) & ~two;       // winner has exactly one supporter
    uint32_t one_pass = lone & ~bytes_of(ge7c<0x02020202u>(pass, one));   // ... and it is the only passing read
    o.single = lone & ~one_pass;                           // several passing reads, all different: needs a re-walk
    o.qual = (two & 0x3C3C3C3Cu) | (cap60_lo(a.qany, one) & one_pass);    // 60 where two or more passing reads agree
    return o;
}

// second walk: quality of the only read supporting the winner
__device__ __forceinline__ uint32_t quad_pick(uint32_t nib, uint32_t q, uint32_t win, uint32_t single)
{
    uint32_t b = nib & bytes_of(pass7_of(q));
    uint32_t same = ~bytes_of(nz7u(b ^ win)) & single;      // win is never 0, so `same` implies passing
    return q & same;
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

// quads of a lane: positions 0-3, 4-7, 8-11, 12-15 (see Quad)
__device__ __forceinline__ void split_row(const LaneRow &r, uint32_t sel[4], uint32_t q[4])
{
    sel[0] = r.s.x; sel[1] = r.s.x >> 16; sel[2] = r.s.y; sel[3] = r.s.y >> 16;     // PRMT reads the low 16 bits
    q[0] = r.q.x; q[1] = r.q.y; q[2] = r.q.z; q[3] = r.q.w;
}
__device__ __forceinline__ void split_row(const LaneRow &r, uint32_t sel[4], uint32_t q[4], uint32_t one)
{
    sel[0] = r.s.x; sel[1] = shr<16>(r.s.x, one); sel[2] = r.s.y; sel[3] = shr<16>(r.s.y, one);
    q[0] = r.q.x; q[1] = r.q.y; q[2] = r.q.z; q[3] = r.q.w;
}

// consensus codes / qualities of the 4 quads -> the lane's 8 sequence bytes and 16 quality bytes
__device__ __forceinline__ void join_row(const uint32_t code[4], const uint32_t qual[4], uint2 &so, uint4 &qo)
{
    uint32_t x0 = code[0] | (code[0] >> 4), x1 = code[1] | (code[1] >> 4);
    uint32_t x2 = code[2] | (code[2] >> 4), x3 = code[3] | (code[3] >> 4);
    so.x = __byte_perm(x0, x1, 0x6420);
    so.y = __byte_perm(x2, x3, 0x6420);
    qo.x = qual[0]; qo.y = qual[1]; qo.z = qual[2]; qo.w = qual[3];
}

// mode of one field over a family held by a lane group (slow path: fields differ inside the family)
// (a tie goes to the value whose first occurrence has the smallest read index - idx(j) - so the member
// list may be in any order)
template <typename Get, typename Idx>
__device__ uint32_t group_mode(Get get, Idx idx, uint32_t n, bool is_flag, int gl, int G, uint32_t gmask)
{
    uint32_t best = 0, best_cnt = 0, counted = 0, best_first = 0xFFFFFFFFu;
    int best_rank = 9;
    for (uint32_t i = 0; counted < n && i < n; i++) {
        uint32_t vi = get(i), cnt = 0, first = 0xFFFFFFFFu;
        bool seen_before = false;
        for (uint32_t j0 = 0; j0 < n; j0 += G) {
            uint32_t j = j0 + gl;
            bool m = j < n && get(j) == vi;
            uint32_t b = __ballot_sync(gmask, m) & gmask;
            if (cnt == 0 && b && j0 + (uint32_t)(__ffs(b) - 1 - (__ffs(gmask) - 1)) < i) { seen_before = true; break; }
            cnt += __popc(b);
            if (m) first = min(first, idx(j));
        }
        if (seen_before) continue;
        for (int d = G >> 1; d >= 1; d >>= 1) first = min(first, __shfl_xor_sync(gmask, first, d));
        counted += cnt;
        int rank = is_flag ? flag_rank(vi) : 4;
        if (cnt > best_cnt || (cnt == best_cnt && (rank < best_rank || (rank == best_rank && first < best_first)))) {
            best = vi; best_cnt = cnt; best_rank = rank; best_first = first;
        }
    }
    return best;
}

// Two-read family, 4 positions, in closed form (the commonest family after singletons):
//   both pass and agree -> base, 60;  one passes -> its base and min(60, q);
//   both pass, differ   -> tie: the earlier base in A,C,G,T order with its own quality,
//                          emitted iff 1/2 >= cutoff;  none passes -> N, 0.
__device__ __forceinline__ void quad_pair(uint32_t nib1, uint32_t q1, uint32_t nib2, uint32_t q2, bool tie_ok,
                                          bool unanimous_ok, uint32_t one, uint32_t &code, uint32_t &qual, uint32_t &err)
{
    uint32_t m1 = bytes_of(fadd(q1, 0x62626262u, one)), m2 = bytes_of(fadd(q2, 0x62626262u, one));   // q < 128, as in quad_add
    uint32_t b1 = nib1 & m1, b2 = nib2 & m2;
    err |= (((b1 | 0x80808080u) - 0x01010101u) & b1) | (((b2 | 0x80808080u) - 0x01010101u) & b2);
    err |= (m1 & ~bytes_of(nz7f(b1, one))) | (m2 & ~bytes_of(nz7f(b2, one)));      // passing base that is none of A,C,G,T
    uint32_t both = m1 & m2, differ = bytes_of(nz7f(b1 ^ b2, one));
    uint32_t agree = both & ~differ, clash = both & differ;
    uint32_t first = bytes_of(ge7f(b2, b1, one));                         // read 1 holds the earlier base
    uint32_t win = sel(b1, b2, (m1 & ~m2) | (clash & first) | agree);
    uint32_t wq = sel(q1, q2, (m1 & ~m2) | (clash & first));
    uint32_t ok = (agree | (m1 ^ m2)) & (unanimous_ok ? 0xFFFFFFFFu : 0u);
    if (tie_ok) ok |= clash;
    code = sel(win, 0x0F0F0F0Fu, ok);
    qual = sel(0x3C3C3C3Cu, cap60_lo(wq, one), agree) & (m1 | m2);
}

// vsize[row] = family size | flags
constexpr uint32_t VSIZE_MIXED = 0x80000000u;     // members' mode fields differ (set by vote_segments_kernel)
constexpr uint32_t VSIZE_QUEUED = 0x40000000u;    // row already queued for the exact kernel
constexpr uint32_t VSIZE_COUNT = 0x3FFFFFFFu;

constexpr int VOTE_CHUNK = 1024;    // SSCS rows a block sorts by family size and then votes

// dynamic shared memory of vote_small_kernel<G, CH, DEPTH>: the lanes' staging rings
constexpr int vote_ring_bytes(int ch, int depth) { return depth * VOTE_ROWS_PER_BATCH * ch * 256 * 24; }

// vstart / vsize: segment of `perm` of every SSCS row (filled by vote_segments_kernel).
// A block takes VOTE_CHUNK consecutive rows, orders the families of n_lo..n_hi reads by size in
// shared memory and hands them to its lane groups in that order, so the groups of a warp walk
// families of (nearly) equal size and their row loops stay in step.
//   G      lanes per family           CH     16-position chunks per lane (G * CH * 16 >= lpad)
//   DEPTH  staging ring slots         FLUSH  families may exceed 15 reads (nibble counters are
//                                            emptied into byte counters every 12 reads)
template <int G, int CH, int DEPTH, bool FLUSH>
__global__ void __launch_bounds__(256, VOTE_BLOCKS_CFG)
vote_small_kernel(const uint8_t *__restrict__ payload, uint32_t stride, uint32_t lpad,
                  const uint32_t *__restrict__ perm, const uint32_t *__restrict__ vstart,
                  uint32_t *vsize, const uint32_t *__restrict__ sscs_list,
                  const uint32_t *__restrict__ row_list, uint32_t n_rows, uint32_t n_lo, uint32_t n_hi,
                  const uint8_t *__restrict__ thr_table, int unanimous_ok, ThrMagic tm,
                  const uint4 *__restrict__ fields, uint8_t *__restrict__ out_rows, uint8_t *__restrict__ fam_mapq,
                  uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
                  uint32_t *__restrict__ err_flag, uint32_t *__restrict__ redo_list, uint32_t *__restrict__ n_redo)
{
    __shared__ uint8_t thr[128];
    __shared__ uint32_t s_start[VOTE_CHUNK];
    __shared__ uint8_t s_size[VOTE_CHUNK];
    __shared__ uint16_t s_order[VOTE_CHUNK];
    __shared__ uint32_t s_row[VOTE_CHUNK];      // SSCS row of every entry (row_list given: only listed rows are voted)
    __shared__ uint32_t s_bin[128];
    constexpr int GROUPS = 32 / G;
    constexpr int BLOCK_GROUPS = 8 * GROUPS;
    constexpr int RB = VOTE_ROWS_PER_BATCH;
    const int tid = threadIdx.x, lane = tid & 31, gl = lane % G, grp = lane / G;
    const uint32_t gmask = (G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u)) << (grp * G);
    const uint32_t half = lpad >> 1;
    const uint32_t row0 = blockIdx.x * VOTE_CHUNK;
    const uint32_t rows_here = min((uint32_t)VOTE_CHUNK, n_rows - row0);

    if (tid < 128) { thr[tid] = thr_table[tid]; s_bin[tid] = 0; }
    __syncthreads();
    for (uint32_t i = tid; i < rows_here; i += 256) {
        const uint32_t r = row_list ? __ldg(&row_list[row0 + i]) : row0 + i;
        const uint32_t v = vsize[r];
        uint32_t n = v & VSIZE_COUNT;
        if (n < n_lo || n > n_hi) n = 0;                   // voted by another launch (other sizes)
        s_row[i] = r;
        s_start[i] = __ldg(&vstart[r]);
        s_size[i] = (uint8_t)(n | ((v >> 31) << 7));       // bit 7: mode fields differ inside the family
        atomicAdd(&s_bin[n], 1u);
    }
    __syncthreads();
    if (tid < 32) {                                        // exclusive prefix over the 128 size bins
        uint32_t v[4], sum = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) { v[k] = s_bin[tid * 4 + k]; sum += v[k]; }
        uint32_t incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
            if (tid >= d) incl += t;
        }
        uint32_t run = incl - sum;
#pragma unroll
        for (int k = 0; k < 4; k++) { s_bin[tid * 4 + k] = run; run += v[k]; }
    }
    __syncthreads();
    const uint32_t skipped = s_bin[1];                     // bin 0 = rows this launch does not vote
    __syncthreads();
    if (skipped == rows_here) return;
    for (uint32_t i = tid; i < rows_here; i += 256) s_order[atomicAdd(&s_bin[s_size[i] & 0x7F], 1u)] = (uint16_t)i;
    __syncthreads();

    const bool tie_ok = 1 >= thr[2];
    uint32_t err = 0;
    const uint32_t one = (uint32_t)unanimous_ok | 1u;      // 1, but only known at run time (see fadd)
    const uint32_t todo = rows_here - skipped;
    const uint32_t my_slot = (tid >> 5) * GROUPS + grp;
    const uint32_t pos0 = gl * (16 * CH);                  // this lane owns CH chunks of 16 positions
    const bool lane_on = pos0 < lpad;

    // Lane-private staging ring in shared memory, filled with cp.async: every lane copies its
    // own 8 sequence bytes + 16 quality bytes per chunk of up to RB rows per batch and keeps
    // DEPTH - 1 batches in flight while it votes the oldest one.  Bytes in flight do not
    // occupy registers.
    extern __shared__ __align__(16) uint8_t ring[];
    constexpr uint32_t CH_Q = 256 * 16, CH_S = 256 * 8;                         // bytes between chunks of a row
    constexpr uint32_t ROW_Q = CH * CH_Q, ROW_S = CH * CH_S;                    // ... between rows of a slot
    constexpr uint32_t SLOT_Q = RB * ROW_Q, SLOT_S = RB * ROW_S;
    uint8_t *const ring_q = ring + tid * 16;                                     // [depth][row][chunk][thread] uint4
    uint8_t *const ring_s = ring + DEPTH * SLOT_Q + tid * 8;                     // [depth][row][chunk][thread] uint2
    const uint32_t ring_q_sh = (uint32_t)__cvta_generic_to_shared(ring_q);
    const uint32_t ring_s_sh = (uint32_t)__cvta_generic_to_shared(ring_s);
    // this lane's first sequence / quality bytes of read 0, kept as opaque 64-bit values so that a
    // row's address is one multiply-add (IMAD.WIDE) instead of offset arithmetic on the kernel argument
    uint64_t lane_src_u = (uint64_t)(payload + (pos0 >> 1)), lane_src_q_u = (uint64_t)(payload + half + pos0);
    asm volatile("" : "+l"(lane_src_u), "+l"(lane_src_q_u));
    const uint8_t *const lane_src = (const uint8_t *)lane_src_u;
    const uint8_t *const lane_src_q = (const uint8_t *)lane_src_q_u;

    // producer cursor: the batch that will be issued next
    uint32_t p_t = my_slot, p_start = 0, p_n = 0, p_j0 = 0, p_slot = 0;
    uint32_t p_idx[RB];
    auto p_load_family = [&]() {
        p_start = 0; p_n = 0; p_j0 = 0;
        if (p_t < todo) {
            uint32_t i = s_order[skipped + p_t];
            p_start = s_start[i];
            p_n = s_size[i] & 0x7F;
        }
    };
    auto p_load_idx = [&]() {                              // member indices of the producer's batch
#pragma unroll
        for (int k = 0; k < RB; k++)
            p_idx[k] = (p_n != 0) ? __ldg(&perm[p_start + min(p_j0 + k, p_n - 1)]) : 0u;
    };
    auto p_issue = [&]() {                                 // copy the producer's batch, then move on
        if (p_n != 0 && lane_on) {
#pragma unroll
            for (int k = 0; k < RB; k++) {
                if (p_j0 + k < p_n) {
                    const uint8_t *src = lane_src + (uint64_t)p_idx[k] * stride;
                    const uint8_t *src_q = lane_src_q + (uint64_t)p_idx[k] * stride;
#pragma unroll
                    for (int h = 0; h < CH; h++) {
                        if (h == 0 || pos0 + 16 * h < lpad) {
                            cp_async_8_sh(ring_s_sh + p_slot * SLOT_S + k * ROW_S + h * CH_S, src + 8 * h);
                            cp_async_16_sh(ring_q_sh + p_slot * SLOT_Q + k * ROW_Q + h * CH_Q, src_q + 16 * h);
                        }
                    }
                }
            }
        }
        cp_async_commit();
        p_slot = p_slot + 1 == DEPTH ? 0 : p_slot + 1;
        p_j0 += RB;
        if (p_j0 >= p_n) { p_t += BLOCK_GROUPS; p_load_family(); }
        p_load_idx();
    };
    p_load_family();
    p_load_idx();
#pragma unroll 1
    for (int d = 0; d < DEPTH - 1; d++) p_issue();

    uint32_t c_slot = 0;                                   // consumer: slot of the oldest batch still in the ring
    auto slot_row = [&](uint32_t slot, int k, int h) {
        return LaneRow{*reinterpret_cast<const uint2 *>(ring_s + slot * SLOT_S + k * ROW_S + h * CH_S),
                       *reinterpret_cast<const uint4 *>(ring_q + slot * SLOT_Q + k * ROW_Q + h * CH_Q)};
    };
    auto next_slot = [&](uint32_t slot) { return slot + 1 == DEPTH ? 0u : slot + 1; };

    for (uint32_t t0 = 0; t0 < todo; t0 += BLOCK_GROUPS) {
        const uint32_t t = t0 + my_slot;
        uint32_t start = 0, n = 0, row = 0;
        bool mixed = false;
        if (t < todo) {
            uint32_t i = s_order[skipped + t];
            start = s_start[i];
            n = s_size[i] & 0x7F;
            mixed = (s_size[i] & 0x80) != 0;
            row = s_row[i];
        }
        uint8_t *out = out_rows + (uint64_t)row * stride;
        CC_CHECK(n == 0 || row_list != nullptr || row < n_rows);       // (with a row list, n_rows counts the list)

        if (n != 0) {
            uint32_t hi = 0, ferr = 0;                     // OR of the family's quality words; its error bits
            if (n == 2) {
                p_issue();
                cp_async_wait<DEPTH - 1>();
                if (lane_on) {
#pragma unroll
                    for (int h = 0; h < CH; h++) {
                        if (h != 0 && pos0 + 16 * h >= lpad) break;
                        LaneRow r1 = slot_row(c_slot, 0, h), r2 = slot_row(c_slot, 1, h);
                        uint32_t n1[4], q1[4], n2[4], q2[4], code[4], qual[4];
                        split_row(r1, n1, q1, one);
                        split_row(r2, n2, q2, one);
                        hi |= q1[0] | q1[1]; hi |= q1[2] | q1[3]; hi |= q2[0] | q2[1]; hi |= q2[2] | q2[3];
#pragma unroll
                        for (int g = 0; g < 4; g++)
                            quad_pair(expand_sel(n1[g], one), q1[g], expand_sel(n2[g], one), q2[g], tie_ok, unanimous_ok != 0, one, code[g], qual[g], ferr);
                        uint2 so;
                        uint4 qo;
                        join_row(code, qual, so, qo);
                        *reinterpret_cast<uint2 *>(out + ((pos0 + 16 * h) >> 1)) = so;
                        *reinterpret_cast<uint4 *>(out + half + pos0 + 16 * h) = qo;
                    }
                }
                c_slot = next_slot(c_slot);
            } else {
                Quad acc[CH][4];
#pragma unroll
                for (int h = 0; h < CH; h++)
#pragma unroll
                    for (int g = 0; g < 4; g++) acc[h][g] = Quad{0, 0, 0, 0, 0, 0, 0, 0};
                const uint32_t first_slot = c_slot;
                for (uint32_t j0 = 0; j0 < n; j0 += RB) {
                    p_issue();
                    cp_async_wait<DEPTH - 1>();
                    if (lane_on) {
#pragma unroll
                        for (int k = 0; k < RB; k++) {
                            if (j0 + k < n) {
#pragma unroll
                                for (int h = 0; h < CH; h++) {
                                    if (h == 0 || pos0 + 16 * h < lpad) {
                                        uint32_t sel4[4], q[4];
                                        split_row(slot_row(c_slot, k, h), sel4, q, one);
                                        hi |= q[0] | q[1]; hi |= q[2] | q[3];
#pragma unroll
                                        for (int g = 0; g < 4; g++) quad_add(acc[h][g], sel4[g], q[g], one);
                                    }
                                }
                            }
                        }
                    }
                    c_slot = next_slot(c_slot);
                    if (FLUSH && (j0 + RB) % 12 == 0 && j0 + RB < n) {
#pragma unroll
                        for (int h = 0; h < CH; h++)
#pragma unroll
                            for (int g = 0; g < 4; g++) quad_flush(acc[h][g]);
                    }
                }
                if (lane_on) {
#pragma unroll
                    for (int h = 0; h < CH; h++) {
                        if (h != 0 && pos0 + 16 * h >= lpad) break;
                        QuadOut o[4];
                        uint32_t any_single = 0;
#pragma unroll
                        for (int g = 0; g < 4; g++) {
                            o[g] = quad_finish<FLUSH>(acc[h][g], thr, unanimous_ok != 0, tm, ferr, one);
                            any_single |= o[g].single;
                        }
                        if (any_single) {                  // lane-local: re-walk the rows for the lone supporters
                            uint32_t pick[4] = {0, 0, 0, 0};
                            const bool in_ring = n <= RB;  // longer families have left the ring: re-read them
                            for (uint32_t j = 0; j < n; j++) {
                                LaneRow rr = in_ring ? slot_row(first_slot, (int)j, h)
                                                     : load_lane_row(payload, stride, half, __ldg(&perm[start + j]), pos0 + 16 * h);
                                uint32_t sel4[4], q[4];
                                split_row(rr, sel4, q);
#pragma unroll
                                for (int g = 0; g < 4; g++) pick[g] |= quad_pick(expand_sel(sel4[g]), q[g], o[g].win, o[g].single);
                            }
#pragma unroll
                            for (int g = 0; g < 4; g++) o[g].qual |= cap60_lo(pick[g]) & o[g].single;
                        }
                        uint32_t code[4], qual[4];
#pragma unroll
                        for (int g = 0; g < 4; g++) { code[g] = o[g].code; qual[g] = o[g].qual; }
                        uint2 so;
                        uint4 qo;
                        join_row(code, qual, so, qo);
                        *reinterpret_cast<uint2 *>(out + ((pos0 + 16 * h) >> 1)) = so;
                        *reinterpret_cast<uint4 *>(out + half + pos0 + 16 * h) = qo;
                    }
                }
            }
            for (uint32_t x = half + lpad + gl * 4; x < stride; x += G * 4) *reinterpret_cast<uint32_t *>(out + x) = 0;
            // A quality byte >= 128 breaks the byte-lane pass test above (and may have raised a false
            // error in this lane): the family goes to vote_big_kernel, which runs after this launch,
            // tests qualities exactly and overwrites the row.
            if (hi & 0x80808080u) {                        // first lane to flag the row queues it
                if (!(atomicOr(&vsize[row], VSIZE_QUEUED) & VSIZE_QUEUED)) redo_list[atomicAdd(n_redo, 1u)] = row;
            } else {
                err |= ferr;
            }
        }

        // ---- mode fields (create_aligned_segment, consensus_helper.py:569-583) -------------------
        // A family whose members all carry the same tlen / flag / mapq / RG keeps the template's
        // values, which family_build_kernel already stored.  Only families flagged by the
        // boundary pass need the real modes; that branch is uniform per lane group.
        if (n != 0 && mixed) {
            auto meta_of = [&](uint32_t j) { return __ldg(&fields[2 * (uint64_t)__ldg(&perm[start + j]) + 1]); };
            bool has_rg = true;
            for (uint32_t j0 = 0; j0 < n; j0 += G) {
                uint32_t j = j0 + gl;
                uint32_t rg_ok = __ballot_sync(gmask, j >= n || (meta_of(j).w >> 24) != 0xFFu) & gmask;
                has_rg = has_rg && rg_ok == gmask;
            }
            auto idx_of = [&](uint32_t j) { return __ldg(&perm[start + j]); };
            uint32_t tlen = group_mode([&](uint32_t j) { return meta_of(j).z; }, idx_of, n, false, gl, G, gmask);
            uint32_t flag = group_mode([&](uint32_t j) { return meta_of(j).w & 0xFFFFu; }, idx_of, n, true, gl, G, gmask);
            uint32_t mapq = group_mode([&](uint32_t j) { return (meta_of(j).w >> 16) & 0xFFu; }, idx_of, n, false, gl, G, gmask);
            uint32_t rg = has_rg ? group_mode([&](uint32_t j) { return meta_of(j).w >> 24; }, idx_of, n, false, gl, G, gmask) : 0xFFu;
            if (gl == 0) {
                const uint32_t c = __ldg(&sscs_list[row]);
                fam_tlen[c] = (int32_t)tlen; fam_flag[c] = (uint16_t)flag;
                fam_mapq[c] = (uint8_t)mapq; fam_rg[c] = (uint8_t)rg;
            }
        }
    }
    cp_async_wait<0>();
    if (err) atomicOr(err_flag, 1u);
}

// segment of every SSCS row, and the lists of rows for the mid-size launch and for vote_big_kernel
__global__ void __launch_bounds__(256)
vote_segments_kernel(const uint32_t *__restrict__ sscs_list, const uint32_t *__restrict__ fam_start,
                     const uint32_t *__restrict__ fam_size, const uint8_t *__restrict__ fam_class, uint32_t n_rows,
                     uint32_t *__restrict__ vstart,
                     uint32_t *__restrict__ vsize, uint32_t *__restrict__ big_list, uint32_t *__restrict__ n_big,
                     uint32_t small_max, uint32_t *__restrict__ mid_list, uint32_t *__restrict__ n_mid)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    uint32_t c = sscs_list[r];
    uint32_t n = fam_size[c];
    vstart[r] = fam_start[c];
    vsize[r] = n | ((uint32_t)(fam_class[c] & 2) << 30);   // bit 31: members' mode fields differ
    if (n > small_max) big_list[atomicAdd(n_big, 1u)] = r;           // one warp per family (vote_big_kernel)
    else if (n > 15) mid_list[atomicAdd(n_mid, 1u)] = r;             // 16..127 reads: the launch with counter flushes
}

// one CTA per large family (its 8 warps share the members), looping over however many the
// device-side counter says
__global__ void __launch_bounds__(256)
vote_big_kernel(const uint4 *__restrict__ fields, const uint8_t *__restrict__ payload, uint32_t stride, uint32_t lpad,
                const uint32_t *__restrict__ perm, const uint32_t *__restrict__ vstart,
                const uint32_t *__restrict__ vsize, const uint32_t *__restrict__ sscs_list,
                const uint32_t *__restrict__ big_list, const uint32_t *__restrict__ n_big, double cutoff,
                uint8_t *__restrict__ out_rows, uint8_t *__restrict__ fam_mapq, uint16_t *__restrict__ fam_flag,
                int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg, uint32_t *__restrict__ err_flag)
{
    __shared__ uint32_t scratch[8 * 32 * VOTE_PART_WORDS];
    const uint32_t total = *n_big;
    for (uint32_t i = blockIdx.x; i < total; i += gridDim.x) {
        const uint32_t row = big_list[i];
        const uint32_t size = vsize[row];
        vote_family_warp<8>(fields, payload, stride, lpad, perm, vstart[row], size & VSIZE_COUNT, cutoff,
                            out_rows + (uint64_t)row * stride, sscs_list[row], fam_mapq, fam_flag, fam_tlen, fam_rg,
                            err_flag, scratch, (size & VSIZE_MIXED) != 0);
    }
}


// ---- K7: duplex_consensus on two rows, 16 positions per thread --------------------------------
// DCS_maker.py:107-119: equal bases -> base, min(60, q1+q2); else N, 0 (no quality filter).
// singleton_correction.py:69-82: additionally both qualities must be > 29 - and then the sum is
// at least 60, so the quality is 60 wherever the base survives.
template <bool SC>
__device__ __forceinline__ void quad_agree(uint32_t n1, uint32_t q1, uint32_t n2, uint32_t q2, uint32_t &code, uint32_t &qual)
{
    uint32_t eq = ~bytes_of(nz7u(n1 ^ n2));
    if (SC) {
        eq &= bytes_of(pass7_of(q1)) & bytes_of(pass7_of(q2));
        qual = eq & 0x3C3C3C3Cu;
    } else {
        qual = eq & cap60(cap60(q1) + cap60(q2));          // min(60, q1+q2) == min(60, min(60,q1) + min(60,q2))
    }
    code = sel(n1, 0x0F0F0F0Fu, eq);
}

template <bool SC>
__device__ __forceinline__ void agree_lane(const uint8_t *__restrict__ r1, const uint8_t *__restrict__ r2,
                                           uint8_t *__restrict__ out, uint32_t lpad, uint32_t stride, uint32_t lane,
                                           uint32_t lanes)
{
    const uint32_t half = lpad >> 1, pos0 = lane * 16;
    if (pos0 < lpad) {
        LaneRow a{__ldg(reinterpret_cast<const uint2 *>(r1 + (pos0 >> 1))), __ldg(reinterpret_cast<const uint4 *>(r1 + half + pos0))};
        LaneRow b{__ldg(reinterpret_cast<const uint2 *>(r2 + (pos0 >> 1))), __ldg(reinterpret_cast<const uint4 *>(r2 + half + pos0))};
        uint32_t n1[4], q1[4], n2[4], q2[4], code[4], qual[4];
        split_row(a, n1, q1);
        split_row(b, n2, q2);
#pragma unroll
        for (int g = 0; g < 4; g++) quad_agree<SC>(expand_sel(n1[g]), q1[g], expand_sel(n2[g]), q2[g], code[g], qual[g]);
        uint2 so;
        uint4 qo;
        join_row(code, qual, so, qo);
        *reinterpret_cast<uint2 *>(out + (pos0 >> 1)) = so;
        *reinterpret_cast<uint4 *>(out + half + pos0) = qo;
    }
    for (uint32_t x = half + lpad + lane * 4; x < stride; x += lanes * 4) *reinterpret_cast<uint32_t *>(out + x) = 0;
}

// threads of the row kernels below take DUPLEX_CHUNKS chunks of 16 positions each (`lanes` = threads per row);
// measured on config 2: 1 chunk 0.242 ms, 2 chunks 0.197 ms (5.85 TB/s, 89 % of the copy peak), 4 chunks 0.284 ms
#ifndef DUPLEX_CHUNKS_CFG
#define DUPLEX_CHUNKS_CFG 2
#endif
constexpr uint32_t DUPLEX_CHUNKS = DUPLEX_CHUNKS_CFG;

template <bool SC>
__device__ __forceinline__ void agree_chunks(const uint8_t *__restrict__ r1, const uint8_t *__restrict__ r2, uint8_t *__restrict__ out,
                                             uint32_t lpad, uint32_t stride, uint32_t lane)
{
    const uint32_t chunks = (lpad + 15) / 16;
#pragma unroll
    for (uint32_t h = 0; h < DUPLEX_CHUNKS; h++)
        if (DUPLEX_CHUNKS * lane + h < chunks) agree_lane<SC>(r1, r2, out, lpad, stride, DUPLEX_CHUNKS * lane + h, chunks);
}

// singleton rescue: row r <-> singleton family corr_list[r]; partner SSCS consensus if the
// partner family has one, else the partner singleton's original read
__global__ void __launch_bounds__(256)
rescue16_kernel(RowSources src, const uint32_t *__restrict__ corr_list, const int32_t *__restrict__ partner,
                uint32_t n_rows, uint32_t lpad, uint32_t lanes, uint8_t *__restrict__ out_rows)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t row = (uint32_t)(t / lanes), lane = (uint32_t)(t % lanes);
    if (row >= n_rows) return;
    uint32_t c = corr_list[row];
    uint32_t p = (uint32_t)partner[c];
    CC_CHECK((int32_t)p >= 0 && (src.fam_size[p] < 2 || src.sscs_row[p] >= 0));
    const uint8_t *mine = src.payload + (uint64_t)src.tmpl[c] * src.stride;
    const uint8_t *other = src.fam_size[p] >= 2 ? src.sscs_rows + (uint64_t)src.sscs_row[p] * src.stride
                                                : src.payload + (uint64_t)src.tmpl[p] * src.stride;
    agree_chunks<true>(mine, other, out_rows + (uint64_t)row * src.stride, lpad, src.stride, lane);
}

// DCS (after_sc = 0: both rows are SSCS consensus) or DCS over SSCS + corrected singletons
__global__ void __launch_bounds__(256)
duplex16_kernel(RowSources src, const uint32_t *__restrict__ lead_list, const int32_t *__restrict__ partner,
                uint32_t n_rows, uint32_t lpad, uint32_t lanes, int after_sc, uint8_t *__restrict__ out_rows,
                const uint8_t *__restrict__ fam_mapq, const uint16_t *__restrict__ fam_flag,
                const int32_t *__restrict__ fam_tlen, const uint8_t *__restrict__ fam_rg,
                uint8_t *__restrict__ o_mapq, uint16_t *__restrict__ o_flag, int32_t *__restrict__ o_tlen,
                uint8_t *__restrict__ o_rg)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t row = (uint32_t)(t / lanes), lane = (uint32_t)(t % lanes);
    if (row >= n_rows) return;
    uint32_t c = lead_list[row];
    uint32_t p = (uint32_t)partner[c];
    CC_CHECK((int32_t)p >= 0);
    auto row_of = [&](uint32_t x) -> const uint8_t * {
        if (!after_sc || src.fam_size[x] >= 2) return src.sscs_rows + (uint64_t)src.sscs_row[x] * src.stride;
        return src.corr_rows + (uint64_t)src.corr_row[x] * src.stride;
    };
    agree_chunks<false>(row_of(c), row_of(p), out_rows + (uint64_t)row * src.stride, lpad, src.stride, lane);
    if (lane == 0) {
        o_mapq[row] = (uint8_t)pair_mode(fam_mapq[c], fam_mapq[p], false);
        o_tlen[row] = (int32_t)pair_mode((uint32_t)fam_tlen[c], (uint32_t)fam_tlen[p], false);
        o_flag[row] = (uint16_t)pair_mode(fam_flag[c], fam_flag[p], true);
        uint32_t ra = fam_rg[c], rb = fam_rg[p];
        o_rg[row] = (ra == 0xFFu || rb == 0xFFu) ? 0xFFu : (uint8_t)pair_mode(ra, rb, false);
    }
}

// stand-alone duplex_consensus over independent row pairs (cc_duplex_consensus)
template <bool SC>
__global__ void __launch_bounds__(256)
agree_pairs16_kernel(const uint8_t *__restrict__ rows1, const uint8_t *__restrict__ rows2, uint64_t n_pairs,
                     uint32_t lpad, uint32_t stride, uint32_t lanes, uint8_t *__restrict__ out_rows)
{
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t row = t / lanes;
    uint32_t lane = (uint32_t)(t % lanes);
    if (row >= n_pairs) return;
    agree_lane<SC>(rows1 + row * stride, rows2 + row * stride, out_rows + row * stride, lpad, stride, lane, lanes);
}

}  // namespace ccgpu
