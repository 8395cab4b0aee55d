// Grouping reads into tag families, tile by tile (sm_100a) - the default grouping path.
//
// read_bam's dictionaries (consensus_helper.py:447-475: `tag in read_dict`, append, size + 1) keyed
// by the exact family tuple, as in group_hash.cuh, but the global table is visited once per
// (tile, family) instead of once per read.  Reads of a family are neighbours in a coordinate-sorted
// input, so a CTA first groups its tile of TILE consecutive reads in shared memory:
//
//   tile_insert   every read finds its tile-local representative (the smallest index of its tuple in the
//                 tile) through a shared-memory table; a representative counts its members and whether
//                 their mode fields differ, then makes ONE trip to the global table for all of them
//                 (claim / lower the slot's read index, subtract 2 x members from the slot's counter).
//                 Per read one word is kept: representative and rank inside the tile's group, or -
//                 for a representative - the global slot.
//   tile_number   (after every tile has inserted) a representative whose slot still names it founds the
//                 family.  One chained scan (decoupled look-back) over the tiles carries two sums - founders
//                 and their family sizes - so the same kernel numbers the families in creation order AND
//                 places their segments: template, size, segment start, class, singleton fields, size
//                 histogram, the bitmap of founders with its word ranks.
//   tile_place    a representative looks its family up (rank query at the founder's bit), reserves
//                 room for its tile's members in the family's segment with one atomic, and the
//                 members write their places.
//
// Members land in `perm` in arbitrary order inside a family; nothing downstream depends on it
// (votes count, lone supporters are unique, mode ties are decided by read index).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "group_hash.cuh"
#include "kernels.cuh"
#include "scan.cuh"

namespace ccgpu {

constexpr int GT_THREADS = 256;
constexpr int GT_PER = 4;                       // reads per thread
constexpr int GT_TILE = GT_THREADS * GT_PER;    // reads per tile
constexpr uint32_t GT_HT = 2 * GT_TILE;         // shared table slots
constexpr uint32_t GT_EMPTY = 0xFFFFFFFFu;
constexpr uint32_t LINK_REP = 0x80000000u;      // link word of a tile representative: flag | global slot
constexpr uint32_t GT_MIXED = 0x80000000u;      // shared counter: members' mode fields differ
constexpr int LINK_RANK_SHIFT = 10;             // link word of a member: representative (10 bits) | rank << 10
static_assert(GT_TILE == 1 << LINK_RANK_SHIFT, "a member's link word holds two tile-local indices");

inline uint64_t gt_tiles(uint64_t n) { return (n + GT_TILE - 1) / GT_TILE; }

__global__ void __launch_bounds__(GT_THREADS)
tile_insert_kernel(const uint4 *__restrict__ fields, uint64_t n, GroupSlot *__restrict__ table, uint64_t mask,
                   uint32_t *__restrict__ link)
{
    __shared__ uint4 key_a[GT_TILE], key_b[GT_TILE];      // the tile's records
    __shared__ uint32_t ht[GT_HT];                        // tuple -> smallest local index seen so far
    __shared__ uint32_t members[GT_TILE];                 // per representative: members besides itself | GT_MIXED
    __shared__ uint16_t rep_list[GT_TILE];               // (static shared memory stays under 48 KB)
    __shared__ uint32_t n_reps;
    const int tid = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * GT_TILE;
    const uint32_t here = (uint32_t)min((uint64_t)GT_TILE, n - base);

    for (int k = tid; k < (int)GT_HT; k += GT_THREADS) ht[k] = GT_EMPTY;
    if (tid == 0) n_reps = 0;
    uint64_t h[GT_PER];
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        members[li] = 0;
        if (li < here) {
            const Rec r = load_rec(fields, (uint32_t)(base + li));
            key_a[li] = r.a;
            key_b[li] = r.b;
            h[k] = rec_hash(r);
        }
    }
    __syncthreads();

    // tile-local table: a slot ends up holding the smallest local index of its tuple
    uint32_t my_slot[GT_PER];
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li >= here) continue;
        const Rec r{key_a[li], key_b[li]};
        uint32_t s = (uint32_t)(h[k] >> 7) & (GT_HT - 1);
        for (;;) {
            uint32_t cur = ht[s];
            if (cur == GT_EMPTY) {
                cur = atomicCAS(&ht[s], GT_EMPTY, li);
                if (cur == GT_EMPTY) break;
            }
            if (same_tag(r, Rec{key_a[cur], key_b[cur]})) {
                if (li < cur) atomicMin(&ht[s], li);
                break;
            }
            s = (s + 1) & (GT_HT - 1);
        }
        my_slot[k] = s;
    }
    __syncthreads();

    // members count themselves at their representative; the rank is their place in the tile's group
    uint32_t lrep[GT_PER], rank[GT_PER];
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li >= here) continue;
        lrep[k] = ht[my_slot[k]];
        rank[k] = 0;
        if (lrep[k] != li) {
            rank[k] = 1 + (atomicAdd(&members[lrep[k]], 1u) & ~GT_MIXED);
            // every member is compared with its representative: no difference anywhere = uniform mode fields
            if (key_b[li].z != key_b[lrep[k]].z || key_b[li].w != key_b[lrep[k]].w) atomicOr(&members[lrep[k]], GT_MIXED);
        } else {
            rep_list[atomicAdd(&n_reps, 1u)] = (uint16_t)li;
        }
    }
    __syncthreads();

    // one trip to the global table per representative, all threads busy
    const uint32_t reps = n_reps;
    for (uint32_t x = tid; x < reps; x += GT_THREADS) {
        const uint32_t li = rep_list[x];
        const uint32_t i = (uint32_t)(base + li);
        const Rec r{key_a[li], key_b[li]};
        const uint32_t m = members[li];
        bool mixed = (m & GT_MIXED) != 0;
        uint64_t slot = (rec_hash(r) * 0xD6E8FEB86659FD93ull >> 20) & mask;
        for (;;) {
            uint32_t rep = ld_relaxed_u32(&table[slot].rep);
            if (rep == 0xFFFFFFFFu) {
                rep = atomicCAS(&table[slot].rep, 0xFFFFFFFFu, i);
                if (rep == 0xFFFFFFFFu) break;                  // claimed: so far this tile founds the family
            }
            const Rec q = load_rec(fields, rep);
            if (same_tag(r, q)) {                                // another tile's representative of the same family
                if (i < rep) atomicMin(&table[slot].rep, i);
                mixed = mixed || r.b.z != q.b.z || r.b.w != q.b.w;   // the comparisons span the family (group_hash.cuh)
                break;
            }
            slot = (slot + 1) & mask;
        }
        atomicSub(&table[slot].cnt, 2u * ((m & ~GT_MIXED) + 1u));
        if (mixed) atomicAnd(&table[slot].cnt, ~1u);
        link[i] = LINK_REP | (uint32_t)slot;
    }
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li < here && lrep[k] != li) link[base + li] = lrep[k] | (rank[k] << LINK_RANK_SHIFT);
    }
}

// status word of the chained scan: state (2 bits) | founders (31 bits) | reads in their families (31 bits)
__device__ __forceinline__ unsigned long long gt_pack(uint32_t state, uint32_t fams, uint32_t reads)
{
    return ((unsigned long long)state << 62) | ((unsigned long long)fams << 31) | reads;
}

__global__ void __launch_bounds__(GT_THREADS)
tile_number_kernel(const uint4 *__restrict__ fields, uint64_t n, const GroupSlot *__restrict__ table,
                   const uint32_t *__restrict__ link, uint32_t *__restrict__ final_rep,
                   unsigned long long *status, uint32_t *ticket,
                   uint32_t *__restrict__ bitmap, uint32_t *__restrict__ word_rank,
                   uint32_t *__restrict__ tmpl, uint32_t *__restrict__ fam_size, uint32_t *__restrict__ fam_start,
                   uint32_t *__restrict__ cursor, uint8_t *__restrict__ is_sscs, uint8_t *__restrict__ fam_mapq,
                   uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
                   int32_t *__restrict__ partner, unsigned long long *__restrict__ hist, uint32_t hist_bins,
                   uint32_t *__restrict__ n_mid, uint32_t *__restrict__ n_fam_out)
{
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp_f[GT_THREADS / 32], s_warp_r[GT_THREADS / 32];
    __shared__ uint32_t s_excl_f, s_excl_r;
    __shared__ uint32_t small_hist[256];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    small_hist[tid] = 0;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);        // tiles are numbered in the order they start: a tile only waits on started ones
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t base = (uint64_t)tile * GT_TILE + (uint64_t)tid * GT_PER;   // a thread owns 4 consecutive reads

    uint32_t lk[GT_PER] = {0, 0, 0, 0};
    if (base + GT_PER <= n) {
        const uint4 q = *reinterpret_cast<const uint4 *>(link + base);
        lk[0] = q.x; lk[1] = q.y; lk[2] = q.z; lk[3] = q.w;
    } else {
#pragma unroll
        for (int k = 0; k < GT_PER; k++) lk[k] = base + k < n ? link[base + k] : 0u;
    }
    uint32_t size[GT_PER], cnt[GT_PER];
    uint32_t founders = 0, reads = 0, bits = 0;
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        size[k] = 0;
        if (base + k < n && (lk[k] & LINK_REP)) {
            const uint2 s = *reinterpret_cast<const uint2 *>(&table[lk[k] & ~LINK_REP]);
            final_rep[base + k] = s.x;                               // tile_place asks for the family of this read
            if (s.x == (uint32_t)(base + k)) {
                cnt[k] = s.y;
                size[k] = (~s.y) >> 1;
                founders++;
                reads += size[k];
                bits |= 1u << k;
            }
        }
    }
    // block-wide exclusive prefixes of (founders, reads)
    uint32_t inc_f = founders, inc_r = reads;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, inc_f, d), b = __shfl_up_sync(0xffffffffu, inc_r, d);
        if (lane >= d) { inc_f += a; inc_r += b; }
    }
    if (lane == 31) { s_warp_f[warp] = inc_f; s_warp_r[warp] = inc_r; }
    __syncthreads();
    uint32_t off_f = 0, off_r = 0, tot_f = 0, tot_r = 0;
#pragma unroll
    for (int k = 0; k < GT_THREADS / 32; k++) {
        if (k < warp) { off_f += s_warp_f[k]; off_r += s_warp_r[k]; }
        tot_f += s_warp_f[k]; tot_r += s_warp_r[k];
    }
    if (warp == 0) {
        // look-back by a whole warp: 32 predecessors per round, nearest first (scan.cuh)
        uint32_t ex_f = 0, ex_r = 0;
        if (tile != 0) {
            if (lane == 0) st_status(&status[tile], gt_pack(1, tot_f, tot_r));
            for (int64_t p = (int64_t)tile - 1;; p -= 32) {
                const int64_t idx = p - lane;
                unsigned long long s = idx >= 0 ? ld_status(&status[idx]) : gt_pack(2, 0, 0);
                while (__any_sync(0xffffffffu, (s >> 62) == 0))
                    if ((s >> 62) == 0) s = ld_status(&status[idx]);
                const uint32_t done = __ballot_sync(0xffffffffu, (s >> 62) == 2);
                const int first = done ? __ffs(done) - 1 : 31;
                uint32_t vf = lane <= first ? (uint32_t)(s >> 31) & 0x7FFFFFFFu : 0u;
                uint32_t vr = lane <= first ? (uint32_t)s & 0x7FFFFFFFu : 0u;
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) {
                    vf += __shfl_xor_sync(0xffffffffu, vf, d);
                    vr += __shfl_xor_sync(0xffffffffu, vr, d);
                }
                ex_f += vf; ex_r += vr;
                if (done) break;
            }
        }
        if (lane == 0) {
            st_status(&status[tile], gt_pack(2, ex_f + tot_f, ex_r + tot_r));
            s_excl_f = ex_f; s_excl_r = ex_r;
            if (tile == gridDim.x - 1) *n_fam_out = ex_f + tot_f;
        }
    }
    __syncthreads();
    uint32_t c = s_excl_f + off_f + (inc_f - founders);          // family number of this thread's first founder
    uint32_t start = s_excl_r + off_r + (inc_r - reads);

    // bitmap of founders: 8 threads x 4 reads = one word; word_rank = founders before the word
    uint32_t word = bits << (4 * (lane & 7));
    word |= __shfl_xor_sync(0xffffffffu, word, 1);
    word |= __shfl_xor_sync(0xffffffffu, word, 2);
    word |= __shfl_xor_sync(0xffffffffu, word, 4);
    if ((lane & 7) == 0 && base < n) {
        bitmap[base >> 5] = word;
        word_rank[base >> 5] = c;
    }
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        if (!(bits & (1u << k))) continue;
        const uint64_t i = base + k;
        const uint32_t sz = size[k];
        tmpl[c] = (uint32_t)i;
        fam_size[c] = sz;
        fam_start[c] = start;
        cursor[c] = start;
        is_sscs[c] = sz >= 2 ? ((cnt[k] & 1u) ? 1 : 3) : 0;      // bit 0: has a consensus row, bit 1: mode fields differ
        const uint4 b = __ldg(&fields[2 * i + 1]);
        fam_tlen[c] = (int32_t)b.z;
        fam_flag[c] = (uint16_t)(b.w & 0xFFFF);
        fam_mapq[c] = (uint8_t)((b.w >> 16) & 0xFF);
        fam_rg[c] = (uint8_t)(b.w >> 24);
        partner[c] = -1;
        const uint32_t bin = sz < hist_bins - 1 ? sz : hist_bins - 1;
        if (bin < 256) atomicAdd(&small_hist[bin], 1u);
        else atomicAdd(&hist[bin], 1ull);
        c++;
        start += sz;
    }
    __syncthreads();
    const uint32_t v = small_hist[tid];
    if (v) atomicAdd(&hist[tid], (unsigned long long)v);
    if (v && tid >= 16 && tid < 128) atomicAdd(n_mid, v);          // families of 16..127 reads
}

// family number of the family founded by read `first` (its bit is set in the bitmap)
__device__ __forceinline__ uint32_t family_of_founder(const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                                                      uint32_t first)
{
    return __ldg(&word_rank[first >> 5]) + __popc(__ldg(&bitmap[first >> 5]) & ((1u << (first & 31)) - 1u));
}

__global__ void __launch_bounds__(GT_THREADS)
tile_place_kernel(uint64_t n, const uint32_t *__restrict__ link, const uint32_t *__restrict__ final_rep,
                  const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                  uint32_t *__restrict__ cursor, uint32_t *__restrict__ perm)
{
    __shared__ uint32_t members[GT_TILE];                  // per representative: members besides itself, then its segment base
    const int tid = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * GT_TILE;
    const uint32_t here = (uint32_t)min((uint64_t)GT_TILE, n - base);
    uint32_t lk[GT_PER];
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        members[li] = 0;
        lk[k] = li < here ? link[base + li] : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li < here && !(lk[k] & LINK_REP)) atomicAdd(&members[lk[k] & (GT_TILE - 1)], 1u);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li < here && (lk[k] & LINK_REP)) {
            const uint32_t c = family_of_founder(bitmap, word_rank, final_rep[base + li]);
            members[li] = atomicAdd(&cursor[c], members[li] + 1u);     // room for the tile's whole group, one atomic
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li >= here) continue;
        const uint32_t at = (lk[k] & LINK_REP) ? members[li] : members[lk[k] & (GT_TILE - 1)] + (lk[k] >> LINK_RANK_SHIFT);
        perm[at] = (uint32_t)(base + li);
    }
}

// The duplex join on the grouping table (group_hash.cuh's group_join_kernel) with the partner's family
// taken from a rank query instead of a read -> family array.
__global__ void __launch_bounds__(256)
tile_join_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ tmpl, uint32_t n_fam,
                 const uint32_t *__restrict__ dup_bc, uint32_t n_bc, const GroupSlot *__restrict__ table, uint64_t mask,
                 const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                 int32_t *__restrict__ partner, const uint32_t *__restrict__ asym)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_fam) return;
    Rec r = load_rec(fields, tmpl[c]);
    const bool symmetric = !*asym;
    if (symmetric && !(r.b.w & 0x40u)) return;      // an R2 family hears from its R1 partner (partner[] starts at -1)
    int32_t found = -1;
    uint32_t d = r.a.x < n_bc ? __ldg(&dup_bc[r.a.x]) : NO_ID;
    if (d != NO_ID) {
        Rec q = r;
        q.a.x = d;                 // barcode halves swapped
        q.b.w ^= 0x40u;            // R1 <-> R2 (only the key bits of meta are compared)
        uint64_t slot = (rec_hash(q) * 0xD6E8FEB86659FD93ull >> 20) & mask;
        for (;;) {
            const uint32_t rep = table[slot].rep;
            if (rep == 0xFFFFFFFFu) break;
            if (same_tag(load_rec(fields, rep), q)) { found = (int32_t)family_of_founder(bitmap, word_rank, rep); break; }
            slot = (slot + 1) & mask;
        }
    }
    partner[c] = found;
    if (symmetric && found >= 0) partner[found] = (int32_t)c;
}

}  // namespace ccgpu
