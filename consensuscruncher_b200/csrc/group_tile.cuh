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
//   tile_mark     (after every tile has inserted) a representative reads its slot once: the slot's final
//                 read index is the family's founder.  Founders are marked in a bitmap (one ballot per 32
//                 reads) next to the number of reads their families hold.
//   (two scans over the bitmap's words: founders before a word = family numbers in creation order,
//    reads before a word = where the segments start)
//   tile_family   founders write the family records: template, size, segment start, class, singleton
//                 fields, size histogram.
//   (A version that fused mark + numbering + records into one chained scan over the tiles ran at 0.43 ms
//    against 0.2 ms for these pieces: every CTA sat at a barrier while its first warp looked back, behind
//    predecessors that were themselves waiting for their table gathers.)
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

#ifndef GT_INSERT_THREADS_CFG
#define GT_INSERT_THREADS_CFG 512
#endif
constexpr int GT_TILE = 1024;                   // reads per tile
constexpr int GT_THREADS = GT_INSERT_THREADS_CFG;   // tile_insert: threads per tile
constexpr int GT_PER = GT_TILE / GT_THREADS;    // ... and reads per thread
constexpr int GP_THREADS = 256;                 // tile_place
constexpr int GP_PER = GT_TILE / GP_THREADS;
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

// (after every tile has inserted) one thread per read, a warp = one word of the founders' bitmap.
// A representative reads its slot once: the slot's final read index names the family's founder; a
// representative that IS that index founds the family and keeps the slot's counter for tile_family.
// Also per word: the number of reads in the families its founders found (their segments' total length).
__global__ void __launch_bounds__(256)
tile_mark_kernel(uint64_t n, const GroupSlot *__restrict__ table, const uint32_t *__restrict__ link,
                 uint32_t *__restrict__ final_rep, uint32_t *__restrict__ founder_cnt,
                 uint32_t *__restrict__ bitmap, uint32_t *__restrict__ word_reads)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool founder = false;
    uint32_t size = 0;
    if (i < n) {
        const uint32_t lk = link[i];
        if (lk & LINK_REP) {
            const uint2 s = *reinterpret_cast<const uint2 *>(&table[lk & ~LINK_REP]);
            final_rep[i] = s.x;                                   // tile_place asks for the family of this read
            if (s.x == (uint32_t)i) {
                founder = true;
                founder_cnt[i] = s.y;
                size = (~s.y) >> 1;
            }
        }
    }
    const uint32_t word = __ballot_sync(0xffffffffu, founder);
    const uint32_t reads = __reduce_add_sync(0xffffffffu, size);
    if ((threadIdx.x & 31) == 0 && i < n) {
        bitmap[i >> 5] = word;
        word_reads[i >> 5] = reads;
    }
}

// (after the two scans over the words: founders before a word, reads before a word) founders write
// the family records in creation order; a family's segment starts where the earlier families' end.
__global__ void __launch_bounds__(256)
tile_family_kernel(const uint4 *__restrict__ fields, uint64_t n, const uint32_t *__restrict__ founder_cnt,
                   const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                   const uint32_t *__restrict__ word_start,
                   uint32_t *__restrict__ tmpl, uint32_t *__restrict__ fam_size, uint32_t *__restrict__ fam_start,
                   uint32_t *__restrict__ cursor, uint8_t *__restrict__ is_sscs, uint8_t *__restrict__ fam_mapq,
                   uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
                   int32_t *__restrict__ partner, unsigned long long *__restrict__ hist, uint32_t hist_bins,
                   uint32_t *__restrict__ n_mid)
{
    __shared__ uint32_t small_hist[256];
    small_hist[threadIdx.x] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    // grid-stride over words (a warp per word): a few CTAs per SM, so the shared histogram is flushed rarely
    const uint64_t words = (n + 31) / 32;
    for (uint64_t w = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < words; w += ((uint64_t)gridDim.x * blockDim.x) >> 5) {
        const uint32_t word = __ldg(&bitmap[w]);
        if (word == 0) continue;
        const bool founder = (word >> lane) & 1u;
        const uint64_t i = w * 32 + lane;
        uint32_t cnt = 0, size = 0;
        if (founder) { cnt = founder_cnt[i]; size = (~cnt) >> 1; }
        uint32_t incl = size;                                     // starts of the word's families: prefix of their sizes
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (!founder) continue;
        const uint32_t c = __ldg(&word_rank[w]) + __popc(word & ((1u << lane) - 1u));
        const uint32_t start = __ldg(&word_start[w]) + incl - size;
        tmpl[c] = (uint32_t)i;
        fam_size[c] = size;
        fam_start[c] = start;
        cursor[c] = start;
        is_sscs[c] = size >= 2 ? ((cnt & 1u) ? 1 : 3) : 0;       // bit 0: has a consensus row, bit 1: mode fields differ
        const uint4 b = __ldg(&fields[2 * i + 1]);
        fam_tlen[c] = (int32_t)b.z;
        fam_flag[c] = (uint16_t)(b.w & 0xFFFF);
        fam_mapq[c] = (uint8_t)((b.w >> 16) & 0xFF);
        fam_rg[c] = (uint8_t)(b.w >> 24);
        partner[c] = -1;
        const uint32_t bin = size < hist_bins - 1 ? size : hist_bins - 1;
        if (bin < 256) atomicAdd(&small_hist[bin], 1u);
        else atomicAdd(&hist[bin], 1ull);
    }
    __syncthreads();
    const uint32_t v = small_hist[threadIdx.x];
    if (v) atomicAdd(&hist[threadIdx.x], (unsigned long long)v);
    if (v && threadIdx.x >= 16 && threadIdx.x < 128) atomicAdd(n_mid, v);          // families of 16..127 reads
}

// family number of the family founded by read `first` (its bit is set in the bitmap)
__device__ __forceinline__ uint32_t family_of_founder(const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                                                      uint32_t first)
{
    return __ldg(&word_rank[first >> 5]) + __popc(__ldg(&bitmap[first >> 5]) & ((1u << (first & 31)) - 1u));
}

__global__ void __launch_bounds__(GP_THREADS)
tile_place_kernel(uint64_t n, const uint32_t *__restrict__ link, const uint32_t *__restrict__ final_rep,
                  const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                  uint32_t *__restrict__ cursor, uint32_t *__restrict__ perm)
{
    __shared__ uint32_t members[GT_TILE];                  // per representative: members besides itself, then its segment base
    const int tid = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * GT_TILE;
    const uint32_t here = (uint32_t)min((uint64_t)GT_TILE, n - base);
    uint32_t lk[GP_PER], fam[GP_PER];
#pragma unroll
    for (int k = 0; k < GP_PER; k++) {
        const uint32_t li = k * GP_THREADS + tid;
        members[li] = 0;
        lk[k] = li < here ? link[base + li] : 0u;
        fam[k] = 0;
        // the representatives' trips to the bitmap start before the members are counted
        if (li < here && (lk[k] & LINK_REP)) fam[k] = family_of_founder(bitmap, word_rank, final_rep[base + li]);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GP_PER; k++) {
        const uint32_t li = k * GP_THREADS + tid;
        if (li < here && !(lk[k] & LINK_REP)) atomicAdd(&members[lk[k] & (GT_TILE - 1)], 1u);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GP_PER; k++) {
        const uint32_t li = k * GP_THREADS + tid;
        if (li < here && (lk[k] & LINK_REP))
            members[li] = atomicAdd(&cursor[fam[k]], members[li] + 1u);     // room for the tile's whole group, one atomic
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GP_PER; k++) {
        const uint32_t li = k * GP_THREADS + tid;
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
