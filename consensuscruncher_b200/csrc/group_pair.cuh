// Grouping by read PAIR, tile by tile (sm_100a) - group_tile.cuh with half the work.
//
// read_bam inserts the two mates of a completed pair together (consensus_helper.py:438-475), and inside
// the reference's domain a pair founds or joins its two families together (SSCS_maker.py:275 emits a pair
// name only when it owns exactly two tags).  So the unit that is grouped here is the pair: pair k = reads
// 2k (first-seen mate) and 2k + 1, keyed by the first mate's family tuple plus the second mate's
// orientation / read-number bits.  A pair family P is two read families of equal size and membership - E
// (the even reads) and O (the odd reads) - which come out as families 2 * rank(P) and 2 * rank(P) + 1 in
// creation order, exactly the numbering the read-level paths produce.  One shared-table probe, one link
// word, one place per PAIR; everything else is the design of group_tile.cuh (interior runs finished in the
// tile, a device-sized global table for the families of runs that touch a tile edge, one chained scan over
// the founders' bitmap, the order the shortcut relies on checked in the same pass).
//
// What the pair view must not hide is checked where it would show:
//   * mates that do not mirror each other (barcode, cigar pair, swapped coordinates): a flag for the host
//   * two pair families on the same barcode / cigar pair / coordinates that agree in exactly ONE of the two
//     mates' bits - the reads of that mate would be one family, those of the other two: the pair does not
//     found / join its two families together (CC_ERR_DOMAIN).  Such variants share a hash chain (the table
//     key leaves the bits out), so whoever is inserted later walks past the earlier one and sees it.
// Either flag makes the host redo the batch with the read-level path, which reproduces the reference's
// families for such input and names the offending pairs.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "group_tile.cuh"

namespace ccgpu {

constexpr int PG_TILE = 512;                    // pairs per tile
constexpr int PG_THREADS = 256;
constexpr int PG_PER = PG_TILE / PG_THREADS;    // pairs per thread
constexpr uint32_t PG_HT = 2 * PG_TILE;         // shared table slots
constexpr uint32_t PG_MIX_E = 0x80000000u, PG_MIX_O = 0x40000000u;      // shared counter: mode fields differ among the even / odd reads
constexpr uint32_t PG_COUNT = 0x3FFFFFFFu;
constexpr uint32_t PG_NONE = 0x3FFu;            // "no partner in the tile" in an interior representative's aux word

// device-side scalars (uint32 block next to the stretch table)
enum { GPC_EDGE = 0, GPC_UNSORTED = 1, GPC_MASK = 2, GPC_COUNT = 4 };

inline uint64_t pg_tiles(uint64_t n_pairs) { return (n_pairs + PG_TILE - 1) / PG_TILE; }

struct PairKey {
    uint4 a;        // bc, cig, tid, pos of the first mate
    uint2 m;        // mtid, mpos
    uint32_t bits;  // key bits of the first mate | key bits of the second mate << 2
};

__device__ __forceinline__ bool same_base(const PairKey &x, const PairKey &y)
{
    return x.a.x == y.a.x && x.a.y == y.a.y && x.a.z == y.a.z && x.a.w == y.a.w && x.m.x == y.m.x && x.m.y == y.m.y;
}

// two pair families on one base key that agree in exactly one mate's bits break the pairing rule
__device__ __forceinline__ bool half_equal(uint32_t b1, uint32_t b2)
{
    return ((b1 & 3u) == (b2 & 3u)) != ((b1 >> 2) == (b2 >> 2));
}

__device__ __forceinline__ uint32_t base_hash(const uint4 &a, const uint2 &m)
{
    uint32_t h = a.x * 0x9E3779B1u ^ a.w * 0x85EBCA6Bu ^ m.y * 0xC2B2AE35u ^ a.y * 0x27D4EB2Fu;
    h ^= (a.z * 31u + m.x) * 0x165667B1u;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return h;
}

__device__ __forceinline__ uint64_t base_hash64(const uint4 &a, const uint2 &m)
{
    return tag_hash(a.x, a.y, a.z, a.w, m.x, m.y, 0u);
}

// pair key of pair p from the field records (global memory)
__device__ __forceinline__ PairKey load_pair_key(const uint4 *__restrict__ fields, uint32_t p)
{
    const Rec r0 = load_rec(fields, 2 * p);
    const uint4 b1 = __ldg(&fields[2 * (2 * (uint64_t)p + 1) + 1]);
    return PairKey{r0.a, make_uint2(r0.b.x, r0.b.y), key_bits(r0.b.w) | (key_bits(b1.w) << 2)};
}

// MODE 0: first pass, one tile per CTA.  MODE 1: runs only if the first pass found the input out of order; then
// nothing is local (every representative is listed); small grid that walks the tiles.
// Per pair: plink (member: representative | rank << 10; representative: LINK_REP [| LINK_LOCAL]), and for
// representatives pstash (members << 2 | even reads' mode fields differ | odd reads' << 1) and paux (interior:
// tile-local partner of the even | of the odd family << 10, PG_NONE for none; listed: later its table slot).
template <int MODE>
__global__ void __launch_bounds__(PG_THREADS)
pair_local_kernel(const uint4 *__restrict__ fields, uint64_t n_pairs, const uint32_t *__restrict__ dup_bc, uint32_t n_bc,
                  uint32_t *__restrict__ gpc, uint32_t *__restrict__ stretch, uint32_t *__restrict__ faults,
                  uint32_t *__restrict__ unmirrored, uint32_t *__restrict__ plink, uint32_t *__restrict__ paux, uint32_t *__restrict__ pstash,
                  uint32_t *__restrict__ edge_list, uint32_t *__restrict__ bitmap, uint32_t *__restrict__ word_pairs,
                  uint32_t *__restrict__ word_sscs)
{
    __shared__ uint4 ka[PG_TILE];                         // bc, cig, tid, pos of the first mate
    __shared__ uint4 kb[PG_TILE];                         // mtid, mpos, tlen, meta of the first mate
    __shared__ uint2 kc[PG_TILE];                         // tlen, meta of the second mate
    __shared__ uint32_t ht[PG_HT];                        // pair key -> smallest local pair index seen so far
    __shared__ uint32_t members[PG_TILE];                 // per representative: members besides itself | PG_MIX_*
    __shared__ uint16_t rep_list[PG_TILE];
    __shared__ uint32_t n_reps, n_edge, edge_base;
    __shared__ int2 neighbour[2];
    __shared__ uint32_t bm[PG_TILE / 32], bm_pairs[PG_TILE / 32], bm_sscs[PG_TILE / 32];
    const int tid = threadIdx.x;
    if (MODE == 1 && !gpc[GPC_UNSORTED]) return;
    uint64_t tile = blockIdx.x;
    do {
    const uint64_t base = tile * PG_TILE;
    const uint32_t here = (uint32_t)min((uint64_t)PG_TILE, n_pairs - base);

    for (int k = tid; k < (int)PG_HT / 4; k += PG_THREADS)
        reinterpret_cast<uint4 *>(ht)[k] = make_uint4(GT_EMPTY, GT_EMPTY, GT_EMPTY, GT_EMPTY);
    if (tid == 0) { n_reps = 0; n_edge = 0; }
    if (tid < PG_TILE / 32) { bm[tid] = 0; bm_pairs[tid] = 0; bm_sscs[tid] = 0; }
    uint32_t h[PG_PER];
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        members[li] = 0;
        if (li < here) {
            const Rec r0 = load_rec(fields, (uint32_t)(2 * (base + li))), r1 = load_rec(fields, (uint32_t)(2 * (base + li) + 1));
            ka[li] = r0.a;
            kb[li] = r0.b;
            kc[li] = make_uint2(r1.b.z, r1.b.w);
            h[k] = base_hash(r0.a, make_uint2(r0.b.x, r0.b.y)) & (PG_HT - 1);
            // the pair view needs mates that mirror each other
            if (r1.a.x != r0.a.x || r1.a.y != r0.a.y || r1.a.z != r0.b.x || r1.a.w != r0.b.y || r1.b.x != r0.a.z || r1.b.y != r0.a.w)
                *unmirrored = 1u;
        }
    }
    // completion coordinates of the nearest one-contig pairs outside the tile, looked for by one warp each
    if (tid < 64) {
        const int lane = tid & 31;
        const bool fwd = tid >= 32;
        const int64_t edge = fwd ? (int64_t)(base + PG_TILE) : (int64_t)base - 1;
        const int64_t p = fwd ? edge + lane : edge - lane;
        const bool exists = p >= 0 && p < (int64_t)n_pairs;
        bool intra = false;
        int2 at = make_int2(NB_NONE, 0);
        if (exists) {
            const Rec r = load_rec(fields, (uint32_t)(2 * p + 1));
            intra = (int)r.a.z == (int)r.b.x;
            at = make_int2((int)r.a.z, (int)r.a.w);
        }
        const uint32_t hit = __ballot_sync(0xffffffffu, intra), off_end = __ballot_sync(0xffffffffu, !exists);
        const int first = hit ? __ffs(hit) - 1 : 32, end_at = off_end ? __ffs(off_end) - 1 : 32;
        if (first < end_at) { if (lane == first) neighbour[fwd] = at; }
        else if (end_at < 32) { if (lane == 0) neighbour[fwd] = make_int2(NB_NONE, 0); }
        else if (lane == 0) { neighbour[fwd] = make_int2(NB_NONE, 0); if (MODE == 0) gpc[GPC_UNSORTED] = 1u; }
    }
    __syncthreads();
    const int2 before = neighbour[0], after = neighbour[1];

    // tile-local table: a slot ends up holding the smallest local index of its pair family; the table key leaves
    // the mates' bits out, so families that differ only there share a chain and meet
    uint32_t my_slot[PG_PER];
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        if (li >= here) continue;
        const PairKey me{ka[li], make_uint2(kb[li].x, kb[li].y), key_bits(kb[li].w) | (key_bits(kc[li].y) << 2)};
        uint32_t s = h[k];
        for (;;) {
            uint32_t cur = ht[s];
            if (cur == GT_EMPTY) {
                cur = atomicCAS(&ht[s], GT_EMPTY, li);
                if (cur == GT_EMPTY) break;
            }
            const PairKey q{ka[cur], make_uint2(kb[cur].x, kb[cur].y), key_bits(kb[cur].w) | (key_bits(kc[cur].y) << 2)};
            if (same_base(me, q)) {
                if (q.bits == me.bits) {
                    if (li < cur) atomicMin(&ht[s], li);
                    break;
                }
                if (half_equal(me.bits, q.bits)) atomicAdd(faults, 1u);
            }
            s = (s + 1) & (PG_HT - 1);
        }
        my_slot[k] = s;
        // what the interior shortcut relies on, checked at every one-contig pair
        if (MODE == 0 && (int)me.a.z == (int)me.m.x) {
            if ((int)me.m.y < (int)me.a.w) gpc[GPC_UNSORTED] = 1u;         // completes at its larger position
            int2 prev = before;                                             // the one-contig pair before this one
            for (uint32_t b = li; b >= 1;) {
                b -= 1;
                if ((int)ka[b].z == (int)kb[b].x) { prev = make_int2((int)kb[b].x, (int)kb[b].y); break; }
            }
            if ((int)me.m.x == prev.x) {
                if ((int)me.m.y < prev.y) gpc[GPC_UNSORTED] = 1u;          // positions never go backwards inside a contig
            } else {
                const uint32_t mark = (uint32_t)me.m.x + 2u;                // a contig begins: its only stretch
                uint32_t s2 = (mark * 2654435761u >> 20) & (STRETCH_SLOTS - 1), tries = 0;
                for (;; s2 = (s2 + 1) & (STRETCH_SLOTS - 1)) {
                    const uint32_t seen = atomicCAS(&stretch[s2], 0u, mark);
                    if (seen == 0u) break;
                    if (seen == mark || ++tries == STRETCH_SLOTS) { gpc[GPC_UNSORTED] = 1u; break; }
                }
            }
        }
    }
    __syncthreads();

    // members count themselves at their representative; the rank is their place in the tile's group
    uint32_t lrep[PG_PER], rank[PG_PER];
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        if (li >= here) continue;
        lrep[k] = ht[my_slot[k]];
        rank[k] = 0;
        if (lrep[k] != li) {
            const uint32_t r = lrep[k];
            rank[k] = 1 + (atomicAdd(&members[r], 1u) & PG_COUNT);
            // every member is compared with its representative: no difference anywhere = uniform mode fields
            uint32_t mix = 0;
            if (kb[li].z != kb[r].z || kb[li].w != kb[r].w) mix |= PG_MIX_E;
            if (kc[li].x != kc[r].x || kc[li].y != kc[r].y) mix |= PG_MIX_O;
            if (mix) atomicOr(&members[r], mix);
        }
    }
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {                     // the representatives queue up, one shared addition per warp
        const uint32_t li = k * PG_THREADS + tid;
        const bool is_rep = li < here && lrep[k] == li;
        const uint32_t reps_here = __ballot_sync(0xffffffffu, is_rep);
        uint32_t at = 0;
        if ((tid & 31) == 0 && reps_here) at = atomicAdd(&n_reps, (uint32_t)__popc(reps_here));
        at = __shfl_sync(0xffffffffu, at, 0);
        if (is_rep) rep_list[at + __popc(reps_here & ((1u << (tid & 31)) - 1u))] = (uint16_t)li;
    }
    __syncthreads();

    // representatives: interior families are finished here, the others are listed for the table
    const uint32_t reps = n_reps;
    uint32_t edge_at[(PG_TILE + PG_THREADS - 1) / PG_THREADS];
    int e = 0;
    for (uint32_t x = tid; x < reps; x += PG_THREADS, e++) {
        const uint32_t li = rep_list[x];
        const uint32_t p = (uint32_t)(base + li);
        const uint32_t m = members[li];
        const uint32_t size = (m & PG_COUNT) + 1u;
        const int2 at = make_int2((int)kb[li].x, (int)kb[li].y);  // where the pair completes: the second mate's coordinate
        const bool one_contig = (int)ka[li].z == at.x;
        const bool local = MODE == 0 && one_contig && !(at.x == before.x && at.y == before.y) && !(at.x == after.x && at.y == after.y);
        pstash[p] = (size << 2) | ((m & PG_MIX_E) ? 1u : 0u) | ((m & PG_MIX_O) ? 2u : 0u);
        edge_at[e] = GT_EMPTY;
        if (local) {
            plink[p] = LINK_REP | LINK_LOCAL;
            atomicOr(&bm[li >> 5], 1u << (li & 31));
            atomicAdd(&bm_pairs[li >> 5], size);
            if (size >= 2) atomicAdd(&bm_sscs[li >> 5], 1u);
            // duplex partners (duplex_tag, consensus_helper.py:605-648) of the even and of the odd family: the pair
            // families on the swapped barcode, same coordinates, whose first / second mate has the other read number.
            // Same coordinates = same run: in this tile or nowhere.
            uint32_t found_e = PG_NONE, found_o = PG_NONE;
            const uint32_t d = ka[li].x < n_bc ? __ldg(&dup_bc[ka[li].x]) : NO_ID;
            if (d != NO_ID) {
                PairKey q{ka[li], make_uint2(kb[li].x, kb[li].y), 0u};
                q.a.x = d;
                const uint32_t mine = key_bits(kb[li].w) | (key_bits(kc[li].y) << 2);
                uint32_t s = base_hash(q.a, q.m) & (PG_HT - 1);
                for (;;) {
                    const uint32_t cur = ht[s];
                    if (cur == GT_EMPTY) break;
                    const PairKey c{ka[cur], make_uint2(kb[cur].x, kb[cur].y), key_bits(kb[cur].w) | (key_bits(kc[cur].y) << 2)};
                    if (same_base(q, c)) {
                        if ((c.bits & 3u) == ((mine & 3u) ^ 2u)) found_e = cur;
                        if ((c.bits >> 2) == ((mine >> 2) ^ 2u)) found_o = cur;
                    }
                    s = (s + 1) & (PG_HT - 1);
                }
            }
            paux[p] = found_e | (found_o << 10);
        } else {
            plink[p] = LINK_REP;
            edge_at[e] = atomicAdd(&n_edge, 1u);
        }
    }
    __syncthreads();
    if (tid == 0 && n_edge) edge_base = atomicAdd(&gpc[GPC_EDGE], n_edge);
    __syncthreads();
    e = 0;
    for (uint32_t x = tid; x < reps; x += PG_THREADS, e++)
        if (edge_at[e] != GT_EMPTY) edge_list[edge_base + edge_at[e]] = (uint32_t)(base + rep_list[x]);
    if (tid < PG_TILE / 32 && (uint32_t)tid * 32 < here) {     // listed founders are added by pair_edge_mark_kernel
        bitmap[(base >> 5) + tid] = bm[tid];
        word_pairs[(base >> 5) + tid] = bm_pairs[tid];
        word_sscs[(base >> 5) + tid] = bm_sscs[tid];
    }
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        if (li < here && lrep[k] != li) plink[base + li] = lrep[k] | (rank[k] << LINK_RANK_SHIFT);
    }
    if (MODE == 1) __syncthreads();
    } while (MODE == 1 && (tile += gridDim.x) * PG_TILE < n_pairs);
}

// start the second pass from an empty list (only if the first pass found the input out of order)
__global__ void pair_redo_kernel(uint32_t *gpc)
{
    if (gpc[GPC_UNSORTED]) gpc[GPC_EDGE] = 0u;
}

// table for the listed representatives: 2 x their number, rounded up to a power of two; a slot holds the smallest
// pair index of its family (0xFFFFFFFF: empty) and members << 2 | mode-field bits, summed / ORed over the tiles
__global__ void __launch_bounds__(256)
pair_table_prepare_kernel(uint32_t *gpc, GroupSlot *__restrict__ table, uint64_t max_slots)
{
    uint64_t slots = 1024;
    const uint64_t want = 2ull * gpc[GPC_EDGE];
    while (slots < want && slots < max_slots) slots <<= 1;
    uint2 *t = reinterpret_cast<uint2 *>(table);
    for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s < slots; s += (uint64_t)gridDim.x * blockDim.x)
        t[s] = make_uint2(0xFFFFFFFFu, 0u);
    if (blockIdx.x == 0 && threadIdx.x == 0) gpc[GPC_MASK] = (uint32_t)(slots - 1);
}

__device__ __forceinline__ uint64_t pair_slot(const PairKey &k, uint64_t mask)
{
    return (base_hash64(k.a, k.m) * 0xD6E8FEB86659FD93ull >> 20) & mask;
}

// one thread per listed representative: claim / lower the slot's pair index, add the tile's group to its counter
__global__ void __launch_bounds__(256)
pair_edge_insert_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ edge_list, const uint32_t *__restrict__ gpc,
                        GroupSlot *__restrict__ table, const uint32_t *__restrict__ pstash, uint32_t *__restrict__ paux,
                        uint32_t *__restrict__ faults)
{
    const uint32_t total = gpc[GPC_EDGE];
    const uint64_t mask = gpc[GPC_MASK];
    for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x) {
        const uint32_t p = edge_list[x];
        const PairKey me = load_pair_key(fields, p);
        uint32_t mine = pstash[p];                              // members << 2 | mode-field bits of the tile's group
        uint64_t slot = pair_slot(me, mask);
        for (;;) {
            uint32_t rep = ld_relaxed_u32(&table[slot].rep);
            if (rep == 0xFFFFFFFFu) {
                rep = atomicCAS(&table[slot].rep, 0xFFFFFFFFu, p);
                if (rep == 0xFFFFFFFFu) break;                  // claimed: so far this tile founds the family
            }
            const PairKey q = load_pair_key(fields, rep);
            if (same_base(me, q)) {
                if (q.bits == me.bits) {                         // another tile's group of the same family
                    if (p < rep) atomicMin(&table[slot].rep, p);
                    // the comparisons span the family (group_hash.cuh): any difference shows on one of them
                    const uint4 b0 = __ldg(&fields[2 * (2 * (uint64_t)p) + 1]), c0 = __ldg(&fields[2 * (2 * (uint64_t)rep) + 1]);
                    const uint4 b1 = __ldg(&fields[2 * (2 * (uint64_t)p + 1) + 1]), c1 = __ldg(&fields[2 * (2 * (uint64_t)rep + 1) + 1]);
                    if (b0.z != c0.z || b0.w != c0.w) mine |= 1u;
                    if (b1.z != c1.z || b1.w != c1.w) mine |= 2u;
                    break;
                }
                if (half_equal(me.bits, q.bits)) atomicAdd(faults, 1u);
            }
            slot = (slot + 1) & mask;
        }
        CC_CHECK(slot <= mask);
        atomicAdd(&table[slot].cnt, mine & ~3u);
        if (mine & 3u) atomicOr(&table[slot].cnt, mine & 3u);
        paux[p] = (uint32_t)slot;
    }
}

// a listed representative whose slot still names it founds its family: its bit, its family's pairs and class join
// the words the tiles wrote; paux turns from the slot into the founder, a founder's pstash into the family's word
__global__ void __launch_bounds__(256)
pair_edge_mark_kernel(const uint32_t *__restrict__ edge_list, const uint32_t *__restrict__ gpc, const GroupSlot *__restrict__ table,
                      uint32_t *__restrict__ paux, uint32_t *__restrict__ pstash, uint32_t *__restrict__ bitmap,
                      uint32_t *__restrict__ word_pairs, uint32_t *__restrict__ word_sscs)
{
    const uint32_t total = gpc[GPC_EDGE];
    for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x) {
        const uint32_t p = edge_list[x];
        const uint2 s = *reinterpret_cast<const uint2 *>(&table[paux[p]]);
        paux[p] = s.x;
        if (s.x == p) {
            const uint32_t size = s.y >> 2;
            pstash[p] = s.y;
            atomicOr(&bitmap[p >> 5], 1u << (p & 31));
            atomicAdd(&word_pairs[p >> 5], size);
            if (size >= 2) atomicAdd(&word_sscs[p >> 5], 1u);
        }
    }
}

// (after word_scan over the pair bitmap: founders / pairs / SSCS pair families before every word) a founding pair
// writes the records of its two read families - E = 2 * rank, O = 2 * rank + 1 - side by side; the pair family's
// segment of 2 * size places holds the even reads first, then the odd reads.
__global__ void __launch_bounds__(256)
pair_family_kernel(const uint4 *__restrict__ fields, uint64_t n_pairs, const uint32_t *__restrict__ pstash,
                   const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                   const uint32_t *__restrict__ word_start, const uint32_t *__restrict__ word_sscs_start,
                   uint32_t *__restrict__ tmpl, uint32_t *__restrict__ fam_size, uint32_t *__restrict__ fam_start,
                   uint32_t *__restrict__ pcursor, uint8_t *__restrict__ is_sscs, uint8_t *__restrict__ fam_mapq,
                   uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
                   unsigned long long *__restrict__ hist, uint32_t hist_bins, uint32_t *__restrict__ n_mid,
                   uint32_t *__restrict__ sscs_list, int32_t *__restrict__ sscs_row, uint32_t *__restrict__ vstart,
                   uint32_t *__restrict__ vsize, uint32_t *__restrict__ big_list, uint32_t *__restrict__ n_big,
                   uint32_t small_max, uint32_t *__restrict__ mid_list, uint32_t *__restrict__ n_midlist)
{
    __shared__ uint32_t small_hist[256];
    small_hist[threadIdx.x] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    constexpr int FW = 4;                                   // words per warp iteration, all their loads issued up front
    const uint64_t words = (n_pairs + 31) / 32;
    const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t w0 = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * FW; w0 < words; w0 += warps * FW) {
        uint32_t word[FW], st[FW];
        uint4 m0[FW], m1[FW];
#pragma unroll
        for (int k = 0; k < FW; k++) word[k] = w0 + k < words ? __ldg(&bitmap[w0 + k]) : 0u;
#pragma unroll
        for (int k = 0; k < FW; k++) {
            const bool founder = (word[k] >> lane) & 1u;
            const uint64_t p = (w0 + k) * 32 + lane;
            st[k] = founder ? pstash[p] : 0u;
            m0[k] = founder ? __ldg(&fields[4 * p + 1]) : make_uint4(0, 0, 0, 0);
            m1[k] = founder ? __ldg(&fields[4 * p + 3]) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int k = 0; k < FW; k++) {
            if (word[k] == 0) continue;
            const uint64_t w = w0 + k, p = w * 32 + lane;
            const bool founder = (word[k] >> lane) & 1u;
            const uint32_t size = st[k] >> 2;
            uint32_t incl = size;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            const uint32_t big = __ballot_sync(0xffffffffu, size >= 2);
            if (!founder) continue;
            const uint32_t P = __ldg(&word_rank[w]) + __popc(word[k] & ((1u << lane) - 1u));
            const uint32_t sp = __ldg(&word_start[w]) + incl - size;       // segment start, in pairs
            CC_CHECK(2ull * P + 1 < 2 * n_pairs && (uint64_t)sp + size <= n_pairs);
            int2 rows = make_int2(-1, -1);
            if (size >= 2) {
                const uint32_t r = 2u * (__ldg(&word_sscs_start[w]) + __popc(big & ((1u << lane) - 1u)));
                rows = make_int2((int)r, (int)r + 1);
                *reinterpret_cast<uint2 *>(sscs_list + r) = make_uint2(2 * P, 2 * P + 1);
                // what the vote reads per consensus row (vote_segments_kernel does the same for the other groupings)
                *reinterpret_cast<uint2 *>(vstart + r) = make_uint2(2 * sp, 2 * sp + size);
                *reinterpret_cast<uint2 *>(vsize + r) = make_uint2(size | ((st[k] & 1u) << 31), size | ((st[k] & 2u) << 30));
                if (size > small_max) *reinterpret_cast<uint2 *>(big_list + atomicAdd(n_big, 2u)) = make_uint2(r, r + 1);
                else if (size > 15) *reinterpret_cast<uint2 *>(mid_list + atomicAdd(n_midlist, 2u)) = make_uint2(r, r + 1);
            }
            *reinterpret_cast<int2 *>(sscs_row + 2 * P) = rows;
            *reinterpret_cast<uint2 *>(tmpl + 2 * P) = make_uint2((uint32_t)(2 * p), (uint32_t)(2 * p + 1));
            *reinterpret_cast<uint2 *>(fam_size + 2 * P) = make_uint2(size, size);
            *reinterpret_cast<uint2 *>(fam_start + 2 * P) = make_uint2(2 * sp, 2 * sp + size);
            pcursor[P] = sp;
            const uint32_t cls_e = size >= 2 ? ((st[k] & 1u) ? 3u : 1u) : 0u, cls_o = size >= 2 ? ((st[k] & 2u) ? 3u : 1u) : 0u;
            *reinterpret_cast<uint16_t *>(is_sscs + 2 * P) = (uint16_t)(cls_e | (cls_o << 8));
            *reinterpret_cast<int2 *>(fam_tlen + 2 * P) = make_int2((int)m0[k].z, (int)m1[k].z);
            *reinterpret_cast<uint32_t *>(fam_flag + 2 * P) = (m0[k].w & 0xFFFFu) | (m1[k].w << 16);
            *reinterpret_cast<uint16_t *>(fam_mapq + 2 * P) = (uint16_t)(((m0[k].w >> 16) & 0xFFu) | (((m1[k].w >> 16) & 0xFFu) << 8));
            *reinterpret_cast<uint16_t *>(fam_rg + 2 * P) = (uint16_t)((m0[k].w >> 24) | ((m1[k].w >> 24) << 8));
            const uint32_t bin = size < hist_bins - 1 ? size : hist_bins - 1;
            if (bin < 256) atomicAdd(&small_hist[bin], 2u);
            else atomicAdd(&hist[bin], 2ull);
        }
    }
    __syncthreads();
    const uint32_t v = small_hist[threadIdx.x];
    if (v) atomicAdd(&hist[threadIdx.x], (unsigned long long)v);
    if (v && threadIdx.x >= 16 && threadIdx.x < 128) atomicAdd(n_mid, v);          // families of 16..127 reads
}

// places: a representative reserves room for its tile's pairs in the pair family's segment (one atomic; none when
// the tile holds the whole family), every pair writes its even read into the first half and its odd read into the
// second half of the segment
__global__ void __launch_bounds__(PG_THREADS)
pair_place_kernel(uint64_t n_pairs, const uint32_t *__restrict__ plink, const uint32_t *__restrict__ paux,
                  const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                  const uint32_t *__restrict__ fam_start, const uint32_t *__restrict__ fam_size,
                  uint32_t *__restrict__ pcursor, uint32_t *__restrict__ perm)
{
    __shared__ uint32_t members[PG_TILE];                  // per representative: members besides itself, then its base (in pairs)
    __shared__ uint32_t seg[PG_TILE], len[PG_TILE];        // per representative: the family's segment start (in pairs) and size
    __shared__ uint32_t t_bits[PG_TILE / 32], t_rank[PG_TILE / 32];
    const int tid = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * PG_TILE;
    const uint32_t here = (uint32_t)min((uint64_t)PG_TILE, n_pairs - base);
    if (tid < PG_TILE / 32 && (uint32_t)tid * 32 < here) {
        t_bits[tid] = bitmap[(base >> 5) + tid];
        t_rank[tid] = word_rank[(base >> 5) + tid];
    }
    uint32_t lk[PG_PER], fam[PG_PER];
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        members[li] = 0;
        lk[k] = li < here ? plink[base + li] : 0u;
        fam[k] = 0;
        if (li < here && (lk[k] & LINK_REP) && !(lk[k] & LINK_LOCAL)) fam[k] = family_of_founder(bitmap, word_rank, paux[base + li]);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        if (li < here && !(lk[k] & LINK_REP)) atomicAdd(&members[lk[k] & (GT_TILE - 1)], 1u);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        if (li >= here || !(lk[k] & LINK_REP)) continue;
        const bool local = (lk[k] & LINK_LOCAL) != 0;
        const uint32_t P = local ? t_rank[li >> 5] + __popc(t_bits[li >> 5] & ((1u << (li & 31)) - 1u)) : fam[k];
        const uint32_t sp = fam_start[2 * P] >> 1;
        seg[li] = sp;
        len[li] = fam_size[2 * P];
        members[li] = local ? sp : atomicAdd(&pcursor[P], members[li] + 1u);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PG_PER; k++) {
        const uint32_t li = k * PG_THREADS + tid;
        if (li >= here) continue;
        const uint32_t r = (lk[k] & LINK_REP) ? li : (lk[k] & (GT_TILE - 1));
        const uint32_t q = members[r] + ((lk[k] & LINK_REP) ? 0u : (lk[k] >> LINK_RANK_SHIFT));    // place among the family's pairs
        const uint32_t at = seg[r] + q;                                                             // = 2 * seg + (q - seg)
        CC_CHECK(q >= seg[r] && q < seg[r] + len[r] && (uint64_t)at + len[r] < 2 * n_pairs);
        perm[at] = (uint32_t)(2 * (base + li));
        perm[at + len[r]] = (uint32_t)(2 * (base + li) + 1);
    }
}

// Duplex partners of a pair family's two read families.  Interior: the tile-local answers of pair_local_kernel;
// otherwise one walk along the chain of the swapped barcode's base key in the table finds both.
__global__ void __launch_bounds__(256)
pair_join_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ tmpl, const uint32_t *__restrict__ n_fam_ptr,
                 const uint32_t *__restrict__ dup_bc, uint32_t n_bc, const GroupSlot *__restrict__ table,
                 const uint32_t *__restrict__ gpc, const uint32_t *__restrict__ plink, const uint32_t *__restrict__ paux,
                 const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank, int32_t *__restrict__ partner)
{
    const uint32_t n_pf = *n_fam_ptr >> 1;
    for (uint32_t P = blockIdx.x * blockDim.x + threadIdx.x; P < n_pf; P += gridDim.x * blockDim.x) {
        const uint32_t p = tmpl[2 * P] >> 1;                      // the founding pair
        int32_t pe = -1, po = -1;
        if (__ldg(&plink[p]) & LINK_LOCAL) {
            const uint32_t aux = __ldg(&paux[p]), tile0 = p & ~(uint32_t)(PG_TILE - 1);
            const uint32_t le = aux & 0x3FFu, lo = (aux >> 10) & 0x3FFu;
            if (le != PG_NONE) pe = (int32_t)(2 * family_of_founder(bitmap, word_rank, tile0 + le));
            if (lo != PG_NONE) po = (int32_t)(2 * family_of_founder(bitmap, word_rank, tile0 + lo) + 1);
        } else {
            PairKey q = load_pair_key(fields, p);
            const uint32_t mine = q.bits;
            const uint32_t d = q.a.x < n_bc ? __ldg(&dup_bc[q.a.x]) : NO_ID;
            if (d != NO_ID) {
                q.a.x = d;
                const uint64_t mask = gpc[GPC_MASK];
                uint64_t slot = pair_slot(q, mask);
                for (;;) {
                    const uint32_t rep = table[slot].rep;
                    if (rep == 0xFFFFFFFFu) break;
                    const PairKey c = load_pair_key(fields, rep);
                    if (same_base(q, c)) {
                        if ((c.bits & 3u) == ((mine & 3u) ^ 2u)) pe = (int32_t)(2 * family_of_founder(bitmap, word_rank, rep));
                        if ((c.bits >> 2) == ((mine >> 2) ^ 2u)) po = (int32_t)(2 * family_of_founder(bitmap, word_rank, rep) + 1);
                    }
                    slot = (slot + 1) & mask;
                }
            }
        }
        *reinterpret_cast<int2 *>(partner + 2 * P) = make_int2(pe, po);
    }
}

}  // namespace ccgpu
