// Grouping reads into tag families, tile by tile (sm_100a) - the default grouping path.
//
// read_bam's dictionaries (consensus_helper.py:447-475: `tag in read_dict`, append, size + 1) keyed by the
// exact family tuple.  The input is in pair-completion order (read 2k+1 is the mate that completed pair k).
// Take the pairs whose mates lie on one contig: in a coordinate-sorted BAM a pair completes at the larger of
// its two positions, that position never goes backwards inside a contig, and it is a function of the family
// key alone - so every member of a family, of its mate family and of its duplex partner (which share both
// coordinates) sits in ONE run of pairs with equal completion coordinate.  A CTA groups a tile of 1024
// consecutive reads in shared memory; a run that lies strictly inside the tile holds whole families, which
// are finished right there (founder, size, mode-field uniformity, duplex partner) without any global
// traffic.  Families of runs that touch a tile edge, and pairs that span two contigs, go through a global
// table, which is sized on the device for exactly those.  Nothing is taken on trust: the same pass checks
// what it relies on - every one-contig pair completes at its larger position, completion positions do not
// decrease from one such pair to the next inside a contig, every contig is one stretch - and otherwise
// raises a flag, upon which the second launch redoes the tiles with every family going through the table.
//
//   tile_local      shared-memory table: every read finds its tile-local representative (smallest index of
//                   its tuple in the tile), representatives count members.  Interior families are final;
//                   the representatives of the others are appended to a list.
//   tile_local (2nd launch)  does nothing unless the flag is up; then every representative goes to the
//                   list (the input is in arbitrary order: nothing is assumed, everything meets in the table).
//   table_prepare   table size = power of two >= 2 x listed representatives (device side), cleared.
//   edge_insert     one thread per listed representative: claim / lower the slot's read index, add the
//                   tile's member count to the slot.
//   edge_mark       a listed representative whose slot still names it founds its family: its bit joins the
//                   interior founders' in the bitmap (one bit per read), next to the per-word sums of family
//                   reads and of families with a consensus row.
//   word_scan       one chained scan over the bitmap's words, three sums: founders before a word = family
//                   number in creation order, reads = start of the segment, SSCS families = consensus row.
//   tile_family     founders write the family records and the SSCS list.
//   tile_place      a representative reserves room for its tile's members in the family's segment with one
//                   atomic; the members write their places.
//   tile_join       duplex partner: found in the tile (interior) or by probing the table (edge families).
//
// Members land in `perm` in arbitrary order inside a family; nothing downstream depends on it
// (votes count, lone supporters are unique, mode ties are decided by read index).
// (A version that fused mark + numbering + records into one chained scan over the tiles ran at 0.43 ms
//  against 0.2 ms for the pieces: every CTA sat at a barrier while its first warp looked back, behind
//  predecessors that were themselves waiting for their gathers.  A version in which every tile
//  representative went to a table sized for n reads spent its time on DRAM: a random 8-byte slot costs a
//  128-byte fetch, 5 M of them per pass.)
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
constexpr int GT_TILE = 1024;                   // reads per tile (even: a pair never straddles tiles)
constexpr int GT_THREADS = GT_INSERT_THREADS_CFG;   // tile_local: threads per tile
constexpr int GT_PER = GT_TILE / GT_THREADS;    // ... and reads per thread
constexpr int GP_THREADS = 256;                 // tile_place
constexpr int GP_PER = GT_TILE / GP_THREADS;
constexpr uint32_t GT_HT = 2 * GT_TILE;         // shared table slots
constexpr uint32_t GT_EMPTY = 0xFFFFFFFFu;
constexpr uint32_t GT_MIXED = 0x80000000u;      // shared counter: members' mode fields differ
// link word of a read: a member keeps its representative (10 bits) | rank in the tile's group << 10;
// a representative keeps LINK_REP, plus LINK_LOCAL when the tile holds its whole family
constexpr uint32_t LINK_REP = 0x80000000u;
constexpr uint32_t LINK_LOCAL = 0x40000000u;
constexpr int LINK_RANK_SHIFT = 10;
static_assert(GT_TILE == 1 << LINK_RANK_SHIFT, "a member's link word holds two tile-local indices");
constexpr uint32_t STRETCH_SLOTS = 4096;        // contigs whose stretch start has been seen (open addressing)
constexpr int NB_NONE = -0x7FFFFFFF;            // "contig" of a neighbour that does not exist

// device-side scalars of the tile path (a block of uint32 in the context's counter buffer)
enum { GTC_EDGE = 0,        // listed (edge) representatives
       GTC_UNSORTED = 1,    // the order tile_local relies on does not hold
       GTC_MASK = 2,        // table slots - 1
       GTC_COUNT = 4 };

inline uint64_t gt_tiles(uint64_t n) { return (n + GT_TILE - 1) / GT_TILE; }

// counter word of a group of `members` reads, as a table slot keeps it (group_hash.cuh): starts at all ones,
// every member subtracts 2, bit 0 is cleared when mode fields differ
__device__ __forceinline__ uint32_t group_cnt(uint32_t members, bool mixed) { return (0xFFFFFFFFu - 2u * members) & (mixed ? ~1u : ~0u); }

// slot of a tuple in the tile's shared table: a short 32-bit mix is enough to spread one tile's families
// (the 64-bit rec_hash, 7 wide multiplies, is kept for the global table)
__device__ __forceinline__ uint32_t tile_hash(const uint4 &a, const uint4 &b)
{
    uint32_t h = a.x * 0x9E3779B1u ^ a.w * 0x85EBCA6Bu ^ b.y * 0xC2B2AE35u ^ a.y * 0x27D4EB2Fu;
    h ^= (a.z * 31u + b.x + key_bits(b.w) * 0x10001u) * 0x165667B1u;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return h & (GT_HT - 1);
}

// MODE 0: first pass, one tile per CTA.  MODE 1: only if the first pass found the input out of order; then nothing is
// local; launched with a small grid that walks the tiles, so it costs nothing when it has nothing to do.
template <int MODE>
__global__ void __launch_bounds__(GT_THREADS, 3)
tile_local_kernel(const uint4 *__restrict__ fields, uint64_t n, const uint32_t *__restrict__ dup_bc, uint32_t n_bc,
                  const uint32_t *__restrict__ asym, uint32_t *__restrict__ gtc, uint32_t *__restrict__ stretch,
                  uint32_t *__restrict__ link, uint32_t *__restrict__ aux, uint32_t *__restrict__ stash,
                  uint32_t *__restrict__ edge_list, uint32_t *__restrict__ bitmap, uint32_t *__restrict__ word_reads,
                  uint32_t *__restrict__ word_sscs)
{
    __shared__ uint4 key_a[GT_TILE], key_b[GT_TILE];      // the tile's records
    __shared__ uint32_t ht[GT_HT];                        // tuple -> smallest local index seen so far
    __shared__ uint32_t members[GT_TILE];                 // per representative: members besides itself | GT_MIXED
    __shared__ uint16_t rep_list[GT_TILE];
    __shared__ uint32_t n_reps, n_edge, edge_base;
    __shared__ int2 neighbour[2];
    __shared__ uint32_t bm[GT_TILE / 32], bm_reads[GT_TILE / 32], bm_sscs[GT_TILE / 32];   // founders known here, per bitmap word
    const int tid = threadIdx.x;
    constexpr int mode = MODE;
    if (mode == 1 && !gtc[GTC_UNSORTED]) return;
    const bool nothing_local = mode == 1 || *asym != 0;   // one-sided duplex partnerships are resolved through the table
    uint64_t tile = blockIdx.x;
    do {
    const uint64_t base = tile * GT_TILE;
    const uint32_t here = (uint32_t)min((uint64_t)GT_TILE, n - base);

    static_assert(GT_HT == 4 * GT_THREADS || GT_HT % (4 * GT_THREADS) == 0, "the shared table is cleared 16 bytes at a time");
    for (int k = tid; k < (int)GT_HT / 4; k += GT_THREADS)
        reinterpret_cast<uint4 *>(ht)[k] = make_uint4(GT_EMPTY, GT_EMPTY, GT_EMPTY, GT_EMPTY);
    if (tid == 0) { n_reps = 0; n_edge = 0; }
    if (tid < GT_TILE / 32) { bm[tid] = 0; bm_reads[tid] = 0; bm_sscs[tid] = 0; }
    uint32_t h[GT_PER];
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        members[li] = 0;
        if (li < here) {
            const Rec r = load_rec(fields, (uint32_t)(base + li));
            key_a[li] = r.a;
            key_b[li] = r.b;
            h[k] = tile_hash(r.a, r.b);
        }
    }
    static_assert(GT_THREADS >= 64, "two warps look for the tile's neighbours");
    // completion coordinates of the nearest one-contig pairs outside the tile, looked for by one warp each
    if (tid < 64) {
        const int lane = tid & 31;
        const bool fwd = tid >= 32;
        const int64_t pairs = (int64_t)(n >> 1), edge = fwd ? (int64_t)((base + GT_TILE) >> 1) : (int64_t)(base >> 1) - 1;
        const int64_t p = fwd ? edge + lane : edge - lane;
        bool exists = p >= 0 && p < pairs, intra = false;
        int2 at = make_int2(NB_NONE, 0);
        if (exists) {
            const Rec r = load_rec(fields, (uint32_t)(2 * p + 1));
            intra = (int)r.a.z == (int)r.b.x;
            at = make_int2((int)r.a.z, (int)r.a.w);
        }
        const uint32_t hit = __ballot_sync(0xffffffffu, intra), off_end = __ballot_sync(0xffffffffu, !exists);
        const int first = hit ? __ffs(hit) - 1 : 32, end_at = off_end ? __ffs(off_end) - 1 : 32;
        // nearest one-contig pair, or the end of the input before any; 32 pairs without either is not order we know
        if (first < end_at) { if (lane == first) neighbour[fwd] = at; }
        else if (end_at < 32) { if (lane == 0) neighbour[fwd] = make_int2(NB_NONE, 0); }
        else if (lane == 0) { neighbour[fwd] = make_int2(NB_NONE, 0); if (mode == 0) gtc[GTC_UNSORTED] = 1u; }
    }
    __syncthreads();
    const int2 before = neighbour[0], after = neighbour[1];

    // tile-local table: a slot ends up holding the smallest local index of its tuple
    uint32_t my_slot[GT_PER];
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li >= here) continue;
        const Rec r{key_a[li], key_b[li]};
        uint32_t s = h[k];
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
        // what this pass relies on, checked at the completing read of every one-contig pair
        if ((li & 1) && mode == 0 && (int)r.a.z == (int)r.b.x) {
            if ((int)r.a.w < (int)r.b.y) gtc[GTC_UNSORTED] = 1u;          // completes at its larger position
            int2 prev = before;                                             // the one-contig pair before this one
            for (uint32_t b = li; b >= 2;) {
                b -= 2;
                if ((int)key_a[b].z == (int)key_b[b].x) { prev = make_int2((int)key_a[b].z, (int)key_a[b].w); break; }
            }
            if ((int)r.a.z == prev.x) {
                if ((int)r.a.w < prev.y) gtc[GTC_UNSORTED] = 1u;          // positions never go backwards inside a contig
            } else {
                // a contig begins: it must be the contig's only stretch (contigs may come in any order)
                const uint32_t mark = (uint32_t)r.a.z + 2u;
                uint32_t s2 = (mark * 2654435761u >> 20) & (STRETCH_SLOTS - 1), tries = 0;
                for (;; s2 = (s2 + 1) & (STRETCH_SLOTS - 1)) {
                    const uint32_t seen = atomicCAS(&stretch[s2], 0u, mark);
                    if (seen == 0u) break;
                    if (seen == mark || ++tries == STRETCH_SLOTS) { gtc[GTC_UNSORTED] = 1u; break; }
                }
            }
        }
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
        }
    }
    // the representatives queue up: one addition to the shared counter per warp, not per representative
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
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
    uint32_t edge_at[(GT_TILE + GT_THREADS - 1) / GT_THREADS];
    int e = 0;
    for (uint32_t x = tid; x < reps; x += GT_THREADS, e++) {
        const uint32_t li = rep_list[x];
        const uint32_t i = (uint32_t)(base + li);
        const uint32_t m = members[li];
        const uint4 odd = key_a[li | 1];                          // the read that completed this pair
        const int2 at = make_int2((int)odd.z, (int)odd.w);
        const bool one_contig = (int)odd.z == (int)key_b[li | 1].x;
        const bool local = !nothing_local && one_contig && !(at.x == before.x && at.y == before.y) &&
                           !(at.x == after.x && at.y == after.y);
        stash[i] = group_cnt((m & ~GT_MIXED) + 1u, (m & GT_MIXED) != 0);
        edge_at[e] = GT_EMPTY;
        if (local) {
            link[i] = LINK_REP | LINK_LOCAL;
            const uint32_t size = (m & ~GT_MIXED) + 1u;            // the whole family: its founder is known right here
            atomicOr(&bm[li >> 5], 1u << (li & 31));
            atomicAdd(&bm_reads[li >> 5], size);
            if (size >= 2) atomicAdd(&bm_sscs[li >> 5], 1u);
            // duplex partner (duplex_tag, consensus_helper.py:605-648): same coordinates, so in this tile or nowhere
            uint32_t found = NO_ID;
            Rec q{key_a[li], key_b[li]};
            if (q.b.w & 0x40u) {                                   // R1 families look, R2 families are told (tile_join)
                const uint32_t d = q.a.x < n_bc ? __ldg(&dup_bc[q.a.x]) : NO_ID;
                if (d != NO_ID) {
                    q.a.x = d;
                    q.b.w ^= 0x40u;
                    uint32_t s = tile_hash(q.a, q.b);
                    for (;;) {
                        const uint32_t cur = ht[s];
                        if (cur == GT_EMPTY) break;
                        if (same_tag(q, Rec{key_a[cur], key_b[cur]})) { found = cur; break; }
                        s = (s + 1) & (GT_HT - 1);
                    }
                }
            }
            aux[i] = found;
        } else {
            link[i] = LINK_REP;
            edge_at[e] = atomicAdd(&n_edge, 1u);
        }
    }
    __syncthreads();
    if (tid == 0 && n_edge) edge_base = atomicAdd(&gtc[GTC_EDGE], n_edge);
    __syncthreads();
    e = 0;
    for (uint32_t x = tid; x < reps; x += GT_THREADS, e++)
        if (edge_at[e] != GT_EMPTY) { CC_CHECK(edge_base + edge_at[e] < n); edge_list[edge_base + edge_at[e]] = (uint32_t)(base + rep_list[x]); }
    if (tid < GT_TILE / 32 && (uint32_t)tid * 32 < here) {     // edge founders are added by edge_mark_kernel
        bitmap[(base >> 5) + tid] = bm[tid];
        word_reads[(base >> 5) + tid] = bm_reads[tid];
        word_sscs[(base >> 5) + tid] = bm_sscs[tid];
    }
#pragma unroll
    for (int k = 0; k < GT_PER; k++) {
        const uint32_t li = k * GT_THREADS + tid;
        if (li < here && lrep[k] != li) link[base + li] = lrep[k] | (rank[k] << LINK_RANK_SHIFT);
    }
    if (MODE == 1) __syncthreads();                        // the shared arrays are free for the next tile
    } while (MODE == 1 && (tile += gridDim.x) * GT_TILE < n);
}

// start the second pass from an empty list (only if the first pass found the input out of order)
__global__ void tile_redo_kernel(uint32_t *gtc)
{
    if (gtc[GTC_UNSORTED]) gtc[GTC_EDGE] = 0u;
}

// table for the listed representatives: 2 x their number, rounded up to a power of two, cleared
__global__ void __launch_bounds__(256)
table_prepare_kernel(uint32_t *gtc, GroupSlot *__restrict__ table, uint64_t max_slots)
{
    uint64_t slots = 1024;
    const uint64_t want = 2ull * gtc[GTC_EDGE];
    while (slots < want && slots < max_slots) slots <<= 1;
    uint2 *t = reinterpret_cast<uint2 *>(table);
    for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s < slots; s += (uint64_t)gridDim.x * blockDim.x)
        t[s] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
    if (blockIdx.x == 0 && threadIdx.x == 0) gtc[GTC_MASK] = (uint32_t)(slots - 1);   // read by the kernels that follow
}

// one thread per listed representative (group_insert_kernel, with the tile's whole group in one trip)
__global__ void __launch_bounds__(256)
edge_insert_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ edge_list, const uint32_t *__restrict__ gtc,
                   GroupSlot *__restrict__ table, const uint32_t *__restrict__ stash, uint32_t *__restrict__ aux)
{
    const uint32_t total = gtc[GTC_EDGE];
    const uint64_t mask = gtc[GTC_MASK];
    for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x) {
        const uint32_t i = edge_list[x];
        CC_CHECK((i & 1u) < 2u && x < total);
        const Rec r = load_rec(fields, i);
        const uint32_t mine = stash[i];                          // group_cnt of the tile's group
        bool mixed = !(mine & 1u);
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
        CC_CHECK(slot <= mask);
        atomicSub(&table[slot].cnt, (~mine) & ~1u);              // 2 x members of the tile's group
        if (mixed) atomicAnd(&table[slot].cnt, ~1u);
        aux[i] = (uint32_t)slot;
    }
}

// (after every listed representative is in the table) a listed representative whose slot still names it
// founds its family: its bit, its family's reads and class are added to the words tile_local wrote.
// aux[] turns from the slot into the family's founder; a founder keeps the family's counter word in stash[].
__global__ void __launch_bounds__(256)
edge_mark_kernel(const uint32_t *__restrict__ edge_list, const uint32_t *__restrict__ gtc, const GroupSlot *__restrict__ table,
                 uint32_t *__restrict__ aux, uint32_t *__restrict__ stash, uint32_t *__restrict__ bitmap,
                 uint32_t *__restrict__ word_reads, uint32_t *__restrict__ word_sscs)
{
    const uint32_t total = gtc[GTC_EDGE];
    for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x) {
        const uint32_t i = edge_list[x];
        const uint2 s = *reinterpret_cast<const uint2 *>(&table[aux[i]]);
        aux[i] = s.x;                                             // tile_place asks for the family of this read
        if (s.x == i) {
            const uint32_t size = (~s.y) >> 1;
            stash[i] = s.y;
            atomicOr(&bitmap[i >> 5], 1u << (i & 31));
            atomicAdd(&word_reads[i >> 5], size);
            if (size >= 2) atomicAdd(&word_sscs[i >> 5], 1u);
        }
    }
}

// One chained scan over the bitmap's words carries three sums: founders (family numbers in creation order),
// reads of their families (segment starts) and families of two or more reads (SSCS rows).  word_reads and
// word_sscs are scanned in place.
constexpr int WS_THREADS = 256, WS_ITEMS = 8, WS_TILE = WS_THREADS * WS_ITEMS;

__global__ void __launch_bounds__(WS_THREADS)
word_scan_kernel(const uint32_t *__restrict__ bitmap, uint32_t *__restrict__ word_rank, uint32_t *word_reads,
                 uint32_t *word_sscs, uint64_t words, unsigned long long *status, uint32_t *ticket,
                 uint32_t *__restrict__ n_fam, uint32_t *__restrict__ n_sscs, uint32_t scale = 1)
{
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp[WS_THREADS / 32][3];
    __shared__ uint32_t s_excl[4];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t base = (uint64_t)tile * WS_TILE + (uint64_t)tid * WS_ITEMS;
    uint32_t v[3][WS_ITEMS];
    if (base + WS_ITEMS <= words) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const uint32_t *src = a == 0 ? bitmap : a == 1 ? word_reads : word_sscs;
            const uint4 x = *reinterpret_cast<const uint4 *>(src + base), y = *reinterpret_cast<const uint4 *>(src + base + 4);
            v[a][0] = x.x; v[a][1] = x.y; v[a][2] = x.z; v[a][3] = x.w; v[a][4] = y.x; v[a][5] = y.y; v[a][6] = y.z; v[a][7] = y.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < WS_ITEMS; k++) {
            const bool in = base + k < words;
            v[0][k] = in ? bitmap[base + k] : 0u; v[1][k] = in ? word_reads[base + k] : 0u; v[2][k] = in ? word_sscs[base + k] : 0u;
        }
    }
    uint32_t own[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < WS_ITEMS; k++) { v[0][k] = __popc(v[0][k]); own[0] += v[0][k]; own[1] += v[1][k]; own[2] += v[2][k]; }
    uint32_t inc[3] = {own[0], own[1], own[2]};
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc[a], d);
            if (lane >= d) inc[a] += t;
        }
    }
    if (lane == 31) { s_warp[warp][0] = inc[0]; s_warp[warp][1] = inc[1]; s_warp[warp][2] = inc[2]; }
    __syncthreads();
    uint32_t off[3] = {0, 0, 0}, tot[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < WS_THREADS / 32; k++) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            if (k < warp) off[a] += s_warp[k][a];
            tot[a] += s_warp[k][a];
        }
    }
    if (warp == 0) {
        uint32_t excl[4];
        lookback_pairs<2>(status, tile, tot, excl, lane);
        if (lane == 0) {
            s_excl[0] = excl[0]; s_excl[1] = excl[1]; s_excl[2] = excl[2];
            if (tile == gridDim.x - 1) { *n_fam = scale * (excl[0] + tot[0]); *n_sscs = scale * (excl[2] + tot[2]); }   // (scale 2: the words count pair families)
        }
    }
    __syncthreads();
    uint32_t run[3];
#pragma unroll
    for (int a = 0; a < 3; a++) run[a] = s_excl[a] + off[a] + inc[a] - own[a];
    uint32_t o[3][WS_ITEMS];
#pragma unroll
    for (int k = 0; k < WS_ITEMS; k++)
#pragma unroll
        for (int a = 0; a < 3; a++) { o[a][k] = run[a]; run[a] += v[a][k]; }
    if (base + WS_ITEMS <= words) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            uint32_t *dst = a == 0 ? word_rank : a == 1 ? word_reads : word_sscs;
            *reinterpret_cast<uint4 *>(dst + base) = make_uint4(o[a][0], o[a][1], o[a][2], o[a][3]);
            *reinterpret_cast<uint4 *>(dst + base + 4) = make_uint4(o[a][4], o[a][5], o[a][6], o[a][7]);
        }
    } else {
#pragma unroll
        for (int k = 0; k < WS_ITEMS; k++)
            if (base + k < words) { word_rank[base + k] = o[0][k]; word_reads[base + k] = o[1][k]; word_sscs[base + k] = o[2][k]; }
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
                   uint32_t *__restrict__ n_mid, const uint32_t *__restrict__ word_sscs_start,
                   uint32_t *__restrict__ sscs_list, int32_t *__restrict__ sscs_row)
{
    __shared__ uint32_t small_hist[256];
    small_hist[threadIdx.x] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    // grid-stride over words, a warp takes FW words at a time and issues all their loads before it uses any:
    // the kernel is bound by the latency of the founders' two gathers (counter word, template record), not by bytes
    constexpr int FW = 4;
    const uint64_t words = (n + 31) / 32;
    const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t w0 = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * FW; w0 < words; w0 += warps * FW) {
        uint32_t word[FW], cnt[FW];
        uint4 meta[FW];
#pragma unroll
        for (int k = 0; k < FW; k++) word[k] = w0 + k < words ? __ldg(&bitmap[w0 + k]) : 0u;
#pragma unroll
        for (int k = 0; k < FW; k++) {
            const bool founder = (word[k] >> lane) & 1u;
            const uint64_t i = (w0 + k) * 32 + lane;
            cnt[k] = founder ? founder_cnt[i] : 0xFFFFFFFFu;
            meta[k] = founder ? __ldg(&fields[2 * i + 1]) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int k = 0; k < FW; k++) {
            if (word[k] == 0) continue;
            const uint64_t w = w0 + k, i = w * 32 + lane;
            const bool founder = (word[k] >> lane) & 1u;
            const uint32_t size = (~cnt[k]) >> 1;
            uint32_t incl = size;                                 // starts of the word's families: prefix of their sizes
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            const uint32_t big = __ballot_sync(0xffffffffu, size >= 2);
            if (!founder) continue;
            const uint32_t c = __ldg(&word_rank[w]) + __popc(word[k] & ((1u << lane) - 1u));
            const uint32_t start = __ldg(&word_start[w]) + incl - size;
            int32_t row = -1;                                      // families of two or more reads own a consensus row
            if (size >= 2) {
                row = (int32_t)(__ldg(&word_sscs_start[w]) + __popc(big & ((1u << lane) - 1u)));
                sscs_list[row] = c;
            }
            sscs_row[c] = row;
            CC_CHECK(c < n && start < n && (uint64_t)start + size <= n && i < n);
            tmpl[c] = (uint32_t)i;
            fam_size[c] = size;
            fam_start[c] = start;
            cursor[c] = start;
            is_sscs[c] = size >= 2 ? ((cnt[k] & 1u) ? 1 : 3) : 0;  // bit 0: has a consensus row, bit 1: mode fields differ
            fam_tlen[c] = (int32_t)meta[k].z;
            fam_flag[c] = (uint16_t)(meta[k].w & 0xFFFF);
            fam_mapq[c] = (uint8_t)((meta[k].w >> 16) & 0xFF);
            fam_rg[c] = (uint8_t)(meta[k].w >> 24);
            partner[c] = -1;
            const uint32_t bin = size < hist_bins - 1 ? size : hist_bins - 1;
            if (bin < 256) atomicAdd(&small_hist[bin], 1u);
            else atomicAdd(&hist[bin], 1ull);
        }
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
tile_place_kernel(uint64_t n, const uint32_t *__restrict__ link, const uint32_t *__restrict__ aux,
                  const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                  uint32_t *__restrict__ cursor, uint32_t *__restrict__ perm)
{
    __shared__ uint32_t members[GT_TILE];                  // per representative: members besides itself, then its segment base
    __shared__ uint32_t t_bits[GT_TILE / 32], t_rank[GT_TILE / 32];     // the tile's own words of the founders' bitmap
    const int tid = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * GT_TILE;
    const uint32_t here = (uint32_t)min((uint64_t)GT_TILE, n - base);
    if (tid < GT_TILE / 32 && (uint32_t)tid * 32 < here) {
        t_bits[tid] = bitmap[(base >> 5) + tid];
        t_rank[tid] = word_rank[(base >> 5) + tid];
    }
    uint32_t lk[GP_PER], fam[GP_PER];
#pragma unroll
    for (int k = 0; k < GP_PER; k++) {
        const uint32_t li = k * GP_THREADS + tid;
        members[li] = 0;
        lk[k] = li < here ? link[base + li] : 0u;
        fam[k] = 0;
        // an edge representative's trip to the founder's bitmap word starts before the members are counted
        if (li < here && (lk[k] & LINK_REP) && !(lk[k] & LINK_LOCAL)) fam[k] = family_of_founder(bitmap, word_rank, aux[base + li]);
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
        if (li >= here || !(lk[k] & LINK_REP)) continue;
        if (lk[k] & LINK_LOCAL) {
            // the tile holds the whole family: its segment is this group's, nobody else asks for room in it
            const uint32_t c = t_rank[li >> 5] + __popc(t_bits[li >> 5] & ((1u << (li & 31)) - 1u));
            members[li] = cursor[c];
        } else {
            members[li] = atomicAdd(&cursor[fam[k]], members[li] + 1u);     // room for the tile's whole group, one atomic
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GP_PER; k++) {
        const uint32_t li = k * GP_THREADS + tid;
        if (li >= here) continue;
        const uint32_t at = (lk[k] & LINK_REP) ? members[li] : members[lk[k] & (GT_TILE - 1)] + (lk[k] >> LINK_RANK_SHIFT);
        CC_CHECK(at < n);
        perm[at] = (uint32_t)(base + li);
    }
}

// The duplex join (duplex_tag, consensus_helper.py:605-648).  An interior family already knows the
// tile-local index of its partner's founder; an edge family probes the table for the complementary tuple
// (group_hash.cuh's group_join_kernel).  Symmetric / asymmetric as join_probe_kernel: with *asym set
// nothing was treated as interior and every family probes.
__global__ void __launch_bounds__(256)
tile_join_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ tmpl, const uint32_t *__restrict__ n_fam_ptr,
                 const uint32_t *__restrict__ dup_bc, uint32_t n_bc, const GroupSlot *__restrict__ table,
                 const uint32_t *__restrict__ gtc, const uint32_t *__restrict__ link, const uint32_t *__restrict__ aux,
                 const uint32_t *__restrict__ bitmap, const uint32_t *__restrict__ word_rank,
                 int32_t *__restrict__ partner, const uint32_t *__restrict__ asym, const uint16_t *__restrict__ fam_flag)
{
    const uint32_t n_fam = *n_fam_ptr;              // the family count lives on the device: a grid-stride loop up to it
    const bool symmetric = !*asym;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < n_fam; c += gridDim.x * blockDim.x) {
    // the read number is part of the family key, so the family's flag (a coalesced 2-byte load) tells R1 from R2
    // without a trip to the template's record
    if (symmetric && !(fam_flag[c] & 0x40u)) continue;   // an R2 family hears from its R1 partner (partner[] starts at -1)
    const uint32_t first = tmpl[c];
    int32_t found = -1;
    if (__ldg(&link[first]) & LINK_LOCAL) {
        const uint32_t p = __ldg(&aux[first]);
        CC_CHECK(p == NO_ID || p < (uint32_t)GT_TILE);
        if (p != NO_ID) found = (int32_t)family_of_founder(bitmap, word_rank, (first & ~(uint32_t)(GT_TILE - 1)) + p);
    } else {
        Rec r = load_rec(fields, first);
        const uint64_t mask = gtc[GTC_MASK];
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
    }
    CC_CHECK(found < (int32_t)n_fam);
    partner[c] = found;
    if (symmetric && found >= 0) partner[found] = (int32_t)c;
    }
}

}  // namespace ccgpu
