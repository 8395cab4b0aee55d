// Grouping reads into tag families with a hash table instead of a sort (sm_100a).
//
// read_bam's dictionaries (consensus_helper.py:447-475: `tag in read_dict`, append, size + 1) map
// directly onto an open-addressing table keyed by the exact family tuple.  A slot holds the
// smallest read index seen so far of its family (any member identifies the tuple, the smallest is
// the family's first read = template) and a member counter.  Reads of a family sit next to each
// other in a coordinate-sorted input, so after the first member the slot, the template's record and
// the counter are cache hits; what reaches DRAM is about one probe per family, not per read.
//
//   insert   every read finds / claims its slot, keeps the slot's read index minimal, counts itself
//   mark     a read that is still its slot's index is the family's first read: one bit per read; every
//            read notes its first read
//   (popcount scan of the bitmap: family number = rank of the first read = creation order)
//   family   first reads write the family records and their own read -> family entry
//   (exclusive scan of the sizes: segment starts)
//   place    every read copies its first read's family number and takes the next free place of the
//            family's segment (atomic cursor per family)
//
// Places come from an atomic cursor, so the order of a family's members in `perm` is arbitrary;
// everything downstream is independent of it (votes count, lone supporters are unique, mode ties
// are decided by read index - see group_mode / warp_mode).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.cuh"

namespace ccgpu {

struct GroupSlot {
    uint32_t rep;       // smallest member index so far (0xFFFFFFFF = empty); in the end the family's first read
    uint32_t cnt;       // starts at 0xFFFFFFFF (the table is cleared to 0xFF bytes); every member subtracts 2,
                        // so the size is (~cnt) >> 1; bit 0 is cleared once two members are seen to differ
                        // in tlen / flag / mapq / RG (the family needs real modes) - that never borrows
};

__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// (Two and four reads per thread, first probes issued together, measured slower: 0.53 / 0.78 ms against
// 0.50 ms - the kernel is bound by the number of dependent trips to L2 / DRAM in flight, which the
// 52 resident warps per SM already saturate.)
__global__ void __launch_bounds__(256)
group_insert_kernel(const uint4 *__restrict__ fields, uint64_t n, GroupSlot *__restrict__ table, uint64_t mask,
                    uint32_t *__restrict__ read_slot)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Rec r = load_rec(fields, (uint32_t)i);
    uint64_t slot = (rec_hash(r) * 0xD6E8FEB86659FD93ull >> 20) & mask;
    for (;;) {
        uint32_t rep = ld_relaxed_u32(&table[slot].rep);
        if (rep == 0xFFFFFFFFu) {
            rep = atomicCAS(&table[slot].rep, 0xFFFFFFFFu, (uint32_t)i);
            if (rep == 0xFFFFFFFFu) break;                      // claimed: this read founds the family
        }
        const Rec q = load_rec(fields, rep);
        if (same_tag(r, q)) {                                    // any member carries the family's tuple
            if ((uint32_t)i < rep) atomicMin(&table[slot].rep, (uint32_t)i);
            // every member but the founder is compared with an earlier one: the comparisons span the
            // family, so no difference on any of them means the mode fields are uniform
            if (r.b.z != q.b.z || r.b.w != q.b.w) atomicAnd(&table[slot].cnt, ~1u);
            break;
        }
        slot = (slot + 1) & mask;
    }
    read_slot[i] = (uint32_t)slot;
    atomicSub(&table[slot].cnt, 2u);                             // result unused: a reduction, nothing to wait for
}

__global__ void __launch_bounds__(256)
group_mark_kernel(const GroupSlot *__restrict__ table, uint32_t *__restrict__ read_slot, uint64_t n,
                  uint32_t *__restrict__ bitmap, uint32_t *__restrict__ first_of)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t rep = 0xFFFFFFFFu;
    if (i < n) {
        // one 8-byte load of the slot; kept per read so that the later kernels need not probe again:
        // first_of[i] = the family's first read, read_slot[i] := the slot's final counter
        const uint2 s = *reinterpret_cast<const uint2 *>(&table[read_slot[i]]);
        first_of[i] = rep = s.x;
        read_slot[i] = s.y;
    }
    const uint32_t word = __ballot_sync(0xffffffffu, i < n && rep == (uint32_t)i);   // 32 consecutive reads = one bitmap word
    if ((threadIdx.x & 31) == 0 && i < n) bitmap[i >> 5] = word;
}

// family records in creation order (what family_build_kernel writes on the sort path)
__global__ void __launch_bounds__(256)
group_family_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ slot_cnt, uint64_t n,
                    const uint32_t *__restrict__ bitmap,
                    const uint32_t *__restrict__ word_rank, uint32_t *__restrict__ read_fam, uint32_t *__restrict__ tmpl,
                    uint32_t *__restrict__ fam_size, uint8_t *__restrict__ is_sscs, uint8_t *__restrict__ fam_mapq,
                    uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
                    unsigned long long *__restrict__ hist, uint32_t hist_bins, uint32_t *__restrict__ n_mid,
                    int32_t *__restrict__ partner)
{
    __shared__ uint32_t small_hist[256];           // sizes < 256 are counted per block first
    small_hist[threadIdx.x] = 0;
    __syncthreads();
    // grid-stride: a few CTAs per SM, so the shared histogram is flushed a few thousand times in all
    // (the global bins are a handful of addresses; one flush per 256 reads would serialise on them)
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t w = bitmap[i >> 5], bit = 1u << (i & 31);
        if (!(w & bit)) continue;
        const uint32_t c = word_rank[i >> 5] + __popc(w & (bit - 1u));
        const uint32_t cnt = slot_cnt[i];                    // left by group_mark_kernel
        const uint32_t size = (~cnt) >> 1;
        read_fam[i] = c;                                     // replaces first_of[i] == i; the other members copy it (group_place_kernel)
        tmpl[c] = (uint32_t)i;
        fam_size[c] = size;
        is_sscs[c] = size >= 2 ? ((cnt & 1u) ? 1 : 3) : 0;   // bit 0: has a consensus row, bit 1: mode fields differ
        const uint4 b = __ldg(&fields[2 * i + 1]);
        fam_tlen[c] = (int32_t)b.z;
        fam_flag[c] = (uint16_t)(b.w & 0xFFFF);
        fam_mapq[c] = (uint8_t)((b.w >> 16) & 0xFF);
        fam_rg[c] = (uint8_t)(b.w >> 24);
        partner[c] = -1;                                     // until the join says otherwise
        const uint32_t bin = size < hist_bins - 1 ? size : hist_bins - 1;
        if (bin < 256) atomicAdd(&small_hist[bin], 1u);
        else atomicAdd(&hist[bin], 1ull);
    }
    __syncthreads();
    const uint32_t v = small_hist[threadIdx.x];
    if (v) atomicAdd(&hist[threadIdx.x], (unsigned long long)v);
    if (v && threadIdx.x >= 16 && threadIdx.x < 128) atomicAdd(n_mid, v);   // families of 16..127 reads
}

__global__ void __launch_bounds__(256)
group_place_kernel(const uint32_t *__restrict__ bitmap, uint64_t n, uint32_t *__restrict__ cursor,
                   uint32_t *__restrict__ perm, uint32_t *read_fam)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // read_fam[] holds the family number at first reads (group_family_kernel) and still the first read's
    // index everywhere else (group_mark_kernel); only the latter entries are rewritten here
    uint32_t c = read_fam[i];
    if (!((bitmap[i >> 5] >> (i & 31)) & 1u)) {
        c = read_fam[c];
        read_fam[i] = c;
    }
    perm[atomicAdd(&cursor[c], 1u)] = (uint32_t)i;             // cursor[] starts as a copy of fam_start[] (a few MB: L2)
}

// The duplex join on the grouping table: a family's partner is the family of the complementary tuple
// (duplex_tag, consensus_helper.py:605-648), and that tuple - like every family's - has a slot here that
// names its first read.  No second table is built.  Symmetric / asymmetric as join_probe_kernel.
__global__ void __launch_bounds__(256)
group_join_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ tmpl, uint32_t n_fam,
                  const uint32_t *__restrict__ dup_bc, uint32_t n_bc, const GroupSlot *__restrict__ table, uint64_t mask,
                  const uint32_t *__restrict__ read_fam, int32_t *__restrict__ partner, const uint32_t *__restrict__ asym)
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
            if (same_tag(load_rec(fields, rep), q)) { found = (int32_t)read_fam[rep]; break; }
            slot = (slot + 1) & mask;
        }
    }
    partner[c] = found;
    if (symmetric && found >= 0) partner[found] = (int32_t)c;
}

}  // namespace ccgpu
