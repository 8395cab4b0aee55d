// Device kernels of the consensus path (sm_100a).  Each kernel names the
// reference code it replaces (paths relative to the reference repository).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "scan.cuh"

// Bounds checks of our own (compute-sanitizer is not available on the GPU pool): a library built with
// -DCCGPU_CHECKS traps (device assert -> the next CUDA call fails) when a kernel is about to index outside an
// array; scripts/gpu_checks.sh runs the parity suite and the soak cases through such a build.
#ifdef CCGPU_CHECKS
#include <cassert>
#define CC_CHECK(cond) assert(cond)
#else
#define CC_CHECK(cond) ((void)0)
#endif

namespace ccgpu {

constexpr uint32_t NO_ID = 0xFFFFFFFFu;
constexpr int Q_SSCS = 30;   // SSCS_maker.py:113  "query_qualities[i] < 30" fails a base
constexpr int Q_SC = 29;     // singleton_correction.py:72  both qualities "> 29"
constexpr int Q_CAP = 60;    // SSCS_maker.py:134, DCS_maker.py:113  quality ceiling

// ---- field records -------------------------------------------------------------
// 32-byte record = two 16-byte halves: a = (bc, cig, tid, pos), b = (mtid, mpos, tlen, meta)
// with meta = flag | mapq << 16 | rg << 24.
struct Rec { uint4 a, b; };

__device__ __forceinline__ Rec load_rec(const uint4 *__restrict__ fields, uint32_t i)
{
    // one 256-bit load (sm_100: LDG.E.256): a gathered record is one sector and one LSU request
    Rec r;
    asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(r.a.x), "=r"(r.a.y), "=r"(r.a.z), "=r"(r.a.w), "=r"(r.b.x), "=r"(r.b.y), "=r"(r.b.z), "=r"(r.b.w)
        : "l"(fields + 2 * (uint64_t)i));
    return r;
}

// orientation (flag 0x10) and read number (flag 0x40) of unique_tag, consensus_helper.py:286-290
__device__ __forceinline__ uint32_t key_bits(uint32_t meta) { return ((meta >> 4) & 1u) | (((meta >> 6) & 1u) << 1); }

__device__ __forceinline__ bool same_tag(const Rec &x, const Rec &y)
{
    return x.a.x == y.a.x && x.a.y == y.a.y && x.a.z == y.a.z && x.a.w == y.a.w &&
           x.b.x == y.b.x && x.b.y == y.b.y && key_bits(x.b.w) == key_bits(y.b.w);
}

__device__ __forceinline__ uint64_t mix64(uint64_t h, uint32_t w)
{
    h = (h ^ w) * 0x9E3779B97F4A7C15ull;
    return h ^ (h >> 32);
}

// 64-bit hash of the unique_tag tuple; low word = sort key, high word = join tag
__device__ __forceinline__ uint64_t tag_hash(uint32_t bc, uint32_t cig, uint32_t tid, uint32_t pos,
                                             uint32_t mtid, uint32_t mpos, uint32_t bits)
{
    uint64_t h = 0x243F6A8885A308D3ull;
    h = mix64(h, bc); h = mix64(h, cig); h = mix64(h, tid); h = mix64(h, pos);
    h = mix64(h, mtid); h = mix64(h, mpos); h = mix64(h, bits);
    h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
    return h;
}

__device__ __forceinline__ uint64_t rec_hash(const Rec &r)
{
    return tag_hash(r.a.x, r.a.y, r.a.z, r.a.w, r.b.x, r.b.y, key_bits(r.b.w));
}

// ---- K1: sort key of every read (unique_tag, consensus_helper.py:293-301) ---------
__global__ void __launch_bounds__(256)
hash_kernel(const uint4 *__restrict__ fields, uint64_t n, uint32_t key_mask, uint32_t *__restrict__ keys)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Rec r = load_rec(fields, (uint32_t)i);
    keys[i] = (uint32_t)rec_hash(r) & key_mask;
}

// ---- K3: exact family boundaries in sorted order -------------------------------------
// head[j] bit 0 set iff read perm[j] starts a family.  Reads with equal sort key but
// different tags ("mixed runs", hash collisions) are queued for fix_mixed_runs.
__global__ void __launch_bounds__(256)
boundary_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ keys,
                const uint32_t *__restrict__ perm, uint64_t n, uint8_t *__restrict__ head,
                uint32_t *__restrict__ mixed_list, uint32_t *__restrict__ mixed_count, uint32_t mixed_cap)
{
    uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    if (j == 0) { head[0] = 1; return; }
    Rec cur = load_rec(fields, perm[j]);
    Rec prv = load_rec(fields, perm[j - 1]);
    bool differs = !same_tag(cur, prv);
    // bit 0: starts a family; bit 1: same family, but tlen / flag / mapq / RG differ from the previous member
    head[j] = differs ? 1 : ((cur.b.z != prv.b.z || cur.b.w != prv.b.w) ? 2 : 0);
    uint32_t k = keys[j];
    if (differs && k == keys[j - 1]) {
        // only the first tag change inside a run reports the run
        uint64_t s = j - 1;
        bool first = true;
        while (s > 0 && keys[s - 1] == k) {
            Rec p2 = load_rec(fields, perm[s - 1]);
            if (!same_tag(prv, p2)) { first = false; break; }
            prv = p2;
            --s;
        }
        if (first) {
            uint32_t slot = atomicAdd(mixed_count, 1u);
            if (slot < mixed_cap) mixed_list[slot] = (uint32_t)s;
        }
    }
}

// One warp per mixed run: stable regrouping by tag, in order of first appearance.
__global__ void __launch_bounds__(128)
fix_mixed_runs_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ keys,
                      uint32_t *__restrict__ perm, uint32_t *__restrict__ tmp, uint64_t n,
                      uint8_t *__restrict__ head, const uint32_t *__restrict__ mixed_list,
                      const uint32_t *__restrict__ n_mixed_ptr, uint32_t mixed_cap)
{
    const int lane = threadIdx.x & 31;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_mixed = min(*n_mixed_ptr, mixed_cap);
    for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < n_mixed; w += warps) {
    const uint64_t s = mixed_list[w];
    const uint32_t k = keys[s];
    uint64_t e = s + 1;
    while (e < n && keys[e] == k) ++e;            // every lane walks the same (short) run
    const uint32_t lt = (1u << lane) - 1u;
    uint64_t out = s;
    while (out < e) {
        // first read not yet regrouped
        uint64_t first = e;
        for (uint64_t k0 = s; k0 < e && first == e; k0 += 32) {
            uint64_t x = k0 + lane;
            bool open = x < e && !(perm[x] >> 31);
            uint32_t b = __ballot_sync(0xffffffffu, open);
            if (b) first = k0 + __ffs(b) - 1;
        }
        Rec t = load_rec(fields, perm[first]);
        uint64_t head_at = out;
        for (uint64_t k0 = s; k0 < e; k0 += 32) {
            uint64_t x = k0 + lane;
            bool take = false;
            uint32_t v = 0;
            if (x < e) {
                v = perm[x];
                if (!(v >> 31)) take = same_tag(load_rec(fields, v), t);
            }
            uint32_t b = __ballot_sync(0xffffffffu, take);
            if (take) {
                uint64_t dst = out + __popc(b & lt);
                tmp[dst] = v;
                head[dst] = dst == head_at ? 1 : 2;      // 2: let the vote kernel take the exact mode path
                perm[x] = v | 0x80000000u;
            }
            out += __popc(b);
        }
        __syncwarp();
    }
    for (uint64_t x = s + lane; x < e; x += 32) perm[x] = tmp[x];
    __syncwarp();
    }
}

// ---- families in creation order -----------------------------------------------------
// A family is created by its first member (stable sort => the first element of its segment).
// One bit per read marks the first members; a family's creation number is the number of marks
// below its own, i.e. a rank query in that bitmap (N / 8 bytes, resident in L2).
__global__ void __launch_bounds__(256)
mark_first_kernel(const uint32_t *__restrict__ hs_start, const uint32_t *__restrict__ perm,
                  uint32_t n_fam, uint32_t *__restrict__ bitmap, uint32_t *__restrict__ first_read)
{
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_fam) return;
    uint32_t first = perm[hs_start[f]];
    first_read[f] = first;
    atomicOr(&bitmap[first >> 5], 1u << (first & 31));
}

// c2hs[creation number of sorted family f] = f
__global__ void __launch_bounds__(256)
creation_rank_kernel(const uint32_t *__restrict__ first_read, const uint32_t *__restrict__ bitmap,
                     const uint32_t *__restrict__ word_rank, uint32_t n_fam, uint32_t *__restrict__ c2hs)
{
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_fam) return;
    uint32_t first = first_read[f];
    uint32_t c = word_rank[first >> 5] + __popc(bitmap[first >> 5] & ((1u << (first & 31)) - 1u));
    c2hs[c] = f;
}

// family c in creation order: segment of perm, size, template, singleton fields,
// "mode fields differ" flag, size histogram
__global__ void __launch_bounds__(256)
family_build_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ hs_start,
                    const uint32_t *__restrict__ first_read, const uint32_t *__restrict__ c2hs,
                    const uint8_t *__restrict__ head, uint32_t n_fam,
                    uint32_t n_reads, uint32_t *__restrict__ tmpl, uint32_t *__restrict__ fam_start,
                    uint32_t *__restrict__ fam_size, uint8_t *__restrict__ is_sscs,
                    uint8_t *__restrict__ fam_mapq, uint16_t *__restrict__ fam_flag,
                    int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
                    unsigned long long *__restrict__ hist, uint32_t hist_bins, uint32_t *__restrict__ n_mid,
                    int32_t *__restrict__ partner)
{
    __shared__ uint32_t small_hist[256];           // sizes < 256 are counted per block first
    small_hist[threadIdx.x] = 0;
    __syncthreads();
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n_fam) {
        uint32_t f = c2hs[c];
        uint32_t first = first_read[f];
        uint32_t s = hs_start[f];
        uint32_t e = (f + 1 < n_fam) ? hs_start[f + 1] : n_reads;
        uint32_t size = e - s;
        uint32_t mixed = 0;                           // do the members' mode fields differ anywhere?
        for (uint32_t j = s + 1; j < e; j++) mixed |= head[j];
        tmpl[c] = first;
        fam_start[c] = s;
        fam_size[c] = size;
        is_sscs[c] = size >= 2 ? ((mixed & 2) ? 3 : 1) : 0;   // bit 0: has a consensus row, bit 1: mode fields differ
        // a singleton keeps its own fields (SSCS_maker.py:278-282); larger families with differing
        // fields are overwritten by the vote kernel's mode fields
        uint4 b = __ldg(&fields[2 * (uint64_t)first + 1]);
        fam_tlen[c] = (int32_t)b.z;
        fam_flag[c] = (uint16_t)(b.w & 0xFFFF);
        fam_mapq[c] = (uint8_t)((b.w >> 16) & 0xFF);
        fam_rg[c] = (uint8_t)(b.w >> 24);
        partner[c] = -1;                              // until the join says otherwise
        uint32_t bin = size < hist_bins - 1 ? size : hist_bins - 1;
        if (bin < 256) atomicAdd(&small_hist[bin], 1u);
        else atomicAdd(&hist[bin], 1ull);
    }
    __syncthreads();
    uint32_t v = small_hist[threadIdx.x];
    if (v) atomicAdd(&hist[threadIdx.x], (unsigned long long)v);
    if (v && threadIdx.x >= 16 && threadIdx.x < 128) atomicAdd(n_mid, v);   // families of 16..127 reads
}

// Pair symmetry check (SSCS_maker.py:275 emits a pair name only with exactly two tags):
// the mate of a family's template must be the template of the neighbouring family, of equal size.
__global__ void __launch_bounds__(256)
domain_check_kernel(const uint32_t *__restrict__ tmpl, const uint32_t *__restrict__ fam_size,
                    const uint32_t *__restrict__ n_fam_ptr, uint32_t *__restrict__ faults)
{
    const uint32_t n_fam = *n_fam_ptr;              // on the device: a grid-stride loop up to it
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < n_fam; c += gridDim.x * blockDim.x) {
        uint32_t o = c ^ 1u;
        bool ok = o < n_fam && tmpl[o] == (tmpl[c] ^ 1u) && fam_size[o] == fam_size[c];
        if (!ok) atomicAdd(faults, 1u);
    }
}

// read -> family (only materialised when a caller fetches CC_READ_FAM)
__global__ void __launch_bounds__(256)
read_fam_kernel(const uint32_t *__restrict__ fam_start, const uint32_t *__restrict__ fam_size,
                const uint32_t *__restrict__ perm, uint32_t n_fam, uint32_t *__restrict__ read_fam)
{
    const int lane = threadIdx.x & 31;
    uint32_t c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= n_fam) return;
    uint32_t s = fam_start[c], n = fam_size[c];
    for (uint32_t j = lane; j < n; j += 32) read_fam[perm[s + j]] = c;
}

// ---- K6: complementary-key hash join (duplex_tag, consensus_helper.py:605-648) -------------
// duplex_tag swaps the barcode halves and R1 <-> R2.  When swapping twice gives the barcode back
// (every dotted barcode, every undotted one of even length) the relation is symmetric: a partner
// of an R1 family is an R2 family and finds it back.  Then only the R2 families go into the table,
// only the R1 families probe, and a hit fills both `partner` entries.  *asym != 0 (some barcode's
// double swap is another barcode: undotted, odd length) keeps the full build and probe.
__global__ void dup_symmetry_kernel(const uint32_t *__restrict__ dup_bc, uint32_t n_bc, uint32_t *__restrict__ asym)
{
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_bc) return;
    uint32_t d = dup_bc[b];
    if (d != NO_ID && (d >= n_bc || dup_bc[d] != b)) *asym = 1u;
}

// table entry: high word = join tag of the family's key hash, low word = family id + 1
__global__ void __launch_bounds__(256)
join_build_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ tmpl, uint32_t n_fam,
                  unsigned long long *__restrict__ table, uint64_t mask, const uint32_t *__restrict__ asym)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_fam) return;
    Rec r = load_rec(fields, tmpl[c]);
    if (!*asym && (r.b.w & 0x40u)) return;          // symmetric join: R1 families only probe
    uint64_t h = rec_hash(r);
    unsigned long long entry = ((h >> 32) << 32) | (unsigned long long)(c + 1);
    uint64_t slot = (h * 0xD6E8FEB86659FD93ull >> 20) & mask;
    while (atomicCAS(&table[slot], 0ull, entry) != 0ull) slot = (slot + 1) & mask;
}

__global__ void __launch_bounds__(256)
join_probe_kernel(const uint4 *__restrict__ fields, const uint32_t *__restrict__ tmpl, uint32_t n_fam,
                  const uint32_t *__restrict__ dup_bc, uint32_t n_bc,
                  const unsigned long long *__restrict__ table, uint64_t mask, int32_t *__restrict__ partner,
                  const uint32_t *__restrict__ asym)
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
        uint64_t h = rec_hash(q);
        uint32_t tag = (uint32_t)(h >> 32);
        uint64_t slot = (h * 0xD6E8FEB86659FD93ull >> 20) & mask;
        for (;;) {
            unsigned long long e = table[slot];
            if (e == 0ull) break;
            if ((uint32_t)(e >> 32) == tag) {
                uint32_t o = (uint32_t)e - 1u;
                if (same_tag(load_rec(fields, tmpl[o]), q)) { found = (int32_t)o; break; }
            }
            slot = (slot + 1) & mask;
        }
    }
    partner[c] = found;
    if (symmetric && found >= 0) partner[found] = (int32_t)c;
}

// ---- cc_join_records: the same join over records that stand for themselves ----------------------------
// (the stand-alone DCS_maker / singleton_correction stage commands: one record per tag, so there are no
// families to build first).  Every record goes into the table and every record probes; of several records
// with the complementary key the one with the smallest index answers.
__global__ void __launch_bounds__(256)
record_build_kernel(const uint4 *__restrict__ fields, uint32_t n, unsigned long long *__restrict__ table, uint64_t mask)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    uint64_t h = rec_hash(load_rec(fields, c));
    unsigned long long entry = ((h >> 32) << 32) | (unsigned long long)(c + 1);
    uint64_t slot = (h * 0xD6E8FEB86659FD93ull >> 20) & mask;
    while (atomicCAS(&table[slot], 0ull, entry) != 0ull) slot = (slot + 1) & mask;
}

__global__ void __launch_bounds__(256)
record_probe_kernel(const uint4 *__restrict__ fields, uint32_t n, const uint32_t *__restrict__ dup_bc, uint32_t n_bc,
                    const unsigned long long *__restrict__ table, uint64_t mask, int32_t *__restrict__ partner)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    Rec q = load_rec(fields, c);
    uint32_t found = 0xFFFFFFFFu;
    uint32_t d = q.a.x < n_bc ? __ldg(&dup_bc[q.a.x]) : NO_ID;
    if (d != NO_ID) {
        q.a.x = d;                 // barcode halves swapped
        q.b.w ^= 0x40u;            // R1 <-> R2
        uint64_t h = rec_hash(q);
        uint32_t tag = (uint32_t)(h >> 32);
        uint64_t slot = (h * 0xD6E8FEB86659FD93ull >> 20) & mask;
        for (;;) {
            unsigned long long e = table[slot];
            if (e == 0ull) break;
            if ((uint32_t)(e >> 32) == tag) {
                uint32_t o = (uint32_t)e - 1u;
                if (o < found && same_tag(load_rec(fields, o), q)) found = o;
            }
            slot = (slot + 1) & mask;
        }
    }
    partner[c] = found == 0xFFFFFFFFu ? -1 : (int32_t)found;
}

// ---- classification of families into output classes ----------------------------------------
// DCS_maker.py:233-253 (lead strand makes the DCS, the other is skipped, no partner ->
// SSCS singleton); singleton_correction.py:258-283; second DCS pass over the merge of
// sscs + sscs.correction + singleton.correction (ConsensusCruncher.py:247-270).
__device__ __forceinline__ bool leads_after_sc(uint32_t c, uint32_t p, bool c_big, bool p_big)
{
    if (c_big != p_big) return c_big;   // sscs records precede correction records in the merge
    return c < p;
}

__global__ void __launch_bounds__(256)
classify_kernel(const uint32_t *__restrict__ fam_size, const int32_t *__restrict__ partner, uint32_t n_fam,
                int scorrect, uint8_t *__restrict__ f_dcs, uint8_t *__restrict__ f_corr,
                uint8_t *__restrict__ f_dcssc, unsigned long long *__restrict__ stats, uint32_t *__restrict__ one_sided)
{
    __shared__ unsigned int s_st[16];
    if (threadIdx.x < 16) s_st[threadIdx.x] = 0;
    __syncthreads();
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n_fam) {
        bool big = fam_size[c] >= 2;
        int32_t p = partner[c];
        // a partner that has another partner (undotted barcode of odd length): which strand leads, and
        // whether the reference survives, depends on file order (DCS_maker.py:231-247) - outside the domain
        if (p >= 0 && partner[p] != (int32_t)c) atomicAdd(one_sided, 1u);
        bool p_big = p >= 0 && fam_size[p] >= 2;
        bool dcs = big && p_big && c < (uint32_t)p;
        bool corr = scorrect && !big && p >= 0;
        bool dcssc = scorrect && p >= 0 && leads_after_sc(c, (uint32_t)p, big, p_big);
        // a pair of two SSCS gives the same record before and after the rescue: its row is shared,
        // only pairs that involve a corrected singleton get a row of their own in the second pass
        f_dcs[c] = dcs; f_corr[c] = corr; f_dcssc[c] = dcssc && !dcs;
        atomicAdd(&s_st[big ? 0 : 1], 1u);
        if (dcs) atomicAdd(&s_st[2], 1u);
        if (big && !p_big) atomicAdd(&s_st[3], 1u);
        if (scorrect) {
            if (!big && p_big) atomicAdd(&s_st[4], 1u);
            if (!big && p >= 0 && !p_big) atomicAdd(&s_st[5], 1u);
            if (!big && p < 0) atomicAdd(&s_st[6], 1u);
            if (dcssc) atomicAdd(&s_st[7], 1u);
            if (big && p < 0) atomicAdd(&s_st[8], 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x < 16 && s_st[threadIdx.x]) atomicAdd(&stats[threadIdx.x], (unsigned long long)s_st[threadIdx.x]);
}

// classify_kernel and the three compactions behind it in one chained scan: every family decides its
// output classes, the running counts of the three classes travel from tile to tile (scan.cuh:
// lookback_pairs), and the lists (row -> family) and maps (family -> row) are written on the spot.
// dcssc_row holds the rank among the second-pass rows, or -2 - (row of the shared DCS) / -1: the rows of the
// second pass follow the n_dcs shared ones, and n_dcs is only known when the scan ends (dcssc_row_finish).
constexpr int CLS_THREADS = 256, CLS_ITEMS = 8, CLS_TILE = CLS_THREADS * CLS_ITEMS;

__global__ void __launch_bounds__(CLS_THREADS)
classify_scan_kernel(const uint32_t *__restrict__ fam_size, const int32_t *__restrict__ partner,
                     const uint32_t *__restrict__ n_fam_ptr, int scorrect,
                     unsigned long long *status, uint32_t *ticket,
                     uint32_t *__restrict__ dcs_list, uint32_t *__restrict__ corr_list, uint32_t *__restrict__ dcssc_list,
                     int32_t *__restrict__ dcs_row, int32_t *__restrict__ corr_row, int32_t *__restrict__ dcssc_row,
                     unsigned long long *__restrict__ stats, uint32_t *__restrict__ one_sided,
                     uint32_t *__restrict__ n_dcs, uint32_t *__restrict__ n_corr, uint32_t *__restrict__ n_extra)
{
    __shared__ unsigned int s_st[16];
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp[CLS_THREADS / 32][3];
    __shared__ uint32_t s_excl[4];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 16) s_st[tid] = 0;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    // the family count lives on the device; the grid covers the host's bound and the tiles beyond the
    // last family leave at once (they come after every tile that matters: nobody looks back at them)
    const uint32_t n_fam = *n_fam_ptr;
    const uint32_t last_tile = n_fam ? (n_fam - 1) / CLS_TILE : 0;
    if (n_fam == 0 || tile > last_tile) return;
    const uint32_t base = tile * CLS_TILE + tid * CLS_ITEMS;
    uint32_t f_dcs = 0, f_corr = 0, f_x = 0;          // bit k: family base + k is in the class
    uint32_t mine[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t size[CLS_ITEMS];
    int32_t par[CLS_ITEMS];
    if (base + CLS_ITEMS <= n_fam) {
        const uint4 a = *reinterpret_cast<const uint4 *>(fam_size + base), b = *reinterpret_cast<const uint4 *>(fam_size + base + 4);
        const int4 c = *reinterpret_cast<const int4 *>(partner + base), d = *reinterpret_cast<const int4 *>(partner + base + 4);
        size[0] = a.x; size[1] = a.y; size[2] = a.z; size[3] = a.w; size[4] = b.x; size[5] = b.y; size[6] = b.z; size[7] = b.w;
        par[0] = c.x; par[1] = c.y; par[2] = c.z; par[3] = c.w; par[4] = d.x; par[5] = d.y; par[6] = d.z; par[7] = d.w;
    } else {
#pragma unroll
        for (int k = 0; k < CLS_ITEMS; k++) {
            size[k] = base + k < n_fam ? fam_size[base + k] : 0u;
            par[k] = base + k < n_fam ? partner[base + k] : -1;
        }
    }
#pragma unroll
    for (int k = 0; k < CLS_ITEMS; k++) {
        const uint32_t c = base + k;
        if (c >= n_fam) continue;
        const bool big = size[k] >= 2;
        const int32_t p = par[k];
        // a partner that has another partner (undotted barcode of odd length): which strand leads, and
        // whether the reference survives, depends on file order (DCS_maker.py:231-247) - outside the domain
        CC_CHECK(p < (int32_t)n_fam);
        if (p >= 0 && __ldg(&partner[p]) != (int32_t)c) atomicAdd(one_sided, 1u);
        const bool p_big = p >= 0 && __ldg(&fam_size[p]) >= 2;
        const bool dcs = big && p_big && c < (uint32_t)p;
        const bool corr = scorrect && !big && p >= 0;
        const bool dcssc = scorrect && p >= 0 && leads_after_sc(c, (uint32_t)p, big, p_big);
        // a pair of two SSCS gives the same record before and after the rescue: its row is shared,
        // only pairs that involve a corrected singleton get a row of their own in the second pass
        f_dcs |= (uint32_t)dcs << k; f_corr |= (uint32_t)corr << k; f_x |= (uint32_t)(dcssc && !dcs) << k;
        mine[big ? 0 : 1]++;
        mine[2] += dcs;
        mine[3] += big && !p_big;
        if (scorrect) {
            mine[4] += !big && p_big;
            mine[5] += !big && p >= 0 && !p_big;
            mine[6] += !big && p < 0;
            mine[7] += dcssc;
            mine[8] += big && p < 0;
        }
    }
#pragma unroll
    for (int k = 0; k < 9; k++) {
        const uint32_t v = __reduce_add_sync(0xffffffffu, mine[k]);
        if (lane == 0 && v) atomicAdd(&s_st[k], v);
    }
    // block-wide exclusive prefixes of the three class counts
    const uint32_t own[3] = {(uint32_t)__popc(f_dcs), (uint32_t)__popc(f_corr), (uint32_t)__popc(f_x)};
    uint32_t inc[3] = {own[0], own[1], own[2]};
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
        for (int v = 0; v < 3; v++) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc[v], d);
            if (lane >= d) inc[v] += t;
        }
    }
    if (lane == 31) { s_warp[warp][0] = inc[0]; s_warp[warp][1] = inc[1]; s_warp[warp][2] = inc[2]; }
    __syncthreads();
    uint32_t off[3] = {0, 0, 0}, tot[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < CLS_THREADS / 32; k++) {
#pragma unroll
        for (int v = 0; v < 3; v++) {
            if (k < warp) off[v] += s_warp[k][v];
            tot[v] += s_warp[k][v];
        }
    }
    if (warp == 0) {
        uint32_t excl[4];
        lookback_pairs<2>(status, tile, tot, excl, lane);
        if (lane == 0) {
            s_excl[0] = excl[0]; s_excl[1] = excl[1]; s_excl[2] = excl[2];
            if (tile == last_tile) { *n_dcs = excl[0] + tot[0]; *n_corr = excl[1] + tot[1]; *n_extra = excl[2] + tot[2]; }
        }
    }
    __syncthreads();
    uint32_t r_dcs = s_excl[0] + off[0] + inc[0] - own[0];
    uint32_t r_corr = s_excl[1] + off[1] + inc[1] - own[1];
    uint32_t r_x = s_excl[2] + off[2] + inc[2] - own[2];
    int32_t o_dcs[CLS_ITEMS], o_corr[CLS_ITEMS], o_x[CLS_ITEMS];
#pragma unroll
    for (int k = 0; k < CLS_ITEMS; k++) {
        const uint32_t c = base + k;
        o_dcs[k] = -1; o_corr[k] = -1; o_x[k] = -1;
        CC_CHECK(r_dcs <= n_fam && r_corr <= n_fam && r_x <= n_fam);
        if ((f_dcs >> k) & 1u) { dcs_list[r_dcs] = c; o_dcs[k] = (int32_t)r_dcs++; }
        if ((f_corr >> k) & 1u) { corr_list[r_corr] = c; o_corr[k] = (int32_t)r_corr++; }
        if ((f_x >> k) & 1u) { dcssc_list[r_x] = c; o_x[k] = (int32_t)r_x++; }
        else if (scorrect && o_dcs[k] >= 0) o_x[k] = -2 - o_dcs[k];        // <= -2: shares DCS row -2 - value
    }
    if (base + CLS_ITEMS <= n_fam) {
        *reinterpret_cast<int4 *>(dcs_row + base) = make_int4(o_dcs[0], o_dcs[1], o_dcs[2], o_dcs[3]);
        *reinterpret_cast<int4 *>(dcs_row + base + 4) = make_int4(o_dcs[4], o_dcs[5], o_dcs[6], o_dcs[7]);
        *reinterpret_cast<int4 *>(corr_row + base) = make_int4(o_corr[0], o_corr[1], o_corr[2], o_corr[3]);
        *reinterpret_cast<int4 *>(corr_row + base + 4) = make_int4(o_corr[4], o_corr[5], o_corr[6], o_corr[7]);
        *reinterpret_cast<int4 *>(dcssc_row + base) = make_int4(o_x[0], o_x[1], o_x[2], o_x[3]);
        *reinterpret_cast<int4 *>(dcssc_row + base + 4) = make_int4(o_x[4], o_x[5], o_x[6], o_x[7]);
    } else {
#pragma unroll
        for (int k = 0; k < CLS_ITEMS; k++)
            if (base + k < n_fam) { dcs_row[base + k] = o_dcs[k]; corr_row[base + k] = o_corr[k]; dcssc_row[base + k] = o_x[k]; }
    }
    __syncthreads();
    if (tid < 16 && s_st[tid]) atomicAdd(&stats[tid], (unsigned long long)s_st[tid]);
}

// dcssc_row as the C ABI describes it: rows of the second pass follow the n_dcs shared rows
__global__ void __launch_bounds__(256)
dcssc_row_finish_kernel(int32_t *__restrict__ dcssc_row, uint32_t n_fam, uint32_t n_dcs)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_fam) return;
    const int32_t v = dcssc_row[c];
    if (v >= 0) dcssc_row[c] = (int32_t)n_dcs + v;
    else if (v <= -2) dcssc_row[c] = -2 - v;
}

// ---- K4 + K5: SSCS vote and family mode fields ------------------------------------------------
// consensus_maker (SSCS_maker.py:103-155): per position, among bases with Q >= 30, the
// first maximum of A,C,G,T wins; quality = min(60, sum of the winner's qualities); the
// winner is emitted iff count / passing >= cutoff (fp64), else N with the same quality.
// Because a passing quality is >= 30, the capped sum is 60 as soon as two reads agree and
// the single read's own quality otherwise, so only counts and an OR of qualities are kept.
//
// One warp per family, lane <-> 4 consecutive positions of a 128-position chunk.
struct VoteAcc {
    uint32_t cnt[4];   // A,C,G,T counts, one byte lane per position (flushed before 256)
    uint32_t qor[4];   // OR of passing qualities per base, one byte lane per position
};

__device__ __forceinline__ uint32_t expand_nibbles(uint32_t s16)
{
    // s16 = two sequence bytes (4 bases, even position in the low nibble) -> one base per byte, position order
    uint32_t lo = s16 & 0x0F0Fu;             // positions 0 and 2
    uint32_t hi = (s16 >> 4) & 0x0F0Fu;      // positions 1 and 3
    return __byte_perm(lo, hi, 0x5140);
}

__device__ __forceinline__ void vote_row(VoteAcc &a, uint32_t q, uint32_t s16, uint32_t &err)
{
    // byte-wise q >= 30  (bit 7 after adding 128-30; values >= 128 pass through the OR)
    uint32_t pass7 = ((((q & 0x7F7F7F7Fu) + 0x62626262u) | q) & 0x80808080u);
    uint32_t pass1 = pass7 >> 7;
    uint32_t b = expand_nibbles(s16) & (pass1 * 15u);
    uint32_t e0 = b & 0x01010101u, e1 = (b >> 1) & 0x01010101u;
    uint32_t e2 = (b >> 2) & 0x01010101u, e3 = (b >> 3) & 0x01010101u;
    err |= (e0 + e1 + e2 + e3) ^ pass1;      // a passing base must be exactly one of A,C,G,T
    a.cnt[0] += e0; a.cnt[1] += e1; a.cnt[2] += e2; a.cnt[3] += e3;
    a.qor[0] |= q & (e0 * 0xFFu); a.qor[1] |= q & (e1 * 0xFFu);
    a.qor[2] |= q & (e2 * 0xFFu); a.qor[3] |= q & (e3 * 0xFFu);
}

// mode of one field over a family, ties -> first appearance (what randint -> 0 selects in
// read_mode, consensus_helper.py:491-495); flags prefer 99 > 83 > 147 > 163 on a tie (:521-531)
__device__ __forceinline__ int flag_rank(uint32_t v) { return v == 99 ? 0 : v == 83 ? 1 : v == 147 ? 2 : v == 163 ? 3 : 4; }

// get(j) = value of member j, idx(j) = its read index.  "First appearing" is decided by read index, not
// by position in the member list, so the list may be in any order.
template <typename Get, typename Idx>
__device__ uint32_t warp_mode(Get get, Idx idx, uint32_t n, bool is_flag, int lane)
{
    uint32_t best = 0, best_cnt = 0, best_first = 0xFFFFFFFFu;
    int best_rank = 9;
    uint32_t counted = 0;
    for (uint32_t i = 0; counted < n && i < n; i++) {
        uint32_t vi = get(i);
        uint32_t cnt = 0, first = 0xFFFFFFFFu;
        bool seen_before = false;
        for (uint32_t j0 = 0; j0 < n; j0 += 32) {
            uint32_t j = j0 + lane;
            bool m = j < n && get(j) == vi;
            uint32_t b = __ballot_sync(0xffffffffu, m);
            if (cnt == 0 && b && j0 + (uint32_t)__ffs(b) - 1u < i) { seen_before = true; break; }
            cnt += __popc(b);
            if (m) first = min(first, idx(j));
        }
        if (seen_before) continue;
        first = __reduce_min_sync(0xffffffffu, first);
        counted += cnt;
        int rank = is_flag ? flag_rank(vi) : 4;
        if (cnt > best_cnt || (cnt == best_cnt && (rank < best_rank || (rank == best_rank && first < best_first)))) {
            best = vi; best_cnt = cnt; best_rank = rank; best_first = first;
        }
    }
    return best;
}

// WARPS warps vote one family of any size: members perm[start .. start+n), consensus row -> out,
// mode fields -> fam_*[c].  Used for families too large for vote_small_kernel (vote.cuh; WARPS = 8:
// the warps of a CTA share the members - counts add, the per-base quality ORs combine - and warp 0
// finishes) and by the single-family entry points (WARPS = 1).  `scratch`: WARPS * 32 * 21 words of
// shared memory when WARPS > 1; every thread of the CTA must make the call.
constexpr int VOTE_PART_WORDS = 21;      // 16 wide counts + 4 quality ORs + error bits per lane

template <int WARPS>
__device__ __forceinline__ void
vote_family_warp(const uint4 *__restrict__ fields, const uint8_t *__restrict__ payload, uint32_t stride,
                 uint32_t lpad, const uint32_t *__restrict__ perm, uint32_t start, uint32_t n, double cutoff,
                 uint8_t *__restrict__ out, uint32_t c, uint8_t *__restrict__ fam_mapq,
                 uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
                 uint32_t *__restrict__ err_flag, uint32_t *scratch = nullptr, bool mixed = true)
{
    const int lane = threadIdx.x & 31;
    const int warp = WARPS > 1 ? (int)(threadIdx.x >> 5) : 0;
    const uint32_t half = lpad >> 1;
    uint32_t err = 0;

    for (uint32_t chunk = 0; chunk * 128 < lpad; chunk++) {
        const uint32_t p0 = chunk * 128 + lane * 4;           // first of this lane's 4 positions
        const bool active = p0 < lpad;
        uint32_t wide[4][4];                                   // [base][position] for families > 255 reads
#pragma unroll
        for (int bi = 0; bi < 4; bi++)
#pragma unroll
            for (int k = 0; k < 4; k++) wide[bi][k] = 0;
        VoteAcc acc;
#pragma unroll
        for (int bi = 0; bi < 4; bi++) { acc.cnt[bi] = 0; acc.qor[bi] = 0; }
        uint32_t since_flush = 0;
        for (uint32_t j0 = 32 * warp; j0 < n; j0 += 32 * WARPS) {
            // lanes fetch 32 member indices at once, then the warp walks the rows together
            uint32_t mine = (j0 + lane < n) ? perm[start + j0 + lane] : 0;
            uint32_t lim = min(32u, n - j0);
            // eight rows at a time: all their loads are issued before the first tally (a family of 10^4 reads is one
            // CTA's work, so the latency of a row must not add up row by row); rows past the end tally nothing (q = 0)
            for (uint32_t jj = 0; jj < lim; jj += 8) {
                uint32_t s16[8], q[8];
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const uint32_t r = __shfl_sync(0xffffffffu, mine, (jj + u) & 31);
                    const bool on = active && jj + u < lim;
                    const uint8_t *src = payload + (uint64_t)r * stride;
                    s16[u] = on ? *reinterpret_cast<const uint16_t *>(src + (p0 >> 1)) : 0u;
                    q[u] = on ? *reinterpret_cast<const uint32_t *>(src + half + p0) : 0u;
                }
#pragma unroll
                for (int u = 0; u < 8; u++) vote_row(acc, q[u], s16[u], err);
                since_flush += 8;
                if (since_flush >= 248) {                      // byte counters hold 255
#pragma unroll
                    for (int bi = 0; bi < 4; bi++) {
#pragma unroll
                        for (int k = 0; k < 4; k++) wide[bi][k] += (acc.cnt[bi] >> (8 * k)) & 0xFF;
                        acc.cnt[bi] = 0;
                    }
                    since_flush = 0;
                }
            }
        }
#pragma unroll
        for (int bi = 0; bi < 4; bi++) {
#pragma unroll
            for (int k = 0; k < 4; k++) wide[bi][k] += (acc.cnt[bi] >> (8 * k)) & 0xFF;
        }
        if (WARPS > 1) {
            // the warps' partial tallies meet in shared memory; warp 0 adds them up and finishes
            uint32_t *mine = scratch + ((size_t)warp * 32 + lane) * VOTE_PART_WORDS;
            __syncthreads();                                   // scratch is free (previous chunk / family done)
            if (warp != 0) {
#pragma unroll
                for (int bi = 0; bi < 4; bi++) {
#pragma unroll
                    for (int k = 0; k < 4; k++) mine[4 * bi + k] = wide[bi][k];
                    mine[16 + bi] = acc.qor[bi];
                }
                mine[20] = err;
            }
            __syncthreads();
            if (warp == 0) {
                for (int w2 = 1; w2 < WARPS; w2++) {
                    const uint32_t *o = scratch + ((size_t)w2 * 32 + lane) * VOTE_PART_WORDS;
#pragma unroll
                    for (int bi = 0; bi < 4; bi++) {
#pragma unroll
                        for (int k = 0; k < 4; k++) wide[bi][k] += o[4 * bi + k];
                        acc.qor[bi] |= o[16 + bi];
                    }
                    err |= o[20];
                }
            }
        }
        if (active && warp == 0) {
            uint32_t seq_out = 0, q_out = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                uint32_t cA = wide[0][k];
                uint32_t cC = wide[1][k];
                uint32_t cG = wide[2][k];
                uint32_t cT = wide[3][k];
                uint32_t best = cA; int w = 0;                  // first maximum in A,C,G,T order
                if (cC > best) { best = cC; w = 1; }
                if (cG > best) { best = cG; w = 2; }
                if (cT > best) { best = cT; w = 3; }
                uint32_t pass = cA + cC + cG + cT;
                uint32_t qw = (acc.qor[w] >> (8 * k)) & 0xFF;
                uint32_t mq = best == 0 ? 0u : (best == 1 ? min(qw, (uint32_t)Q_CAP) : (uint32_t)Q_CAP);
                uint32_t code = 15u;
                if (pass != 0 && ((double)best / (double)pass) >= cutoff) code = 1u << w;
                // positions 0,1 -> first sequence byte (low, high nibble); 2,3 -> second byte
                seq_out |= code << ((k >> 1) * 8 + ((k & 1) ? 4 : 0));
                q_out |= mq << (8 * k);
            }
            *reinterpret_cast<uint16_t *>(out + (p0 >> 1)) = (uint16_t)seq_out;
            *reinterpret_cast<uint32_t *>(out + half + p0) = q_out;
        }
    }
    if (warp != 0) return;                                     // (after the last barrier)
    for (uint32_t x = half + lpad + lane * 4; x < stride; x += 128) *reinterpret_cast<uint32_t *>(out + x) = 0;
    if (err) atomicOr(err_flag, 1u);
    if (!mixed) return;                                        // the grouping found one value per field: the template's, already written

    // ---- mode fields of the family (create_aligned_segment, consensus_helper.py:569-583)
    auto meta_of = [&](uint32_t j) { return __ldg(&fields[2 * (uint64_t)perm[start + j] + 1]); };
    uint4 m0 = meta_of(0);
    bool uniform = true, has_rg = true;
    for (uint32_t j0 = 0; j0 < n; j0 += 32) {
        uint32_t j = j0 + lane;
        uint4 m = j < n ? meta_of(j) : m0;
        uniform = uniform && __all_sync(0xffffffffu, m.z == m0.z && m.w == m0.w);
        has_rg = has_rg && __all_sync(0xffffffffu, (m.w >> 24) != 0xFFu);
    }
    uint32_t tlen = m0.z, flag = m0.w & 0xFFFF, mapq = (m0.w >> 16) & 0xFF, rg = m0.w >> 24;
    if (!uniform) {
        auto idx_of = [&](uint32_t j) { return perm[start + j]; };
        tlen = warp_mode([&](uint32_t j) { return meta_of(j).z; }, idx_of, n, false, lane);
        flag = warp_mode([&](uint32_t j) { return meta_of(j).w & 0xFFFFu; }, idx_of, n, true, lane);
        mapq = warp_mode([&](uint32_t j) { return (meta_of(j).w >> 16) & 0xFFu; }, idx_of, n, false, lane);
        rg = has_rg ? warp_mode([&](uint32_t j) { return meta_of(j).w >> 24; }, idx_of, n, false, lane) : 0xFFu;
    }
    if (lane == 0) {
        fam_tlen[c] = (int32_t)tlen; fam_flag[c] = (uint16_t)flag;
        fam_mapq[c] = (uint8_t)mapq; fam_rg[c] = (uint8_t)rg;
    }
}

__global__ void __launch_bounds__(256)
vote_kernel(const uint4 *__restrict__ fields, const uint8_t *__restrict__ payload, uint32_t stride,
            uint32_t lpad, const uint32_t *__restrict__ perm, const uint32_t *__restrict__ fam_start,
            const uint32_t *__restrict__ fam_size, const uint32_t *__restrict__ sscs_list, uint32_t n_rows,
            double cutoff, uint8_t *__restrict__ out_rows, uint8_t *__restrict__ fam_mapq,
            uint16_t *__restrict__ fam_flag, int32_t *__restrict__ fam_tlen, uint8_t *__restrict__ fam_rg,
            uint32_t *__restrict__ err_flag)
{
    const uint32_t row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= n_rows) return;
    const uint32_t c = sscs_list[row];
    vote_family_warp<1>(fields, payload, stride, lpad, perm, fam_start[c], fam_size[c], cutoff,
                     out_rows + (uint64_t)row * stride, c, fam_mapq, fam_flag, fam_tlen, fam_rg, err_flag);
}

// mode of two records (create_aligned_segment on [lead, partner]): equal -> value, else the
// lead's value (first appearance); flags prefer 99 > 83 > 147 > 163
__device__ __forceinline__ uint32_t pair_mode(uint32_t a, uint32_t b, bool is_flag)
{
    if (a == b) return a;
    if (is_flag && flag_rank(b) < flag_rank(a)) return b;
    return a;
}

struct RowSources {
    const uint8_t *payload;       // original reads
    const uint8_t *sscs_rows;
    const uint8_t *corr_rows;
    const uint32_t *tmpl;
    const uint32_t *fam_size;
    const int32_t *sscs_row;
    const int32_t *corr_row;
    uint32_t stride;
};

// ---- compact input rows (cc_load_reads_q4): lpad/2 base bytes + lpad/2 bytes of 4-bit quality
// codes -> the standard row (lpad/2 base bytes + lpad quality bytes, zero tail).  One thread
// expands 16 positions; a CTA first builds byte -> two-qualities from the 16-entry table.
struct QualTable { uint8_t q[16]; };

__global__ void __launch_bounds__(256)
expand_q4_kernel(const uint8_t *__restrict__ in, uint32_t stride_in, uint8_t *__restrict__ out, uint32_t stride_out,
                 uint32_t lpad, uint64_t n, QualTable table)
{
    __shared__ uint16_t pair_of[256];
    pair_of[threadIdx.x] = (uint16_t)(table.q[threadIdx.x & 15] | (table.q[threadIdx.x >> 4] << 8));   // low nibble = even position
    __syncthreads();
    const uint32_t lanes = lpad / 16, half = lpad / 2;
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t row = t / lanes;
    const uint32_t lane = (uint32_t)(t % lanes);
    if (row >= n) return;
    const uint8_t *src = in + row * stride_in;
    uint8_t *dst = out + row * stride_out;
    const uint2 bases = *reinterpret_cast<const uint2 *>(src + lane * 8);
    const uint2 codes = *reinterpret_cast<const uint2 *>(src + half + lane * 8);
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t c = (k < 2 ? codes.x : codes.y) >> (16 * (k & 1));
        w[k] = pair_of[c & 0xFF] | ((uint32_t)pair_of[(c >> 8) & 0xFF] << 16);
    }
    *reinterpret_cast<uint2 *>(dst + lane * 8) = bases;
    *reinterpret_cast<uint4 *>(dst + half + lane * 16) = make_uint4(w[0], w[1], w[2], w[3]);
    for (uint32_t x = half + lpad + lane * 4; x < stride_out; x += lanes * 4) *reinterpret_cast<uint32_t *>(dst + x) = 0;
}

// ---- compact upload (cc_load_reads_compact) ---------------------------------------------------------
// Pair records: the two mates of a pair share barcode and cigar pair and mirror each other's coordinates,
// so 40 bytes describe both reads' 32-byte field records.
struct PairRec {
    uint32_t bc, cig;
    int32_t tid, pos, mtid, mpos;      // of the first-seen mate
    int32_t tlen1, tlen2;
    uint32_t flags;                    // flag1 | flag2 << 16
    uint32_t meta;                     // mapq1 | mapq2 << 8 | rg1 << 16 | rg2 << 24
};
static_assert(sizeof(PairRec) == 40, "cc_pair is 40 bytes");

__global__ void __launch_bounds__(256)
expand_pairs_kernel(const uint2 *__restrict__ pairs, uint64_t n_pairs, uint4 *__restrict__ fields)
{
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pairs) return;
    const uint2 a = pairs[5 * k], b = pairs[5 * k + 1], c = pairs[5 * k + 2], d = pairs[5 * k + 3], e = pairs[5 * k + 4];
    // a = (bc, cig)  b = (tid, pos)  c = (mtid, mpos)  d = (tlen1, tlen2)  e = (flags, meta)
    const uint32_t m1 = (e.x & 0xFFFFu) | ((e.y & 0xFFu) << 16) | (((e.y >> 16) & 0xFFu) << 24);
    const uint32_t m2 = (e.x >> 16) | (((e.y >> 8) & 0xFFu) << 16) | ((e.y >> 24) << 24);
    fields[4 * k] = make_uint4(a.x, a.y, b.x, b.y);
    fields[4 * k + 1] = make_uint4(c.x, c.y, d.x, m1);
    fields[4 * k + 2] = make_uint4(a.x, a.y, c.x, c.y);
    fields[4 * k + 3] = make_uint4(b.x, b.y, d.y, m2);
}

// ---- result rows in bit planes (cc_pack_rows) ---------------------------------------------------------
// The inverse of expand_planes_kernel for the rows the path produces.  A position's key is
// quality << 8 | literal, literal = 0xFF for a plain A/C/G/T and the 4-bit base code otherwise (N, the padding
// after the read's end).  row_alphabet_kernel marks the keys that occur (65 536 bits); with at most 16 of them
// pack_rows_kernel writes 2-bit bases + QBITS-bit codes in the layout of cc_load_reads_compact's planes.
constexpr int ALPHA_WORDS = 65536 / 32;

__device__ __forceinline__ uint32_t row_key(uint32_t nib, uint32_t q)
{
    const bool plain = nib == 1u || nib == 2u || nib == 4u || nib == 8u;
    return (q << 8) | (plain ? 0xFFu : nib);
}

__global__ void __launch_bounds__(256)
row_alphabet_kernel(const uint8_t *__restrict__ rows, uint32_t stride, uint32_t lpad, uint64_t n_rows, uint32_t *__restrict__ seen)
{
    __shared__ uint32_t s_seen[ALPHA_WORDS];
    for (int k = threadIdx.x; k < ALPHA_WORDS; k += blockDim.x) s_seen[k] = 0;
    __syncthreads();
    const uint32_t lanes = lpad / 16, half = lpad / 2;
    const uint64_t total = n_rows * lanes;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t row = t / lanes;
        const uint32_t lane = (uint32_t)(t % lanes);
        const uint8_t *src = rows + row * stride;
        const uint2 b = *reinterpret_cast<const uint2 *>(src + lane * 8);
        const uint4 q = *reinterpret_cast<const uint4 *>(src + half + lane * 16);
        const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
        uint32_t last = 0xFFFFFFFFu;
#pragma unroll
        for (int p = 0; p < 16; p++) {
            const uint32_t nib = ((p < 8 ? b.x : b.y) >> (4 * (p & 7))) & 15u;
            const uint32_t key = row_key(nib, (qw[p >> 2] >> (8 * (p & 3))) & 0xFFu);
            if (key == last) continue;                                  // runs of one key are the rule
            last = key;
            if (!((s_seen[key >> 5] >> (key & 31)) & 1u)) atomicOr(&s_seen[key >> 5], 1u << (key & 31));
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < ALPHA_WORDS; k += blockDim.x)
        if (s_seen[k]) atomicOr(&seen[k], s_seen[k]);
}

struct RowKeys { uint16_t key[16]; uint32_t count; };

template <int QBITS>
__global__ void __launch_bounds__(256)
pack_rows_kernel(const uint8_t *__restrict__ rows, uint32_t stride, uint32_t lpad, uint64_t n_rows,
                 uint8_t *__restrict__ out, uint32_t stride_c, RowKeys keys, uint32_t *__restrict__ misfit)
{
    __shared__ uint8_t s_code[17][256];                 // [literal, 16 = plain A/C/G/T][quality] -> code, 0xFF = none
    for (int k = threadIdx.x; k < 17 * 256 / 4; k += blockDim.x) reinterpret_cast<uint32_t *>(&s_code[0][0])[k] = 0xFFFFFFFFu;
    __syncthreads();
    if (threadIdx.x < keys.count) {
        const uint32_t key = keys.key[threadIdx.x], lit = key & 0xFFu;
        s_code[lit == 0xFFu ? 16 : (lit & 15u)][key >> 8] = (uint8_t)threadIdx.x;
    }
    __syncthreads();
    const uint32_t lanes = lpad / 16, half = lpad / 2;
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t row = t / lanes;
    const uint32_t lane = (uint32_t)(t % lanes);
    if (row >= n_rows) return;
    const uint8_t *src = rows + row * stride;
    uint8_t *dst = out + row * stride_c;
    const uint2 b = *reinterpret_cast<const uint2 *>(src + lane * 8);
    const uint4 q = *reinterpret_cast<const uint4 *>(src + half + lane * 16);
    const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
    uint32_t b2 = 0;
    unsigned long long codes = 0;
    bool bad = false;
#pragma unroll
    for (int p = 0; p < 16; p++) {
        const uint32_t nib = ((p < 8 ? b.x : b.y) >> (4 * (p & 7))) & 15u;
        const uint32_t qv = (qw[p >> 2] >> (8 * (p & 3))) & 0xFFu;
        const bool plain = nib == 1u || nib == 2u || nib == 4u || nib == 8u;
        const uint32_t code = s_code[plain ? 16 : nib][qv];
        bad |= code == 0xFFu;
        if (plain) b2 |= (uint32_t)(__ffs(nib) - 1) << (2 * p);
        codes |= (unsigned long long)(code & ((1u << QBITS) - 1u)) << (p * QBITS);
    }
    if (bad) atomicAdd(misfit, 1u);
    *reinterpret_cast<uint32_t *>(dst + lane * 4) = b2;
    uint16_t *qdst = reinterpret_cast<uint16_t *>(dst + lpad / 4 + lane * 2 * QBITS);
#pragma unroll
    for (int k = 0; k < QBITS; k++) qdst[k] = (uint16_t)(codes >> (16 * k));
    for (uint32_t x = lpad / 4 + lpad * QBITS / 8 + lane * 4; x < stride_c; x += lanes * 4) *reinterpret_cast<uint32_t *>(dst + x) = 0;
}

// Short pair records (cc_load_reads_classes): what many pairs have in common - cigar pair, contigs, flags, MAPQs,
// read groups - sits once in a class table; a pair keeps its barcode, its two positions, the first mate's TLEN
// (the second mate's is its negative) and the number of its class.  16 bytes describe both 32-byte field records.
struct PairClass {
    uint32_t cig;
    int32_t tid, mtid;
    uint32_t flags;                    // flag1 | flag2 << 16
    uint32_t meta;                     // mapq1 | mapq2 << 8 | rg1 << 16 | rg2 << 24
    uint32_t reserved[3];
};
static_assert(sizeof(PairClass) == 32, "cc_pair_class is 32 bytes");

__global__ void __launch_bounds__(256)
expand_pairs16_kernel(const uint4 *__restrict__ pairs, const uint4 *__restrict__ classes, uint32_t n_classes,
                      uint64_t n_pairs, uint4 *__restrict__ fields, uint32_t *__restrict__ bad_class)
{
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pairs) return;
    const uint4 p = pairs[k];                       // bc, pos, mpos, tlen1 (i16) | cls << 16
    uint32_t cls = p.w >> 16;
    if (cls >= n_classes) { atomicAdd(bad_class, 1u); cls = 0; }
    const uint4 c0 = __ldg(&classes[2 * cls]);      // cig, tid, mtid, flags
    const uint32_t meta = __ldg(&classes[2 * cls + 1]).x;
    const int32_t tlen1 = (int32_t)(int16_t)(p.w & 0xFFFFu);
    const uint32_t m1 = (c0.w & 0xFFFFu) | ((meta & 0xFFu) << 16) | (((meta >> 16) & 0xFFu) << 24);
    const uint32_t m2 = (c0.w >> 16) | (((meta >> 8) & 0xFFu) << 16) | ((meta >> 24) << 24);
    fields[4 * k] = make_uint4(p.x, c0.x, c0.y, p.y);
    fields[4 * k + 1] = make_uint4(c0.z, p.z, (uint32_t)tlen1, m1);
    fields[4 * k + 2] = make_uint4(p.x, c0.x, c0.z, p.z);
    fields[4 * k + 3] = make_uint4(c0.y, p.y, (uint32_t)(-tlen1), m2);
}

// Rows in bit planes: lpad/4 bytes of 2-bit bases (A,C,G,T = 0..3, position p at bits 2p..2p+1) followed by
// lpad * QBITS / 8 bytes of QBITS-bit codes (position p at bit p * QBITS, little endian).  A code names
// a quality and, where base_of[code] != 0xFF, the literal 4-bit base code of the position (N, a pad, an
// IUPAC letter) instead of the 2-bit base.  One thread expands 16 positions into the standard row.
struct CodeTable { uint8_t qual[16]; uint8_t base[16]; };

template <int QBITS>
__global__ void __launch_bounds__(256)
expand_planes_kernel(const uint8_t *__restrict__ in, uint32_t stride_in, uint8_t *__restrict__ out, uint32_t stride_out,
                     uint32_t lpad, uint64_t n, CodeTable table)
{
    __shared__ uint8_t s_qual[16], s_base[16];
    if (threadIdx.x < 16) { s_qual[threadIdx.x] = table.qual[threadIdx.x]; s_base[threadIdx.x] = table.base[threadIdx.x]; }
    __syncthreads();
    const uint32_t lanes = lpad / 16, half = lpad / 2;
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t row = t / lanes;
    const uint32_t lane = (uint32_t)(t % lanes);
    if (row >= n) return;
    const uint8_t *src = in + row * stride_in;
    uint8_t *dst = out + row * stride_out;
    const uint32_t b2 = *reinterpret_cast<const uint32_t *>(src + lane * 4);              // 16 bases
    const uint16_t *qsrc = reinterpret_cast<const uint16_t *>(src + lpad / 4 + lane * 2 * QBITS);
    unsigned long long codes = 0;                                                          // 16 codes, QBITS bits each
#pragma unroll
    for (int k = 0; k < QBITS; k++) codes |= (unsigned long long)qsrc[k] << (16 * k);
    uint32_t seq[2] = {0, 0}, q[4] = {0, 0, 0, 0};
#pragma unroll
    for (int p = 0; p < 16; p++) {
        const uint32_t code = (uint32_t)(codes >> (p * QBITS)) & ((1u << QBITS) - 1u);
        const uint32_t lit = s_base[code];
        const uint32_t base = lit != 0xFFu ? lit : (1u << ((b2 >> (2 * p)) & 3u));
        seq[p >> 3] |= base << (4 * (p & 7));                       // even position in the low nibble
        q[p >> 2] |= (uint32_t)s_qual[code] << (8 * (p & 3));
    }
    *reinterpret_cast<uint2 *>(dst + lane * 8) = make_uint2(seq[0], seq[1]);
    *reinterpret_cast<uint4 *>(dst + half + lane * 16) = make_uint4(q[0], q[1], q[2], q[3]);
    for (uint32_t x = half + lpad + lane * 4; x < stride_out; x += lanes * 4) *reinterpret_cast<uint32_t *>(dst + x) = 0;
}

// ---- coordinate sort of output records (sort_index, ConsensusCruncher.py:10-29) ----------------
// low key: position then strand, as `samtools sort` orders records of one contig
__global__ void coord_pos_key_kernel(const int32_t *__restrict__ pos, const uint16_t *__restrict__ flag, uint64_t n,
                                     uint32_t *__restrict__ key)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) key[i] = ((uint32_t)(pos[i] + 1) << 1) | ((flag[i] >> 4) & 1u);     // pos >= -1 (checked on the host)
}

// high key, read through the order the first sort left: contig index, records without one last
__global__ void coord_tid_key_kernel(const int32_t *__restrict__ tid, const uint32_t *__restrict__ order, uint64_t n,
                                     uint32_t no_contig, uint32_t *__restrict__ key)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        int32_t t = tid[order[i]];
        key[i] = t < 0 ? no_contig : (uint32_t)t;
    }
}

__global__ void iota_kernel(uint32_t *p, uint32_t n)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}

}  // namespace ccgpu
