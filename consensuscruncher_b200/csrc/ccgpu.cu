// libccgpu.so - host side of the C ABI declared in include/ccgpu.h.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include "../../include/ccgpu.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <nccl.h>          // types and constants only: the library is loaded at run time (see NcclApi)

#include "kernels.cuh"
#include "radix_onesweep.cuh"
#include "scan.cuh"
#include "vote.cuh"
#include "group_hash.cuh"
#include "group_tile.cuh"
#include "group_pair.cuh"

using namespace ccgpu;

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

struct PhaseTimer {
    std::vector<cudaEvent_t> ev;   // start/stop pairs
    size_t used = 0;
    double ms = 0.0;
    uint64_t runs = 0;
};

// NCCL is bound with dlopen when the first communicator call is made, so libccgpu.so has no link-time
// dependency on it: single-GPU users never load it, and a process that already carries a libnccl.so.2
// (for instance the one bundled with PyTorch) shares that copy instead of loading a second one.
struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    std::string error;
};

NcclApi &nccl_api()
{
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {getenv("CCGPU_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char *nm : names) {
            if (!nm || !*nm) continue;
            api.handle = dlopen(nm, RTLD_NOW | RTLD_LOCAL);
            if (api.handle) break;
            api.error = dlerror();
        }
        if (!api.handle) { api.error = "cannot load NCCL (" + api.error + "); set CCGPU_NCCL_LIB"; return; }
        auto sym = [&](const char *nm) { void *p = dlsym(api.handle, nm); if (!p) api.error = std::string("NCCL lacks ") + nm; return p; };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
        if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy || !api.GetErrorString) {
            dlclose(api.handle);
            api.handle = nullptr;
        } else {
            api.error.clear();
        }
    });
    return api;
}

enum Ctr { CT_NFAM = 0, CT_MIXED = 1, CT_NSSCS = 2, CT_NDCS = 3, CT_NCORR = 4, CT_NDCSSC = 5,
           CT_FAULTS = 6, CT_ERR = 7, CT_TICKET0 = 8, CT_NMIDLIST = 28, CT_NMID = 29, CT_NBIG = 30, CT_NREDO = 31, CT_ASYM = 32, CT_ONESIDED = 33, CT_UNMIRRORED = 34, CT_BADCLASS = 35, CT_COUNT = 40 };

}  // namespace

struct cc_ctx {
    int device = 0;
    cc_params params{0.7, 1};
    std::string err;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    // input
    const uint4 *fields = nullptr;
    const uint8_t *payload = nullptr;
    const uint32_t *dup_bc = nullptr;
    uint64_t n = 0;
    uint32_t lpad = 0, stride = 0, n_bc = 0;
    DevBuf in_fields, in_payload, in_dup, in_q4, in_pairs, in_classes;
    // grouping scratch
    DevBuf key_a, key_b, val_a, val_b, head, radix_counts, scan_status, counters, hist, stats;
    uint32_t *perm = nullptr, *sorted_keys = nullptr, *spare_keys = nullptr, *spare_vals = nullptr;
    uint32_t next_ticket = 0;
    // families
    uint32_t n_fam = 0, n_sscs = 0, n_dcs = 0, n_corr = 0, n_dcssc = 0;   // n_dcssc = n_dcs + n_extra when rescuing
    uint32_t n_extra = 0;                                                 // second-pass pairs with a row of their own
    DevBuf bitmap, word_rank, first_read;
    DevBuf tmpl, c2hs, fam_start, fam_size, is_sscs, fam_mapq, fam_flag, fam_tlen, fam_rg;
    DevBuf partner, table, f_dcs, f_corr, f_dcssc;
    DevBuf sscs_list, dcs_list, corr_list, dcssc_list, sscs_row, dcs_row, corr_row, dcssc_row;
    DevBuf sscs_rows, corr_rows, dcs_rows;                 // dcs_rows: pool of all duplex rows (see do_duplex)
    DevBuf dcs_mapq, dcs_flag, dcs_tlen, dcs_rg;
    DevBuf vstart, vsize, big_list, mid_list, thr_table;
    uint8_t thr_host[128];
    ThrMagic thr_magic{0, 0, 1};
    bool thr_valid = false;
    DevBuf tmp_a, tmp_b, tmp_c;   // stand-alone primitives
    DevBuf cs_in, cs_k0, cs_k1, cs_v0, cs_v1;   // cc_sort_coordinates
    DevBuf gtable;                              // hash grouping (group_hash.cuh, group_tile.cuh)
    DevBuf cursor;                              // tile grouping: next free place of every family's segment
    DevBuf word_reads, word_start;              // ... per bitmap word: reads in the families founded there, and their prefix
    DevBuf gt_state;                            // ... device-side scalars + the contig stretch table (group_tile.cuh)
    enum GroupMode { GROUP_PAIR, GROUP_TILE, GROUP_HASH, GROUP_SORT };
    GroupMode group_mode = GROUP_PAIR;
    bool pair_join = false;                     // the loaded batch was grouped pair by pair (group_pair.cuh)
    bool tile_join = false;
    uint64_t fam_bound = 0;                     // grids / buffers of family-level work cover this many families: the count itself
                                                // (paths that learn it while grouping) or n (tile path: it stays on the device
                                                // until the one synchronisation that also brings the row counts)                     // the loaded batch was grouped by the tile path (bitmap + word ranks name families)
    uint64_t gtable_mask = 0;                   // non-zero while gtable describes the loaded batch
    bool have_reads = false, grouped = false, listed = false, voted = false, joined = false, rescued = false;
    bool read_fam_ready = false;
    bool dcssc_row_ready = true;                // CC_DCSSC_ROW holds ranks until somebody asks for it (materialise)
    // cc_pack_rows: the produced rows once more, in bit planes
    DevBuf planes_sscs, planes_corr, planes_dcs, alphabet;
    bool packed = false;
    uint32_t pack_qbits = 0, pack_stride = 0;
    uint8_t pack_qual[16] = {}, pack_base[16] = {};
    bool by_classes = false;                    // loaded through cc_load_reads_classes: in_classes ends with a "bad class index" word
    uint64_t redo_off = 0;                      // big_list[redo_off ..]: rows the bulk vote kernel hands to the exact one
    bool stats_clean = false;                   // the 16 stage counters are zero (no classification has added to them)
    bool segments_ready = false;                // vstart / vsize / big_list / mid_list were written by the grouping, no vote has touched them
    // counters read back while later kernels run (cc_run): a side stream, "counters final" / "copy done" events, pinned words
    cudaStream_t side = nullptr;
    cudaEvent_t snap_src[2] = {nullptr, nullptr}, snap_done[2] = {nullptr, nullptr};
    cudaEvent_t big_src = nullptr, big_done = nullptr;     // the vote of the largest families runs on the side stream
    uint32_t *pin_ctr = nullptr;                // [3][CT_COUNT]: two snapshots, read_counters
    uint32_t host_ctr[CT_COUNT];
    // bookkeeping
    bool profile = false;
    PhaseTimer phase[CC_PH_COUNT];
    uint64_t launches = 0;
    uint64_t dev_bytes = 0;
    // multi-GPU: the communicator of cc_comm_init (one rank per context)
    ncclComm_t comm = nullptr;
    int comm_rank = 0, comm_nranks = 1;
};

namespace {

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
            return CC_ERR_CUDA;                                                                    \
        }                                                                                          \
    } while (0)

#define TRY(expr)                                                                                  \
    do {                                                                                           \
        int rc_ = (expr);                                                                          \
        if (rc_ != CC_OK) return rc_;                                                              \
    } while (0)

#define LAUNCH(kernel, grid, block, ...)                                                           \
    do {                                                                                           \
        kernel<<<(grid), (block), 0, ctx->stream>>>(__VA_ARGS__);                                  \
        ctx->launches++;                                                                           \
        CK(cudaGetLastError());                                                                    \
    } while (0)

int fail(cc_ctx *ctx, int code, const char *msg)
{
    ctx->err = msg;
    return code;
}

int ensure(cc_ctx *ctx, DevBuf &b, size_t bytes)
{
    if (bytes == 0) bytes = 256;
    if (b.cap >= bytes) return CC_OK;
    if (b.p) {
        // a buffer that has to grow grows by a quarter more than asked, so that a stream of slowly growing
        // batches (region chunks of one sample) does not pay the synchronise + free + allocate every time
        bytes += bytes / 4;
        CK(cudaStreamSynchronize(ctx->stream));
        CK(cudaFree(b.p));
        ctx->dev_bytes -= b.cap;
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = (bytes + 255) & ~(size_t)255;
    CK(cudaMalloc(&b.p, want));
    b.cap = want;
    ctx->dev_bytes += want;
    return CC_OK;
}

void release(cc_ctx *ctx, DevBuf &b)
{
    if (b.p) cudaFree(b.p);
    ctx->dev_bytes -= b.cap;
    b.p = nullptr;
    b.cap = 0;
}

template <typename T> T *ptr(DevBuf &b) { return reinterpret_cast<T *>(b.p); }

inline unsigned blocks_for(uint64_t n, unsigned per) { return (unsigned)((n + per - 1) / per); }

int phase_begin(cc_ctx *ctx, int ph)
{
    if (!ctx->profile) return CC_OK;
    PhaseTimer &t = ctx->phase[ph];
    if (t.used + 2 > t.ev.size()) {
        cudaEvent_t a, b;
        CK(cudaEventCreate(&a));
        CK(cudaEventCreate(&b));
        t.ev.push_back(a);
        t.ev.push_back(b);
    }
    CK(cudaEventRecord(t.ev[t.used], ctx->stream));
    return CC_OK;
}

int phase_end(cc_ctx *ctx, int ph)
{
    if (!ctx->profile) return CC_OK;
    PhaseTimer &t = ctx->phase[ph];
    CK(cudaEventRecord(t.ev[t.used + 1], ctx->stream));
    t.used += 2;
    return CC_OK;
}

int collect_phases(cc_ctx *ctx)
{
    CK(cudaStreamSynchronize(ctx->stream));
    for (int ph = 0; ph < CC_PH_COUNT; ph++) {
        PhaseTimer &t = ctx->phase[ph];
        for (size_t i = 0; i + 1 < t.used; i += 2) {
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, t.ev[i], t.ev[i + 1]));
            t.ms += ms;
            t.runs++;
        }
        t.used = 0;
    }
    return CC_OK;
}

// ---- prefix-sum helpers ---------------------------------------------------------------
// Every chained scan draws its tile numbers from a counter of its own ("ticket"), zeroed with the other counters
// when a batch is grouped.  A batch whose later stages are run again and again (cc_set_params + the stage calls)
// uses the pool up: the scans queued so far are complete by the time anything queued now runs, so the pool is
// zeroed in stream order and handed out again.
int ticket_room(cc_ctx *ctx)
{
    if (ctx->next_ticket < CT_NMIDLIST - CT_TICKET0) return CC_OK;
    CK(cudaMemsetAsync(ptr<uint32_t>(ctx->counters) + CT_TICKET0, 0, (CT_NMIDLIST - CT_TICKET0) * sizeof(uint32_t), ctx->stream));
    ctx->next_ticket = 0;
    return CC_OK;
}

int scan_prepare(cc_ctx *ctx, uint64_t n, uint32_t **ticket)
{
    uint64_t tiles = scan_tiles(n);
    TRY(ensure(ctx, ctx->scan_status, (tiles + 1) * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(ctx->scan_status.p, 0, (tiles + 1) * sizeof(unsigned long long), ctx->stream));
    TRY(ticket_room(ctx));
    *ticket = ptr<uint32_t>(ctx->counters) + CT_TICKET0 + ctx->next_ticket++;
    return CC_OK;
}

int scan_exclusive(cc_ctx *ctx, const uint32_t *in, uint32_t *out, uint64_t n)
{
    if (n == 0) return CC_OK;
    uint32_t *ticket;
    TRY(scan_prepare(ctx, n, &ticket));
    auto kernel = scan_kernel<SCAN_MODE_EXCL, uint32_t>;
    LAUNCH(kernel, (unsigned)scan_tiles(n), SCAN_THREADS, in, out, (uint32_t *)nullptr, n,
           ptr<unsigned long long>(ctx->scan_status), ticket, (uint32_t *)nullptr, (int32_t *)nullptr,
           (const int32_t *)nullptr, (const uint32_t *)nullptr);
    return CC_OK;
}

// list[rank] = index of every non-zero flag; *total (device) = how many
// optionally also the inverse map: row_out[i] = *row_base + rank if flagged, else row_alt[i] (-1 without row_alt)
int compact_flags(cc_ctx *ctx, const uint8_t *flags, uint64_t n, uint32_t *list, uint32_t *total,
                  int32_t *row_out = nullptr, const int32_t *row_alt = nullptr, const uint32_t *row_base = nullptr)
{
    if (n == 0) return CC_OK;   // *total was zeroed with the counter block
    uint32_t *ticket;
    TRY(scan_prepare(ctx, n, &ticket));
    auto kernel = scan_kernel<SCAN_MODE_COMPACT, uint8_t>;
    LAUNCH(kernel, (unsigned)scan_tiles(n), SCAN_THREADS, flags, list, (uint32_t *)nullptr,
           n, ptr<unsigned long long>(ctx->scan_status), ticket, total, row_out, row_alt, row_base);
    return CC_OK;
}

// out[w] = number of set bits in words[0 .. w)
int scan_popcounts(cc_ctx *ctx, const uint32_t *words, uint32_t *out, uint64_t n, uint32_t *total = nullptr)
{
    if (n == 0) return CC_OK;
    uint32_t *ticket;
    TRY(scan_prepare(ctx, n, &ticket));
    auto kernel = scan_kernel<SCAN_MODE_POPC, uint32_t>;
    LAUNCH(kernel, (unsigned)scan_tiles(n), SCAN_THREADS, words, out, (uint32_t *)nullptr, n,
           ptr<unsigned long long>(ctx->scan_status), ticket, total, (int32_t *)nullptr, (const int32_t *)nullptr,
           (const uint32_t *)nullptr);
    return CC_OK;
}

int read_counters(cc_ctx *ctx)
{
    uint32_t *pinned = ctx->pin_ctr + 2 * CT_COUNT;          // (slots 0 and 1 belong to the side stream's snapshots)
    CK(cudaMemcpyAsync(pinned, ctx->counters.p, sizeof(ctx->host_ctr), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    memcpy(ctx->host_ctr, pinned, sizeof(ctx->host_ctr));
    if (ctx->host_ctr[CT_BADCLASS])
        return fail(ctx, CC_ERR_ARG, "cc_load_reads_classes: a pair record names a class beyond n_classes");
    if (ctx->host_ctr[CT_ERR])
        return fail(ctx, CC_ERR_BASE, "a base other than A/C/G/T carries Q>=30 inside a family "
                                      "(the reference raises IndexError/ValueError, SSCS_maker.py:119-122)");
    return CC_OK;
}

int sort_bits_for(uint64_t n)
{
    int bits = 8;
    while (bits < 32 && (1ull << bits) < n * 16) bits += 8;   // ~4 spare bits keep mixed runs rare
    return bits;
}

// ---- stage 1: grouping ----------------------------------------------------------------------
// Stable LSD sort of n (u32 key, u32 value) pairs on the low 8 * passes key bits (radix_onesweep.cuh).
// vals_in == nullptr: values start as 0..n-1.  The pairs ping-pong between (k0, v0) and (k1, v1);
// *k_out / *v_out say where the sorted ones ended up, *k_free / *v_free name the other buffer of each.
int onesweep_pairs(cc_ctx *ctx, uint32_t *k0, uint32_t *k1, const uint32_t *vals_in, uint32_t *v0, uint32_t *v1,
                   uint64_t n, int passes, uint32_t **k_out, uint32_t **v_out, uint32_t **k_free, uint32_t **v_free)
{
    const uint32_t ntiles = (uint32_t)os_tiles(n);
    TRY(ensure(ctx, ctx->radix_counts, ((size_t)ntiles * OS_BINS + OS_MAX_PASSES * OS_BINS + 64) * 4));
    uint32_t *ghist = ptr<uint32_t>(ctx->radix_counts);                 // [passes][256] digit totals, then bases
    uint32_t *tickets = ghist + OS_MAX_PASSES * OS_BINS;                // one tile ticket per pass
    uint32_t *status = tickets + 64;                                    // [tiles][256] look-back words
    CK(cudaMemsetAsync(ghist, 0, (OS_MAX_PASSES * OS_BINS + 64) * sizeof(uint32_t), ctx->stream));
    LAUNCH(os_histogram_kernel, 148 * 4, 256, k0, n, passes, ghist);
    LAUNCH(os_scan_hist_kernel, 1, 32 * OS_MAX_PASSES, ghist, passes);
    uint32_t *kin = k0, *kout = k1;
    const uint32_t *vin = vals_in;
    uint32_t *vout = v0, *vother = v1;
    if (vals_in == v0) { vout = v1; vother = v0; }
    for (int pass = 0; pass < passes; pass++) {
        CK(cudaMemsetAsync(status, 0, (size_t)ntiles * OS_BINS * sizeof(uint32_t), ctx->stream));
        LAUNCH(os_pass_kernel, ntiles, OS_THREADS, kin, vin, kout, vout, n, 8 * pass, ghist + pass * OS_BINS, status,
               tickets + pass);
        uint32_t *t = kin; kin = kout; kout = t;
        vin = vout; t = vout; vout = vother; vother = t;
    }
    *k_out = kin; *k_free = kout;
    *v_out = const_cast<uint32_t *>(vin); *v_free = vout;
    return CC_OK;
}

// Grouping through the hash table of group_hash.cuh; leaves exactly what the sort path leaves
// (families in creation order, segments of `perm`, templates, singleton fields, class bytes,
// size histogram, SSCS list) plus read -> family, which the sort path only builds on demand.
int do_group_hash(cc_ctx *ctx)
{
    const uint64_t n = ctx->n;
    uint32_t *ctr = ptr<uint32_t>(ctx->counters);
    TRY(ensure(ctx, ctx->key_a, n * 4)); TRY(ensure(ctx, ctx->key_b, n * 4));
    TRY(ensure(ctx, ctx->val_a, n * 4)); TRY(ensure(ctx, ctx->val_b, n * 4));
    uint32_t *read_slot = ptr<uint32_t>(ctx->key_a), *cursor = ptr<uint32_t>(ctx->key_b);   // cursor: one per family
    uint64_t slots = 1024;
    while (slots < n + n / 4) slots <<= 1;                  // load factor <= 0.8 even if every read is a family
    TRY(ensure(ctx, ctx->gtable, slots * sizeof(GroupSlot)));
    GroupSlot *table = ptr<GroupSlot>(ctx->gtable);
    const uint64_t words = (n + 31) / 32;
    TRY(ensure(ctx, ctx->bitmap, words * 4)); TRY(ensure(ctx, ctx->word_rank, words * 4));

    TRY(phase_begin(ctx, CC_PH_HASH));
    CK(cudaMemsetAsync(table, 0xFF, slots * sizeof(GroupSlot), ctx->stream));
    LAUNCH(group_insert_kernel, blocks_for(n, 256), 256, ctx->fields, n, table, slots - 1, read_slot);
    TRY(phase_end(ctx, CC_PH_HASH));

    TRY(phase_begin(ctx, CC_PH_SEGMENT));
    LAUNCH(group_mark_kernel, blocks_for(n, 256), 256, table, read_slot, n, ptr<uint32_t>(ctx->bitmap),
           ptr<uint32_t>(ctx->val_b));
    TRY(scan_popcounts(ctx, ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank), words, ctr + CT_NFAM));
    TRY(read_counters(ctx));
    const uint32_t nf = ctx->n_fam = ctx->host_ctr[CT_NFAM];

    TRY(ensure(ctx, ctx->tmpl, (size_t)nf * 4));
    TRY(ensure(ctx, ctx->fam_start, (size_t)nf * 4)); TRY(ensure(ctx, ctx->fam_size, (size_t)nf * 4));
    TRY(ensure(ctx, ctx->is_sscs, (size_t)nf + 4)); TRY(ensure(ctx, ctx->fam_mapq, nf)); TRY(ensure(ctx, ctx->fam_rg, nf));
    TRY(ensure(ctx, ctx->fam_flag, (size_t)nf * 2)); TRY(ensure(ctx, ctx->fam_tlen, (size_t)nf * 4));
    TRY(ensure(ctx, ctx->sscs_list, (size_t)nf * 4)); TRY(ensure(ctx, ctx->partner, (size_t)nf * 4));
    LAUNCH(group_family_kernel, (unsigned)std::min<uint64_t>(blocks_for(n, 256), 148 * 16), 256, ctx->fields, read_slot, n,
           ptr<uint32_t>(ctx->bitmap),
           ptr<uint32_t>(ctx->word_rank), ptr<uint32_t>(ctx->val_b), ptr<uint32_t>(ctx->tmpl), ptr<uint32_t>(ctx->fam_size),
           ptr<uint8_t>(ctx->is_sscs), ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag),
           ptr<int32_t>(ctx->fam_tlen), ptr<uint8_t>(ctx->fam_rg), ptr<unsigned long long>(ctx->hist),
           (uint32_t)CC_HIST_BINS, ctr + CT_NMID, ptr<int32_t>(ctx->partner));
    TRY(scan_exclusive(ctx, ptr<uint32_t>(ctx->fam_size), ptr<uint32_t>(ctx->fam_start), nf));
    ctx->perm = ptr<uint32_t>(ctx->val_a);
    ctx->spare_vals = ptr<uint32_t>(ctx->val_b);           // read -> family
    ctx->sorted_keys = ctx->spare_keys = nullptr;
    CK(cudaMemcpyAsync(cursor, ctx->fam_start.p, (size_t)nf * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    LAUNCH(group_place_kernel, blocks_for(n, 256), 256, ptr<uint32_t>(ctx->bitmap), n, cursor, ctx->perm, ctx->spare_vals);
    ctx->read_fam_ready = true;
    ctx->gtable_mask = slots - 1;
    ctx->fam_bound = nf;
    LAUNCH(domain_check_kernel, blocks_for(nf, 256), 256, ptr<uint32_t>(ctx->tmpl), ptr<uint32_t>(ctx->fam_size), ctr + CT_NFAM,
           ctr + CT_FAULTS);
    TRY(ensure(ctx, ctx->sscs_row, (size_t)nf * 4));
    TRY(compact_flags(ctx, ptr<uint8_t>(ctx->is_sscs), nf, ptr<uint32_t>(ctx->sscs_list), ctr + CT_NSSCS,
                      ptr<int32_t>(ctx->sscs_row)));
    TRY(phase_end(ctx, CC_PH_SEGMENT));
    ctx->grouped = true;
    return CC_OK;
}

// Grouping tile by tile (group_tile.cuh): the default.  Leaves what the other two paths leave.  Family-level
// arrays written before the family count reaches the host are sized by their bound, n.
int do_group_tile(cc_ctx *ctx)
{
    const uint64_t n = ctx->n;
    uint32_t *ctr = ptr<uint32_t>(ctx->counters);
    TRY(ensure(ctx, ctx->key_a, n * 4)); TRY(ensure(ctx, ctx->key_b, n * 4));
    TRY(ensure(ctx, ctx->val_a, n * 4)); TRY(ensure(ctx, ctx->val_b, n * 4));
    uint32_t *link = ptr<uint32_t>(ctx->key_a), *aux = ptr<uint32_t>(ctx->key_b);
    uint32_t *stash = ptr<uint32_t>(ctx->val_b);            // group / family counter words; dead once the family records are written
    uint32_t *edge_list = ptr<uint32_t>(ctx->val_a);        // dead before tile_place writes `perm` there
    uint64_t slots = 1024;
    while (slots < n + n / 4) slots <<= 1;                  // the most the device may use: every read a listed family, load <= 0.8
    TRY(ensure(ctx, ctx->gtable, slots * sizeof(GroupSlot)));
    GroupSlot *table = ptr<GroupSlot>(ctx->gtable);
    const uint64_t words = (n + 31) / 32, tiles = gt_tiles(n);
    TRY(ensure(ctx, ctx->bitmap, words * 4)); TRY(ensure(ctx, ctx->word_rank, words * 4));
    TRY(ensure(ctx, ctx->word_reads, words * 4)); TRY(ensure(ctx, ctx->word_start, words * 4));
    TRY(ensure(ctx, ctx->tmpl, n * 4)); TRY(ensure(ctx, ctx->fam_start, n * 4)); TRY(ensure(ctx, ctx->fam_size, n * 4));
    TRY(ensure(ctx, ctx->cursor, n * 4)); TRY(ensure(ctx, ctx->is_sscs, n + 4)); TRY(ensure(ctx, ctx->fam_mapq, n));
    TRY(ensure(ctx, ctx->fam_rg, n)); TRY(ensure(ctx, ctx->fam_flag, n * 2)); TRY(ensure(ctx, ctx->fam_tlen, n * 4));
    TRY(ensure(ctx, ctx->partner, n * 4));
    TRY(ensure(ctx, ctx->sscs_list, (n / 2 + 1) * 4)); TRY(ensure(ctx, ctx->sscs_row, n * 4));
    const uint64_t ws_tiles = (words + WS_TILE - 1) / WS_TILE;
    TRY(ensure(ctx, ctx->scan_status, (2 * ws_tiles + 2) * sizeof(unsigned long long)));
    TRY(ensure(ctx, ctx->gt_state, (GTC_COUNT + STRETCH_SLOTS) * sizeof(uint32_t)));
    uint32_t *gtc = ptr<uint32_t>(ctx->gt_state), *stretch = gtc + GTC_COUNT;

    TRY(phase_begin(ctx, CC_PH_HASH));
    CK(cudaMemsetAsync(gtc, 0, (GTC_COUNT + STRETCH_SLOTS) * sizeof(uint32_t), ctx->stream));
    if (ctx->n_bc) LAUNCH(dup_symmetry_kernel, blocks_for(ctx->n_bc, 256), 256, ctx->dup_bc, ctx->n_bc, ctr + CT_ASYM);
    LAUNCH(tile_local_kernel<0>, (unsigned)tiles, GT_THREADS, ctx->fields, n, ctx->dup_bc, ctx->n_bc, ctr + CT_ASYM, gtc, stretch,
           link, aux, stash, edge_list, ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_reads), ptr<uint32_t>(ctx->word_start));
    // input out of order (flag raised on the device): the same tiles again, every family through the table
    LAUNCH(tile_redo_kernel, 1, 1, gtc);
    LAUNCH(tile_local_kernel<1>, (unsigned)std::min<uint64_t>(tiles, 148 * 4), GT_THREADS, ctx->fields, n, ctx->dup_bc, ctx->n_bc, ctr + CT_ASYM, gtc, stretch,
           link, aux, stash, edge_list, ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_reads), ptr<uint32_t>(ctx->word_start));
    LAUNCH(table_prepare_kernel, 148 * 4, 256, gtc, table, slots);
    LAUNCH(edge_insert_kernel, 148 * 8, 256, ctx->fields, edge_list, gtc, table, stash, aux);
    TRY(phase_end(ctx, CC_PH_HASH));

    TRY(phase_begin(ctx, CC_PH_SEGMENT));
    // word_reads -> segment starts, word_start (until here: SSCS families per word) -> consensus rows, both in place
    uint32_t *word_sscs = ptr<uint32_t>(ctx->word_start);
    LAUNCH(edge_mark_kernel, 148 * 8, 256, edge_list, gtc, table, aux, stash, ptr<uint32_t>(ctx->bitmap),
           ptr<uint32_t>(ctx->word_reads), word_sscs);
    CK(cudaMemsetAsync(ctx->scan_status.p, 0, (2 * ws_tiles + 2) * sizeof(unsigned long long), ctx->stream));
    TRY(ticket_room(ctx));
    LAUNCH(word_scan_kernel, (unsigned)ws_tiles, WS_THREADS, ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank),
           ptr<uint32_t>(ctx->word_reads), word_sscs, words, ptr<unsigned long long>(ctx->scan_status),
           ctr + CT_TICKET0 + ctx->next_ticket++, ctr + CT_NFAM, ctr + CT_NSSCS);
    LAUNCH(tile_family_kernel, (unsigned)std::min<uint64_t>(blocks_for(n, 256), 148 * 16), 256, ctx->fields, n, stash,
           ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank), ptr<uint32_t>(ctx->word_reads),
           ptr<uint32_t>(ctx->tmpl), ptr<uint32_t>(ctx->fam_size), ptr<uint32_t>(ctx->fam_start), ptr<uint32_t>(ctx->cursor),
           ptr<uint8_t>(ctx->is_sscs), ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag), ptr<int32_t>(ctx->fam_tlen),
           ptr<uint8_t>(ctx->fam_rg), ptr<int32_t>(ctx->partner), ptr<unsigned long long>(ctx->hist), (uint32_t)CC_HIST_BINS,
           ctr + CT_NMID, word_sscs, ptr<uint32_t>(ctx->sscs_list), ptr<int32_t>(ctx->sscs_row));
    ctx->perm = ptr<uint32_t>(ctx->val_a);
    ctx->spare_vals = ptr<uint32_t>(ctx->val_b);           // read -> family, materialised when somebody fetches it
    ctx->sorted_keys = ctx->spare_keys = nullptr;
    LAUNCH(tile_place_kernel, (unsigned)tiles, GP_THREADS, n, link, aux, ptr<uint32_t>(ctx->bitmap),
           ptr<uint32_t>(ctx->word_rank), ptr<uint32_t>(ctx->cursor), ctx->perm);
    ctx->n_fam = 0;                                          // known after do_lists
    ctx->fam_bound = n;
    ctx->read_fam_ready = false;
    ctx->gtable_mask = slots - 1;
    ctx->tile_join = true;
    LAUNCH(domain_check_kernel, (unsigned)std::min<uint64_t>(blocks_for(n, 256), 148 * 32), 256, ptr<uint32_t>(ctx->tmpl), ptr<uint32_t>(ctx->fam_size), ctr + CT_NFAM,
           ctr + CT_FAULTS);
    TRY(phase_end(ctx, CC_PH_SEGMENT));
    ctx->grouped = true;
    return CC_OK;
}

// Grouping pair by pair (group_pair.cuh): the default.  Leaves what the read-level paths leave.
int do_group_pair(cc_ctx *ctx)
{
    const uint64_t n = ctx->n, np = n / 2;
    uint32_t *ctr = ptr<uint32_t>(ctx->counters);
    TRY(ensure(ctx, ctx->key_a, n * 4)); TRY(ensure(ctx, ctx->key_b, n * 4));
    TRY(ensure(ctx, ctx->val_a, n * 4)); TRY(ensure(ctx, ctx->val_b, n * 4));
    uint32_t *plink = ptr<uint32_t>(ctx->key_a), *paux = ptr<uint32_t>(ctx->key_b);
    uint32_t *pstash = ptr<uint32_t>(ctx->val_b);           // group / family words; dead once the family records are written
    uint32_t *edge_list = ptr<uint32_t>(ctx->val_a);        // dead before pair_place writes `perm` there
    uint64_t slots = 1024;
    while (slots < np + np / 4) slots <<= 1;                // the most the device may use: every pair a listed family
    TRY(ensure(ctx, ctx->gtable, slots * sizeof(GroupSlot)));
    GroupSlot *table = ptr<GroupSlot>(ctx->gtable);
    const uint64_t words = (np + 31) / 32, tiles = pg_tiles(np);
    TRY(ensure(ctx, ctx->bitmap, words * 4)); TRY(ensure(ctx, ctx->word_rank, words * 4));
    TRY(ensure(ctx, ctx->word_reads, words * 4)); TRY(ensure(ctx, ctx->word_start, words * 4));
    TRY(ensure(ctx, ctx->tmpl, n * 4)); TRY(ensure(ctx, ctx->fam_start, n * 4)); TRY(ensure(ctx, ctx->fam_size, n * 4));
    TRY(ensure(ctx, ctx->cursor, np * 4 + 4)); TRY(ensure(ctx, ctx->is_sscs, n + 4)); TRY(ensure(ctx, ctx->fam_mapq, n + 4));
    TRY(ensure(ctx, ctx->fam_rg, n + 4)); TRY(ensure(ctx, ctx->fam_flag, n * 2 + 4)); TRY(ensure(ctx, ctx->fam_tlen, n * 4));
    TRY(ensure(ctx, ctx->partner, n * 4));
    TRY(ensure(ctx, ctx->sscs_list, (n / 2 + 2) * 4)); TRY(ensure(ctx, ctx->sscs_row, n * 4));
    // the vote's per-row segments and its two lists of larger families are written with the family records
    TRY(ensure(ctx, ctx->vstart, (n / 2 + 2) * 4)); TRY(ensure(ctx, ctx->vsize, (n / 2 + 2) * 4));
    TRY(ensure(ctx, ctx->big_list, 2 * (n / 2 + 2) * 4)); TRY(ensure(ctx, ctx->mid_list, (n / 16 + 4) * 4));
    const uint64_t ws_tiles = (words + WS_TILE - 1) / WS_TILE;
    TRY(ensure(ctx, ctx->scan_status, (2 * ws_tiles + 2) * sizeof(unsigned long long)));
    TRY(ensure(ctx, ctx->gt_state, (GTC_COUNT + STRETCH_SLOTS) * sizeof(uint32_t)));
    uint32_t *gpc = ptr<uint32_t>(ctx->gt_state), *stretch = gpc + GTC_COUNT;
    uint32_t *word_pairs = ptr<uint32_t>(ctx->word_reads), *word_sscs = ptr<uint32_t>(ctx->word_start);

    TRY(phase_begin(ctx, CC_PH_HASH));
    CK(cudaMemsetAsync(gpc, 0, (GTC_COUNT + STRETCH_SLOTS) * sizeof(uint32_t), ctx->stream));
    LAUNCH(pair_local_kernel<0>, (unsigned)tiles, PG_THREADS, ctx->fields, np, ctx->dup_bc, ctx->n_bc, gpc, stretch, ctr + CT_FAULTS,
           ctr + CT_UNMIRRORED, plink, paux, pstash, edge_list, ptr<uint32_t>(ctx->bitmap), word_pairs, word_sscs);
    // input out of order (flag raised on the device): the same tiles again, every family through the table
    LAUNCH(pair_redo_kernel, 1, 1, gpc);
    LAUNCH(pair_local_kernel<1>, (unsigned)std::min<uint64_t>(tiles, 148 * 8), PG_THREADS, ctx->fields, np, ctx->dup_bc, ctx->n_bc, gpc,
           stretch, ctr + CT_FAULTS, ctr + CT_UNMIRRORED, plink, paux, pstash, edge_list, ptr<uint32_t>(ctx->bitmap), word_pairs, word_sscs);
    LAUNCH(pair_table_prepare_kernel, 148 * 4, 256, gpc, table, slots);
    LAUNCH(pair_edge_insert_kernel, 148 * 8, 256, ctx->fields, edge_list, gpc, table, pstash, paux, ctr + CT_FAULTS);
    TRY(phase_end(ctx, CC_PH_HASH));

    TRY(phase_begin(ctx, CC_PH_SEGMENT));
    LAUNCH(pair_edge_mark_kernel, 148 * 8, 256, edge_list, gpc, table, paux, pstash, ptr<uint32_t>(ctx->bitmap), word_pairs, word_sscs);
    CK(cudaMemsetAsync(ctx->scan_status.p, 0, (2 * ws_tiles + 2) * sizeof(unsigned long long), ctx->stream));
    TRY(ticket_room(ctx));
    // word_pairs -> segment starts (in pairs), word_sscs -> consensus row pairs, both in place; the totals count read families
    LAUNCH(word_scan_kernel, (unsigned)ws_tiles, WS_THREADS, ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank), word_pairs,
           word_sscs, words, ptr<unsigned long long>(ctx->scan_status), ctr + CT_TICKET0 + ctx->next_ticket++, ctr + CT_NFAM,
           ctr + CT_NSSCS, 2u);
    LAUNCH(pair_family_kernel, (unsigned)std::min<uint64_t>(blocks_for(np, 256), 148 * 16), 256, ctx->fields, np, pstash,
           ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank), word_pairs, word_sscs, ptr<uint32_t>(ctx->tmpl),
           ptr<uint32_t>(ctx->fam_size), ptr<uint32_t>(ctx->fam_start), ptr<uint32_t>(ctx->cursor), ptr<uint8_t>(ctx->is_sscs),
           ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag), ptr<int32_t>(ctx->fam_tlen), ptr<uint8_t>(ctx->fam_rg),
           ptr<unsigned long long>(ctx->hist), (uint32_t)CC_HIST_BINS, ctr + CT_NMID, ptr<uint32_t>(ctx->sscs_list),
           ptr<int32_t>(ctx->sscs_row), ptr<uint32_t>(ctx->vstart), ptr<uint32_t>(ctx->vsize), ptr<uint32_t>(ctx->big_list),
           ctr + CT_NBIG, ctx->lpad <= 512 ? VOTE_SMALL_MAX : 0u, ptr<uint32_t>(ctx->mid_list), ctr + CT_NMIDLIST);
    ctx->segments_ready = true;
    ctx->perm = ptr<uint32_t>(ctx->val_a);
    ctx->spare_vals = ptr<uint32_t>(ctx->val_b);           // read -> family, materialised when somebody fetches it
    ctx->sorted_keys = ctx->spare_keys = nullptr;
    LAUNCH(pair_place_kernel, (unsigned)tiles, PG_THREADS, np, plink, paux, ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank),
           ptr<uint32_t>(ctx->fam_start), ptr<uint32_t>(ctx->fam_size), ptr<uint32_t>(ctx->cursor), ctx->perm);
    ctx->n_fam = 0;                                          // known after do_lists
    ctx->fam_bound = n;
    ctx->read_fam_ready = false;
    ctx->gtable_mask = slots - 1;
    ctx->pair_join = true;
    TRY(phase_end(ctx, CC_PH_SEGMENT));
    ctx->grouped = true;
    return CC_OK;
}

int do_group(cc_ctx *ctx)
{
    const uint64_t n = ctx->n;
    ctx->packed = false;
    ctx->grouped = ctx->listed = ctx->voted = ctx->joined = ctx->rescued = ctx->read_fam_ready = false;
    ctx->n_fam = ctx->n_sscs = ctx->n_dcs = ctx->n_corr = ctx->n_dcssc = ctx->n_extra = 0;
    ctx->fam_bound = 0;
    ctx->next_ticket = 0;
    ctx->gtable_mask = 0;
    ctx->tile_join = false;
    ctx->pair_join = false;
    ctx->segments_ready = false;
    TRY(ensure(ctx, ctx->counters, CT_COUNT * sizeof(uint32_t)));
    // [16 stage counters | CC_HIST_BINS family-size bins]: one buffer, so that cc_stats_allreduce is one collective
    TRY(ensure(ctx, ctx->stats, (16 + (size_t)CC_HIST_BINS) * sizeof(unsigned long long)));
    ctx->hist.p = ptr<unsigned long long>(ctx->stats) + 16;                // a view, never released on its own
    ctx->hist.cap = 0;
    CK(cudaMemsetAsync(ctx->counters.p, 0, CT_COUNT * sizeof(uint32_t), ctx->stream));
    if (ctx->by_classes)        // what the expansion of the short pair records found (the word behind the class table)
        CK(cudaMemcpyAsync(ptr<uint32_t>(ctx->counters) + CT_BADCLASS, ptr<uint8_t>(ctx->in_classes) + ctx->in_classes.cap - 4, 4,
                           cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaMemsetAsync(ctx->stats.p, 0, (16 + (size_t)CC_HIST_BINS) * sizeof(unsigned long long), ctx->stream));
    ctx->stats_clean = true;
    memset(ctx->host_ctr, 0, sizeof(ctx->host_ctr));
    if (n == 0) { ctx->grouped = true; return CC_OK; }
    if (ctx->group_mode == cc_ctx::GROUP_PAIR) return do_group_pair(ctx);
    if (ctx->group_mode == cc_ctx::GROUP_TILE) return do_group_tile(ctx);
    if (ctx->group_mode == cc_ctx::GROUP_HASH) return do_group_hash(ctx);

    TRY(ensure(ctx, ctx->key_a, n * 4)); TRY(ensure(ctx, ctx->key_b, n * 4));
    TRY(ensure(ctx, ctx->val_a, n * 4)); TRY(ensure(ctx, ctx->val_b, n * 4));
    TRY(ensure(ctx, ctx->head, n));
    const int bits = sort_bits_for(n);
    const uint32_t key_mask = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1u);
    TRY(phase_begin(ctx, CC_PH_HASH));
    LAUNCH(hash_kernel, blocks_for(n, 256), 256, ctx->fields, n, key_mask, ptr<uint32_t>(ctx->key_a));
    TRY(phase_end(ctx, CC_PH_HASH));

    TRY(phase_begin(ctx, CC_PH_SORT));
    TRY(onesweep_pairs(ctx, ptr<uint32_t>(ctx->key_a), ptr<uint32_t>(ctx->key_b), nullptr, ptr<uint32_t>(ctx->val_a),
                       ptr<uint32_t>(ctx->val_b), n, bits / 8, &ctx->sorted_keys, &ctx->perm, &ctx->spare_keys,
                       &ctx->spare_vals));
    TRY(phase_end(ctx, CC_PH_SORT));

    TRY(phase_begin(ctx, CC_PH_SEGMENT));
    uint32_t *ctr = ptr<uint32_t>(ctx->counters);
    uint32_t *mixed_list = ctx->spare_keys;        // dead after the sort; at most n/2 runs
    LAUNCH(boundary_kernel, blocks_for(n, 256), 256, ctx->fields, ctx->sorted_keys, ctx->perm, n,
           ptr<uint8_t>(ctx->head), mixed_list, ctr + CT_MIXED, (uint32_t)(n / 2 + 1));
    // collided runs (if any) are regrouped by however many warps the device-side count asks for
    LAUNCH(fix_mixed_runs_kernel, 592, 128, ctx->fields, ctx->sorted_keys, ctx->perm, ctx->spare_vals, n,
           ptr<uint8_t>(ctx->head), mixed_list, ctr + CT_MIXED, (uint32_t)(n / 2 + 1));
    uint32_t *hs_start = ctx->spare_keys;          // reused again: start of every family in sorted order
    TRY(compact_flags(ctx, ptr<uint8_t>(ctx->head), n, hs_start, ctr + CT_NFAM));
    TRY(read_counters(ctx));
    const uint32_t nf = ctx->n_fam = ctx->host_ctr[CT_NFAM];

    TRY(ensure(ctx, ctx->tmpl, (size_t)nf * 4));
    TRY(ensure(ctx, ctx->fam_start, (size_t)nf * 4)); TRY(ensure(ctx, ctx->fam_size, (size_t)nf * 4));
    TRY(ensure(ctx, ctx->is_sscs, nf)); TRY(ensure(ctx, ctx->fam_mapq, nf)); TRY(ensure(ctx, ctx->fam_rg, nf));
    TRY(ensure(ctx, ctx->fam_flag, (size_t)nf * 2)); TRY(ensure(ctx, ctx->fam_tlen, (size_t)nf * 4));
    TRY(ensure(ctx, ctx->sscs_list, (size_t)nf * 4)); TRY(ensure(ctx, ctx->partner, (size_t)nf * 4));

    // creation order: bitmap of first members + rank queries (see family_build_kernel)
    const uint64_t words = (n + 31) / 32;
    TRY(ensure(ctx, ctx->bitmap, words * 4)); TRY(ensure(ctx, ctx->word_rank, words * 4));
    TRY(ensure(ctx, ctx->first_read, (size_t)nf * 4));
    CK(cudaMemsetAsync(ctx->bitmap.p, 0, words * 4, ctx->stream));
    LAUNCH(mark_first_kernel, blocks_for(nf, 256), 256, hs_start, ctx->perm, nf, ptr<uint32_t>(ctx->bitmap),
           ptr<uint32_t>(ctx->first_read));
    TRY(scan_popcounts(ctx, ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank), words));
    TRY(ensure(ctx, ctx->c2hs, (size_t)nf * 4));
    LAUNCH(creation_rank_kernel, blocks_for(nf, 256), 256, ptr<uint32_t>(ctx->first_read), ptr<uint32_t>(ctx->bitmap),
           ptr<uint32_t>(ctx->word_rank), nf, ptr<uint32_t>(ctx->c2hs));
    LAUNCH(family_build_kernel, blocks_for(nf, 256), 256, ctx->fields, hs_start, ptr<uint32_t>(ctx->first_read),
           ptr<uint32_t>(ctx->c2hs), ptr<uint8_t>(ctx->head), nf, (uint32_t)n,
           ptr<uint32_t>(ctx->tmpl), ptr<uint32_t>(ctx->fam_start), ptr<uint32_t>(ctx->fam_size),
           ptr<uint8_t>(ctx->is_sscs), ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag),
           ptr<int32_t>(ctx->fam_tlen), ptr<uint8_t>(ctx->fam_rg), ptr<unsigned long long>(ctx->hist),
           (uint32_t)CC_HIST_BINS, ctr + CT_NMID, ptr<int32_t>(ctx->partner));
    ctx->fam_bound = nf;
    LAUNCH(domain_check_kernel, blocks_for(nf, 256), 256, ptr<uint32_t>(ctx->tmpl), ptr<uint32_t>(ctx->fam_size), ctr + CT_NFAM,
           ctr + CT_FAULTS);
    TRY(ensure(ctx, ctx->sscs_row, (size_t)nf * 4));
    TRY(compact_flags(ctx, ptr<uint8_t>(ctx->is_sscs), nf, ptr<uint32_t>(ctx->sscs_list), ctr + CT_NSSCS,
                      ptr<int32_t>(ctx->sscs_row)));
    TRY(phase_end(ctx, CC_PH_SEGMENT));
    ctx->grouped = true;
    return CC_OK;
}

// ---- stage 2: join + classification (index work only) -----------------------------------------
int do_classify(cc_ctx *ctx);

int do_join_index(cc_ctx *ctx)
{
    const uint64_t bound = ctx->fam_bound;
    const uint32_t nf = ctx->n_fam;
    uint32_t *ctr = ptr<uint32_t>(ctx->counters);
    TRY(ensure(ctx, ctx->partner, bound * 4));
    TRY(ensure(ctx, ctx->dcs_list, bound * 4)); TRY(ensure(ctx, ctx->corr_list, bound * 4));
    TRY(ensure(ctx, ctx->dcssc_list, bound * 4));
    if (bound == 0) return CC_OK;
    TRY(phase_begin(ctx, CC_PH_JOIN));
    if (ctx->gtable_mask && ctx->pair_join) {
        LAUNCH(pair_join_kernel, (unsigned)std::min<uint64_t>(blocks_for(bound / 2 + 1, 256), 148 * 32), 256, ctx->fields,
               ptr<uint32_t>(ctx->tmpl), ctr + CT_NFAM, ctx->dup_bc, ctx->n_bc, ptr<GroupSlot>(ctx->gtable), ptr<uint32_t>(ctx->gt_state),
               ptr<uint32_t>(ctx->key_a), ptr<uint32_t>(ctx->key_b), ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank),
               ptr<int32_t>(ctx->partner));
        TRY(phase_end(ctx, CC_PH_JOIN));
        return do_classify(ctx);
    }
    if (ctx->gtable_mask && ctx->tile_join) {
        // (the barcode swap's symmetry was established before the tiles were grouped)
        LAUNCH(tile_join_kernel, (unsigned)std::min<uint64_t>(blocks_for(bound, 256), 148 * 32), 256, ctx->fields, ptr<uint32_t>(ctx->tmpl), ctr + CT_NFAM, ctx->dup_bc, ctx->n_bc,
               ptr<GroupSlot>(ctx->gtable), ptr<uint32_t>(ctx->gt_state), ptr<uint32_t>(ctx->key_a), ptr<uint32_t>(ctx->key_b),
               ptr<uint32_t>(ctx->bitmap), ptr<uint32_t>(ctx->word_rank), ptr<int32_t>(ctx->partner), ctr + CT_ASYM,
               ptr<uint16_t>(ctx->fam_flag));
        TRY(phase_end(ctx, CC_PH_JOIN));
        return do_classify(ctx);
    }
    CK(cudaMemsetAsync(ctr + CT_ASYM, 0, 4, ctx->stream));
    if (ctx->n_bc) LAUNCH(dup_symmetry_kernel, blocks_for(ctx->n_bc, 256), 256, ctx->dup_bc, ctx->n_bc, ctr + CT_ASYM);
    if (ctx->gtable_mask) {
        // the grouping table already maps every family tuple to its first read
        LAUNCH(group_join_kernel, blocks_for(nf, 256), 256, ctx->fields, ptr<uint32_t>(ctx->tmpl), nf, ctx->dup_bc, ctx->n_bc,
               ptr<GroupSlot>(ctx->gtable), ctx->gtable_mask, ctx->spare_vals, ptr<int32_t>(ctx->partner), ctr + CT_ASYM);
        TRY(phase_end(ctx, CC_PH_JOIN));
        return do_classify(ctx);
    }
    uint64_t slots = 64;
    while (slots < 2ull * nf) slots <<= 1;
    TRY(ensure(ctx, ctx->table, slots * 8));
    CK(cudaMemsetAsync(ctx->table.p, 0, slots * 8, ctx->stream));
    LAUNCH(join_build_kernel, blocks_for(nf, 256), 256, ctx->fields, ptr<uint32_t>(ctx->tmpl), nf,
           ptr<unsigned long long>(ctx->table), slots - 1, ctr + CT_ASYM);
    LAUNCH(join_probe_kernel, blocks_for(nf, 256), 256, ctx->fields, ptr<uint32_t>(ctx->tmpl), nf, ctx->dup_bc,
           ctx->n_bc, ptr<unsigned long long>(ctx->table), slots - 1, ptr<int32_t>(ctx->partner), ctr + CT_ASYM);
    TRY(phase_end(ctx, CC_PH_JOIN));
    return do_classify(ctx);
}

int do_classify(cc_ctx *ctx)
{
    const uint64_t nf = ctx->fam_bound;
    uint32_t *ctr = ptr<uint32_t>(ctx->counters);
    TRY(phase_begin(ctx, CC_PH_CLASSIFY));
    // one chained scan classifies the families and writes the three lists with their family -> row maps;
    // in the second DCS pass the SSCS-SSCS pairs keep their DCS row and the pairs with a corrected
    // singleton follow in the same pool (dcssc_row is finished in do_lists, once n_dcs is known)
    TRY(ensure(ctx, ctx->dcs_row, (size_t)nf * 4)); TRY(ensure(ctx, ctx->corr_row, (size_t)nf * 4));
    TRY(ensure(ctx, ctx->dcssc_row, (size_t)nf * 4));
    const uint64_t tiles = ((uint64_t)nf + CLS_TILE - 1) / CLS_TILE;
    TRY(ensure(ctx, ctx->scan_status, (2 * tiles + 2) * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(ctx->scan_status.p, 0, (2 * tiles + 2) * sizeof(unsigned long long), ctx->stream));
    if (!ctx->stats_clean)      // cc_dcs_join again on one batch (after cc_set_params): the stage counters are counted afresh
        CK(cudaMemsetAsync(ctx->stats.p, 0, 16 * sizeof(unsigned long long), ctx->stream));
    ctx->stats_clean = false;
    TRY(ticket_room(ctx));
    uint32_t *ticket = ctr + CT_TICKET0 + ctx->next_ticket++;
    LAUNCH(classify_scan_kernel, (unsigned)tiles, CLS_THREADS, ptr<uint32_t>(ctx->fam_size), ptr<int32_t>(ctx->partner), ctr + CT_NFAM,
           (int)ctx->params.scorrect, ptr<unsigned long long>(ctx->scan_status), ticket, ptr<uint32_t>(ctx->dcs_list),
           ptr<uint32_t>(ctx->corr_list), ptr<uint32_t>(ctx->dcssc_list), ptr<int32_t>(ctx->dcs_row), ptr<int32_t>(ctx->corr_row),
           ptr<int32_t>(ctx->dcssc_row), ptr<unsigned long long>(ctx->stats), ctr + CT_ONESIDED, ctr + CT_NDCS, ctr + CT_NCORR,
           ctr + CT_NDCSSC);
    TRY(phase_end(ctx, CC_PH_CLASSIFY));
    return CC_OK;
}

int do_join_index(cc_ctx *ctx);

// Counters as they stand at this point of the stream, copied to pinned memory by the side stream while the
// main stream carries on; snapshot_wait hands them to host_ctr.
int snapshot_begin(cc_ctx *ctx, int k)
{
    CK(cudaEventRecord(ctx->snap_src[k], ctx->stream));
    CK(cudaStreamWaitEvent(ctx->side, ctx->snap_src[k], 0));
    CK(cudaMemcpyAsync(ctx->pin_ctr + k * CT_COUNT, ctx->counters.p, CT_COUNT * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->side));
    CK(cudaEventRecord(ctx->snap_done[k], ctx->side));
    return CC_OK;
}

int snapshot_wait(cc_ctx *ctx, int k)
{
    CK(cudaEventSynchronize(ctx->snap_done[k]));
    memcpy(ctx->host_ctr, ctx->pin_ctr + k * CT_COUNT, sizeof(ctx->host_ctr));
    return CC_OK;
}

int finish_lists(cc_ctx *ctx, bool with_join);

int do_lists(cc_ctx *ctx, bool with_join)
{
    TRY(read_counters(ctx));
    return finish_lists(ctx, with_join);
}

// host_ctr holds the counters: fall back / fail / take the row counts
int finish_lists(cc_ctx *ctx, bool with_join)
{
    if (ctx->pair_join && (ctx->host_ctr[CT_UNMIRRORED] || ctx->host_ctr[CT_FAULTS])) {
        // The pair view does not hold for this batch (mates that do not mirror each other, or pair families that
        // share the reads of one mate only): group it read by read, which reproduces the reference's families
        // for such input - and, for the second kind, reports the pairs through the usual domain error.
        const cc_ctx::GroupMode asked = ctx->group_mode;
        ctx->group_mode = cc_ctx::GROUP_TILE;
        int rc = do_group(ctx);
        if (rc == CC_OK && with_join) rc = do_join_index(ctx);
        ctx->group_mode = asked;
        TRY(rc);
        TRY(read_counters(ctx));
    }
    ctx->n_fam = ctx->host_ctr[CT_NFAM];                 // (before the checks: the family arrays stay fetchable after a fault)
    if (ctx->host_ctr[CT_FAULTS])
        return fail(ctx, CC_ERR_DOMAIN, "a read pair does not found/join its two families together "
                                        "(SSCS_maker.py:275 would not emit this pair name): outside the GPU path's domain");
    if (ctx->host_ctr[CT_ONESIDED])
        return fail(ctx, CC_ERR_DOMAIN, "a duplex partnership is not mutual (undotted barcodes of odd length: duplex_tag "
                                        "applied twice gives another barcode, consensus_helper.py:629-639); what the "
                                        "reference writes then depends on file order: outside the GPU path's domain");
    ctx->n_sscs = ctx->host_ctr[CT_NSSCS];               // family -> row maps were written by the compactions
    if (with_join) {
        ctx->n_dcs = ctx->host_ctr[CT_NDCS];
        ctx->n_corr = ctx->host_ctr[CT_NCORR];
        ctx->n_extra = ctx->host_ctr[CT_NDCSSC];
        ctx->n_dcssc = ctx->params.scorrect ? ctx->n_dcs + ctx->n_extra : 0;
        ctx->dcssc_row_ready = false;                        // the rows of the second pass follow the n_dcs shared ones: added on demand
    }
    ctx->listed = true;
    return CC_OK;
}

// thr[p] = smallest count c with (double)c / (double)p >= cutoff: the comparison of
// SSCS_maker.py:143-144, evaluated once per possible (count, passing) pair with IEEE division
int upload_thresholds(cc_ctx *ctx)
{
    TRY(ensure(ctx, ctx->thr_table, 128));
    const double cutoff = ctx->params.cutoff;
    ctx->thr_host[0] = 255;
    for (int p = 1; p < 128; p++) {
        int t = 255;
        for (int c = 0; c <= p; c++)
            if ((double)c / (double)p >= cutoff) { t = c; break; }
        ctx->thr_host[p] = (uint8_t)t;
    }
    CK(cudaMemcpyAsync(ctx->thr_table.p, ctx->thr_host, 128, cudaMemcpyHostToDevice, ctx->stream));
    // closed form of the table for the sizes vote_small_kernel handles: thr[p] == (p*k + r) >> 8
    ctx->thr_magic = ThrMagic{0, 0, 1};
    for (uint32_t k = 0; k <= 512 && ctx->thr_magic.use_table; k++)
        for (uint32_t r = 0; r < 256; r++) {
            bool ok = true;
            for (uint32_t p2 = 1; p2 <= VOTE_SMALL_MAX && ok; p2++) ok = ((p2 * k + r) >> 8) == ctx->thr_host[p2];
            if (ok) { ctx->thr_magic = ThrMagic{k, r | (r << 16), 0}; break; }
        }
    ctx->thr_valid = true;
    return CC_OK;
}

// one launch of vote_small_kernel over the families of n_lo..n_hi reads
template <int G, int CH, int DEPTH, bool FLUSH>
int launch_vote_small(cc_ctx *ctx, uint32_t rows, uint32_t n_lo, uint32_t n_hi, const uint32_t *row_list)
{
    auto kernel = vote_small_kernel<G, CH, DEPTH, FLUSH>;
    constexpr int ring = vote_ring_bytes(CH, DEPTH);
    static bool configured[64] = {};     // per template instance and per device (the attribute is per device)
    if (ctx->device < 0 || ctx->device >= 64 || !configured[ctx->device]) {
        CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ring));
        if (ctx->device >= 0 && ctx->device < 64) configured[ctx->device] = true;
    }
    kernel<<<blocks_for(rows, VOTE_CHUNK), 256, ring, ctx->stream>>>(
        ctx->payload, ctx->stride, ctx->lpad, ctx->perm, ptr<uint32_t>(ctx->vstart), ptr<uint32_t>(ctx->vsize),
        ptr<uint32_t>(ctx->sscs_list), row_list, rows, n_lo, n_hi, ptr<uint8_t>(ctx->thr_table),
        (1.0 >= ctx->params.cutoff) ? 1 : 0, ctx->thr_magic, ctx->fields, ptr<uint8_t>(ctx->sscs_rows),
        ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag), ptr<int32_t>(ctx->fam_tlen), ptr<uint8_t>(ctx->fam_rg),
        ptr<uint32_t>(ctx->counters) + CT_ERR, ptr<uint32_t>(ctx->big_list) + ctx->redo_off, ptr<uint32_t>(ctx->counters) + CT_NREDO);
    ctx->launches++;
    CK(cudaGetLastError());
    return CC_OK;
}

// Families of 2..15 reads (the bulk) and families of 16..127 reads (nibble counters need flushing)
// are voted by separate launches; the second one is skipped when the batch has no such family.
template <int G>
int launch_vote_pair(cc_ctx *ctx, uint32_t rows)
{
    TRY((launch_vote_small<G, 1, VOTE_DEPTH_CFG, false>(ctx, rows, 2, 15, nullptr)));
    const uint32_t n_mid = ctx->host_ctr[CT_NMID];       // counted by family_build_kernel, listed by vote_segments_kernel
    if (n_mid) TRY((launch_vote_small<G, 1, VOTE_DEPTH_CFG, true>(ctx, n_mid, 16, VOTE_SMALL_MAX, ptr<uint32_t>(ctx->mid_list))));
    return CC_OK;
}

int do_vote(cc_ctx *ctx)
{
    const uint32_t rows = ctx->n_sscs;
    ctx->packed = false;
    TRY(ensure(ctx, ctx->sscs_rows, (size_t)rows * ctx->stride));
    TRY(ensure(ctx, ctx->vstart, (size_t)rows * 4)); TRY(ensure(ctx, ctx->vsize, (size_t)rows * 4));
    // (two lists in one buffer: families of more than 127 reads, and behind them the rows the bulk kernel hands over;
    //  the pair grouping has written the first one already - its buffer is sized for any count, so this never regrows it)
    TRY(ensure(ctx, ctx->big_list, 2 * ((size_t)rows + 2) * 4));
    ctx->redo_off = (size_t)rows + 2;
    TRY(ensure(ctx, ctx->mid_list, ((size_t)ctx->host_ctr[CT_NMID] + 1) * 4));
    if (!ctx->thr_valid) TRY(upload_thresholds(ctx));
    uint32_t *ctr = ptr<uint32_t>(ctx->counters);
    if (rows && !ctx->segments_ready) {
        // (a second vote of one batch comes here too: the first one marked rows in vsize and appended to big_list)
        CK(cudaMemsetAsync(ctr + CT_NBIG, 0, 8, ctx->stream));       // and CT_NREDO behind it
        CK(cudaMemsetAsync(ctr + CT_NMIDLIST, 0, 4, ctx->stream));
        TRY(phase_begin(ctx, CC_PH_VOTE_AUX));
        LAUNCH(vote_segments_kernel, blocks_for(rows, 256), 256, ptr<uint32_t>(ctx->sscs_list),
               ptr<uint32_t>(ctx->fam_start), ptr<uint32_t>(ctx->fam_size), ptr<uint8_t>(ctx->is_sscs), rows,
               ptr<uint32_t>(ctx->vstart),
               ptr<uint32_t>(ctx->vsize), ptr<uint32_t>(ctx->big_list), ctr + CT_NBIG,
               ctx->lpad <= 512 ? VOTE_SMALL_MAX : 0u,    // a lane group covers at most 512 positions
               ptr<uint32_t>(ctx->mid_list), ctr + CT_NMIDLIST);
        TRY(phase_end(ctx, CC_PH_VOTE_AUX));
    }
    ctx->segments_ready = false;
    if (rows) {
        // The families of more than 127 reads (one CTA each, a 10^4-read family is a long chain of row loads) are
        // voted on the side stream while the bulk kernel runs on this one: different families, different rows.
        // What the bulk kernel hands over (rows with a quality byte >= 128) is a list of its own, voted after both.
        CK(cudaEventRecord(ctx->big_src, ctx->stream));
        CK(cudaStreamWaitEvent(ctx->side, ctx->big_src, 0));
        vote_big_kernel<<<148 * 8, 256, 0, ctx->side>>>(
            ctx->fields, ctx->payload, ctx->stride, ctx->lpad, ctx->perm, ptr<uint32_t>(ctx->vstart), ptr<uint32_t>(ctx->vsize),
            ptr<uint32_t>(ctx->sscs_list), ptr<uint32_t>(ctx->big_list), ctr + CT_NBIG, ctx->params.cutoff,
            ptr<uint8_t>(ctx->sscs_rows), ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag), ptr<int32_t>(ctx->fam_tlen),
            ptr<uint8_t>(ctx->fam_rg), ctr + CT_ERR);
        ctx->launches++;
        CK(cudaGetLastError());
        CK(cudaEventRecord(ctx->big_done, ctx->side));
    }
    TRY(phase_begin(ctx, CC_PH_VOTE));
    if (rows && ctx->lpad <= 512) {
        const uint32_t lanes = (ctx->lpad + 15) / 16;
        if (lanes <= 2) TRY(launch_vote_pair<2>(ctx, rows));
        else if (lanes <= 4) TRY(launch_vote_pair<4>(ctx, rows));
        else if (lanes <= 8) TRY(launch_vote_pair<8>(ctx, rows));
        else if (lanes <= 16) TRY(launch_vote_pair<16>(ctx, rows));
        else TRY(launch_vote_pair<32>(ctx, rows));
    }
    TRY(phase_end(ctx, CC_PH_VOTE));
    if (rows) {
        TRY(phase_begin(ctx, CC_PH_VOTE_AUX));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->big_done, 0));
        LAUNCH(vote_big_kernel, 148 * 8, 256, ctx->fields, ctx->payload, ctx->stride, ctx->lpad, ctx->perm,
               ptr<uint32_t>(ctx->vstart), ptr<uint32_t>(ctx->vsize), ptr<uint32_t>(ctx->sscs_list),
               ptr<uint32_t>(ctx->big_list) + ctx->redo_off, ctr + CT_NREDO, ctx->params.cutoff, ptr<uint8_t>(ctx->sscs_rows),
               ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag), ptr<int32_t>(ctx->fam_tlen),
               ptr<uint8_t>(ctx->fam_rg), ctr + CT_ERR);
        TRY(phase_end(ctx, CC_PH_VOTE_AUX));
    }
    ctx->voted = true;
    return CC_OK;
}

RowSources sources(cc_ctx *ctx)
{
    RowSources s;
    s.payload = ctx->payload;
    s.sscs_rows = ptr<uint8_t>(ctx->sscs_rows);
    s.corr_rows = ptr<uint8_t>(ctx->corr_rows);
    s.tmpl = ptr<uint32_t>(ctx->tmpl);
    s.fam_size = ptr<uint32_t>(ctx->fam_size);
    s.sscs_row = ptr<int32_t>(ctx->sscs_row);
    s.corr_row = ptr<int32_t>(ctx->corr_row);
    s.stride = ctx->stride;
    return s;
}

inline uint32_t row_groups(const cc_ctx *ctx) { return ((ctx->lpad + 15) / 16 + DUPLEX_CHUNKS - 1) / DUPLEX_CHUNKS; }   // threads per row

// One pool holds every duplex row: [0, n_dcs) the SSCS-SSCS pairs (DCS_maker's first run; the
// second run reproduces them unchanged), [n_dcs, n_dcs + n_extra) the pairs of the second run
// that involve a corrected singleton.
int do_duplex(cc_ctx *ctx, bool after_sc)
{
    const uint32_t total = ctx->n_dcs + ctx->n_extra;
    const uint32_t rows = after_sc ? ctx->n_extra : ctx->n_dcs;
    const uint32_t first = after_sc ? ctx->n_dcs : 0;
    ctx->packed = false;
    TRY(ensure(ctx, ctx->dcs_rows, (size_t)total * ctx->stride));
    TRY(ensure(ctx, ctx->dcs_mapq, total)); TRY(ensure(ctx, ctx->dcs_rg, total));
    TRY(ensure(ctx, ctx->dcs_flag, (size_t)total * 2)); TRY(ensure(ctx, ctx->dcs_tlen, (size_t)total * 4));
    const int ph = after_sc ? CC_PH_DCS_SC : CC_PH_DCS;
    TRY(phase_begin(ctx, ph));
    if (rows) {
        const uint32_t g = row_groups(ctx);
        LAUNCH(duplex16_kernel, blocks_for((uint64_t)rows * g, 256), 256, sources(ctx),
               ptr<uint32_t>(after_sc ? ctx->dcssc_list : ctx->dcs_list), ptr<int32_t>(ctx->partner), rows, ctx->lpad, g,
               after_sc ? 1 : 0, ptr<uint8_t>(ctx->dcs_rows) + (size_t)first * ctx->stride,
               ptr<uint8_t>(ctx->fam_mapq), ptr<uint16_t>(ctx->fam_flag), ptr<int32_t>(ctx->fam_tlen),
               ptr<uint8_t>(ctx->fam_rg), ptr<uint8_t>(ctx->dcs_mapq) + first, ptr<uint16_t>(ctx->dcs_flag) + first,
               ptr<int32_t>(ctx->dcs_tlen) + first, ptr<uint8_t>(ctx->dcs_rg) + first);
    }
    TRY(phase_end(ctx, ph));
    return CC_OK;
}

int do_rescue(cc_ctx *ctx)
{
    const uint32_t rows = ctx->n_corr;
    ctx->packed = false;
    TRY(ensure(ctx, ctx->corr_rows, (size_t)rows * ctx->stride));
    TRY(phase_begin(ctx, CC_PH_RESCUE));
    if (rows) {
        const uint32_t g = row_groups(ctx);
        LAUNCH(rescue16_kernel, blocks_for((uint64_t)rows * g, 256), 256, sources(ctx), ptr<uint32_t>(ctx->corr_list),
               ptr<int32_t>(ctx->partner), rows, ctx->lpad, g, ptr<uint8_t>(ctx->corr_rows));
    }
    TRY(phase_end(ctx, CC_PH_RESCUE));
    return CC_OK;
}

int check_geometry(cc_ctx *ctx, const void *fields, const void *payload, uint64_t n, uint32_t lpad, uint32_t stride)
{
    if (n && (!fields || !payload)) return fail(ctx, CC_ERR_ARG, "NULL fields/payload");
    if (lpad == 0 || lpad % 32 || stride % 64 || stride < lpad / 2 + lpad)
        return fail(ctx, CC_ERR_ARG, "need lpad % 32 == 0, stride % 64 == 0, stride >= 3*lpad/2");
    if (n >= (1ull << 30)) return fail(ctx, CC_ERR_TOO_LARGE, "a batch holds at most 2^30-1 reads");
    if (n & 1) return fail(ctx, CC_ERR_ARG, "reads come in pairs (2k, 2k+1)");
    return CC_OK;
}


}  // namespace

// ==================================== C ABI ==========================================
extern "C" {

cc_ctx *cc_create(int device, const cc_params *params)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || device < 0 || device >= count) {
        g_create_error = e != cudaSuccess ? std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e)
                                          : "no such CUDA device (this library has no CPU fallback)";
        return nullptr;
    }
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess || prop.major != 10) {
        g_create_error = "device is not an sm_100 (B200) GPU; libccgpu.so carries sm_100a code only";
        return nullptr;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        g_create_error = "cudaSetDevice failed";
        return nullptr;
    }
    cc_ctx *ctx = new cc_ctx();
    if (const char *g = getenv("CCGPU_GROUP"))         // "tile": read by read in tiles, "hash": one table trip per read, "sort": radix sort
        ctx->group_mode = !strcmp(g, "sort") ? cc_ctx::GROUP_SORT : !strcmp(g, "hash") ? cc_ctx::GROUP_HASH
                        : !strcmp(g, "tile") ? cc_ctx::GROUP_TILE : cc_ctx::GROUP_PAIR;
    ctx->device = device;
    if (params) ctx->params = *params;
    if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        g_create_error = "cudaStreamCreate failed";
        delete ctx;
        return nullptr;
    }
    ctx->stream = ctx->own_stream;
    bool ok = cudaStreamCreateWithFlags(&ctx->side, cudaStreamNonBlocking) == cudaSuccess &&
              cudaMallocHost(&ctx->pin_ctr, 3 * CT_COUNT * sizeof(uint32_t)) == cudaSuccess;
    for (int k = 0; k < 2 && ok; k++)
        ok = cudaEventCreateWithFlags(&ctx->snap_src[k], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->snap_done[k], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&ctx->big_src, cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&ctx->big_done, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        g_create_error = "cudaStreamCreate / cudaMallocHost / cudaEventCreate failed";
        cc_destroy(ctx);
        return nullptr;
    }
    return ctx;
}

void cc_destroy(cc_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->side) cudaStreamSynchronize(ctx->side);
    DevBuf *all[] = {&ctx->in_fields, &ctx->in_payload, &ctx->in_dup, &ctx->in_q4, &ctx->in_pairs, &ctx->in_classes, &ctx->planes_sscs, &ctx->planes_corr, &ctx->planes_dcs, &ctx->alphabet, &ctx->key_a, &ctx->key_b, &ctx->val_a, &ctx->val_b,
                     &ctx->head, &ctx->radix_counts, &ctx->scan_status, &ctx->counters, &ctx->stats,
                     &ctx->tmpl, &ctx->c2hs, &ctx->fam_start, &ctx->fam_size, &ctx->is_sscs, &ctx->fam_mapq,
                     &ctx->fam_flag, &ctx->fam_tlen, &ctx->fam_rg, &ctx->partner, &ctx->table, &ctx->f_dcs, &ctx->f_corr,
                     &ctx->f_dcssc, &ctx->sscs_list, &ctx->dcs_list, &ctx->corr_list, &ctx->dcssc_list, &ctx->sscs_row,
                     &ctx->dcs_row, &ctx->corr_row, &ctx->dcssc_row, &ctx->sscs_rows, &ctx->corr_rows, &ctx->dcs_rows,
                     &ctx->dcs_mapq, &ctx->dcs_flag, &ctx->dcs_tlen, &ctx->dcs_rg, &ctx->tmp_a, &ctx->tmp_b, &ctx->tmp_c,
                     &ctx->vstart, &ctx->vsize, &ctx->big_list, &ctx->mid_list, &ctx->thr_table, &ctx->bitmap, &ctx->word_rank,
                     &ctx->first_read, &ctx->cs_in, &ctx->cs_k0, &ctx->cs_k1, &ctx->cs_v0, &ctx->cs_v1, &ctx->gtable, &ctx->cursor, &ctx->word_reads, &ctx->word_start, &ctx->gt_state};
    for (DevBuf *b : all) release(ctx, *b);
    for (auto &t : ctx->phase)
        for (cudaEvent_t ev : t.ev) cudaEventDestroy(ev);
    if (ctx->comm) nccl_api().CommDestroy(ctx->comm);
    for (int k = 0; k < 2; k++) {
        if (ctx->snap_src[k]) cudaEventDestroy(ctx->snap_src[k]);
        if (ctx->snap_done[k]) cudaEventDestroy(ctx->snap_done[k]);
    }
    if (ctx->big_src) cudaEventDestroy(ctx->big_src);
    if (ctx->big_done) cudaEventDestroy(ctx->big_done);
    if (ctx->pin_ctr) cudaFreeHost(ctx->pin_ctr);
    if (ctx->side) cudaStreamDestroy(ctx->side);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

const char *cc_last_error(cc_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int cc_set_params(cc_ctx *ctx, const cc_params *params)
{
    if (!ctx || !params) return CC_ERR_ARG;
    // the row lists, consensus rows and joins of the loaded batch were made with the old parameters:
    // later stages must not hand them out as current (the grouping does not depend on either value)
    if (params->cutoff != ctx->params.cutoff || (params->scorrect != 0) != (ctx->params.scorrect != 0))
        ctx->voted = ctx->joined = ctx->rescued = ctx->packed = false;
    ctx->params = *params;
    ctx->thr_valid = false;
    return CC_OK;
}

int cc_set_stream(cc_ctx *ctx, void *stream)
{
    if (!ctx) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->stream = stream ? reinterpret_cast<cudaStream_t>(stream) : ctx->own_stream;
    return CC_OK;
}

int cc_load_reads(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload, uint64_t n, uint32_t lpad,
                  uint32_t stride, const uint32_t *dup_bc, uint32_t n_bc)
{
    if (!ctx) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    TRY(check_geometry(ctx, fields, payload, n, lpad, stride));
    TRY(ensure(ctx, ctx->in_fields, n * sizeof(cc_fields)));
    TRY(ensure(ctx, ctx->in_payload, n * (size_t)stride));
    TRY(ensure(ctx, ctx->in_dup, (size_t)n_bc * 4));
    TRY(phase_begin(ctx, CC_PH_H2D));
    if (n) {
        CK(cudaMemcpyAsync(ctx->in_fields.p, fields, n * sizeof(cc_fields), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->in_payload.p, payload, n * (size_t)stride, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (n_bc) CK(cudaMemcpyAsync(ctx->in_dup.p, dup_bc, (size_t)n_bc * 4, cudaMemcpyHostToDevice, ctx->stream));
    TRY(phase_end(ctx, CC_PH_H2D));
    ctx->fields = ptr<uint4>(ctx->in_fields);
    ctx->payload = ptr<uint8_t>(ctx->in_payload);
    ctx->dup_bc = ptr<uint32_t>(ctx->in_dup);
    ctx->n = n; ctx->lpad = lpad; ctx->stride = stride; ctx->n_bc = n_bc;
    ctx->have_reads = true;
    ctx->by_classes = false;
    ctx->grouped = ctx->listed = ctx->voted = ctx->joined = ctx->rescued = false;
    return CC_OK;
}

int cc_load_reads_q4(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload4, uint64_t n, uint32_t lpad,
                     uint32_t stride4, const uint8_t qual_table[16], const uint32_t *dup_bc, uint32_t n_bc)
{
    if (!ctx || !qual_table) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (lpad == 0 || lpad % 32 || stride4 % 16 || stride4 < lpad)
        return fail(ctx, CC_ERR_ARG, "need lpad % 32 == 0, stride4 % 16 == 0, stride4 >= lpad");
    const uint32_t stride = (lpad / 2 + lpad + 63) / 64 * 64;           // the standard row the kernels read
    TRY(check_geometry(ctx, fields, payload4, n, lpad, stride));
    if (qual_table[0] != 0) return fail(ctx, CC_ERR_ARG, "qual_table[0] must be 0 (code 0 pads the rows)");
    TRY(ensure(ctx, ctx->in_fields, n * sizeof(cc_fields)));
    TRY(ensure(ctx, ctx->in_q4, n * (size_t)stride4));
    TRY(ensure(ctx, ctx->in_payload, n * (size_t)stride));
    TRY(ensure(ctx, ctx->in_dup, (size_t)n_bc * 4));
    TRY(phase_begin(ctx, CC_PH_H2D));
    if (n) {
        CK(cudaMemcpyAsync(ctx->in_fields.p, fields, n * sizeof(cc_fields), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->in_q4.p, payload4, n * (size_t)stride4, cudaMemcpyHostToDevice, ctx->stream));
        QualTable table;
        memcpy(table.q, qual_table, 16);
        LAUNCH(expand_q4_kernel, blocks_for(n * (lpad / 16), 256), 256, ptr<uint8_t>(ctx->in_q4), stride4,
               ptr<uint8_t>(ctx->in_payload), stride, lpad, n, table);
    }
    if (n_bc) CK(cudaMemcpyAsync(ctx->in_dup.p, dup_bc, (size_t)n_bc * 4, cudaMemcpyHostToDevice, ctx->stream));
    TRY(phase_end(ctx, CC_PH_H2D));
    ctx->fields = ptr<uint4>(ctx->in_fields);
    ctx->payload = ptr<uint8_t>(ctx->in_payload);
    ctx->dup_bc = ptr<uint32_t>(ctx->in_dup);
    ctx->n = n; ctx->lpad = lpad; ctx->stride = stride; ctx->n_bc = n_bc;
    ctx->have_reads = true;
    ctx->by_classes = false;
    ctx->grouped = ctx->listed = ctx->voted = ctx->joined = ctx->rescued = false;
    return CC_OK;
}

// cc_load_reads_compact / cc_load_reads_classes: pair records (40-byte ones, or 16-byte ones + a class table) and
// rows in bit planes, both expanded on the device into the layout every kernel reads
static int load_planes(cc_ctx *ctx, const char *who, const void *pairs, size_t pair_bytes, const cc_pair_class *classes,
                       uint32_t n_classes, const uint8_t *planes, uint64_t n, uint32_t lpad, uint32_t stride_c, uint32_t qbits,
                       const uint8_t qual_table[16], const uint8_t base_table[16], const uint32_t *dup_bc, uint32_t n_bc)
{
    if (!ctx || !qual_table || !base_table) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (qbits < 1 || qbits > 4) return fail(ctx, CC_ERR_ARG, (std::string(who) + ": qbits must be 1..4").c_str());
    if (lpad == 0 || lpad % 32 || stride_c % 16 || stride_c < lpad / 4 + lpad * qbits / 8)
        return fail(ctx, CC_ERR_ARG, (std::string(who) + ": need lpad % 32 == 0, stride_c % 16 == 0, stride_c >= lpad/4 + lpad*qbits/8").c_str());
    const bool by_classes = pair_bytes == sizeof(cc_pair16);
    if (by_classes && n && (!classes || n_classes == 0 || n_classes > 65536))
        return fail(ctx, CC_ERR_ARG, (std::string(who) + ": need 1..65536 classes").c_str());
    const uint32_t stride = (lpad / 2 + lpad + 63) / 64 * 64;           // the standard row the kernels read
    TRY(check_geometry(ctx, pairs, planes, n, lpad, stride));
    TRY(ensure(ctx, ctx->in_fields, n * sizeof(cc_fields)));
    TRY(ensure(ctx, ctx->in_pairs, (n / 2) * pair_bytes));
    if (by_classes) TRY(ensure(ctx, ctx->in_classes, (size_t)n_classes * sizeof(cc_pair_class) + 64));
    TRY(ensure(ctx, ctx->in_q4, n * (size_t)stride_c));
    TRY(ensure(ctx, ctx->in_payload, n * (size_t)stride));
    TRY(ensure(ctx, ctx->in_dup, (size_t)n_bc * 4));
    TRY(phase_begin(ctx, CC_PH_H2D));
    ctx->by_classes = false;
    if (n) {
        CK(cudaMemcpyAsync(ctx->in_pairs.p, pairs, (n / 2) * pair_bytes, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->in_q4.p, planes, n * (size_t)stride_c, cudaMemcpyHostToDevice, ctx->stream));
        if (by_classes) {
            uint32_t *bad = reinterpret_cast<uint32_t *>(ptr<uint8_t>(ctx->in_classes) + ctx->in_classes.cap - 4);
            CK(cudaMemcpyAsync(ctx->in_classes.p, classes, (size_t)n_classes * sizeof(cc_pair_class), cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemsetAsync(bad, 0, 4, ctx->stream));
            LAUNCH(expand_pairs16_kernel, blocks_for(n / 2, 256), 256, ptr<uint4>(ctx->in_pairs), ptr<uint4>(ctx->in_classes), n_classes,
                   n / 2, ptr<uint4>(ctx->in_fields), bad);
            ctx->by_classes = true;
        } else {
            LAUNCH(expand_pairs_kernel, blocks_for(n / 2, 256), 256, ptr<uint2>(ctx->in_pairs), n / 2, ptr<uint4>(ctx->in_fields));
        }
        CodeTable table;
        memcpy(table.qual, qual_table, 16);
        memcpy(table.base, base_table, 16);
        const unsigned grid = blocks_for(n * (lpad / 16), 256);
        uint8_t *in = ptr<uint8_t>(ctx->in_q4), *out = ptr<uint8_t>(ctx->in_payload);
        switch (qbits) {
        case 1: LAUNCH(expand_planes_kernel<1>, grid, 256, in, stride_c, out, stride, lpad, n, table); break;
        case 2: LAUNCH(expand_planes_kernel<2>, grid, 256, in, stride_c, out, stride, lpad, n, table); break;
        case 3: LAUNCH(expand_planes_kernel<3>, grid, 256, in, stride_c, out, stride, lpad, n, table); break;
        default: LAUNCH(expand_planes_kernel<4>, grid, 256, in, stride_c, out, stride, lpad, n, table); break;
        }
    }
    if (n_bc) CK(cudaMemcpyAsync(ctx->in_dup.p, dup_bc, (size_t)n_bc * 4, cudaMemcpyHostToDevice, ctx->stream));
    TRY(phase_end(ctx, CC_PH_H2D));
    ctx->fields = ptr<uint4>(ctx->in_fields);
    ctx->payload = ptr<uint8_t>(ctx->in_payload);
    ctx->dup_bc = ptr<uint32_t>(ctx->in_dup);
    ctx->n = n; ctx->lpad = lpad; ctx->stride = stride; ctx->n_bc = n_bc;
    ctx->have_reads = true;
    ctx->grouped = ctx->listed = ctx->voted = ctx->joined = ctx->rescued = false;
    return CC_OK;
}

int cc_load_reads_compact(cc_ctx *ctx, const cc_pair *pairs, const uint8_t *planes, uint64_t n, uint32_t lpad,
                          uint32_t stride_c, uint32_t qbits, const uint8_t qual_table[16], const uint8_t base_table[16],
                          const uint32_t *dup_bc, uint32_t n_bc)
{
    return load_planes(ctx, "cc_load_reads_compact", pairs, sizeof(cc_pair), nullptr, 0, planes, n, lpad, stride_c, qbits,
                       qual_table, base_table, dup_bc, n_bc);
}

int cc_load_reads_classes(cc_ctx *ctx, const cc_pair16 *pairs, const cc_pair_class *classes, uint32_t n_classes,
                          const uint8_t *planes, uint64_t n, uint32_t lpad, uint32_t stride_c, uint32_t qbits,
                          const uint8_t qual_table[16], const uint8_t base_table[16], const uint32_t *dup_bc, uint32_t n_bc)
{
    return load_planes(ctx, "cc_load_reads_classes", pairs, sizeof(cc_pair16), classes, n_classes, planes, n, lpad, stride_c,
                       qbits, qual_table, base_table, dup_bc, n_bc);
}

int cc_load_reads_device(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload, uint64_t n, uint32_t lpad,
                         uint32_t stride, const uint32_t *dup_bc, uint32_t n_bc)
{
    if (!ctx) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    TRY(check_geometry(ctx, fields, payload, n, lpad, stride));
    if (((uintptr_t)fields & 31) || ((uintptr_t)payload & 15))
        return fail(ctx, CC_ERR_ARG, "device buffers: fields must be 32-byte aligned, payload 16-byte aligned");
    ctx->fields = reinterpret_cast<const uint4 *>(fields);
    ctx->payload = payload;
    ctx->dup_bc = dup_bc;
    ctx->n = n; ctx->lpad = lpad; ctx->stride = stride; ctx->n_bc = n_bc;
    ctx->have_reads = true;
    ctx->by_classes = false;
    ctx->grouped = ctx->listed = ctx->voted = ctx->joined = ctx->rescued = false;
    return CC_OK;
}

int cc_group(cc_ctx *ctx)
{
    if (!ctx) return CC_ERR_ARG;
    if (!ctx->have_reads) return fail(ctx, CC_ERR_ARG, "cc_group before cc_load_reads");
    CK(cudaSetDevice(ctx->device));
    TRY(do_group(ctx));
    return do_lists(ctx, false);
}

int cc_sscs_vote(cc_ctx *ctx)
{
    if (!ctx) return CC_ERR_ARG;
    if (!ctx->grouped || !ctx->listed) return fail(ctx, CC_ERR_ARG, "cc_sscs_vote before cc_group");
    CK(cudaSetDevice(ctx->device));
    TRY(do_vote(ctx));
    return read_counters(ctx);
}

int cc_dcs_join(cc_ctx *ctx)
{
    if (!ctx) return CC_ERR_ARG;
    if (!ctx->voted) return fail(ctx, CC_ERR_ARG, "cc_dcs_join before cc_sscs_vote");
    CK(cudaSetDevice(ctx->device));
    TRY(do_join_index(ctx));
    TRY(do_lists(ctx, true));
    TRY(do_duplex(ctx, false));
    ctx->joined = true;
    return CC_OK;
}

int cc_sc_rescue(cc_ctx *ctx)
{
    if (!ctx) return CC_ERR_ARG;
    if (!ctx->joined) return fail(ctx, CC_ERR_ARG, "cc_sc_rescue before cc_dcs_join");
    if (!ctx->params.scorrect) return fail(ctx, CC_ERR_ARG, "cc_sc_rescue with params.scorrect == 0");
    CK(cudaSetDevice(ctx->device));
    TRY(do_rescue(ctx));
    TRY(do_duplex(ctx, true));
    ctx->rescued = true;
    return CC_OK;
}

int cc_run(cc_ctx *ctx)
{
    if (!ctx) return CC_ERR_ARG;
    if (!ctx->have_reads) return fail(ctx, CC_ERR_ARG, "cc_run before cc_load_reads");
    CK(cudaSetDevice(ctx->device));
    TRY(do_group(ctx));
    if (ctx->pair_join && ctx->n) {
        // Everything the vote's launch needs (consensus rows, families of 16+ reads, the grouping's verdict on the
        // input) is final before the join: the host reads it while the join and the classification run, queues the
        // vote behind them, and reads the duplex counts while the vote runs.
        TRY(snapshot_begin(ctx, 0));
        TRY(do_join_index(ctx));
        TRY(snapshot_begin(ctx, 1));
        TRY(snapshot_wait(ctx, 0));
        if (ctx->host_ctr[CT_UNMIRRORED] || ctx->host_ctr[CT_FAULTS]) {
            TRY(do_lists(ctx, true));                        // regroups read by read, or reports the pairs
            TRY(do_vote(ctx));
        } else {
            ctx->n_fam = ctx->host_ctr[CT_NFAM];
            ctx->n_sscs = ctx->host_ctr[CT_NSSCS];
            TRY(do_vote(ctx));
            TRY(snapshot_wait(ctx, 1));
            const int rc = finish_lists(ctx, true);
            if (rc != CC_OK) { ctx->voted = false; return rc; }    // (a partnership that is not mutual: as if the vote had not run)
        }
    } else {
        TRY(do_join_index(ctx));
        TRY(do_lists(ctx, true));
        TRY(do_vote(ctx));
    }
    if (ctx->params.scorrect) TRY(do_rescue(ctx));
    TRY(do_duplex(ctx, false));
    ctx->joined = true;
    if (ctx->params.scorrect) {
        TRY(do_duplex(ctx, true));
        ctx->rescued = true;
    }
    return read_counters(ctx);     // synchronises; surfaces a base error raised by the vote kernel
}

int cc_run_host(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload, uint64_t n, uint32_t lpad,
                uint32_t stride, const uint32_t *dup_bc, uint32_t n_bc)
{
    TRY(cc_load_reads(ctx, fields, payload, n, lpad, stride, dup_bc, n_bc));
    return cc_run(ctx);
}

int cc_run_host_q4(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload4, uint64_t n, uint32_t lpad,
                   uint32_t stride4, const uint8_t qual_table[16], const uint32_t *dup_bc, uint32_t n_bc)
{
    TRY(cc_load_reads_q4(ctx, fields, payload4, n, lpad, stride4, qual_table, dup_bc, n_bc));
    return cc_run(ctx);
}

int cc_run_host_compact(cc_ctx *ctx, const cc_pair *pairs, const uint8_t *planes, uint64_t n, uint32_t lpad,
                        uint32_t stride_c, uint32_t qbits, const uint8_t qual_table[16], const uint8_t base_table[16],
                        const uint32_t *dup_bc, uint32_t n_bc)
{
    TRY(cc_load_reads_compact(ctx, pairs, planes, n, lpad, stride_c, qbits, qual_table, base_table, dup_bc, n_bc));
    return cc_run(ctx);
}

int cc_run_host_classes(cc_ctx *ctx, const cc_pair16 *pairs, const cc_pair_class *classes, uint32_t n_classes,
                        const uint8_t *planes, uint64_t n, uint32_t lpad, uint32_t stride_c, uint32_t qbits,
                        const uint8_t qual_table[16], const uint8_t base_table[16], const uint32_t *dup_bc, uint32_t n_bc)
{
    TRY(cc_load_reads_classes(ctx, pairs, classes, n_classes, planes, n, lpad, stride_c, qbits, qual_table, base_table, dup_bc, n_bc));
    return cc_run(ctx);
}

int cc_counts(cc_ctx *ctx, uint64_t counts[8])
{
    if (!ctx || !counts) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    TRY(read_counters(ctx));
    memset(counts, 0, 8 * sizeof(uint64_t));
    counts[CC_N_READS] = ctx->n; counts[CC_N_FAM] = ctx->n_fam; counts[CC_N_SSCS] = ctx->n_sscs;
    counts[CC_N_CORR] = ctx->n_corr; counts[CC_N_DCS] = ctx->n_dcs; counts[CC_N_DCSSC] = ctx->n_dcssc;
    return CC_OK;
}

static const void *array_ptr(cc_ctx *ctx, int what, uint64_t *bytes)
{
    const uint64_t nf = ctx->n_fam, S = ctx->stride;
    *bytes = 0;
    if (!ctx->grouped) return nullptr;
    switch (what) {
    case CC_READ_FAM: *bytes = ctx->n * 4; return ctx->spare_vals;
    case CC_FAM_SIZE: *bytes = nf * 4; return ctx->fam_size.p;
    case CC_FAM_TMPL: *bytes = nf * 4; return ctx->tmpl.p;
    case CC_FAM_MAPQ: *bytes = nf; return ctx->fam_mapq.p;
    case CC_FAM_FLAG: *bytes = nf * 2; return ctx->fam_flag.p;
    case CC_FAM_TLEN: *bytes = nf * 4; return ctx->fam_tlen.p;
    case CC_FAM_RG: *bytes = nf; return ctx->fam_rg.p;
    case CC_PARTNER: if (!ctx->joined) return nullptr; *bytes = nf * 4; return ctx->partner.p;
    case CC_SSCS_ROW: if (!ctx->listed) return nullptr; *bytes = nf * 4; return ctx->sscs_row.p;
    case CC_CORR_ROW: if (!ctx->joined) return nullptr; *bytes = nf * 4; return ctx->corr_row.p;
    case CC_DCS_ROW: if (!ctx->joined) return nullptr; *bytes = nf * 4; return ctx->dcs_row.p;
    case CC_DCSSC_ROW: if (!ctx->joined) return nullptr; *bytes = nf * 4; return ctx->dcssc_row.p;
    case CC_SSCS_ROWS: if (!ctx->voted) return nullptr; *bytes = ctx->n_sscs * S; return ctx->sscs_rows.p;
    case CC_CORR_ROWS: if (!ctx->rescued) return nullptr; *bytes = ctx->n_corr * S; return ctx->corr_rows.p;
    case CC_DCS_ROWS: if (!ctx->joined) return nullptr; *bytes = ctx->n_dcs * S; return ctx->dcs_rows.p;
    case CC_DCSSC_ROWS: if (!ctx->rescued) return nullptr; *bytes = ctx->n_dcssc * S; return ctx->dcs_rows.p;
    case CC_DCS_MAPQ: if (!ctx->joined) return nullptr; *bytes = ctx->n_dcs; return ctx->dcs_mapq.p;
    case CC_DCS_FLAG: if (!ctx->joined) return nullptr; *bytes = ctx->n_dcs * 2ull; return ctx->dcs_flag.p;
    case CC_DCS_TLEN: if (!ctx->joined) return nullptr; *bytes = ctx->n_dcs * 4ull; return ctx->dcs_tlen.p;
    case CC_DCS_RG: if (!ctx->joined) return nullptr; *bytes = ctx->n_dcs; return ctx->dcs_rg.p;
    case CC_DCSSC_MAPQ: if (!ctx->rescued) return nullptr; *bytes = ctx->n_dcssc; return ctx->dcs_mapq.p;
    case CC_DCSSC_FLAG: if (!ctx->rescued) return nullptr; *bytes = ctx->n_dcssc * 2ull; return ctx->dcs_flag.p;
    case CC_DCSSC_TLEN: if (!ctx->rescued) return nullptr; *bytes = ctx->n_dcssc * 4ull; return ctx->dcs_tlen.p;
    case CC_DCSSC_RG: if (!ctx->rescued) return nullptr; *bytes = ctx->n_dcssc; return ctx->dcs_rg.p;
    case CC_SSCS_PLANES: if (!ctx->packed) return nullptr; *bytes = (uint64_t)ctx->n_sscs * ctx->pack_stride; return ctx->planes_sscs.p;
    case CC_CORR_PLANES: if (!ctx->packed || !ctx->rescued) return nullptr; *bytes = (uint64_t)ctx->n_corr * ctx->pack_stride; return ctx->planes_corr.p;
    case CC_DCS_PLANES: if (!ctx->packed || !ctx->joined) return nullptr; *bytes = (uint64_t)ctx->n_dcs * ctx->pack_stride; return ctx->planes_dcs.p;
    case CC_DCSSC_PLANES: if (!ctx->packed || !ctx->rescued) return nullptr; *bytes = (uint64_t)ctx->n_dcssc * ctx->pack_stride; return ctx->planes_dcs.p;
    case CC_STATS: *bytes = 16 * 8; return ctx->stats.p;
    case CC_FAM_HIST: *bytes = CC_HIST_BINS * 8ull; return ctx->hist.p;
    default: return nullptr;
    }
}

uint64_t cc_array_bytes(cc_ctx *ctx, int what)
{
    if (!ctx) return 0;
    uint64_t bytes = 0;
    array_ptr(ctx, what, &bytes);
    return bytes;
}

static int materialise(cc_ctx *ctx, int what)
{
    if (what == CC_DCSSC_ROW && !ctx->dcssc_row_ready && ctx->n_fam) {
        LAUNCH(dcssc_row_finish_kernel, blocks_for(ctx->n_fam, 256), 256, ptr<int32_t>(ctx->dcssc_row), ctx->n_fam, ctx->n_dcs);
        ctx->dcssc_row_ready = true;
    }
    if (what == CC_READ_FAM && !ctx->read_fam_ready && ctx->n_fam) {
        LAUNCH(read_fam_kernel, blocks_for((uint64_t)ctx->n_fam * 32, 256), 256, ptr<uint32_t>(ctx->fam_start),
               ptr<uint32_t>(ctx->fam_size), ctx->perm, ctx->n_fam, ctx->spare_vals);
        ctx->read_fam_ready = true;
    }
    return CC_OK;
}

const void *cc_device_ptr(cc_ctx *ctx, int what)
{
    if (!ctx) return nullptr;
    uint64_t bytes = 0;
    const void *p = array_ptr(ctx, what, &bytes);
    if (!p || bytes == 0) return nullptr;
    if (cudaSetDevice(ctx->device) != cudaSuccess || materialise(ctx, what) != CC_OK) return nullptr;
    return p;
}

int cc_pack_rows(cc_ctx *ctx, uint32_t *qbits, uint32_t *stride_c, uint8_t qual_table[16], uint8_t base_table[16])
{
    if (!ctx || !qbits || !stride_c || !qual_table || !base_table) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (!ctx->voted) return fail(ctx, CC_ERR_ARG, "cc_pack_rows before cc_sscs_vote");
    *qbits = 0; *stride_c = 0;
    memset(qual_table, 0, 16); memset(base_table, 0xFF, 16);
    struct Part { DevBuf *src; DevBuf *dst; uint64_t rows; };
    const Part parts[3] = {{&ctx->sscs_rows, &ctx->planes_sscs, ctx->n_sscs},
                           {&ctx->corr_rows, &ctx->planes_corr, ctx->rescued ? ctx->n_corr : 0},
                           {&ctx->dcs_rows, &ctx->planes_dcs, ctx->rescued ? ctx->n_dcssc : ctx->joined ? ctx->n_dcs : 0}};
    const uint32_t lpad = ctx->lpad, lanes = lpad / 16;
    if (!ctx->packed) {
        // which (quality, literal base) combinations occur: 65 536 presence bits (+ the pack kernels' "does not fit" word)
        TRY(ensure(ctx, ctx->alphabet, (ALPHA_WORDS + 1) * sizeof(uint32_t)));
        uint32_t *seen = ptr<uint32_t>(ctx->alphabet);
        CK(cudaMemsetAsync(seen, 0, (ALPHA_WORDS + 1) * sizeof(uint32_t), ctx->stream));
        for (const Part &p : parts)
            if (p.rows)
                LAUNCH(row_alphabet_kernel, (unsigned)std::min<uint64_t>(blocks_for(p.rows * lanes, 256), 148 * 8), 256,
                       ptr<uint8_t>(*p.src), ctx->stride, lpad, p.rows, seen);
        std::vector<uint32_t> host(ALPHA_WORDS);
        CK(cudaMemcpyAsync(host.data(), seen, ALPHA_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        RowKeys keys;
        memset(&keys, 0, sizeof(keys));
        uint32_t found = 0;
        for (uint32_t w = 0; w < ALPHA_WORDS; w++)
            for (uint32_t bits = host[w]; bits; bits &= bits - 1) {
                if (found < 16) keys.key[found] = (uint16_t)(w * 32 + (uint32_t)__builtin_ctz(bits));
                found++;
            }
        if (found > 16) return CC_OK;                         // *qbits == 0: the rows do not fit, fetch them as they are
        keys.count = found;
        uint32_t qb = 1;
        while ((1u << qb) < found) qb++;
        ctx->pack_qbits = qb;
        ctx->pack_stride = (lpad / 4 + lpad * qb / 8 + 15) / 16 * 16;
        memset(ctx->pack_qual, 0, 16); memset(ctx->pack_base, 0xFF, 16);
        for (uint32_t k = 0; k < found; k++) { ctx->pack_qual[k] = (uint8_t)(keys.key[k] >> 8); ctx->pack_base[k] = (uint8_t)(keys.key[k] & 0xFF); }
        for (const Part &p : parts) {
            if (!p.rows) continue;
            TRY(ensure(ctx, *p.dst, p.rows * ctx->pack_stride));
            const unsigned grid = blocks_for(p.rows * lanes, 256);
            const uint8_t *in = ptr<uint8_t>(*p.src);
            uint8_t *out = ptr<uint8_t>(*p.dst);
            switch (qb) {
            case 1: LAUNCH(pack_rows_kernel<1>, grid, 256, in, ctx->stride, lpad, p.rows, out, ctx->pack_stride, keys, seen + ALPHA_WORDS); break;
            case 2: LAUNCH(pack_rows_kernel<2>, grid, 256, in, ctx->stride, lpad, p.rows, out, ctx->pack_stride, keys, seen + ALPHA_WORDS); break;
            case 3: LAUNCH(pack_rows_kernel<3>, grid, 256, in, ctx->stride, lpad, p.rows, out, ctx->pack_stride, keys, seen + ALPHA_WORDS); break;
            default: LAUNCH(pack_rows_kernel<4>, grid, 256, in, ctx->stride, lpad, p.rows, out, ctx->pack_stride, keys, seen + ALPHA_WORDS); break;
            }
        }
        ctx->packed = true;
    }
    *qbits = ctx->pack_qbits; *stride_c = ctx->pack_stride;
    memcpy(qual_table, ctx->pack_qual, 16); memcpy(base_table, ctx->pack_base, 16);
    return CC_OK;
}

int cc_fetch(cc_ctx *ctx, int what, void *host_dst, uint64_t bytes)
{
    if (!ctx) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    uint64_t have = 0;
    const void *p = array_ptr(ctx, what, &have);
    if (!p && have == 0 && !(ctx->grouped && bytes == 0)) {
        if (bytes == 0) return CC_OK;
        return fail(ctx, CC_ERR_ARG, "cc_fetch: array not produced yet");
    }
    if (bytes != have) return fail(ctx, CC_ERR_ARG, "cc_fetch: size mismatch (use cc_array_bytes)");
    if (bytes == 0) return CC_OK;
    if (!host_dst) return fail(ctx, CC_ERR_ARG, "cc_fetch: NULL destination");
    TRY(materialise(ctx, what));
    CK(cudaMemcpyAsync(host_dst, p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return CC_OK;
}

int cc_fetch_many(cc_ctx *ctx, uint32_t count, const int *what, void *const *host_dst, const uint64_t *bytes)
{
    if (!ctx || (count && (!what || !host_dst || !bytes))) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    for (uint32_t k = 0; k < count; k++) {
        uint64_t have = 0;
        const void *p = array_ptr(ctx, what[k], &have);
        if (bytes[k] != have) return fail(ctx, CC_ERR_ARG, "cc_fetch_many: size mismatch (use cc_array_bytes)");
        if (have == 0) continue;
        if (!p) return fail(ctx, CC_ERR_ARG, "cc_fetch_many: array not produced yet");
        if (!host_dst[k]) return fail(ctx, CC_ERR_ARG, "cc_fetch_many: NULL destination");
        TRY(materialise(ctx, what[k]));
        CK(cudaMemcpyAsync(host_dst[k], p, have, cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));                  // one wait for all the copies
    return CC_OK;
}

int cc_consensus_maker(cc_ctx *ctx, const uint8_t *rows, uint32_t n, uint32_t lpad, uint32_t stride, double cutoff,
                       uint8_t *out_row)
{
    if (!ctx || !rows || !out_row || n == 0) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (lpad == 0 || lpad % 32 || stride % 64 || stride < lpad / 2 + lpad)
        return fail(ctx, CC_ERR_ARG, "need lpad % 32 == 0, stride % 64 == 0, stride >= 3*lpad/2");
    // one family: perm = identity, one segment [0, n), mode fields are discarded
    const size_t meta = 4096;
    TRY(ensure(ctx, ctx->tmp_a, (size_t)n * stride));
    TRY(ensure(ctx, ctx->tmp_b, (size_t)n * 4 + meta + (size_t)n * 32));
    TRY(ensure(ctx, ctx->tmp_c, (size_t)stride + 64));
    uint8_t *base = ptr<uint8_t>(ctx->tmp_b);
    uint32_t *perm = reinterpret_cast<uint32_t *>(base + meta + (size_t)n * 32);
    uint32_t *small = reinterpret_cast<uint32_t *>(base);         // start, size, list, err, out meta
    uint4 *fields = reinterpret_cast<uint4 *>(base + meta);
    CK(cudaMemcpyAsync(ctx->tmp_a.p, rows, (size_t)n * stride, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemsetAsync(base, 0, meta + (size_t)n * 32, ctx->stream));
    uint32_t hdr[2] = {0, n};
    CK(cudaMemcpyAsync(small, hdr, sizeof(hdr), cudaMemcpyHostToDevice, ctx->stream));
    LAUNCH(iota_kernel, blocks_for(n, 256), 256, perm, n);
    LAUNCH(vote_kernel, 1, 32, fields, ptr<uint8_t>(ctx->tmp_a), stride, lpad, perm, small + 0, small + 1, small + 2, 1u,
           cutoff, ptr<uint8_t>(ctx->tmp_c), reinterpret_cast<uint8_t *>(small + 16), reinterpret_cast<uint16_t *>(small + 20),
           reinterpret_cast<int32_t *>(small + 24), reinterpret_cast<uint8_t *>(small + 28), small + 8);
    uint32_t err = 0;
    CK(cudaMemcpyAsync(out_row, ctx->tmp_c.p, stride, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(&err, small + 8, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (err)
        return fail(ctx, CC_ERR_BASE, "a base other than A/C/G/T carries Q>=30 "
                                      "(the reference raises IndexError/ValueError, SSCS_maker.py:119-122)");
    return CC_OK;
}

int cc_duplex_consensus(cc_ctx *ctx, const uint8_t *rows1, const uint8_t *rows2, uint64_t n_pairs, uint32_t lpad,
                        uint32_t stride, int sc, uint8_t *out_rows)
{
    if (!ctx || !rows1 || !rows2 || !out_rows) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (lpad == 0 || lpad % 32 || stride % 64 || stride < lpad / 2 + lpad)
        return fail(ctx, CC_ERR_ARG, "need lpad % 32 == 0, stride % 64 == 0, stride >= 3*lpad/2");
    if (n_pairs == 0) return CC_OK;
    const size_t bytes = (size_t)n_pairs * stride;
    TRY(ensure(ctx, ctx->tmp_a, bytes)); TRY(ensure(ctx, ctx->tmp_b, bytes)); TRY(ensure(ctx, ctx->tmp_c, bytes));
    CK(cudaMemcpyAsync(ctx->tmp_a.p, rows1, bytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->tmp_b.p, rows2, bytes, cudaMemcpyHostToDevice, ctx->stream));
    const uint32_t g = (lpad + 15) / 16;
    auto kernel = sc ? agree_pairs16_kernel<true> : agree_pairs16_kernel<false>;
    LAUNCH(kernel, blocks_for(n_pairs * g, 256), 256, ptr<uint8_t>(ctx->tmp_a),
           ptr<uint8_t>(ctx->tmp_b), n_pairs, lpad, stride, g, ptr<uint8_t>(ctx->tmp_c));
    CK(cudaMemcpyAsync(out_rows, ctx->tmp_c.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return CC_OK;
}

int cc_join_records(cc_ctx *ctx, const cc_fields *records, uint64_t n, const uint32_t *dup_bc, uint32_t n_bc, int32_t *partner)
{
    if (!ctx || (n && (!records || !partner)) || (n_bc && !dup_bc)) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (n == 0) return CC_OK;
    if (n >= (1ull << 31)) return fail(ctx, CC_ERR_TOO_LARGE, "cc_join_records: at most 2^31 - 1 records");
    uint64_t slots = 64;
    while (slots < 2 * n) slots <<= 1;
    // staged: records, then the barcode table; the hash table; the answer
    TRY(ensure(ctx, ctx->tmp_a, n * sizeof(cc_fields) + (size_t)n_bc * 4 + 256));
    TRY(ensure(ctx, ctx->tmp_b, slots * 8));
    TRY(ensure(ctx, ctx->tmp_c, n * 4));
    uint32_t *d_dup = reinterpret_cast<uint32_t *>(ptr<uint8_t>(ctx->tmp_a) + n * sizeof(cc_fields));
    CK(cudaMemcpyAsync(ctx->tmp_a.p, records, n * sizeof(cc_fields), cudaMemcpyHostToDevice, ctx->stream));
    if (n_bc) CK(cudaMemcpyAsync(d_dup, dup_bc, (size_t)n_bc * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemsetAsync(ctx->tmp_b.p, 0, slots * 8, ctx->stream));
    LAUNCH(record_build_kernel, blocks_for(n, 256), 256, ptr<uint4>(ctx->tmp_a), (uint32_t)n, ptr<unsigned long long>(ctx->tmp_b), slots - 1);
    LAUNCH(record_probe_kernel, blocks_for(n, 256), 256, ptr<uint4>(ctx->tmp_a), (uint32_t)n, d_dup, n_bc,
           ptr<unsigned long long>(ctx->tmp_b), slots - 1, ptr<int32_t>(ctx->tmp_c));
    CK(cudaMemcpyAsync(partner, ctx->tmp_c.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return CC_OK;
}

int cc_sort_coordinates(cc_ctx *ctx, const int32_t *tid, const int32_t *pos, const uint16_t *flag, uint64_t n,
                        uint32_t *order)
{
    if (!ctx || (n && (!tid || !pos || !flag || !order))) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (n == 0) return CC_OK;
    if (n >= (1ull << 31)) return fail(ctx, CC_ERR_TOO_LARGE, "cc_sort_coordinates: at most 2^31 - 1 records");
    int32_t max_pos = -1, max_tid = -1;
    bool no_contig = false;
    for (uint64_t i = 0; i < n; i++) {
        if (pos[i] < -1 || pos[i] == INT32_MAX) return fail(ctx, CC_ERR_ARG, "cc_sort_coordinates: position outside [-1, 2^31 - 2]");
        max_pos = pos[i] > max_pos ? pos[i] : max_pos;
        max_tid = tid[i] > max_tid ? tid[i] : max_tid;
        no_contig |= tid[i] < 0;
    }
    auto passes_for = [](uint64_t max_key) { int p = 1; while (p < 4 && (max_key >> (8 * p))) p++; return p; };
    const int passes_lo = passes_for((((uint64_t)max_pos + 1) << 1) | 1);
    const uint32_t tid_last = (uint32_t)(max_tid + 1);                   // key of records without a contig
    const int passes_hi = (max_tid <= 0 && !no_contig) ? 0 : passes_for(tid_last);
    // staged input: tid[n], pos[n] (i32) then flag[n] (u16)
    TRY(ensure(ctx, ctx->cs_in, n * 10));
    TRY(ensure(ctx, ctx->cs_k0, n * 4)); TRY(ensure(ctx, ctx->cs_k1, n * 4));
    TRY(ensure(ctx, ctx->cs_v0, n * 4)); TRY(ensure(ctx, ctx->cs_v1, n * 4));
    int32_t *d_tid = ptr<int32_t>(ctx->cs_in), *d_pos = d_tid + n;
    uint16_t *d_flag = reinterpret_cast<uint16_t *>(d_pos + n);
    CK(cudaMemcpyAsync(d_tid, tid, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_pos, pos, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_flag, flag, n * 2, cudaMemcpyHostToDevice, ctx->stream));
    uint32_t *k0 = ptr<uint32_t>(ctx->cs_k0), *k1 = ptr<uint32_t>(ctx->cs_k1);
    uint32_t *v0 = ptr<uint32_t>(ctx->cs_v0), *v1 = ptr<uint32_t>(ctx->cs_v1);
    uint32_t *ks, *vs, *kf, *vf;
    LAUNCH(coord_pos_key_kernel, blocks_for(n, 256), 256, d_pos, d_flag, n, k0);
    TRY(onesweep_pairs(ctx, k0, k1, nullptr, v0, v1, n, passes_lo, &ks, &vs, &kf, &vf));
    if (passes_hi) {
        // LSD over the two words: the second sort is stable, so position order survives inside a contig
        LAUNCH(coord_tid_key_kernel, blocks_for(n, 256), 256, d_tid, vs, n, tid_last, kf);
        uint32_t *kin = kf, *kother = ks, *vin = vs, *vother = vf;
        TRY(onesweep_pairs(ctx, kin, kother, vin, vin, vother, n, passes_hi, &ks, &vs, &kf, &vf));
    }
    CK(cudaMemcpyAsync(order, vs, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return CC_OK;
}

int cc_profile(cc_ctx *ctx, int enable)
{
    if (!ctx) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    TRY(collect_phases(ctx));
    for (auto &t : ctx->phase) { t.ms = 0.0; t.runs = 0; }
    ctx->profile = enable != 0;
    return CC_OK;
}

int cc_phase_ms(cc_ctx *ctx, int phase, double *ms, uint64_t *launches)
{
    if (!ctx || phase < 0 || phase >= CC_PH_COUNT) return CC_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    TRY(collect_phases(ctx));
    if (ms) *ms = ctx->phase[phase].ms;
    if (launches) *launches = ctx->phase[phase].runs;
    return CC_OK;
}

// ---- multi-GPU: the path's one collective ---------------------------------------------------------
int cc_comm_unique_id(uint8_t id[CC_COMM_ID_BYTES])
{
    if (!id) return CC_ERR_ARG;
    NcclApi &api = nccl_api();
    if (!api.handle) { g_create_error = api.error; return CC_ERR_COMM; }
    static_assert(CC_COMM_ID_BYTES == sizeof(ncclUniqueId), "CC_COMM_ID_BYTES must be NCCL's unique-id size");
    ncclUniqueId uid;
    ncclResult_t r = api.GetUniqueId(&uid);
    if (r != ncclSuccess) { g_create_error = std::string("ncclGetUniqueId: ") + api.GetErrorString(r); return CC_ERR_COMM; }
    memcpy(id, &uid, sizeof(uid));
    return CC_OK;
}

int cc_comm_init(cc_ctx *ctx, const uint8_t id[CC_COMM_ID_BYTES], int rank, int nranks)
{
    if (!ctx || !id || nranks < 1 || rank < 0 || rank >= nranks) return CC_ERR_ARG;
    NcclApi &api = nccl_api();
    if (!api.handle) return fail(ctx, CC_ERR_COMM, api.error.c_str());
    CK(cudaSetDevice(ctx->device));
    if (ctx->comm) { api.CommDestroy(ctx->comm); ctx->comm = nullptr; }
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    ncclResult_t r = api.CommInitRank(&ctx->comm, nranks, uid, rank);
    if (r != ncclSuccess) {
        ctx->comm = nullptr;
        ctx->err = std::string("ncclCommInitRank: ") + api.GetErrorString(r);
        return CC_ERR_COMM;
    }
    ctx->comm_rank = rank;
    ctx->comm_nranks = nranks;
    return CC_OK;
}

int cc_comm_destroy(cc_ctx *ctx)
{
    if (!ctx) return CC_ERR_ARG;
    if (ctx->comm) {
        CK(cudaSetDevice(ctx->device));
        CK(cudaStreamSynchronize(ctx->stream));
        nccl_api().CommDestroy(ctx->comm);
        ctx->comm = nullptr;
    }
    ctx->comm_rank = 0;
    ctx->comm_nranks = 1;
    return CC_OK;
}

int cc_stats_allreduce(cc_ctx *ctx)
{
    if (!ctx) return CC_ERR_ARG;
    if (!ctx->grouped) return fail(ctx, CC_ERR_ARG, "cc_stats_allreduce before cc_group / cc_run");
    if (!ctx->comm) return ctx->comm_nranks == 1 ? CC_OK : fail(ctx, CC_ERR_ARG, "cc_stats_allreduce before cc_comm_init");
    CK(cudaSetDevice(ctx->device));
    // [16 counters | CC_HIST_BINS bins] is one buffer (do_group): one in-place sum over the ranks, on the
    // context's stream, ordered after the kernels that filled it
    ncclResult_t r = nccl_api().AllReduce(ctx->stats.p, ctx->stats.p, 16 + (size_t)CC_HIST_BINS, ncclInt64, ncclSum, ctx->comm,
                                          ctx->stream);
    if (r != ncclSuccess) {
        ctx->err = std::string("ncclAllReduce: ") + nccl_api().GetErrorString(r);
        return CC_ERR_COMM;
    }
    return CC_OK;
}

int cc_comm_info(cc_ctx *ctx, int *rank, int *nranks, int *nccl_version)
{
    if (!ctx) return CC_ERR_ARG;
    if (rank) *rank = ctx->comm_rank;
    if (nranks) *nranks = ctx->comm_nranks;
    if (nccl_version) {
        *nccl_version = 0;
        NcclApi &api = nccl_api();
        if (api.handle && api.GetVersion) api.GetVersion(nccl_version);
    }
    return CC_OK;
}

uint64_t cc_kernel_launches(cc_ctx *ctx) { return ctx ? ctx->launches : 0; }
uint64_t cc_device_bytes(cc_ctx *ctx) { return ctx ? ctx->dev_bytes : 0; }

}  // extern "C"

// ======================= allocation-free consensus_maker ==========================================
// Scratch of cc_consensus_maker_static: a fixed array in the library's device image, shared by
// every call on a device (see include/ccgpu.h).  Device code may name it directly; host code
// must not - there the symbol is only a stand-in, so the address is looked up once per device.
namespace {

struct StaticVoteScratch {
    uint4 fields[2 * CC_STATIC_MAX_READS];   // zeroed field records: every member has the same mode fields
    uint32_t perm[CC_STATIC_MAX_READS];      // member list 0..n-1
    uint32_t header[16];                     // [0] segment start, [1] family size, [2] list entry, [3] error flag
    uint32_t meta_out[16];                   // mode fields the kernel writes and nobody reads
};

static __device__ __align__(256) unsigned char g_static_vote_scratch[sizeof(StaticVoteScratch)];

constexpr int kMaxDevices = 64;
std::atomic<void *> g_static_vote_ptr[kMaxDevices];   // device address of the array, per device (nullptr: not looked up yet)

// Everything a call relies on is rewritten here, because the array still holds the previous
// call's contents: member list, header, the error flag and the field records of this family.
__global__ void static_vote_prepare_kernel(uint32_t n)
{
    StaticVoteScratch *s = reinterpret_cast<StaticVoteScratch *>(g_static_vote_scratch);
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        s->perm[i] = i;
        s->fields[2 * i] = make_uint4(0, 0, 0, 0);
        s->fields[2 * i + 1] = make_uint4(0, 0, 0, 0);
    }
    if (i < 16) {
        s->header[i] = (i == 1) ? n : 0u;
        s->meta_out[i] = 0;
    }
}

int static_fail(int code, const std::string &msg)
{
    g_create_error = msg;
    return code;
}

}  // namespace

extern "C" int cc_consensus_maker_static(const uint8_t *d_rows, uint32_t n, uint32_t lpad, uint32_t stride,
                                         double cutoff, uint8_t *d_out_row, void *stream)
{
    if (!d_rows || !d_out_row || n == 0) return static_fail(CC_ERR_ARG, "cc_consensus_maker_static: NULL pointer or empty family");
    if (n > CC_STATIC_MAX_READS)
        return static_fail(CC_ERR_TOO_LARGE, "cc_consensus_maker_static: family of " + std::to_string(n) +
                                                 " reads exceeds CC_STATIC_MAX_READS (" +
                                                 std::to_string(CC_STATIC_MAX_READS) + "); use cc_consensus_maker");
    if (lpad == 0 || lpad % 32 || stride % 64 || stride < lpad / 2 + lpad)
        return static_fail(CC_ERR_ARG, "cc_consensus_maker_static: need lpad % 32 == 0, stride % 64 == 0, stride >= 3*lpad/2");
    if (((uintptr_t)d_rows & 3) || ((uintptr_t)d_out_row & 3))
        return static_fail(CC_ERR_ARG, "cc_consensus_maker_static: device rows must be 4-byte aligned");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return static_fail(CC_ERR_CUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices)
        return static_fail(CC_ERR_ARG, "cc_consensus_maker_static: device index " + std::to_string(dev) +
                                           " is outside the " + std::to_string(kMaxDevices) + " devices this entry point keeps an address for");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    // the call copies the error flag back and waits for it: neither is legal on a capturing stream,
    // and trying would invalidate the caller's capture
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    e = cudaStreamIsCapturing(st, &cap);
    if (e != cudaSuccess) return static_fail(CC_ERR_CUDA, std::string("cudaStreamIsCapturing: ") + cudaGetErrorString(e));
    if (cap != cudaStreamCaptureStatusNone)
        return static_fail(CC_ERR_ARG, "cc_consensus_maker_static: the stream is capturing; this call synchronises and cannot be captured");
    void *scratch = g_static_vote_ptr[dev].load(std::memory_order_acquire);
    if (scratch == nullptr) {
        // once per device, with that device current: the symbol's device address.  Two host threads
        // making the first call together both look it up and publish the same value.
        e = cudaGetSymbolAddress(&scratch, g_static_vote_scratch);
        if (e != cudaSuccess || scratch == nullptr)
            return static_fail(CC_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e));
        g_static_vote_ptr[dev].store(scratch, std::memory_order_release);
    }
    StaticVoteScratch *s = reinterpret_cast<StaticVoteScratch *>(scratch);
    static_vote_prepare_kernel<<<(n + 255) / 256 > 0 ? (n + 255) / 256 : 1, 256, 0, st>>>(n);
    e = cudaGetLastError();
    if (e != cudaSuccess) return static_fail(CC_ERR_CUDA, std::string("prepare launch: ") + cudaGetErrorString(e));
    vote_kernel<<<1, 32, 0, st>>>(s->fields, d_rows, stride, lpad, s->perm, &s->header[0], &s->header[1], &s->header[2], 1u,
                                  cutoff, d_out_row, reinterpret_cast<uint8_t *>(&s->meta_out[0]),
                                  reinterpret_cast<uint16_t *>(&s->meta_out[4]), reinterpret_cast<int32_t *>(&s->meta_out[8]),
                                  reinterpret_cast<uint8_t *>(&s->meta_out[12]), &s->header[3]);
    e = cudaGetLastError();
    if (e != cudaSuccess) return static_fail(CC_ERR_CUDA, std::string("vote launch: ") + cudaGetErrorString(e));
    uint32_t err = 0;
    e = cudaMemcpyAsync(&err, &s->header[3], sizeof(err), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return static_fail(CC_ERR_CUDA, std::string("cc_consensus_maker_static: ") + cudaGetErrorString(e));
    if (err)
        return static_fail(CC_ERR_BASE, "a base other than A/C/G/T carries Q>=30 "
                                        "(the reference raises IndexError/ValueError, SSCS_maker.py:119-122)");
    return CC_OK;
}
