/*
 * ccgpu.h - C ABI of libccgpu.so: the B200 (sm_100a) implementation of
 * ConsensusCruncher's consensus hot path.
 *
 * The reference has no FFI: its consensus path is Python that keeps reads in
 * dictionaries.  These entry points are what a binding
 * for that path would call instead; every one names the reference code it
 * replaces (paths relative to the reference repository).  Plain pointers and
 * sizes only - no torch / numpy types.  INTEGRATION.md shows the ctypes
 * binding a maintainer of the reference would add.
 *
 * Conventions
 *   - return 0 on success, a negative cc_status on failure; cc_last_error()
 *     gives the message (pass NULL for a failed cc_create).
 *   - the caller owns every host buffer; the library owns device memory until
 *     cc_destroy().  One context per GPU; a context is not thread-safe.
 *   - there is no CPU fallback: cc_create fails without an sm_100 device.
 *   - all work is issued on the context's stream (cc_set_stream); calls that
 *     return counts synchronise that stream.
 *
 * Input layout ("packed reads", produced on host by consensuscruncher_b200/pack.py,
 * i.e. the filter / mate-pairing half of read_bam, consensus_helper.py:376-436):
 *   reads are in pair-completion order, read 2k = first-seen mate of pair k,
 *   read 2k+1 = second-seen mate.
 *   fields   n x 32-byte cc_fields records
 *   payload  n x stride bytes: lpad/2 bytes of BAM 4-bit base codes, two per byte with
 *            the EVEN position in the LOW nibble (the nibbles of BAM's own bytes swapped),
 *            then lpad quality bytes, zero padded; lpad % 32 == 0, stride % 64 == 0,
 *            stride >= 3*lpad/2
 *   dup_bc   n_bc x u32: barcode id of the complementary strand's barcode
 *            (duplex_tag, consensus_helper.py:629-639), 0xffffffff if absent
 */
#ifndef CCGPU_H
#define CCGPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cc_ctx cc_ctx;

/* One read's key and record fields (unique_tag, consensus_helper.py:293-301, plus
 * the fields create_aligned_segment takes modes of, :569-581). */
typedef struct {
    uint32_t bc;      /* barcode dictionary id                         */
    uint32_t cig;     /* ordered cigar-pair dictionary id (cigar_order) */
    int32_t tid, pos, mtid, mpos;
    int32_t tlen;
    uint16_t flag;    /* 0x10 orientation and 0x40/0x80 read number are key fields */
    uint8_t mapq;
    uint8_t rg;       /* read-group dictionary id, 0xff = no RG tag     */
} cc_fields;

typedef struct {
    double cutoff;    /* SSCS_maker.py --cutoff: fraction of passing bases the winner needs */
    int32_t scorrect; /* ConsensusCruncher.py --scorrect: run the rescue + second DCS pass  */
} cc_params;

enum cc_status {
    CC_OK = 0,
    CC_ERR_CUDA = -1,        /* CUDA runtime error (message has the call)               */
    CC_ERR_ARG = -2,         /* bad geometry / NULL pointer / wrong call order          */
    CC_ERR_NO_DEVICE = -3,   /* no sm_100 device                                        */
    CC_ERR_BASE = -4,        /* N or IUPAC base with Q>=30 inside a family: the reference
                                raises IndexError / ValueError (SSCS_maker.py:119-122)   */
    CC_ERR_DOMAIN = -5,      /* a read pair does not found/join its two families together:
                                outside the domain where a consensus pair name owns exactly
                                two equal families (SSCS_maker.py:275) */
    CC_ERR_TOO_LARGE = -6,
    CC_ERR_COMM = -7         /* NCCL could not be loaded, or a communicator call failed   */
};

/* Result arrays readable with cc_fetch (element types in brackets).  Per-family
 * arrays are indexed by family id in order of family creation. */
enum cc_array {
    CC_READ_FAM = 0,     /* [u32 x n_reads]  family of each read                      */
    CC_FAM_SIZE = 1,     /* [u32 x n_fam]    tag_dict[tag]                            */
    CC_FAM_TMPL = 2,     /* [u32 x n_fam]    first member = template read             */
    CC_FAM_MAPQ = 3,     /* [u8]  read_mode('mapping_quality')                        */
    CC_FAM_FLAG = 4,     /* [u16] consensus_flag                                      */
    CC_FAM_TLEN = 5,     /* [i32] read_mode('template_length')                        */
    CC_FAM_RG = 6,       /* [u8]  read_mode("get_tag('RG')"), 0xff = none             */
    CC_PARTNER = 7,      /* [i32] family with the duplex_tag key, -1 = none           */
    CC_SSCS_ROW = 8,     /* [i32] row in CC_SSCS_ROWS or -1 (singleton)               */
    CC_CORR_ROW = 9,     /* [i32] row in CC_CORR_ROWS or -1                           */
    CC_DCS_ROW = 10,     /* [i32] row in CC_DCS_ROWS or -1 (set on the leading strand) */
    CC_DCSSC_ROW = 11,   /* [i32] row in CC_DCSSC_ROWS or -1                          */
    CC_SSCS_ROWS = 12,   /* [u8 x n_sscs x stride]  consensus_maker output            */
    CC_CORR_ROWS = 13,   /* [u8 x n_corr x stride]  corrected singletons              */
    CC_DCS_ROWS = 14,    /* [u8 x n_dcs x stride]   duplex_consensus (DCS_maker)      */
    CC_DCSSC_ROWS = 15,  /* [u8 x n_dcssc x stride] DCS after singleton correction; its first n_dcs rows
                            ARE the CC_DCS_ROWS (two SSCS pair up identically in both DCS_maker runs),
                            the pairs that involve a corrected singleton follow */
    CC_DCS_MAPQ = 16, CC_DCS_FLAG = 17, CC_DCS_TLEN = 18, CC_DCS_RG = 19,         /* per DCS row   */
    CC_DCSSC_MAPQ = 20, CC_DCSSC_FLAG = 21, CC_DCSSC_TLEN = 22, CC_DCSSC_RG = 23, /* per DCS-SC row */
    CC_STATS = 24,       /* [i64 x 16] see cc_stat                                    */
    CC_FAM_HIST = 25,    /* [u64 x CC_HIST_BINS] families per size (last bin: >= CC_HIST_BINS-1) */
    /* the four row arrays once more in bit planes, after cc_pack_rows: [u8 x rows x stride_c] */
    CC_SSCS_PLANES = 26, CC_CORR_PLANES = 27, CC_DCS_PLANES = 28, CC_DCSSC_PLANES = 29
};

#define CC_HIST_BINS 65536

/* Indices into CC_STATS: the counters of the reference's stats.txt blocks
 * (SSCS_maker.py:312-318, DCS_maker.py:264-270, singleton_correction.py:294-300). */
enum cc_stat {
    CC_ST_SSCS = 0, CC_ST_SINGLETONS = 1, CC_ST_DCS = 2, CC_ST_SSCS_SINGLETONS = 3,
    CC_ST_SC_BY_SSCS = 4, CC_ST_SC_BY_SINGLETON = 5, CC_ST_UNCORRECTED = 6,
    CC_ST_DCS_SC = 7, CC_ST_SSCS_SC_SINGLETONS = 8
};

enum cc_count { CC_N_READS = 0, CC_N_FAM = 1, CC_N_SSCS = 2, CC_N_CORR = 3, CC_N_DCS = 4, CC_N_DCSSC = 5 };

/* Timed phases (cc_profile / cc_phase_ms). */
enum cc_phase {
    CC_PH_H2D = 0, CC_PH_HASH = 1, CC_PH_SORT = 2, CC_PH_SEGMENT = 3, CC_PH_JOIN = 4,
    CC_PH_CLASSIFY = 5, CC_PH_VOTE = 6, CC_PH_RESCUE = 7, CC_PH_DCS = 8, CC_PH_DCS_SC = 9,
    CC_PH_VOTE_AUX = 10,   /* segment list + the warp-per-family kernel for families > 127 reads */
    CC_PH_COUNT = 11
};

/* Context on `device` (must be an sm_100 GPU).  Replaces the per-stage dictionary
 * set-up of SSCS_maker.py:211-214 / DCS_maker.py:165-178 / singleton_correction.py:145-155. */
cc_ctx *cc_create(int device, const cc_params *params);
void cc_destroy(cc_ctx *ctx);
const char *cc_last_error(cc_ctx *ctx);
int cc_set_params(cc_ctx *ctx, const cc_params *params);

/* Issue all later work on `stream` (a cudaStream_t); NULL restores the context's own stream. */
int cc_set_stream(cc_ctx *ctx, void *stream);

/* Hand over one batch of packed reads from HOST memory (copied to the device on the
 * context's stream).  Replaces the read_dict/pair_dict hand-over at the end of
 * read_bam's pairing step (consensus_helper.py:418-441). */
int cc_load_reads(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload, uint64_t n,
                  uint32_t lpad, uint32_t stride, const uint32_t *dup_bc, uint32_t n_bc);

/* Same, for buffers that already live in DEVICE memory (not copied, must stay valid);
 * fields 32-byte aligned, payload 16-byte aligned (any cudaMalloc'd buffer is). */
int cc_load_reads_device(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload, uint64_t n,
                         uint32_t lpad, uint32_t stride, const uint32_t *dup_bc, uint32_t n_bc);

/* Family grouping by the exact family tuple: pair by pair (both mates' families at once) and tile by tile in
 * shared memory, with a small global table for the families at tile edges (CCGPU_GROUP in the environment at
 * cc_create selects the other paths - tile: read by read in tiles, hash: one table trip per read, sort: key hash
 * + stable radix sort + exact boundaries - same results); families in creation order, sizes, size histogram.  Replaces read_dict / tag_dict / csn_pair_dict
 * construction (consensus_helper.py:438-475) and Counter(tag_dict.values())
 * (SSCS_maker.py:361). */
int cc_group(cc_ctx *ctx);

/* Per-position vote for every family of size >= 2 and the family's mode fields.
 * Replaces consensus_maker (SSCS_maker.py:75-157), read_mode / consensus_flag
 * (consensus_helper.py:483-535) and the size-1 / size>=2 split (SSCS_maker.py:274-292). */
int cc_sscs_vote(cc_ctx *ctx);

/* Complementary-key join over the families and duplex consensus of paired SSCS.
 * Replaces duplex_tag (consensus_helper.py:605-648), the probes and duplex_consensus of
 * DCS_maker.py:226-259, :99-121 and the two-record modes of create_aligned_segment. */
int cc_dcs_join(cc_ctx *ctx);

/* Singleton rescue (by SSCS, else by singleton) and the DCS pass over
 * SSCS + corrected singletons.  Replaces singleton_correction.py:249-285, :61-104 and the
 * second DCS_maker run (ConsensusCruncher.py:247-270).  Needs cc_dcs_join first. */
int cc_sc_rescue(cc_ctx *ctx);

/* cc_group, cc_sscs_vote, cc_dcs_join and (if params.scorrect) cc_sc_rescue in one call.  The host needs the
 * row counts to size the output rows: it reads them from pinned memory while later kernels run (a side stream
 * copies the counters once the grouping has placed the reads and again after the classification), so the call
 * waits for the device once, at its end.
 * The stage calls may be repeated on a loaded batch (cc_set_params with another cutoff, then cc_sscs_vote,
 * cc_dcs_join, cc_sc_rescue again): rows, lists and CC_STATS are made afresh every time. */
int cc_run(cc_ctx *ctx);

/* Whole path from host buffers to host results: cc_load_reads + cc_run; afterwards
 * fetch what is needed with cc_fetch. */
int cc_run_host(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload, uint64_t n,
                uint32_t lpad, uint32_t stride, const uint32_t *dup_bc, uint32_t n_bc);

/* Compact input for instruments that bin their qualities (at most 15 distinct non-zero values,
 * e.g. the 4 of a NovaSeq): a row is lpad/2 base bytes, as in `payload`, followed by lpad/2 bytes of
 * 4-bit quality codes (even position in the low nibble), stride4 % 16 == 0, stride4 >= lpad;
 * qual_table[code] is the quality byte, qual_table[0] == 0 (code 0 pads the rows).  The rows are
 * expanded on the device into the standard layout (stride = 3*lpad/2 rounded up to 64); everything
 * after that, results included, is as with cc_load_reads / cc_run_host.  Same reference code as
 * cc_load_reads (read_bam, consensus_helper.py:376-436); it moves 29 % fewer bytes over PCIe at
 * lpad = 128. */
int cc_load_reads_q4(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload4, uint64_t n, uint32_t lpad,
                     uint32_t stride4, const uint8_t qual_table[16], const uint32_t *dup_bc, uint32_t n_bc);
int cc_run_host_q4(cc_ctx *ctx, const cc_fields *fields, const uint8_t *payload4, uint64_t n, uint32_t lpad,
                   uint32_t stride4, const uint8_t qual_table[16], const uint32_t *dup_bc, uint32_t n_bc);

/* The most compact input: half the bytes of cc_load_reads for binned instruments.  Every byte of a batch
 * crosses PCIe once, so the bytes a read costs there set the end-to-end rate.
 *   pairs   n/2 x 40-byte cc_pair: the mates of a pair share barcode and cigar pair and mirror each other's
 *           coordinates (read 2k+1 has tid/pos and mtid/mpos of read 2k swapped), so one record describes
 *           both cc_fields records.  A batch whose pairs do not mirror must use cc_load_reads.
 *   planes  n x stride_c bytes: lpad/4 bytes of 2-bit bases (A,C,G,T = 0..3; position p in bits 2p, 2p+1 of the
 *           plane, little endian) then lpad*qbits/8 bytes of qbits-bit codes (position p at bit p*qbits);
 *           stride_c % 16 == 0.  qual_table[code] is the position's quality; base_table[code] is 0xff for
 *           "take the 2-bit base" or the literal 4-bit BAM base code of the position (15 for N, 0 for the
 *           padding after the read's end, an IUPAC code).  The host picks the smallest qbits (1..4) that
 *           names every (quality, literal base) combination of the batch; none fits: cc_load_reads.
 * Expanded on the device into cc_fields records and standard rows; everything after that is as with
 * cc_load_reads.  Same reference code as cc_load_reads (read_bam, consensus_helper.py:376-436).
 * At lpad = 128 with 7 codes: 20 + 80 = 100 bytes per read against 224. */
typedef struct {
    uint32_t bc, cig;
    int32_t tid, pos, mtid, mpos;      /* of the first-seen mate (read 2k)                   */
    int32_t tlen1, tlen2;
    uint16_t flag1, flag2;
    uint8_t mapq1, mapq2, rg1, rg2;
} cc_pair;
int cc_load_reads_compact(cc_ctx *ctx, const cc_pair *pairs, const uint8_t *planes, uint64_t n, uint32_t lpad,
                          uint32_t stride_c, uint32_t qbits, const uint8_t qual_table[16], const uint8_t base_table[16],
                          const uint32_t *dup_bc, uint32_t n_bc);
int cc_run_host_compact(cc_ctx *ctx, const cc_pair *pairs, const uint8_t *planes, uint64_t n, uint32_t lpad,
                        uint32_t stride_c, uint32_t qbits, const uint8_t qual_table[16], const uint8_t base_table[16],
                        const uint32_t *dup_bc, uint32_t n_bc);

/* cc_load_reads_compact with 16-byte pair records: what many pairs have in common sits once in a class table
 * (at most 65 536 classes), a pair keeps its barcode, its two positions, the first-seen mate's TLEN and the
 * number of its class; the second mate's TLEN is the negative of the first's.  A batch with a pair whose TLENs
 * are not each other's negatives, with |TLEN| > 32767 or with more classes must use cc_load_reads_compact.
 * planes / qbits / tables as there.  A class index >= n_classes makes the run return CC_ERR_ARG.
 * At lpad = 128 with 7 codes: 8 + 80 = 88 bytes per read against 224.  Same reference code as
 * cc_load_reads (read_bam, consensus_helper.py:376-436). */
typedef struct {
    uint32_t bc;
    int32_t pos, mpos;                 /* of the first-seen mate (read 2k)                   */
    int16_t tlen1;                     /* TLEN of read 2k; read 2k+1 has -tlen1              */
    uint16_t cls;                      /* index into the class table                         */
} cc_pair16;
typedef struct {
    uint32_t cig;
    int32_t tid, mtid;                 /* of the first-seen mate                             */
    uint16_t flag1, flag2;
    uint8_t mapq1, mapq2, rg1, rg2;
    uint32_t reserved[3];              /* 32-byte records                                    */
} cc_pair_class;
int cc_load_reads_classes(cc_ctx *ctx, const cc_pair16 *pairs, const cc_pair_class *classes, uint32_t n_classes,
                          const uint8_t *planes, uint64_t n, uint32_t lpad, uint32_t stride_c, uint32_t qbits,
                          const uint8_t qual_table[16], const uint8_t base_table[16], const uint32_t *dup_bc, uint32_t n_bc);
int cc_run_host_classes(cc_ctx *ctx, const cc_pair16 *pairs, const cc_pair_class *classes, uint32_t n_classes,
                        const uint8_t *planes, uint64_t n, uint32_t lpad, uint32_t stride_c, uint32_t qbits,
                        const uint8_t qual_table[16], const uint8_t base_table[16], const uint32_t *dup_bc, uint32_t n_bc);

/* counts[CC_N_*] after the stage that produces them (synchronises). */
int cc_counts(cc_ctx *ctx, uint64_t counts[8]);

/* Copy result array `what` (enum cc_array) into host memory; `bytes` must be the exact size. */
int cc_fetch(cc_ctx *ctx, int what, void *host_dst, uint64_t bytes);
/* Several result arrays at once: every copy is issued before the one wait (cc_fetch waits per array).
 * what[k] / host_dst[k] / bytes[k] as for cc_fetch; arrays of size 0 are skipped. */
int cc_fetch_many(cc_ctx *ctx, uint32_t count, const int *what, void *const *host_dst, const uint64_t *bytes);
/* Result rows in bit planes, for a host that pays for every byte it reads back.  Finds the (quality, literal
 * base) combinations that occur in the rows produced so far; with at most 16 of them every row is written once
 * more in the layout of cc_load_reads_compact's `planes` - lpad/4 bytes of 2-bit bases, lpad*qbits/8 bytes of
 * codes, stride_c bytes per row; qual_table[code] / base_table[code] as there (0xff: take the 2-bit base; 15: N;
 * 0: the padding after the read's end) - and CC_*_PLANES can be fetched instead of CC_*_ROWS (96 instead of 192
 * bytes per row at lpad = 128).  *qbits == 0 on return: more than 16 combinations, fetch the rows as they are.
 * Consensus qualities are 60, 0 or a lone supporter's own quality (SSCS_maker.py:139-153, DCS_maker.py:107-119),
 * so binned input gives few of them.  Synchronises; a later stage call or cc_set_params discards the planes. */
int cc_pack_rows(cc_ctx *ctx, uint32_t *qbits, uint32_t *stride_c, uint8_t qual_table[16], uint8_t base_table[16]);
/* Size in bytes of result array `what` (0 if not produced yet). */
uint64_t cc_array_bytes(cc_ctx *ctx, int what);
/* Device address of result array `what` (NULL if not produced yet); valid until the next load. */
const void *cc_device_ptr(cc_ctx *ctx, int what);

/* Stand-alone primitives on host buffers, mirroring the reference's function signatures.
 * consensus_maker(readList, cutoff) (SSCS_maker.py:75): rows = n x stride payload rows of
 * one family, out = one row. */
int cc_consensus_maker(cc_ctx *ctx, const uint8_t *rows, uint32_t n, uint32_t lpad, uint32_t stride,
                       double cutoff, uint8_t *out_row);
/* duplex_consensus(read1, read2): DCS_maker.py:99-121 (sc = 0) or
 * singleton_correction.py:61-84 (sc = 1), for n_pairs independent pairs of rows. */
int cc_duplex_consensus(cc_ctx *ctx, const uint8_t *rows1, const uint8_t *rows2, uint64_t n_pairs,
                        uint32_t lpad, uint32_t stride, int sc, uint8_t *out_rows);

/* The duplex join over n records that stand for themselves (one record per tag, as in the SSCS and singleton
 * files the stand-alone stage commands read): partner[i] = the record whose family key - barcode, cigar pair,
 * contig / position of read and mate, strand and read-number bits - is record i's with the barcode halves swapped
 * (dup_bc) and R1 <-> R2, or -1; of several such records the first.  Replaces duplex_tag
 * (consensus_helper.py:605-648) and the dictionary probes of DCS_maker.py:229-247 and
 * singleton_correction.py:253-270 in those commands.  Host buffers. */
int cc_join_records(cc_ctx *ctx, const cc_fields *records, uint64_t n, const uint32_t *dup_bc, uint32_t n_bc, int32_t *partner);

/* Coordinate sort of output records: the `samtools view -bu | samtools sort` of sort_index
 * (ConsensusCruncher.py:10-29) and, applied to concatenated inputs, `samtools merge`
 * (ConsensusCruncher.py:248,282).  order[k] = index of the record of rank k under the stable
 * order (contig index, position, strand bit 0x10 of flag); records without a contig (tid < 0)
 * go last.  Host buffers; position must lie in [-1, 2^31 - 2]. */
int cc_sort_coordinates(cc_ctx *ctx, const int32_t *tid, const int32_t *pos, const uint16_t *flag, uint64_t n,
                        uint32_t *order);

/* Largest family cc_consensus_maker_static accepts (its scratch is sized for this). */
#define CC_STATIC_MAX_READS 4096

/* This entry point has no counterpart in the reference.  It is cc_consensus_maker for
 * callers that are not allowed to allocate device memory while they run - code inside a
 * framework whose memory pool owns the whole GPU, a real-time or plugin host that forbids
 * driver allocations after start-up, a caller between other libraries' graph captures where
 * a cudaMalloc would synchronise the device.  Such callers already hold device buffers, so
 * both the family's rows and the result row are DEVICE pointers here, and the working
 * memory the vote kernel needs (member list, segment header, zeroed field records, error
 * flag; about 145 KB) lives in a fixed array inside the library's device image instead of
 * coming from cudaMalloc.  Every GPU that loads libccgpu.so carries that array until exit.
 *
 *   d_rows     n x stride bytes, payload rows of ONE family, in device memory
 *   d_out_row  stride bytes in device memory: the consensus row
 *   stream     cudaStream_t of the calling thread's current device (NULL = default stream)
 *
 * Runs on the current device.  Calls no cudaMalloc / cudaFree and no allocating library.
 * The first call on each device looks up the array's device address; make that first call
 * before any stream capture begins.  Every call copies an error flag back and waits for its
 * own work before returning, so it must never be made on a stream that is capturing: the call
 * checks (cudaStreamIsCapturing) and returns CC_ERR_ARG there instead of breaking the capture.
 * The address look-up may be made by several host threads at once (they publish the same value).
 * All calls on a device share the one array: two calls whose GPU work overlaps (different
 * streams or threads) race; issue them one after the other.
 * Returns CC_ERR_TOO_LARGE if n > CC_STATIC_MAX_READS, CC_ERR_BASE as cc_consensus_maker
 * does; the message is read with cc_last_error(NULL). */
int cc_consensus_maker_static(const uint8_t *d_rows, uint32_t n, uint32_t lpad, uint32_t stride,
                              double cutoff, uint8_t *d_out_row, void *stream);

/* ---- multi-GPU (one process or thread per GPU, one context per rank) -------------------------
 * The path shards by genomic region: families, duplex partners and rescue partners of a read all
 * start at that read's position, so ranks that own disjoint sets of the reference's bed regions
 * (bed_separator, consensus_helper.py:38-54; the region loops SSCS_maker.py:229-257,
 * DCS_maker.py:185-211, singleton_correction.py:175-213) never exchange reads.  The one exchange is
 * the sum of what the reference prints once per stage: the stats.txt counters
 * (SSCS_maker.py:312-318, DCS_maker.py:264-270, singleton_correction.py:294-300) and
 * Counter(tag_dict.values()) (SSCS_maker.py:360-365).  It is a direct ncclAllReduce(sum, int64) of
 * [16 counters | CC_HIST_BINS bins] - one buffer, one collective - on the context's stream.
 * NCCL is loaded with dlopen("libnccl.so.2") at the first of these calls (CCGPU_NCCL_LIB names
 * another file); nothing else in the library needs it. */
#define CC_COMM_ID_BYTES 128
/* Rank 0 creates the id (ncclGetUniqueId) and hands the 128 bytes to every rank by whatever
 * means the host has (a file, a socket, MPI, a launcher's key-value store). */
int cc_comm_unique_id(uint8_t id[CC_COMM_ID_BYTES]);
/* Collective over all ranks: joins the communicator (ncclCommInitRank) on the context's device. */
int cc_comm_init(cc_ctx *ctx, const uint8_t id[CC_COMM_ID_BYTES], int rank, int nranks);
/* After cc_run / cc_group (+ later stages): CC_STATS and CC_FAM_HIST become the sums over all ranks,
 * in place, asynchronously on the context's stream (cc_fetch synchronises).  Every rank must call it
 * the same number of times.  Without a communicator (single GPU) it does nothing.  Call it once per
 * run or stage, as the reference writes its stats, not once per batch. */
int cc_stats_allreduce(cc_ctx *ctx);
int cc_comm_destroy(cc_ctx *ctx);
/* rank / nranks of the context's communicator (0 / 1 without one) and NCCL's version code (0 if not loaded). */
int cc_comm_info(cc_ctx *ctx, int *rank, int *nranks, int *nccl_version);

/* Phase timing with CUDA events on the context's stream. */
int cc_profile(cc_ctx *ctx, int enable);
/* Milliseconds spent in `phase` (enum cc_phase) since the last cc_profile(ctx, 1) and the
 * number of times it ran; synchronises. */
int cc_phase_ms(cc_ctx *ctx, int phase, double *ms, uint64_t *launches);
/* Number of kernels this library has launched on this context. */
uint64_t cc_kernel_launches(cc_ctx *ctx);
/* Device bytes currently held by the context. */
uint64_t cc_device_bytes(cc_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* CCGPU_H */
