/*
 * TEST INFRASTRUCTURE - CPU restatement (plain C) of the reference's consensus
 * hot path, used only as the checker for the CUDA path and as the timed CPU
 * baseline ("port").  Nothing in the product links or calls this file.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this code against
 * tests/golden/ (json.gz files), which oracle/make_golden.py produced by running the
 * unmodified reference scripts (oracle/run_reference.py) in the build
 * container, and against the reference's own known-answer vectors
 * (SURVEY.md App. D).
 *
 * What is restated, on the packed SoA of consensuscruncher_b200/pack.py
 * (reads in pair-completion order; 32-byte field records; 4-bit bases + u8
 * qualities per row):
 *
 *   grouping            consensus_helper.py:438-475  read_dict / tag_dict: one family per
 *                       distinct `unique_tag` (:293-301), members in arrival order,
 *                       families numbered in order of creation
 *   SSCS vote           SSCS_maker.py:103-155        consensus_maker
 *   family mode fields  consensus_helper.py:483-497  read_mode (mapq, tlen, RG)
 *                       consensus_helper.py:500-535  consensus_flag
 *                       ties: first-appearing value (what randint -> 0 selects; App. B Q1)
 *   SSCS / singleton    SSCS_maker.py:274-298        size 1 -> singleton, else consensus
 *   duplex key          consensus_helper.py:629-648  duplex_tag: barcode halves swapped, R1<->R2
 *   DCS                 DCS_maker.py:226-259, :107-119 duplex_consensus (no quality filter)
 *   singleton rescue    singleton_correction.py:249-285, :69-82 (both Q > 29)
 *   DCS after rescue    DCS_maker.py on sscs + sscs.correction + singleton.correction
 *                       (ConsensusCruncher.py:247-270); which strand leads a pair follows
 *                       the merge order of those three files, then family creation order
 *
 * In-domain assumption shared with the CUDA path (checked by cco_domain_faults):
 * the two mates of a pair always found/join their two families together, so a
 * consensus pair name owns exactly two families of equal size
 * (SSCS_maker.py:275).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    uint32_t bc, cig;
    int32_t tid, pos, mtid, mpos;
    int32_t tlen;
    uint16_t flag;
    uint8_t mapq, rg;
} cco_fields;

enum { CCO_OK = 0, CCO_ENOMEM = -1, CCO_EBASE_N = -2, CCO_EBASE_IUPAC = -3, CCO_EARG = -4 };

typedef struct {
    /* inputs (borrowed) */
    const cco_fields *f;
    const uint8_t *payload;
    const uint32_t *dup_bc;
    uint64_t n;
    uint32_t n_bc, lpad, stride;
    double cutoff;
    int q_sscs, q_sc, q_cap;
    /* grouping */
    uint64_t n_fam;
    uint32_t *read_fam;   /* N */
    uint32_t *fam_size, *fam_tmpl, *fam_start, *members;
    uint8_t *fam_mapq, *fam_rg;
    uint16_t *fam_flag;
    int32_t *fam_tlen;
    int32_t *partner, *sscs_row, *corr_row, *dcs_row, *dcssc_row;
    uint64_t n_sscs, n_corr, n_dcs, n_dcssc;
    uint8_t *sscs_rows, *corr_rows, *dcs_rows, *dcssc_rows;
    uint8_t *dcs_mapq, *dcs_rg, *dcssc_mapq, *dcssc_rg;
    uint16_t *dcs_flag, *dcssc_flag;
    int32_t *dcs_tlen, *dcssc_tlen;
    int64_t stats[16];
    /* tuple -> family table */
    uint32_t *table;
    uint64_t tmask;
} cco_ctx;

/* ---- key helpers ----------------------------------------------------- */
typedef struct { uint32_t bc, cig; int32_t tid, pos, mtid, mpos; uint32_t bits; } cco_key;

static cco_key key_of(const cco_fields *r)
{
    cco_key k;
    k.bc = r->bc; k.cig = r->cig; k.tid = r->tid; k.pos = r->pos; k.mtid = r->mtid; k.mpos = r->mpos;
    /* orientation (0x10) and read number (0x40 first / 0x80 second in pair) */
    k.bits = ((r->flag & 0x10) ? 1u : 0u) | ((r->flag & 0x40) ? 2u : 0u);
    return k;
}

static int key_eq(const cco_key *a, const cco_key *b)
{
    return a->bc == b->bc && a->cig == b->cig && a->tid == b->tid && a->pos == b->pos &&
           a->mtid == b->mtid && a->mpos == b->mpos && a->bits == b->bits;
}

static uint64_t key_hash(const cco_key *k)
{
    uint64_t h = 0x9E3779B97F4A7C15ull;
    const uint32_t w[7] = { k->bc, k->cig, (uint32_t)k->tid, (uint32_t)k->pos,
                            (uint32_t)k->mtid, (uint32_t)k->mpos, k->bits };
    for (int i = 0; i < 7; i++) {
        h ^= w[i];
        h *= 0xff51afd7ed558ccdull;
        h ^= h >> 29;
    }
    return h;
}

static int64_t table_find(const cco_ctx *c, const cco_key *k)
{
    uint64_t s = key_hash(k) & c->tmask;
    for (;;) {
        uint32_t v = c->table[s];
        if (v == 0xFFFFFFFFu) return -1;
        cco_key o = key_of(&c->f[c->fam_tmpl[v]]);
        if (key_eq(&o, k)) return (int64_t)v;
        s = (s + 1) & c->tmask;
    }
}

/* ---- grouping: consensus_helper.py:438-475 --------------------------------- */
static int group(cco_ctx *c)
{
    uint64_t n = c->n, cap = 16;
    while (cap < 2 * n + 2) cap <<= 1;
    c->tmask = cap - 1;
    c->table = (uint32_t *)malloc(cap * sizeof(uint32_t));
    c->read_fam = (uint32_t *)malloc((n + 1) * sizeof(uint32_t));
    c->fam_size = (uint32_t *)calloc(n + 1, sizeof(uint32_t));
    c->fam_tmpl = (uint32_t *)malloc((n + 1) * sizeof(uint32_t));
    if (!c->table || !c->read_fam || !c->fam_size || !c->fam_tmpl) return CCO_ENOMEM;
    memset(c->table, 0xFF, cap * sizeof(uint32_t));
    uint64_t nf = 0;
    for (uint64_t i = 0; i < n; i++) {
        cco_key k = key_of(&c->f[i]);
        uint64_t s = key_hash(&k) & c->tmask;
        for (;;) {
            uint32_t v = c->table[s];
            if (v == 0xFFFFFFFFu) {             /* new tag: family created by this read */
                c->table[s] = (uint32_t)nf;
                c->fam_tmpl[nf] = (uint32_t)i;
                c->fam_size[nf] = 1;
                c->read_fam[i] = (uint32_t)nf;
                nf++;
                break;
            }
            cco_key o = key_of(&c->f[c->fam_tmpl[v]]);
            if (key_eq(&o, &k)) {               /* existing tag: append (PCR duplicate) */
                c->fam_size[v]++;
                c->read_fam[i] = v;
                break;
            }
            s = (s + 1) & c->tmask;
        }
    }
    c->n_fam = nf;
    /* member lists in arrival order */
    c->fam_start = (uint32_t *)malloc((nf + 1) * sizeof(uint32_t));
    c->members = (uint32_t *)malloc((n + 1) * sizeof(uint32_t));
    uint32_t *cur = (uint32_t *)malloc((nf + 1) * sizeof(uint32_t));
    if (!c->fam_start || !c->members || !cur) return CCO_ENOMEM;
    uint32_t acc = 0;
    for (uint64_t g = 0; g < nf; g++) { c->fam_start[g] = acc; cur[g] = acc; acc += c->fam_size[g]; }
    c->fam_start[nf] = acc;
    for (uint64_t i = 0; i < n; i++) c->members[cur[c->read_fam[i]]++] = (uint32_t)i;
    free(cur);
    return CCO_OK;
}

/* ---- payload access ------------------------------------------------------ */
static inline int base_at(const uint8_t *row, uint32_t i)
{
    uint8_t b = row[i >> 1];
    return (i & 1) ? (b >> 4) : (b & 15);          /* even position in the low nibble */
}

static inline void set_base(uint8_t *row, uint32_t i, int code)
{
    uint8_t *b = &row[i >> 1];
    if (i & 1) *b = (uint8_t)((*b & 0x0F) | (code << 4));
    else *b = (uint8_t)((*b & 0xF0) | code);
}

/* ---- SSCS vote: SSCS_maker.py:103-155 ------------------------------------------ */
static int vote(const cco_ctx *c, uint64_t g, uint8_t *out)
{
    const uint32_t lpad = c->lpad, half = lpad / 2;
    const uint32_t n = c->fam_size[g];
    const uint32_t *mem = c->members + c->fam_start[g];
    static const int code_of[5] = { 1, 2, 4, 8, 15 };    /* A C G T N in BAM's 4-bit alphabet */
    memset(out, 0, c->stride);
    for (uint32_t i = 0; i < lpad; i++) {
        uint32_t cnt[5] = { 0, 0, 0, 0, 0 };
        uint64_t qs[4] = { 0, 0, 0, 0 };
        uint32_t fail = 0;
        for (uint32_t j = 0; j < n; j++) {
            const uint8_t *row = c->payload + (uint64_t)mem[j] * c->stride;
            int q = row[half + i];
            int b = base_at(row, i);
            int idx = b == 1 ? 0 : b == 2 ? 1 : b == 4 ? 2 : b == 8 ? 3 : b == 15 ? 4 : -1;
            if (q < c->q_sscs) { fail++; continue; }   /* the reference never looks at a failed base again */
            if (idx < 0) return CCO_EBASE_IUPAC;      /* nuc_lst.index -> ValueError */
            if (idx == 4) return CCO_EBASE_N;         /* quality_score[4] -> IndexError */
            cnt[idx]++;
            qs[idx] += (uint64_t)q;
        }
        int w = 0;
        for (int k = 1; k < 5; k++) if (cnt[k] > cnt[w]) w = k;   /* first maximum, A C G T N */
        uint64_t mq = w < 4 ? qs[w] : 0;
        if (mq > (uint64_t)c->q_cap) mq = (uint64_t)c->q_cap;
        uint32_t pass = n - fail;
        int code = 15;
        if (pass != 0 && ((double)cnt[w] / (double)pass) >= c->cutoff) code = code_of[w];
        set_base(out, i, code);
        out[half + i] = (uint8_t)mq;              /* an N keeps the winner's capped quality sum */
    }
    return CCO_OK;
}

/* ---- read_mode / consensus_flag over a family ---------------------------------- */
static int64_t member_val(const cco_fields *r, int what)
{
    switch (what) { case 0: return r->mapq; case 1: return r->tlen; case 2: return r->rg; default: return r->flag; }
}

static int flag_rank(int64_t v) { return v == 99 ? 0 : v == 83 ? 1 : v == 147 ? 2 : v == 163 ? 3 : 4; }

static int64_t family_mode(const cco_ctx *c, uint64_t g, int what)
{
    const uint32_t n = c->fam_size[g];
    const uint32_t *mem = c->members + c->fam_start[g];
    int64_t best = 0; uint32_t best_cnt = 0; int best_rank = 9;
    for (uint32_t i = 0; i < n; i++) {
        int64_t v = member_val(&c->f[mem[i]], what);
        uint32_t cnt = 0; int first = 1;
        for (uint32_t j = 0; j < n; j++) {
            if (member_val(&c->f[mem[j]], what) == v) { if (j < i) { first = 0; break; } cnt++; }
        }
        if (!first) continue;
        int rank = what == 3 ? flag_rank(v) : 4;
        /* most_common() keeps first-appearance order among equal counts; flags get 99>83>147>163 */
        if (cnt > best_cnt || (cnt == best_cnt && rank < best_rank)) { best = v; best_cnt = cnt; best_rank = rank; }
    }
    return best;
}

static uint8_t family_rg(const cco_ctx *c, uint64_t g)
{
    const uint32_t n = c->fam_size[g];
    const uint32_t *mem = c->members + c->fam_start[g];
    for (uint32_t i = 0; i < n; i++) if (c->f[mem[i]].rg == 0xFF) return 0xFF;   /* KeyError -> no RG at all */
    return (uint8_t)family_mode(c, g, 2);
}

/* mode of two records (create_aligned_segment on [t, d]) */
static int64_t pair_mode(int64_t a, int64_t b, int is_flag)
{
    if (a == b) return a;
    if (is_flag && flag_rank(b) < flag_rank(a)) return b;
    return a;
}

/* ---- duplex_consensus variants ---------------------------------------------------------- */
static void agree(const cco_ctx *c, const uint8_t *r1, const uint8_t *r2, int qmin, uint8_t *out)
{
    const uint32_t lpad = c->lpad, half = lpad / 2;
    memset(out, 0, c->stride);
    for (uint32_t i = 0; i < lpad; i++) {
        int b1 = base_at(r1, i), b2 = base_at(r2, i);
        int q1 = r1[half + i], q2 = r2[half + i];
        if (b1 == b2 && (qmin < 0 || (q1 > qmin && q2 > qmin))) {
            int q = q1 + q2;
            if (q > c->q_cap) q = c->q_cap;
            set_base(out, i, b1);
            out[half + i] = (uint8_t)q;
        } else {
            set_base(out, i, 15);
            out[half + i] = 0;
        }
    }
}

/* row that represents family g in sscs.sc (SSCS consensus or corrected singleton) */
static const uint8_t *sc_row(const cco_ctx *c, uint64_t g)
{
    if (c->fam_size[g] >= 2) return c->sscs_rows + (uint64_t)c->sscs_row[g] * c->stride;
    return c->corr_rows + (uint64_t)c->corr_row[g] * c->stride;
}

static int leads(const cco_ctx *c, uint64_t g, uint64_t p, int after_sc)
{
    /* which of two complementary families is met first by DCS_maker's loop */
    if (after_sc) {
        int gs = c->fam_size[g] >= 2, ps = c->fam_size[p] >= 2;
        if (gs != ps) return gs;              /* sscs file precedes the correction files in the merge */
    }
    return g < p;
}

void *cco_create(const void *fields, const uint8_t *payload, uint64_t n, uint32_t lpad, uint32_t stride,
                 const uint32_t *dup_bc, uint32_t n_bc, double cutoff)
{
    cco_ctx *c = (cco_ctx *)calloc(1, sizeof(cco_ctx));
    if (!c) return NULL;
    c->f = (const cco_fields *)fields; c->payload = payload; c->n = n; c->lpad = lpad; c->stride = stride;
    c->dup_bc = dup_bc; c->n_bc = n_bc; c->cutoff = cutoff; c->q_sscs = 30; c->q_sc = 29; c->q_cap = 60;
    return c;
}

#define ALLOC(p, cnt, T) do { (p) = (T *)calloc((cnt) + 1, sizeof(T)); if (!(p)) return CCO_ENOMEM; } while (0)

/* scorrect != 0 also runs the rescue and the second DCS pass */
int cco_run(void *ctx, int scorrect)
{
    cco_ctx *c = (cco_ctx *)ctx;
    if (c->lpad % 2 || c->stride < c->lpad / 2 + c->lpad) return CCO_EARG;
    int rc = group(c);
    if (rc) return rc;
    const uint64_t nf = c->n_fam;
    ALLOC(c->fam_mapq, nf, uint8_t); ALLOC(c->fam_rg, nf, uint8_t); ALLOC(c->fam_flag, nf, uint16_t);
    ALLOC(c->fam_tlen, nf, int32_t); ALLOC(c->partner, nf, int32_t);
    ALLOC(c->sscs_row, nf, int32_t); ALLOC(c->corr_row, nf, int32_t);
    ALLOC(c->dcs_row, nf, int32_t); ALLOC(c->dcssc_row, nf, int32_t);

    /* complementary family of each family (duplex_tag) */
    for (uint64_t g = 0; g < nf; g++) {
        cco_key k = key_of(&c->f[c->fam_tmpl[g]]);
        c->partner[g] = -1;
        uint32_t d = k.bc < c->n_bc ? c->dup_bc[k.bc] : 0xFFFFFFFFu;
        if (d == 0xFFFFFFFFu) continue;
        k.bc = d;
        k.bits ^= 2u;                                        /* R1 <-> R2 */
        c->partner[g] = (int32_t)table_find(c, &k);
    }
    /* classify, number the output rows in family creation order.  A pair of two SSCS yields the
     * same record in both DCS_maker runs, so the second run's rows are laid out as
     * [the nd DCS rows | pairs that involve a corrected singleton]. */
    uint64_t ns = 0, nc = 0, nd = 0, ndsc = 0;
    int64_t *st = c->stats;
    for (uint64_t g = 0; g < nf; g++) {
        int32_t p = c->partner[g];
        int big = c->fam_size[g] >= 2;
        c->sscs_row[g] = big ? (int32_t)ns++ : -1;
        c->dcs_row[g] = (big && p >= 0 && c->fam_size[p] >= 2 && leads(c, g, (uint64_t)p, 0)) ? (int32_t)nd++ : -1;
        c->corr_row[g] = -1; c->dcssc_row[g] = -1;
        if (scorrect && !big && p >= 0) c->corr_row[g] = (int32_t)nc++;
        if (big) st[0]++; else st[1]++;                          /* SSCS reads / singletons */
        if (big && !(p >= 0 && c->fam_size[p] >= 2)) st[3]++;    /* SSCS singletons (no duplex) */
        if (scorrect) {
            if (!big && p >= 0 && c->fam_size[p] >= 2) st[4]++;  /* rescued by SSCS */
            if (!big && p >= 0 && c->fam_size[p] < 2) st[5]++;   /* rescued by singleton */
            if (!big && p < 0) st[6]++;                          /* uncorrected */
            if (big && p < 0) st[8]++;                           /* SSCS SC singletons */
        }
    }
    if (scorrect) {
        ndsc = nd;
        for (uint64_t g = 0; g < nf; g++) {
            int32_t p = c->partner[g];
            if (p < 0 || !leads(c, g, (uint64_t)p, 1)) continue;
            c->dcssc_row[g] = c->dcs_row[g] >= 0 ? c->dcs_row[g] : (int32_t)ndsc++;
        }
    }
    st[2] = (int64_t)nd; st[7] = (int64_t)ndsc;
    c->n_sscs = ns; c->n_corr = nc; c->n_dcs = nd; c->n_dcssc = ndsc;
    const uint64_t S = c->stride;
    ALLOC(c->sscs_rows, ns * S, uint8_t); ALLOC(c->corr_rows, nc * S, uint8_t);
    ALLOC(c->dcs_rows, nd * S, uint8_t); ALLOC(c->dcssc_rows, ndsc * S, uint8_t);
    ALLOC(c->dcs_mapq, nd, uint8_t); ALLOC(c->dcs_rg, nd, uint8_t); ALLOC(c->dcs_flag, nd, uint16_t);
    ALLOC(c->dcs_tlen, nd, int32_t);
    ALLOC(c->dcssc_mapq, ndsc, uint8_t); ALLOC(c->dcssc_rg, ndsc, uint8_t); ALLOC(c->dcssc_flag, ndsc, uint16_t);
    ALLOC(c->dcssc_tlen, ndsc, int32_t);

    /* SSCS stage */
    for (uint64_t g = 0; g < nf; g++) {
        if (c->fam_size[g] >= 2) {
            rc = vote(c, g, c->sscs_rows + (uint64_t)c->sscs_row[g] * S);
            if (rc) return rc;
            c->fam_mapq[g] = (uint8_t)family_mode(c, g, 0);
            c->fam_tlen[g] = (int32_t)family_mode(c, g, 1);
            c->fam_flag[g] = (uint16_t)family_mode(c, g, 3);
            c->fam_rg[g] = family_rg(c, g);
        } else {
            const cco_fields *r = &c->f[c->fam_tmpl[g]];
            c->fam_mapq[g] = r->mapq; c->fam_tlen[g] = r->tlen; c->fam_flag[g] = r->flag; c->fam_rg[g] = r->rg;
        }
    }
    /* singleton rescue: partner SSCS first, else the partner singleton's original read */
    for (uint64_t g = 0; g < nf && scorrect; g++) {
        if (c->corr_row[g] < 0) continue;
        uint64_t p = (uint64_t)c->partner[g];
        const uint8_t *mine = c->payload + (uint64_t)c->fam_tmpl[g] * S;
        const uint8_t *other = c->fam_size[p] >= 2 ? c->sscs_rows + (uint64_t)c->sscs_row[p] * S
                                                   : c->payload + (uint64_t)c->fam_tmpl[p] * S;
        agree(c, mine, other, c->q_sc, c->corr_rows + (uint64_t)c->corr_row[g] * S);
    }
    /* DCS, then DCS over sscs + corrected singletons */
    for (int pass = 0; pass < (scorrect ? 2 : 1); pass++) {
        for (uint64_t g = 0; g < nf; g++) {
            int32_t row = pass ? c->dcssc_row[g] : c->dcs_row[g];
            if (row < 0) continue;
            uint64_t p = (uint64_t)c->partner[g];
            const uint8_t *a = pass ? sc_row(c, g) : c->sscs_rows + (uint64_t)c->sscs_row[g] * S;
            const uint8_t *b = pass ? sc_row(c, p) : c->sscs_rows + (uint64_t)c->sscs_row[p] * S;
            agree(c, a, b, -1, (pass ? c->dcssc_rows : c->dcs_rows) + (uint64_t)row * S);
            uint8_t *mq = pass ? c->dcssc_mapq : c->dcs_mapq, *rg = pass ? c->dcssc_rg : c->dcs_rg;
            uint16_t *fl = pass ? c->dcssc_flag : c->dcs_flag;
            int32_t *tl = pass ? c->dcssc_tlen : c->dcs_tlen;
            mq[row] = (uint8_t)pair_mode(c->fam_mapq[g], c->fam_mapq[p], 0);
            tl[row] = (int32_t)pair_mode(c->fam_tlen[g], c->fam_tlen[p], 0);
            fl[row] = (uint16_t)pair_mode(c->fam_flag[g], c->fam_flag[p], 1);
            rg[row] = (c->fam_rg[g] == 0xFF || c->fam_rg[p] == 0xFF) ? 0xFF
                                                                   : (uint8_t)pair_mode(c->fam_rg[g], c->fam_rg[p], 0);
        }
    }
    return CCO_OK;
}

/* Families whose pair partner does not found / share the partner family in step:
 * non-zero means the input leaves the domain where a consensus pair name owns
 * exactly two equally sized families. */
uint64_t cco_domain_faults(void *ctx)
{
    cco_ctx *c = (cco_ctx *)ctx;
    uint64_t bad = 0;
    for (uint64_t g = 0; g < c->n_fam; g++) {
        uint32_t t = c->fam_tmpl[g], m = t ^ 1u;
        if (m >= c->n) { bad++; continue; }
        uint32_t h = c->read_fam[m];
        if (c->fam_tmpl[h] != m || c->fam_size[h] != c->fam_size[g]) bad++;
    }
    /* a duplex partner that has another partner (duplex_tag is not its own inverse for undotted
     * barcodes of odd length): the reference's DCS loop then depends on file order and can die on
     * read_dict[ds] (DCS_maker.py:231-247) */
    if (c->partner)
        for (uint64_t g = 0; g < c->n_fam; g++) {
            int32_t p = c->partner[g];
            if (p >= 0 && c->partner[p] != (int32_t)g) bad++;
        }
    return bad;
}

uint64_t cco_count(void *ctx, int what)
{
    cco_ctx *c = (cco_ctx *)ctx;
    switch (what) { case 0: return c->n_fam; case 1: return c->n_sscs; case 2: return c->n_corr;
                    case 3: return c->n_dcs; case 4: return c->n_dcssc; default: return 0; }
}

const void *cco_array(void *ctx, int what)
{
    cco_ctx *c = (cco_ctx *)ctx;
    switch (what) {
    case 0: return c->read_fam; case 1: return c->fam_size; case 2: return c->fam_tmpl;
    case 3: return c->fam_mapq; case 4: return c->fam_flag; case 5: return c->fam_tlen; case 6: return c->fam_rg;
    case 7: return c->partner; case 8: return c->sscs_row; case 9: return c->corr_row;
    case 10: return c->dcs_row; case 11: return c->dcssc_row;
    case 12: return c->sscs_rows; case 13: return c->corr_rows; case 14: return c->dcs_rows; case 15: return c->dcssc_rows;
    case 16: return c->dcs_mapq; case 17: return c->dcs_flag; case 18: return c->dcs_tlen; case 19: return c->dcs_rg;
    case 20: return c->dcssc_mapq; case 21: return c->dcssc_flag; case 22: return c->dcssc_tlen; case 23: return c->dcssc_rg;
    case 24: return c->stats;
    default: return NULL;
    }
}

void cco_destroy(void *ctx)
{
    cco_ctx *c = (cco_ctx *)ctx;
    if (!c) return;
    void *p[] = { c->read_fam, c->fam_size, c->fam_tmpl, c->fam_start, c->members, c->fam_mapq, c->fam_rg,
                  c->fam_flag, c->fam_tlen, c->partner, c->sscs_row, c->corr_row, c->dcs_row, c->dcssc_row,
                  c->sscs_rows, c->corr_rows, c->dcs_rows, c->dcssc_rows, c->dcs_mapq, c->dcs_rg, c->dcssc_mapq,
                  c->dcssc_rg, c->dcs_flag, c->dcssc_flag, c->dcs_tlen, c->dcssc_tlen, c->table };
    for (unsigned i = 0; i < sizeof(p) / sizeof(p[0]); i++) free(p[i]);
    free(c);
}

/* Stand-alone primitives for the known-answer vectors (one family / one pair). */
int cco_consensus_maker(const uint8_t *rows, uint32_t n, uint32_t lpad, uint32_t stride, double cutoff, uint8_t *out)
{
    cco_ctx c; memset(&c, 0, sizeof(c));
    uint32_t start[2] = { 0, n }, size = n;
    uint32_t *mem = (uint32_t *)malloc((n + 1) * sizeof(uint32_t));
    if (!mem) return CCO_ENOMEM;
    for (uint32_t i = 0; i < n; i++) mem[i] = i;
    c.payload = rows; c.lpad = lpad; c.stride = stride; c.cutoff = cutoff; c.q_sscs = 30; c.q_cap = 60;
    c.fam_size = &size; c.fam_start = start; c.members = mem;
    int rc = vote(&c, 0, out);
    free(mem);
    return rc;
}

void cco_duplex_consensus(const uint8_t *r1, const uint8_t *r2, uint32_t lpad, uint32_t stride, int sc, uint8_t *out)
{
    cco_ctx c; memset(&c, 0, sizeof(c));
    c.lpad = lpad; c.stride = stride; c.q_cap = 60;
    agree(&c, r1, r2, sc ? 29 : -1, out);
}
