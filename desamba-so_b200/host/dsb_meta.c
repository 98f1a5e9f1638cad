/* dsb_meta.c -- meta_analysis: per-taxon tally of a SAM text (reference src/cly_mt.c:590-670 taxonTree_rank,
 * :687-813 getOneSAM, :902-961 ana_get_tid, :848-900 ana_meta_loop_fprint, :1083-1235 meta_analysis_core,
 * :1329-1414 meta_analysis).  SURVEY.md 8(f) rank 1.  Host-side text processing, no GPU involvement.
 *
 * The reference round-trips the SAM text through three tmpfiles and parses with strtok(); its answer depends on
 * strtok's exact token rules, so the record parser below tokenises a private copy of each line with a small
 * re-entrant tokenizer that has those rules (skip leading delimiters, end the token at the first delimiter and
 * overwrite it with NUL), issued in the same order.  Two consequences that are kept on purpose:
 *   * the AS:i:<n> score is never read (the token taken for it is the literal "AS"), so the per-record score is 0;
 *   * ana_get_tid therefore stops at the first following record: EVERY SAM record, secondary and supplementary
 *     lines included, is tallied as one "read" (weight 1, or the length of its SEQ field -- 1 for "*").
 * Where the reference has undefined behaviour we choose a defined one and say so at the place.
 */
#include "dsb_host.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>
#include <sys/mman.h>

#define NO_TID 0xFFFFFFFFu
#define NAME_MAX_L 200

typedef struct dsb_taxonomy {
    uint32_t max_tid;              /* tid of the LAST line of nodes.dmp + 1 000 000 (src/cly_mt.c:606-612) */
    uint32_t *p_tid;
    char (*rank)[20];
    char (*name)[NAME_MAX_L + 1];
} dsb_taxonomy;

/* strtok() with explicit state */
typedef struct tokz { char *save; } tokz;
static char *tok(tokz *z, char *s, const char *delim)
{
    if (!s) s = z->save;
    if (!s) return NULL;
    s += strspn(s, delim);
    if (*s == 0) { z->save = s; return NULL; }
    char *start = s;
    s += strcspn(s, delim);
    if (*s) { *s = 0; z->save = s + 1; }
    else z->save = s;
    return start;
}

static char *read_line(FILE *fp, char **buf, size_t *cap)
{
    return getline(buf, cap, fp) > 0 ? *buf : NULL;
}

dsb_taxonomy *dsb_taxonomy_load(const char *dir)
{
    char path[4096];
    snprintf(path, sizeof path, "%s/nodes.dmp", dir);
    FILE *fp = fopen(path, "r");
    if (!fp) return NULL;          /* the reference aborts here; we defer the failure to meta_analysis() */
    dsb_taxonomy *t = (dsb_taxonomy *)dsb_xcalloc(1, sizeof *t);
    char *line = NULL;
    size_t cap = 0;
    tokz z;
    uint32_t last = 0;
    while (read_line(fp, &line, &cap)) {
        char *a = tok(&z, line, "\t|");
        if (a) last = (uint32_t)strtoul(a, NULL, 10);
    }
    t->max_tid = last + 1000000u;
    const size_t n = (size_t)t->max_tid + 1;
    t->p_tid = (uint32_t *)dsb_xmalloc(n * sizeof(uint32_t));
    t->rank = (char (*)[20])dsb_xcalloc(n, 20);
    t->name = (char (*)[NAME_MAX_L + 1])dsb_xcalloc(n, NAME_MAX_L + 1);
    for (size_t i = 0; i < n; i++) t->p_tid[i] = NO_TID;
    rewind(fp);
    while (read_line(fp, &line, &cap)) {
        char *a = tok(&z, line, "\t|");
        if (!a) continue;
        const uint32_t tid = (uint32_t)strtoul(a, NULL, 10);
        char *p = tok(&z, NULL, "\t|");
        char *r = tok(&z, NULL, "\t|");
        if (tid > t->max_tid || !p || !r) continue;     /* the reference writes out of bounds / crashes */
        t->p_tid[tid] = (uint32_t)strtoul(p, NULL, 10);
        strncpy(t->rank[tid], r, 19);
    }
    fclose(fp);
    t->p_tid[1] = NO_TID;                               /* root of the taxonomy */
    t->p_tid[0] = NO_TID;                               /* a second one-node tree for unclassified reads */
    strcpy(t->rank[0], "no rank");
    strcpy(t->name[0], "CLY_FAIL");
    snprintf(path, sizeof path, "%s/names.dmp", dir);
    fp = fopen(path, "r");
    if (fp) {
        while (read_line(fp, &line, &cap)) {
            char *a = tok(&z, line, "|\t");
            if (!a) continue;
            const uint32_t tid = (uint32_t)strtoul(a, NULL, 10);
            char *nm = tok(&z, NULL, "\t|");
            tok(&z, NULL, "|");                         /* unique-name column */
            char *cls = tok(&z, NULL, "|");
            if (tid <= t->max_tid && nm && cls && strncmp("\tscien", cls, 6) == 0) {
                strncpy(t->name[tid], nm, NAME_MAX_L);
                t->name[tid][NAME_MAX_L] = 0;
            }
        }
        fclose(fp);
    }
    free(line);
    return t;
}

/* one SAM record as meta_analysis sees it (RST, src/cly_mt.h:55-63) */
typedef struct rec {
    const char *name;
    size_t name_l;
    char cls;                       /* 'C' classified, 'U' unclassified */
    uint32_t tid, read_length, score;
    uint8_t mapq;
    const char *seq;
} rec;

/* getOneSAM: `line` is a private, writable, NUL-terminated copy of one record.  Returns 0, or -1 when the line
 * is too short to hold the fields the reference dereferences (it would crash there). */
static int parse_record(char *line, rec *r)
{
    tokz z;
    char *cur = line, *t;
#define FIELD() do { t = tok(&z, cur, "\t"); if (!t) return -1; cur += strlen(t) + 1; } while (0)
    FIELD(); r->name = t; r->name_l = strlen(t);
    FIELD();                                            /* FLAG */
    r->read_length = 0;
    r->score = 0;
    FIELD();                                            /* RNAME */
    char *seq;
    if (t[0] == '*') {
        r->cls = 'U'; r->tid = 0; r->mapq = 0;
        for (int k = 0; k < 6; k++) FIELD();            /* POS MAPQ CIGAR RNEXT PNEXT TLEN */
        FIELD(); seq = t;
    } else {
        r->cls = 'C';
        char *ref = t;
        FIELD();                                        /* POS */
        FIELD(); r->mapq = (uint8_t)strtoul(t, NULL, 10);
        for (int k = 0; k < 4; k++) FIELD();            /* CIGAR RNEXT PNEXT TLEN */
        FIELD(); seq = t;
        FIELD();                                        /* QUAL */
        t = tok(&z, NULL, ":");
        if (t && ((t[0] == 'A' && t[1] == 'S') || (t[0] == 'N' && t[1] == 'M'))) {
            tok(&z, NULL, ":");
            t = tok(&z, cur, "\t");                     /* restarts at the tag: yields "AS" / "NM", never the number */
            if (t) { cur += strlen(t) + 1; r->score = (uint32_t)strtoul(t, NULL, 10); }
            t = tok(&z, NULL, ":");
            if (t && t[0] == 'm' && t[1] == 's') {      /* unreachable after the restart above; kept for shape */
                tok(&z, NULL, ":");
                t = tok(&z, cur, "\t");
                if (t) { cur += strlen(t) + 1; r->score = (uint32_t)strtoul(t, NULL, 10); }
            }
        }
        /* "tid|<taxid>|..." : the second '|' token, found by stepping over the first token's NUL */
        char *rc = ref;
        tok(&z, ref, "|");
        rc += strlen(rc) + 1;
        t = tok(&z, rc, "|");
        r->tid = t ? (uint32_t)strtoul(t, NULL, 10) : 0;
    }
#undef FIELD
    r->read_length = (uint32_t)strlen(seq);
    r->seq = seq;
    return 0;
}

/* The same record WITHOUT writing into the text, for lines of the usual shape: at least eleven non-empty tab-separated
 * fields in front of the line's '\n' (strtok merges empty fields, so those lines take the tokenising path above), RNAME
 * either starting with '*' or of the form <no '|'>...|<digit>..., MAPQ starting with a digit.  For such a line
 * parse_record yields exactly: name = field 1, tid = the number after RNAME's first '|', mapq = field 5 as uint8,
 * seq = field 10, score = 0 (the AS tag is never read, see the header).  Returns 0 when the line is not of that shape. */
static int parse_record_fast(const char *p, const char *nl, rec *r)
{
    const char *f[11];
    size_t fl[11];
    const char *c = p;
    for (int k = 0; k < 11; k++) {
        const char *t = (const char *)memchr(c, '\t', (size_t)(nl - c));
        if (!t || t == c) return 0;
        f[k] = c; fl[k] = (size_t)(t - c); c = t + 1;
    }
    r->name = f[0]; r->name_l = fl[0];
    r->score = 0;
    r->seq = f[9]; r->read_length = (uint32_t)fl[9];
    if (fl[9] > 0xffffffffu) return 0;
    if (f[2][0] == '*') {
        r->cls = 'U'; r->tid = 0; r->mapq = 0;
        return 1;
    }
    r->cls = 'C';
    {   /* MAPQ: strtoul of a short digit string */
        if (f[4][0] < '0' || f[4][0] > '9' || fl[4] > 9) return 0;
        unsigned long v = 0;
        for (size_t i = 0; i < fl[4] && f[4][i] >= '0' && f[4][i] <= '9'; i++) v = v * 10 + (unsigned long)(f[4][i] - '0');
        r->mapq = (uint8_t)v;
    }
    {   /* "tid|<taxid>|...": strtok(ref, "|") then the token behind its NUL */
        const char *ref = f[2], *re = f[2] + fl[2];
        if (ref[0] == '|') return 0;
        const char *bar = (const char *)memchr(ref, '|', fl[2]);
        if (!bar || bar + 1 >= re || bar[1] < '0' || bar[1] > '9') return 0;
        unsigned long v = 0;
        size_t nd = 0;
        for (const char *q = bar + 1; q < re && *q >= '0' && *q <= '9'; q++, nd++) v = v * 10 + (unsigned long)(*q - '0');
        if (nd > 9) return 0;
        r->tid = (uint32_t)v;
    }
    return 1;
}

/* ---- tally (meta_analysis_core) ------------------------------------------------------------- */
typedef struct count_sort { uint32_t tid; long count; } count_sort;
static int cmp_count_sort(const void *a_, const void *b_)
{
    /* the reference's comparator returns only 0 / 1 (src/cly_mt.c:584-587); with the host libc's qsort --
     * the same one the reference calls -- that is a stable descending sort */
    const count_sort *a = (const count_sort *)a_, *b = (const count_sort *)b_;
    return a->count < b->count;
}

/* zeroed tables over the whole taxonomy id range of which only a few entries are touched: anonymous mappings are
 * zero pages until written (calloc may hand out recycled heap memory and clear all of it: 40 MB per call here) */
static void *lazy_zeros(size_t bytes)
{
    void *p = mmap(NULL, bytes ? bytes : 1, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) dsb_fatal("meta_analysis: cannot map %zu bytes", bytes);
    return p;
}
static void lazy_free(void *p, size_t bytes) { munmap(p, bytes ? bytes : 1); }

static int cmp_u32(const void *a, const void *b)
{
    const uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

typedef struct node { uint64_t weight; uint32_t child_begin; } node;
typedef struct child { uint32_t tid, next; } child;

typedef struct tally {
    const dsb_taxonomy *tax;
    node *tab;
    child *kids;
    dsb_strbuf *out;
    uint64_t total_weight;
} tally;

static void print_leaves(tally *t, uint32_t node_ID)
{
    const dsb_taxonomy *tax = t->tax;
    const node *nd = t->tab + node_ID;
    if (nd->weight == 0) return;
    const float rate = (float)nd->weight / t->total_weight;
    if (nd->child_begin != 0) {
        uint32_t c = nd->child_begin;
        for (;;) {
            print_leaves(t, t->kids[c].tid);
            if (t->kids[c].next == 0) break;
            c = t->kids[c].next;
        }
        return;
    }
    const char *type = "microbe";
    const uint32_t leaf = node_ID;
    if (leaf == 0 || leaf == 1)
        type = "no_match";
    else {
        for (uint32_t n = node_ID; n != NO_TID; n = tax->p_tid[n]) {
            if (n == 9606) { type = "human"; break; }
            if (n == 33208 || n == 33090) { type = "animal_and_plant"; break; }
        }
    }
    dsb_strbuf_printf(t->out, "%s\t%s|%s\tnull\t%f\n", type, tax->name[leaf], tax->rank[leaf], rate);
}

/* ana_get_tid: the taxon of the read whose first record is recs[*i]; leaves *i at the first record that is
 * not consumed (== n at the end of the list) */
static uint32_t read_tid(const dsb_taxonomy *tax, const rec *recs, size_t n, size_t *i)
{
    const rec *first = recs + *i;
    if (first->cls != 'C') { (*i)++; return 0; }
    uint32_t tid = 0, score = 0;
    if (first->tid <= tax->max_tid) { tid = first->tid; score = first->score; }
    for (;;) {
        (*i)++;
        if (*i >= n) return tid;
        const rec *r = recs + *i;
        if (first->name_l != r->name_l || memcmp(first->name, r->name, r->name_l) != 0) break;
        if (score == 0) break;                           /* always: see the header */
        if (r->score != score) continue;
        if (r->tid > tax->max_tid) continue;
        for (uint32_t p = r->tid;;) {                    /* equal score: prefer a descendant of the current pick */
            if (p == tid) { tid = r->tid; break; }
            if (p < 1 || p == NO_TID) break;
            p = tax->p_tid[p];
        }
    }
    return tid;
}

typedef struct meta_row { char *type, *species, *tech; float rate; } meta_row;
static int cmp_meta_row(const void *a, const void *b)
{
    const float x = ((const meta_row *)a)->rate, y = ((const meta_row *)b)->rate;
    return x > y ? -1 : (x < y ? 1 : 0);
}

/* ---- threaded copy + tokenising of the SAM text ------------------------------------------------ */
#define META_MAX_THREADS 16
typedef struct parse_job {
    const char *lo, *hi;                                 /* scan phase: bytes [lo, hi) */
    const char *first_nul;
    const char *beg, *end;                               /* parse phase: whole lines in [beg, end) */
    rec *recs; size_t n_rec, m_rec;
    dsb_strbuf human; uint64_t snap_cap;
    char **priv; size_t n_priv, m_priv;                  /* private copies of the lines that had to be tokenised in place */
    size_t n_slow;
    int bad;
} parse_job;

static int meta_threads(uint64_t input_n)
{
    const char *e = getenv("DSB_META_THREADS");
    long n = e && *e ? atol(e) : 0;
    if (n <= 0) {
        n = sysconf(_SC_NPROCESSORS_ONLN);
        const char *w = getenv("LOCAL_WORLD_SIZE");
        if (w && atol(w) > 1) n /= atol(w);
    }
    if (n > META_MAX_THREADS) n = META_MAX_THREADS;
    if (n < 1 || input_n < ((uint64_t)4 << 20)) n = 1;
    return (int)n;
}

static void run_jobs(pthread_t *th, parse_job *jobs, int n, void *(*fn)(void *))
{
    int started[META_MAX_THREADS] = {0};
    for (int t = 1; t < n; t++) started[t] = pthread_create(&th[t], NULL, fn, &jobs[t]) == 0;
    fn(&jobs[0]);
    for (int t = 1; t < n; t++) {
        if (started[t]) pthread_join(th[t], NULL);
        else fn(&jobs[t]);
    }
}

static void *scan_nul_worker(void *arg)
{
    parse_job *j = (parse_job *)arg;
    j->first_nul = (const char *)memchr(j->lo, 0, (size_t)(j->hi - j->lo));
    return NULL;
}

static void *parse_worker(void *arg)
{
    parse_job *j = (parse_job *)arg;
    const char *p = j->beg, *end = j->end;
    j->m_rec = 1024;
    j->recs = (rec *)dsb_xmalloc(j->m_rec * sizeof(rec));
    dsb_strbuf_reserve(&j->human, 1);
    while (p < end) {
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        const char *next = nl ? nl + 1 : end;
        rec r;
        if (!(nl && parse_record_fast(p, nl, &r))) {
            /* unusual line (or the text's last line without '\n'): the strtok-faithful parser, on a private NUL-terminated
             * copy that lives as long as the records (getline keeps the '\n' in the buffer, so does the copy) */
            const size_t n = (size_t)(next - p);
            char *line = (char *)dsb_xmalloc(n + 1);
            memcpy(line, p, n);
            line[n] = 0;
            if (j->n_priv == j->m_priv) { j->m_priv = j->m_priv ? 2 * j->m_priv : 16; j->priv = (char **)dsb_xrealloc(j->priv, j->m_priv * sizeof(char *)); }
            j->priv[j->n_priv++] = line;
            j->n_slow++;
            if (parse_record(line, &r) != 0) { j->bad = 1; return NULL; }
        }
        if (r.seq[0] != '*' && (r.tid == 9606 || r.tid == 63221 || r.tid == 741158) &&
            (j->human.l < j->snap_cap || j->human.l == 0)) {
            /* only the first max_snapshot_len bases of the whole text are ever returned (but "some human bases" must stay
             * distinguishable from "none": with max_snapshot_len 0 the reference returns an empty block, not NULL) */
            dsb_strbuf_reserve(&j->human, r.read_length);
            memcpy(j->human.s + j->human.l, r.seq, r.read_length);
            j->human.l += r.read_length;
        }
        if (j->n_rec == j->m_rec) { j->m_rec *= 2; j->recs = (rec *)dsb_xrealloc(j->recs, j->m_rec * sizeof(rec)); }
        j->recs[j->n_rec++] = r;
        p = next;
    }
    return NULL;
}

static void free_priv(parse_job *jobs)
{
    for (int k = 0; k < META_MAX_THREADS; k++) {
        for (size_t i = 0; i < jobs[k].n_priv; i++) free(jobs[k].priv[i]);
        free(jobs[k].priv);
    }
}

/* Everything behind the record list: tally per taxon, propagation to the ancestors, report and snapshot.  `recs` and
 * `human.s` are consumed (freed).  Shared by the text path below and by dsb_meta_analysis_lines. */
static void meta_tail(dsb_taxonomy *tax, rec *recs, size_t n_rec, dsb_strbuf human, char **output, uint64_t *output_n, int flag,
                      uint64_t max_snapshot_len, char **human_snapshot, uint64_t *human_snapshot_n)
{
    if (n_rec == 0) {
        /* the reference returns with *output never assigned and then dereferences it; we return an empty answer */
        *output = (char *)dsb_xcalloc(1, 1);
        *output_n = 0;
        *human_snapshot = NULL;
        *human_snapshot_n = 0;
        free(recs); free(human.s);
        return;
    }
    const int tv_ = getenv("DSB_VERBOSE") != NULL;
    const double tt0_ = dsb_now();
    /* tally per taxon */
    const size_t n_tid = (size_t)tax->max_tid + 1;
    /* the tables span the taxonomy's whole id range (millions); calloc hands them out as untouched zero pages, so only
     * the ids that occur are ever written or read: they are listed as they first appear instead of scanning the range */
    uint64_t *node_count = (uint64_t *)lazy_zeros(n_tid * sizeof(uint64_t));
    uint64_t total_weight = 0;
    size_t n_seen = 0, m_seen = 256;
    uint32_t *seen = (uint32_t *)dsb_xmalloc(m_seen * sizeof(uint32_t));
    for (size_t i = 0; i < n_rec;) {
        const uint32_t w = ((flag & 0x1) == 0) ? 1u : recs[i].read_length;
        total_weight += w;
        const uint32_t tid = read_tid(tax, recs, n_rec, &i);
        if (node_count[tid] == 0 && w != 0) {
            if (n_seen == m_seen) { m_seen *= 2; seen = (uint32_t *)dsb_xrealloc(seen, m_seen * sizeof(uint32_t)); }
            seen[n_seen++] = tid;
        }
        node_count[tid] += w;
    }
    const double tt1_ = dsb_now();
    qsort(seen, n_seen, sizeof(uint32_t), cmp_u32);       /* ascending id: the order the range scan produced (ties of the count sort) */
    size_t n_sort = 0;
    count_sort *srt = (count_sort *)dsb_xmalloc((n_seen + 1) * sizeof(count_sort));
    for (size_t i = 0; i < n_seen; i++)
        if (node_count[seen[i]]) { srt[n_sort].tid = seen[i]; srt[n_sort++].count = (long)node_count[seen[i]]; }
    free(seen);
    qsort(srt, n_sort, sizeof(count_sort), cmp_count_sort);
    /* propagate the weights to the ancestors, registering each child once, in order of first appearance */
    tally t;
    t.tax = tax;
    t.tab = (node *)lazy_zeros(n_tid * sizeof(node));
    t.kids = (child *)lazy_zeros((2 * n_tid + 2) * sizeof(child));
    t.total_weight = total_weight;
    uint32_t n_kids = 1;
    for (size_t i = 0; i < n_sort; i++) {
        uint32_t c = srt[i].tid;
        for (;;) {
            const uint32_t par = tax->p_tid[c];
            t.tab[c].weight += node_count[srt[i].tid];
            if (par == NO_TID) break;
            if (par > tax->max_tid) break;               /* a parent outside the table: the reference reads out of bounds */
            if (t.tab[par].child_begin == 0) {
                t.tab[par].child_begin = n_kids;
                t.kids[n_kids++].tid = c;
            } else {
                uint32_t k = t.tab[par].child_begin;
                while (t.kids[k].tid != c && t.kids[k].next != 0) k = t.kids[k].next;
                if (t.kids[k].tid != c) {
                    t.kids[k].next = n_kids;
                    t.kids[n_kids++].tid = c;
                }
            }
            c = par;
        }
    }
    const double tt2_ = dsb_now();
    dsb_strbuf lines = {NULL, 0, 0};
    dsb_strbuf_reserve(&lines, 1);
    lines.s[0] = 0;
    t.out = &lines;
    print_leaves(&t, 0);
    print_leaves(&t, 1);
    const double tt3_ = dsb_now();
    /* ---- meta_analysis proper: snapshot + top-3 table ---------------------------------------- */
    if (human.l == 0) {
        *human_snapshot = NULL;
        *human_snapshot_n = 0;
    } else {
        human.s[human.l] = 0;
        const uint64_t nb = strlen(human.s);
        *human_snapshot_n = nb < max_snapshot_len ? nb : max_snapshot_len;
        *human_snapshot = (char *)dsb_xmalloc(*human_snapshot_n + 1);
        memcpy(*human_snapshot, human.s, *human_snapshot_n);
        (*human_snapshot)[*human_snapshot_n] = 0;
    }
    /* the rates go through their "%f" text (six decimals), exactly as the reference re-reads its own output */
    size_t n_rows = 0, m_rows = 16;
    meta_row *rows = (meta_row *)dsb_xmalloc(m_rows * sizeof(meta_row));
    float no_match_rate = 0;
    tokz zl;
    for (char *ln = tok(&zl, lines.s, "\n"); ln; ln = tok(&zl, NULL, "\n")) {
        tokz zf;
        meta_row r;
        r.type = tok(&zf, ln, "\t");
        r.species = tok(&zf, NULL, "\t");
        r.tech = tok(&zf, NULL, "\t");
        char *rate = tok(&zf, NULL, "\t");
        if (!r.type || !r.species || !r.tech || !rate) continue;
        r.rate = strtof(rate, NULL);
        if (strcmp("no_match", r.type) == 0) { no_match_rate += r.rate; continue; }
        if (n_rows == m_rows) { m_rows *= 2; rows = (meta_row *)dsb_xrealloc(rows, m_rows * sizeof(meta_row)); }
        rows[n_rows++] = r;
    }
    dsb_strbuf res = {NULL, 0, 0};
    dsb_strbuf_reserve(&res, 64);
    res.s[0] = 0;
    if (no_match_rate > 0.95) {
        dsb_strbuf_printf(&res, "no_match\tnull|null\tnull\t0\n");
    } else {
        for (size_t i = 0; i < n_rows; i++) rows[i].rate = rows[i].rate / (1 - no_match_rate);
        qsort(rows, n_rows, sizeof(meta_row), cmp_meta_row);
        for (size_t i = 0; i < n_rows; i++)
            if (i < 3 || (strcmp("human", rows[i].type) == 0 && rows[i].rate > 0.05))
                dsb_strbuf_printf(&res, "%s\t%s\t%s\t%f\n", rows[i].type, rows[i].species, rows[i].tech, rows[i].rate);
    }
    *output = res.s;                                      /* malloc'd, caller frees (desamba.h) */
    *output_n = res.l;
    const double tt4_ = dsb_now();
    free(rows); free(lines.s); free(srt); free(recs); free(human.s);
    lazy_free(t.tab, n_tid * sizeof(node));
    lazy_free(t.kids, (2 * n_tid + 2) * sizeof(child));
    lazy_free(node_count, n_tid * sizeof(uint64_t));
    if (tv_)
        fprintf(stderr, "[dsb]   meta tail: tally %.2f ms, sort + propagate %.2f ms, leaves %.2f ms, report %.2f ms, free %.2f ms\n",
                1e3 * (tt1_ - tt0_), 1e3 * (tt2_ - tt1_), 1e3 * (tt3_ - tt2_), 1e3 * (tt4_ - tt3_), 1e3 * (dsb_now() - tt4_));
}

/* RNAME as parse_record_fast reads it (and parse_record agrees on names of this shape) */
int dsb_meta_rname_tid(const char *ref_name, uint32_t *tid)
{
    const size_t n = strlen(ref_name);
    if (n == 0 || ref_name[0] == '*' || ref_name[0] == '|') return 0;
    for (size_t i = 0; i < n; i++)
        if (ref_name[i] == '\t' || ref_name[i] == '\n') return 0;
    const char *bar = (const char *)memchr(ref_name, '|', n), *re = ref_name + n;
    if (!bar || bar + 1 >= re || bar[1] < '0' || bar[1] > '9') return 0;
    unsigned long v = 0;
    size_t nd = 0;
    for (const char *q = bar + 1; q < re && *q >= '0' && *q <= '9'; q++, nd++) v = v * 10 + (unsigned long)(*q - '0');
    if (nd > 9) return 0;
    *tid = (uint32_t)v;
    return 1;
}

/* meta_analysis from the line records the batcher kept for the text it returned: the records are what
 * parse_worker would produce from `text`, line by line, so everything behind them is shared with the text path. */
void dsb_meta_analysis_lines(dsb_taxonomy *tax, const dsb_meta_line *lines, size_t n_lines, const char *text, char **output,
                             uint64_t *output_n, int flag, uint64_t max_snapshot_len, char **human_snapshot,
                             uint64_t *human_snapshot_n)
{
    if (!tax) dsb_fatal("meta_analysis: the index directory has no nodes.dmp / names.dmp");
    const double t_0 = dsb_now();
    rec *recs = (rec *)dsb_xmalloc((n_lines ? n_lines : 1) * sizeof(rec));
    dsb_strbuf human = {NULL, 0, 0};
    dsb_strbuf_reserve(&human, 1);
    for (size_t i = 0; i < n_lines; i++) {
        const dsb_meta_line *l = lines + i;
        rec r;
        r.name = ""; r.name_l = 0;                      /* read_tid never gets past its `score == 0` test: names are not compared */
        r.cls = (char)l->cls; r.tid = l->tid; r.read_length = l->read_length; r.score = 0; r.mapq = 0;
        r.seq = l->has_seq ? text + l->seq_off : "*";
        /* the taxon first: the text (2 bytes per base, written with streaming stores) is only touched for human reads */
        if ((r.tid == 9606 || r.tid == 63221 || r.tid == 741158) && r.seq[0] != '*' && (human.l < max_snapshot_len || human.l == 0)) {
            dsb_strbuf_reserve(&human, r.read_length);
            memcpy(human.s + human.l, r.seq, r.read_length);
            human.l += r.read_length;
        }
        recs[i] = r;
    }
    const double t_1 = dsb_now();
    meta_tail(tax, recs, n_lines, human, output, output_n, flag, max_snapshot_len, human_snapshot, human_snapshot_n);
    if (getenv("DSB_VERBOSE"))
        fprintf(stderr, "[dsb] meta_analysis: %zu line records kept by read_classify (no text parsed): records %.2f ms, tally + report %.2f ms\n",
                n_lines, 1e3 * (t_1 - t_0), 1e3 * (dsb_now() - t_1));
}

void dsb_meta_analysis_impl(dsb_taxonomy *tax, const char *input, uint64_t input_n, char **output, uint64_t *output_n,
                            int flag, uint64_t max_snapshot_len, char **human_snapshot, uint64_t *human_snapshot_n)
{
    if (input_n == 0) {
        *output_n = 0;
        *human_snapshot_n = 0;
        return;
    }
    if (!tax) dsb_fatal("meta_analysis: the index directory has no nodes.dmp / names.dmp");
    /* The text is read in place by several threads (contiguous pieces cut at line ends; records and human bases are
     * joined in text order, so the result is that of one pass).  Lines of the usual shape are parsed without being
     * modified (parse_record_fast); any other line is copied and tokenised by the strtok-faithful parse_record.  The first
     * version copied the whole text and tokenised it on one thread: 165-210 ms for 104 MB of SAM_FULL, 9 x the
     * classification that produced it. */
    const int verbose = getenv("DSB_VERBOSE") != NULL;
    const double t_0 = dsb_now();
    int n_thr = meta_threads(input_n);
    parse_job jobs[META_MAX_THREADS];
    pthread_t th[META_MAX_THREADS];
    memset(jobs, 0, sizeof jobs);
    for (int t = 0; t < n_thr; t++) {
        jobs[t].lo = input + (size_t)(input_n * (uint64_t)t / (uint64_t)n_thr);
        jobs[t].hi = input + (size_t)(input_n * (uint64_t)(t + 1) / (uint64_t)n_thr);
        jobs[t].snap_cap = max_snapshot_len;
    }
    run_jobs(th, jobs, n_thr, scan_nul_worker);
    const double t_1 = dsb_now();
    const char *end = input + input_n;                   /* like getline on a text file: stops at an embedded NUL */
    for (int t = 0; t < n_thr; t++)
        if (jobs[t].first_nul) { end = jobs[t].first_nul; break; }
    const char *p = input;
    while (p < end && *p == '@') {                       /* skip_sam_head: @SQ / @PG lines */
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        p = nl ? nl + 1 : end;
    }
    /* pieces of the body, each starting at a line start */
    {
        const size_t body = (size_t)(end - p);
        if (body < ((size_t)1 << 20)) n_thr = 1;
        const char *cut = p;
        for (int t = 0; t < n_thr; t++) {
            jobs[t].beg = cut;
            const char *want = p + body * (size_t)(t + 1) / (size_t)n_thr;
            if (t + 1 == n_thr || want >= end) cut = end;
            else {
                if (want < cut) want = cut;
                const char *nl = (const char *)memchr(want, '\n', (size_t)(end - want));
                cut = nl ? nl + 1 : end;
            }
            jobs[t].end = cut;
        }
    }
    run_jobs(th, jobs, n_thr, parse_worker);
    const double t_2 = dsb_now();
    size_t n_rec = 0;
    for (int t = 0; t < n_thr; t++) {
        if (jobs[t].bad) dsb_fatal("meta_analysis: malformed SAM record (the reference crashes here)");
        n_rec += jobs[t].n_rec;
    }
    rec *recs = (rec *)dsb_xmalloc((n_rec ? n_rec : 1) * sizeof(rec));
    dsb_strbuf human = {NULL, 0, 0};
    dsb_strbuf_reserve(&human, 1);
    n_rec = 0;
    for (int t = 0; t < n_thr; t++) {
        memcpy(recs + n_rec, jobs[t].recs, jobs[t].n_rec * sizeof(rec));
        n_rec += jobs[t].n_rec;
        if (jobs[t].human.l) {
            dsb_strbuf_reserve(&human, jobs[t].human.l);
            memcpy(human.s + human.l, jobs[t].human.s, jobs[t].human.l);
            human.l += jobs[t].human.l;
        }
        free(jobs[t].recs);
        free(jobs[t].human.s);
    }
    meta_tail(tax, recs, n_rec, human, output, output_n, flag, max_snapshot_len, human_snapshot, human_snapshot_n);
    const double t_3 = dsb_now();
    size_t n_slow = 0;
    for (int k = 0; k < META_MAX_THREADS; k++) n_slow += jobs[k].n_slow;
    free_priv(jobs);
    if (verbose)
        fprintf(stderr, "[dsb] meta_analysis: %d threads, NUL scan %.1f ms, records %.1f ms, tally + report %.1f ms, free %.1f ms (%zu records, %zu through the tokenising parser)\n",
                n_thr, 1e3 * (t_1 - t_0), 1e3 * (t_2 - t_1), 1e3 * (t_3 - t_2), 1e3 * (dsb_now() - t_3), n_rec, n_slow);
}
