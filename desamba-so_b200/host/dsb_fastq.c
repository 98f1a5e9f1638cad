/* dsb_fastq.c -- FASTQ / FASTA record reader with the semantics of the reference's kseq
 * (reference src/lib/utils.c:841-977): the header starts at the next '>' or '@' byte, the name
 * ends at the first isspace() byte, sequence lines are concatenated until a line starts with
 * '>', '+' or '@', the quality is read until it is at least as long as the sequence.
 * Input is a gzip-transparent stream (file) or a memory block (the API's in-memory mode, which
 * the reference round-trips through tmpfile()+gzdopen(), src/cly_mt.c:1048-1062).
 */
#include "dsb_host.h"
#include <ctype.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <zlib.h>
#include <pthread.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>

#define FQ_BUFSIZE (1 << 20)

struct dsb_fq {
    gzFile gz;                 /* stream mode */
    FILE *tmp;                 /* gzip-in-memory goes through a temp file like the reference */
    const unsigned char *mem;  /* memory mode: `buf` IS the caller's block (no copy), begin/end span it */
    uint64_t mem_n;
    void *map;                 /* memory mode over a mapped plain file (dsb_fq_open_path) */
    size_t map_n;
    unsigned char *buf;
    size_t begin, end;
    int is_eof;
    int last_char;
    /* records found ahead of time by dsb_fq_prescan (memory mode) */
    dsb_rec_view *pre;
    size_t pre_n, pre_i;
    int scan_threads;          /* > 0: the block is pre-scanned piece by piece, the next piece when `pre` runs out */
    size_t scan_piece;
    int scan_pieces;           /* pieces scanned so far (tests) */
    /* the reference's read batches (read_reads, src/cly_mt.c:22-43: 5000 records or 10 Mbases), only as far as they decide
     * what happens at a record kseq rejects: see dsb_fq_next_view */
    uint64_t rb_count, rb_bases;
    int rb_empty;
    /* pending record when the caller's buffer was too small */
    int have_pending;
    dsb_strbuf p_name, p_seq, p_qual;
    int p_is_fastq;
    int64_t p_ret;
};

static size_t fq_fill(dsb_fq *f)
{
    if (!f->gz) {               /* memory mode has everything in view from the start */
        f->is_eof = 1;
        return 0;
    }
    f->begin = 0;
    int n = gzread(f->gz, f->buf, FQ_BUFSIZE);
    f->end = n < 0 ? 0 : (size_t)n;
    if (f->end < FQ_BUFSIZE) f->is_eof = 1;
    return f->end;
}

static inline int fq_getc(dsb_fq *f)
{
    if (f->is_eof && f->begin >= f->end) return -1;
    if (f->begin >= f->end) {
        if (fq_fill(f) == 0) return -1;
    }
    return (int)f->buf[f->begin++];
}

static void sb_need(dsb_strbuf *s, size_t extra)
{
    if (s->m - s->l < extra + 1) {
        size_t m = s->l + extra + 1;
        m = m + (m >> 1) + 64;
        s->s = (char *)dsb_xrealloc(s->s, m);
        s->m = m;
    }
}

/* delimiter: 0 = isspace, otherwise the byte.  returns length or -1 at end of input */
static int64_t fq_getuntil(dsb_fq *f, int delimiter, dsb_strbuf *str, int *dret, int append)
{
    if (dret) *dret = 0;
    if (!append) str->l = 0;
    if (f->begin >= f->end && f->is_eof) return -1;
    for (;;) {
        size_t i;
        if (f->begin >= f->end) {
            if (!f->is_eof) {
                if (fq_fill(f) == 0) break;
            } else
                break;
        }
        if (delimiter) {
            unsigned char *p = (unsigned char *)memchr(f->buf + f->begin, delimiter, (size_t)(f->end - f->begin));
            i = p ? (size_t)(p - f->buf) : f->end;
        } else {
            for (i = f->begin; i < f->end; ++i)
                if (isspace(f->buf[i])) break;
        }
        sb_need(str, (size_t)(i - f->begin));
        memcpy(str->s + str->l, f->buf + f->begin, (size_t)(i - f->begin));
        str->l += (size_t)(i - f->begin);
        f->begin = i + 1;
        if (i < f->end) {
            if (dret) *dret = f->buf[i];
            break;
        }
    }
    sb_need(str, 0);
    str->s[str->l] = 0;
    return (int64_t)str->l;
}

static dsb_fq *fq_new(int stream)
{
    dsb_fq *f = (dsb_fq *)dsb_xcalloc(1, sizeof *f);
    if (stream) f->buf = (unsigned char *)dsb_xmalloc(FQ_BUFSIZE);
    return f;
}

dsb_fq *dsb_fq_open_path(const char *path)
{
    /* A plain (not gzip) regular file is mapped and read in memory mode: zero-copy record views and the threaded
     * pre-scan instead of a single-threaded stream reader (1.2-1.6 Gbases/s, less than one GPU classifies).  Records
     * are the same either way (tests/test_host.py).  DSB_NO_MMAP=1 keeps the stream reader. */
    if (!getenv("DSB_NO_MMAP")) {
        int fd = open(path, O_RDONLY);
        struct stat st;
        if (fd >= 0 && fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size >= 2) {
            unsigned char magic[2] = {0, 0};
            if (pread(fd, magic, 2, 0) == 2 && !(magic[0] == 0x1f && magic[1] == 0x8b)) {
                void *m = mmap(NULL, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
                if (m != MAP_FAILED) {
                    close(fd);
                    madvise(m, (size_t)st.st_size, MADV_WILLNEED);
                    dsb_fq *f = dsb_fq_open_mem((const char *)m, (uint64_t)st.st_size);
                    f->map = m;
                    f->map_n = (size_t)st.st_size;
                    return f;
                }
            }
        }
        if (fd >= 0) close(fd);
    }
    dsb_fq *f = fq_new(1);
    f->gz = gzopen(path, "r");
    if (!f->gz) dsb_fatal("cannot open reads file %s", path);
    gzbuffer(f->gz, 1 << 20);
    return f;
}

dsb_fq *dsb_fq_open_mem(const char *buf, uint64_t n)
{
    const int gz = (n >= 2 && (unsigned char)buf[0] == 0x1f && (unsigned char)buf[1] == 0x8b);
    dsb_fq *f = fq_new(gz);
    if (gz) {
        f->tmp = tmpfile();
        if (!f->tmp) dsb_fatal("tmpfile() failed");
        if (fwrite(buf, 1, n, f->tmp) != n) dsb_fatal("short write to temp file");
        rewind(f->tmp);
        f->gz = gzdopen(dup(fileno(f->tmp)), "r");
        if (!f->gz) dsb_fatal("gzdopen failed");
    } else {
        f->mem = (const unsigned char *)buf;
        f->mem_n = n;
        f->buf = (unsigned char *)buf;     /* never written: every access below only reads buf */
        f->begin = 0;
        f->end = (size_t)n;
        f->is_eof = 1;
    }
    return f;
}

void dsb_fq_close(dsb_fq *f)
{
    if (!f) return;
    if (f->gz) gzclose(f->gz);
    if (f->tmp) fclose(f->tmp);
    if (!f->mem) free(f->buf);
    if (f->map) munmap(f->map, f->map_n);
    free(f->pre);
    free(f->p_name.s);
    free(f->p_seq.s);
    free(f->p_qual.s);
    free(f);
}

/* kseq_read (src/lib/utils.c:939-977) into the pending buffers */
static int64_t fq_read_record(dsb_fq *f)
{
    int c;
    f->p_is_fastq = 0;
    if (f->last_char == 0) {
        while ((c = fq_getc(f)) != -1 && c != '>' && c != '@') {}
        if (c == -1) return -1;
        f->last_char = c;
    }
    f->p_seq.l = f->p_qual.l = 0;
    if (fq_getuntil(f, 0, &f->p_name, &c, 0) < 0) return -1;
    if (c != '\n') {
        /* comment: skipped (the reference stores it but never prints it) */
        dsb_strbuf junk = {NULL, 0, 0};
        fq_getuntil(f, '\n', &junk, NULL, 0);
        free(junk.s);
    }
    while ((c = fq_getc(f)) != -1 && c != '>' && c != '+' && c != '@') {
        sb_need(&f->p_seq, 1);
        f->p_seq.s[f->p_seq.l++] = (char)c;
        fq_getuntil(f, '\n', &f->p_seq, NULL, 1);
    }
    if (c == '>' || c == '@') f->last_char = c;
    sb_need(&f->p_seq, 0);
    f->p_seq.s[f->p_seq.l] = 0;
    if (c != '+') return (int64_t)f->p_seq.l;         /* FASTA */
    f->p_is_fastq = 1;
    while ((c = fq_getc(f)) != -1 && c != '\n') {}
    if (c == -1) return -2;
    while (fq_getuntil(f, '\n', &f->p_qual, NULL, 1) >= 0 && f->p_qual.l < f->p_seq.l) {}
    f->last_char = 0;
    if (f->p_seq.l != f->p_qual.l) return -2;
    return (int64_t)f->p_seq.l;
}

int64_t dsb_fq_next(dsb_fq *f, dsb_strbuf *name, char *seq_out, uint64_t cap, dsb_strbuf *qual, int *is_fastq)
{
    if (!f->have_pending) {
        f->p_ret = fq_read_record(f);
        if (f->p_ret < 0) return f->p_ret;
        f->have_pending = 1;
    }
    if ((uint64_t)f->p_seq.l > cap) return -3;
    f->have_pending = 0;
    name->l = 0;
    sb_need(name, f->p_name.l);
    memcpy(name->s, f->p_name.s, f->p_name.l + 1);
    name->l = f->p_name.l;
    memcpy(seq_out, f->p_seq.s, f->p_seq.l);
    if (qual) {
        qual->l = 0;
        sb_need(qual, f->p_qual.l);
        if (f->p_is_fastq) memcpy(qual->s, f->p_qual.s, f->p_qual.l);
        qual->l = f->p_is_fastq ? f->p_qual.l : 0;
        qual->s[qual->l] = 0;
    }
    if (is_fastq) *is_fastq = f->p_is_fastq;
    return f->p_ret;
}

/* ---- zero-copy record views ---------------------------------------------------------------
 * In memory mode the common record shape -- header line, ONE sequence line, '+' line, ONE quality
 * line of the same length -- is recognised directly in the caller's block and returned as pointers
 * into it; anything else (FASTA, wrapped lines, short qualities, end-of-input oddities) goes through
 * fq_read_record() above, i.e. the kseq restatement, from the unchanged reader state.  The fast
 * path accepts a record only where kseq_read would produce exactly the same name / seq / qual. */
static int fq_try_fast(dsb_fq *f, dsb_rec_view *v)
{
    const unsigned char *b = f->buf;
    size_t p = f->begin;
    const size_t e = f->end;
    if (f->last_char == 0) {
        while (p < e && b[p] != '>' && b[p] != '@') p++;
        if (p >= e) return 0;
        p++;
    }
    const size_t ns = p;
    while (p < e && !isspace(b[p])) p++;
    if (p >= e) return 0;
    const size_t nl = p - ns;
    const unsigned char *q;
    if (b[p] != '\n') {
        q = (const unsigned char *)memchr(b + p, '\n', e - p);
        if (!q) return 0;
        p = (size_t)(q - b);
    }
    p++;
    if (p >= e) return 0;
    if (b[p] == '>' || b[p] == '+' || b[p] == '@' || b[p] == '\n') return 0;
    const size_t ss = p;
    q = (const unsigned char *)memchr(b + p, '\n', e - p);
    if (!q) return 0;
    const size_t sl = (size_t)(q - b) - ss;
    p = (size_t)(q - b) + 1;
    if (p >= e || b[p] != '+') return 0;
    q = (const unsigned char *)memchr(b + p, '\n', e - p);
    if (!q) return 0;
    p = (size_t)(q - b) + 1;
    const size_t qs = p;
    size_t ql;
    q = (p < e) ? (const unsigned char *)memchr(b + p, '\n', e - p) : NULL;
    if (q) { ql = (size_t)(q - b) - qs; p = (size_t)(q - b) + 1; }
    else { ql = e - qs; p = e; }
    if (ql != sl) return 0;
    v->name = (const char *)b + ns; v->name_l = nl;
    v->seq = (const char *)b + ss; v->seq_l = sl;
    v->qual = (const char *)b + qs; v->is_fastq = 1;
    v->stable = 1;
    f->begin = p;
    f->last_char = 0;
    return 1;
}

/* What the reference does at a record kseq_read rejects (-2: quality string of another length, or none): read_reads ends
 * its batch there (`if (rst < 0) break;`, src/cly_mt.c:36-37), the records read so far are classified, and the next batch
 * carries on BEHIND the rejected record.  If the batch was still empty (the rejected record is the first of a reference
 * batch: the first record overall, the one after 5000 records or 10 Mbases, or the one right after another rejected
 * record), step 0 of the pipeline reports "no reads" and THAT pipeline worker ends (src/cly_mt.c:367,
 * src/lib/kthread.c:136-141); the other workers keep reading, so the run only ends with the third empty batch
 * (PIPELINE_T_NUM = 3, src/cly_mt.c:461).  So: a rejected record is dropped and reading continues, until the third one
 * that opens a batch. */
#define REF_PIPELINE_WORKERS 3
#define REF_BATCH_RECORDS 5000u
#define REF_BATCH_BASES 10000000u
static void ref_batch_note(dsb_fq *f, uint64_t seq_l)
{
    if (f->rb_count >= REF_BATCH_RECORDS || f->rb_bases >= REF_BATCH_BASES) f->rb_count = f->rb_bases = 0;
    f->rb_count++;
    f->rb_bases += seq_l;
}

static void fq_prescan_piece(dsb_fq *f);

int64_t dsb_fq_next_view(dsb_fq *f, dsb_rec_view *v)
{
    if (f->have_pending) {              /* a record fetched by dsb_fq_next() that did not fit its buffer */
        f->have_pending = 0;
    } else {
        if (f->pre_i >= f->pre_n && f->scan_threads > 0) fq_prescan_piece(f);
        if (f->pre_i < f->pre_n) { *v = f->pre[f->pre_i++]; ref_batch_note(f, v->seq_l); return (int64_t)v->seq_l; }
        if (f->mem && fq_try_fast(f, v)) { ref_batch_note(f, v->seq_l); return (int64_t)v->seq_l; }
        for (;;) {
            f->p_ret = fq_read_record(f);
            if (f->p_ret != -2) break;
            if (f->rb_count >= REF_BATCH_RECORDS || f->rb_bases >= REF_BATCH_BASES) f->rb_count = f->rb_bases = 0;
            if (f->rb_count == 0 && ++f->rb_empty >= REF_PIPELINE_WORKERS) return -2;   /* the last pipeline worker ends */
            f->rb_count = f->rb_bases = 0;              /* the batch ends; the next one starts behind the rejected record */
        }
        if (f->p_ret < 0) return f->p_ret;
    }
    ref_batch_note(f, f->p_seq.l);
    v->name = f->p_name.s; v->name_l = f->p_name.l;
    v->seq = f->p_seq.s; v->seq_l = f->p_seq.l;
    v->qual = f->p_is_fastq ? f->p_qual.s : NULL;
    v->is_fastq = f->p_is_fastq;
    v->stable = 0;
    return f->p_ret;
}


/* ---- parallel pre-scan of an in-memory block --------------------------------------------------
 * The block is cut into n_threads pieces.  Each piece starts at a GUESSED record start: an '@' at the beginning
 * of a line whose third line starts with '+'.  Every thread runs the fast-path recogniser from its start until it
 * reaches the next thread's start.  The guesses are then verified left to right: thread t's records are accepted
 * only if thread t-1 ended EXACTLY where thread t began, i.e. the sequential parser would have arrived at that
 * byte in the same state.  Whatever is not accepted is parsed by the ordinary sequential path afterwards, so the
 * result is the sequential result by construction. */
typedef struct scan_job {
    const dsb_fq *f;
    size_t start, limit, end;
    int ok;                         /* ended by reaching `limit` (not by an unrecognised record) */
    dsb_rec_view *v;
    size_t n, m;
} scan_job;

static void *scan_worker(void *arg)
{
    scan_job *j = (scan_job *)arg;
    dsb_fq tmp = *j->f;             /* private cursor over the shared, read-only block */
    tmp.begin = j->start;
    tmp.last_char = 0;
    j->ok = 0;
    for (;;) {
        size_t q = tmp.begin;
        const unsigned char *b = tmp.buf;
        while (q < tmp.end && b[q] != '>' && b[q] != '@') q++;
        if (q >= j->limit) { j->end = q; j->ok = 1; break; }
        tmp.begin = q;
        dsb_rec_view v;
        if (!fq_try_fast(&tmp, &v)) { j->end = q; break; }
        if (j->n == j->m) {
            j->m = j->m ? j->m * 2 : 4096;
            j->v = (dsb_rec_view *)dsb_xrealloc(j->v, j->m * sizeof(dsb_rec_view));
        }
        j->v[j->n++] = v;
    }
    return NULL;
}

/* The block is scanned PIECE BY PIECE (scan_piece bytes at a time, all threads on one piece): the first batch can be staged
 * and launched after the first piece instead of after the whole text (53 ms for 8 GB on 32 cores, with every GPU idle), and
 * the record loop, which is far ahead of the devices anyway, scans the next piece when it runs out of records. */
void dsb_fq_prescan(dsb_fq *f, int n_threads)
{
    if (n_threads > 32) n_threads = 32;
    f->scan_threads = n_threads;
    f->scan_piece = (size_t)n_threads * ((size_t)32 << 20);
    if (f->scan_piece < ((size_t)64 << 20)) f->scan_piece = (size_t)64 << 20;
    const char *pe = getenv("DSB_SCAN_PIECE");                        /* tests: pieces of a few MB */
    if (pe && atoll(pe) > 0) f->scan_piece = (size_t)atoll(pe);
    fq_prescan_piece(f);
}

static void fq_prescan_piece(dsb_fq *f)
{
    const int n_threads = f->scan_threads;
    f->scan_threads = 0;                                                /* set again below if there is more to scan */
    if (!f->mem || f->have_pending || f->last_char != 0 || f->pre_i < f->pre_n) return;
    const size_t lo = f->begin;
    const size_t hi = (f->end - lo > f->scan_piece + (f->scan_piece >> 2)) ? lo + f->scan_piece : f->end;
    if (n_threads < 1 || hi - lo < ((size_t)1 << 20)) return;          /* small inputs: not worth the threads */
    scan_job jobs[32];
    pthread_t th[32];
    const unsigned char *b = f->buf;
    int n = 0;
    size_t prev = lo;
    for (int t = 0; t < n_threads; t++) {
        size_t s = lo + (hi - lo) / (size_t)n_threads * (size_t)t;
        if (t > 0) {
            if (s < prev) s = prev;
            /* next line start holding '@' whose line + 2 starts with '+' */
            size_t found = hi;
            for (size_t p = s; p + 1 < hi;) {
                const unsigned char *nl = (const unsigned char *)memchr(b + p, '\n', hi - p);
                if (!nl) break;
                const size_t h = (size_t)(nl - b) + 1;
                if (h >= hi) break;
                if (b[h] == '@') {
                    const unsigned char *e1 = (const unsigned char *)memchr(b + h, '\n', hi - h);
                    const unsigned char *e2 = e1 ? (const unsigned char *)memchr(e1 + 1, '\n', hi - (size_t)(e1 + 1 - b)) : NULL;
                    if (e2 && (size_t)(e2 + 1 - b) < hi && e2[1] == '+') { found = h; break; }
                }
                p = h;
            }
            if (found >= hi) break;
            s = found;
            if (s <= prev && n > 0) continue;
        }
        memset(&jobs[n], 0, sizeof jobs[n]);
        jobs[n].f = f;
        jobs[n].start = s;
        prev = s + 1;
        n++;
    }
    for (int t = 0; t < n; t++) jobs[t].limit = (t + 1 < n) ? jobs[t + 1].start : hi;
    int started[32] = {0};
    for (int t = 1; t < n; t++) started[t] = pthread_create(&th[t], NULL, scan_worker, &jobs[t]) == 0;
    scan_worker(&jobs[0]);
    for (int t = 1; t < n; t++) {
        if (started[t]) pthread_join(th[t], NULL);
        else jobs[t].ok = -1;                       /* no thread: its piece (and what follows) goes to the sequential parser */
    }
    f->scan_pieces++;
    /* accept the verified prefix */
    size_t total = 0;
    int k = 0;
    for (; k < n; k++) {
        if (jobs[k].ok == -1) break;
        total += jobs[k].n;
        if (!(jobs[k].ok == 1 && k + 1 < n && jobs[k].end == jobs[k + 1].start)) { k++; break; }
    }
    free(f->pre);
    f->pre = (dsb_rec_view *)dsb_xmalloc((total ? total : 1) * sizeof(dsb_rec_view));
    f->pre_n = f->pre_i = 0;
    for (int t = 0; t < k; t++) {
        memcpy(f->pre + f->pre_n, jobs[t].v, jobs[t].n * sizeof(dsb_rec_view));
        f->pre_n += jobs[t].n;
    }
    if (k > 0) { f->begin = jobs[k - 1].end; f->last_char = 0; }
    /* everything up to the piece's end was recognised: the next piece is scanned when these records have been handed out;
     * otherwise the sequential parser takes over from f->begin for the rest of the input, as it always did */
    if (k == n && n > 0 && jobs[n - 1].ok == 1 && f->begin < f->end) f->scan_threads = n_threads;
    for (int t = 0; t < n; t++) free(jobs[t].v);
}

int dsb_fq_scan_pieces(const dsb_fq *f) { return f->scan_pieces; }

/* Input bytes not yet handed out (memory mode: exact; stream mode: unknown = UINT64_MAX).  The batcher sizes its
 * last batches with it. */
uint64_t dsb_fq_remaining_bytes(const dsb_fq *f)
{
    if (f->gz || !f->mem) return UINT64_MAX;
    if (f->pre && f->pre_i < f->pre_n) {
        const unsigned char *next = (const unsigned char *)f->pre[f->pre_i].name;
        if (next >= f->mem && next <= f->mem + f->mem_n) return (uint64_t)(f->mem + f->mem_n - next);
    }
    return f->end > f->begin ? (uint64_t)(f->end - f->begin) : 0;
}

