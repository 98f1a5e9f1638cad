/* dsb_batch.c -- CUDA-stream read batcher: replaces the reference's pthread pool and 3-stage
 * pipeline (reference src/cly_mt.c:29-43 read_reads, :352-381 worker_for / classify_pipeline,
 * src/lib/kthread.c kt_for / kt_pipeline).
 *
 * Threads of one dsb_session_run():
 *   producer (the caller's thread): record loop over the pre-scanned text (kseq semantics), history bytes and the
 *            2G / 3G class of every read in read order (dsb_history.c), descriptors into the slot's pinned staging;
 *   submit : gathers the bases of the batch into the pinned block (helper pool), dsb_slot_submit = async H2D, five
 *            kernels, D2H on the slot's stream;
 *   drain  : dsb_slot_wait in submission order, formats the record fields (helper pool) -- pass 1 of the output;
 *   splice : (SAM_FULL written straight into the caller's block) moves fields + SEQ + QUAL to their final place -- pass 2,
 *            pipelined behind the drain thread over two banks of field text;
 *   prefault: up to four threads touch the pages of the caller's output block ahead of the writers.
 *
 * Every GPU holds a replica of the index and SLOTS_PER_GPU slots; batches go to the devices round-robin, results are
 * collected in submission order, so the output order is the input order and no device-to-device communication exists
 * (SURVEY.md 8e).  Batch boundaries do not change results: the history-dependent bytes are computed on
 * the host, in read order, before a read is handed to any device; the one value that depends on RESULTS (max_read_l,
 * reference src/cly.c:2884-2885,2953) is resolved on the host too, and the producer waits for the batches in flight
 * only when the class of a short read hangs on them.
 */
#include "dsb_host.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

/* Bulk copies of the pipeline (bases into the pinned staging block, SEQ and QUAL into the caller's output block) write
 * memory nobody reads again soon: the GPU's copy engine and the caller come much later, when a 33 MB batch has long left
 * the caches.  Streaming stores skip the read-for-ownership of every destination line (1/3 of such a copy's memory traffic)
 * and leave the caches to the text being parsed.  From four GPUs on the host's memory bandwidth is the bound of the whole
 * call.  DSB_NT_COPY=0 goes back to memcpy. */
static int g_nt_copy = -1;
/* returns non-zero when the copied bytes hold a NUL (the batcher's line records are only good for texts without one) */
static inline int copy_stream(char *dst, const char *src, size_t n)
{
    int nul = 0;
#if defined(__SSE2__)
    if (g_nt_copy && n >= 256) {
        const size_t head = (size_t)(-(uintptr_t)dst & 15u);
        if (head) { nul |= memchr(src, 0, head) != NULL; memcpy(dst, src, head); dst += head; src += head; n -= head; }
        size_t k = n >> 6;
        const __m128i zero = _mm_setzero_si128();
        __m128i acc = zero;
        while (k--) {
            const __m128i a = _mm_loadu_si128((const __m128i *)src), b = _mm_loadu_si128((const __m128i *)(src + 16));
            const __m128i c = _mm_loadu_si128((const __m128i *)(src + 32)), d = _mm_loadu_si128((const __m128i *)(src + 48));
            _mm_stream_si128((__m128i *)dst, a); _mm_stream_si128((__m128i *)(dst + 16), b);
            _mm_stream_si128((__m128i *)(dst + 32), c); _mm_stream_si128((__m128i *)(dst + 48), d);
            acc = _mm_or_si128(acc, _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(a, zero), _mm_cmpeq_epi8(b, zero)),
                                                 _mm_or_si128(_mm_cmpeq_epi8(c, zero), _mm_cmpeq_epi8(d, zero))));
            src += 64; dst += 64;
        }
        nul |= _mm_movemask_epi8(acc) != 0;
        n &= 63;
    }
#endif
    if (n) { nul |= memchr(src, 0, n) != NULL; memcpy(dst, src, n); }
    return nul;
}
static inline void copy_stream_done(void)
{
#if defined(__SSE2__)
    _mm_sfence();          /* streaming stores become visible before the pool reports the part as done */
#endif
}

#define SLOTS_PER_GPU 4          /* default; DSB_SLOTS overrides (1..8) */
#define MAX_FMT_THREADS 32

uint32_t dsb_history_max(const dsb_history *h);
int dsb_history_class(const dsb_history *h, uint32_t len);
void dsb_history_note_pending(dsb_history *h, uint32_t len);
void dsb_history_note_result(dsb_history *h, uint32_t len, int had_chain);
void dsb_history_all_drained(dsb_history *h);

struct dsb_engine {
    dsb_loaded_index *ix;
    int n_gpu;
    dsb_index **dev;
    uint64_t batch_bases;
    uint64_t batch_first;      /* size of the first batch of a run (DSB_BATCH_FIRST) */
    int batch_taper;           /* shrink the last batches when the input size is known (DSB_BATCH_TAPER, default 0) */
    uint64_t batch_small;      /* batch size of a call whose input is known to be small (a few batches in all) */
    uint32_t batch_reads;
    int fmt_threads;
    int scan_threads;          /* FASTQ pre-scan: nothing else runs yet, so it may use every core of the rank's share */
    uint32_t *meta_tid;        /* per reference: the taxon meta_analysis reads from its name */
    int meta_names_ok;         /* every reference name has the usual "<name>|<taxid>..." shape */
};

typedef struct batch_meta {
    uint32_t n_reads;
    uint64_t n_bases;
    dsb_strbuf names;       /* NUL-separated */
    uint64_t *name_off;
    /* quality strings: a pointer into the caller's input block when the record view is stable
     * (zero copy), else an offset into `quals`; both "none" for FASTA records */
    const char **qual_ptr;
    uint64_t *qual_off;
    dsb_strbuf quals;
} batch_meta;

/* fork-join helpers for the text formatter (the serial step 2 of the reference's pipeline,
 * src/cly_mt.c:333-348, is the slowest host stage once the classifier runs on the GPU) */
typedef struct fmt_pool {
    pthread_t th[MAX_FMT_THREADS];
    int n;
    pthread_mutex_t mu;
    pthread_cond_t cv_go, cv_done;
    uint64_t gen;
    int pending, quit;
    void (*fn)(void *arg, int part, int n_parts);
    void *arg;
} fmt_pool;

typedef struct fmt_worker_arg { fmt_pool *p; int id; } fmt_worker_arg;

struct dsb_session {
    dsb_engine *e;
    int n_slots;
    dsb_slot **slots;
    batch_meta *meta;
    dsb_history *hist;
    fmt_pool pool;                  /* formatter helpers (used by the drain thread) */
    fmt_worker_arg pool_arg[MAX_FMT_THREADS];
    fmt_pool spool;                 /* splice helpers (used by the splice thread: SAM_FULL text into the caller's block) */
    fmt_worker_arg spool_arg[MAX_FMT_THREADS];
    fmt_pool cpool;                 /* staging-copy helpers (used by the producer) */
    fmt_worker_arg cpool_arg[MAX_FMT_THREADS];
    const char ***copy_src;         /* [slot][read]: where the read's bases still have to be copied from (NULL: already in staging) */
    /* two banks: while the splice thread moves batch b's text, the drain thread formats the fields of batch b + 1 */
    dsb_strbuf fmt_buf[2][MAX_FMT_THREADS + 1];
    size_t *fmt_ins[2], *fmt_end[2];   /* per read of the batch being formatted (SAM_FULL, direct mode): splice point / end of its fields */
    size_t *fmt_seqpos[2];             /* per read: where its SEQ went, relative to the batch's first byte */
    /* line records of the SAM_FULL text of the last direct call, for a meta_analysis() on that very text */
    dsb_meta_line *ml;
    size_t ml_n, ml_m;
    int ml_ok;                         /* every line of this call's text is of the shape the records describe */
    const char *ml_text;
    uint64_t ml_text_n, ml_guard;
    /* a record that did not fit the previous batch */
    int have_carry;
    dsb_rec_view carry;
    /* the carried read's filter class hangs on batches still on the device: drain them before the next record loop */
    int need_drain;
    uint64_t n_class_stalls;
};

static uint64_t env_u64(const char *name, uint64_t dflt)
{
    const char *v = getenv(name);
    if (!v || !*v) return dflt;
    return strtoull(v, NULL, 10);
}

dsb_engine *dsb_engine_create(dsb_loaded_index *ix, int n_gpu_wanted)
{
    dsb_engine *e = (dsb_engine *)dsb_xcalloc(1, sizeof *e);
    e->ix = ix;
    int have = dsb_gpu_count();
    if (have <= 0) dsb_fatal("no CUDA device available (%s); this build has no CPU path", dsb_last_error());
    int want = n_gpu_wanted > 0 ? n_gpu_wanted : (int)env_u64("DSB_GPUS", (uint64_t)have);
    if (want > have) want = have;
    e->n_gpu = want;
    e->dev = (dsb_index **)dsb_xcalloc((size_t)want, sizeof(dsb_index *));
    /* DSB_GPU_FIRST: first device ordinal (one-process-per-GPU launchers give each rank its own) */
    int first = (int)env_u64("DSB_GPU_FIRST", 0);
    if (first + want > have) first = 0;
    /* the first GPU gets the index from host memory; the others copy it from the first GPU's HBM (DSB_INDEX_CLONE=0:
     * every GPU uploads from the host, for timing comparisons) */
    const int clone = (int)env_u64("DSB_INDEX_CLONE", 1);
    for (int g = 0; g < want; g++) {
        const double t0 = dsb_now();
        if (g == 0 || !clone) {
            if (dsb_index_upload(&ix->h, first + g, &e->dev[g]) != 0)
                dsb_fatal("index upload to GPU %d failed: %s", first + g, dsb_last_error());
        } else if (dsb_index_clone(e->dev[0], first + g, &e->dev[g]) != 0)
            dsb_fatal("index copy GPU %d -> GPU %d failed: %s", first, first + g, dsb_last_error());
        if (getenv("DSB_VERBOSE"))
            fprintf(stderr, "[dsb] index on GPU %d in %.1f ms (%s, %.1f MB)\n", first + g, 1e3 * (dsb_now() - t0),
                    (g == 0 || !clone) ? "host upload + k_fm_reblock" : "peer copy", dsb_index_device_bytes(e->dev[g]) / 1e6);
    }
    /* Slot capacity = the batch size of long inputs.  Measured on the 198 Mbase / 10 kb-read set: 32 Mbase batches 70 ms,
     * 16 Mbase 80 ms, 64 Mbase 71 ms per call (a batch needs more reads than the device has warp slots to keep them
     * busy); a 50 Mbase call is fastest with three 16 Mbase batches (DSB_BATCH_SMALL). */
    e->batch_bases = env_u64("DSB_BATCH_BASES", (uint64_t)32 << 20);
    e->batch_small = env_u64("DSB_BATCH_SMALL", (uint64_t)16 << 20);
    if (e->batch_small > e->batch_bases) e->batch_small = e->batch_bases;
    e->batch_reads = (uint32_t)env_u64("DSB_BATCH_READS", 65536);
    /* short calls start with a half-size batch (8, 16, 16, ... Mbases): the device starts ~1 ms earlier.  Same-box sweeps
     * (tools/e2e_sweep.py, 50 Mbase call): first batch 16 Mbases 17.8-18.5 ms, 8 Mbases 17.0-17.4, 4 Mbases ~17.8, 2 Mbases
     * 18.1-18.5; long inputs start with full batches */
    e->batch_first = env_u64("DSB_BATCH_FIRST", (uint64_t)8 << 20);
    if (e->batch_first > e->batch_bases || e->batch_first == 0) e->batch_first = e->batch_bases;
    e->batch_taper = (int)env_u64("DSB_BATCH_TAPER", 0);
    long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
    if (ncpu < 1) ncpu = 1;
    /* one-process-per-GPU launchers (torchrun sets LOCAL_WORLD_SIZE) share the host cores between the ranks */
    long share = (long)env_u64("LOCAL_WORLD_SIZE", 1);
    if (share > 1) ncpu = ncpu / share > 0 ? ncpu / share : 1;
    /* formatter / copy helpers: all but a few cores (producer, submit and drain threads, the CUDA driver's own) up to 12:
     * 16 cores: 8 -> 12 helpers took 1.4 ms off a 22 ms call */
    if (g_nt_copy < 0) g_nt_copy = env_u64("DSB_NT_COPY", 1) != 0;
    e->meta_tid = (uint32_t *)dsb_xcalloc((size_t)ix->h.n_ref + 1, sizeof(uint32_t));
    e->meta_names_ok = 1;
    for (uint64_t i = 0; i < ix->h.n_ref; i++)
        if (!memchr(ix->ref_info[i].ref_name, 0, sizeof ix->ref_info[i].ref_name) ||
            !dsb_meta_rname_tid(ix->ref_info[i].ref_name, &e->meta_tid[i])) e->meta_names_ok = 0;
    /* one process driving four or more GPUs: the splice stage alone has to move ~4 bytes per base of every device's output
     * (12 helpers: 2.9 ms per 33.5 Mbase batch = 11.7 Gbases/s, the bound of the 4-GPU trace); up to 24 helpers there */
    const long cap = e->n_gpu >= 4 ? 24 : 12;
    e->fmt_threads = (int)env_u64("DSB_FORMAT_THREADS", (uint64_t)(ncpu > 12 ? (ncpu - 4 > cap ? cap : ncpu - 4) : (ncpu > 8 ? 8 : ncpu)));
    e->scan_threads = (int)(ncpu > 32 ? 32 : ncpu);
    if (e->fmt_threads < 1) e->fmt_threads = 1;
    if (e->fmt_threads > MAX_FMT_THREADS) e->fmt_threads = MAX_FMT_THREADS;
    return e;
}

void dsb_engine_destroy(dsb_engine *e)
{
    if (!e) return;
    for (int g = 0; g < e->n_gpu; g++) dsb_index_free(e->dev[g]);
    free(e->dev);
    free(e->meta_tid);
    free(e);
}

int dsb_engine_n_gpu(const dsb_engine *e) { return e->n_gpu; }

/* ---- formatter pool ------------------------------------------------------------------- */
static void *fmt_worker(void *argp)
{
    fmt_worker_arg *wa = (fmt_worker_arg *)argp;
    fmt_pool *p = wa->p;
    uint64_t seen = 0;
    for (;;) {
        pthread_mutex_lock(&p->mu);
        while (!p->quit && p->gen == seen) pthread_cond_wait(&p->cv_go, &p->mu);
        if (p->quit) { pthread_mutex_unlock(&p->mu); return NULL; }
        seen = p->gen;
        pthread_mutex_unlock(&p->mu);
        p->fn(p->arg, wa->id, p->n + 1);
        pthread_mutex_lock(&p->mu);
        if (--p->pending == 0) pthread_cond_signal(&p->cv_done);
        pthread_mutex_unlock(&p->mu);
    }
}

/* run fn(arg, part, n_parts) for part = 0..n_parts-1; the caller does the last part itself */
static void pool_run(fmt_pool *p, void (*fn)(void *, int, int), void *arg)
{
    if (p->n > 0) {
        pthread_mutex_lock(&p->mu);
        p->fn = fn;
        p->arg = arg;
        p->pending = p->n;
        p->gen++;
        pthread_cond_broadcast(&p->cv_go);
        pthread_mutex_unlock(&p->mu);
    }
    fn(arg, p->n, p->n + 1);
    if (p->n > 0) {
        pthread_mutex_lock(&p->mu);
        while (p->pending) pthread_cond_wait(&p->cv_done, &p->mu);
        pthread_mutex_unlock(&p->mu);
    }
}

static void pool_start(fmt_pool *p, fmt_worker_arg *args, int n_workers)
{
    pthread_mutex_init(&p->mu, NULL);
    pthread_cond_init(&p->cv_go, NULL);
    pthread_cond_init(&p->cv_done, NULL);
    p->n = n_workers;
    for (int i = 0; i < p->n; i++) {
        args[i].p = p;
        args[i].id = i;
        if (pthread_create(&p->th[i], NULL, fmt_worker, &args[i]) != 0) { p->n = i; break; }
    }
}

static void pool_stop(fmt_pool *p)
{
    pthread_mutex_lock(&p->mu);
    p->quit = 1;
    pthread_cond_broadcast(&p->cv_go);
    pthread_mutex_unlock(&p->mu);
    for (int i = 0; i < p->n; i++) pthread_join(p->th[i], NULL);
}

dsb_session *dsb_session_create(dsb_engine *e)
{
    dsb_session *s = (dsb_session *)dsb_xcalloc(1, sizeof *s);
    s->e = e;
    int per_gpu = (int)env_u64("DSB_SLOTS", SLOTS_PER_GPU);
    if (per_gpu < 1) per_gpu = 1;
    if (per_gpu > 8) per_gpu = 8;
    s->n_slots = e->n_gpu * per_gpu;
    s->slots = (dsb_slot **)dsb_xcalloc((size_t)s->n_slots, sizeof(dsb_slot *));
    s->meta = (batch_meta *)dsb_xcalloc((size_t)s->n_slots, sizeof(batch_meta));
    for (int k = 0; k < s->n_slots; k++) {
        /* slot k lives on GPU k % n_gpu, so consecutive batches go to different devices */
        if (dsb_slot_create(e->dev[k % e->n_gpu], e->batch_bases, e->batch_reads, &s->slots[k]) != 0)
            dsb_fatal("slot creation failed: %s", dsb_last_error());
        s->meta[k].name_off = (uint64_t *)dsb_xmalloc((size_t)e->batch_reads * 8);
        s->meta[k].qual_ptr = (const char **)dsb_xmalloc((size_t)e->batch_reads * sizeof(char *));
        s->meta[k].qual_off = (uint64_t *)dsb_xmalloc((size_t)e->batch_reads * 8);
    }
    s->hist = dsb_history_new();
    for (int b = 0; b < 2; b++) {
        s->fmt_ins[b] = (size_t *)dsb_xmalloc((size_t)e->batch_reads * sizeof(size_t));
        s->fmt_seqpos[b] = (size_t *)dsb_xmalloc((size_t)e->batch_reads * sizeof(size_t));
        s->fmt_end[b] = (size_t *)dsb_xmalloc((size_t)e->batch_reads * sizeof(size_t));
    }
    pool_start(&s->pool, s->pool_arg, e->fmt_threads - 1);
    pool_start(&s->spool, s->spool_arg, e->fmt_threads - 1);
    {
        const int ccap = e->n_gpu >= 4 ? 12 : 6;       /* the submit thread's gather: 2.0 ms per batch with 6, next in line after the splice */
        pool_start(&s->cpool, s->cpool_arg, (e->fmt_threads > ccap ? ccap : e->fmt_threads) - 1);
    }
    s->copy_src = (const char ***)dsb_xcalloc((size_t)s->n_slots, sizeof(char **));
    for (int k = 0; k < s->n_slots; k++) s->copy_src[k] = (const char **)dsb_xmalloc((size_t)e->batch_reads * sizeof(char *));
    return s;
}

/* the reference rebuilds a thread_id's buffers when the caller changes thread_num (src/cly_mt.c:1286-1295): what earlier
 * reads left in the worker buffer is gone, max_read_l starts again */
void dsb_session_reset_history(dsb_session *s)
{
    dsb_history_free(s->hist);
    s->hist = dsb_history_new();
}

void dsb_session_destroy(dsb_session *s)
{
    if (!s) return;
    pool_stop(&s->pool);
    pool_stop(&s->spool);
    pool_stop(&s->cpool);
    for (int k = 0; k < s->n_slots; k++) free(s->copy_src[k]);
    free(s->copy_src);
    free(s->ml);
    for (int k = 0; k < s->n_slots; k++) {
        dsb_slot_free(s->slots[k]);
        free(s->meta[k].names.s);
        free(s->meta[k].quals.s);
        free(s->meta[k].name_off);
        free(s->meta[k].qual_ptr);
        free(s->meta[k].qual_off);
    }
    for (int b = 0; b < 2; b++) {
        for (int i = 0; i <= MAX_FMT_THREADS; i++) free(s->fmt_buf[b][i].s);
        free(s->fmt_ins[b]);
        free(s->fmt_seqpos[b]);
        free(s->fmt_end[b]);
    }
    free(s->slots);
    free(s->meta);
    dsb_history_free(s->hist);
    free(s);
}

typedef struct copy_job { const char **src; char *seq; const dsb_read_desc *desc; uint32_t n; uint64_t bases; } copy_job;

static void copy_seq_part(void *arg, int part, int n_parts)
{
    const copy_job *j = (const copy_job *)arg;
    const uint64_t lo = j->bases * (uint64_t)part / (uint64_t)n_parts, hi = j->bases * (uint64_t)(part + 1) / (uint64_t)n_parts;
    /* reads whose first base falls into [lo, hi): found by bisection on the ascending offsets */
    uint32_t a = 0, b = j->n;
    while (a < b) { const uint32_t mid = (a + b) / 2; if (j->desc[mid].seq_off < lo) a = mid + 1; else b = mid; }
    for (uint32_t i = a; i < j->n && j->desc[i].seq_off < hi; i++)
        if (j->src[i]) copy_stream(j->seq + j->desc[i].seq_off, j->src[i], j->desc[i].len);
    copy_stream_done();
}

/* Record loop of one batch: descriptors, history bytes, names, quality views of up to one batch go into slot k's staging,
 * in read order.  The bases of stable (zero-copy) records are NOT moved here: the submit thread copies them with the copy
 * pool and launches the batch while this thread already runs the record loop of the next one.  Returns the number of
 * reads (0 = input exhausted). */
static uint32_t fill_slot(dsb_session *s, int k, dsb_fq *f, int want_qual, int *eof, uint64_t limit_bases)
{
    dsb_engine *e = s->e;
    batch_meta *m = &s->meta[k];
    char *seq;
    dsb_read_desc *desc;
    uint8_t *edge;
    if (dsb_slot_staging(s->slots[k], &seq, &desc, &edge) != 0) dsb_fatal("staging: %s", dsb_last_error());
    m->names.l = 0;
    m->quals.l = 0;
    uint32_t n = 0;
    uint64_t bases = 0;
    while (n < e->batch_reads) {
        dsb_rec_view v;
        if (s->have_carry) {
            v = s->carry;
            s->have_carry = 0;
        } else {
            int64_t r = dsb_fq_next_view(f, &v);
            if (r < 0) { *eof = 1; break; }
        }
        const uint64_t L = v.seq_l;
        /* the read's 9-mer table on the device holds 24-bit positions; the reference's own table stops at 1 Mi nodes per
         * strand, i.e. reads of about 0.9 Mbases (src/cly_mt.c:537-538, src/cly.c:2179-2182) */
        if (L >= (1u << 24)) dsb_fatal("a read of %llu bases: reads of 2^24 bases or more are not supported", (unsigned long long)L);
        if (n > 0 && bases + L > limit_bases && L <= e->batch_bases) {   /* this batch is meant to be a small one */
            s->carry = v;
            s->have_carry = 1;
            break;
        }
        if (L > e->batch_bases - bases) {
            if (n == 0) dsb_fatal("a single read is longer than the batch capacity (%llu bases); raise DSB_BATCH_BASES",
                                  (unsigned long long)e->batch_bases);
            s->carry = v;                      /* stays valid: no further reader call before the next batch */
            s->have_carry = 1;
            break;
        }
        /* 2G / 3G filter class (src/cly.c:2953-2981): known now, or only once the reads before this one are back */
        const int cls = L >= 40 ? dsb_history_class(s->hist, (uint32_t)L) : 0;
        if (cls == 2) {
            s->carry = v;
            s->have_carry = 1;
            s->need_drain = 1;
            s->n_class_stalls++;
            break;
        }
        /* a stable view's bases stay where they are until the copy helpers move the whole batch at once; the
         * history model only needs to READ them, from wherever they live */
        const char *src = v.seq;
        if (v.stable) s->copy_src[k][n] = v.seq;
        else { memcpy(seq + bases, v.seq, L); s->copy_src[k][n] = NULL; src = seq + bases; }
        desc[n].seq_off = bases;
        desc[n].len = (uint32_t)L;
        desc[n].flags = dsb_history_next(s->hist, src, (uint32_t)L, edge + (size_t)n * (DSB_HEAD_BYTES + DSB_TAIL_BYTES)) |
                        (cls == 1 ? DSB_READ_2G : 0u);
        dsb_history_note_pending(s->hist, (uint32_t)L);
        m->name_off[n] = m->names.l;
        dsb_strbuf_reserve(&m->names, v.name_l + 1);
        memcpy(m->names.s + m->names.l, v.name, v.name_l);
        m->names.s[m->names.l + v.name_l] = 0;
        /* lines the text parser of meta_analysis would read differently (empty fields, a header-like first line, a NUL):
         * no line records for this call's text */
        if (v.name_l == 0 || v.name[0] == '@' || L == 0 || memchr(v.name, 0, v.name_l)) __atomic_store_n(&s->ml_ok, 0, __ATOMIC_RELAXED);
        m->names.l += v.name_l + 1;
        m->qual_ptr[n] = NULL;
        m->qual_off[n] = UINT64_MAX;
        if (want_qual && v.is_fastq) {
            if (v.stable)
                m->qual_ptr[n] = v.qual;
            else {
                dsb_strbuf_reserve(&m->quals, L + 1);
                memcpy(m->quals.s + m->quals.l, v.qual, L);
                m->qual_off[n] = m->quals.l;
                m->quals.l += L;
            }
        }
        bases += L;
        n++;
    }
    m->n_reads = n;
    m->n_bases = bases;
    return n;
}

static void add_counters(dsb_counters *a, const dsb_counters *b)
{
    uint64_t *x = (uint64_t *)a;
    const uint64_t *y = (const uint64_t *)b;
    for (size_t i = 0; i < sizeof(dsb_counters) / 8; i++) x[i] += y[i];
}

typedef struct fmt_job {
    dsb_session *s;
    const batch_meta *m;
    const dsb_out_opt *o;
    const char *seq;
    const dsb_read_desc *desc;
    const dsb_read_result *res;
    const dsb_hit *hits;
    uint32_t cut[MAX_FMT_THREADS + 2];     /* part p formats reads [cut[p], cut[p+1]) */
    char *dst;                             /* direct mode: where part p's text goes = dst + off[p] */
    size_t dst_off;                        /* dst - start of the caller's block */
    int nul[MAX_FMT_THREADS + 2];          /* part p copied a NUL byte */
    size_t off[MAX_FMT_THREADS + 2];
    size_t len[MAX_FMT_THREADS + 2];       /* SAM_FULL direct mode: bytes part p will write */
    int bank;                              /* which fmt_buf / fmt_ins / fmt_end bank holds this batch's fields */
    int parts;
} fmt_job;

static void fmt_part(void *arg, int part, int n_parts)
{
    (void)n_parts;
    fmt_job *j = (fmt_job *)arg;
    const batch_meta *m = j->m;
    dsb_strbuf *text = &j->s->fmt_buf[j->bank][part];
    text->l = 0;
    for (uint32_t i = j->cut[part]; i < j->cut[part + 1]; i++) {
        const char *q = m->qual_ptr[i];
        if (!q && m->qual_off[i] != UINT64_MAX) q = m->quals.s + m->qual_off[i];
        dsb_format_read(text, j->s->e->ix, j->o, m->names.s + m->name_off[i], j->seq + j->desc[i].seq_off, q, j->desc[i].len,
                        &j->res[i], j->hits + j->res[i].hit_off);
    }
}

/* SAM_FULL in direct mode.  SEQ and QUAL are 99 % of a record and already sit in memory (pinned staging, caller's input):
 * pass 1 formats only the record FIELDS and thereby learns every record's exact size, pass 2 writes fields, SEQ, QUAL
 * straight into the caller's output block -- the bases and qualities are moved once, not twice. */
static void fmt_fields_part(void *arg, int part, int n_parts)
{
    (void)n_parts;
    fmt_job *j = (fmt_job *)arg;
    const batch_meta *m = j->m;
    dsb_strbuf *text = &j->s->fmt_buf[j->bank][part];
    text->l = 0;
    size_t total = 0;
    for (uint32_t i = j->cut[part]; i < j->cut[part + 1]; i++) {
        const int has_q = m->qual_ptr[i] || m->qual_off[i] != UINT64_MAX;
        dsb_format_sam_full_fields(text, j->s->e->ix, j->o, m->names.s + m->name_off[i], j->desc[i].len, &j->res[i],
                                   j->hits + j->res[i].hit_off, &j->s->fmt_ins[j->bank][i]);
        j->s->fmt_end[j->bank][i] = text->l;
        total += (size_t)j->desc[i].len + 1 + (has_q ? (size_t)j->desc[i].len : 6);
    }
    j->len[part] = text->l + total;
}

static void splice_part(void *arg, int part, int n_parts)
{
    (void)n_parts;
    fmt_job *j = (fmt_job *)arg;
    const batch_meta *m = j->m;
    const char *text = j->s->fmt_buf[j->bank][part].s;
    char *d = j->dst + j->off[part];
    size_t from = 0;
    int nul = 0;
    for (uint32_t i = j->cut[part]; i < j->cut[part + 1]; i++) {
        const size_t ins = j->s->fmt_ins[j->bank][i], end = j->s->fmt_end[j->bank][i];
        const uint32_t L = j->desc[i].len;
        const char *q = m->qual_ptr[i];
        if (!q && m->qual_off[i] != UINT64_MAX) q = m->quals.s + m->qual_off[i];
        memcpy(d, text + from, ins - from); d += ins - from;
        j->s->fmt_seqpos[j->bank][i] = (size_t)(d - j->dst);
        nul |= copy_stream(d, j->seq + j->desc[i].seq_off, L); d += L;
        *d++ = '\t';
        if (q) { nul |= copy_stream(d, q, L); d += L; }
        else { memcpy(d, "(null)", 6); d += 6; }
        memcpy(d, text + ins, end - ins); d += end - ins;
        from = end;
    }
    j->nul[part] = nul;
    copy_stream_done();
}

/* The lines this batch added to the text, as meta_analysis would read them back (the same walk as format_sam, dsb_sam.c).
 * Runs on the splice thread, batches in order. */
static void meta_lines_append(dsb_session *s, const fmt_job *j)
{
    if (!__atomic_load_n(&s->ml_ok, __ATOMIC_RELAXED)) return;
    for (int p = 0; p < j->parts; p++)
        if (j->nul[p]) { __atomic_store_n(&s->ml_ok, 0, __ATOMIC_RELAXED); return; }
    const uint32_t *mt = s->e->meta_tid;
    const uint32_t n = j->m->n_reads;
    if (s->ml_n + (size_t)n * 8 > s->ml_m) {
        s->ml_m = (s->ml_n + (size_t)n * 8) * 2;
        s->ml = (dsb_meta_line *)dsb_xrealloc(s->ml, s->ml_m * sizeof(dsb_meta_line));
    }
    for (uint32_t i = 0; i < n; i++) {
        const dsb_read_result *r = &j->res[i];
        const dsb_hit *h = j->hits + r->hit_off;
        if (s->ml_n + (size_t)r->n_hit + 1 > s->ml_m) {
            s->ml_m = (s->ml_n + (size_t)r->n_hit + 1) * 2;
            s->ml = (dsb_meta_line *)dsb_xrealloc(s->ml, s->ml_m * sizeof(dsb_meta_line));
        }
        dsb_meta_line *l = &s->ml[s->ml_n++];
        l->read_length = j->desc[i].len;
        l->seq_off = (uint64_t)(j->dst_off + s->fmt_seqpos[j->bank][i]);
        l->has_seq = 1;
        if (r->n_hit == 0) { l->cls = 'U'; l->tid = 0; continue; }
        l->cls = 'C'; l->tid = mt[h[0].ref_ID];
        for (int loop = 0; loop <= 1; loop++)
            for (uint32_t k = 1; k < r->n_hit; k++) {
                const dsb_hit *c = h + k;
                const int show = (loop == 0) ? (c->pri_index == 0) : (c->pri_index > 0 && c->pri_index <= (uint32_t)j->o->max_sec_N);
                if (!show) continue;
                dsb_meta_line *x = &s->ml[s->ml_n++];
                x->cls = 'C'; x->tid = mt[c->ref_ID]; x->read_length = 1; x->seq_off = 0; x->has_seq = 0;
            }
    }
}

static uint64_t meta_guard(const char *text, uint64_t n)
{
    /* 64 windows of 64 bytes spread over the text: a caller that hands back another text of the same size at the same
     * address is told apart (the records are only used for the very text they describe) */
    uint64_t h = 1469598103934665603ull;
    const uint64_t step = n / 64 ? n / 64 : 1;
    for (uint64_t k = 0; k < 64; k++) {
        const uint64_t at = k * step;
        if (at >= n) break;
        const uint64_t len = n - at < 64 ? n - at : 64;
        for (uint64_t i = 0; i < len; i++) h = (h ^ (unsigned char)text[at + i]) * 1099511628211ull;
    }
    return h ^ n;
}

void dsb_session_meta_seal(dsb_session *s, const char *text, uint64_t n)
{
    s->ml_text = NULL;
    if (!__atomic_load_n(&s->ml_ok, __ATOMIC_RELAXED)) return;
    s->ml_text = text;
    s->ml_text_n = n;
    s->ml_guard = meta_guard(text, n);
}

int dsb_session_meta_lines(dsb_session *s, const char *text, uint64_t n, const dsb_meta_line **lines, size_t *n_lines)
{
    if (!s || !s->ml_text || text != s->ml_text || (n != s->ml_text_n && n != s->ml_text_n + 1)) return 0;
    if (meta_guard(text, s->ml_text_n) != s->ml_guard) return 0;
    *lines = s->ml;
    *n_lines = s->ml_n;
    return 1;
}

static void copy_part(void *arg, int part, int n_parts)
{
    (void)n_parts;
    fmt_job *j = (fmt_job *)arg;
    const dsb_strbuf *text = &j->s->fmt_buf[j->bank][part];
    if (text->l) memcpy(j->dst + j->off[part], text->s, text->l);
}

/* ---- the pipeline: producer (caller's thread) parses, a submit thread stages + launches, a drain thread waits for the
 * results and formats the record fields, and -- for SAM_FULL straight into the caller's block -- a splice thread moves
 * fields, SEQ and QUAL to their final place while the drain thread is already on the next batch ---- */
typedef struct run_ctx {
    dsb_session *s;
    const dsb_out_opt *o;
    dsb_sink_fn sink;
    void *ctx;
    dsb_run_stats *stats;
    pthread_mutex_t mu;
    pthread_cond_t cv;
    uint64_t filled, submitted, drained;
    int fill_done, done;
    /* splice stage (direct SAM_FULL mode only) */
    int splice_mode;
    double t_splice;               /* seconds the splice thread worked (under mu) */
    fmt_job sj[2];                  /* bank b & 1 holds batch b between its field pass and its splice */
    uint64_t formatted, spliced;
    int format_done;
    /* direct mode: the caller's output block is fresh memory; helper threads touch its pages in the background */
    char *prefault_lo;
    size_t prefault_n;
    volatile int prefault_stop;
    int prefault_threads;
    pthread_t prefault_th[8];
    double t_start;
    int verbose;
} run_ctx;

typedef struct prefault_arg { run_ctx *r; int id; } prefault_arg;

/* Touch the pages of the output block front to back, PREFAULT_CHUNK at a time, thread t taking every n-th chunk.  The touch
 * is an atomic OR of zero: it allocates the page without changing a byte, so it is safe next to the splice threads that are
 * already writing text into the same block. */
#define PREFAULT_CHUNK ((size_t)2 << 20)
static void *prefault_thread(void *argp)
{
    prefault_arg *a = (prefault_arg *)argp;
    run_ctx *r = a->r;
    const size_t page = 4096, n_chunks = (r->prefault_n + PREFAULT_CHUNK - 1) / PREFAULT_CHUNK;
    for (size_t c = (size_t)a->id; c < n_chunks && !r->prefault_stop; c += (size_t)r->prefault_threads) {
        const size_t lo = c * PREFAULT_CHUNK, hi = lo + PREFAULT_CHUNK < r->prefault_n ? lo + PREFAULT_CHUNK : r->prefault_n;
        for (size_t i = lo; i < hi; i += page) __atomic_fetch_or(r->prefault_lo + i, 0, __ATOMIC_RELAXED);
    }
    return NULL;
}

static void prefault_join(run_ctx *r)
{
    r->prefault_stop = 1;
    for (int i = 0; i < r->prefault_threads; i++) pthread_join(r->prefault_th[i], NULL);
    r->prefault_threads = 0;
}

/* wait for batch b in slot k, format its records; direct SAM_FULL mode hands the second pass to the splice thread */
static void drain_slot(run_ctx *r, uint64_t b, int k)
{
    dsb_session *s = r->s;
    const dsb_out_opt *o = r->o;
    dsb_run_stats *st = r->stats;
    batch_meta *m = &s->meta[k];
    const int bank = (int)(b & 1);
    if (r->splice_mode) {                          /* bank b & 1 is free once batch b - 2 has been spliced */
        pthread_mutex_lock(&r->mu);
        while (b >= r->spliced + 2) pthread_cond_wait(&r->cv, &r->mu);
        pthread_mutex_unlock(&r->mu);
    }
    fmt_job *jp = &r->sj[bank];
    fmt_job local;
    if (!r->splice_mode) jp = &local;
    uint32_t n_hits;
    dsb_counters c;
    double t0 = dsb_now();
    if (dsb_slot_wait(s->slots[k], &jp->res, &jp->hits, &n_hits, &c) != 0) dsb_fatal("GPU batch failed: %s", dsb_last_error());
    double t1 = dsb_now();
    char *seq;
    dsb_read_desc *desc;
    uint8_t *edge;
    dsb_slot_staging(s->slots[k], &seq, &desc, &edge);
    /* the reference worker's max_read_l, in read order: batches are drained in submission order */
    for (uint32_t i = 0; i < m->n_reads; i++)
        if (desc[i].len >= 510) dsb_history_note_result(s->hist, desc[i].len, (jp->res[i].fast_classify & DSB_RES_HAD_CHAIN) != 0);
    jp->s = s; jp->m = m; jp->o = o; jp->seq = seq; jp->desc = desc; jp->bank = bank;
    /* contiguous read ranges of about equal size in bases, one per formatter thread, sunk in order */
    const int parts = s->pool.n + 1;
    jp->parts = parts;
    jp->cut[0] = 0;
    {
        uint32_t i = 0;
        for (int p = 1; p < parts; p++) {
            const uint64_t target = m->n_bases * (uint64_t)p / (uint64_t)parts;
            while (i < m->n_reads && desc[i].seq_off < target) i++;
            jp->cut[p] = i;
        }
        jp->cut[parts] = m->n_reads;
    }
    pool_run(&s->pool, r->splice_mode ? fmt_fields_part : fmt_part, jp);
    const double t_f = dsb_now();
    if (r->splice_mode) {
        dsb_strbuf *out = (dsb_strbuf *)r->ctx;
        size_t total = 0;
        for (int p = 0; p < parts; p++) { jp->off[p] = total; total += jp->len[p]; }
        if (out->l + total + 1 > out->m) {
            /* the block must grow (the size of the input was unknown or the estimate short): nobody may be writing into it */
            prefault_join(r);
            pthread_mutex_lock(&r->mu);
            while (r->spliced < b) pthread_cond_wait(&r->cv, &r->mu);
            pthread_mutex_unlock(&r->mu);
            dsb_strbuf_reserve(out, total);
        }
        jp->dst = out->s + out->l;
        jp->dst_off = out->l;
        out->l += total;
        pthread_mutex_lock(&r->mu);
        r->formatted = b + 1;
        pthread_cond_broadcast(&r->cv);
        pthread_mutex_unlock(&r->mu);
    } else if (r->sink) {
        for (int p = 0; p < parts; p++)
            if (s->fmt_buf[bank][p].l) r->sink(r->ctx, s->fmt_buf[bank][p].s, s->fmt_buf[bank][p].l);
    } else {
        /* direct mode: ctx is the caller's dsb_strbuf; every thread copies its own part */
        dsb_strbuf *out = (dsb_strbuf *)r->ctx;
        size_t total = 0;
        for (int p = 0; p < parts; p++) { jp->off[p] = total; total += s->fmt_buf[bank][p].l; }
        if (out->l + total + 1 > out->m) prefault_join(r);
        dsb_strbuf_reserve(out, total);
        jp->dst = out->s + out->l;
        pool_run(&s->pool, copy_part, jp);
        out->l += total;
    }
    double t2 = dsb_now();
    if (st) {
        float a = 0, bb = 0, d = 0;
        dsb_slot_timing(s->slots[k], &a, &bb, &d);
        st->gpu_ms_total += a; st->gpu_ms_exist += bb; st->gpu_ms_classify += d;
        st->t_gpu_wait += t1 - t0;
        st->t_format += t2 - t1;
        if (r->verbose)
            fprintf(stderr, "[dsb]   slot %d: wait %.1f ms, format %.1f ms, sink %.1f ms | gpu %.2f ms | t: waited from %.2f, results at %.2f, text at %.2f, sunk at %.2f\n",
                    k, 1e3 * (t1 - t0), 1e3 * (t_f - t1), 1e3 * (t2 - t_f), a, 1e3 * (t0 - r->t_start), 1e3 * (t1 - r->t_start), 1e3 * (t_f - r->t_start), 1e3 * (t2 - r->t_start));
        add_counters(&st->counters, &c);
    }
}

static void *drain_thread(void *argp)
{
    run_ctx *r = (run_ctx *)argp;
    for (uint64_t b = 0;; b++) {
        pthread_mutex_lock(&r->mu);
        while (r->submitted <= b && !r->done) pthread_cond_wait(&r->cv, &r->mu);
        const int have = r->submitted > b;
        pthread_mutex_unlock(&r->mu);
        if (!have) break;
        drain_slot(r, b, (int)(b % (uint64_t)r->s->n_slots));
        if (!r->splice_mode) {
            pthread_mutex_lock(&r->mu);
            r->drained = b + 1;
            pthread_cond_broadcast(&r->cv);
            pthread_mutex_unlock(&r->mu);
        }
    }
    pthread_mutex_lock(&r->mu);
    r->format_done = 1;
    pthread_cond_broadcast(&r->cv);
    pthread_mutex_unlock(&r->mu);
    return NULL;
}

/* second pass of the direct SAM_FULL mode: fields, SEQ and QUAL of batch b to their final place in the caller's block; the
 * slot (its pinned staging holds the bases) is free for the producer once this is done */
static void *splice_thread(void *argp)
{
    run_ctx *r = (run_ctx *)argp;
    for (uint64_t b = 0;; b++) {
        pthread_mutex_lock(&r->mu);
        while (r->formatted <= b && !r->format_done) pthread_cond_wait(&r->cv, &r->mu);
        const int have = r->formatted > b;
        pthread_mutex_unlock(&r->mu);
        if (!have) return NULL;
        const double t0 = dsb_now();
        pool_run(&r->s->spool, splice_part, &r->sj[b & 1]);
        meta_lines_append(r->s, &r->sj[b & 1]);
        if (r->verbose)
            fprintf(stderr, "[dsb]   batch %llu spliced %.2f..%.2f ms\n", (unsigned long long)b, 1e3 * (t0 - r->t_start), 1e3 * (dsb_now() - r->t_start));
        const double dt = dsb_now() - t0;
        pthread_mutex_lock(&r->mu);
        r->t_splice += dt;                           /* added to stats->t_format when the run ends (the drain thread owns that field) */
        r->spliced = b + 1;
        r->drained = b + 1;
        pthread_cond_broadcast(&r->cv);
        pthread_mutex_unlock(&r->mu);
    }
}

/* batch b's record loop is done -> copy its bases into the pinned staging (copy pool), launch it, tell the drain thread */
static void *submit_thread(void *argp)
{
    run_ctx *r = (run_ctx *)argp;
    dsb_session *s = r->s;
    for (uint64_t b = 0;; b++) {
        pthread_mutex_lock(&r->mu);
        while (r->filled <= b && !r->fill_done) pthread_cond_wait(&r->cv, &r->mu);
        const int have = r->filled > b;
        pthread_mutex_unlock(&r->mu);
        if (!have) break;
        const int k = (int)(b % (uint64_t)s->n_slots);
        const batch_meta *m = &s->meta[k];
        char *seq;
        dsb_read_desc *desc;
        uint8_t *edge;
        if (dsb_slot_staging(s->slots[k], &seq, &desc, &edge) != 0) dsb_fatal("staging: %s", dsb_last_error());
        const double t0 = dsb_now();
        copy_job cj = {s->copy_src[k], seq, desc, m->n_reads, m->n_bases};
        pool_run(&s->cpool, copy_seq_part, &cj);
        const double t1 = dsb_now();
        if (dsb_slot_submit(s->slots[k], m->n_reads, m->n_bases) != 0) dsb_fatal("submit failed: %s", dsb_last_error());
        if (r->verbose)
            fprintf(stderr, "[dsb]   batch %llu -> slot %d: base copy %.2f..%.2f, submitted at %.2f ms\n", (unsigned long long)b, k,
                    1e3 * (t0 - r->t_start), 1e3 * (t1 - r->t_start), 1e3 * (dsb_now() - r->t_start));
        pthread_mutex_lock(&r->mu);
        r->submitted = b + 1;
        pthread_cond_broadcast(&r->cv);
        pthread_mutex_unlock(&r->mu);
    }
    pthread_mutex_lock(&r->mu);
    r->done = 1;
    pthread_cond_broadcast(&r->cv);
    pthread_mutex_unlock(&r->mu);
    return NULL;
}

void dsb_session_run(dsb_session *s, dsb_fq *f, const dsb_out_opt *o, dsb_sink_fn sink, void *ctx, dsb_run_stats *stats)
{
    dsb_run_stats local;
    if (!stats) stats = &local;
    memset(stats, 0, sizeof *stats);
    const double t_start = dsb_now();
    const int want_qual = (o->format == DSB_OUT_SAM_FULL);
    run_ctx r;
    memset(&r, 0, sizeof r);
    r.s = s; r.o = o; r.sink = sink; r.ctx = ctx; r.stats = stats;
    r.t_start = t_start;
    r.verbose = getenv("DSB_VERBOSE") != NULL;
    pthread_mutex_init(&r.mu, NULL);
    pthread_cond_init(&r.cv, NULL);
    pthread_t th, th_submit, th_splice;
    int eof = 0;
    r.splice_mode = (!sink && o->format == DSB_OUT_SAM_FULL);
    /* line records for meta_analysis: only for the text this call writes itself; DSB_META_TEXT=1 keeps
     * meta_analysis on the text parser */
    s->ml_n = 0;
    s->ml_text = NULL;
    __atomic_store_n(&s->ml_ok, (r.splice_mode && s->e->meta_names_ok && !getenv("DSB_META_TEXT")) ? 1 : 0, __ATOMIC_RELAXED);
    s->have_carry = 0;
    /* a call that is only a few batches long (in-memory input of known size, ~2 bytes per base) uses the small batch
     * size: the device starts earlier and the last batch's text is shorter */
    uint64_t call_batch = s->e->batch_bases;
    {
        const uint64_t bytes = dsb_fq_remaining_bytes(f);
        if (bytes != UINT64_MAX && bytes / 2 <= 4 * s->e->batch_small) call_batch = s->e->batch_small;
    }
    {
        double t0 = dsb_now();
        dsb_fq_prescan(f, s->e->scan_threads);
        stats->t_parse += dsb_now() - t0;
        if (r.verbose) fprintf(stderr, "[dsb]   prescan done at %.2f ms\n", 1e3 * (dsb_now() - t_start));
    }
    prefault_arg pf_arg[8];
    if (!sink && s->e->fmt_threads > 1) {
        /* direct mode: the caller's output block is fresh memory.  A few threads touch its pages front to back in the
         * background, so that neither the first drain waits for 4 GB of page faults (64 ms on the 2-GPU trace of round 2)
         * nor the splice threads take them one by one */
        dsb_strbuf *out = (dsb_strbuf *)ctx;
        if (out->m - out->l >= ((size_t)8 << 20)) {
            r.prefault_lo = out->s + out->l;
            r.prefault_n = out->m - out->l;
            int n = s->e->fmt_threads / 3;
            if (n < 1) n = 1;
            if (n > 4) n = 4;
            for (int i = 0; i < n; i++) {
                pf_arg[i].r = &r;
                pf_arg[i].id = i;
                r.prefault_threads = n;
                if (pthread_create(&r.prefault_th[i], NULL, prefault_thread, &pf_arg[i]) != 0) { r.prefault_threads = i; break; }
            }
        }
    }
    /* the helpers start once the pre-scan (which uses every core) is over */
    if (r.splice_mode && pthread_create(&th_splice, NULL, splice_thread, &r) != 0) dsb_fatal("pthread_create failed");
    if (pthread_create(&th, NULL, drain_thread, &r) != 0) dsb_fatal("pthread_create failed");
    if (pthread_create(&th_submit, NULL, submit_thread, &r) != 0) dsb_fatal("pthread_create failed");
    for (uint64_t b = 0; !eof; b++) {
        const int k = (int)(b % (uint64_t)s->n_slots);
        /* batch b reuses the slot of batch b - n_slots: wait until that one has been formatted */
        pthread_mutex_lock(&r.mu);
        while (b >= r.drained + (uint64_t)s->n_slots) pthread_cond_wait(&r.cv, &r.mu);
        pthread_mutex_unlock(&r.mu);
        if (s->need_drain) {
            /* the next read's filter class depends on results that are still on a device: wait for every batch */
            pthread_mutex_lock(&r.mu);
            while (r.drained < b) pthread_cond_wait(&r.cv, &r.mu);
            pthread_mutex_unlock(&r.mu);
            dsb_history_all_drained(s->hist);
            s->need_drain = 0;
        }
        /* the history borrows sequence pointers from the staging of earlier batches: before any
         * staging is overwritten it must own what it still needs */
        dsb_history_end_batch(s->hist);
        /* Batch-size schedule of a short call: DSB_BATCH_FIRST (default: half a batch) doubles up to the call's batch
         * size; DSB_BATCH_TAPER=1 (off: no gain measured) halves towards the end when the amount of input left is known. */
        uint64_t limit = (call_batch == s->e->batch_small) ? (s->e->batch_first << (b < 16 ? b : 16)) : call_batch;
        if (limit > call_batch || limit == 0) limit = call_batch;
        {
            const uint64_t left = dsb_fq_remaining_bytes(f);
            if (left != UINT64_MAX && s->e->batch_taper) {
                const uint64_t half = left / 4;                 /* ~2 input bytes per base: half of what is left */
                if (half < limit) limit = half > s->e->batch_first ? half : s->e->batch_first;
            }
        }
        double t0 = dsb_now();
        uint32_t n = fill_slot(s, k, f, want_qual, &eof, limit);
        stats->t_parse += dsb_now() - t0;
        if (n == 0) {
            if (s->have_carry && s->need_drain) { b--; continue; }     /* nothing but a read whose filter class must wait */
            break;
        }
        stats->n_reads += n;
        stats->n_bases += s->meta[k].n_bases;
        stats->n_batches++;
        if (r.verbose)
            fprintf(stderr, "[dsb]   batch %llu -> slot %d: %u reads %llu bases, record loop %.2f..%.2f ms\n", (unsigned long long)b, k, n,
                    (unsigned long long)s->meta[k].n_bases, 1e3 * (t0 - t_start), 1e3 * (dsb_now() - t_start));
        pthread_mutex_lock(&r.mu);
        r.filled = b + 1;
        pthread_cond_broadcast(&r.cv);
        pthread_mutex_unlock(&r.mu);
    }
    pthread_mutex_lock(&r.mu);
    r.fill_done = 1;
    pthread_cond_broadcast(&r.cv);
    pthread_mutex_unlock(&r.mu);
    pthread_join(th_submit, NULL);
    pthread_join(th, NULL);
    if (r.splice_mode) pthread_join(th_splice, NULL);
    if (r.stats) r.stats->t_format += r.t_splice;
    prefault_join(&r);
    dsb_history_end_batch(s->hist);
    dsb_history_all_drained(s->hist);
    pthread_mutex_destroy(&r.mu);
    pthread_cond_destroy(&r.cv);
    stats->t_total = dsb_now() - t_start;
    if (getenv("DSB_VERBOSE"))
        fprintf(stderr, "[dsb] %llu reads %llu bases %llu batches: parse %.1f ms, gpu-wait %.1f ms, format %.1f ms, total %.1f ms\n",
                (unsigned long long)stats->n_reads, (unsigned long long)stats->n_bases, (unsigned long long)stats->n_batches,
                1e3 * stats->t_parse, 1e3 * stats->t_gpu_wait, 1e3 * stats->t_format, 1e3 * stats->t_total);
}
