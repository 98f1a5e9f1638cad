/* dsb_batch.c -- CUDA-stream read batcher: replaces the reference's pthread pool and 3-stage
 * pipeline (reference src/cly_mt.c:29-43 read_reads, :352-381 worker_for / classify_pipeline,
 * src/lib/kthread.c kt_for / kt_pipeline).
 *
 *   parse (host, kseq semantics) -> pinned staging of slot k -> dsb_slot_submit (async H2D,
 *   kernels, D2H on the slot's stream) -> while the GPU works, parse the next batch into slot
 *   k+1 -> dsb_slot_wait(k) -> format records in input order -> sink.
 *
 * Every GPU holds a replica of the index and two slots (double buffering); batches go to the
 * devices round-robin, results are collected in submission order, so the output order is the
 * input order and no device-to-device communication exists.  Batch boundaries
 * do not change results: the history-dependent bytes are computed on the host,
 * in read order, before a read is handed to any device (dsb_history.c).
 */
#include "dsb_host.h"
#include <stdlib.h>
#include <string.h>

#define SLOTS_PER_GPU 2

uint32_t dsb_history_max(const dsb_history *h);
void dsb_history_note_max(dsb_history *h, uint32_t len);

struct dsb_engine {
    dsb_loaded_index *ix;
    int n_gpu;
    dsb_index **dev;
    uint64_t batch_bases;
    uint32_t batch_reads;
};

typedef struct batch_meta {
    uint32_t n_reads;
    uint64_t n_bases;
    dsb_strbuf names;       /* NUL-separated */
    dsb_strbuf quals;       /* concatenated, same offsets as seq */
    uint64_t *name_off;
    uint8_t *has_qual;
    int busy;
} batch_meta;

struct dsb_session {
    dsb_engine *e;
    int n_slots;
    dsb_slot **slots;
    batch_meta *meta;
    dsb_history *hist;
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
    for (int g = 0; g < want; g++)
        if (dsb_index_upload(&ix->h, first + g, &e->dev[g]) != 0)
            dsb_fatal("index upload to GPU %d failed: %s", first + g, dsb_last_error());
    e->batch_bases = env_u64("DSB_BATCH_BASES", (uint64_t)32 << 20);
    e->batch_reads = (uint32_t)env_u64("DSB_BATCH_READS", 65536);
    return e;
}

void dsb_engine_destroy(dsb_engine *e)
{
    if (!e) return;
    for (int g = 0; g < e->n_gpu; g++) dsb_index_free(e->dev[g]);
    free(e->dev);
    free(e);
}

int dsb_engine_n_gpu(const dsb_engine *e) { return e->n_gpu; }

dsb_session *dsb_session_create(dsb_engine *e)
{
    dsb_session *s = (dsb_session *)dsb_xcalloc(1, sizeof *s);
    s->e = e;
    s->n_slots = e->n_gpu * SLOTS_PER_GPU;
    s->slots = (dsb_slot **)dsb_xcalloc((size_t)s->n_slots, sizeof(dsb_slot *));
    s->meta = (batch_meta *)dsb_xcalloc((size_t)s->n_slots, sizeof(batch_meta));
    for (int k = 0; k < s->n_slots; k++) {
        /* slot k lives on GPU k % n_gpu, so consecutive batches go to different devices */
        if (dsb_slot_create(e->dev[k % e->n_gpu], e->batch_bases, e->batch_reads, &s->slots[k]) != 0)
            dsb_fatal("slot creation failed: %s", dsb_last_error());
        s->meta[k].name_off = (uint64_t *)dsb_xmalloc((size_t)e->batch_reads * 8);
        s->meta[k].has_qual = (uint8_t *)dsb_xmalloc((size_t)e->batch_reads);
    }
    s->hist = dsb_history_new();
    return s;
}

void dsb_session_destroy(dsb_session *s)
{
    if (!s) return;
    for (int k = 0; k < s->n_slots; k++) {
        dsb_slot_free(s->slots[k]);
        free(s->meta[k].names.s);
        free(s->meta[k].quals.s);
        free(s->meta[k].name_off);
        free(s->meta[k].has_qual);
    }
    free(s->slots);
    free(s->meta);
    dsb_history_free(s->hist);
    free(s);
}

/* parse up to one batch into slot k; returns number of reads (0 = input exhausted) */
static uint32_t fill_slot(dsb_session *s, int k, dsb_fq *f, int want_qual, int *eof)
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
    dsb_strbuf name = {NULL, 0, 0}, qual = {NULL, 0, 0};
    while (n < e->batch_reads) {
        int is_fq = 0;
        int64_t L = dsb_fq_next(f, &name, seq + bases, e->batch_bases - bases, want_qual ? &qual : NULL, &is_fq);
        if (L == -3) {
            if (n == 0) dsb_fatal("a single read is longer than the batch capacity (%llu bases); raise DSB_BATCH_BASES",
                                  (unsigned long long)e->batch_bases);
            break;
        }
        if (L < 0) { *eof = 1; break; }
        if (L > 0x7fffffff) dsb_fatal("read too long");
        desc[n].seq_off = bases;
        desc[n].len = (uint32_t)L;
        desc[n].flags = dsb_history_next(s->hist, seq + bases, (uint32_t)L, edge + (size_t)n * (DSB_HEAD_BYTES + DSB_TAIL_BYTES));
        if (L >= 40) dsb_history_note_max(s->hist, (uint32_t)L);
        m->name_off[n] = m->names.l;
        dsb_strbuf_reserve(&m->names, name.l + 1);
        memcpy(m->names.s + m->names.l, name.s, name.l + 1);
        m->names.l += name.l + 1;
        if (want_qual) {
            dsb_strbuf_reserve(&m->quals, bases + (uint64_t)L + 1 - m->quals.l);
            m->has_qual[n] = (uint8_t)is_fq;
            if (is_fq) memcpy(m->quals.s + bases, qual.s, (size_t)L);
            m->quals.l = bases + (uint64_t)L;
        }
        bases += (uint64_t)L;
        n++;
    }
    free(name.s);
    free(qual.s);
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

static void drain_slot(dsb_session *s, int k, const dsb_out_opt *o, dsb_sink_fn sink, void *ctx, dsb_strbuf *text,
                       dsb_run_stats *st)
{
    batch_meta *m = &s->meta[k];
    const dsb_read_result *res;
    const dsb_hit *hits;
    uint32_t n_hits;
    dsb_counters c;
    double t0 = dsb_now();
    if (dsb_slot_wait(s->slots[k], &res, &hits, &n_hits, &c) != 0) dsb_fatal("GPU batch failed: %s", dsb_last_error());
    double t1 = dsb_now();
    char *seq;
    dsb_read_desc *desc;
    uint8_t *edge;
    dsb_slot_staging(s->slots[k], &seq, &desc, &edge);
    text->l = 0;
    for (uint32_t i = 0; i < m->n_reads; i++) {
        const char *q = NULL;
        if (o->format == DSB_OUT_SAM_FULL && m->has_qual[i]) q = m->quals.s + desc[i].seq_off;
        dsb_format_read(text, s->e->ix, o, m->names.s + m->name_off[i], seq + desc[i].seq_off, q, desc[i].len, &res[i],
                        hits + res[i].hit_off);
        if (text->l > ((size_t)8 << 20)) { sink(ctx, text->s, text->l); text->l = 0; }
    }
    if (text->l) sink(ctx, text->s, text->l);
    double t2 = dsb_now();
    if (st) {
        float a = 0, b = 0, d = 0;
        dsb_slot_timing(s->slots[k], &a, &b, &d);
        st->gpu_ms_total += a; st->gpu_ms_exist += b; st->gpu_ms_classify += d;
        st->t_gpu_wait += t1 - t0;
        st->t_format += t2 - t1;
        add_counters(&st->counters, &c);
    }
    m->busy = 0;
}

void dsb_session_run(dsb_session *s, dsb_fq *f, const dsb_out_opt *o, dsb_sink_fn sink, void *ctx, dsb_run_stats *stats)
{
    dsb_run_stats local;
    if (!stats) stats = &local;
    memset(stats, 0, sizeof *stats);
    const double t_start = dsb_now();
    dsb_strbuf text = {NULL, 0, 0};
    const int want_qual = (o->format == DSB_OUT_SAM_FULL);
    int eof = 0, k = 0;
    while (!eof) {
        if (s->meta[k].busy) drain_slot(s, k, o, sink, ctx, &text, stats);
        /* the slot's sequence staging is about to be overwritten: the history must own what it still needs */
        dsb_history_end_batch(s->hist);
        double t0 = dsb_now();
        uint32_t n = fill_slot(s, k, f, want_qual, &eof);
        stats->t_parse += dsb_now() - t0;
        if (n == 0) break;
        stats->n_reads += n;
        stats->n_bases += s->meta[k].n_bases;
        stats->n_batches++;
        if (dsb_slot_submit(s->slots[k], n, s->meta[k].n_bases) != 0) dsb_fatal("submit failed: %s", dsb_last_error());
        s->meta[k].busy = 1;
        k = (k + 1) % s->n_slots;
    }
    /* drain in submission order: oldest first */
    for (int i = 0; i < s->n_slots; i++) {
        int j = (k + i) % s->n_slots;
        if (s->meta[j].busy) drain_slot(s, j, o, sink, ctx, &text, stats);
    }
    dsb_history_end_batch(s->hist);
    free(text.s);
    stats->t_total = dsb_now() - t_start;
}
