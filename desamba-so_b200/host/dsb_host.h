/* dsb_host.h -- internal interfaces of the C host side of libdesamba.so / deSAMBA classify.
 *
 * The host keeps what the reference keeps on the host (file formats, FASTQ parsing, SAM text,
 * the three API entry points, the CLI) and hands the per-read classifier to the CUDA layer
 * through include/dsb_cuda.h.  Nothing in here classifies a read.
 */
#ifndef DSB_HOST_H
#define DSB_HOST_H
#include <stdint.h>
#include <stdio.h>
#include <stddef.h>
#include "dsb_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* fatal error: message + abort, the reference's only error style (src/lib/utils.h:111) */
void dsb_fatal(const char *fmt, ...) __attribute__((noreturn, format(printf, 1, 2)));
void *dsb_xmalloc(size_t n);
void *dsb_xcalloc(size_t n, size_t sz);
void *dsb_xrealloc(void *p, size_t n);
double dsb_now(void);

/* ---- index files (reference src/idx.c:1103-1160, src/bwt.c:68-104) ---------------------- */
typedef struct dsb_ref_info {      /* on-disk REF_INFO, 144 bytes (src/idx.h:22-26) */
    char ref_name[128];
    uint64_t seq_l;
    uint64_t seq_offset;
} dsb_ref_info;

typedef struct dsb_loaded_index {
    dsb_host_index h;              /* views for the CUDA layer */
    /* owned storage behind the views */
    uint8_t *bwt_occ;
    uint64_t *hash_index;
    uint64_t *sa;
    uint8_t *ek0, *ek1;
    uint64_t *unitig;
    uint64_t *ref_pos;
    uint8_t *ref_bin;
    dsb_ref_info *ref_info;
    uint64_t *ref_off, *ref_len;
    int32_t q_mem[2000];
    int32_t q_lv[400];
    int ek_len;
    /* the big arrays are views into read-only mappings of the index files (no host copy is made): the views of `h`
     * point into these; dsb_index_release_bulk unmaps them */
    struct { void *p; size_t n; } maps[8];
    int n_maps;
    void *prefault;                /* helper threads that touch the mappings' pages ahead of the upload */
} dsb_loaded_index;

/* read the ten deSAMBA.* files of `dir`; min_len/min_score are MAP_opt.L_min_matching / min_score */
void dsb_index_load(dsb_loaded_index *ix, const char *dir, int min_len, int min_score);
/* drop the big host arrays once they are resident on the GPUs (names / offsets are kept) */
void dsb_index_release_bulk(dsb_loaded_index *ix);
void dsb_index_destroy(dsb_loaded_index *ix);
/* ---- meta_analysis on hit records ----------------------------------------
 * One line of the SAM_FULL text read_classify() returned, as meta_analysis reads it (RST, reference src/cly_mt.h:55-63):
 * built by the batcher from the hit records while it writes the text, so that a meta_analysis() call on that very text
 * does not have to parse 2 bytes per base again. */
typedef struct dsb_meta_line {
    uint32_t tid;                  /* taxon of the line's RNAME ("tid|<taxid>|..."), 0 for an unclassified read */
    uint32_t read_length;          /* length of the line's SEQ field (1 for the "*" of SUP / SEC lines) */
    uint64_t seq_off;              /* where SEQ starts in the text (lines that carry the read's bases) */
    uint8_t cls;                   /* 'C' or 'U' */
    uint8_t has_seq;
} dsb_meta_line;
/* taxon of a reference name as the text parser reads it; 0 = the name is not of the usual "<name>|<digits>..." shape
 * (then only the text path is exact) */
int dsb_meta_rname_tid(const char *ref_name, uint32_t *tid);

/* MAPQ tables (reference src/cly_mt.c:396-420) */
void dsb_mapq_tables(int32_t *q_mem, int32_t *q_lv, double p_e, uint64_t l_ref);

/* ---- FASTQ / FASTA input with kseq semantics (reference src/lib/utils.c:841-977) --------- */
typedef struct dsb_fq dsb_fq;
dsb_fq *dsb_fq_open_path(const char *path);                 /* plain or gzip */
dsb_fq *dsb_fq_open_mem(const char *buf, uint64_t n);       /* plain or gzip bytes */
void dsb_fq_close(dsb_fq *f);
uint64_t dsb_fq_remaining_bytes(const dsb_fq *f);           /* UINT64_MAX when unknown (streams) */

typedef struct dsb_strbuf { char *s; size_t l, m; } dsb_strbuf;
/* next record: >= 0 sequence length, -1 end of input, -2 malformed quality (kseq_read codes).
 * seq bytes are appended to seq_out (cap bytes available at seq_out) when they fit; the record
 * is kept pending and -3 returned when they do not (call again with a bigger/empty buffer). */
int64_t dsb_fq_next(dsb_fq *f, dsb_strbuf *name, char *seq_out, uint64_t cap, dsb_strbuf *qual, int *is_fastq);

/* A record as pointers (not NUL-terminated).  stable = 1: they point into the caller's memory block
 * and stay valid for the life of the reader; stable = 0: into the reader's buffers, valid until
 * the next call.  qual is NULL for FASTA records. */
typedef struct dsb_rec_view {
    const char *name, *seq, *qual;
    size_t name_l, seq_l;
    int is_fastq, stable;
} dsb_rec_view;
/* next record as a view: >= 0 sequence length, -1 end of input, -2 malformed quality */
int64_t dsb_fq_next_view(dsb_fq *f, dsb_rec_view *v);
/* memory mode: find the records of the whole block with n_threads threads ahead of time (same records, same order
 * as the sequential reader; see dsb_fastq.c); a no-op for streams */
void dsb_fq_prescan(dsb_fq *f, int n_threads);
int dsb_fq_scan_pieces(const dsb_fq *f);     /* pieces the pre-scan has gone through so far */

/* ---- worker-buffer history ------------------------------------------- */
typedef struct dsb_history dsb_history;
dsb_history *dsb_history_new(void);
void dsb_history_free(dsb_history *h);
/* For the next read of this worker (reads MUST be presented in input order): fill
 * edge[0..DSB_HEAD_BYTES) = bytes before F[0], edge[DSB_HEAD_BYTES..+DSB_TAIL_BYTES) = bytes after
 * R[L-1], and return the descriptor flags.  `seq` must stay valid until dsb_history_end_batch(). */
uint32_t dsb_history_next(dsb_history *h, const char *seq, uint32_t len, uint8_t *edge);
/* the batch's sequence storage is about to be recycled: take private copies of what is still visible */
void dsb_history_end_batch(dsb_history *h);

/* ---- output (reference src/cly_mt.c:47-348) --------------------------------------------- */
#define DSB_OUT_SAM 1
#define DSB_OUT_SAM_FULL 2
#define DSB_OUT_DES 3
#define DSB_OUT_DES_FULL 4
typedef struct dsb_out_opt { int format; int max_sec_N; } dsb_out_opt;
/* append one read's records to `out` */
void dsb_format_read(dsb_strbuf *out, const dsb_loaded_index *ix, const dsb_out_opt *o, const char *name,
                     const char *seq, const char *qual, uint32_t read_len, const dsb_read_result *r,
                     const dsb_hit *hits);
void dsb_format_sam_full_fields(dsb_strbuf *out, const dsb_loaded_index *ix, const dsb_out_opt *o, const char *name,
                                uint32_t read_len, const dsb_read_result *r, const dsb_hit *hits, size_t *ins);
void dsb_strbuf_reserve(dsb_strbuf *b, size_t extra);
void dsb_strbuf_printf(dsb_strbuf *b, const char *fmt, ...) __attribute__((format(printf, 2, 3)));

/* ---- batching engine (replaces reference src/cly_mt.c:29-43,352-381 + lib/kthread.c) ------ */
typedef struct dsb_engine dsb_engine;
/* one engine = the index resident on n_gpu devices */
dsb_engine *dsb_engine_create(dsb_loaded_index *ix, int n_gpu_wanted);
void dsb_engine_destroy(dsb_engine *e);
int dsb_engine_n_gpu(const dsb_engine *e);

typedef struct dsb_session dsb_session;   /* = one reference "worker buffer set" (a thread_id) */
/* after a direct SAM_FULL run: the text the records describe (start and length of the caller's block) */
void dsb_session_meta_seal(dsb_session *s, const char *text, uint64_t n);
/* the records, if `text` is that very text (same address, same length or one more for the NUL, same sampled bytes) */
int dsb_session_meta_lines(dsb_session *s, const char *text, uint64_t n, const dsb_meta_line **lines, size_t *n_lines);
dsb_session *dsb_session_create(dsb_engine *e);
void dsb_session_destroy(dsb_session *s);
void dsb_session_reset_history(dsb_session *s);
typedef void (*dsb_sink_fn)(void *ctx, const char *text, size_t n);
typedef struct dsb_run_stats {
    uint64_t n_reads, n_bases, n_batches;
    double t_parse, t_gpu_wait, t_format, t_total;
    dsb_counters counters;
    double gpu_ms_total, gpu_ms_exist, gpu_ms_classify;
} dsb_run_stats;
/* classify every record of `f`, in order, writing text to sink; sink == NULL: ctx is a dsb_strbuf* the
 * text is appended to (parallel copy) */
void dsb_session_run(dsb_session *s, dsb_fq *f, const dsb_out_opt *o, dsb_sink_fn sink, void *ctx,
                     dsb_run_stats *stats);

#ifdef __cplusplus
}
#endif
#endif
