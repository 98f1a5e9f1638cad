/* dsb_api.c -- the three exported entry points of libdesamba.so (include/desamba.h).
 *
 * Mirrors reference src/cly_mt.c:1238-1316: load_index() loads once and keeps the index
 * resident (here: on every GPU), read_classify() may be called many times and from several
 * caller threads with distinct thread_ids; each thread_id owns one session = one reference
 * "RM_buffer" (stream slots, pinned staging, and the worker-buffer history that the reference
 * carries across calls in Classify_buff_pool).
 */
#include "desamba.h"
#include "dsb_host.h"
#include <pthread.h>
#include <sys/mman.h>
#include <stdlib.h>
#include <string.h>

typedef struct dsb_taxonomy dsb_taxonomy;
dsb_taxonomy *dsb_taxonomy_load(const char *dir);
void dsb_meta_analysis_impl(dsb_taxonomy *tax, const char *input, uint64_t input_n, char **output, uint64_t *output_n,
                            int flag, uint64_t max_snapshot_len, char **human_snapshot, uint64_t *human_snapshot_n);
void dsb_meta_analysis_lines(dsb_taxonomy *tax, const dsb_meta_line *lines, size_t n_lines, const char *text, char **output,
                             uint64_t *output_n, int flag, uint64_t max_snapshot_len, char **human_snapshot,
                             uint64_t *human_snapshot_n);

typedef struct api_index {
    uint32_t magic;
    dsb_loaded_index ix;
    dsb_engine *engine;
    dsb_taxonomy *tax;
    pthread_mutex_t mu;
    int n_sessions, m_sessions;
    int *session_tid;
    int *session_thread_num;        /* thread_num of the thread_id's last read_classify call */
    dsb_session **sessions;
} api_index;

#define API_MAGIC 0xD5A3B200u

void load_index(void **idx, const char *dirPath)
{
    api_index *a = (api_index *)dsb_xcalloc(1, sizeof *a);
    a->magic = API_MAGIC;
    /* fixed options of the API (reference src/cly_mt.c:1257-1259) */
    dsb_index_load(&a->ix, dirPath, 170, 64);
    a->engine = dsb_engine_create(&a->ix, 0);
    dsb_index_release_bulk(&a->ix);
    a->tax = dsb_taxonomy_load(dirPath);
    pthread_mutex_init(&a->mu, NULL);
    *idx = a;
}

static dsb_session *session_for(api_index *a, int thread_id, int thread_num)
{
    dsb_session *s = NULL;
    pthread_mutex_lock(&a->mu);
    for (int i = 0; i < a->n_sessions; i++)
        if (a->session_tid[i] == thread_id) {
            s = a->sessions[i];
            /* a different thread_num makes the reference throw the thread_id's buffers away and start afresh
             * (src/cly_mt.c:1286-1295): the buffer history of earlier calls ends here */
            if (thread_num != -1 && a->session_thread_num[i] != thread_num) {
                dsb_session_reset_history(s);
                a->session_thread_num[i] = thread_num;
            }
            break;
        }
    if (!s) {
        if (a->n_sessions == a->m_sessions) {
            a->m_sessions = a->m_sessions ? a->m_sessions * 2 : 8;
            a->session_tid = (int *)dsb_xrealloc(a->session_tid, (size_t)a->m_sessions * sizeof(int));
            a->session_thread_num = (int *)dsb_xrealloc(a->session_thread_num, (size_t)a->m_sessions * sizeof(int));
            a->sessions = (dsb_session **)dsb_xrealloc(a->sessions, (size_t)a->m_sessions * sizeof(dsb_session *));
        }
        s = dsb_session_create(a->engine);
        a->session_tid[a->n_sessions] = thread_id;
        a->session_thread_num[a->n_sessions] = thread_num;
        a->sessions[a->n_sessions++] = s;
    }
    pthread_mutex_unlock(&a->mu);
    return s;
}

void read_classify(void *idx, char *input, uint64_t input_n, char **output, uint64_t *output_n, int thread_id,
                   int thread_num)
{
    (void)thread_num;   /* results are those of thread_num == 1, the reference's only reproducible mode */
    if (input_n == 0) {
        *output_n = 0;
        return;
    }
    api_index *a = (api_index *)idx;
    if (!a || a->magic != API_MAGIC) dsb_fatal("read_classify: bad index handle");
    dsb_session *s = session_for(a, thread_id, thread_num);
    dsb_fq *f = (input_n == (uint64_t)-1) ? dsb_fq_open_path(input) : dsb_fq_open_mem(input, input_n);
    dsb_out_opt o = {DSB_OUT_SAM_FULL, 5};          /* reference src/cly_mt.c:970-976 */
    dsb_strbuf out = {NULL, 0, 0};
    /* SAM_FULL repeats SEQ and QUAL of every read: about the size of the input plus the record fields */
    {
        /* a path that names a plain file is mapped, so its size is known as well */
        const uint64_t bytes = (input_n == (uint64_t)-1) ? dsb_fq_remaining_bytes(f) : input_n;
        if (bytes == UINT64_MAX)
            dsb_strbuf_reserve(&out, 1);
        else {                                      /* exactly this much (dsb_strbuf_reserve would add half again) */
            out.m = (size_t)(bytes + bytes / 16 + 4096);
            out.s = (char *)dsb_xmalloc(out.m);
        }
    }
#ifdef MADV_HUGEPAGE
    if (out.m >= ((size_t)4 << 20)) {
        /* a fresh block this large comes straight from mmap: ask for huge pages so that filling it
         * costs tens of page faults instead of tens of thousands */
        uintptr_t a = ((uintptr_t)out.s + ((size_t)2 << 20) - 1) & ~(uintptr_t)(((size_t)2 << 20) - 1);
        uintptr_t e = ((uintptr_t)out.s + out.m) & ~(uintptr_t)(((size_t)2 << 20) - 1);
        if (e > a) madvise((void *)a, (size_t)(e - a), MADV_HUGEPAGE);
    }
#endif
    dsb_session_run(s, f, &o, NULL, &out, NULL);        /* NULL sink: append straight to `out` */
    dsb_fq_close(f);
    out.s[out.l] = 0;
    dsb_session_meta_seal(s, out.s, out.l);           /* a meta_analysis() on this very block reads the hit records, not the text */
    /* caller free()s: hand over a plain malloc block */
    *output = out.s;
    *output_n = out.l;
}

/* which path the calling thread's last meta_analysis() took: 1 = hit records kept by read_classify, 0 = text parser (tests) */
static __thread int g_meta_path;
int dsb_meta_last_path(void) { return g_meta_path; }

void meta_analysis(void *idx, char *input, uint64_t input_n, char **output, uint64_t *output_n, int thread_id, int flag,
                   uint64_t max_snapshot_len, char **human_snapshot, uint64_t *human_snapshot_n)
{
    api_index *a = (api_index *)idx;
    if (!a || a->magic != API_MAGIC) dsb_fatal("meta_analysis: bad index handle");
    if (input_n != 0) {
        /* the block read_classify() returned to this thread_id comes back unchanged -> its hit records */
        dsb_session *s = NULL;
        pthread_mutex_lock(&a->mu);
        for (int i = 0; i < a->n_sessions; i++)
            if (a->session_tid[i] == thread_id) { s = a->sessions[i]; break; }
        pthread_mutex_unlock(&a->mu);
        const dsb_meta_line *lines;
        size_t n_lines;
        g_meta_path = 0;
        if (s && dsb_session_meta_lines(s, input, input_n, &lines, &n_lines)) {
            g_meta_path = 1;
            dsb_meta_analysis_lines(a->tax, lines, n_lines, input, output, output_n, flag, max_snapshot_len, human_snapshot,
                                    human_snapshot_n);
            return;
        }
    }
    dsb_meta_analysis_impl(a->tax, input, input_n, output, output_n, flag, max_snapshot_len, human_snapshot, human_snapshot_n);
}
