/* dsb_history.c -- model of the reference worker's `bin_read` buffer across reads.
 *
 * The reference classifier reads a few dozen bytes beyond both ends of the encoded read
 * (reference src/cly.c:2365-2367 forward, :2414-2416 backward) inside a per-worker buffer that
 * is reused for every read (src/cly.c:1236,1242,1250): layout [F (L) | R (L) | older bytes].
 * With one worker (-t 1 / thread_num 1, the only deterministic mode) those bytes
 * are a function of the reads processed before.  This file computes them on the host, in read
 * order, so that the GPU may classify reads in any order and on any device:
 *
 *   tail[p - 2L], p in [2L, 2L+64): the byte written by the most recent earlier read j with
 *       2*L_j > p  (F_j[p] if p < L_j else R_j[p - L_j]); bytes nobody wrote yet are fresh heap
 *       memory (0 from the kernel); bytes beyond the allocation (m = 2*L_max_so_far + 20,
 *       BUFF_REALLOC src/lib/utils.h:117-122) are the neighbouring heap chunk: 0xFF = no match.
 *   head[64 + q], q in [-64, -1]: the 8 bytes before the buffer are glibc's chunk size field
 *       (little endian, low bits = flags); anything further is unknown: 0xFF = no match.
 *   flags bit0: `buff->max_read_l < 510` after this read (src/cly.c:2953-2955).
 */
#include "dsb_host.h"
#include <stdlib.h>
#include <string.h>

typedef struct hist_entry {
    const char *seq;   /* ASCII bases of read j (borrowed from the batch, or owned) */
    char *owned;
    uint32_t len;
} hist_entry;

struct dsb_history {
    hist_entry *st;    /* skyline: lengths strictly decrease from bottom (oldest) to top (newest) */
    int n, m;
    uint64_t m_bin_read;   /* current allocation size of bin_read */
    uint32_t max_read_l;
    int n_borrowed;
};

static const uint8_t enc_tab_init[4] = {0, 1, 2, 3};
static uint8_t enc_tab[256];
static int enc_ready;

static void enc_setup(void)
{
    /* CLY_Bit (src/cly.c:16-34): A/a 0, G/g 2, T/t 3, everything else (incl. C, N) 1 */
    for (int i = 0; i < 256; i++) enc_tab[i] = 1;
    enc_tab['A'] = enc_tab['a'] = 0;
    enc_tab['G'] = enc_tab['g'] = 2;
    enc_tab['T'] = enc_tab['t'] = 3;
    (void)enc_tab_init;
    enc_ready = 1;
}

dsb_history *dsb_history_new(void)
{
    if (!enc_ready) enc_setup();
    dsb_history *h = (dsb_history *)dsb_xcalloc(1, sizeof *h);
    return h;
}

void dsb_history_free(dsb_history *h)
{
    if (!h) return;
    for (int i = 0; i < h->n; i++) free(h->st[i].owned);
    free(h->st);
    free(h);
}

static inline uint8_t buffer_byte_of(const hist_entry *e, uint64_t p)
{
    if (p < e->len) return enc_tab[(unsigned char)e->seq[p]];
    return (uint8_t)(3 - enc_tab[(unsigned char)e->seq[e->len - 1 - (p - e->len)]]);
}

uint32_t dsb_history_next(dsb_history *h, const char *seq, uint32_t len, uint8_t *edge)
{
    uint8_t *head = edge, *tail = edge + DSB_HEAD_BYTES;
    if (len < 40) {
        /* classify_seq returns before getIsland (src/cly.c:3084): the buffer is untouched */
        memset(edge, 0xFF, DSB_HEAD_BYTES + DSB_TAIL_BYTES);
        return (h->max_read_l < 510) ? DSB_READ_2G : 0;
    }
    /* BUFF_REALLOC(bin_read, m, 2L) */
    if ((uint64_t)2 * len > h->m_bin_read) h->m_bin_read = (uint64_t)2 * len + 20;
    /* head: chunk size field.  request2size(m) with PREV_INUSE|NON_MAIN_ARENA (the buffer is
     * (re)allocated by a pipeline thread, src/lib/kthread.c:122-137) */
    memset(head, 0xFF, DSB_HEAD_BYTES);
    {
        uint64_t csz = (h->m_bin_read + 8 + 15) & ~(uint64_t)15;
        if (csz < 32) csz = 32;
        csz |= 0x5;
        for (int b = 0; b < 8; b++) head[DSB_HEAD_BYTES - 8 + b] = (uint8_t)(csz >> (8 * b));
    }
    /* tail */
    const uint64_t base = (uint64_t)2 * len;
    for (int t = 0; t < DSB_TAIL_BYTES; t++) {
        const uint64_t p = base + (uint64_t)t;
        uint8_t v;
        if (p >= h->m_bin_read) {
            v = 0xFF;
        } else {
            v = 0;
            for (int k = h->n - 1; k >= 0; k--) {
                if ((uint64_t)2 * h->st[k].len > p) { v = buffer_byte_of(&h->st[k], p); break; }
            }
        }
        tail[t] = v;
    }
    /* this read now overwrites [0, 2L): everything not longer than it disappears */
    while (h->n > 0 && h->st[h->n - 1].len <= len) {
        free(h->st[h->n - 1].owned);
        h->n--;
    }
    if (h->n == h->m) {
        h->m = h->m ? h->m * 2 : 16;
        h->st = (hist_entry *)dsb_xrealloc(h->st, (size_t)h->m * sizeof(hist_entry));
    }
    h->st[h->n].seq = seq;
    h->st[h->n].owned = NULL;
    h->st[h->n].len = len;
    h->n++;
    /* max_read_l: see dsb_batch.c for the exact rule (only reads that reach the filter update it) */
    uint32_t flags = 0;
    uint32_t mx = h->max_read_l > len ? h->max_read_l : len;
    if (mx < 510) flags |= DSB_READ_2G;
    return flags;
}

void dsb_history_note_max(dsb_history *h, uint32_t len)
{
    if (len > h->max_read_l) h->max_read_l = len;
}

uint32_t dsb_history_max(const dsb_history *h) { return h->max_read_l; }

void dsb_history_end_batch(dsb_history *h)
{
    for (int i = 0; i < h->n; i++) {
        if (!h->st[i].owned) {
            h->st[i].owned = (char *)dsb_xmalloc(h->st[i].len ? h->st[i].len : 1);
            memcpy(h->st[i].owned, h->st[i].seq, h->st[i].len);
            h->st[i].seq = h->st[i].owned;
        }
    }
}
