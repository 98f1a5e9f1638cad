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
 *   flags bit0: `buff->max_read_l < 510` after this read (src/cly.c:2953-2955).  max_read_l is raised only by reads
 *       that reach delete_small_score_rst WITH a chain (src/cly.c:2884-2885 returns before :2953), which the host
 *       learns when a batch comes back (DSB_RES_HAD_CHAIN).  So for a read shorter than 510 bases there are three
 *       cases (dsb_history_class): an earlier read of >= 510 bases is KNOWN to have had a chain -> long-read filters;
 *       no earlier read of >= 510 bases can still turn out to have had one -> 2G filter; otherwise the answer hangs on
 *       batches still on the device: the batcher closes the batch before that read and waits for them.
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
    uint32_t max_known;    /* the reference's max_read_l over the reads whose results are back */
    uint32_t max_pending;  /* longest read handed to a device whose DSB_RES_HAD_CHAIN bit is not known yet */
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
        return 0;
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
    return 0;
}

/* Which filter class a read of `len` bases gets (src/cly.c:2953-2981): 0 = max_read_l >= 510 for certain,
 * 1 = max_read_l < 510 for certain ("2G"), 2 = depends on reads whose results are not back yet. */
int dsb_history_class(const dsb_history *h, uint32_t len)
{
    if (len >= 510 || __atomic_load_n(&h->max_known, __ATOMIC_ACQUIRE) >= 510) return 0;
    if (h->max_pending < 510) return 1;
    return 2;
}

/* a read goes to a device (producer thread) */
void dsb_history_note_pending(dsb_history *h, uint32_t len)
{
    if (len >= 40 && len > h->max_pending) h->max_pending = len;
}

/* a read's result is back (drain thread, in read order) */
void dsb_history_note_result(dsb_history *h, uint32_t len, int had_chain)
{
    if (had_chain && len > __atomic_load_n(&h->max_known, __ATOMIC_RELAXED)) __atomic_store_n(&h->max_known, len, __ATOMIC_RELEASE);
}

/* every batch handed out so far has been drained (producer thread, after waiting for the drain thread) */
void dsb_history_all_drained(dsb_history *h) { h->max_pending = 0; }

uint32_t dsb_history_max(const dsb_history *h) { return h->max_known; }

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
