/* dsb_sam.c -- text records for classified reads, byte-identical to the reference writer
 * (reference src/cly_mt.c:47-92 print_hit, :144-227 DES / DES_FULL, :229-327 SAM / SAM_FULL).
 * Every record ends "\t\n"; POS is t_st as stored; the third CIGAR-like field may be negative
 * (SURVEY.md Q31).  Input is the structured hit list returned by the CUDA layer.
 */
#include "dsb_host.h"
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

void dsb_strbuf_reserve(dsb_strbuf *b, size_t extra)
{
    if (b->m - b->l < extra + 1) {
        size_t m = b->l + extra + 1;
        m += m >> 1;
        if (m < 4096) m = 4096;
        b->s = (char *)dsb_xrealloc(b->s, m);
        b->m = m;
    }
}

void dsb_strbuf_printf(dsb_strbuf *b, const char *fmt, ...)
{
    va_list ap;
    dsb_strbuf_reserve(b, 512);
    for (;;) {
        va_start(ap, fmt);
        int n = vsnprintf(b->s + b->l, b->m - b->l, fmt, ap);
        va_end(ap);
        if (n < 0) dsb_fatal("vsnprintf failed");
        if ((size_t)n < b->m - b->l) { b->l += (size_t)n; return; }
        dsb_strbuf_reserve(b, (size_t)n + 1);
    }
}

static void sb_puts(dsb_strbuf *b, const char *s, size_t n)
{
    dsb_strbuf_reserve(b, n);
    memcpy(b->s + b->l, s, n);
    b->l += n;
    b->s[b->l] = 0;
}

/* ins != NULL: SEQ "\t" QUAL are left out and *ins receives the offset in `out` where they belong (the caller puts the
 * three pieces straight into the final text, dsb_batch.c) */
static void format_sam(dsb_strbuf *out, const dsb_loaded_index *ix, const dsb_out_opt *o, const char *name,
                       const char *seq_s, size_t seq_n, const char *qual_s, size_t qual_n, uint32_t read_l,
                       const dsb_read_result *r, const dsb_hit *h, size_t *ins)
{
    if (r->n_hit == 0) {
        dsb_strbuf_printf(out, "%s\t4\t*\t0\t0\t*\t*\t0\t0\t", name);
        if (ins) *ins = out->l;
        else {
            sb_puts(out, seq_s, seq_n);
            sb_puts(out, "\t", 1);
            sb_puts(out, qual_s, qual_n);
        }
        sb_puts(out, "\t\n", 2);
        return;
    }
    const dsb_hit *c_s = h;
    int flag = c_s->direction ? 0 : 0x10;
    int mapQ_PRI;
    if (r->n_hit == 1 || (c_s->sum_score - c_s[1].sum_score > 5))
        mapQ_PRI = 30;
    else
        mapQ_PRI = (int)((c_s->sum_score - c_s[1].sum_score) << 2);
    dsb_strbuf_printf(out, "%s\t%d\t%s\t%d\t%d\t%dS%dM%dS\t*\t0\t0\t", name, flag, ix->ref_info[c_s->ref_ID].ref_name,
                      (int)c_s->t_st, mapQ_PRI, (int)c_s->q_st, (int)(c_s->q_ed - c_s->q_st), (int)(read_l - c_s->q_ed));
    if (ins) *ins = out->l;
    else {
        sb_puts(out, seq_s, seq_n);
        sb_puts(out, "\t", 1);
        sb_puts(out, qual_s, qual_n);
    }
    dsb_strbuf_printf(out, "\tAS:i:%d\t\n", (int)c_s->sum_score);
    for (int loop = 0; loop <= 1; loop++) {
        for (uint32_t i = 1; i < r->n_hit; i++) {
            const dsb_hit *c = h + i;
            int show = 0, fl = c->direction ? 0 : 0x10, mapQ = 0;
            if (loop == 0 && c->pri_index == 0) {
                show = 1;
                fl += 0x800;
                mapQ = mapQ_PRI < 30 ? mapQ_PRI : 30;
            } else if (loop == 1 && c->pri_index > 0 && c->pri_index <= o->max_sec_N) {
                show = 1;
                fl += 0x100;
            }
            if (show) {
                char hs = (loop == 0) ? 'H' : 'S';
                dsb_strbuf_printf(out, "%s\t%d\t%s\t%d\t%d\t%d%c%dM%d%c\t*\t0\t0\t*\t*\tAS:i:%d\t\n", name, fl,
                                  ix->ref_info[c->ref_ID].ref_name, (int)c->t_st, mapQ, (int)c->q_st, hs,
                                  (int)(c->q_ed - c->q_st), (int)(read_l - c->q_ed), hs, (int)c->sum_score);
            }
        }
    }
}

static void print_hit(dsb_strbuf *out, const dsb_loaded_index *ix, const dsb_hit *c, int rst_cnt)
{
    static const char primary_string[3][4] = {"PRI", "SEC", "SUP"};
    dsb_strbuf_printf(out, "%3d %s %s %20s ts:%-10d te:%-10d qs:%-10d qe:%-10d %-5d\t%d\t\n", rst_cnt,
                      primary_string[(c->primary - 1) % 3], c->direction ? "F" : "R", ix->ref_info[c->ref_ID].ref_name,
                      (int)c->t_st, (int)c->t_ed, (int)c->q_st, (int)c->q_ed, (int)c->sum_score, (int)c->indel);
}

static void format_des(dsb_strbuf *out, const dsb_loaded_index *ix, const dsb_out_opt *o, const char *name,
                       uint32_t read_l, const dsb_read_result *r, const dsb_hit *h, int full)
{
    dsb_strbuf_printf(out, "%s\t%s\t%s\t%ld\tn_rst:[%ld]\tn_anc:[%ld]\t\n", name, r->n_hit ? "CLASSIFY" : "UNCLASSIFY",
                      (r->fast_classify & DSB_RES_FAST) ? "FAST" : "SLOW", (long)read_l, (long)r->n_hit, (long)r->n_anchor);
    int rst_cnt = 0;
    for (uint32_t i = 0; i < r->n_hit; i++)
        if (h[i].pri_index == 0) print_hit(out, ix, h + i, rst_cnt++);
    for (uint32_t i = 0; i < r->n_hit; i++)
        if (h[i].pri_index > 0 && (full || h[i].pri_index <= o->max_sec_N)) print_hit(out, ix, h + i, rst_cnt++);
    sb_puts(out, "\n", 1);
}

void dsb_format_read(dsb_strbuf *out, const dsb_loaded_index *ix, const dsb_out_opt *o, const char *name,
                     const char *seq, const char *qual, uint32_t read_len, const dsb_read_result *r,
                     const dsb_hit *hits)
{
    switch (o->format) {
    case DSB_OUT_SAM:
        format_sam(out, ix, o, name, "*", 1, "*", 1, read_len, r, hits, NULL);
        break;
    case DSB_OUT_SAM_FULL: {
        /* FASTA records have no quality: the reference prints whatever the recycled kseq slot
         * holds, "(null)" for a never-used slot; we always print "(null)" */
        const char *q = qual ? qual : "(null)";
        format_sam(out, ix, o, name, seq, read_len, q, qual ? read_len : 6, read_len, r, hits, NULL);
        break;
    }
    case DSB_OUT_DES:
        format_des(out, ix, o, name, read_len, r, hits, 0);
        break;
    default:
        format_des(out, ix, o, name, read_len, r, hits, 1);
        break;
    }
}

/* SAM_FULL without SEQ and QUAL: the record fields go to `out`, *ins = where "SEQ\tQUAL" has to be spliced in */
void dsb_format_sam_full_fields(dsb_strbuf *out, const dsb_loaded_index *ix, const dsb_out_opt *o, const char *name,
                                uint32_t read_len, const dsb_read_result *r, const dsb_hit *hits, size_t *ins)
{
    format_sam(out, ix, o, name, NULL, 0, NULL, 0, read_len, r, hits, ins);
}
