/* dsb_index.c -- read the reference's on-disk index (ten deSAMBA.* files) into host arrays.
 *
 * File formats follow reference src/idx.c:1103-1160 (load_idx) and src/bwt.c:68-104 (load_bwt):
 *   .bwt   u64 byteLen | byteLen bytes of 168-byte FM blocks | 5 x u64 rank | (2^26+1) x u64 LUT
 *   .sa    u64 n | n x {u32 unitig_ID, u32 offset}
 *   .exki  u64 table bytes        .exk0 / .exk1  raw bit tables
 *   .unv   u64 n | n x {u32 ref_list, u32 length}     (+ sentinel synthesised, src/idx.c:1127)
 *   .ref_b u64 n | n bytes (4 bases / byte)
 *   .ref_i u64 n | n x 144-byte REF_INFO
 *   .ref_p u64 n | n x 8-byte REF_POS
 * The .acg popcount table is not needed: the device counts nibbles arithmetically.
 */
#include "dsb_host.h"
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

void dsb_fatal(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "[desamba-b200] fatal: ");
    vfprintf(stderr, fmt, ap);
    fprintf(stderr, "\n");
    va_end(ap);
    abort();
}

void *dsb_xmalloc(size_t n)
{
    void *p = malloc(n ? n : 1);
    if (!p) dsb_fatal("out of memory (%zu bytes)", n);
    return p;
}

void *dsb_xcalloc(size_t n, size_t sz)
{
    void *p = calloc(n ? n : 1, sz ? sz : 1);
    if (!p) dsb_fatal("out of memory (%zu x %zu bytes)", n, sz);
    return p;
}

void *dsb_xrealloc(void *p, size_t n)
{
    void *q = realloc(p, n ? n : 1);
    if (!q) dsb_fatal("out of memory (%zu bytes)", n);
    return q;
}

double dsb_now(void)
{
    struct timeval tv;
    gettimeofday(&tv, NULL);
    return (double)tv.tv_sec + 1e-6 * (double)tv.tv_usec;
}

static FILE *open_part(const char *dir, const char *suffix)
{
    char path[4096];
    snprintf(path, sizeof path, "%s/deSAMBA%s", dir, suffix);
    FILE *fp = fopen(path, "rb");
    if (!fp) dsb_fatal("cannot open index file %s", path);
    return fp;
}

static void read_exact(void *dst, size_t size, size_t count, FILE *fp, const char *what)
{
    /* fread in <= 1 GiB pieces: some libc builds mis-handle very large single requests */
    size_t total = size * count, done = 0;
    while (done < total) {
        size_t want = total - done;
        if (want > ((size_t)1 << 30)) want = (size_t)1 << 30;
        size_t got = fread((char *)dst + done, 1, want, fp);
        if (got != want) dsb_fatal("index file truncated while reading %s", what);
        done += got;
    }
}

static uint64_t read_u64(FILE *fp, const char *what)
{
    uint64_t v;
    read_exact(&v, 8, 1, fp, what);
    return v;
}

/* k and hash mask from the table size (reference src/idx.c:966-982, set_ekmer_par) */
static void ekmer_params(uint64_t table_bytes, int *k, uint64_t *mask_bits)
{
    int bits = 37, len = 20;
    switch (table_bytes) {
    case 0x8000000ull: bits = 30; len = 16; break;
    case 0x10000000ull: bits = 31; len = 17; break;
    case 0x20000000ull: bits = 32; len = 17; break;
    case 0x40000000ull: bits = 33; len = 18; break;
    case 0x80000000ull: bits = 34; len = 18; break;
    case 0x100000000ull: bits = 35; len = 19; break;
    case 0x200000000ull: bits = 36; len = 19; break;
    case 0x400000000ull: bits = 37; len = 20; break;
    }
    *k = len;
    *mask_bits = (uint64_t)bits;
}

int dsb_ekmer_len(uint64_t table_bytes)
{
    int k;
    uint64_t b;
    ekmer_params(table_bytes, &k, &b);
    return k;
}

void dsb_mapq_tables(int32_t *q_mem, int32_t *q_lv, double P_E, uint64_t L_REF)
{
    /* same expressions, same types and the same libm as the reference (src/cly_mt.c:396-420) */
    double REF_SIZE_PUNALTY = -10 * log(L_REF) / log(10);
    double MATCH_SCORE = -10 * log(0.25 / (1 - P_E)) / log(10);
    double MISMATCH_PUNALTY = -10 * log(0.75 / (P_E)) / log(10);
    for (int i = 0; i < 2000; i++) q_mem[i] = REF_SIZE_PUNALTY + i * MATCH_SCORE + 0.5;
    for (int j = 0; j < 20; j++) {
        for (int i = 0; i < 20; i++) {
            int v = (j - i) * MATCH_SCORE + i * MISMATCH_PUNALTY + 0.5;
            if (j < 5) v += 15;
            if (v < -8) v = -8;
            q_lv[i * 20 + j] = v;
        }
    }
}

void dsb_index_load(dsb_loaded_index *ix, const char *dir, int min_len, int min_score)
{
    memset(ix, 0, sizeof *ix);
    FILE *fp;
    /* FM index */
    fp = open_part(dir, ".bwt");
    ix->h.bwt_bytes = read_u64(fp, ".bwt length");
    ix->bwt_occ = (uint8_t *)dsb_xmalloc(ix->h.bwt_bytes + 168);
    read_exact(ix->bwt_occ, 1, ix->h.bwt_bytes, fp, ".bwt blocks");
    memset(ix->bwt_occ + ix->h.bwt_bytes, 0, 168);
    read_exact(ix->h.rank, 8, 5, fp, ".bwt rank");
    ix->h.rank[5] = ix->h.rank[0] - 1;
    const uint64_t lut_n = ((uint64_t)1 << 26) + 1;
    ix->hash_index = (uint64_t *)dsb_xmalloc(lut_n * 8);
    read_exact(ix->hash_index, 8, lut_n, fp, ".bwt prefix table");
    fclose(fp);
    fp = open_part(dir, ".sa");
    ix->h.sa_n = read_u64(fp, ".sa count");
    ix->sa = (uint64_t *)dsb_xmalloc(ix->h.sa_n * 8);
    read_exact(ix->sa, 8, ix->h.sa_n, fp, ".sa");
    fclose(fp);
    /* exist k-mer tables */
    fp = open_part(dir, ".exki");
    ix->h.e_kmer_size = read_u64(fp, ".exki");
    fclose(fp);
    ix->ek_len = dsb_ekmer_len(ix->h.e_kmer_size);
    ix->ek0 = (uint8_t *)dsb_xmalloc(ix->h.e_kmer_size);
    ix->ek1 = (uint8_t *)dsb_xmalloc(ix->h.e_kmer_size);
    fp = open_part(dir, ".exk0");
    read_exact(ix->ek0, 1, ix->h.e_kmer_size, fp, ".exk0");
    fclose(fp);
    fp = open_part(dir, ".exk1");
    read_exact(ix->ek1, 1, ix->h.e_kmer_size, fp, ".exk1");
    fclose(fp);
    /* unitigs (+ sentinel) */
    fp = open_part(dir, ".unv");
    uint64_t n_uni = read_u64(fp, ".unv count");
    if (n_uni < 2) dsb_fatal("index has %llu unitigs", (unsigned long long)n_uni);
    ix->unitig = (uint64_t *)dsb_xmalloc((n_uni + 1) * 8);
    read_exact(ix->unitig, 8, n_uni, fp, ".unv");
    fclose(fp);
    {
        uint32_t last_ref_list = (uint32_t)ix->unitig[n_uni - 1];
        uint32_t last_len = (uint32_t)(ix->unitig[n_uni - 1] >> 32);
        uint32_t sentinel_ref_list = last_ref_list + 1 + last_len;      /* src/idx.c:1127 */
        ix->unitig[n_uni] = (uint64_t)sentinel_ref_list;                /* length is uninitialised in the reference */
    }
    ix->h.n_unitig = n_uni + 1;
    ix->h.dollar_pos = n_uni - 1 - 1;                                   /* src/idx.c:1128 */
    /* references */
    fp = open_part(dir, ".ref_b");
    ix->h.ref_bin_n = read_u64(fp, ".ref_b count");
    ix->ref_bin = (uint8_t *)dsb_xmalloc(ix->h.ref_bin_n + 1);
    read_exact(ix->ref_bin, 1, ix->h.ref_bin_n, fp, ".ref_b");
    fclose(fp);
    fp = open_part(dir, ".ref_i");
    ix->h.n_ref = read_u64(fp, ".ref_i count");
    ix->ref_info = (dsb_ref_info *)dsb_xmalloc(ix->h.n_ref * sizeof(dsb_ref_info));
    read_exact(ix->ref_info, sizeof(dsb_ref_info), ix->h.n_ref, fp, ".ref_i");
    fclose(fp);
    ix->ref_off = (uint64_t *)dsb_xmalloc(ix->h.n_ref * 8);
    ix->ref_len = (uint64_t *)dsb_xmalloc(ix->h.n_ref * 8);
    for (uint64_t i = 0; i < ix->h.n_ref; i++) {
        ix->ref_off[i] = ix->ref_info[i].seq_offset;
        ix->ref_len[i] = ix->ref_info[i].seq_l;
        ix->ref_info[i].ref_name[127] = 0;
    }
    fp = open_part(dir, ".ref_p");
    ix->h.n_ref_pos = read_u64(fp, ".ref_p count");
    ix->ref_pos = (uint64_t *)dsb_xmalloc((ix->h.n_ref_pos + 1) * 8);
    read_exact(ix->ref_pos, 8, ix->h.n_ref_pos, fp, ".ref_p");
    ix->ref_pos[ix->h.n_ref_pos] = 0;
    fclose(fp);
    /* score tables and filters (src/cly_mt.c:518-524) */
    dsb_mapq_tables(ix->q_mem, ix->q_lv, 0.15, ix->h.ref_bin_n * 4);
    ix->h.filter_min_length = min_len;
    ix->h.filter_min_score = min_score;
    ix->h.filter_min_score_lv3 = min_score + 10;
    /* views */
    ix->h.bwt_occ = ix->bwt_occ;
    ix->h.hash_index = ix->hash_index;
    ix->h.sa_taxon = ix->sa;
    ix->h.e_kmer0 = ix->ek0;
    ix->h.e_kmer1 = ix->ek1;
    ix->h.unitig = ix->unitig;
    ix->h.ref_pos = ix->ref_pos;
    ix->h.ref_bin = ix->ref_bin;
    ix->h.ref_seq_offset = ix->ref_off;
    ix->h.ref_seq_len = ix->ref_len;
    ix->h.q_mem = ix->q_mem;
    ix->h.q_lv = ix->q_lv;
}

void dsb_index_release_bulk(dsb_loaded_index *ix)
{
    free(ix->bwt_occ); ix->bwt_occ = NULL; ix->h.bwt_occ = NULL;
    free(ix->hash_index); ix->hash_index = NULL; ix->h.hash_index = NULL;
    free(ix->sa); ix->sa = NULL; ix->h.sa_taxon = NULL;
    free(ix->ek0); ix->ek0 = NULL; ix->h.e_kmer0 = NULL;
    free(ix->ek1); ix->ek1 = NULL; ix->h.e_kmer1 = NULL;
    free(ix->unitig); ix->unitig = NULL; ix->h.unitig = NULL;
    free(ix->ref_pos); ix->ref_pos = NULL; ix->h.ref_pos = NULL;
    free(ix->ref_bin); ix->ref_bin = NULL; ix->h.ref_bin = NULL;
}

void dsb_index_destroy(dsb_loaded_index *ix)
{
    dsb_index_release_bulk(ix);
    free(ix->ref_info);
    free(ix->ref_off);
    free(ix->ref_len);
    memset(ix, 0, sizeof *ix);
}
