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
#include <fcntl.h>
#include <pthread.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>

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

/* ---- mapped index files ------------------------------------------------------------------
 * The reference reads every file into a malloc'd array (src/idx.c:1103-1160, src/bwt.c:68-104).  Here the arrays go to
 * the GPUs and the host never needs them again, so the files are only MAPPED: the views handed to dsb_index_upload point
 * into the page cache, and a helper thread per file touches the pages ahead of the upload's staging copies. */
typedef struct prefault_job { const volatile char *p; size_t n; pthread_t th; } prefault_job;
typedef struct prefault_set { prefault_job job[8]; int n; } prefault_set;

static void *prefault_thread(void *arg)
{
    prefault_job *j = (prefault_job *)arg;
    char sink = 0;
    for (size_t i = 0; i < j->n; i += 4096) sink ^= j->p[i];
    (void)sink;
    return NULL;
}

static const uint8_t *map_part(dsb_loaded_index *ix, const char *dir, const char *suffix, uint64_t *size)
{
    char path[4096];
    snprintf(path, sizeof path, "%s/deSAMBA%s", dir, suffix);
    int fd = open(path, O_RDONLY);
    if (fd < 0) dsb_fatal("cannot open index file %s", path);
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size < 8) dsb_fatal("index file %s is truncated", path);
    void *p = mmap(NULL, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (p == MAP_FAILED) dsb_fatal("cannot map index file %s", path);
    madvise(p, (size_t)st.st_size, MADV_WILLNEED);
    if (ix->n_maps >= 8) dsb_fatal("too many mapped index files");
    ix->maps[ix->n_maps].p = p;
    ix->maps[ix->n_maps].n = (size_t)st.st_size;
    ix->n_maps++;
    prefault_set *ps = (prefault_set *)ix->prefault;
    if (ps && ps->n < 8 && st.st_size > (16 << 20)) {
        prefault_job *j = &ps->job[ps->n];
        j->p = (const volatile char *)p;
        j->n = (size_t)st.st_size;
        if (pthread_create(&j->th, NULL, prefault_thread, j) == 0) ps->n++;
    }
    *size = (uint64_t)st.st_size;
    return (const uint8_t *)p;
}

static uint64_t counted(const uint8_t *m, uint64_t size, uint64_t elem, const char *what)
{
    uint64_t n;
    memcpy(&n, m, 8);
    if (size < 8 || n > (size - 8) / elem) dsb_fatal("index file truncated while reading %s", what);
    return n;
}

void dsb_index_load(dsb_loaded_index *ix, const char *dir, int min_len, int min_score)
{
    memset(ix, 0, sizeof *ix);
    ix->prefault = dsb_xcalloc(1, sizeof(prefault_set));
    uint64_t sz;
    const uint8_t *m;
    FILE *fp;
    /* FM index: u64 byteLen | blocks | 5 x u64 rank | (2^26 + 1) x u64 prefix table */
    m = map_part(ix, dir, ".bwt", &sz);
    memcpy(&ix->h.bwt_bytes, m, 8);
    const uint64_t lut_n = ((uint64_t)1 << 26) + 1;
    if (ix->h.bwt_bytes > sz || sz - ix->h.bwt_bytes < 8 + 40 + lut_n * 8) dsb_fatal("index file truncated while reading .bwt");
    ix->h.bwt_occ = m + 8;
    memcpy(ix->h.rank, m + 8 + ix->h.bwt_bytes, 40);
    ix->h.rank[5] = ix->h.rank[0] - 1;
    ix->h.hash_index = (const uint64_t *)(m + 8 + ix->h.bwt_bytes + 40);
    m = map_part(ix, dir, ".sa", &sz);
    ix->h.sa_n = counted(m, sz, 8, ".sa");
    ix->h.sa_taxon = (const uint64_t *)(m + 8);
    /* exist k-mer tables */
    fp = open_part(dir, ".exki");
    ix->h.e_kmer_size = read_u64(fp, ".exki");
    fclose(fp);
    ix->ek_len = dsb_ekmer_len(ix->h.e_kmer_size);
    ix->h.e_kmer0 = map_part(ix, dir, ".exk0", &sz);
    if (sz < ix->h.e_kmer_size) dsb_fatal("index file truncated while reading .exk0");
    ix->h.e_kmer1 = map_part(ix, dir, ".exk1", &sz);
    if (sz < ix->h.e_kmer_size) dsb_fatal("index file truncated while reading .exk1");
    /* unitigs (+ sentinel): the one array that gets an entry the file does not hold, so it is copied */
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
    ix->h.unitig = ix->unitig;
    ix->h.n_unitig = n_uni + 1;
    ix->h.dollar_pos = n_uni - 1 - 1;                                   /* src/idx.c:1128 */
    /* references */
    m = map_part(ix, dir, ".ref_b", &sz);
    ix->h.ref_bin_n = counted(m, sz, 1, ".ref_b");
    ix->h.ref_bin = m + 8;
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
    m = map_part(ix, dir, ".ref_p", &sz);
    ix->h.n_ref_pos = counted(m, sz, 8, ".ref_p");
    ix->h.ref_pos = (const uint64_t *)(m + 8);
    /* score tables and filters (src/cly_mt.c:518-524) */
    dsb_mapq_tables(ix->q_mem, ix->q_lv, 0.15, ix->h.ref_bin_n * 4);
    ix->h.filter_min_length = min_len;
    ix->h.filter_min_score = min_score;
    ix->h.filter_min_score_lv3 = min_score + 10;
    ix->h.ref_seq_offset = ix->ref_off;
    ix->h.ref_seq_len = ix->ref_len;
    ix->h.q_mem = ix->q_mem;
    ix->h.q_lv = ix->q_lv;
}

void dsb_index_release_bulk(dsb_loaded_index *ix)
{
    prefault_set *ps = (prefault_set *)ix->prefault;
    if (ps) {
        for (int i = 0; i < ps->n; i++) pthread_join(ps->job[i].th, NULL);
        free(ps);
        ix->prefault = NULL;
    }
    for (int i = 0; i < ix->n_maps; i++) munmap(ix->maps[i].p, ix->maps[i].n);
    ix->n_maps = 0;
    ix->h.bwt_occ = NULL; ix->h.hash_index = NULL; ix->h.sa_taxon = NULL; ix->h.e_kmer0 = NULL; ix->h.e_kmer1 = NULL;
    ix->h.ref_pos = NULL; ix->h.ref_bin = NULL;
    free(ix->unitig); ix->unitig = NULL; ix->h.unitig = NULL;
}

void dsb_index_destroy(dsb_loaded_index *ix)
{
    dsb_index_release_bulk(ix);
    free(ix->ref_info);
    free(ix->ref_off);
    free(ix->ref_len);
    memset(ix, 0, sizeof *ix);
}
