/* kat_harness.c -- TEST INFRASTRUCTURE ONLY.  Our own file; it #includes the reference's
 * src/cly.c IN PLACE (path given by -DREF_CLY_C, nothing is copied into this repository) so that
 * the reference's static functions can be called directly, and writes known-answer vectors that
 * pin oracle/oracle_*.c (tests/test_oracle.py) and the device primitives (tests/test_gpu_*.py).
 *
 *   kat_harness <index_dir> <reads.fq> <out_dir> [n_reads]
 *
 * Files written (little endian, arrays back to back):
 *   occ.bin    u64 n | u64 r[n] | u8 c_in[n] | u8 c_out[n] | u64 result[n]           occ()            src/bwt.c:43-65
 *   hash.bin   u64 n | u64 key[n] | u64 h1[n] | u64 h2[n]                            hash64_1/2       src/lib/utils.c:1067-1091
 *   lv.bin     u64 n | i32 len[n] | u8 ref[n][12] | u8 qry[n][12] | i32 result[n]    lv_extd()        src/cly.c:505-604
 *   mapq.bin   u64 l_ref | i32 q_mem[2000] | i32 q_lv[400]                           calculate_MAPQ_TABLE src/cly_mt.c:396-420
 *   reads.bin  u64 n_reads | per read: u32 L | u8 bin[2L] (F then R) | u8 exist[2][L-k+1] |
 *              per strand (0 fwd, 1 rev): u32 total_score | u32 n_seeds | {u32 offset,len,top}[n_seeds]
 *              (store_kmers + get_exist_kmer + getIsland, src/cly.c:359-397, 951-967, 1231-1263)
 *   getref.bin u64 n | u64 off[n] | u32 len[n] | u8 fwd[n] | u8 out[n][16]           get_ref()        src/cly.c:434-461
 */
#include REF_CLY_C
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

extern void calculate_MAPQ_TABLE(int *Q_MEM, int (*Q_LV)[MAX_LV_R_LEN], double P_E, uint64_t L_REF);

static uint64_t rng_s = 0x9E3779B97F4A7C15ull;
static uint64_t rnd(void)
{
    rng_s ^= rng_s << 13; rng_s ^= rng_s >> 7; rng_s ^= rng_s << 17;
    return rng_s;
}
static FILE *out_open(const char *dir, const char *name)
{
    char p[4096];
    snprintf(p, sizeof p, "%s/%s", dir, name);
    FILE *f = fopen(p, "wb");
    if (!f) { perror(p); exit(1); }
    return f;
}
#define W(f, ptr, n) fwrite(ptr, sizeof(*(ptr)), n, f)

int main(int argc, char **argv)
{
    if (argc < 4) { fprintf(stderr, "usage: kat_harness <index_dir> <reads.fq> <out_dir> [n_reads]\n"); return 1; }
    const char *out = argv[3];
    int want_reads = argc > 4 ? atoi(argv[4]) : 16;
    DA_IDX *idx = (DA_IDX *)calloc(1, sizeof(DA_IDX));
    load_idx(idx, argv[1]);
    bwt *bt = &idx->bt;
    const uint64_t n_sym = (bt->byteLen / 168) * 256;

    {   /* occ */
        const uint64_t n = 20000;
        uint64_t *r = malloc(n * 8), *res = malloc(n * 8);
        uint8_t *ci = malloc(n), *co = malloc(n);
        for (uint64_t i = 0; i < n; i++) {
            r[i] = rnd() % (n_sym - 256);
            if (i < 512) r[i] = (r[i] & ~255ull) + (i & 255);          /* every in-block offset */
            ci[i] = (i % 5 == 4) ? 0xff : (uint8_t)(rnd() & 3);
            uint8_t c = ci[i];
            res[i] = occ(bt, r[i], &c);
            co[i] = c;
        }
        FILE *f = out_open(out, "occ.bin");
        W(f, &n, 1); W(f, r, n); W(f, ci, n); W(f, co, n); W(f, res, n);
        fclose(f);
    }
    {   /* hashes */
        const uint64_t n = 1000;
        uint64_t k[1000], h1[1000], h2[1000];
        for (uint64_t i = 0; i < n; i++) { k[i] = rnd() >> (i % 40); h1[i] = hash64_1(k[i]); h2[i] = hash64_2(k[i]); }
        FILE *f = out_open(out, "hash.bin");
        W(f, &n, 1); W(f, k, n); W(f, h1, n); W(f, h2, n);
        fclose(f);
    }
    {   /* lv_extd: strings sit between guard bytes that can never match, so the result does not
         * depend on what the reference happens to find around its buffers */
        const uint64_t n = 13000;
        int32_t *len = malloc(n * 4), *res = malloc(n * 4);
        uint8_t *rf = calloc(n, 12), *qr = calloc(n, 12);
        for (uint64_t i = 0; i < n; i++) {
            uint8_t rb[40], qb[40];
            memset(rb, 0xE0, sizeof rb);
            memset(qb, 0xD0, sizeof qb);
            const int l = (int)(i % 13);
            len[i] = l;
            const int mode = (int)(rnd() % 4);         /* 0 random, 1..3 = a few edits */
            for (int j = 0; j < l; j++) rb[8 + j] = (uint8_t)(rnd() & 3);
            if (mode == 0) for (int j = 0; j < l; j++) qb[8 + j] = (uint8_t)(rnd() & 3);
            else {
                int o = 0;
                for (int j = 0; j < l && o < l; j++) {
                    const uint64_t e = rnd() % 24;
                    if (e == 0) continue;                                    /* deletion */
                    if (e == 1) { qb[8 + o++] = (uint8_t)(rnd() & 3); if (o >= l) break; }   /* insertion */
                    qb[8 + o++] = (e == 2) ? (uint8_t)((rb[8 + j] + 1) & 3) : rb[8 + j];
                }
                while (o < l) qb[8 + o++] = (uint8_t)(rnd() & 3);
            }
            memcpy(rf + i * 12, rb + 8, 12 > l ? (size_t)l : 12);
            memcpy(qr + i * 12, qb + 8, 12 > l ? (size_t)l : 12);
            res[i] = lv_extd(rb + 8, l, qb + 8, l);
        }
        FILE *f = out_open(out, "lv.bin");
        W(f, &n, 1); W(f, len, n); W(f, rf, n * 12); W(f, qr, n * 12); W(f, res, n);
        fclose(f);
    }
    {   /* MAPQ tables */
        static int q_mem[Q_MEM_MAX];
        static int q_lv[MAX_LV_WRONG][MAX_LV_R_LEN];
        uint64_t l_ref = idx->ref_bin.n * 4;
        calculate_MAPQ_TABLE(q_mem, q_lv, 0.15, l_ref);
        FILE *f = out_open(out, "mapq.bin");
        W(f, &l_ref, 1); W(f, q_mem, 2000);
        for (int i = 0; i < 20; i++) W(f, q_lv[i], 20);
        fclose(f);
    }
    {   /* get_ref */
        const uint64_t n = 2000;
        uint64_t off[2000];
        uint32_t len[2000];
        uint8_t fwd[2000], o[2000][16];
        memset(o, 0, sizeof o);
        for (uint64_t i = 0; i < n; i++) {
            off[i] = 64 + rnd() % (idx->ref_bin.n * 4 - 128);
            len[i] = (uint32_t)(rnd() % 17);
            fwd[i] = (uint8_t)(rnd() & 1);
            get_ref(idx->ref_bin.a, o[i], off[i], len[i], fwd[i]);
        }
        FILE *f = out_open(out, "getref.bin");
        W(f, &n, 1); W(f, off, n); W(f, len, n); W(f, fwd, n); W(f, &o[0][0], n * 16);
        fclose(f);
    }
    {   /* reads: encode, exist bits, islands through getIsland */
        FILE *fq = fopen(argv[2], "r");
        if (!fq) { perror(argv[2]); return 1; }
        FILE *f = out_open(out, "reads.bin");
        uint64_t n_reads = 0;
        W(f, &n_reads, 1);
        char *line = NULL, *seq = NULL;
        size_t cap = 0, scap = 0;
        const int k = idx->ek.len_e_kmer;
        Classify_buff_pool buff;
        memset(&buff, 0, sizeof buff);
        while ((int)n_reads < want_reads && getline(&line, &cap, fq) > 0) {
            if (getline(&seq, &scap, fq) <= 0) break;
            uint32_t L = (uint32_t)strlen(seq);
            while (L && (seq[L - 1] == '\n' || seq[L - 1] == '\r')) seq[--L] = 0;
            getline(&line, &cap, fq);
            getline(&line, &cap, fq);
            if (L < 40) continue;
            kseq_t ks;
            memset(&ks, 0, sizeof ks);
            ks.seq.s = seq;
            ks.seq.l = L;
            SEARCH_DIR sd[2];
            memset(sd, 0, sizeof sd);
            getIsland(&ks, &buff, &idx->ek, sd);
            W(f, &L, 1);
            W(f, buff.bin_read, 2 * L);
            const uint32_t nk = L - k + 1;
            uint8_t *ex = malloc(nk);
            for (int s = 0; s < 2; s++) {
                uint64_t *kb = buff.kmer_buff + (s ? L : 0);
                for (uint32_t i = 0; i < nk; i++)
                    ex[i] = (uint8_t)get_exist_kmer(idx->ek.e_kmer0, idx->ek.e_kmer1, kb[i], idx->ek.e_kmer_hash_mask);
                W(f, ex, nk);
            }
            free(ex);
            for (int want = 0; want < 2; want++) {           /* strand 0 = FORWARD first */
                SEARCH_DIR *d = (sd[0].direction == (want == 0 ? FORWARD : REVERSE)) ? &sd[0] : &sd[1];
                W(f, &d->total_score, 1);
                W(f, &d->l_seed_v_f, 1);
                for (uint32_t i = 0; i < d->l_seed_v_f; i++) {
                    uint32_t t[3] = {d->seed_v_f[i].offset, d->seed_v_f[i].len, d->seed_v_f[i].top};
                    W(f, t, 3);
                }
            }
            n_reads++;
        }
        fseek(f, 0, SEEK_SET);
        W(f, &n_reads, 1);
        fclose(f);
        fclose(fq);
    }
    fprintf(stderr, "kat_harness: vectors written to %s\n", out);
    return 0;
}
