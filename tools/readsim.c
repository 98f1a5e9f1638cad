/* readsim.c -- fast seeded read simulator for the large bench / parity data sets (test + bench infrastructure).
 *
 * tools/synth.py simulates reads in Python (0.7 ms per read); BASELINE.json configs[2] wants 10^5 .. 10^6 reads of
 * 10 kb, so the large data sets use this program instead: same model (uniform start, 50 % reverse strand, per-base
 * substitution / insertion / deletion errors, junk reads, length jitter), its own generator.  Every read is seeded from
 * (seed, read index), so the output does not depend on the number of threads and the first n reads of a larger set are
 * the set of n reads.
 *
 *   readsim ref.fa out.fq n_reads read_len err seed junk sub,ins,del jitter [prefix]
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t splitmix(uint64_t *s)
{
    uint64_t z = (*s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
static double uni(uint64_t *s) { return (double)(splitmix(s) >> 11) * (1.0 / 9007199254740992.0); }

typedef struct { char *seq; uint64_t len; } genome;

int main(int argc, char **argv)
{
    if (argc < 10) {
        fprintf(stderr, "usage: readsim ref.fa out.fq n_reads read_len err seed junk sub,ins,del jitter [prefix]\n");
        return 2;
    }
    const char *ref = argv[1], *out = argv[2];
    const long n_reads = atol(argv[3]);
    const long read_len = atol(argv[4]);
    const double err = atof(argv[5]);
    const uint64_t seed = strtoull(argv[6], NULL, 10);
    const double junk = atof(argv[7]);
    double ws = 1, wi = 1, wd = 1;
    sscanf(argv[8], "%lf,%lf,%lf", &ws, &wi, &wd);
    const double jitter = atof(argv[9]);
    const char *prefix = argc > 10 ? argv[10] : "r";
    const double wt = ws + wi + wd;
    const double ps = err * ws / wt, pi = err * wi / wt, pd = err * wd / wt;

    /* load the FASTA; sequences are compacted in place */
    FILE *fp = fopen(ref, "rb");
    if (!fp) { perror(ref); return 1; }
    fseek(fp, 0, SEEK_END);
    long fsz = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    char *buf = (char *)malloc((size_t)fsz + 1);
    if (fread(buf, 1, (size_t)fsz, fp) != (size_t)fsz) { fprintf(stderr, "short read\n"); return 1; }
    fclose(fp);
    buf[fsz] = 0;
    genome *g = NULL;
    size_t ng = 0, cap = 0;
    char *w = buf;
    for (long p = 0; p < fsz;) {
        if (buf[p] == '>') {
            while (p < fsz && buf[p] != '\n') p++;
            p++;
            if (ng == cap) { cap = cap ? cap * 2 : 1024; g = (genome *)realloc(g, cap * sizeof *g); }
            g[ng].seq = w;
            g[ng].len = 0;
            ng++;
            continue;
        }
        char *nl = (char *)memchr(buf + p, '\n', (size_t)(fsz - p));
        long e = nl ? nl - buf : fsz;
        if (ng) { memmove(w, buf + p, (size_t)(e - p)); w += e - p; g[ng - 1].len += (uint64_t)(e - p); }
        p = e + 1;
    }
    uint64_t total = 0;
    uint64_t *cum = (uint64_t *)malloc((ng + 1) * 8);
    for (size_t i = 0; i < ng; i++) { cum[i] = total; total += g[i].len; }
    cum[ng] = total;
    if (!ng || !total) { fprintf(stderr, "no sequence in %s\n", ref); return 1; }

    FILE *fo = fopen(out, "wb");
    if (!fo) { perror(out); return 1; }
    setvbuf(fo, NULL, _IOFBF, 1 << 22);
    static const char B[4] = {'A', 'C', 'G', 'T'};
    const long batch = 4096;
    char **text = (char **)calloc((size_t)batch, sizeof *text);
    size_t *tlen = (size_t *)calloc((size_t)batch, sizeof *tlen);
    size_t *tcap = (size_t *)calloc((size_t)batch, sizeof *tcap);
    uint64_t bases = 0;
    for (long b0 = 0; b0 < n_reads; b0 += batch) {
        long nb = n_reads - b0 < batch ? n_reads - b0 : batch;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : bases)
        for (long k = 0; k < nb; k++) {
            long r = b0 + k;
            uint64_t s = seed * 0x2545f4914f6cdd1dull + (uint64_t)r * 0x9e3779b97f4a7c15ull + 12345;
            splitmix(&s);
            long L = read_len;
            if (jitter > 0) {
                L = (long)(read_len * (1.0 + jitter * (2 * uni(&s) - 1)));
                if (L < 50) L = 50;
            }
            size_t need = (size_t)L * 4 + 256;     /* every base may insert one: 2 L bases + 2 L qualities */
            if (tcap[k] < need) { tcap[k] = need; text[k] = (char *)realloc(text[k], need); }
            char *t = text[k];
            int n = sprintf(t, "@%s%ld", prefix, r);
            long m = 0;                             /* bases emitted */
            char *sq;
            if (uni(&s) < junk) {
                n += sprintf(t + n, "_junk\n");
                sq = t + n;
                for (long i = 0; i < L; i++) sq[m++] = B[splitmix(&s) & 3];
            } else {
                /* genome chosen by length, start uniform, clear of the reference ends (the reference crashes on some
                 * chains there, reference src/cly.c:2719-2722) */
                uint64_t at = (uint64_t)(uni(&s) * (double)total);
                size_t lo = 0, hi = ng;
                while (lo + 1 < hi) { size_t mid = (lo + hi) / 2; if (cum[mid] <= at) lo = mid; else hi = mid; }
                const genome *gg = &g[lo];
                const long edge = 1000;
                long span = L;
                if ((uint64_t)(span + 2 * edge) > gg->len) span = (long)gg->len - 2 * edge;
                if (span < 50) span = gg->len < 50 ? (long)gg->len : 50;
                uint64_t room = gg->len > (uint64_t)(span + 2 * edge) ? gg->len - (uint64_t)(span + 2 * edge) + 1 : 1;
                uint64_t st = (gg->len >= (uint64_t)(span + 2 * edge) ? (uint64_t)edge : 0) + (uint64_t)(uni(&s) * (double)room);
                if (st + (uint64_t)span > gg->len) st = gg->len - (uint64_t)span;
                int rev = (splitmix(&s) >> 17) & 1;
                n += sprintf(t + n, "_g%zu_%llu_%c\n", lo, (unsigned long long)st, rev ? '-' : '+');
                sq = t + n;
                for (long i = 0; i < span; i++) {
                    char c;
                    if (!rev) c = gg->seq[st + (uint64_t)i];
                    else {
                        char x = gg->seq[st + (uint64_t)(span - 1 - i)];
                        c = x == 'A' ? 'T' : x == 'C' ? 'G' : x == 'G' ? 'C' : x == 'T' ? 'A' : x;
                    }
                    double u = uni(&s);
                    if (u < ps) {                                   /* substitution: one of the three other bases */
                        int cur = c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : 3;
                        sq[m++] = B[(cur + 1 + (int)(splitmix(&s) % 3)) & 3];
                    } else if (u < ps + pi) {                       /* insertion after the base */
                        sq[m++] = c;
                        sq[m++] = B[splitmix(&s) & 3];
                    } else if (u < ps + pi + pd) {                  /* deletion */
                    } else sq[m++] = c;
                }
                if (m == 0) sq[m++] = 'A';
            }
            char *e = sq + m;
            *e++ = '\n';
            *e++ = '+';
            *e++ = '\n';
            memset(e, 'I', (size_t)m);
            e += m;
            *e++ = '\n';
            tlen[k] = (size_t)(e - t);
            bases += (uint64_t)m;
        }
        for (long k = 0; k < nb; k++) fwrite(text[k], 1, tlen[k], fo);
    }
    fclose(fo);
    printf("reads=%ld bases=%llu\n", n_reads, (unsigned long long)bases);
    return 0;
}
