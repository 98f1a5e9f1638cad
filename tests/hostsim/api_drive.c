/* ASAN / UBSAN driver of the three-symbol API over the host simulator (test infrastructure, not shipped) */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "desamba.h"
int dsb_meta_last_path(void);
static char *slurp(const char *p, uint64_t *n)
{
    FILE *f = fopen(p, "rb");
    if (!f) { perror(p); exit(1); }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    char *b = (char *)malloc((size_t)sz + 1);
    if (fread(b, 1, (size_t)sz, f) != (size_t)sz) exit(1);
    fclose(f); b[sz] = 0; *n = (uint64_t)sz;
    return b;
}
int main(int argc, char **argv)
{
    void *idx = NULL;
    load_index(&idx, argv[1]);
    uint64_t n;
    char *fq = slurp(argv[2], &n);
    for (int rep = 0; rep < 3; rep++) {
        char *out = NULL; uint64_t out_n = 0;
        read_classify(idx, fq, n, &out, &out_n, 3, 1);
        char *m1 = NULL, *h1 = NULL, *m2 = NULL, *h2 = NULL; uint64_t m1n = 0, h1n = 0, m2n = 0, h2n = 0;
        meta_analysis(idx, out, out_n, &m1, &m1n, 3, 1, 65536, &h1, &h1n);
        const int path1 = dsb_meta_last_path();
        char *copy = (char *)malloc(out_n + 1); memcpy(copy, out, out_n + 1);
        meta_analysis(idx, copy, out_n, &m2, &m2n, 3, 1, 65536, &h2, &h2n);
        const int path2 = dsb_meta_last_path();
        printf("rep %d: %llu bytes, meta paths %d %d, equal %d\n", rep, (unsigned long long)out_n, path1, path2,
               m1n == m2n && memcmp(m1, m2, m1n) == 0 && h1n == h2n && (!h1n || memcmp(h1, h2, h1n) == 0));
        free(out); free(copy); free(m1); free(h1); free(m2); free(h2);
    }
    free(fq);
    return 0;
}
