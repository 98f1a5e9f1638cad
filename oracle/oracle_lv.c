/* oracle_lv.c -- TEST INFRASTRUCTURE ONLY (see oracle_port.h).  Landau-Vishkin extension, 2-bit
 * reference fetch, MAPQ tables and the glibc merge-sort model, restated in plain C. */
#include "oracle_port.h"
#include <math.h>
#include <string.h>

#define LV_ERROR 4

/* reference src/cly.c:505-604.  Equal lengths (every call site passes len, len).  The reference
 * plants '#' / '$' after the strings in place; here both strings are copied into buffers whose
 * surroundings can never compare equal, which is the defined part of the reference's behaviour
 * (tests pin it with vectors whose reference result does not depend on neighbouring bytes). */
int32_t orc_lv_extd(const uint8_t *ref_in, const uint8_t *query_in, int32_t len)
{
    if (len <= 0) return 0;
    uint8_t rb[8 + 32], qb[8 + 32];
    memset(rb, 0xE0, sizeof rb);
    memset(qb, 0xD0, sizeof qb);
    uint8_t *ref = rb + 8, *query = qb + 8;
    memcpy(ref, ref_in, (size_t)len);
    memcpy(query, query_in, (size_t)len);
    ref[len] = '#';
    query[len] = '$';
    int32_t mn_d[16], ed_d[16];
    int32_t *mn = mn_d + LV_ERROR + 1, *ed = ed_d + LV_ERROR + 1;
    for (int i = -LV_ERROR - 1; i <= LV_ERROR + 2; i++) { mn[i] = -1; ed[i] = i > 0 ? i : -i; }
    int32_t best = len;
    for (int i = 0; i <= LV_ERROR; i++) {
        int32_t prev_mn = -1, cur_mn = i - 1, next_mn = mn[-i + 1];
        int32_t prev_ed = i + 1, cur_ed = i, next_ed = ed[-i + 1];
        for (int j = -i; j <= LV_ERROR; j++) {
            if (cur_mn + j < len - 1) {
                int m = cur_mn + 1 - cur_ed;
                mn[j] = cur_mn + 1; ed[j] = cur_ed + 1;
                if (m < next_mn + 1 - next_ed) { mn[j] = next_mn + 1; ed[j] = next_ed + 1; m = next_mn - next_ed; }
                if (m < prev_mn - prev_ed) { mn[j] = prev_mn + 1; ed[j] = prev_ed + 1; }
            } else {
                int m = cur_mn - cur_ed;
                mn[j] = cur_mn; ed[j] = cur_ed + 1;
                if (m < prev_mn - prev_ed) { mn[j] = prev_mn; ed[j] = prev_ed + 1; m = prev_mn - prev_ed; }
                if (m < next_mn + 1 - next_ed) { mn[j] = next_mn + 1; ed[j] = next_ed + 1; }
            }
            int mn_j = mn[j] < len ? mn[j] : len;
            if (mn_j > len - j) mn_j = len - j;
            while (ref[mn_j + j] == query[mn_j]) mn_j++;
            mn[j] = mn_j;
            if (query[mn_j] == '$' || ref[mn_j + j] == '#') {
                if (ed[j] - 1 < best) best = ed[j] - 1;
                if (j <= i + 1) return best;
            }
            prev_mn = cur_mn; cur_mn = next_mn; next_mn = mn[j + 2];
            prev_ed = cur_ed; cur_ed = next_ed; next_ed = ed[j + 2];
        }
    }
    return best;
}

/* reference src/cly.c:434-461: 4 bases per byte, first base in the top two bits */
void orc_get_ref(const uint8_t *ref_bin, uint8_t *dst, uint64_t off, uint32_t len, int forward)
{
    for (uint32_t k = 0; k < len; k++) {
        const uint64_t p = forward ? off + k : off - k;
        dst[k] = (uint8_t)((ref_bin[p >> 2] >> ((3 - (p & 3)) << 1)) & 3);
    }
}

/* reference src/cly_mt.c:396-420, same expressions in double */
void orc_mapq_tables(int32_t *q_mem, int32_t *q_lv, double P_E, uint64_t L_REF)
{
    double REF_SIZE_PUNALTY = -10 * log(L_REF) / log(10);
    double MATCH_SCORE = -10 * log(0.25 / (1 - P_E)) / log(10);
    double MISMATCH_PUNALTY = -10 * log(0.75 / (P_E)) / log(10);
    for (int i = 0; i < 2000; i++) q_mem[i] = REF_SIZE_PUNALTY + i * MATCH_SCORE + 0.5;
    for (int j = 0; j < 20; j++)
        for (int i = 0; i < 20; i++) {
            int v = (j - i) * MATCH_SCORE + i * MISMATCH_PUNALTY + 0.5;
            if (j < 5) v += 15;
            if (v < -8) v = -8;
            q_lv[i * 20 + j] = v;
        }
}

/* glibc 2.39 stdlib/msort.c msort_with_tmp: n1 = n/2, n2 = n - n1, recurse on both halves, merge
 * taking the left element when cmp(left, right) <= 0. */
static void msort_rec(uint32_t *b, uint32_t *tmp, uint32_t n, int (*cmp)(const void *, uint32_t, uint32_t),
                      const void *ctx)
{
    if (n <= 1) return;
    const uint32_t n1 = n / 2, n2 = n - n1;
    uint32_t *b1 = b, *b2 = b + n1;
    msort_rec(b1, tmp, n1, cmp, ctx);
    msort_rec(b2, tmp, n2, cmp, ctx);
    uint32_t i1 = 0, i2 = 0, o = 0;
    while (i1 < n1 && i2 < n2) {
        if (cmp(ctx, b1[i1], b2[i2]) <= 0) tmp[o++] = b1[i1++];
        else tmp[o++] = b2[i2++];
    }
    while (i1 < n1) tmp[o++] = b1[i1++];
    memcpy(b, tmp, (size_t)o * sizeof(uint32_t));
}

void orc_msort_idx(uint32_t *idx, uint32_t *tmp, uint32_t n, int (*cmp)(const void *ctx, uint32_t i, uint32_t j),
                   const void *ctx)
{
    for (uint32_t i = 0; i < n; i++) idx[i] = i;
    msort_rec(idx, tmp, n, cmp, ctx);
}
