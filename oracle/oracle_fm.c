/* oracle_fm.c -- TEST INFRASTRUCTURE ONLY (see oracle_port.h).  FM-index rank, prefix-LUT backward
 * search, exist-k-mer probes and island scan, restated in plain C on the reference's layouts. */
#include "oracle_port.h"
#include <string.h>

/* reference src/bwt.c:43-65.  Block = 5 x u64 checkpoints (A,C,G,T,#) + 128 bytes of nibbles,
 * symbol i of the block in nibble (i&1) of byte i/2 (low nibble first, src/bwt.c:52). */
uint64_t orc_occ(const uint8_t *bwt_occ, uint64_t dollar_pos, uint64_t r, uint8_t *c)
{
    const uint8_t *blk = bwt_occ + (r >> 8) * 168;
    const uint8_t *data = blk + 40;
    const uint32_t in = (uint32_t)(r & 255);
    if (*c == 0xff) {
        *c = (uint8_t)((data[in >> 1] >> ((in & 1) << 2)) & 0xf);
        if (*c == 5) return dollar_pos;
    }
    uint64_t base;
    memcpy(&base, blk + 8 * (*c), 8);
    uint64_t k = 0;
    for (uint32_t i = 0; i < in; i++)
        if (((data[i >> 1] >> ((i & 1) << 2)) & 0xf) == *c) k++;
    return base + k;
}

uint64_t orc_hash64_1(uint64_t key)   /* reference src/lib/utils.c:1067-1078 */
{
    key = (~key + (key << 21));
    key = key ^ key >> 24;
    key = ((key + (key << 3)) + (key << 8));
    key = key ^ key >> 14;
    key = ((key + (key << 2)) + (key << 4));
    key = key ^ key >> 28;
    key = (key + (key << 31));
    return key;
}

uint64_t orc_hash64_2(uint64_t key)   /* reference src/lib/utils.c:1080-1091 */
{
    key += ~(key << 32);
    key ^= (key >> 22);
    key += ~(key << 13);
    key ^= (key >> 8);
    key += (key << 3);
    key ^= (key >> 15);
    key += ~(key << 27);
    key ^= (key >> 31);
    return key;
}

/* store_kmers (src/cly.c:359-397): a k-mer whose window holds >= 0.8k copies of one base becomes 0;
 * get_exist_kmer (src/cly.c:951-967): k-mer 0 does not exist; bit h&mask of table 0 (MSB-first
 * inside a byte) and, if set, bit of table 1. */
int64_t orc_exist_bits(const uint8_t *s, int64_t n, int k, const uint8_t *ek0, const uint8_t *ek1, uint64_t mask,
                       uint8_t *out)
{
    const int single_max = (int)(0.8 * k);      /* src/idx.c:981 */
    const uint64_t kmask = (k == 32) ? ~0ull : ((1ull << (2 * k)) - 1ull);
    int64_t nk = n - k + 1;
    if (nk <= 0) return 0;
    int cnt[4] = {0, 0, 0, 0};
    uint64_t v = 0;
    for (int64_t i = 0; i < n; i++) {
        const uint8_t b = s[i] & 3;
        v = ((v << 2) | b) & kmask;
        cnt[b]++;
        if (i >= k) cnt[s[i - k] & 3]--;
        if (i + 1 < k) continue;
        const int64_t p = i + 1 - k;
        uint64_t kmer = v;
        if (cnt[0] >= single_max || cnt[1] >= single_max || cnt[2] >= single_max || cnt[3] >= single_max) kmer = 0;
        uint8_t bit = 0;
        if (kmer != 0) {
            const uint64_t h1 = orc_hash64_1(kmer) & mask;
            if ((ek0[h1 >> 3] >> (7 - (h1 & 7))) & 1) {
                const uint64_t h2 = orc_hash64_2(kmer) & mask;
                bit = (uint8_t)((ek1[h2 >> 3] >> (7 - (h2 & 7))) & 1);
            }
        }
        out[p] = bit;
    }
    return nk;
}

/* search_exist_kmer_M2 (src/cly.c:1066-1155) replayed over precomputed exist bits, then the `top`
 * marking of get_seed_vector_M2 (src/cly.c:1157-1229). */
int64_t orc_islands(const uint8_t *ex, int64_t n, int strand, uint32_t *seeds, uint32_t *total_score)
{
    int64_t ns = 0;
    if (strand == 0) {
        for (int64_t i = 2; i < n; i += 3) {
            if (!ex[i]) continue;
            int64_t off = i, len = 1;
            for (int j = 1; j < 3; j++) { if (ex[i - j]) { off--; len++; } else break; }
            for (int64_t j = 1; i + j < n; j++) { if (ex[i + j]) { len++; if (len > 60) break; } else break; }
            seeds[3 * ns] = (uint32_t)off; seeds[3 * ns + 1] = (uint32_t)len; seeds[3 * ns + 2] = 0;
            ns++;
            i = off + len;
        }
    } else {
        for (int64_t i = n - 3; i >= 0; i -= 3) {
            if (!ex[i]) continue;
            int64_t off = i, len = 1;
            for (int j = 1; j < 3; j++) { if (ex[i + j]) { off++; len++; } else break; }
            for (int64_t j = 1; j <= i; j++) { if (ex[i - j]) { len++; if (len > 60) break; } else break; }
            seeds[3 * ns] = (uint32_t)(off - len + 1); seeds[3 * ns + 1] = (uint32_t)len; seeds[3 * ns + 2] = 0;
            ns++;
            i = off - len;
        }
    }
    uint32_t total = 0, max_len = 0, index_end = 100;
    int64_t max_idx = 0;
    for (int64_t m = 0; m < ns; m++) {
        const uint32_t key = (strand == 0) ? seeds[3 * m] : (uint32_t)(n - seeds[3 * m] - seeds[3 * m + 1]);
        if (key < index_end) {
            if (max_len < seeds[3 * m + 1]) { max_len = seeds[3 * m + 1]; max_idx = m; }
            seeds[3 * max_idx + 2] = 0;
        } else {
            seeds[3 * max_idx + 2] = 1;
            index_end += 100;
            total += max_len;
            max_idx = m;
            max_len = seeds[3 * m + 1];
        }
    }
    if (ns > 0) seeds[3 * max_idx + 2] = 1;
    total += max_len;
    *total_score = total;
    return ns;
}
