/* oracle_port.h -- TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatement of the index-access and
 * seed-extension primitives of the reference's classify hot path, on the REFERENCE's own data
 * layouts (168-byte FM blocks, bit tables, 2-bit packed references).  Only tests/, bench.py's
 * cpu_baseline leg and __graft_entry__.smoke() may load this; the product never links it.
 *
 * Parity status: PINNED.  Every function below is checked (tests/test_oracle.py) against known-
 * answer vectors produced by the unmodified reference code itself (oracle/kat_harness.c includes
 * the reference's src/cly.c in place; vectors committed under tests/golden/ by tools/make_golden.py).
 * The end-to-end oracle is the unmodified reference binary oracle/_ref/deSAMBA (`classify -t 1`).
 */
#ifndef ORACLE_PORT_H
#define ORACLE_PORT_H
#include <stdint.h>

/* reference src/bwt.c:43-65 (occ) on the reference block layout src/bwt.c:32-41.
 * *c == 0xff: use the symbol at row r (returned through *c); returns dollar_pos if it is '$'.
 * Counts symbol *c in BWT[blk*256, r) plus the block checkpoint; the caller adds rank[*c]. */
uint64_t orc_occ(const uint8_t *bwt_occ, uint64_t dollar_pos, uint64_t r, uint8_t *c);

/* reference src/lib/utils.c:1067-1091 */
uint64_t orc_hash64_1(uint64_t key);
uint64_t orc_hash64_2(uint64_t key);

/* reference src/cly.c:359-397 (store_kmers: rolling k-mer + low-complexity mask) followed by
 * src/cly.c:951-967 (get_exist_kmer): exist bit of the k-mer starting at every position of the
 * 2-bit coded string s[0..n).  out[i] for i in [0, n-k].  Returns number of k-mers. */
int64_t orc_exist_bits(const uint8_t *s, int64_t n, int k, const uint8_t *ek0, const uint8_t *ek1, uint64_t mask,
                       uint8_t *out);

/* reference src/cly.c:1066-1155 + :1157-1229: islands of one strand from the exist bits
 * (strand 0 scans left-to-right, strand 1 right-to-left).  seeds = triplets {offset,len,top}.
 * Returns the number of islands; *total_score as get_seed_vector_M2 computes it. */
int64_t orc_islands(const uint8_t *ex, int64_t n_kmer, int strand, uint32_t *seeds, uint32_t *total_score);

/* reference src/cly.c:505-604 with equal lengths and private copies (sentinels never match,
 * bytes outside the strings never match). */
int32_t orc_lv_extd(const uint8_t *ref, const uint8_t *query, int32_t len);

/* reference src/cly.c:434-461 */
void orc_get_ref(const uint8_t *ref_bin, uint8_t *dst, uint64_t off, uint32_t len, int forward);

/* reference src/cly_mt.c:396-420 */
void orc_mapq_tables(int32_t *q_mem /*2000*/, int32_t *q_lv /*400*/, double p_e, uint64_t l_ref);

/* glibc 2.39 qsort model (msort.c: top-down merge sort, left element wins when cmp <= 0).
 * Sorts idx[0..n) = 0..n-1 with cmp(ctx, i, j). */
void orc_msort_idx(uint32_t *idx, uint32_t *tmp, uint32_t n, int (*cmp)(const void *ctx, uint32_t i, uint32_t j),
                   const void *ctx);

#endif
