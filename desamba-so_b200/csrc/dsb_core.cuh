// dsb_core.cuh -- the per-read classifier as warp-uniform device code (sm_100a).
//
// Execution model: ONE WARP PER READ.  Every function below is written so that all 32 lanes of
// the warp execute the same instruction stream on the same values ("warp-uniform"): loads of a
// common address are one broadcast transaction, stores of a common value to a common address
// are benign, and control flow never diverges.  Inner loops that are worth spreading over the
// lanes are specialised under DSB_DEVICE_CODE (ballots / shuffles) and fall back to the plain
// loop otherwise.  Because the uniform form is ordinary C++, the same header also compiles with
// g++ for the host-side simulator in tests/hostsim (test infrastructure: it lets parity be
// debugged where no GPU exists; the product never links it).
//
// What it computes: the reference's classify_seq (reference src/cly.c:3059-3127) and everything
// below it, with the reference's C integer semantics kept type-for-type.
// Each function cites the reference lines it restates.  Data structures are our own:
// index-linked fixed arenas instead of pointers/kvec, a 64-symbol FM block read with 128-bit
// loads instead of the 168-byte block + 64 KiB popcount LUT, a direct (head,next) 9-mer table.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define DSB_HD __host__ __device__ __forceinline__
#define DSB_HDN __host__ __device__ __noinline__
// helpers that are called from many places: kept out of line on the device so that the classifier's
// instruction footprint stays small (ncu r01: 52 % of stall samples were instruction-cache misses)
#if defined(DSB_INLINE_ALL)
#define DSB_HDO __host__ __device__ __forceinline__
#else
#define DSB_HDO __host__ __device__ __noinline__
#endif
#else
#define DSB_HD static inline
#define DSB_HDN static
#define DSB_HDO static inline
#endif
#if defined(__CUDA_ARCH__)
#define DSB_DEVICE_CODE 1
#endif

// ------------------------------------------------------------------------------------------
// constants of the reference algorithm
// ------------------------------------------------------------------------------------------
#define DSB_FORWARD 1            // src/lib/utils.h:66
#define DSB_REVERSE 0            // src/lib/utils.h:67
#define DSB_L_PRE_IDX 13         // src/idx.h:58
#define DSB_PRE_IDX_MASK 0x3FFFFFFu
#define DSB_SA_MASK 0x7          // src/bwt.h:7
#define DSB_SA_OFF 3
#define DSB_MIN_UNI_L 35         // src/idx.h:40
#define DSB_LV_L 12              // src/cly.c:699
#define DSB_LV_ERROR 4           // src/cly.c:493
#define DSB_MIN_S_1 12           // src/cly.c:691
#define DSB_MIN_S_2 20           // src/cly.c:692
#define DSB_STEP_EK 3            // src/cly.c:969
#define DSB_SEED_RANGE 100       // src/cly.c (SEED_RANGE)
#define DSB_SP_SET_MAX 500       // src/cly.c:1482-1483
#define DSB_KMER9 9              // S_A_KEMR_L src/cly.c:1707
#define DSB_MIN_SCORE_MEM 12     // src/cly.c:1708
#define DSB_OVER_SEARCH 50       // OVER_SEARCH_M2 src/cly.c:2221
#define DSB_SMS_OVERLAP 6        // src/cly.c:2437-2438
#define DSB_PRIMARY 1
#define DSB_SECONDARY 2
#define DSB_SUPPLEMENTARY 3

#define DSB_HEAD_PAD 64          // bytes kept before F[0] (chunk-header model)
#define DSB_TAIL_PAD 64          // bytes kept after R[L-1] (buffer-history model)

// ------------------------------------------------------------------------------------------
// device-resident index view
// ------------------------------------------------------------------------------------------
// FM block: 64 BWT symbols in 64 bytes.  cnt[c] = number of symbol c (A,C,G,T) in BWT[0, 64*blk);
// plane[b] bit i = bit b of symbol i's 4-bit code (the reference's nibbles, src/bwt.c:52, transposed
// into bit planes at upload): "symbols equal to c among the first r" is then four XOR/ORs and one
// popcount on 64-bit words instead of nibble arithmetic on four words.
struct __attribute__((aligned(64))) DsbFmBlock {
    uint64_t cnt[4];
    uint64_t plane[4];
};

struct DsbIndexView {
    const DsbFmBlock *fm;
    uint64_t n_fm_blocks;
    uint64_t rank[6];
    uint64_t dollar_pos;          // what occ returns for '$' (src/bwt.c:53-54): DOLLOR_POS = unitig count - 2, NOT a BWT row
    uint64_t dollar_row;          // the BWT row that holds '$' (found while re-blocking); ~0 if none
    const uint64_t *lut;          // 2^26+1
    const uint64_t *sa;           // lo32 unitig_ID, hi32 offset
    uint64_t sa_n;
    const uint8_t *ek0, *ek1;
    uint64_t ek_mask;
    int32_t ek_len;               // k
    int32_t single_base_max;
    const uint64_t *unitig;       // lo32 ref_list, hi32 length
    uint64_t n_unitig;
    const uint64_t *ref_pos;      // bits 0-39 global_offset, 40-62 ref_ID, 63 direction
    const uint8_t *ref_bin;
    uint64_t ref_bin_n;
    const uint64_t *ref_off;
    const uint64_t *ref_len;
    uint64_t n_ref;
    const int32_t *q_mem;         // [2000]
    const int32_t *q_lv;          // [20][20]
    int32_t filter_min_length, filter_min_score, filter_min_score_lv3;
};

// ------------------------------------------------------------------------------------------
// per-read working structures (our layout; field meaning follows src/cly.h)
// ------------------------------------------------------------------------------------------
struct DsbSeed {                  // CLY_seed src/cly.h:27-32
    uint32_t offset;
    uint32_t len;
    uint32_t top;
};

struct DsbAnchor {                // Anchor + Anchor_map src/cly.h:34-61 (fields that are read back)
    uint32_t ref_ID;
    uint32_t ref_offset;
    uint32_t index_in_read;
    int32_t  pre;                 // chain_anchor_pre as an index, -1 = NULL
    uint16_t mtch_len;
    int16_t  score;
    uint8_t  direction;
    uint8_t  useless;             // anchor_useless
    uint8_t  duplicate;
    uint8_t  pad;
};

struct DsbChain {                 // chain_item src/cly.h:69-89
    uint32_t ref_ID;
    int32_t  q_t_dis;
    uint32_t sum_score;
    uint32_t anchor_number;
    uint32_t t_st, t_ed, q_st, q_ed;
    uint32_t indel;
    int32_t  cur;                 // chain_anchor_cur as an index
    uint8_t  direction;
    uint8_t  with_top;
    uint8_t  primary;
    uint8_t  pri_index;
};

struct DsbSms {                   // spd_match src/cly.h:128-134
    uint32_t t_pos, q_pos, len, score;
};

struct DsbMem {                   // MEM_rst src/cly.c:614-622
    int32_t  match_len;
    int32_t  sa_sp_l;
    uint64_t sp;
    uint64_t sa_sp;
    int32_t  read_offset;
    int32_t  pad;
};

struct DsbScHash {                // seed_con_hash src/cly.h:120-125
    uint16_t next;
    uint16_t seed_ID;             // 15 bits
    uint8_t  s_or_e;
    uint8_t  pad;
};

// One top island handed to one lane by the fast seeding stage (see dsb_fast_seeding)
struct DsbJob {
    uint32_t strand_ci;           // bit 31 strand (0 forward, 1 reverse), bits 0-30 island index
    uint32_t start;               // first anchor of this island in the owning lane's staging area
    uint32_t count;               // anchors it produced
    uint16_t lane;                // which lane's staging area
    uint16_t skip;                // the island ended with max_score > 512: the serial loop skips island ci+1
};

// Split pipeline (device): k_islands lists every read's top islands in one batch-wide job list, k_seed runs them ONE
// LANE PER ISLAND with lanes of a warp working on islands of different reads, k_classify picks the anchors up again.
struct DsbSeedJob {
    uint32_t read;                // read index in the batch
    uint32_t strand_ci;           // as DsbJob
    uint32_t offset, len;         // the island (DsbSeed)
    uint32_t a_off, count;        // its anchors in the batch's anchor pool (written by k_seed)
    uint32_t skip;                // as DsbJob
    uint32_t pad;
};
struct DsbReadSeeds {             // per read, written by k_islands
    uint32_t job_base, n_jobs;    // its jobs, in the serial order (dir0's islands, then dir1's)
    uint32_t total_score[2];
};

#define DSB_LANE_MEMS 8           // MEM slots of a lane working alone (fast mode keeps <= 2 rows per search)
#define DSB_LANE_PRIV_BYTES (512 * 8 + DSB_LANE_MEMS * 32)

struct DsbStats {
    unsigned long long occ, lut, sa, unitig, refpos, ref_bytes, mem_search, map_seed, sdp_win, slow;
    uint32_t ex0, ex1;            // exist-table probes of the on-demand island scan (the warp's)
    uint32_t slow_serial;         // strands of the slow path that took the warp-serial form as the fallback
    unsigned long long cyc[8];    // SM cycles per phase (device only; see dsb_counters.phase_cycles)
};

#if defined(__CUDA_ARCH__)
#define DSB_PHASE_BEGIN() long long dsb_t0_ = clock64()
#define DSB_PHASE_BEGIN2() dsb_t0_ = clock64()
#define DSB_PHASE_END(w, ph) do { const long long t_ = clock64(); (w)->st.cyc[ph] += (unsigned long long)(t_ - dsb_t0_); dsb_t0_ = t_; } while (0)
#define DSB_SUB_BEGIN() long long dsb_s0_ = clock64()
#define DSB_SUB_BEGIN2() dsb_s0_ = clock64()
#define DSB_SUB_END(w, ph) (w)->st.cyc[ph] += (unsigned long long)(clock64() - dsb_s0_)
#else
#define DSB_SUB_BEGIN() ((void)0)
#define DSB_SUB_BEGIN2() ((void)0)
#define DSB_SUB_END(w, ph) ((void)0)
#define DSB_PHASE_BEGIN() ((void)0)
#define DSB_PHASE_BEGIN2() ((void)0)
#define DSB_PHASE_END(w, ph) ((void)0)
#endif

// Everything one warp needs while it owns a read.  Pointers reference the warp's private slice
// of the global scratch arena (L1/L2 resident while in use).
struct DsbWork {
    const DsbIndexView *ix;
    // the read: bin[-DSB_HEAD_PAD .. 2L+DSB_TAIL_PAD), F at 0, R at L (src/cly.c:1242,1250)
    uint8_t *bin;
    uint32_t L;
    uint32_t is_2g;               // buff->max_read_l < 510 after this read (src/cly.c:2953-2955)
    const uint8_t *exist[2];      // [0] forward strand k-mer exist bytes, [1] reverse
    // islands
    DsbSeed *seeds[2];            // [0] forward, [1] reverse
    uint32_t n_seeds[2];
    uint32_t total_score[2];
    // anchors / chains
    DsbAnchor *anchors, *anchors_tmp;
    uint32_t n_anchor, cap_anchor;
    DsbChain *chains, *chains_tmp;
    uint32_t n_chain, cap_chain;
    uint32_t *sort_a, *sort_b;    // index buffers for the merge sort, cap_anchor each
    int32_t *score_v;             // 1024 ints (M3 DP)
    uint64_t *sp_set;             // 500
    DsbMem *mems;                 // 256
    // scoring
    DsbSms *sms;
    uint32_t n_sms, cap_sms;
    DsbSms *sms_stage;            // 32 lanes x DSB_SMS_LANE_CAP (dsb_sdp_match)
    uint32_t *h_head[2];          // [0] forward strand table, [1] reverse; 2^key_len entries
    uint32_t *h_ent[2];           // L entries: (next position + 1) << 8 | (9-mer >> key_len); one load per chain step
    DsbScHash *sc_hash;           // 256 + 2*400
    uint8_t *ref_mid;             // 2000+64  : sdp_middle window  (persistent per warp)
    uint8_t *ref_ext;             // 1000+64+64 : right / left window
    uint32_t fast_classify;
    uint32_t had_chain;           // the read reached delete_small_score_rst with >= 1 chain (src/cly.c:2884-2885)
    uint32_t overflow;            // set when an arena was too small -> batch is re-run bigger
    // fast seeding: islands are independent work items, one lane each (dsb_fast_seeding)
    uint32_t solo;                // 1 while this lane runs an island on its own (no warp-collective code)
    uint32_t *warp_ctr;           // the warp's job counter (shared memory on the device)
    int *smax;                    // 32 ints of shared memory for the lock-step map_seed (device)
    uint32_t slow_lane_mems;      // MEM slots of a lane of the slow path (0 = default)
    DsbJob *jobs;
    uint32_t cap_jobs;
    DsbAnchor *stage;             // 32 lanes x cap_stage anchors
    uint32_t cap_stage;
    uint8_t *lane_priv;           // 32 x DSB_LANE_PRIV_BYTES: a lane's own SP_SET + MEM slots
    DsbStats st;
};

// host-simulator tracing (tests/hostsim built with -DDSB_TRACE_ON): compiled out everywhere else
#if defined(DSB_TRACE_ON) && !defined(__CUDACC__)
#include <stdio.h>
#define DSB_TRACE(...) fprintf(stderr, __VA_ARGS__)
#else
#define DSB_TRACE(...) ((void)0)
#endif

#define DSB_MAX(a, b) (((a) > (b)) ? (a) : (b))     // src/lib/utils.h:61-64 (same macro shapes:
#define DSB_MIN(a, b) (((a) < (b)) ? (a) : (b))     //  mixed signed/unsigned operands convert
#define DSB_ABS(a) (((a) > 0) ? (a) : (-(a)))       //  exactly as in the reference)
#define DSB_ABS_U(a, b) (((a) > (b)) ? ((a) - (b)) : ((b) - (a)))

// ------------------------------------------------------------------------------------------
// FM index: rank query on the 64-symbol block (restates occ, src/bwt.c:43-65)
// ------------------------------------------------------------------------------------------
DSB_HD uint32_t dsb_popc64(uint64_t x)
{
#if defined(DSB_DEVICE_CODE)
    return (uint32_t)__popcll(x);
#else
    return (uint32_t)__builtin_popcountll(x);
#endif
}

// occ(bt, r, &c): *c == 0xff means "use the symbol at row r" (written back to *c).
// Returns the count of *c in BWT[0, r)  (r exclusive), or dollar_pos when
// the symbol at r is '$'.  The caller adds rank[*c].
DSB_HD uint64_t dsb_occ_ix(const DsbIndexView *ix, uint64_t r, uint8_t *c)
{
    const uint64_t blk = r >> 6;
    const uint32_t in = (uint32_t)(r & 63);
    const DsbFmBlock *b = ix->fm + blk;
    uint64_t p0, p1, p2, p3, cnt_c = 0, cnt_sum = 0;
    uint32_t cc = *c;
#if defined(DSB_DEVICE_CODE)
    const uint4 *p = reinterpret_cast<const uint4 *>(b);
    const uint4 v2 = __ldg(p + 2), v3 = __ldg(p + 3);
    p0 = ((uint64_t)v2.y << 32) | v2.x; p1 = ((uint64_t)v2.w << 32) | v2.z;
    p2 = ((uint64_t)v3.y << 32) | v3.x; p3 = ((uint64_t)v3.w << 32) | v3.z;
    if (cc == 0xff || cc >= 4) {
        // the count to use depends on the symbol found: fetch all four with the planes (one 64-byte line)
        const uint4 v0 = __ldg(p), v1 = __ldg(p + 1);
        cc = (cc == 0xff) ? (uint32_t)(((p0 >> in) & 1) | (((p1 >> in) & 1) << 1) | (((p2 >> in) & 1) << 2) | (((p3 >> in) & 1) << 3)) : cc;
        const uint64_t c0 = ((uint64_t)v0.y << 32) | v0.x, c1 = ((uint64_t)v0.w << 32) | v0.z;
        const uint64_t c2 = ((uint64_t)v1.y << 32) | v1.x, c3 = ((uint64_t)v1.w << 32) | v1.z;
        cnt_c = (cc == 0) ? c0 : (cc == 1) ? c1 : (cc == 2) ? c2 : c3;
        cnt_sum = c0 + c1 + c2 + c3;
    } else {
        cnt_c = __ldg(&b->cnt[cc]);
    }
#else
    p0 = b->plane[0]; p1 = b->plane[1]; p2 = b->plane[2]; p3 = b->plane[3];
    if (cc == 0xff)
        cc = (uint32_t)(((p0 >> in) & 1) | (((p1 >> in) & 1) << 1) | (((p2 >> in) & 1) << 2) | (((p3 >> in) & 1) << 3));
    cnt_c = b->cnt[cc & 3];
    cnt_sum = b->cnt[0] + b->cnt[1] + b->cnt[2] + b->cnt[3];
#endif
    if (*c == 0xff) {
        *c = (uint8_t)cc;
        if (cc == 5) return ix->dollar_pos;
    }
    uint64_t base;
    if (cc < 4) {
        base = cnt_c;
    } else {
        // '#': everything before the block that is not A/C/G/T/$.  The second LF walk of map_seed (src/cly.c:776-781)
        // does not stop at '#', so this rank IS consumed (rarely) and has to equal the reference's stored count.
        const uint64_t start = blk << 6;
        base = start - cnt_sum - ((ix->dollar_row < start) ? 1 : 0);
    }
    const uint64_t x = (p0 ^ (0ull - (uint64_t)(cc & 1))) | (p1 ^ (0ull - (uint64_t)((cc >> 1) & 1))) |
                       (p2 ^ (0ull - (uint64_t)((cc >> 2) & 1))) | (p3 ^ (0ull - (uint64_t)((cc >> 3) & 1)));
    const uint64_t low = in ? (~0ull >> (64 - in)) : 0ull;
    return base + dsb_popc64(~x & low);
}

// the classifier's out-of-line copy (counts the query)
DSB_HDO uint64_t dsb_occ(DsbWork *w, uint64_t r, uint8_t *c)
{
    w->st.occ++;
    return dsb_occ_ix(w->ix, r, c);
}

// ------------------------------------------------------------------------------------------
// 2-bit reference fetch (restates get_ref, src/cly.c:434-461)
// forward: bases uni_offset .. uni_offset+length-1; reverse: uni_offset down to uni_offset-length+1
// ------------------------------------------------------------------------------------------
DSB_HD uint8_t dsb_ref_base(const DsbIndexView *ix, uint64_t pos)
{
    const uint64_t byte = pos >> 2;
    // the reference reads whatever lies beyond the array; we define those bases as 'A'
    uint8_t v = (byte < ix->ref_bin_n) ? ix->ref_bin[byte] : (uint8_t)0;
    return (uint8_t)((v >> ((3 - (uint32_t)(pos & 3)) << 1)) & 3);
}

DSB_HDO void dsb_get_ref(DsbWork *w, uint8_t *dst, uint64_t uni_offset, uint32_t length, bool is_forward)
{
    const DsbIndexView *ix = w->ix;
    w->st.ref_bytes += (length + 3) >> 2;
    if (length == 0) return;
    // One packed byte = four bases: walk the SOURCE bytes (b0, b0+1, ... or b0, b0-1, ...) and scatter
    // each byte's bases to their places in dst.  Bytes outside the array read as 0 ('A'), which is also
    // what a wrapped (negative) position of the reference's unsigned arithmetic gives here.
    const uint32_t first = (uint32_t)(uni_offset & 3);
    const int64_t b0 = (int64_t)(uni_offset >> 2);
    const uint32_t nb = (((is_forward ? first : 3u - first) + length - 1u) >> 2) + 1u;
    uint32_t j0 = 0, jstep = 1;
#if defined(DSB_DEVICE_CODE)
    // Short fetches (the 12-base LV windows) land in thread-private arrays, so every lane fills its
    // own copy; long windows always land in the warp's scratch slice and are spread over the lanes.
    const bool spread = !(length <= 16 || w->solo);
    if (spread) { j0 = threadIdx.x & 31; jstep = 32; }
#endif
    for (uint32_t j = j0; j < nb; j += jstep) {
        const int64_t sb = is_forward ? b0 + (int64_t)j : b0 - (int64_t)j;
        const uint32_t v = (sb >= 0 && (uint64_t)sb < ix->ref_bin_n) ? (uint32_t)ix->ref_bin[sb] : 0u;
        for (uint32_t r = 0; r < 4; r++) {
            const uint32_t k = is_forward ? (4u * j + r - first) : (4u * j + first - r);     // wraps when out of range
            if (k < length) dst[k] = (uint8_t)((v >> ((3u - r) << 1)) & 3u);
        }
    }
#if defined(DSB_DEVICE_CODE)
    if (spread) __syncwarp();
#endif
}

// ------------------------------------------------------------------------------------------
// Landau-Vishkin semi-global extension with at most 4 edits (restates lv_extd, src/cly.c:505-604)
// The reference plants '#' / '$' sentinels in place and restores them; we work on private copies.
// Both strings always have the same length here (every call site passes len, len).
// Returns the edit distance, or `len` when no alignment with <= 4 edits exists.
//
// query_m1 = the byte BEFORE query[0] as the reference sees it.  The diagonal j = i + 1 starts at
// mn_j = -1 and compares ref[j - 1] with query[-1] (src/cly.c:589); when the query is a pointer into the
// read (the suffix side) that byte is the previous read base -- real data that decides e.g. whether a
// 1-base extension with a mismatch returns 0 or 1 -- so callers pass it.  When the query is one of
// map_seed's stack arrays the byte belongs to another stack object (reference-undefined): callers pass
// DSB_LV_NO_BYTE, which matches nothing, like the bytes before ref[0].
// ------------------------------------------------------------------------------------------
#define DSB_LV_NO_BYTE 0xD0
// Q_MEM has 2000 entries (src/idx.h:61); an exact match of 2000 bases or more makes the reference read past its table
// (whatever follows it on the heap; reference-undefined, and modelling glibc's chunk layout there did not
// reproduce it).  Here such a seed scores like a 1999-base one instead of faulting on the device.
#define DSB_QMEM_IX(l) ((uint32_t)(l) < 2000u ? (uint32_t)(l) : 1999u)
DSB_HDO int32_t dsb_lv_extd(const uint8_t *ref_in, const uint8_t *query_in, int32_t len, uint8_t query_m1)
{
    if (len <= 0) return 0;                   // returns 0 whatever the bytes are
    uint8_t rbuf[8 + DSB_LV_L + 2], qbuf[8 + DSB_LV_L + 2];
    uint8_t *ref = rbuf + 8, *query = qbuf + 8;
    for (int i = 0; i < 8; i++) { rbuf[i] = 0xE0; qbuf[i] = DSB_LV_NO_BYTE; }   // never equal to anything
    qbuf[7] = query_m1;
    for (int i = 0; i < len; i++) { ref[i] = ref_in[i]; query[i] = query_in[i]; }
    const int32_t ref_length = len, query_length = len;
    ref[ref_length] = '#';
    query[query_length] = '$';
    int32_t mn_d[14], ed_d[14];
    int32_t *mn = mn_d + DSB_LV_ERROR + 1, *ed = ed_d + DSB_LV_ERROR + 1;
    int32_t prev_mn, cur_mn, next_mn, prev_ed, cur_ed, next_ed;
    int32_t best_score = query_length;
    for (int i = -DSB_LV_ERROR - 1; i <= DSB_LV_ERROR + 1; i++) {
        mn[i] = -1;
        ed[i] = (i > 0) ? i : -i;
    }
    mn[DSB_LV_ERROR + 2] = -1; ed[DSB_LV_ERROR + 2] = DSB_LV_ERROR + 2;   // read but never used
#if defined(DSB_DEVICE_CODE)
#pragma unroll 1                              // 25 unrolled cells were 26 KB of code: the instruction cache matters more
#endif
    for (int i = 0; i <= DSB_LV_ERROR; i++) {
        prev_mn = -1;
        cur_mn = i - 1;
        next_mn = mn[-i + 1];
        prev_ed = i + 1;
        cur_ed = i;
        next_ed = ed[-i + 1];
#if defined(DSB_DEVICE_CODE)
#pragma unroll 1
#endif
        for (int j = -i; j <= DSB_LV_ERROR; j++) {
            if (cur_mn + j < ref_length - 1) {
                int m = cur_mn + 1 - cur_ed;
                mn[j] = cur_mn + 1;
                ed[j] = cur_ed + 1;
                if (m < next_mn + 1 - next_ed) {
                    mn[j] = next_mn + 1;
                    ed[j] = next_ed + 1;
                    m = next_mn - next_ed;
                }
                if (m < prev_mn - prev_ed) {
                    mn[j] = prev_mn + 1;
                    ed[j] = prev_ed + 1;
                }
            } else {
                int m = cur_mn - cur_ed;
                mn[j] = cur_mn;
                ed[j] = cur_ed + 1;
                if (m < prev_mn - prev_ed) {
                    mn[j] = prev_mn;
                    ed[j] = prev_ed + 1;
                    m = prev_mn - prev_ed;
                }
                if (m < next_mn + 1 - next_ed) {
                    mn[j] = next_mn + 1;
                    ed[j] = next_ed + 1;
                }
            }
            int mn_j = DSB_MIN(mn[j], query_length);
            mn_j = DSB_MIN(mn_j, ref_length - j);
            for (; ref[mn_j + j] == query[mn_j]; mn_j++) {}
            mn[j] = mn_j;
            if (query[mn_j] == '$' || ref[mn_j + j] == '#') {
                best_score = DSB_MIN(ed[j] - 1, best_score);
                if (j <= i + 1) return best_score;
            }
            prev_mn = cur_mn; cur_mn = next_mn; next_mn = mn[j + 2];
            prev_ed = cur_ed; cur_ed = next_ed; next_ed = ed[j + 2];
        }
    }
    return best_score;
}

// ------------------------------------------------------------------------------------------
// exist-k-mer probe (restates store_kmers src/cly.c:359-397, get_exist_kmer src/cly.c:951-967,
// hash64_1/2 src/lib/utils.c:1067-1091).  Evaluated for EVERY k-mer position by the exist
// kernel (one thread per position); the island scan below then replays the reference's
// data-dependent probe order over the precomputed bits, which is equivalent because the probe
// is a pure function of the k-mer.
// ------------------------------------------------------------------------------------------
DSB_HD uint64_t dsb_hash64_1(uint64_t key)
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

DSB_HD uint64_t dsb_hash64_2(uint64_t key)
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

// exist bit of the k-mer starting at s[0]; probes counts first/second table probes.
DSB_HD uint8_t dsb_exist_kmer(const DsbIndexView *ix, const uint8_t *s, uint32_t *probes0, uint32_t *probes1)
{
    const int k = ix->ek_len;
    int cnt[4] = {0, 0, 0, 0};
    uint64_t kmer = 0;
    for (int i = 0; i < k; i++) {
        const uint8_t b = s[i] & 3;
        cnt[b]++;
        kmer = (kmer << 2) | b;
    }
    const int m = ix->single_base_max;
    if (cnt[0] >= m || cnt[1] >= m || cnt[2] >= m || cnt[3] >= m) return 0;   // low complexity -> k-mer 0
    if (kmer == 0) return 0;
    const uint64_t h1 = dsb_hash64_1(kmer) & ix->ek_mask;
    (*probes0)++;
    if (((ix->ek0[h1 >> 3] >> (7 - (h1 & 7))) & 1) == 0) return 0;
    const uint64_t h2 = dsb_hash64_2(kmer) & ix->ek_mask;
    (*probes1)++;
    return (uint8_t)((ix->ek1[h2 >> 3] >> (7 - (h2 & 7))) & 1);
}

// ------------------------------------------------------------------------------------------
// islands (restates search_exist_kmer_M2 src/cly.c:1066-1155 and the `top` marking of
// get_seed_vector_M2 src/cly.c:1157-1229).  ex[i] = exist bit of k-mer i, n = L-k+1.
// ------------------------------------------------------------------------------------------
// top 1 island per 100 bp window of the ns islands in w->seeds[strand] (the `top` marking of get_seed_vector_M2
// src/cly.c:1157-1229), and the strand's score
DSB_HDO void dsb_islands_mark(DsbWork *w, int strand, uint32_t ns)
{
    const uint32_t n = w->L - (uint32_t)w->ix->ek_len + 1;
    DsbSeed *sv = w->seeds[strand];
    // top 1 island per 100 bp window (replayed literally)
    uint32_t total = 0, max_index = 0, max_length = 0, index_end = DSB_SEED_RANGE;
    for (uint32_t m = 0; m < ns; m++) {
        sv[m].top = 0;
        const uint32_t key = (strand == 0) ? sv[m].offset : (n - sv[m].offset - sv[m].len);
        if (key < index_end) {
            if (max_length < sv[m].len) { max_length = sv[m].len; max_index = m; }
            sv[max_index].top = 0;
        } else {
            sv[max_index].top = 1;
            index_end += DSB_SEED_RANGE;
            total += max_length;
            max_index = m;
            max_length = sv[m].len;
        }
    }
    sv[max_index].top = 1;      // also executed with ns == 0 in the reference (harmless there too)
    total += max_length;
    w->n_seeds[strand] = ns;
    w->total_score[strand] = total;
}

DSB_HDO void dsb_islands(DsbWork *w, int strand /*0 fwd, 1 rev*/)
{
    const uint8_t *ex = w->exist[strand];
    const uint32_t n = w->L - (uint32_t)w->ix->ek_len + 1;
    DsbSeed *sv = w->seeds[strand];
    uint32_t ns = 0;
#if defined(DSB_DEVICE_CODE)
    {
        // The scan visits every 3rd k-mer and restarts behind each island, so the visited positions are
        // data dependent -- but between islands it is a plain strided search for the next set byte: 32
        // candidates per step (ballot), and the run extension of a hit is a ballot over its neighbours.
        const int lane = (int)(threadIdx.x & 31);
        const int nn = (int)n;
        if (strand == 0) {
            for (int i = DSB_STEP_EK - 1; i < nn;) {
                const int c = i + DSB_STEP_EK * lane;
                const uint32_t m = __ballot_sync(0xffffffffu, c < nn && ex[c] == 1);
                if (!m) { i += DSB_STEP_EK * 32; continue; }
                i += DSB_STEP_EK * (__ffs((int)m) - 1);
                const int b1 = (ex[i - 1] == 1) ? 1 : 0;
                const int b2 = (b1 && ex[i - 2] == 1) ? 1 : 0;
                const int len0 = 1 + b1 + b2;
                int run = 0;
                for (int jb = 1;; jb += 32) {
                    const int p = i + jb + lane;
                    const uint32_t mm = __ballot_sync(0xffffffffu, p < nn && ex[p] == 1);
                    if (mm != 0xffffffffu) { run += __ffs((int)~mm) - 1; break; }
                    run += 32;
                    if (run >= 61 - len0) break;
                }
                const int len = len0 + DSB_MIN(run, 61 - len0);      // the reference stops right after len reaches 61
                const int offset = i - b1 - b2;
                sv[ns].offset = (uint32_t)offset;
                sv[ns].len = (uint32_t)len;
                ns++;
                i = offset + len + DSB_STEP_EK;
            }
        } else {
            for (int i = nn - DSB_STEP_EK; i >= 0;) {
                const int c = i - DSB_STEP_EK * lane;
                const uint32_t m = __ballot_sync(0xffffffffu, c >= 0 && ex[c] == 1);
                if (!m) { i -= DSB_STEP_EK * 32; continue; }
                i -= DSB_STEP_EK * (__ffs((int)m) - 1);
                const int b1 = (ex[i + 1] == 1) ? 1 : 0;
                const int b2 = (b1 && ex[i + 2] == 1) ? 1 : 0;
                const int len0 = 1 + b1 + b2;
                int run = 0;
                for (int jb = 1;; jb += 32) {
                    const int p = i - jb - lane;
                    const uint32_t mm = __ballot_sync(0xffffffffu, p >= 0 && ex[p] == 1);
                    if (mm != 0xffffffffu) { run += __ffs((int)~mm) - 1; break; }
                    run += 32;
                    if (run >= 61 - len0) break;
                }
                const int len = len0 + DSB_MIN(run, 61 - len0);
                const int offset = i + b1 + b2;
                sv[ns].offset = (uint32_t)(offset - len + 1);
                sv[ns].len = (uint32_t)len;
                ns++;
                i = offset - len - DSB_STEP_EK;
            }
        }
        __syncwarp();
    }
#else
    if (strand == 0) {
        for (uint32_t i = DSB_STEP_EK - 1; i < n; i += DSB_STEP_EK) {
            if (ex[i] != 1) continue;
            uint32_t offset = i, len = 1;
            for (int j = 1; j < DSB_STEP_EK; ++j) {
                if (ex[i - j] == 1) { offset--; len++; } else break;
            }
            for (uint32_t j = 1; i + j < n; ++j) {
                if (ex[i + j] == 1) { len++; if (len > 60) break; } else break;
            }
            sv[ns].offset = offset;
            sv[ns].len = len;
            ns++;
            i = offset + len;
        }
    } else {
        for (int i = (int)(n - DSB_STEP_EK); i >= 0; i -= DSB_STEP_EK) {
            if (ex[i] != 1) continue;
            uint32_t offset = (uint32_t)i, len = 1;
            for (int j = 1; j < DSB_STEP_EK; ++j) {
                if (ex[i + j] == 1) { offset++; len++; } else break;
            }
            for (int j = 1; j <= i; ++j) {
                if (ex[i - j] == 1) { len++; if (len > 60) break; } else break;
            }
            sv[ns].offset = offset - len + 1;
            sv[ns].len = len;
            ns++;
            i = (int)(offset - len);
        }
    }
#endif
    dsb_islands_mark(w, strand, ns);
}

#if defined(DSB_DEVICE_CODE)
// The island scan of BOTH strands with the exist tables probed on demand: no exist byte is computed for a k-mer the scan
// never looks at.  The reference probes every third k-mer and, around a hit, its neighbours (search_exist_kmer_M2
// src/cly.c:1066-1155); a probe is a pure function of the k-mer, so which positions are probed besides those does not
// change the islands.  Half a warp per strand, both halves in lock step (one converged probe per round):
//   lattice round : 16 candidates u, u+3, ..., u+45 -- none set: u += 48
//   first run round: the two k-mers before the hit and the 14 behind it
//   more run rounds: 16 further k-mers each, until one is clear or the island is 61 long
// u counts k-mers in scan direction: the reverse strand is scanned from its last k-mer downwards (u = n-1-i), which makes
// its scan the forward scan of the mirrored exist row.
__device__ __noinline__ void dsb_islands_probe(DsbWork *w)
{
    const int lane = (int)(threadIdx.x & 31);
    const int h = lane >> 4, tl = lane & 15;
    const int nn = (int)(w->L - (uint32_t)w->ix->ek_len + 1);
    const uint8_t *s = w->bin + (h ? w->L : 0);
    DsbSeed *sv = w->seeds[h];
    uint32_t ns = 0, p0 = 0, p1 = 0;
    int u = DSB_STEP_EK - 1;
    int mode = (u < nn) ? 0 : 3;          // 0 lattice, 1 first run round, 2 further run rounds, 3 done
    int ui = 0, b1 = 0, b2 = 0, run = 0;
    while (__any_sync(0xffffffffu, mode != 3)) {
        int q = -1;
        if (mode == 0) q = u + DSB_STEP_EK * tl;
        else if (mode == 1) q = (tl < 2) ? ui - 1 - tl : ui + tl - 1;
        else if (mode == 2) q = ui + 1 + run + tl;
        uint8_t e = 0;
        if (q >= 0 && q < nn) e = dsb_exist_kmer(w->ix, s + (h ? nn - 1 - q : q), &p0, &p1);
        const uint32_t hb = (__ballot_sync(0xffffffffu, e == 1) >> (16 * h)) & 0xffffu;
        bool close = false;
        if (mode == 0) {
            if (!hb) { u += DSB_STEP_EK * 16; if (u >= nn) mode = 3; }
            else { ui = u + DSB_STEP_EK * (__ffs((int)hb) - 1); mode = 1; }
        } else if (mode == 1) {
            b1 = (int)(hb & 1u);
            b2 = b1 & (int)((hb >> 1) & 1u);
            const uint32_t fw = hb >> 2;                  // 14 k-mers behind the hit
            if (fw == 0x3fffu) { run = 14; mode = 2; }
            else { run = __ffs((int)~fw) - 1; close = true; }
        } else if (mode == 2) {
            if (hb == 0xffffu) { run += 16; close = run >= 61 - (1 + b1 + b2); }
            else { run += __ffs((int)~hb) - 1; close = true; }
        }
        if (close) {
            const int len0 = 1 + b1 + b2;
            const int len = len0 + DSB_MIN(run, 61 - len0);       // the reference stops right after len reaches 61
            const int uoff = ui - b1 - b2;
            if (tl == 0) {
                sv[ns].offset = (uint32_t)(h ? nn - uoff - len : uoff);
                sv[ns].len = (uint32_t)len;
            }
            ns++;
            u = uoff + len + DSB_STEP_EK;
            mode = (u < nn) ? 0 : 3;
        }
    }
    w->st.ex0 += __reduce_add_sync(0xffffffffu, p0);      // the warp's probes, the same in every lane
    w->st.ex1 += __reduce_add_sync(0xffffffffu, p1);
    __syncwarp();
    const uint32_t ns0 = __shfl_sync(0xffffffffu, ns, 0), ns1 = __shfl_sync(0xffffffffu, ns, 16);
    dsb_islands_mark(w, 0, ns0);
    dsb_islands_mark(w, 1, ns1);
}
#endif

// islands of both strands: from the exist bytes where the caller computed them, else by probing on demand
DSB_HDO void dsb_islands_both(DsbWork *w)
{
#if defined(DSB_DEVICE_CODE)
    if (!w->exist[0]) { dsb_islands_probe(w); return; }
#endif
    dsb_islands(w, 0);
    dsb_islands(w, 1);
}

// ------------------------------------------------------------------------------------------
// FM seeding (restates sp_set_insert src/cly.c:1281-1293, bwt_single_search :1339-1378,
// bwt_MEM_search :1383-1442)
// ------------------------------------------------------------------------------------------
struct DsbSpSet { uint64_t *set; int l, m; int solo; };

// insert fails (returns 0) when the row is already present; the list silently empties when
// full
DSB_HDO int dsb_sp_set_insert(uint64_t node, DsbSpSet *s)
{
    if (s->l == s->m) s->l = 0;
    uint64_t *set = s->set;
#if defined(DSB_DEVICE_CODE)
    if (s->solo) {
        for (int i = 0; i < s->l; i++)
            if (set[i] == node) return 0;
        set[s->l] = node;
    } else {
        const uint32_t lane = threadIdx.x & 31;
        bool found = false;
        for (int i = lane; i < s->l; i += 32) found |= (set[i] == node);
        if (__any_sync(0xffffffffu, found)) return 0;
        if (lane == 0) set[s->l] = node;
        __syncwarp();
    }
#else
    for (int i = 0; i < s->l; i++)
        if (set[i] == node) return 0;
    set[s->l] = node;
#endif
    s->l++;
    return 1;
}

// single-row LF walk to the left while the BWT symbol equals the next read base.
// `str` points at the next base to compare (moving towards index 0).
DSB_HDO void dsb_single_search(DsbWork *w, uint64_t sp, const uint8_t *str, int max_match_len,
                              DsbSpSet *sp_set, DsbMem *m)
{
    const uint64_t *rank = w->ix->rank;
    uint64_t new_sp, sa_sp = 0xffffffffffffffffull;
    int match_len = 0, sa_sp_l = 0;
    while (1) {
        if (match_len >= max_match_len) break;
        if ((sp & DSB_SA_MASK) == 0) { sa_sp = sp; sa_sp_l = 0; }
        else sa_sp_l--;
        uint8_t c = 0xff;
        new_sp = dsb_occ(w, sp, &c);
        new_sp += rank[c];
        if (c != *str) break;
        match_len++;
        str--;
        if (dsb_sp_set_insert(new_sp, sp_set) == 0) { m->match_len = -1000; return; }
        sp = new_sp;
    }
    m->sp = sp;
    m->match_len = match_len;
    m->sa_sp = sa_sp;
    m->sa_sp_l = sa_sp_l;
}

// `str` points at the LAST base of the exist-k-mer; pre_v is the value of its last 13 bases.
DSB_HDO int dsb_mem_search(DsbWork *w, const uint8_t *str, uint64_t pre_v, int max_rst, int l_min_mth,
                          int l_max_mth, DsbSpSet *sp_set, DsbMem *mem)
{
    const uint64_t *rank = w->ix->rank;
    int n_rst = 0;
    w->st.mem_search++;
    w->st.lut++;
    uint64_t sp = w->ix->lut[pre_v], ep = w->ix->lut[pre_v + 1], new_sp, new_ep;
    str -= DSB_L_PRE_IDX;
    int match_len = DSB_L_PRE_IDX;
    uint8_t c;
    while (1) {
        c = *str;
        str--;
        new_sp = rank[c] + dsb_occ(w, sp, &c);
        new_ep = rank[c] + dsb_occ(w, ep, &c);
        if (match_len >= l_min_mth - 1) {
            if (new_sp + (uint64_t)(int64_t)max_rst >= new_ep) break;
            if (match_len >= l_max_mth) return 0;
        }
        if (new_sp + 1 >= new_ep) break;
        match_len++;
        sp = new_sp;
        ep = new_ep;
    }
    if (new_sp >= new_ep) return 0;
    if (new_sp + 1 == new_ep) {
        if (dsb_sp_set_insert(new_sp, sp_set) == 0) return 0;
        dsb_single_search(w, new_sp, str, DSB_MAX(0, l_max_mth - match_len), sp_set, mem + n_rst);
        mem[n_rst].match_len += match_len + 1;
        if (mem[n_rst].match_len >= l_min_mth) n_rst++;
    } else {
        for (uint64_t c_sp = new_sp; c_sp < new_ep; c_sp++) {
            if (dsb_sp_set_insert(c_sp, sp_set) == 0) continue;
            dsb_single_search(w, c_sp, str, DSB_MAX(0, l_max_mth - match_len), sp_set, mem + n_rst);
            mem[n_rst].match_len += match_len + 1;
            if (mem[n_rst].match_len >= l_min_mth) n_rst++;
        }
    }
    return n_rst;
}

// ------------------------------------------------------------------------------------------
// seed -> unitig -> reference anchors (restates get_uni src/cly.c:466-491, get_new_ed :624-689,
// map_seed :701-934)
// ------------------------------------------------------------------------------------------
// returns the unitig index; *global_offset / *uni_offset as in the reference
DSB_HDO uint64_t dsb_get_uni(DsbWork *w, uint64_t bwt_pos, int search_l, uint64_t *global_offset, uint32_t *uni_offset_)
{
    const DsbIndexView *ix = w->ix;
    const uint64_t sa = ix->sa[bwt_pos >> DSB_SA_OFF];
    w->st.sa++;
    uint64_t uni = (uint32_t)sa;
    uint32_t uni_offset = (uint32_t)(sa >> 32) + (uint32_t)search_l + 1;
    if (search_l > 0) {
        for (;;) {
            const uint32_t ulen = (uint32_t)(ix->unitig[uni] >> 32);
            w->st.unitig++;
            if (uni_offset < ulen) break;
            uni_offset -= (ulen + 1);
            uni++;
        }
    }
    // the reference's other branch (`uni_offset < 0` on an unsigned) is dead code
    const uint32_t ref_list = (uint32_t)ix->unitig[uni];
    w->st.unitig++;
    w->st.refpos++;
    *global_offset = (ix->ref_pos[ref_list] & 0xFFFFFFFFFFull) + uni_offset;
    *uni_offset_ = uni_offset;
    return uni;
}

// q_b = strand base pointer (index 0 = first base of the strand)
DSB_HDO void dsb_get_new_ed(DsbWork *w, uint32_t *e_d, uint32_t *len_, uint32_t *l_mem_ext, int32_t q_off,
                           uint64_t t_off, uint32_t l_read, const uint8_t *q_b, bool is_fwd)
{
    uint8_t q_buff[13], t_buff[13];
    const uint8_t *q = q_buff;
    uint8_t *t = t_buff;
    uint32_t len, max_len;
    if (is_fwd) {
        if (q_off < 0) q_off = 0;
        max_len = (uint32_t)q_off;
        len = DSB_MIN(12u, max_len);
        for (uint8_t k = 0; k < len; k++) q_buff[k] = q_b[q_off - k];
    } else {
        max_len = l_read - (uint32_t)q_off;
        len = DSB_MIN(12u, max_len);
        q = q_b + q_off;
    }
    dsb_get_ref(w, t, t_off, len, !is_fwd);
    if (len > 0 && t[0] == q[0]) {
        uint32_t mtc;
        do {
            for (mtc = 0; mtc < len; mtc++)
                if (t[mtc] != q[mtc]) break;
            if (mtc > 0) {
                *l_mem_ext += mtc;
                max_len -= mtc;
                len = DSB_MIN(12u, max_len);
                if (is_fwd) {
                    q_off -= (int32_t)mtc;
                    t_off -= mtc;
                    for (uint8_t k = 0; k < len; k++) q_buff[k] = q_b[q_off - k];
                } else {
                    t_off += mtc;
                    q += mtc;
                }
                dsb_get_ref(w, t, t_off, len, !is_fwd);
            }
        } while (mtc > 0);
    }
    *e_d = (uint32_t)dsb_lv_extd(t, q, (int32_t)len, is_fwd ? (uint8_t)DSB_LV_NO_BYTE : q[-1]);
    *len_ = len;
}

DSB_HD void dsb_push_anchor(DsbWork *w, const DsbAnchor *a)
{
    if (w->n_anchor >= w->cap_anchor) { w->overflow = 1; return; }
    w->anchors[w->n_anchor++] = *a;
}

// returns the best anchor score produced by this MEM
DSB_HDN int32_t dsb_map_seed(DsbWork *w, DsbMem *m_r, const uint8_t *q_b, uint32_t read_L, uint8_t direction)
{
    const DsbIndexView *ix = w->ix;
    const int32_t *Q_MEM = ix->q_mem;
    const int32_t *Q_LV = ix->q_lv;       // Q_LV[d][l] = Q_LV[d*20 + l]
    uint64_t b_p = m_r->sp;
    int32_t q_off = m_r->read_offset;
    uint32_t l_m = (uint32_t)m_r->match_len;
    bool have_uni = false;
    uint64_t uni = 0;
    uint32_t uni_len = 0;
    uint32_t u_off = 0;
    uint64_t t_off = 0;
    uint32_t l_pre = 0, l_suf = 0, d_pre = 0, d_suf = 0;
    int32_t s = 0, max_s = 0;
    w->st.map_seed++;
    do {
        uint8_t q_pre[DSB_LV_L + 1], t_pre[DSB_LV_L + 1], t_suf[DSB_LV_L + 1];
        const uint8_t *q_suf;
        // the reference leaves q_pre uninitialised when the MEM starts at read base 0 and then
        // compares q_pre[0] (src/cly.c:727-757 with l_pre = 0 -> s_l); we define it as a mismatch
        for (int k = 0; k <= DSB_LV_L; k++) { q_pre[k] = 0xF1; t_pre[k] = 0xF2; t_suf[k] = 0xF3; }
        l_pre = (uint32_t)DSB_MIN(q_off + 1, DSB_LV_L);
        for (uint8_t k = 0; k < l_pre; k++) q_pre[k] = q_b[q_off - k];
        int s_l = 0;
        if (m_r->sa_sp != 0xffffffffffffffffull) {
            uni = dsb_get_uni(w, m_r->sa_sp, m_r->sa_sp_l, &t_off, &u_off);
            have_uni = true;
        } else {
            uint8_t c;
            uint64_t new_sp;
            while (1) {
                if ((b_p & DSB_SA_MASK) == 0) break;
                c = 0xff;
                new_sp = dsb_occ(w, b_p, &c);
                new_sp += ix->rank[c];
                if (c == 4) break;
                t_pre[s_l++] = c;
                b_p = new_sp;
                if (s_l >= (int)l_pre) break;
            }
            if ((b_p & DSB_SA_MASK) == 0) {
                uni = dsb_get_uni(w, b_p, s_l, &t_off, &u_off);
                have_uni = true;
            } else
                l_pre = (uint32_t)s_l;
        }
        if (have_uni) {
            uni_len = (uint32_t)(ix->unitig[uni] >> 32);
            if (uni_len < DSB_MIN_UNI_L) break;
            l_pre = DSB_MIN(l_pre, u_off);
            dsb_get_ref(w, t_pre, t_off - 1, l_pre, false);
        }
        d_pre = (uint32_t)dsb_lv_extd(t_pre, q_pre, (int32_t)l_pre, DSB_LV_NO_BYTE);
        s = Q_MEM[DSB_QMEM_IX(l_m)] + Q_LV[d_pre * 20 + l_pre];
        if (s < DSB_MIN_S_1 && l_pre == DSB_LV_L && !have_uni) { s = 0; break; }
        if (!have_uni) {
            while (b_p & DSB_SA_MASK) {
                uint8_t c = 0xff;
                b_p = dsb_occ(w, b_p, &c);
                b_p += ix->rank[c];
                s_l++;
            }
            uni = dsb_get_uni(w, b_p, s_l, &t_off, &u_off);
            have_uni = true;
            uni_len = (uint32_t)(ix->unitig[uni] >> 32);
            if (uni_len < DSB_MIN_UNI_L) { s = 0; break; }
        }
        int32_t q_off_r = q_off + (int32_t)l_m + 1;
        uint32_t l_max_suf = DSB_MIN(uni_len - u_off - l_m, read_L - (uint32_t)q_off_r);
        if (l_max_suf != 0) {
            l_suf = DSB_MIN(l_max_suf, (uint32_t)DSB_LV_L);
            q_suf = q_b + q_off_r;
            dsb_get_ref(w, t_suf, t_off + l_m, l_suf, true);
            if (t_suf[0] == q_suf[0]) {
                uint32_t mtc;
                do {
                    for (mtc = 0; mtc < l_suf; mtc++)
                        if (t_suf[mtc] != q_suf[mtc]) break;
                    if (mtc > 0) {
                        l_m += mtc;
                        s = Q_MEM[DSB_QMEM_IX(l_m)] + Q_LV[d_pre * 20 + l_pre];
                        l_max_suf -= mtc;
                        l_suf = DSB_MIN(l_max_suf, (uint32_t)DSB_LV_L);
                        q_suf += mtc;
                        dsb_get_ref(w, t_suf, t_off + l_m, l_suf, true);
                    }
                } while (mtc > 0);
            }
            d_suf = (uint32_t)dsb_lv_extd(t_suf, q_suf, (int32_t)l_suf, q_suf[-1]);
            s += Q_LV[d_suf * 20 + l_suf];
        } else
            l_suf = d_suf = 0;
        if (s <= DSB_MIN_S_2 && l_suf == DSB_LV_L) { s = 0; break; }
    } while (0);

    if (s > 0) {
        // Anchor_map a_m = {l_m, s, l_pre, d_pre, l_suf, d_suf} with the reference's field widths
        uint16_t am_mtch = (uint16_t)l_m;
        int16_t am_score = (int16_t)s;
        uint8_t am_left_len = (uint8_t)l_pre, am_left_ED = (uint8_t)d_pre;
        uint8_t am_rigt_len = (uint8_t)l_suf, am_rigt_ED = (uint8_t)d_suf;
        const uint32_t rp_s = (uint32_t)ix->unitig[uni], rp_e = (uint32_t)ix->unitig[uni + 1];
        w->st.unitig++;
        const bool ref_search_l = (l_pre < DSB_LV_L || d_pre == 0);
        const bool ref_search_r = (l_suf < DSB_LV_L || d_suf == 0);
        if ((int64_t)rp_e - (int64_t)rp_s > 50) {
            if (!((int64_t)rp_e - (int64_t)rp_s < 1000)) return 50;
        }
        for (uint32_t rp = rp_s; rp < rp_e; rp++) {
            const uint64_t rpv = ix->ref_pos[rp];
            w->st.refpos++;
            const uint64_t r_global = rpv & 0xFFFFFFFFFFull;
            const uint32_t r_ref_ID = (uint32_t)((rpv >> 40) & 0x7FFFFF);
            uint32_t ed_l, ed_r, len_l, len_r;
            uint32_t l_m_ext_l = 0, l_m_ext_r;
            if (ref_search_l || ref_search_r) {
                if (ref_search_l) {
                    dsb_get_new_ed(w, &ed_l, &len_l, &l_m_ext_l, q_off, r_global + u_off - 1, read_L, q_b, true);
                    am_left_len = (uint8_t)len_l;
                    am_left_ED = (uint8_t)ed_l;
                }
                am_mtch = (uint16_t)(l_m + l_m_ext_l);
                if (ref_search_r) {
                    l_m_ext_r = 0;
                    dsb_get_new_ed(w, &ed_r, &len_r, &l_m_ext_r, q_off + (int32_t)l_m + 1, r_global + u_off + l_m,
                                   read_L, q_b, false);
                    am_rigt_len = (uint8_t)len_r;
                    am_rigt_ED = (uint8_t)ed_r;
                    am_mtch = (uint16_t)(am_mtch + l_m_ext_r);
                }
                am_score = (int16_t)(Q_MEM[DSB_QMEM_IX(am_mtch)] + Q_LV[am_left_ED * 20 + am_left_len] +
                                     Q_LV[am_rigt_ED * 20 + am_rigt_len]);
                if (am_score < DSB_MIN_S_2) continue;
            }
            max_s = DSB_MAX(max_s, (int32_t)am_score);
            DsbAnchor a;
            a.direction = direction;
            a.index_in_read = (uint32_t)(q_off + 1) - l_m_ext_l;
            const uint64_t g = r_global + u_off - l_m_ext_l;
            a.ref_ID = r_ref_ID;
            a.ref_offset = (uint32_t)(g - ix->ref_off[r_ref_ID]);
            a.mtch_len = am_mtch;
            a.score = am_score;
            a.pre = -1;
            a.useless = 0;
            a.duplicate = 0;
            a.pad = 0;
            dsb_push_anchor(w, &a);
        }
    }
    return max_s;
}

// ------------------------------------------------------------------------------------------
// seeding drivers (restates fast_classify src/cly.c:1473-1541, slow_classify :1545-1606)
// ------------------------------------------------------------------------------------------
// value of the 13 bases ending at str[0] (== kmer & PRE_IDX_MASK for every k-mer that passed the
// complexity filter, which all island k-mers did)
DSB_HD uint64_t dsb_prefix13(const uint8_t *str)
{
    uint64_t v = 0;
    for (int i = DSB_L_PRE_IDX - 1; i >= 0; i--) v = (v << 2) | (uint64_t)(str[-i] & 3);
    return v;
}

DSB_HD void dsb_mark_useless(DsbWork *w, uint32_t a_b, uint32_t a_e)
{
    int top_score = 35;
    for (uint32_t i = a_b; i < a_e; i++) top_score = DSB_MAX(top_score, (int)w->anchors[i].score);
    for (uint32_t i = a_b; i < a_e; i++) w->anchors[i].useless = (w->anchors[i].score < top_score) ? 1 : 0;
}

// One top island of fast_classify (the body of the reference's loop over `c_i`, src/cly.c:1489-1538).
// Islands are independent of each other: SP_SET restarts empty for each one and the anchors are
// only appended.  The single coupling is `max_score > 512` making the serial loop skip island
// ci+1; it is returned as a flag and applied when the islands' anchors are merged in order.
// Pushes into w->anchors / w->n_anchor (whatever they currently point at).
DSB_HDN uint32_t dsb_fast_island(DsbWork *w, int strand, uint32_t ci, DsbSpSet *sp_set, DsbMem *m_r)
{
    const int l_ek = w->ix->ek_len;
    const int min_index = 21 - l_ek;                      // MIN_MEM_LEN_FAST - l_ek
    const uint8_t *bin_read = w->bin + (strand ? w->L : 0);
    const uint8_t direction = strand ? DSB_REVERSE : DSB_FORWARD;
    const DsbSeed sd = w->seeds[strand][ci];
    uint32_t skip_next = 0;
    sp_set->l = 0;
    const uint32_t a_b_idx = w->n_anchor;
    for (int j = (int)sd.len - 1; j >= min_index;) {
        const int kmer_index = (int)sd.offset + j;
        const int string_index = kmer_index + l_ek - 1;
        const uint64_t pre_v = dsb_prefix13(bin_read + string_index);
        const int n_mem = dsb_mem_search(w, bin_read + string_index, pre_v, 2, 21 - 1, string_index, sp_set, m_r);
        DSB_TRACE("FAST strand=%d island=%u j=%d sidx=%d n_mem=%d\n", strand, ci, j, string_index, n_mem);
        if (n_mem == 0) { j -= 2; continue; }
        j -= 3;
        int max_score = 0;
        for (int mi = 0; mi < n_mem; mi++) {
            m_r[mi].read_offset = string_index - m_r[mi].match_len;
            const int c_score = dsb_map_seed(w, m_r + mi, bin_read, w->L, direction);
            DSB_TRACE("  MEM len=%d roff=%d sp=%llu score=%d\n", m_r[mi].match_len, m_r[mi].read_offset,
                      (unsigned long long)m_r[mi].sp, c_score);
            max_score = DSB_MAX(c_score, max_score);
        }
        if (max_score > 35) j -= 7;
        if (max_score > 256) {
            if (max_score > 512) skip_next = 1;
            break;
        }
    }
    dsb_mark_useless(w, a_b_idx, w->n_anchor);
    return skip_next;
}

// fast_classify over one or two strands (src/cly.c:1473-1541, called at :3093-3095).
// Stage 1 lists the top islands in the serial order (dir0's, then dir1's).  Stage 2 runs them as
// independent jobs, ONE LANE PER ISLAND on the device (each lane pulls the next job from the warp's
// counter and works on private SP_SET / MEM slots / anchor staging; 32 dependent FM walks are in
// flight per warp instead of one); the host build runs the same jobs one after the other, cycling
// through the 32 staging areas so that the merge below is exercised identically.  Stage 3 appends
// the staged anchors in job order and drops the islands the serial loop would have skipped.
DSB_HDN void dsb_fast_seeding(DsbWork *w, int dir0, int dir1 /* -1: one strand only */)
{
    DsbJob *jobs = w->jobs;
    uint32_t n_jobs = 0;
    for (int pass = 0; pass < 2; pass++) {
        const int strand = pass ? dir1 : dir0;
        if (strand < 0) continue;
        const DsbSeed *sv = w->seeds[strand];
        const uint32_t nsv = w->n_seeds[strand];
        for (uint32_t ci = 0; ci < nsv; ci++) {
            if (sv[ci].top == 0) continue;
            if (n_jobs >= w->cap_jobs) { w->overflow = 1; return; }
            jobs[n_jobs].strand_ci = ((uint32_t)strand << 31) | ci;
            n_jobs++;
        }
    }
    if (n_jobs == 0) return;
    DsbAnchor *const main_anchors = w->anchors;
    const uint32_t main_n = w->n_anchor, main_cap = w->cap_anchor;
#if defined(DSB_DEVICE_CODE)
    {
        const uint32_t lane = threadIdx.x & 31;
        __syncwarp();                                     // the job list (written by every lane alike) is complete
        if (lane == 0) *w->warp_ctr = 0;
        __syncwarp();
        const DsbStats snap = w->st;
        uint8_t *priv = w->lane_priv + (size_t)lane * DSB_LANE_PRIV_BYTES;
        DsbSpSet sp_set = {(uint64_t *)priv, 0, DSB_SP_SET_MAX, 1};
        DsbMem *m_r = (DsbMem *)(priv + 512 * 8);
        w->anchors = w->stage + (size_t)lane * w->cap_stage;
        w->cap_anchor = w->cap_stage;
        w->n_anchor = 0;
        w->solo = 1;
        // The lanes run their islands in lock-step ROUNDS (one round = one bwt_MEM_search of the island's
        // k-mer loop plus the map_seed calls it triggers, i.e. one iteration of the loop of
        // dsb_fast_island): all lanes with work enter dsb_mem_search / dsb_map_seed together, so the
        // instruction stream is shared instead of 32 private copies fighting for the instruction cache,
        // while the 32 FM walks still overlap their memory latency.  A lane whose island is finished
        // pulls the next job at the round boundary.
        const int l_ek = w->ix->ek_len;
        const int min_index = 21 - l_ek;
        bool have = false;
        uint32_t jb = 0, j_strand = 0, j_ci = 0, start = 0, skip_next = 0;
        int j = 0;
        DsbSeed sd = {0, 0, 0};
        for (;;) {
            if (!have) {
                jb = atomicAdd(w->warp_ctr, 1u);
                if (jb < n_jobs) {
                    const uint32_t key = jobs[jb].strand_ci;
                    j_strand = key >> 31;
                    j_ci = key & 0x7fffffffu;
                    sd = w->seeds[j_strand][j_ci];
                    j = (int)sd.len - 1;
                    start = w->n_anchor;
                    skip_next = 0;
                    sp_set.l = 0;
                    have = true;
                }
            }
            if (!__any_sync(0xffffffffu, have)) break;
            bool finish = false;
            int n_mem = 0, string_index = 0;
            const uint8_t *bin_read = w->bin + (j_strand ? w->L : 0);
            if (have) {
                if (j < min_index) finish = true;
                else {
                    string_index = (int)sd.offset + j + l_ek - 1;
                    const uint64_t pre_v = dsb_prefix13(bin_read + string_index);
                    n_mem = dsb_mem_search(w, bin_read + string_index, pre_v, 2, 21 - 1, string_index, &sp_set, m_r);
                }
            }
            __syncwarp();
            int max_score = 0;
            for (int mi = 0; mi < 2; mi++) {
                if (have && mi < n_mem) {
                    m_r[mi].read_offset = string_index - m_r[mi].match_len;
                    const int c_score = dsb_map_seed(w, m_r + mi, bin_read, w->L, j_strand ? DSB_REVERSE : DSB_FORWARD);
                    max_score = DSB_MAX(c_score, max_score);
                }
                __syncwarp();
            }
            if (have && !finish) {
                if (n_mem == 0) j -= 2;
                else {
                    j -= 3;
                    if (max_score > 35) j -= 7;
                    if (max_score > 256) {
                        if (max_score > 512) skip_next = 1;
                        finish = true;
                    }
                }
            }
            if (have && finish) {
                dsb_mark_useless(w, start, w->n_anchor);
                jobs[jb].start = start;
                jobs[jb].count = w->n_anchor - start;
                jobs[jb].lane = (uint16_t)lane;
                jobs[jb].skip = (uint16_t)skip_next;
                have = false;
            }
        }
        w->solo = 0;
        __syncwarp();
        // every lane counted its own work: fold the counters back into the warp-uniform totals
#define DSB_FOLD(f) w->st.f = snap.f + __reduce_add_sync(0xffffffffu, (unsigned)(w->st.f - snap.f))
        DSB_FOLD(occ); DSB_FOLD(lut); DSB_FOLD(sa); DSB_FOLD(unitig); DSB_FOLD(refpos); DSB_FOLD(ref_bytes);
        DSB_FOLD(mem_search); DSB_FOLD(map_seed);
#undef DSB_FOLD
        w->overflow = __any_sync(0xffffffffu, w->overflow != 0) ? 1u : 0u;
    }
#else
    {
        uint32_t n_priv[32];
        for (int l = 0; l < 32; l++) n_priv[l] = 0;
        DsbSpSet sp_set = {(uint64_t *)w->lane_priv, 0, DSB_SP_SET_MAX, 1};
        DsbMem *m_r = (DsbMem *)(w->lane_priv + 512 * 8);
        for (uint32_t j = 0; j < n_jobs; j++) {
            const uint32_t lane = (j * 7u) & 31u;         // any assignment must give the same result
            const uint32_t key = jobs[j].strand_ci;
            w->anchors = w->stage + (size_t)lane * w->cap_stage;
            w->cap_anchor = w->cap_stage;
            w->n_anchor = n_priv[lane];
            const uint32_t start = w->n_anchor;
            const uint32_t skip = dsb_fast_island(w, (int)(key >> 31), key & 0x7fffffffu, &sp_set, m_r);
            jobs[j].start = start;
            jobs[j].count = w->n_anchor - start;
            jobs[j].lane = (uint16_t)lane;
            jobs[j].skip = (uint16_t)skip;
            n_priv[lane] = w->n_anchor;
        }
    }
#endif
    w->anchors = main_anchors;
    w->n_anchor = main_n;
    w->cap_anchor = main_cap;
    if (w->overflow) return;
    uint32_t skip_key = 0xffffffffu;
    for (uint32_t j = 0; j < n_jobs; j++) {
        const DsbJob jb = jobs[j];
        if (jb.strand_ci == skip_key) { skip_key = 0xffffffffu; continue; }   // `ci++` of src/cly.c:1533
        skip_key = jb.skip ? jb.strand_ci + 1 : 0xffffffffu;
        if (jb.count == 0) continue;
        if (w->n_anchor + jb.count > w->cap_anchor) { w->overflow = 1; return; }
        const DsbAnchor *src = w->stage + (size_t)jb.lane * w->cap_stage + jb.start;
        DsbAnchor *dst = w->anchors + w->n_anchor;
#if defined(DSB_DEVICE_CODE)
        for (uint32_t k = threadIdx.x & 31; k < jb.count; k += 32) dst[k] = src[k];
#else
        for (uint32_t k = 0; k < jb.count; k++) dst[k] = src[k];
#endif
        w->n_anchor += jb.count;
    }
#if defined(DSB_DEVICE_CODE)
    __syncwarp();
#endif
}

// stable insertion sort of the MEM list by match_len descending: equivalent to the reference's
// qsort (glibc merge sort, consistent comparator MEM_rst_cmp_by_match_len src/cly.c:1325-1328)
DSB_HD void dsb_sort_mems(DsbMem *m, int n)
{
    for (int i = 1; i < n; i++) {
        DsbMem key = m[i];
        int j = i - 1;
        while (j >= 0 && m[j].match_len < key.match_len) { m[j + 1] = m[j]; j--; }
        m[j + 1] = key;
    }
}

#if defined(__CUDACC__)
// slow_classify with one lane per island, in lock step (dsb_cuda.cu, next to the lock-step seeding primitives it uses);
// false = the strand has to go through the serial form below, nothing was changed
#define DSB_SLOW_LANE_MEMS 120
__device__ bool dsb_slow_seeding_lanes(DsbWork *w, int strand);
#endif

DSB_HDN void dsb_slow_classify(DsbWork *w, int strand)
{
    const int l_ek = w->ix->ek_len;
    const uint8_t *bin_read = w->bin + (strand ? w->L : 0);
    const uint8_t direction = strand ? DSB_REVERSE : DSB_FORWARD;
#if defined(DSB_DEVICE_CODE) && !defined(DSB_SLOW_SERIAL)
    if (dsb_slow_seeding_lanes(w, strand)) { w->fast_classify = 0; return; }
    w->st.slow_serial++;
#endif
    DsbSpSet sp_set = {w->sp_set, 0, DSB_SP_SET_MAX, 0};
    DsbMem *mem_rst = w->mems;
    const DsbSeed *sv = w->seeds[strand];
    const uint32_t nsv = w->n_seeds[strand];
    for (uint32_t i = 0; i < nsv; i++) {
        if ((int)(sv[i].len) < 3 && sv[0].top == 0) continue;          // first seed's top
        const int min_match_len = DSB_MIN(20 - 1, l_ek + 1);
        sp_set.l = 0;
        int mem_rst_num = 0;
        for (int j = (int)sv[i].len - 1; j >= 1; j -= 2) {
            const int k_idx = (int)sv[i].offset + j;
            const int s_idx = k_idx + l_ek - 1;
            const uint64_t pre_v = dsb_prefix13(bin_read + s_idx);
            const int n_mem = dsb_mem_search(w, bin_read + s_idx, pre_v, 8, min_match_len, s_idx, &sp_set,
                                             mem_rst + mem_rst_num);
            for (int t = mem_rst_num; t < mem_rst_num + n_mem; t++)
                mem_rst[t].read_offset = k_idx + l_ek - 1 - mem_rst[t].match_len;
            mem_rst_num += n_mem;
        }
        if (mem_rst_num == 0) continue;
        if (mem_rst_num > 1) dsb_sort_mems(mem_rst, mem_rst_num);
        const uint32_t a_b_idx = w->n_anchor;
        const int max_search = DSB_MIN(mem_rst_num, 8);
        for (int t = 0; t < max_search; t++) dsb_map_seed(w, mem_rst + t, bin_read, w->L, direction);
        dsb_mark_useless(w, a_b_idx, w->n_anchor);
    }
    w->fast_classify = 0;
}

// ------------------------------------------------------------------------------------------
// glibc-2.39 qsort replica.  The reference sorts with comparators that are not all consistent
// (SURVEY.md 0-9, 7.3-2), so the RESULT depends on the algorithm: glibc's qsort here is the
// top-down merge sort of msort.c -- split n1 = n/2, n2 = n-n1, sort both halves, merge taking
// the LEFT element when cmp(left, right) <= 0.  We sort an index permutation with exactly that
// recursion tree (explicit stack) and the caller applies the permutation.
//   cmp(ctx, i, j) must return the reference comparator's value for elements i and j.
// ------------------------------------------------------------------------------------------
template <typename Cmp>
DSB_HD void dsb_msort_idx(uint32_t *idx, uint32_t *tmp, uint32_t n, Cmp cmp)
{
    for (uint32_t i = 0; i < n; i++) idx[i] = i;
    if (n <= 1) return;
    // post-order walk of the split tree
    uint32_t st_b[40], st_n[40];
    uint8_t st_state[40];
    int sp = 0;
    st_b[0] = 0; st_n[0] = n; st_state[0] = 0;
    while (sp >= 0) {
        const uint32_t b = st_b[sp], nn = st_n[sp];
        if (nn <= 1) { sp--; continue; }
        const uint32_t n1 = nn / 2, n2 = nn - n1;
        if (st_state[sp] == 0) {
            st_state[sp] = 1;
            sp++; st_b[sp] = b; st_n[sp] = n1; st_state[sp] = 0;
            continue;
        }
        if (st_state[sp] == 1) {
            st_state[sp] = 2;
            sp++; st_b[sp] = b + n1; st_n[sp] = n2; st_state[sp] = 0;
            continue;
        }
        // merge [b, b+n1) and [b+n1, b+nn)
        uint32_t i1 = b, e1 = b + n1, i2 = b + n1, e2 = b + nn, o = 0;
        while (i1 < e1 && i2 < e2) {
            if (cmp(idx[i1], idx[i2]) <= 0) tmp[o++] = idx[i1++];
            else tmp[o++] = idx[i2++];
        }
        while (i1 < e1) tmp[o++] = idx[i1++];
        for (uint32_t k = 0; k < o; k++) idx[b + k] = tmp[k];   // the rest of the right run is in place
        sp--;
    }
}

// ------------------------------------------------------------------------------------------
// chaining (restates chain_insert_meta src/cly.c:71-111, chain_insert_M2 :200-223,
// chain_insert_M3 :237-322, chain_cmp_by_score :37-51, resolve_tree :325-348)
// ------------------------------------------------------------------------------------------
DSB_HD DsbChain *dsb_push_chain(DsbWork *w)
{
    if (w->n_chain >= w->cap_chain) { w->overflow = 1; return w->chains + (w->cap_chain - 1); }
    return w->chains + (w->n_chain++);
}

DSB_HD void dsb_chain_insert_meta(DsbWork *w, uint32_t ai, DsbChain *c, bool new_chain, int dis_minus)
{
    DsbAnchor *anchor = w->anchors + ai;
    const uint32_t ref_l = anchor->ref_offset;
    const uint32_t ref_r = ref_l + anchor->mtch_len;
    const uint32_t read_l = anchor->index_in_read;
    const uint32_t read_r = read_l + anchor->mtch_len;
    if (new_chain) {
        anchor->pre = -1;
        c->ref_ID = anchor->ref_ID;
        c->direction = anchor->direction;
        c->q_t_dis = (int32_t)(anchor->ref_offset - anchor->index_in_read);
        c->t_st = ref_l;
        c->t_ed = ref_r;
        c->q_st = read_l;
        c->q_ed = read_r;
        c->with_top = !anchor->useless;
        c->anchor_number = 1;
        c->sum_score = (anchor->duplicate) ? 1 : (uint32_t)(int32_t)anchor->score;
        c->indel = 0;
        c->cur = (int32_t)ai;
    } else {
        c->with_top |= (!anchor->useless);
        if (c->q_ed >= read_r) return;
        c->t_ed = DSB_MAX(ref_r, c->t_ed);
        c->q_ed = read_r;
        anchor->pre = c->cur;
        c->cur = (int32_t)ai;
        c->q_t_dis = (int32_t)(anchor->ref_offset - anchor->index_in_read);
        c->indel += (uint32_t)dis_minus;
        c->anchor_number++;
        c->sum_score += (anchor->duplicate) ? 1 : (uint32_t)(int32_t)anchor->score;
    }
}

DSB_HD void dsb_chain_insert_M2(DsbWork *w, uint32_t ai)
{
    const DsbAnchor *anchor = w->anchors + ai;
    const uint8_t direction = anchor->direction;
    const uint32_t ref_ID = anchor->ref_ID;
    const int32_t dis = (int32_t)(anchor->ref_offset - anchor->index_in_read);
    int dis_minus = 0;
    for (uint32_t ci = 0; ci < w->n_chain; ci++) {
        DsbChain *c_s = w->chains + ci;
        if (c_s->direction == direction && c_s->ref_ID == ref_ID &&
            (dis_minus = DSB_ABS(dis - c_s->q_t_dis)) < 30 &&
            DSB_ABS_U(c_s->t_ed, anchor->ref_offset) < 400) {
            dsb_chain_insert_meta(w, ai, c_s, false, dis_minus);
            return;
        }
    }
    DsbChain *nc = dsb_push_chain(w);
    nc->primary = 0; nc->pri_index = 0;
    dsb_chain_insert_meta(w, ai, nc, true, dis_minus);
}

DSB_HDN void dsb_chain_insert_M3(DsbWork *w)
{
    const uint32_t n = w->n_anchor;
    // stable sort by (ref_ID, direction, ref_offset): Anchor_cmp_by_chr_ID_and_pos returns a>b
#if defined(DSB_DEVICE_CODE)
    {
        // The comparator is a consistent order on (ref_ID, direction, ref_offset), so the reference's merge
        // sort is simply a STABLE sort: equal to sorting the unique pairs (key, input index) with any
        // algorithm.  Bitonic network over the warp, keys in the (idle) chains_tmp area.
        const uint32_t lane = threadIdx.x & 31;
        uint32_t P = 64;
        while (P < n) P <<= 1;
        uint64_t *key = (uint64_t *)w->chains_tmp;          // cap_anchor * sizeof(DsbChain) >= 2 * cap_anchor * 8
        uint32_t *idx = w->sort_a;                          // 2 * cap_anchor entries
        for (uint32_t i = lane; i < P; i += 32) {
            uint64_t k = ~0ull;
            if (i < n) {
                const DsbAnchor *a = w->anchors + i;
                k = ((uint64_t)a->ref_ID << 33) | ((uint64_t)(a->direction ? 1u : 0u) << 32) | (uint64_t)a->ref_offset;
            }
            key[i] = k;
            idx[i] = i;
        }
        __syncwarp();
        for (uint32_t kk = 2; kk <= P; kk <<= 1) {
            for (uint32_t j = kk >> 1; j > 0; j >>= 1) {
                for (uint32_t t = lane; t < (P >> 1); t += 32) {
                    const uint32_t lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));     // pair (lo, lo | j)
                    const uint32_t hi = lo | j;
                    const bool up = ((lo & kk) == 0);
                    const uint64_t ka = key[lo], kb = key[hi];
                    const uint32_t ia = idx[lo], ib = idx[hi];
                    const bool a_gt_b = (ka > kb) || (ka == kb && ia > ib);
                    if (a_gt_b == up) { key[lo] = kb; key[hi] = ka; idx[lo] = ib; idx[hi] = ia; }
                }
                __syncwarp();
            }
        }
        for (uint32_t i = lane; i < n; i += 32) w->anchors_tmp[i] = w->anchors[idx[i]];
        __syncwarp();
        for (uint32_t i = lane; i < n; i += 32) w->anchors[i] = w->anchors_tmp[i];
        __syncwarp();
    }
#else
    {
        const DsbAnchor *A = w->anchors;
        dsb_msort_idx(w->sort_a, w->sort_b, n, [A](uint32_t i, uint32_t j) -> int {
            const DsbAnchor *a = A + i, *b = A + j;
            if (a->ref_ID != b->ref_ID) return a->ref_ID > b->ref_ID;
            if (a->direction != b->direction) return a->direction > b->direction;
            return a->ref_offset > b->ref_offset;
        });
        for (uint32_t i = 0; i < n; i++) w->anchors_tmp[i] = w->anchors[w->sort_a[i]];
        for (uint32_t i = 0; i < n; i++) w->anchors[i] = w->anchors_tmp[i];
    }
#endif
    DsbAnchor *A = w->anchors;
    int32_t *score_v = w->score_v;
    for (uint32_t chr_st = 0; chr_st < n;) {
        uint32_t chr_ed = chr_st + 1;
        const uint32_t ref_ID = A[chr_st].ref_ID;
        const uint32_t direction = A[chr_st].direction;
        for (; chr_ed < n && A[chr_ed].ref_ID == ref_ID && A[chr_ed].direction == direction &&
               A[chr_ed].ref_offset - A[chr_ed - 1].ref_offset < 2000; chr_ed++) {}
        if (chr_ed - chr_st > 1024) chr_ed = chr_st + 1024;
        int32_t max_anchor = -1;
        int max_score = 0, anchor_max_score;
        for (uint32_t c_a = chr_st; c_a < chr_ed; c_a++) {
            A[c_a].pre = -1;
            anchor_max_score = A[c_a].score;
            const uint32_t max_t = A[c_a].ref_offset + 3;
            const uint32_t max_q = A[c_a].index_in_read + 3;
#if defined(DSB_DEVICE_CODE)
            {
                // 32 predecessors per step, scanned from c_a-1 downwards.  The reference takes the FIRST
                // predecessor reaching the maximum (strict >) and stops at the first one that passes both
                // overlap tests and lies more than 1000 bases back.
                const int32_t lane = (int32_t)(threadIdx.x & 31);
                const uint32_t c_len = A[c_a].mtch_len;
                int32_t best_pre = -1;
                for (int32_t top = (int32_t)c_a - 1; top >= (int32_t)chr_st; top -= 32) {
                    const int32_t pre = top - lane;
                    bool brk = false;
                    int cand = -2147483647 - 1;
                    if (pre >= (int32_t)chr_st) {
                        const DsbAnchor pv = A[pre];
                        if (!(pv.index_in_read + pv.mtch_len > max_q) && !(pv.ref_offset + pv.mtch_len > max_t)) {
                            if (pv.index_in_read + 1000 < max_q || pv.ref_offset + 1000 < max_t) brk = true;
                            else {
                                const int indel = (int)(pv.index_in_read - pv.ref_offset - (max_q - max_t));
                                const int ABS_indel = DSB_ABS(indel);
                                if (!(ABS_indel > 200))
                                    cand = (int)(score_v[pre - (int32_t)chr_st] + c_len - (ABS_indel >> 4) -
                                                 ((max_q - pv.index_in_read) >> 8));
                            }
                        }
                    }
                    const uint32_t bm = __ballot_sync(0xffffffffu, brk);
                    if (bm) {
                        const int32_t cut = __ffs((int)bm) - 1;      // lanes below `cut` were scanned before the break
                        if (lane >= cut) cand = -2147483647 - 1;
                    }
                    const int gmax = __reduce_max_sync(0xffffffffu, cand);
                    if (gmax > anchor_max_score) {
                        const uint32_t who = __ballot_sync(0xffffffffu, cand == gmax);
                        anchor_max_score = gmax;
                        best_pre = top - (__ffs((int)who) - 1);      // lowest lane = highest index = met first
                    }
                    if (bm) break;
                }
                if (best_pre >= 0) A[c_a].pre = best_pre;
                __syncwarp();
            }
#else
            for (int32_t pre = (int32_t)c_a - 1; pre >= (int32_t)chr_st; pre--) {
                const DsbAnchor *p = A + pre;
                if (p->index_in_read + p->mtch_len > max_q) continue;
                if (p->ref_offset + p->mtch_len > max_t) continue;
                if (p->index_in_read + 1000 < max_q) break;
                if (p->ref_offset + 1000 < max_t) break;
                const int indel = (int)(p->index_in_read - p->ref_offset - (max_q - max_t));
                const int ABS_indel = DSB_ABS(indel);
                if (ABS_indel > 200) continue;
                const int new_score = (int)(score_v[pre - (int32_t)chr_st] + A[c_a].mtch_len - (ABS_indel >> 4) -
                                            ((max_q - p->index_in_read) >> 8));
                if (new_score > anchor_max_score) {
                    anchor_max_score = new_score;
                    A[c_a].pre = pre;
                }
            }
#endif
            score_v[c_a - chr_st] = anchor_max_score;
#if defined(DSB_DEVICE_CODE)
            __syncwarp();                                  // the next anchor's lanes read this entry
#endif
            if (max_score < anchor_max_score) {
                max_score = anchor_max_score;
                max_anchor = (int32_t)c_a;
            }
        }
        // the reference dereferences max_anchor unconditionally; every anchor score is >= 20 so
        // max_anchor is always set
        if (max_anchor < 0) max_anchor = (int32_t)chr_st;
        int sum_INDEL = 0, anchor_number = 1;
        int32_t pre = max_anchor;
        int sum_score = (A[max_anchor].duplicate) ? 1 : A[max_anchor].score;
        bool with_top = !A[max_anchor].useless;
        for (; A[pre].pre != -1; anchor_number++) {               // replayed literally
            const int32_t pre_ = A[pre].pre;
            sum_INDEL += (int)((A[pre].index_in_read - A[pre_].index_in_read) - (A[pre].ref_offset - A[pre_].ref_offset));
            with_top |= (!A[pre].useless);
            sum_score += (A[pre].duplicate) ? 1 : A[pre].score;
            pre = pre_;
        }
        DsbChain *nc = dsb_push_chain(w);
        nc->ref_ID = ref_ID;
        nc->direction = (uint8_t)direction;
        nc->q_t_dis = (int32_t)(A[max_anchor].ref_offset - A[max_anchor].index_in_read);
        nc->t_st = A[pre].ref_offset;
        nc->t_ed = A[max_anchor].ref_offset + A[max_anchor].mtch_len;
        nc->q_st = A[pre].index_in_read;
        nc->q_ed = A[max_anchor].index_in_read + A[max_anchor].mtch_len;
        nc->with_top = with_top;
        nc->anchor_number = (uint32_t)anchor_number;
        nc->sum_score = (uint32_t)sum_score;
        nc->indel = (uint32_t)sum_INDEL;
        nc->cur = max_anchor;
        nc->primary = 0; nc->pri_index = 0;
        chr_st = chr_ed;
    }
}

DSB_HD void dsb_apply_chain_perm(DsbWork *w, uint32_t n)
{
    for (uint32_t i = 0; i < n; i++) w->chains_tmp[i] = w->chains[w->sort_a[i]];
    for (uint32_t i = 0; i < n; i++) w->chains[i] = w->chains_tmp[i];
}

DSB_HDN void dsb_resolve_tree(DsbWork *w)
{
    w->n_chain = 0;
    if (w->n_anchor < 50)
        for (uint32_t ai = 0; ai < w->n_anchor; ai++) dsb_chain_insert_M2(w, ai);
    else
        dsb_chain_insert_M3(w);
    if (w->n_chain > 1) {
        const DsbChain *C = w->chains;
        dsb_msort_idx(w->sort_a, w->sort_b, w->n_chain, [C](uint32_t i, uint32_t j) -> int {
            const DsbChain *a = C + i, *b = C + j;
            if (a->with_top != b->with_top) return (a->with_top) ? (-1) : (1);
            int score_a = (int)(a->sum_score + ((a->q_ed - a->q_st) << 1));
            score_a -= (int)(a->indel << 2);
            int score_b = (int)(b->sum_score + ((b->q_ed - b->q_st) << 1));
            score_b -= (int)(b->indel << 2);
            if (score_a < score_b) return 1;
            if (score_a > score_b) return -1;
            return 0;
        });
        dsb_apply_chain_perm(w, w->n_chain);
    }
    uint32_t rst_num = DSB_MIN(5u, w->n_chain);
    while (rst_num < w->n_chain && w->chains[rst_num].with_top == 1) rst_num++;
    w->n_chain = rst_num;
}

// ------------------------------------------------------------------------------------------
// read 9-mer table (restates build_hash_table_M2 src/cly.c:2168-2219).  The reference appends
// each position to the tail of its bucket chain, so a bucket lists positions in ascending
// order; we keep that order with head/next arrays filled from the highest position down.
// Bucket key = low key_len bits of the 18-bit 9-mer; entries store pos+1, 0 terminates.
// ------------------------------------------------------------------------------------------
// Bucket-index width of the read's 9-mer table.  The reference takes the smallest 2^k >= L, k in 10..17
// (src/cly.c:2179-2182); the width only decides how the positions with equal 9-mers are FOUND, never which ones or in what
// order (every chain lists them in ascending position and compares the whole 9-mer), so it is a free parameter here.
#ifndef DSB_KEYLEN_BIAS
#define DSB_KEYLEN_BIAS 0
#endif
DSB_HD int dsb_read_hash_key_len(uint32_t L)
{
    int key_len = 10;
    for (; key_len < 18; key_len++)
        if ((1u << key_len) >= L) break;
    key_len += DSB_KEYLEN_BIAS;
    if (key_len < 10) key_len = 10;
    if (key_len > 18) key_len = 18;
    return key_len;
}

// The read's 9-mer table is BLOCKED BY READ POSITION: segment g holds the positions [g << DSB_SEG_BITS, (g + 1) << DSB_SEG_BITS)
// in its own 2^DSB_SEG_BITS buckets (key = low bits of the 9-mer, the tag in the entry = the remaining bits, so a tag match is
// an exact 9-mer match).  The reference's table (src/cly.c:2168-2219) is one hash over the whole read; its results depend only
// on, per window 9-mer, the matching read positions inside [q_bg, q_ed] in ASCENDING order (chains are appended at the tail,
// :2206-2215) -- the bucket width is free.  A scoring window only asks for positions within ~2000 bases of the chain's end
// (sdp_right / sdp_left / sdp_middle), i.e. 2-3 segments: its loads stay inside ~6 KB per segment (heads 2 B, entries 4 B)
// instead of being spread over the whole 70 KB table of a 10 kb read, which every resident warp evicted from L1 / L2 between
// two uses (round 1: 5.7 GB of DRAM traffic per step for 0.27 GB of algorithmic bytes).  Same-box A/B on the 1 Gbp / 10 kb
// workload (profiles/ab_r02.txt): k_classify 171.6 -> 165.3 ms with 2048-position segments (1024: 166.5, 4096: 167.5), output
// identical.  Tried on top and dropped, all slower in the same runs: first entry's tag + a has-next bit inlined in 32-bit
// heads (one load instead of head -> entry for most look-ups: 190 ms, the heads double and the cursor grows), 4 / 8
// segments built in an interleaved fashion (181 / 188 ms: register arrays under the kernel's 80-register cap), word-wide
// exact-run extension of the hits (166.8 ms), compaction of the window k-mers that have a hit before the divergent extension
// phase (166.9 ms).
//   head16[g * SEG + key] = (position inside the segment) + 1, 0 = empty bucket
//   ent[p]                = (next position in the bucket + 1) << 8 | tag       (positions ascending along a chain)
#ifndef DSB_SEG_BITS
#define DSB_SEG_BITS 11
#endif
#define DSB_SEG (1u << DSB_SEG_BITS)
#define DSB_SEG_MASK (DSB_SEG - 1u)
#if DSB_SEG_BITS < 10 || DSB_SEG_BITS > 15
#error "DSB_SEG_BITS: the tag (18 - DSB_SEG_BITS bits) must fit 8 bits and a local position + 1 must fit 16 bits"
#endif

// cursor over the hits of one window 9-mer: ascending read positions p with 9-mer(p) == kmer and q_bg <= p <= q_ed
struct DsbTabCur { uint32_t node, seg, seg_hi, key, tag; };

// advance from the candidate c->node (0 = bucket exhausted) to the next entry that matches; returns position + 1 or 0
DSB_HD uint32_t dsb_tab_settle(const uint16_t *head, const uint32_t *ent, DsbTabCur *c, uint32_t q_bg, uint32_t q_ed)
{
    uint32_t node = c->node;
    for (;;) {
        while (node != 0) {
            const uint32_t qp = node - 1;
            if (qp > q_ed) { node = 0; c->seg = c->seg_hi; break; }      // chains ascend: nothing further can be in range
            const uint32_t e = ent[qp];
            if ((e & 0xffu) == c->tag && qp >= q_bg) { c->node = node; return node; }
            node = e >> 8;
        }
        if (c->seg >= c->seg_hi) { c->node = 0; return 0; }
        c->seg++;
        const uint32_t h = head[(c->seg << DSB_SEG_BITS) + c->key];
        node = h ? (c->seg << DSB_SEG_BITS) + h : 0u;
    }
}

// first hit of `kmer` in [q_bg, q_ed]; n_pos = number of positions the table holds
DSB_HD uint32_t dsb_tab_first(const uint16_t *head, const uint32_t *ent, DsbTabCur *c, uint32_t kmer, uint32_t q_bg, uint32_t q_ed,
                              uint32_t n_pos)
{
    c->key = kmer & DSB_SEG_MASK;
    c->tag = kmer >> DSB_SEG_BITS;
    c->node = 0;
    if (q_bg > q_ed || q_bg >= n_pos) { c->seg = c->seg_hi = 0; return 0; }     // a wrapped q_bg matches nothing
    c->seg = q_bg >> DSB_SEG_BITS;
    c->seg_hi = (q_ed < n_pos ? q_ed : n_pos - 1) >> DSB_SEG_BITS;
    const uint32_t h = head[(c->seg << DSB_SEG_BITS) + c->key];
    c->node = h ? (c->seg << DSB_SEG_BITS) + h : 0u;
    return dsb_tab_settle(head, ent, c, q_bg, q_ed);
}

DSB_HD uint32_t dsb_tab_next(const uint16_t *head, const uint32_t *ent, DsbTabCur *c, uint32_t q_bg, uint32_t q_ed)
{
    c->node = ent[c->node - 1] >> 8;
    return dsb_tab_settle(head, ent, c, q_bg, q_ed);
}

DSB_HD uint32_t dsb_kmer9(const uint8_t *s)
{
    uint32_t v = 0;
    for (int i = 0; i < DSB_KMER9; i++) v = (v << 2) | (uint32_t)(s[i] & 3);
    return v;
}

DSB_HDN int dsb_build_read_hash(DsbWork *w)
{
    int both_dir = 0;
    for (uint32_t i = 0; i < w->n_chain; i++) {
        both_dir |= (w->chains[i].direction == DSB_FORWARD) ? 0x2 : 0x1;
        if (both_dir == 3) break;
    }
    const int q_len = (int)w->L;
    const int key_len = DSB_SEG_BITS;
    const uint32_t n_pos = (uint32_t)(q_len - DSB_KMER9 + 1);
    const uint32_t n_head = ((n_pos + DSB_SEG_MASK) >> DSB_SEG_BITS) << DSB_SEG_BITS;
    if (w->L >= (1u << 24)) { w->overflow = 16; return key_len; }    // entries hold 24-bit positions (the reference's own table stops at 1 Mi nodes); its own flag: the host rejects such reads before they get here
    for (int c_dir = 2; c_dir >= 1; c_dir--) {
        if ((c_dir & both_dir) == 0) continue;
        const int strand = (c_dir == 2) ? 0 : 1;
        const uint8_t *q = w->bin + (strand ? w->L : 0);
        uint16_t *head = (uint16_t *)w->h_head[strand];
        uint32_t *ent = w->h_ent[strand];
#if defined(DSB_DEVICE_CODE)
        const uint32_t lane = threadIdx.x & 31;
        for (uint32_t i = lane; i < (n_head >> 1); i += 32) ((uint32_t *)head)[i] = 0;
        __syncwarp();
        // Buckets must list positions in ascending order (the reference appends to the tail).  32
        // positions per step, from the highest chunk down (a chunk never straddles a segment): inside a chunk,
        // lanes with the same key (match_any) link to the next higher lane of their group; the group's highest
        // lane links to the bucket head left by the chunks above, its lowest lane becomes the new head.
        for (int32_t cs = (int32_t)((n_pos - 1) & ~31u); cs >= 0; cs -= 32) {
            const uint32_t p = (uint32_t)cs + lane;
            const bool valid = p < n_pos;
            const uint32_t km = valid ? dsb_kmer9(q + p) : 0u;
            const uint32_t seg_base = (uint32_t)cs & ~DSB_SEG_MASK;
            const uint32_t key = valid ? (km & DSB_SEG_MASK) : (0x80000000u | lane);
            const uint32_t peers = __match_any_sync(0xffffffffu, key);
            const uint32_t higher = peers & ~((2u << lane) - 1u);
            uint32_t nx = 0;
            if (valid) {
                if (higher) nx = (uint32_t)cs + (uint32_t)__ffs((int)higher) - 1u + 1u;
                else { const uint32_t h = head[seg_base + key]; nx = h ? seg_base + h : 0u; }
            }
            __syncwarp();
            if (valid) {
                ent[p] = (nx << 8) | (km >> DSB_SEG_BITS);
                if ((peers & ((1u << lane) - 1u)) == 0) head[seg_base + key] = (uint16_t)((p & DSB_SEG_MASK) + 1u);
            }
            __syncwarp();
        }
#else
        for (uint32_t i = 0; i < n_head; i++) head[i] = 0;
        for (uint32_t p = n_pos; p-- > 0;) {
            const uint32_t km = dsb_kmer9(q + p);
            const uint32_t seg_base = p & ~DSB_SEG_MASK, key = km & DSB_SEG_MASK;
            const uint32_t h = head[seg_base + key];
            ent[p] = ((h ? seg_base + h : 0u) << 8) | (km >> DSB_SEG_BITS);
            head[seg_base + key] = (uint16_t)((p & DSB_SEG_MASK) + 1u);
        }
#endif
    }
    return key_len;
}

// ------------------------------------------------------------------------------------------
// match blocks between a reference window and the read (restates MEM_search src/cly.c:1805-1813
// and sdp_match :2330-2435).  q_str = strand base pointer; t_str = window with readable margins.
// ------------------------------------------------------------------------------------------
DSB_HD int dsb_run_len(const uint8_t *q, const uint8_t *t, bool forward, int max)
{
    int len = 0;
#if defined(DSB_DEVICE_CODE)
    if (max > 8) {
        // 32 positions per step: the first lane whose bases differ (or that lies beyond `max`) ends the run
        const int lane = (int)(threadIdx.x & 31);
        for (;;) {
            const int idx = len + lane;
            bool ok = false;
            if (idx < max) ok = forward ? (q[idx] == t[idx]) : (q[-idx] == t[-idx]);
            const uint32_t m = __ballot_sync(0xffffffffu, ok);
            if (m != 0xffffffffu) return len + __ffs((int)~m) - 1;
            len += 32;
        }
    }
#endif
    if (forward) for (; len < max && *q++ == *t++; len++) {}
    else for (; len < max && *q-- == *t--; len++) {}
    return len;
}

// the same, one lane on its own
DSB_HD int dsb_run_len_solo(const uint8_t *q, const uint8_t *t, bool forward, int max)
{
    int len = 0;
    if (forward) for (; len < max && *q++ == *t++; len++) {}
    else for (; len < max && *q-- == *t--; len++) {}
    return len;
}

#define DSB_SMS_LANE_CAP 8       // match blocks one lane may stage per window k-mer before the serial path takes over

DSB_HD void dsb_push_sms(DsbWork *w, uint32_t t_pos, uint32_t q_pos, uint32_t len)
{
    if (w->n_sms >= w->cap_sms) { w->overflow = 1; return; }
    DsbSms *p = w->sms + (w->n_sms++);
    p->t_pos = t_pos; p->q_pos = q_pos; p->len = len;
}

// One 9-mer hit (window 9-mer number i at offset c_off of the window, read position q_pos): the two exact-run
// extensions and the acceptance rules of sdp_match (src/cly.c:2341-2378 forward, :2390-2430 backward), by
// one lane on its own.  Returns whether a match block is emitted.
DSB_HD bool dsb_extend_hit(const uint8_t *q_str, const uint8_t *t_str, uint32_t t_len, uint32_t q_ed, uint32_t t_st,
                           bool is_forward, int i_l, long off_l, uint32_t q_pos, uint32_t *o_t, uint32_t *o_q, uint32_t *o_len)
{
    const uint8_t *c_t = t_str + off_l;
    if (is_forward) {
        const int back_len = dsb_run_len_solo(q_str + q_pos - 1, c_t - 1, false, 4);
        if (back_len < 4 || i_l == 4) {
            uint32_t max_search = q_ed - q_pos - 1;
            max_search = DSB_MIN(max_search, t_len - (uint32_t)i_l - 1) + DSB_OVER_SEARCH;
            const int forward_len = dsb_run_len_solo(q_str + q_pos + DSB_KMER9, c_t + DSB_KMER9, true, (int)max_search);
            const int total_len = back_len + forward_len + 1;
            if (total_len >= 4) {
                *o_t = (uint32_t)(i_l - back_len) + t_st; *o_q = q_pos - (uint32_t)back_len; *o_len = (uint32_t)total_len;
                return true;
            }
        }
    } else {
        const int forward_len = dsb_run_len_solo(q_str + q_pos + DSB_KMER9, c_t + DSB_KMER9, true, 4);
        if (forward_len < 4 || i_l == 4) {
            uint32_t max_search = q_pos;
            max_search = (uint32_t)(DSB_MIN((long)max_search, off_l)) + DSB_OVER_SEARCH;
            const int back_len = dsb_run_len_solo(q_str + q_pos - 1, c_t - 1, false, (int)max_search);
            const int total_len = back_len + forward_len + 1;
            if (total_len >= 4) {
                *o_t = (uint32_t)(off_l - back_len + (long)t_st); *o_q = q_pos - (uint32_t)back_len; *o_len = (uint32_t)total_len;
                return true;
            }
        }
    }
    return false;
}

#define DSB_SDP_STAGE 16         // staging slots per lane (dsb_sdp_match)

DSB_HDN void dsb_sdp_match(DsbWork *w, uint32_t q_bg, uint32_t q_ed, const uint8_t *q_str, const uint8_t *t_str,
                           uint32_t t_len, int key_len, int strand, uint32_t t_st, bool is_forward)
{
    (void)key_len;
    const uint32_t t_kmer_num = t_len - DSB_KMER9 + 1;
    const uint16_t *head = (const uint16_t *)w->h_head[strand];
    const uint32_t *ent = w->h_ent[strand];
    const uint32_t n_pos = w->L - DSB_KMER9 + 1;
    w->st.sdp_win++;
#if defined(DSB_DEVICE_CODE)
    {
        // Lane-parallel form: lane l owns window k-mer number i = 4 + 4*(base + l).  Every lane walks its
        // own bucket chain to its first hit (32 dependent-load chains in flight instead of one); hits are
        // then consumed in (lane, chain) order = the serial order, with all lanes cooperating on the
        // exact-run extension, so the match-block list is identical to the reference's.
        const uint32_t lane = threadIdx.x & 31;
        const uint32_t n_i = (t_kmer_num > 4) ? ((t_kmer_num - 4 + 3) >> 2) : 0;     // i = 4, 8, ... < t_kmer_num
        for (uint32_t base = 0; base < n_i; base += 32) {
            const uint32_t mine = base + lane;
            const int i_l = 4 + 4 * (int)mine;
            const long off_l = is_forward ? (long)i_l : ((long)t_len - DSB_KMER9 - (long)i_l);
            uint32_t node = 0;
            DsbTabCur cur = {0, 0, 0, 0, 0};
            if (mine < n_i) node = dsb_tab_first(head, ent, &cur, dsb_kmer9(t_str + off_l), q_bg, q_ed, n_pos);
            if (!__any_sync(0xffffffffu, node != 0)) continue;
            // Every hit is an independent computation (two exact-run extensions); only the ORDER of the
            // match-block list couples them (window position, then bucket order).  So each lane extends
            // all hits of its own k-mer alone into a small staging area, and the warp then appends the
            // staged blocks in lane order.  A lane with more hits than the staging holds (a repeat)
            // sends the whole group of 32 through the serial consumption below instead.
            {
                DsbSms *my = w->sms_stage + lane * DSB_SMS_LANE_CAP;
                uint32_t cnt = 0, nd = node;
                DsbTabCur c2 = cur;                           // the serial path below restarts from `cur`
                while (nd != 0 && cnt <= DSB_SMS_LANE_CAP) {
                    const uint32_t q_pos = nd - 1;
                    uint32_t o_t = 0, o_q = 0, o_len = 0;
                    const bool emit = dsb_extend_hit(q_str, t_str, t_len, q_ed, t_st, is_forward, i_l, off_l, q_pos, &o_t, &o_q, &o_len);
                    if (emit) {
                        if (cnt < DSB_SMS_LANE_CAP) { my[cnt].t_pos = o_t; my[cnt].q_pos = o_q; my[cnt].len = o_len; }
                        cnt++;
                    }
                    nd = dsb_tab_next(head, ent, &c2, q_bg, q_ed);
                }
                if (!__any_sync(0xffffffffu, cnt > DSB_SMS_LANE_CAP)) {
                    uint32_t pre = cnt;                       // inclusive prefix sum of the lanes' counts
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t v = __shfl_up_sync(0xffffffffu, pre, o);
                        if ((int)lane >= o) pre += v;
                    }
                    const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
                    if (w->n_sms + total > w->cap_sms) { w->overflow = 1; return; }
                    DsbSms *dst = w->sms + w->n_sms + (pre - cnt);
                    for (uint32_t k = 0; k < cnt; k++) { dst[k].t_pos = my[k].t_pos; dst[k].q_pos = my[k].q_pos; dst[k].len = my[k].len; }
                    w->n_sms += total;
                    __syncwarp();
                    continue;
                }
            }
            uint32_t act = __ballot_sync(0xffffffffu, node != 0);
            while (act) {
                const int l = __ffs((int)act) - 1;
                const int i = 4 + 4 * (int)(base + (uint32_t)l);
                const long c_off = is_forward ? (long)i : ((long)t_len - DSB_KMER9 - (long)i);
                const uint8_t *c_t_str = t_str + c_off;
                for (;;) {
                    const uint32_t n_l = __shfl_sync(0xffffffffu, node, l);
                    if (n_l == 0) break;
                    const uint32_t q_pos = n_l - 1;
                    if (is_forward) {
                        const int back_len = dsb_run_len(q_str + q_pos - 1, c_t_str - 1, false, 4);
                        if (back_len < 4 || i == 4) {
                            uint32_t max_search = q_ed - q_pos - 1;
                            max_search = DSB_MIN(max_search, t_len - (uint32_t)i - 1) + DSB_OVER_SEARCH;
                            const int forward_len = dsb_run_len(q_str + q_pos + DSB_KMER9, c_t_str + DSB_KMER9, true, (int)max_search);
                            const int total_len = back_len + forward_len + 1;
                            if (total_len >= 4)
                                dsb_push_sms(w, (uint32_t)(i - back_len) + t_st, q_pos - (uint32_t)back_len, (uint32_t)total_len);
                        }
                    } else {
                        const int forward_len = dsb_run_len(q_str + q_pos + DSB_KMER9, c_t_str + DSB_KMER9, true, 4);
                        if (forward_len < 4 || i == 4) {
                            uint32_t max_search = q_pos;
                            max_search = (uint32_t)(DSB_MIN((long)max_search, c_off)) + DSB_OVER_SEARCH;
                            const int back_len = dsb_run_len(q_str + q_pos - 1, c_t_str - 1, false, (int)max_search);
                            const int total_len = back_len + forward_len + 1;
                            if (total_len >= 4)
                                dsb_push_sms(w, (uint32_t)(c_off - back_len + (long)t_st), q_pos - (uint32_t)back_len, (uint32_t)total_len);
                        }
                    }
                    if ((int)lane == l) node = dsb_tab_next(head, ent, &cur, q_bg, q_ed);   // the owner moves to its next hit
                    __syncwarp();
                }
                act &= ~(1u << l);
            }
        }
        return;
    }
#endif
    if (is_forward) {
        for (int i = 4; (uint32_t)i < t_kmer_num; i += 4) {
            const uint8_t *c_t_str = t_str + i;
            DsbTabCur cur;
            uint32_t node = dsb_tab_first(head, ent, &cur, dsb_kmer9(c_t_str), q_bg, q_ed, n_pos);
            while (node != 0) {
                const uint32_t q_pos = node - 1;
                {
                    const int back_len = dsb_run_len(q_str + q_pos - 1, c_t_str - 1, false, 4);
                    if (back_len < 4 || i == 4) {
                        uint32_t max_search = q_ed - q_pos - 1;
                        max_search = DSB_MIN(max_search, t_len - (uint32_t)i - 1) + DSB_OVER_SEARCH;
                        const int forward_len = dsb_run_len(q_str + q_pos + DSB_KMER9, c_t_str + DSB_KMER9, true, (int)max_search);
                        const int total_len = back_len + forward_len + 1;
                        if (total_len >= 4)
                            dsb_push_sms(w, (uint32_t)(i - back_len) + t_st, q_pos - (uint32_t)back_len, (uint32_t)total_len);
                    }
                }
                node = dsb_tab_next(head, ent, &cur, q_bg, q_ed);
            }
        }
    } else {
        for (int i = 4; (uint32_t)i < t_kmer_num; i += 4) {
            const long c_off = (long)t_len - DSB_KMER9 - (long)i;       // c_t_str - t_str
            const uint8_t *c_t_str = t_str + c_off;
            DsbTabCur cur;
            uint32_t node = dsb_tab_first(head, ent, &cur, dsb_kmer9(c_t_str), q_bg, q_ed, n_pos);
            while (node != 0) {
                const uint32_t q_pos = node - 1;
                {
                    const int forward_len = dsb_run_len(q_str + q_pos + DSB_KMER9, c_t_str + DSB_KMER9, true, 4);
                    if (forward_len < 4 || i == 4) {
                        uint32_t max_search = q_pos;
                        max_search = (uint32_t)(DSB_MIN((long)max_search, c_off)) + DSB_OVER_SEARCH;
                        const int back_len = dsb_run_len(q_str + q_pos - 1, c_t_str - 1, false, (int)max_search);
                        const int total_len = back_len + forward_len + 1;
                        if (total_len >= 4)
                            dsb_push_sms(w, (uint32_t)(c_off - back_len + (long)t_st), q_pos - (uint32_t)back_len, (uint32_t)total_len);
                    }
                }
                node = dsb_tab_next(head, ent, &cur, q_bg, q_ed);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// diagonal hash of chain ends (restates sc_hash_idx src/cly.c:1686-1705) and chain merging
// (combine_chain :1758-1803)
// ------------------------------------------------------------------------------------------
DSB_HDO void dsb_sc_hash_idx(DsbWork *w)
{
    DsbScHash *h = w->sc_hash;
    for (int i = 0; i < 256; i++) { h[i].next = 0; h[i].seed_ID = 0; h[i].s_or_e = 0; }
    int con = 256;
    for (uint32_t ci = 0; ci < w->n_chain; ci++) {
        const DsbChain *c_h = w->chains + ci;
        for (int i = 1; i >= 0; i--) {
            uint16_t c_key = (uint16_t)(((i == 1) ? (c_h->t_st - c_h->q_st) : (c_h->t_ed - c_h->q_ed)) & 0xff);
            while (h[c_key].next != 0) c_key = h[c_key].next;
            h[c_key].seed_ID = (uint16_t)((ci + 1) & 0x7fff);
            h[c_key].s_or_e = (uint8_t)i;
            h[c_key].next = (uint16_t)con;
            h[con++].next = 0;
        }
    }
}

// returns the index of the absorbed chain or -1
DSB_HDO int dsb_combine_chain(DsbWork *w, int chain_ID, int dis, bool isleft, int c_q_pos)
{
    const DsbScHash *h = w->sc_hash;
    uint16_t key = (uint16_t)(dis & 0xff);
    DsbChain *c_h = w->chains + chain_ID;
    while (h[key].next != 0) {
        const uint16_t seed_ID = h[key].seed_ID;
        DsbChain *c = w->chains + ((int)seed_ID - 1);
        const int dis_con = (int)((isleft) ? (c->t_ed - c->q_ed) : (c->t_st - c->q_st));
        const int q_pos_con = (int)((!isleft) ? (c->q_st) : (c->q_ed - DSB_KMER9));
        if (dis == dis_con && c_h != c && (int)isleft != (int)h[key].s_or_e &&
            DSB_ABS_U(c_q_pos, q_pos_con) < 8 && c_h->ref_ID == c->ref_ID &&
            c_h->direction == c->direction && c->sum_score != 0 && (int)seed_ID - 1 > chain_ID) {
            c_h->sum_score += c->sum_score;
            c_h->anchor_number += c->anchor_number;
            c_h->indel += c->indel;
            c_h->q_st = DSB_MIN(c_h->q_st, c->q_st);
            c_h->t_st = DSB_MIN(c_h->t_st, c->t_st);
            c_h->q_ed = DSB_MAX(c_h->q_ed, c->q_ed);
            c_h->t_ed = DSB_MAX(c_h->t_ed, c->t_ed);
            c->sum_score = 0;
            c->t_st = c->t_ed = c->q_st = c->q_ed = 0;
            return (int)seed_ID - 1;
        }
        key = h[key].next;
    }
    return -1;
}

// ------------------------------------------------------------------------------------------
// chain re-scoring by sparse DP over match blocks (restates sdp_middle_M2 src/cly.c:2439-2525,
// sdp_right_M2 :2527-2672, sdp_left_M2 :2674-2814, get_score_M2 :2816-2844)
//
// Bytes the reference reads without having written them (window tails on its stack,
// src/cly.c:2467,2537,2683) are defined here as 0xFE = "never equal to a base", which is what
// the reference source does when built with -ftrivial-auto-var-init=pattern (oracle/Makefile
// target ref_pinit); DESIGN.md explains how reads that depend on those bytes are identified.
// ------------------------------------------------------------------------------------------
#define DSB_UNINIT 0xFE

DSB_HD void dsb_fill(uint8_t *p, uint32_t n, uint8_t v)
{
#if defined(DSB_DEVICE_CODE)
    for (uint32_t i = threadIdx.x & 31; i < n; i += 32) p[i] = v;
    __syncwarp();
#else
    for (uint32_t i = 0; i < n; i++) p[i] = v;
#endif
}

DSB_HDN int dsb_sdp_middle(DsbWork *w, int32_t c_a, const uint8_t *q_str, int strand, int key_len)
{
    const DsbIndexView *ix = w->ix;
    const DsbAnchor *A = w->anchors;
    int score = 10000;
    const uint64_t t_offset = ix->ref_off[A[c_a].ref_ID];
    while (c_a != -1) {
        const int32_t pre_a = A[c_a].pre;
        if (pre_a != -1) {
            const int pre_mch = A[pre_a].mtch_len;
            const int pre_refoffset = (int)(A[pre_a].ref_offset - 3);
            const int total_ref_len = (int)(A[c_a].ref_offset - (uint32_t)(pre_refoffset + pre_mch) + 3);
            w->n_sms = 0;
            dsb_push_sms(w, A[pre_a].ref_offset, A[pre_a].index_in_read, (uint32_t)(A[pre_a].mtch_len - DSB_KMER9 + 1));
            w->sms[0].score = (uint32_t)score;
            if (total_ref_len > 12) {
                if (total_ref_len >= 2000) { w->overflow = 2; return 0; }    // reference: xassert -> abort (src/cly.c:2468)
                uint8_t *ref = w->ref_mid;
                const uint64_t ref_offset = (uint64_t)(int64_t)(pre_refoffset) + t_offset + (uint64_t)(int64_t)pre_mch;
                dsb_get_ref(w, ref, ref_offset, (uint32_t)total_ref_len, true);
                dsb_fill(ref + total_ref_len, 64, DSB_UNINIT);
                dsb_sdp_match(w, A[pre_a].index_in_read + (uint32_t)pre_mch - 8, A[c_a].index_in_read - 1, q_str, ref,
                              (uint32_t)total_ref_len, key_len, strand, (uint32_t)(pre_refoffset + pre_mch), true);
            }
            dsb_push_sms(w, A[c_a].ref_offset, A[c_a].index_in_read, (uint32_t)(A[c_a].mtch_len - DSB_KMER9 + 1));
            if (w->n_sms > 1) {
                DsbSms *S = w->sms;
                for (uint32_t ci = 1; ci < w->n_sms; ci++) {
                    DsbSms *c_spd = S + ci;
                    int max_score = (int)c_spd->len;
                    const uint32_t max_q = c_spd->q_pos + DSB_SMS_OVERLAP;
                    const uint32_t max_t = c_spd->t_pos + DSB_SMS_OVERLAP;
#if defined(DSB_DEVICE_CODE)
                    {   // predecessors spread over the lanes; max is order-independent
                        const int32_t lane = (int32_t)(threadIdx.x & 31);
                        const uint32_t c_len = c_spd->len, c_q = c_spd->q_pos, c_t = c_spd->t_pos;
                        for (int32_t top = (int32_t)ci - 1; top >= 0; top -= 32) {
                            const int32_t pi = top - lane;
                            if (pi < 0) continue;
                            const DsbSms pv = S[pi];
                            const int pre_q_ed = (int)(pv.q_pos + pv.len + DSB_KMER9 - 1);
                            const int pre_t_ed = (int)(pv.t_pos + pv.len + DSB_KMER9 - 1);
                            if ((uint32_t)pre_q_ed > max_q) continue;
                            if ((uint32_t)pre_t_ed > max_t) continue;
                            const int indel = (int)(pv.q_pos - pv.t_pos - (max_q - max_t));
                            const int ABS_indel = DSB_ABS(indel);
                            if (ABS_indel > 200) continue;
                            int new_score = (int)(pv.score + c_len - (uint32_t)(ABS_indel >> 3));
                            if ((uint32_t)pre_q_ed > c_q || (uint32_t)pre_t_ed > c_t) {
                                const int overlap_q = (int)((uint32_t)pre_q_ed - c_q);
                                const int overlap_t = (int)((uint32_t)pre_t_ed - c_t);
                                new_score -= DSB_MAX(overlap_q, overlap_t);
                            }
                            max_score = DSB_MAX(max_score, new_score);
                        }
                        max_score = __reduce_max_sync(0xffffffffu, max_score);
                    }
#else
                    for (int32_t pi = (int32_t)ci - 1; pi >= 0; pi--) {
                        const DsbSms *p = S + pi;
                        const int pre_q_ed = (int)(p->q_pos + p->len + DSB_KMER9 - 1);
                        const int pre_t_ed = (int)(p->t_pos + p->len + DSB_KMER9 - 1);
                        if ((uint32_t)pre_q_ed > max_q) continue;
                        if ((uint32_t)pre_t_ed > max_t) continue;
                        const int indel = (int)(p->q_pos - p->t_pos - (max_q - max_t));
                        const int ABS_indel = DSB_ABS(indel);
                        if (ABS_indel > 200) continue;
                        int new_score = (int)(p->score + c_spd->len - (uint32_t)(ABS_indel >> 3));
                        if ((uint32_t)pre_q_ed > c_spd->q_pos || (uint32_t)pre_t_ed > c_spd->t_pos) {
                            const int overlap_q = (int)((uint32_t)pre_q_ed - c_spd->q_pos);
                            const int overlap_t = (int)((uint32_t)pre_t_ed - c_spd->t_pos);
                            new_score -= DSB_MAX(overlap_q, overlap_t);
                        }
                        max_score = DSB_MAX(max_score, new_score);
                    }
#endif
                    score = DSB_MAX(max_score, score);
                    c_spd->score = (uint32_t)max_score;
                }
            }
        } else
            score += A[c_a].mtch_len - DSB_KMER9 + 1;
        c_a = pre_a;
    }
    return score - 10000;
}

DSB_HDN int dsb_sdp_right(DsbWork *w, const uint8_t *q_str, int strand, int key_len, int chain_ID, uint32_t l_read,
                          int score_ori)
{
    const DsbIndexView *ix = w->ix;
    score_ori += 10000;
    int total_max_score = score_ori;
    int max_sms_id = 0;
    DsbChain *c_h = w->chains + chain_ID;
    w->n_sms = 0;
    uint8_t *ref = w->ref_ext;
    dsb_fill(ref, 1000, DSB_UNINIT);                                   // uint8_t ref[1000] (src/cly.c:2537)
    dsb_push_sms(w, c_h->t_ed, c_h->q_ed, (uint32_t)(1 - DSB_KMER9));
    w->sms[0].score = (uint32_t)score_ori;
    uint32_t current_sms = 1;
    const uint64_t t_offset_global = ix->ref_off[c_h->ref_ID];
    const uint64_t t_length = ix->ref_len[c_h->ref_ID];
    uint32_t c_t_offset = c_h->t_ed - 3;
    int last_search = 0;
    DSB_TRACE("RBEG chain=%d ref=%u q_st=%u q_ed=%u t_st=%u t_ed=%u score=%d\n", chain_ID, c_h->ref_ID, c_h->q_st, c_h->q_ed,
              c_h->t_st, c_h->t_ed, score_ori - 10000);
    while (1) {
        if (w->overflow) break;
        if (w->n_sms == current_sms) {
            const uint32_t next_step = (uint32_t)(t_length - c_t_offset);
            if (next_step < DSB_MIN_SCORE_MEM) break;
            uint32_t max_search_ref;
            if (l_read - c_h->q_ed < 600) {
                if (last_search) break;
                last_search = 1;
                max_search_ref = l_read - c_h->q_ed + 60;
            } else
                max_search_ref = (uint32_t)(t_length - c_t_offset);
            max_search_ref = DSB_MIN(600u, max_search_ref);
            dsb_get_ref(w, ref, c_t_offset + t_offset_global, max_search_ref + DSB_OVER_SEARCH, true);
            int search_q_ed = (int)w->sms[max_sms_id].q_pos + 1000;
            search_q_ed = (int)DSB_MIN((uint32_t)search_q_ed, l_read);         // MIN(int, uint32): unsigned
            const int search_q_st = (int)DSB_MAX((uint32_t)(search_q_ed - 2000), c_h->q_st - 8);   // MAX(int, uint32): unsigned
            dsb_sdp_match(w, (uint32_t)search_q_st, (uint32_t)search_q_ed, q_str, ref, max_search_ref, key_len, strand,
                          c_t_offset, true);
            c_t_offset += max_search_ref - DSB_KMER9 - 3;
            if (w->n_sms == current_sms) break;
            if (w->sms[current_sms].t_pos > w->sms[max_sms_id].t_pos + 1000) break;
        }
        DsbSms *c_sms = w->sms + current_sms++;
        int max_score = (int)c_sms->len;
        const uint32_t max_pre_q = c_sms->q_pos + DSB_SMS_OVERLAP;
        const uint32_t max_pre_t = c_sms->t_pos + DSB_SMS_OVERLAP;
#if defined(DSB_DEVICE_CODE)
        {   // lanes take 32 predecessors at a time; the reference's `break` cuts the scan at the first
            // (highest-index) predecessor that passes both overlap tests and lies > 600 bp back
            const int32_t lane = (int32_t)(threadIdx.x & 31);
            const uint32_t c_len = c_sms->len, c_q = c_sms->q_pos, c_t = c_sms->t_pos;
            for (int32_t top = (int32_t)current_sms - 2; top >= 0; top -= 32) {
                const int32_t pi = top - lane;
                bool brk = false;
                int cand = -2147483647 - 1;
                if (pi >= 0) {
                    const DsbSms pv = w->sms[pi];
                    const int pre_q_ed = (int)(pv.q_pos + pv.len + DSB_KMER9 - 1);
                    const int pre_t_ed = (int)(pv.t_pos + pv.len + DSB_KMER9 - 1);
                    if (!((uint32_t)pre_q_ed > max_pre_q) && !((uint32_t)pre_t_ed > max_pre_t)) {
                        if (pv.t_pos + 600 < max_pre_t) brk = true;
                        else {
                            const int indel = (int)(pv.q_pos - pv.t_pos - (max_pre_q - max_pre_t));
                            const int ABS_indel = DSB_ABS(indel);
                            if (!(ABS_indel > 200)) {
                                int new_score = (int)(pv.score + c_len - (uint32_t)(ABS_indel >> 3));
                                if ((uint32_t)pre_q_ed > c_q || (uint32_t)pre_t_ed > c_t) {
                                    const int overlap_q = (int)((uint32_t)pre_q_ed - c_q);
                                    const int overlap_t = (int)((uint32_t)pre_t_ed - c_t);
                                    new_score -= DSB_MAX(overlap_q, overlap_t);
                                }
                                cand = new_score;
                            }
                        }
                    }
                }
                const uint32_t bm = __ballot_sync(0xffffffffu, brk);
                if (bm) {
                    const int32_t cut = __ffs((int)bm) - 1;          // lanes below `cut` were scanned before the break
                    if (lane < cut) max_score = DSB_MAX(max_score, cand);
                    break;
                }
                max_score = DSB_MAX(max_score, cand);
            }
            max_score = __reduce_max_sync(0xffffffffu, max_score);
        }
#else
        for (int32_t pi = (int32_t)current_sms - 2; pi >= 0; pi--) {
            const DsbSms *p = w->sms + pi;
            const int pre_q_ed = (int)(p->q_pos + p->len + DSB_KMER9 - 1);
            const int pre_t_ed = (int)(p->t_pos + p->len + DSB_KMER9 - 1);
            if ((uint32_t)pre_q_ed > max_pre_q) continue;
            if ((uint32_t)pre_t_ed > max_pre_t) continue;
            if (p->t_pos + 600 < max_pre_t) break;
            const int indel = (int)(p->q_pos - p->t_pos - (max_pre_q - max_pre_t));
            const int ABS_indel = DSB_ABS(indel);
            if (ABS_indel > 200) continue;
            int new_score = (int)(p->score + c_sms->len - (uint32_t)(ABS_indel >> 3));
            if ((uint32_t)pre_q_ed > c_sms->q_pos || (uint32_t)pre_t_ed > c_sms->t_pos) {
                const int overlap_q = (int)((uint32_t)pre_q_ed - c_sms->q_pos);
                const int overlap_t = (int)((uint32_t)pre_t_ed - c_sms->t_pos);
                new_score -= DSB_MAX(overlap_q, overlap_t);
            }
            max_score = DSB_MAX(max_score, new_score);
        }
#endif
        c_sms->score = (uint32_t)max_score;
        if (c_sms->len >= 8) {
            const uint32_t cs_len = c_sms->len;     // c_sms is invalidated by the nested sdp_middle
            const int comb = dsb_combine_chain(w, chain_ID, (int)(c_sms->t_pos - c_sms->q_pos), false, (int)c_sms->q_pos);
            if (comb >= 0) {
                total_max_score = (int)((uint32_t)DSB_MAX(score_ori, max_score) - cs_len +
                                        (uint32_t)dsb_sdp_middle(w, w->chains[comb].cur, q_str, strand, key_len));
                score_ori = total_max_score;
                max_sms_id = 0;
                w->n_sms = 0;
                dsb_push_sms(w, c_h->t_ed, c_h->q_ed, (uint32_t)(-DSB_KMER9));
                w->sms[0].score = (uint32_t)total_max_score;
                current_sms = 1;
                c_t_offset = c_h->t_ed;
                continue;
            }
        }
        if (total_max_score < max_score) {
            total_max_score = max_score;
            max_sms_id = (int)current_sms - 1;
        }
        if (c_sms->t_pos > w->sms[max_sms_id].t_pos + 1000) break;
    }
    c_h->q_ed = w->sms[max_sms_id].q_pos + w->sms[max_sms_id].len + DSB_KMER9;
    c_h->t_ed = w->sms[max_sms_id].t_pos + w->sms[max_sms_id].len + DSB_KMER9;
    DSB_TRACE("REND chain=%d ref=%u q_ed=%u t_ed=%u score=%d maxid=%d n=%d\n", chain_ID, c_h->ref_ID, c_h->q_ed, c_h->t_ed,
              total_max_score - 10000, max_sms_id, (int)w->n_sms);
    return total_max_score - 10000;
}

DSB_HDN int dsb_sdp_left(DsbWork *w, const uint8_t *q_str, int strand, int key_len, int chain_ID, uint32_t l_read,
                         int score_ori)
{
    const DsbIndexView *ix = w->ix;
    (void)l_read;
    score_ori += 10000;
    int total_max_score = score_ori;
    int max_sms_id = 0;
    DsbChain *c_h = w->chains + chain_ID;
    w->n_sms = 0;
    uint8_t *ref = w->ref_ext;
    dsb_fill(ref, 1000, DSB_UNINIT);                                   // uint8_t ref[1000] (src/cly.c:2683)
    dsb_push_sms(w, c_h->t_st, c_h->q_st, DSB_UNINIT * 0x01010101u);   // len is never written nor read there
    w->sms[0].score = (uint32_t)score_ori;
    uint32_t current_sms = 1;
    const uint64_t t_offset_global = ix->ref_off[c_h->ref_ID];
    uint32_t c_t_offset = c_h->t_st + 3;
    int last_search = 0;
    while (1) {
        if (w->overflow) break;
        if (w->n_sms == current_sms) {
            const uint32_t next_step = c_t_offset;
            if (next_step < DSB_MIN_SCORE_MEM) break;
            uint32_t max_search_ref;
            if (c_h->q_st < 600) {
                if (last_search) break;
                last_search = 1;
                max_search_ref = c_h->q_st + 60;
            } else
                max_search_ref = c_t_offset;
            max_search_ref = DSB_MIN(600u, max_search_ref);
            if (t_offset_global == 0 && c_t_offset < DSB_OVER_SEARCH + max_search_ref)   // "bug" branch, SURVEY.md Q24
                dsb_get_ref(w, ref, c_t_offset + t_offset_global - max_search_ref, max_search_ref, true);
            else
                dsb_get_ref(w, ref, c_t_offset + t_offset_global - max_search_ref - DSB_OVER_SEARCH,
                            max_search_ref + DSB_OVER_SEARCH, true);
            int search_q_st = (int)w->sms[max_sms_id].q_pos - 1000;
            search_q_st = DSB_MAX(search_q_st, 0);
            const int search_q_ed = (int)DSB_MIN((uint32_t)(search_q_st + 2000), c_h->q_st - 1);   // MIN(int, uint32): unsigned
            dsb_sdp_match(w, (uint32_t)search_q_st, (uint32_t)search_q_ed, q_str, ref + DSB_OVER_SEARCH, max_search_ref,
                          key_len, strand, c_t_offset - max_search_ref, false);
            c_t_offset = c_t_offset - max_search_ref + DSB_KMER9 + 3;
            if (w->n_sms == current_sms) break;
            if (w->sms[current_sms].t_pos + 1000 < w->sms[max_sms_id].t_pos) break;
        }
        DsbSms *c_sms = w->sms + current_sms++;
        int max_score = (int)c_sms->len;
        const uint32_t min_pre_q = c_sms->q_pos + c_sms->len - DSB_SMS_OVERLAP + DSB_KMER9 - 1;
        const uint32_t min_pre_t = c_sms->t_pos + c_sms->len - DSB_SMS_OVERLAP + DSB_KMER9 - 1;
#if defined(DSB_DEVICE_CODE)
        {
            const int32_t lane = (int32_t)(threadIdx.x & 31);
            const uint32_t c_len = c_sms->len;
            for (int32_t top = (int32_t)current_sms - 2; top >= 0; top -= 32) {
                const int32_t pi = top - lane;
                bool brk = false;
                int cand = -2147483647 - 1;
                if (pi >= 0) {
                    const DsbSms pv = w->sms[pi];
                    if (!(pv.q_pos < min_pre_q) && !(pv.t_pos < min_pre_t)) {
                        if (min_pre_t + 600 < pv.t_pos) brk = true;
                        else {
                            const int indel = (int)(pv.q_pos - pv.t_pos - (min_pre_q - min_pre_t));
                            const int ABS_indel = DSB_ABS(indel);
                            if (!(ABS_indel > 200)) {
                                int new_score = (int)(pv.score + c_len - (uint32_t)(ABS_indel >> 3));
                                if (min_pre_q + DSB_SMS_OVERLAP > pv.q_pos || min_pre_t + DSB_SMS_OVERLAP > pv.t_pos) {
                                    const int overlap_q = (int)(min_pre_q + DSB_SMS_OVERLAP - pv.q_pos);
                                    const int overlap_t = (int)(min_pre_t + DSB_SMS_OVERLAP - pv.t_pos);
                                    new_score -= DSB_MAX(overlap_q, overlap_t);
                                }
                                cand = new_score;
                            }
                        }
                    }
                }
                const uint32_t bm = __ballot_sync(0xffffffffu, brk);
                if (bm) {
                    const int32_t cut = __ffs((int)bm) - 1;
                    if (lane < cut) max_score = DSB_MAX(max_score, cand);
                    break;
                }
                max_score = DSB_MAX(max_score, cand);
            }
            max_score = __reduce_max_sync(0xffffffffu, max_score);
        }
#else
        for (int32_t pi = (int32_t)current_sms - 2; pi >= 0; pi--) {
            const DsbSms *p = w->sms + pi;
            if (p->q_pos < min_pre_q) continue;
            if (p->t_pos < min_pre_t) continue;
            if (min_pre_t + 600 < p->t_pos) break;
            const int indel = (int)(p->q_pos - p->t_pos - (min_pre_q - min_pre_t));
            const int ABS_indel = DSB_ABS(indel);
            if (ABS_indel > 200) continue;
            int new_score = (int)(p->score + c_sms->len - (uint32_t)(ABS_indel >> 3));
            if (min_pre_q + DSB_SMS_OVERLAP > p->q_pos || min_pre_t + DSB_SMS_OVERLAP > p->t_pos) {
                const int overlap_q = (int)(min_pre_q + DSB_SMS_OVERLAP - p->q_pos);
                const int overlap_t = (int)(min_pre_t + DSB_SMS_OVERLAP - p->t_pos);
                new_score -= DSB_MAX(overlap_q, overlap_t);
            }
            max_score = DSB_MAX(max_score, new_score);
        }
#endif
        c_sms->score = (uint32_t)max_score;
        if (c_sms->len >= 8) {
            const uint32_t cs_len = c_sms->len;
            const int comb = dsb_combine_chain(w, chain_ID, (int)(c_sms->t_pos - c_sms->q_pos), true,
                                               (int)(c_sms->q_pos + c_sms->len));
            if (comb >= 0) {
                total_max_score = (int)((uint32_t)DSB_MAX(score_ori, max_score) - cs_len +
                                        (uint32_t)dsb_sdp_middle(w, w->chains[comb].cur, q_str, strand, key_len));
                score_ori = total_max_score;
                max_sms_id = 0;
                w->n_sms = 0;
                dsb_push_sms(w, c_h->t_st, c_h->q_st, DSB_UNINIT * 0x01010101u);
                w->sms[0].score = (uint32_t)total_max_score;
                current_sms = 1;
                c_t_offset = c_h->t_st;
                continue;
            }
        }
        if (total_max_score < max_score) {
            total_max_score = max_score;
            max_sms_id = (int)current_sms - 1;
        }
        if (c_sms->t_pos + 1000 < w->sms[max_sms_id].t_pos) break;
    }
    c_h->q_st = w->sms[max_sms_id].q_pos;
    c_h->t_st = w->sms[max_sms_id].t_pos;
    return total_max_score - 10000;
}

DSB_HDN void dsb_get_score(DsbWork *w)
{
    DSB_PHASE_BEGIN();
    const int key_len = dsb_build_read_hash(w);
    DSB_PHASE_END(w, 4);
    for (uint32_t i = 0; i < w->n_chain; i++) {
        if (w->chains[i].sum_score == 0) continue;
        const int strand = (w->chains[i].direction == DSB_FORWARD) ? 0 : 1;
        const uint8_t *q_str = w->bin + (strand ? w->L : 0);
        DSB_TRACE("CH %u ref=%u dir=%d q=%u-%u t=%u-%u ss=%u an=%u indel=%u top=%d\n", i, w->chains[i].ref_ID,
                  w->chains[i].direction, w->chains[i].q_st, w->chains[i].q_ed, w->chains[i].t_st, w->chains[i].t_ed,
                  w->chains[i].sum_score, w->chains[i].anchor_number, w->chains[i].indel, w->chains[i].with_top);
        int score = dsb_sdp_middle(w, w->chains[i].cur, q_str, strand, key_len);
        DSB_TRACE("MID %u score=%d\n", i, score);
        score = dsb_sdp_right(w, q_str, strand, key_len, (int)i, w->L, score);
        score = dsb_sdp_left(w, q_str, strand, key_len, (int)i, w->L, score);
        w->chains[i].sum_score = (uint32_t)score;
        if (w->overflow) return;
    }
    DSB_PHASE_END(w, 5);
}

// ------------------------------------------------------------------------------------------
// filter / merge / rank (restates delete_small_score_rst src/cly.c:2878-2988 with its
// comparators chain_cmp_by_pos :2848-2865 and chain_cmp_by_MEM_score :53-63)
// ------------------------------------------------------------------------------------------
DSB_HDN void dsb_delete_small_score_rst(DsbWork *w)
{
    // only a read that arrives here WITH chains raises the worker's max_read_l (src/cly.c:2884-2885 return, :2953 update):
    // the host needs that bit to evaluate `max_read_l < 510` for the reads that follow (dsb_history.c)
    w->had_chain = w->n_chain != 0;
    if (w->n_chain == 0) return;
    if (w->n_chain > 200) {
        uint32_t rst_num = 200;
        for (; rst_num < w->n_chain && w->chains[rst_num].sum_score > 50; rst_num++) {}
        w->n_chain = rst_num;
    }
    w->n_chain = DSB_MIN(400u, w->n_chain);
    const uint32_t l_read = w->L;
    dsb_sc_hash_idx(w);
    dsb_get_score(w);
    if (w->overflow) return;
    DSB_PHASE_BEGIN();
    const uint32_t n = w->n_chain;
    DsbChain *C = w->chains;
    if (n > 1) {
        const DsbChain *CC = C;
        dsb_msort_idx(w->sort_a, w->sort_b, n, [CC](uint32_t i, uint32_t j) -> int {
            const DsbChain *a = CC + i, *b = CC + j;
            if (a->ref_ID > b->ref_ID) return 1;
            if (a->ref_ID < b->ref_ID) return -1;
            if (a->t_st > b->t_st) return 1;
            if (a->t_st < b->t_st) return -1;
            if (a->sum_score < b->sum_score) return 1;
            if (a->sum_score > b->sum_score) return -1;
            return 0;
        });
        dsb_apply_chain_perm(w, n);
    }
    for (uint32_t ci = 0; ci + 1 < n; ci++) {
        DsbChain *c_c = C + ci;
        if (c_c->sum_score == 0) continue;
        for (uint32_t ni = ci + 1; ni < n; ni++) {
            DsbChain *next_c = C + ni;
            if (c_c->ref_ID == next_c->ref_ID) {
                if (c_c->direction != next_c->direction) continue;
                if (next_c->sum_score == 0) continue;
                if (next_c->t_st < c_c->t_st + 5 && next_c->q_st < c_c->q_st + 5 && next_c->sum_score < c_c->sum_score + 5) {
                    next_c->sum_score = 0;
                    next_c->q_ed = next_c->q_st;
                    next_c->t_ed = next_c->t_st;
                    continue;
                }
                const int dis_t = (int)(next_c->t_st - c_c->t_ed);
                const int dis_q = (int)(next_c->q_st - c_c->q_ed);
                const int dis_t_q = DSB_ABS(dis_t - dis_q);
                if ((dis_t > -20 && dis_t < 1000 && dis_q > -20 && dis_q < 1000) && dis_t_q < 200) {
                    c_c->t_ed = DSB_MAX(c_c->t_ed, next_c->t_ed);
                    c_c->q_ed = DSB_MAX(c_c->q_ed, next_c->q_ed);
                    c_c->sum_score += next_c->sum_score;
                    next_c->sum_score = 0;
                    next_c->q_ed = next_c->q_st;
                    next_c->t_ed = next_c->t_st;
                }
            } else
                break;
        }
    }
    // buff->max_read_l is a running maximum over the worker's history; the host evaluates it in
    // read order and ships the outcome of `max_read_l < 510` with the read
    if (w->is_2g) {
        for (uint32_t ci = 0; ci < n; ci++) {
            const int score = (int)(C[ci].sum_score + ((C[ci].q_ed - C[ci].q_st) >> 5));
            if (score < 26) C[ci].sum_score = 0;
        }
    } else if (l_read < 310) {
        for (uint32_t ci = 0; ci < n; ci++) {
            const int score = (int)(C[ci].sum_score + ((C[ci].q_ed - C[ci].q_st) >> 5));
            if (score < 30) C[ci].sum_score = 0;
        }
    } else {
        const DsbIndexView *ix = w->ix;
        for (uint32_t ci = 0; ci < n; ci++) {
            const int score = (int)(C[ci].sum_score + ((C[ci].q_ed - C[ci].q_st) >> 5));
            // `q_ed - q_st < filter_min_length` compares uint32 with int: unsigned
            if (score < ix->filter_min_score_lv3 &&
                ((C[ci].q_ed - C[ci].q_st) < (uint32_t)ix->filter_min_length || score < ix->filter_min_score))
                C[ci].sum_score = 0;
        }
    }
    if (n > 1) {
        const DsbChain *CC = C;
        dsb_msort_idx(w->sort_a, w->sort_b, n, [CC](uint32_t i, uint32_t j) -> int {
            const DsbChain *a = CC + i, *b = CC + j;
            const int score_a = (int)(a->sum_score << 5);
            const int score_b = (int)(b->sum_score << 5);
            if (score_a < score_b) return 1;
            if (score_a > score_b) return -1;
            return (int)(a->sum_score % 2);
        });
        dsb_apply_chain_perm(w, n);
    }
    uint32_t k = 0;
    for (; k < n; k++)
        if (C[k].sum_score == 0) break;
    w->n_chain = k;
    DSB_PHASE_END(w, 6);
}

// ------------------------------------------------------------------------------------------
// PRIMARY / SECONDARY / SUPPLEMENTARY (restates detect_primary src/cly.c:2990-3053)
// ------------------------------------------------------------------------------------------
DSB_HDN void dsb_detect_primary(DsbWork *w)
{
    const uint32_t n_hit = w->n_chain;
    if (n_hit == 0) return;
    DsbChain *hit = w->chains;
    const uint32_t read_len = w->L;
    // primary_v[800] / primary_v_idx[800] of the reference live in the (now free) sort buffers
    int32_t *primary_v = (int32_t *)w->sort_a;
    uint8_t *primary_v_idx = (uint8_t *)w->sort_b;
    int n_primary_v = 1;
    hit[0].pri_index = primary_v_idx[0] = 0;
    primary_v[0] = 0;
    hit[0].primary = DSB_PRIMARY;
    for (uint32_t i = 0; i < n_hit; i++)
        if (hit[i].q_st > 4294960000u) hit[i].q_st = 0;
    for (uint32_t ci = 1; ci < n_hit; ci++) {
        DsbChain *c_hit = hit + ci;
        bool overlap = false;
        for (int i = 0; i < n_primary_v; i++) {
            const DsbChain *P = hit + primary_v[i];
            int primary_st, primary_ed;
            if (P->direction == c_hit->direction) {
                primary_st = (int)P->q_st;
                primary_ed = (int)P->q_ed;
            } else {
                primary_st = (int)(read_len - P->q_ed);
                primary_ed = (int)(read_len - P->q_st);
            }
            const uint32_t overlap_st = DSB_MAX(c_hit->q_st, (uint32_t)primary_st);   // MAX(uint32,int): unsigned
            const uint32_t overlap_ed = DSB_MIN(c_hit->q_ed, (uint32_t)primary_ed);
            if ((overlap_st < overlap_ed) && (((overlap_ed - overlap_st) << 1) >= (c_hit->q_ed - c_hit->q_st)))
                overlap = true;
            if (overlap) {
                c_hit->primary = DSB_SECONDARY;
                c_hit->pri_index = ++primary_v_idx[i];
                const int max_gap = (int)DSB_MAX((P->sum_score >> 6), 5u);
                if (c_hit->sum_score + (uint32_t)max_gap > P->sum_score) c_hit->pri_index = 1;
                if (primary_v_idx[i] == 255) primary_v_idx[i] = 254;
                break;
            }
        }
        if (!overlap) {
            c_hit->primary = DSB_SUPPLEMENTARY;
            c_hit->pri_index = primary_v_idx[n_primary_v] = 0;
            primary_v[n_primary_v++] = (int32_t)ci;
            if (n_primary_v > 750) n_primary_v = 750;
        }
    }
}

// ------------------------------------------------------------------------------------------
// per-read driver (restates getIsland's strand ordering src/cly.c:1256-1261 and classify_seq
// src/cly.c:3059-3127).  On entry w->bin / w->exist / arenas are set; on exit w->chains[0..n_chain)
// are the reported hits in output order.
// ------------------------------------------------------------------------------------------
// everything after fast seeding (src/cly.c:3096-3127); have_islands = w->seeds hold this read's islands
DSB_HDN void dsb_classify_tail(DsbWork *w, int dir0, int dir1, bool both_direction, bool have_islands)
{
    DSB_PHASE_BEGIN();
    dsb_resolve_tree(w);
    DSB_PHASE_END(w, 2);
    bool run_slow_mode = false;
    if (w->n_chain <= 0)
        run_slow_mode = true;
    else if (w->chains[0].anchor_number < 5) {                              // super_repeat is always 0
        run_slow_mode = true;
        if (w->L <= 300 && w->chains[0].sum_score > 200) run_slow_mode = false;
    }
    if (run_slow_mode && !w->overflow) {
        w->st.slow++;
        w->n_anchor = 0;
        if (!have_islands) dsb_islands_both(w);                             // split pipeline: the islands stayed in k_islands
        DSB_SUB_BEGIN();
        dsb_slow_classify(w, dir0);
        DSB_SUB_END(w, 7);                                                  // phase 7 = the seeding part of phase 3
        dsb_resolve_tree(w);
        if (both_direction || w->n_chain <= 0 || (w->chains[0].anchor_number < 5)) {
            DSB_SUB_BEGIN2();
            dsb_slow_classify(w, dir1);                                     // appends
            DSB_SUB_END(w, 7);
            dsb_resolve_tree(w);
        }
        DSB_PHASE_END(w, 3);
    }
    if (w->overflow) return;
    dsb_delete_small_score_rst(w);
    if (w->overflow) return;
    DSB_PHASE_BEGIN2();
    dsb_detect_primary(w);
    DSB_PHASE_END(w, 6);
}

DSB_HDN void dsb_classify_read(DsbWork *w)
{
    w->n_anchor = 0;
    w->n_chain = 0;
    w->fast_classify = 1;
    w->had_chain = 0;
    if (w->L < 40) return;                                                  // MIN_READ_LEN
    DSB_PHASE_BEGIN();
    dsb_islands_both(w);
    DSB_PHASE_END(w, 0);
    // search_dir[0] = forward unless the reverse strand scores strictly higher
    int dir0 = 0, dir1 = 1;
    if (w->total_score[0] < w->total_score[1]) { dir0 = 1; dir1 = 0; }
    const uint32_t s0 = w->total_score[dir0], s1 = w->total_score[dir1];
    const bool both_direction = ((s0 - s1) <= (s0 >> 3));
    dsb_fast_seeding(w, dir0, both_direction ? dir1 : -1);
    DSB_PHASE_END(w, 1);
    dsb_classify_tail(w, dir0, dir1, both_direction, true);
}

#if defined(__CUDACC__)
// Split pipeline, last kernel: the read's islands were searched by k_seed; append their anchors in island order and
// drop the islands the serial loop would have skipped (the merge of dsb_fast_seeding), then carry on as usual.
__device__ __noinline__ void dsb_classify_seeded(DsbWork *w, const DsbReadSeeds *rs, const DsbSeedJob *jobs, const DsbAnchor *pool)
{
    w->n_anchor = 0;
    w->n_chain = 0;
    w->fast_classify = 1;
    w->had_chain = 0;
    if (w->L < 40) return;
    w->total_score[0] = rs->total_score[0];
    w->total_score[1] = rs->total_score[1];
    int dir0 = 0, dir1 = 1;
    if (w->total_score[0] < w->total_score[1]) { dir0 = 1; dir1 = 0; }
    const uint32_t s0 = w->total_score[dir0], s1 = w->total_score[dir1];
    const bool both_direction = ((s0 - s1) <= (s0 >> 3));
    const DsbSeedJob *jv = jobs + rs->job_base;
    uint32_t skip_key = 0xffffffffu;
    for (uint32_t j = 0; j < rs->n_jobs; j++) {
        const DsbSeedJob jb = jv[j];
        if (jb.strand_ci == skip_key) { skip_key = 0xffffffffu; continue; }   // `ci++` of src/cly.c:1533
        skip_key = jb.skip ? jb.strand_ci + 1 : 0xffffffffu;
        if (jb.count == 0) continue;
        if (w->n_anchor + jb.count > w->cap_anchor) { w->overflow = 1; return; }
        const DsbAnchor *src = pool + jb.a_off;
        DsbAnchor *dst = w->anchors + w->n_anchor;
        for (uint32_t k = threadIdx.x & 31; k < jb.count; k += 32) dst[k] = src[k];
        w->n_anchor += jb.count;
    }
    __syncwarp();
    dsb_classify_tail(w, dir0, dir1, both_direction, false);
}
#endif
