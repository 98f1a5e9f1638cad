// dsb_layout.cuh -- memory layout shared by the CUDA layer and the host simulator:
//   * how a batch is laid out in HBM (encoded reads with their head/tail margins, exist bytes)
//   * how one warp's private scratch slice is carved into the arenas of DsbWork
//   * the FM re-blocking done once at index upload
#pragma once
#include <stddef.h>
#include <string.h>
#include "dsb_core.cuh"

// ---------------------------------------------------------------------------------------
// batch layout: read i owns
//   bin  : [HEAD_PAD | F (L) | R (L) | TAIL_PAD], start = bin_base + bin_off(i)
//   exist: forward k-mer bytes at ex_base + 2*seq_off(i), reverse at + L
// bin_off(i) = 2*seq_off(i) + i*(HEAD_PAD+TAIL_PAD): no per-read offset table is needed.
// ---------------------------------------------------------------------------------------
DSB_HD uint64_t dsb_bin_off(uint64_t seq_off, uint32_t read_idx)
{
    return 2 * seq_off + (uint64_t)read_idx * (DSB_HEAD_PAD + DSB_TAIL_PAD);
}
DSB_HD uint64_t dsb_bin_bytes(uint64_t n_bases, uint32_t n_reads)
{
    return 2 * n_bases + (uint64_t)n_reads * (DSB_HEAD_PAD + DSB_TAIL_PAD) + 64;
}

// CLY_Bit (src/cly.c:16-34): A/a 0, G/g 2, T/t 3, every other byte (C, N, ...) 1
DSB_HD uint8_t dsb_encode_base(uint8_t ch)
{
    const uint8_t u = ch & 0xDF;            // fold case for letters
    if (ch == 'A' || ch == 'a') return 0;
    if (ch == 'G' || ch == 'g') return 2;
    if (ch == 'T' || ch == 't') return 3;
    (void)u;
    return 1;
}

// ---------------------------------------------------------------------------------------
// per-warp scratch
// ---------------------------------------------------------------------------------------
struct DsbScratchDims {
    uint32_t Lmax;          // longest read of the batch
    uint32_t cap_anchor;    // anchors == chains == sort buffers capacity
    uint32_t cap_sms;
    uint32_t head_entries;  // 16-bit bucket heads of the position-blocked 9-mer table: Lmax rounded up to whole segments
    uint32_t cap_jobs;      // top islands of both strands (<= L/100 + 1 each)
    uint32_t cap_stage;     // anchors one lane may stage during fast seeding
    uint32_t slow_lane_mems; // MEM slots of a lane of the slow path (0 = DSB_SLOW_LANE_MEMS; tests shrink it to force the serial fallback)
    // byte offsets inside a slice
    size_t o_seeds0, o_seeds1, o_anchors, o_anchors_tmp, o_chains, o_chains_tmp, o_sort_a, o_sort_b, o_score_v,
        o_sp_set, o_mems, o_jobs, o_stage, o_lane_priv, o_sms_stage, o_sms, o_head0, o_head1, o_ent0, o_ent1, o_sc_hash, o_ref_mid, o_ref_ext;
    size_t bytes;           // slice size (multiple of 256)
};

static inline size_t dsb_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static inline void dsb_scratch_dims(DsbScratchDims *d, uint32_t Lmax, uint32_t cap_anchor, uint32_t cap_sms)
{
    memset(d, 0, sizeof *d);
    if (Lmax < 64) Lmax = 64;
    d->Lmax = Lmax;
    d->cap_anchor = cap_anchor;
    d->cap_sms = cap_sms;
    d->head_entries = ((Lmax + DSB_SEG_MASK) >> DSB_SEG_BITS) << DSB_SEG_BITS;
    size_t o = 0;
#define DSB_TAKE(field, nbytes) d->field = o; o = dsb_align_up(o + (nbytes), 64)
    const size_t n_seed = (size_t)(Lmax / 4) + 8;       // < L/4 islands per strand
    DSB_TAKE(o_seeds0, n_seed * sizeof(DsbSeed));
    DSB_TAKE(o_seeds1, n_seed * sizeof(DsbSeed));
    DSB_TAKE(o_anchors, (size_t)cap_anchor * sizeof(DsbAnchor));
    DSB_TAKE(o_anchors_tmp, (size_t)cap_anchor * sizeof(DsbAnchor));
    DSB_TAKE(o_chains, (size_t)cap_anchor * sizeof(DsbChain));
    DSB_TAKE(o_chains_tmp, (size_t)cap_anchor * sizeof(DsbChain));
    DSB_TAKE(o_sort_a, (size_t)cap_anchor * 8);          // 2 x cap: the bitonic anchor sort pads to a power of two
    DSB_TAKE(o_sort_b, (size_t)cap_anchor * 4);
    DSB_TAKE(o_score_v, 1024 * 4);
    DSB_TAKE(o_sp_set, 512 * 8);
    DSB_TAKE(o_mems, 256 * sizeof(DsbMem));
    d->cap_jobs = 2 * (Lmax / DSB_SEED_RANGE + 2) + 8;
    d->cap_stage = cap_anchor / 8;
    DSB_TAKE(o_jobs, (size_t)d->cap_jobs * sizeof(DsbJob));
    DSB_TAKE(o_stage, (size_t)32 * d->cap_stage * sizeof(DsbAnchor));
    DSB_TAKE(o_lane_priv, (size_t)32 * DSB_LANE_PRIV_BYTES);
    DSB_TAKE(o_sms_stage, (size_t)32 * DSB_SDP_STAGE * sizeof(DsbSms));
    DSB_TAKE(o_sms, (size_t)cap_sms * sizeof(DsbSms));
    DSB_TAKE(o_head0, (size_t)d->head_entries * 2 + 64);
    DSB_TAKE(o_head1, (size_t)d->head_entries * 2 + 64);
    DSB_TAKE(o_ent0, (size_t)Lmax * 4);
    DSB_TAKE(o_ent1, (size_t)Lmax * 4);
    DSB_TAKE(o_sc_hash, (256 + 2 * 400 + 8) * sizeof(DsbScHash));
    DSB_TAKE(o_ref_mid, 2000 + 128);
    DSB_TAKE(o_ref_ext, 1000 + 192);
#undef DSB_TAKE
    d->bytes = dsb_align_up(o, 256);
}

DSB_HD void dsb_bind_scratch(DsbWork *w, uint8_t *base, const DsbScratchDims *d)
{
    w->seeds[0] = (DsbSeed *)(base + d->o_seeds0);
    w->seeds[1] = (DsbSeed *)(base + d->o_seeds1);
    w->anchors = (DsbAnchor *)(base + d->o_anchors);
    w->anchors_tmp = (DsbAnchor *)(base + d->o_anchors_tmp);
    w->chains = (DsbChain *)(base + d->o_chains);
    w->chains_tmp = (DsbChain *)(base + d->o_chains_tmp);
    w->cap_anchor = d->cap_anchor;
    w->cap_chain = d->cap_anchor;
    w->sort_a = (uint32_t *)(base + d->o_sort_a);
    w->sort_b = (uint32_t *)(base + d->o_sort_b);
    w->score_v = (int32_t *)(base + d->o_score_v);
    w->sp_set = (uint64_t *)(base + d->o_sp_set);
    w->mems = (DsbMem *)(base + d->o_mems);
    w->jobs = (DsbJob *)(base + d->o_jobs);
    w->cap_jobs = d->cap_jobs;
    w->stage = (DsbAnchor *)(base + d->o_stage);
    w->cap_stage = d->cap_stage;
    w->slow_lane_mems = d->slow_lane_mems;
    w->lane_priv = base + d->o_lane_priv;
    w->solo = 0;
    w->sms = (DsbSms *)(base + d->o_sms);
    w->sms_stage = (DsbSms *)(base + d->o_sms_stage);
    w->cap_sms = d->cap_sms;
    w->h_head[0] = (uint32_t *)(base + d->o_head0);
    w->h_head[1] = (uint32_t *)(base + d->o_head1);
    w->h_ent[0] = (uint32_t *)(base + d->o_ent0);
    w->h_ent[1] = (uint32_t *)(base + d->o_ent1);
    w->sc_hash = (DsbScHash *)(base + d->o_sc_hash);
    w->ref_mid = base + d->o_ref_mid;
    w->ref_ext = base + d->o_ref_ext;
}

// ---------------------------------------------------------------------------------------
// FM re-blocking (host, once per upload).  Reference block b (168 B): 5 x u64 counts of
// A,C,G,T,# before symbol 256*b, then 128 B of nibbles (src/bwt.c:32-41, :214-220).
// Ours: four 64-symbol blocks per reference block plus one terminal block so that r == n works.
// Returns the number of DsbFmBlock written (= 4*nb + 1); out must hold that many.
// ---------------------------------------------------------------------------------------
static inline uint64_t dsb_fm_blocks_needed(uint64_t bwt_bytes) { return (bwt_bytes / 168) * 4 + 1; }

// *dollar_row = the row whose symbol is '$' (nibble 5), ~0 if there is none; *mismatch_hash_count = reference blocks whose
// stored '#' count differs from start - (A+C+G+T) - ['$' before start], the value dsb_occ_ix derives (must be 0).
static inline uint64_t dsb_fm_reblock(const uint8_t *bwt_occ, uint64_t bwt_bytes, uint64_t *dollar_row, DsbFmBlock *out,
                                      uint64_t *mismatch_hash_count)
{
    const uint64_t nb = bwt_bytes / 168;
    uint64_t bad = 0;
    uint64_t d_row = ~0ull;
    for (uint64_t b = 0; b < nb && d_row == ~0ull; b++) {
        const uint8_t *data = bwt_occ + b * 168 + 40;
        for (int i = 0; i < 128; i++) {
            if ((data[i] & 0xf) == 5) { d_row = b * 256 + 2 * (uint64_t)i; break; }
            if ((data[i] >> 4) == 5) { d_row = b * 256 + 2 * (uint64_t)i + 1; break; }
        }
    }
    if (dollar_row) *dollar_row = d_row;
    uint64_t run[4] = {0, 0, 0, 0};
    for (uint64_t b = 0; b < nb; b++) {
        const uint8_t *blk = bwt_occ + b * 168;
        uint64_t base[5];
        memcpy(base, blk, 40);
        const uint8_t *data = blk + 40;
        uint64_t cnt[4] = {base[0], base[1], base[2], base[3]};
        // cross-check our derived '#' count against the stored one
        {
            const uint64_t start = b * 256;
            const uint64_t hash = start - (cnt[0] + cnt[1] + cnt[2] + cnt[3]) - ((d_row < start) ? 1 : 0);
            if (hash != base[4]) bad++;
        }
        for (int s = 0; s < 4; s++) {
            DsbFmBlock *o = out + b * 4 + s;
            for (int c = 0; c < 4; c++) o->cnt[c] = cnt[c];
            for (int pl = 0; pl < 4; pl++) o->plane[pl] = 0;
            for (int i = 0; i < 64; i++) {              // symbol i: low nibble of byte i/2 for even i (src/bwt.c:52)
                const uint8_t nib = (uint8_t)((data[32 * s + (i >> 1)] >> ((i & 1) << 2)) & 0xf);
                for (int pl = 0; pl < 4; pl++) o->plane[pl] |= (uint64_t)((nib >> pl) & 1) << i;
            }
            for (int i = 0; i < 32; i++) {
                const uint8_t v = data[32 * s + i];
                const uint8_t lo = v & 0xf, hi = v >> 4;
                if (lo < 4) cnt[lo]++;
                if (hi < 4) cnt[hi]++;
            }
        }
        for (int c = 0; c < 4; c++) run[c] = cnt[c];
    }
    DsbFmBlock *t = out + nb * 4;
    for (int c = 0; c < 4; c++) t->cnt[c] = run[c];
    memset(t->plane, 0, 32);
    if (mismatch_hash_count) *mismatch_hash_count = bad;
    return nb * 4 + 1;
}

// k and mask of the exist-k-mer tables from their size (src/idx.c:966-982)
static inline void dsb_ekmer_params(uint64_t table_bytes, int *k, uint64_t *mask)
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
    *mask = (bits == 64) ? ~0ull : ((1ull << bits) - 1ull);
}
