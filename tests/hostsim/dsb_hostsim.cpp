// dsb_hostsim.cpp -- TEST INFRASTRUCTURE, NOT A PRODUCT PATH.
//
// Implements the C-ABI of include/dsb_cuda.h by compiling the device algorithm
// (desamba-so_b200/csrc/dsb_core.cuh) for the host with g++ and running it serially.
// Purpose: there is no GPU in the development container, so bit-exactness of the kernel
// source against the reference is debugged here; the -m gpu tests then run the very same source
// as sm_100a kernels.  libdesamba.so / deSAMBA never link this file (they link libdsb_cuda.so
// and abort when no CUDA device exists); only tests/ build and load it.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include <string>
#include "dsb_cuda.h"
#include "dsb_core.cuh"
#include "dsb_layout.cuh"

static std::string g_err;
static int g_launches = 0;

struct dsb_index {
    DsbIndexView v;
    std::vector<DsbFmBlock> fm;
    std::vector<uint64_t> lut, sa, unitig, ref_pos, ref_off, ref_len;
    std::vector<uint8_t> ek0, ek1, ref_bin;
    std::vector<int32_t> q_mem, q_lv;
    uint64_t bytes;
};

struct dsb_slot {
    dsb_index *ix;
    uint64_t max_bases;
    uint32_t max_reads;
    std::vector<char> seq;
    std::vector<dsb_read_desc> desc;
    std::vector<uint8_t> edge;
    std::vector<dsb_read_result> results;
    std::vector<dsb_hit> hits;
    dsb_counters counters;
    uint32_t n_reads;
    uint64_t n_bases;
};

extern "C" {

const char *dsb_last_error(void) { return g_err.c_str(); }
int dsb_kernel_launch_count(void) { return g_launches; }
int dsb_gpu_count(void)
{
    const char *e = getenv("DSB_HOSTSIM_GPUS");
    return e ? atoi(e) : 1;
}

int dsb_index_upload(const dsb_host_index *h, int dev, dsb_index **out)
{
    (void)dev;
    dsb_index *ix = new dsb_index();
    const uint64_t nblk = dsb_fm_blocks_needed(h->bwt_bytes);
    ix->fm.resize(nblk);
    uint64_t bad = 0;
    uint64_t dollar_row = ~0ull;
    dsb_fm_reblock(h->bwt_occ, h->bwt_bytes, &dollar_row, ix->fm.data(), &bad);
    if (bad) fprintf(stderr, "[hostsim] warning: %llu FM blocks where the derived '#' count differs from the stored one\n",
                     (unsigned long long)bad);
    const uint64_t lut_n = (1ull << 26) + 1;
    ix->lut.assign(h->hash_index, h->hash_index + lut_n);
    ix->sa.assign(h->sa_taxon, h->sa_taxon + h->sa_n);
    ix->ek0.assign(h->e_kmer0, h->e_kmer0 + h->e_kmer_size);
    ix->ek1.assign(h->e_kmer1, h->e_kmer1 + h->e_kmer_size);
    ix->unitig.assign(h->unitig, h->unitig + h->n_unitig);
    ix->ref_pos.assign(h->ref_pos, h->ref_pos + h->n_ref_pos);
    ix->ref_pos.push_back(0);
    ix->ref_bin.assign(h->ref_bin, h->ref_bin + h->ref_bin_n);
    ix->ref_off.assign(h->ref_seq_offset, h->ref_seq_offset + h->n_ref);
    ix->ref_len.assign(h->ref_seq_len, h->ref_seq_len + h->n_ref);
    ix->q_mem.assign(h->q_mem, h->q_mem + 2000);
    ix->q_lv.assign(h->q_lv, h->q_lv + 400);
    DsbIndexView &v = ix->v;
    memset(&v, 0, sizeof v);
    v.fm = ix->fm.data();
    v.n_fm_blocks = nblk;
    memcpy(v.rank, h->rank, sizeof v.rank);
    v.dollar_pos = h->dollar_pos;
    v.dollar_row = dollar_row;
    v.lut = ix->lut.data();
    v.sa = ix->sa.data();
    v.sa_n = h->sa_n;
    v.ek0 = ix->ek0.data();
    v.ek1 = ix->ek1.data();
    int k;
    uint64_t mask;
    dsb_ekmer_params(h->e_kmer_size, &k, &mask);
    v.ek_mask = mask;
    v.ek_len = k;
    v.single_base_max = (int)(0.8 * k);
    v.unitig = ix->unitig.data();
    v.n_unitig = h->n_unitig;
    v.ref_pos = ix->ref_pos.data();
    v.ref_bin = ix->ref_bin.data();
    v.ref_bin_n = h->ref_bin_n;
    v.ref_off = ix->ref_off.data();
    v.ref_len = ix->ref_len.data();
    v.n_ref = h->n_ref;
    v.q_mem = ix->q_mem.data();
    v.q_lv = ix->q_lv.data();
    v.filter_min_length = h->filter_min_length;
    v.filter_min_score = h->filter_min_score;
    v.filter_min_score_lv3 = h->filter_min_score_lv3;
    ix->bytes = nblk * 64 + lut_n * 8 + h->sa_n * 8 + 2 * h->e_kmer_size + h->ref_bin_n;
    *out = ix;
    return 0;
}

void dsb_index_free(dsb_index *ix) { delete ix; }

int dsb_index_clone(const dsb_index *src, int dev, dsb_index **out)
{
    (void)dev;
    dsb_index *ix = new dsb_index(*src);                  // simulated "devices" share nothing: a deep copy, pointers rebound
    DsbIndexView &v = ix->v;
    v.fm = ix->fm.data(); v.lut = ix->lut.data(); v.sa = ix->sa.data(); v.ek0 = ix->ek0.data(); v.ek1 = ix->ek1.data();
    v.unitig = ix->unitig.data(); v.ref_pos = ix->ref_pos.data(); v.ref_bin = ix->ref_bin.data();
    v.ref_off = ix->ref_off.data(); v.ref_len = ix->ref_len.data(); v.q_mem = ix->q_mem.data(); v.q_lv = ix->q_lv.data();
    *out = ix;
    return 0;
}
uint64_t dsb_index_device_bytes(const dsb_index *ix) { return ix->bytes; }

int dsb_slot_create(dsb_index *ix, uint64_t max_bases, uint32_t max_reads, dsb_slot **out)
{
    dsb_slot *s = new dsb_slot();
    s->ix = ix;
    s->max_bases = max_bases;
    s->max_reads = max_reads;
    s->seq.resize(max_bases + 64);
    s->desc.resize(max_reads);
    s->edge.resize((size_t)max_reads * (DSB_HEAD_BYTES + DSB_TAIL_BYTES));
    s->n_reads = 0;
    *out = s;
    return 0;
}

void dsb_slot_free(dsb_slot *s) { delete s; }

int dsb_slot_staging(dsb_slot *s, char **seq, dsb_read_desc **desc, uint8_t **edge)
{
    *seq = s->seq.data();
    *desc = s->desc.data();
    *edge = s->edge.data();
    return 0;
}

static int run_batch(dsb_slot *s, uint32_t cap_anchor, uint32_t cap_sms)
{
    const DsbIndexView *ix = &s->ix->v;
    const uint32_t n = s->n_reads;
    uint32_t Lmax = 0;
    for (uint32_t i = 0; i < n; i++) Lmax = s->desc[i].len > Lmax ? s->desc[i].len : Lmax;
    DsbScratchDims dims;
    dsb_scratch_dims(&dims, Lmax, cap_anchor, cap_sms);
    std::vector<uint8_t> scratch(dims.bytes);
    std::vector<uint8_t> bin(dsb_bin_bytes(s->n_bases, n));
    std::vector<uint8_t> exist(2 * s->n_bases + 64);
    s->results.assign(n, dsb_read_result());
    s->hits.clear();
    memset(&s->counters, 0, sizeof s->counters);
    const int k = ix->ek_len;
    g_launches += 3;
    for (uint32_t i = 0; i < n; i++) {
        const dsb_read_desc &d = s->desc[i];
        const uint32_t L = d.len;
        uint8_t *b = bin.data() + dsb_bin_off(d.seq_off, i);
        const uint8_t *edge = s->edge.data() + (size_t)i * (DSB_HEAD_BYTES + DSB_TAIL_BYTES);
        // "encode kernel"
        memcpy(b, edge, DSB_HEAD_PAD);
        uint8_t *F = b + DSB_HEAD_PAD, *R = F + L;
        for (uint32_t p = 0; p < L; p++) {
            F[p] = dsb_encode_base((uint8_t)s->seq[d.seq_off + p]);
            R[L - 1 - p] = (uint8_t)(3 - F[p]);
        }
        memcpy(R + L, edge + DSB_HEAD_BYTES, DSB_TAIL_PAD);
        // "exist kernel"
        uint8_t *ex = exist.data() + 2 * d.seq_off;
        if (L >= (uint32_t)k) {
            for (uint32_t p = 0; p + k <= L; p++) {
                uint32_t p0 = 0, p1 = 0;
                ex[p] = dsb_exist_kmer(ix, F + p, &p0, &p1);
                s->counters.exist_probes0 += p0; s->counters.exist_probes1 += p1;
                p0 = p1 = 0;
                ex[L + p] = dsb_exist_kmer(ix, R + p, &p0, &p1);
                s->counters.exist_probes0 += p0; s->counters.exist_probes1 += p1;
            }
        }
        // "classify kernel"
        DsbWork w;
        memset(&w, 0, sizeof w);
        w.ix = ix;
        w.bin = F;
        w.L = L;
        w.is_2g = (d.flags & DSB_READ_2G) ? 1 : 0;
        w.exist[0] = ex;
        w.exist[1] = ex + L;
        dsb_bind_scratch(&w, scratch.data(), &dims);
        dsb_classify_read(&w);
        if (w.overflow) return (int)w.overflow;
        dsb_read_result &r = s->results[i];
        r.hit_off = (uint32_t)s->hits.size();
        r.n_hit = w.n_chain;
        r.fast_classify = (w.fast_classify ? DSB_RES_FAST : 0u) | (w.had_chain ? DSB_RES_HAD_CHAIN : 0u);
        r.n_anchor = w.n_anchor;
        for (uint32_t c = 0; c < w.n_chain; c++) {
            const DsbChain &ch = w.chains[c];
            dsb_hit h;
            memset(&h, 0, sizeof h);
            h.ref_ID = ch.ref_ID; h.t_st = ch.t_st; h.t_ed = ch.t_ed; h.q_st = ch.q_st; h.q_ed = ch.q_ed;
            h.sum_score = ch.sum_score; h.indel = ch.indel; h.anchor_number = ch.anchor_number;
            h.direction = ch.direction; h.primary = ch.primary; h.pri_index = ch.pri_index; h.with_top_anchor = ch.with_top;
            s->hits.push_back(h);
        }
        s->counters.occ_queries += w.st.occ; s->counters.lut_lookups += w.st.lut; s->counters.sa_lookups += w.st.sa;
        s->counters.unitig_reads += w.st.unitig; s->counters.refpos_reads += w.st.refpos;
        s->counters.ref_window_bytes += w.st.ref_bytes; s->counters.mem_searches += w.st.mem_search;
        s->counters.map_seeds += w.st.map_seed; s->counters.sdp_windows += w.st.sdp_win;
        s->counters.anchors += w.n_anchor; s->counters.slow_reads += w.st.slow;
    }
    return 0;
}

int dsb_slot_submit(dsb_slot *s, uint32_t n_reads, uint64_t n_bases)
{
    s->n_reads = n_reads;
    s->n_bases = n_bases;
    uint32_t cap_anchor = 8192, cap_sms = 16384;
    for (int attempt = 0; attempt < 6; attempt++) {
        int rc = run_batch(s, cap_anchor, cap_sms);
        if (rc == 0) return 0;
        if (rc == 2) { g_err = "reference would abort: sdp_middle window >= 2000 (src/cly.c:2468)"; return -2; }
        cap_anchor *= 4;
        cap_sms *= 4;
    }
    g_err = "arena overflow";
    return -1;
}

int dsb_slot_wait(dsb_slot *s, const dsb_read_result **results, const dsb_hit **hits, uint32_t *n_hits,
                  dsb_counters *counters)
{
    if (s->hits.empty()) s->hits.resize(1);
    *results = s->results.data();
    *hits = s->hits.data();
    *n_hits = (uint32_t)s->hits.size();
    if (counters) *counters = s->counters;
    return 0;
}

int dsb_slot_timing_stages(dsb_slot *, float *a, float *b, float *c) { *a = *b = *c = 0.f; return 0; }
int dsb_slot_timing(dsb_slot *s, float *a, float *b, float *c)
{
    (void)s;
    *a = *b = *c = 0;
    return 0;
}

int dsb_slot_rerun_device(dsb_slot *s) { return dsb_slot_submit(s, s->n_reads, s->n_bases); }

int dsb_fm_rank_batch(dsb_index *ix, const uint64_t *row, const uint8_t *c, uint64_t n, uint64_t *out, uint8_t *c_out,
                      int repeat, float *ms)
{
    (void)repeat;
    DsbWork w;
    memset(&w, 0, sizeof w);
    w.ix = &ix->v;
    for (uint64_t i = 0; i < n; i++) {
        uint8_t cc = c[i];
        out[i] = dsb_occ(&w, row[i], &cc);
        c_out[i] = cc;
    }
    if (ms) *ms = 0;
    g_launches++;
    return 0;
}

int dsb_exist_batch(dsb_index *ix, const uint8_t *codes, uint64_t n, uint8_t *exist, int repeat, float *ms)
{
    (void)repeat;
    const uint64_t k = (uint64_t)ix->v.ek_len;
    if (n < k) return -1;
    uint32_t p0 = 0, p1 = 0;
    for (uint64_t i = 0; i + k <= n; i++) exist[i] = dsb_exist_kmer(&ix->v, codes + i, &p0, &p1);
    if (ms) *ms = 0;
    g_launches++;
    return 0;
}

}  // extern "C"

extern "C" int dsb_random_access_probe(int dev, uint64_t table_bytes, uint32_t chunk_bytes, uint64_t n_access, int repeat, float *ms)
{
    (void)dev; (void)table_bytes; (void)chunk_bytes; (void)n_access; (void)repeat;
    if (ms) *ms = 0;
    g_err = "the host simulator has no memory system to probe";
    return -1;
}
