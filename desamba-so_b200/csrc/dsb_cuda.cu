// dsb_cuda.cu -- sm_100a kernels + the C-ABI of include/dsb_cuda.h (libdsb_cuda.so).
//
// Five kernels per batch, all integer / byte work bounded by random HBM traffic:
//   k_encode   : ASCII -> 2-bit codes, both strands, plus the head/tail margins of the history model
//   k_exist    : one thread per k-mer position and strand: k-mer, low-complexity mask, two hash probes into the
//                exist-k-mer bit tables (reference src/cly.c:951-967)
//   k_islands  : warp per read: the island scan of both strands over the exist bytes, `top` marking, one job per top
//                island into the batch-wide job list (src/cly.c:1066-1229)
//   k_seed     : persistent grid, ONE LANE PER ISLAND of the whole batch, lanes in lock step: FM backward search, locate,
//                Landau-Vishkin anchors into the batch-wide anchor pool (src/cly.c:1339-1541, 701-934)
//   k_classify : persistent grid, ONE WARP PER READ pulling reads from an atomic queue: chaining, sparse match-block
//                scoring, filters, PRI/SEC/SUP; the slow path (no usable chain) re-seeds the read inside the warp
// DSB_SPLIT=0 runs islands + seeding inside k_classify (round 1's form, kept for the parity tests of both forms);
// DSB_EXIST_ON_DEMAND=1 drops k_exist and lets the island scan probe the tables itself.  The index builder's kernels live
// in dsb_build.cu.
// There is no CPU path in this file: every entry point fails with an error when CUDA fails.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <mutex>
#include <atomic>
#include <algorithm>
#include "dsb_cuda.h"
#include "dsb_core.cuh"
#include "dsb_layout.cuh"

#ifndef DSB_CLASSIFY_THREADS
#define DSB_CLASSIFY_THREADS 128
#endif

static thread_local std::string g_err;
static std::atomic<int> g_launches{0};

static int fail(const char *what, cudaError_t e)
{
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return -1;
}
#define CU(call)                                        \
    do {                                                \
        cudaError_t e_ = (call);                        \
        if (e_ != cudaSuccess) return fail(#call, e_);  \
    } while (0)

// ------------------------------------------------------------------------------------------
// device-side control block of a batch
// ------------------------------------------------------------------------------------------
struct DsbCtrl {
    uint32_t next_read;      // work queue head of k_classify
    uint32_t n_hits;         // hit arena allocator
    uint32_t overflow;       // bit0 arena too small, bit1 reference would abort, bit2 hit arena too small, bit3 anchor pool / job list too small, bit4 a read of >= 2^24 bases
    uint32_t next_island_read;   // work queue head of k_islands
    uint32_t n_jobs;         // job list allocator (k_islands), then the job count k_seed works through
    uint32_t next_job;       // work queue head of k_seed
    uint32_t n_pool;         // anchor pool allocator (k_seed)
    uint32_t pad;
    dsb_counters counters;
};

struct dsb_index {
    int dev;
    DsbIndexView v;                 // pointers are DEVICE pointers
    std::vector<void *> allocs;
    std::vector<size_t> alloc_bytes;   // same order as allocs (dsb_index_clone copies array by array)
    uint64_t bytes;
    int sm_count;
};

struct dsb_slot {
    dsb_index *ix;
    cudaStream_t stream;
    uint64_t max_bases;
    uint32_t max_reads;
    // pinned host staging
    char *h_seq;
    dsb_read_desc *h_desc;
    uint8_t *h_edge;
    dsb_read_result *h_results;
    dsb_hit *h_hits;
    DsbCtrl *h_ctrl;
    uint32_t *h_order;
    // device
    char *d_seq;
    dsb_read_desc *d_desc;
    uint8_t *d_edge;
    uint8_t *d_bin;
    uint8_t *d_exist;
    uint32_t *d_h1, *d_h2;      // the two exist hashes of every k-mer (sliced probing)
    int exist_slices;           // > 0: k_exist_hash + 2 x exist_slices launches of k_exist_pass instead of k_exist
    uint32_t exist_slice_shift;
    dsb_read_result *d_results;
    dsb_hit *d_hits;
    DsbCtrl *d_ctrl;
    uint32_t *d_order;
    uint8_t *d_scratch;
    size_t scratch_bytes;
    // split pipeline (k_islands -> k_seed -> k_classify)
    bool split;
    bool exist_bytes;        // k_exist writes one byte per k-mer first (default); DSB_EXIST_ON_DEMAND=1: the island scan probes the tables itself
    DsbReadSeeds *d_rinfo;
    DsbSeedJob *d_jobs;
    DsbAnchor *d_pool;
    uint32_t cap_jobs, cap_pool;
    uint32_t cap_hits, h_cap_hits;
    uint32_t cap_anchor, cap_sms;
    DsbScratchDims dims;
    int n_warps, blocks, threads;
    int seed_blocks;                // grid of k_seed (its own occupancy)
    // state of the batch in flight
    uint32_t n_reads;
    uint64_t n_bases;
    uint32_t Lmax;
    bool in_flight;
    cudaEvent_t ev[8];              // 0 start, 1 after H2D+encode, 2 after exist, 3 after classify, 4 end, 6 after islands, 7 after seeding
    float ms_total, ms_exist, ms_classify, ms_islands, ms_seed;
};

// ------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------
// one block per read: 2-bit codes of both strands + the margins shipped by the host
__global__ void __launch_bounds__(256) k_encode(const char *__restrict__ seq, const dsb_read_desc *__restrict__ desc,
                                                const uint8_t *__restrict__ edge, uint8_t *__restrict__ bin,
                                                uint32_t n_reads)
{
    for (uint32_t r = blockIdx.x; r < n_reads; r += gridDim.x) {
        const dsb_read_desc d = desc[r];
        const uint32_t L = d.len;
        uint8_t *b = bin + dsb_bin_off(d.seq_off, r);
        const uint8_t *e = edge + (size_t)r * (DSB_HEAD_BYTES + DSB_TAIL_BYTES);
        uint8_t *F = b + DSB_HEAD_PAD, *R = F + L;
        if (threadIdx.x < DSB_HEAD_PAD) b[threadIdx.x] = e[threadIdx.x];
        if (threadIdx.x < DSB_TAIL_PAD) R[L + threadIdx.x] = e[DSB_HEAD_BYTES + threadIdx.x];
        const char *s = seq + d.seq_off;
        for (uint32_t p = threadIdx.x; p < L; p += blockDim.x) {
            const uint8_t c = dsb_encode_base((uint8_t)s[p]);
            F[p] = c;
            R[L - 1 - p] = (uint8_t)(3 - c);
        }
    }
}

// one block per read, one thread per (strand, k-mer position)
__global__ void __launch_bounds__(256) k_exist(const __grid_constant__ DsbIndexView ix,
                                               const dsb_read_desc *__restrict__ desc, const uint8_t *__restrict__ bin,
                                               uint8_t *__restrict__ exist, uint32_t n_reads, DsbCtrl *ctrl)
{
    uint32_t p0 = 0, p1 = 0;
    const uint32_t k = (uint32_t)ix.ek_len;
    for (uint32_t r = blockIdx.x; r < n_reads; r += gridDim.x) {
        const dsb_read_desc d = desc[r];
        const uint32_t L = d.len;
        if (L < 40) continue;                                   // never classified (src/cly.c:3084)
        const uint8_t *F = bin + dsb_bin_off(d.seq_off, r) + DSB_HEAD_PAD;
        uint8_t *ex = exist + 2 * d.seq_off;
        const uint32_t n = L - k + 1;
        for (uint32_t t = threadIdx.x; t < 2 * n; t += blockDim.x) {
            const uint32_t strand = (t >= n) ? 1u : 0u;
            const uint32_t p = t - strand * n;
            ex[strand * L + p] = dsb_exist_kmer(&ix, F + strand * L + p, &p0, &p1);
        }
    }
    // block-level reduction of the probe counters
    for (int o = 16; o > 0; o >>= 1) {
        p0 += __shfl_down_sync(0xffffffffu, p0, o);
        p1 += __shfl_down_sync(0xffffffffu, p1, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (p0) atomicAdd((unsigned long long *)&ctrl->counters.exist_probes0, (unsigned long long)p0);
        if (p1) atomicAdd((unsigned long long *)&ctrl->counters.exist_probes1, (unsigned long long)p1);
    }
}

// ------------------------------------------------------------------------------------------
// The exist probes in HASH-RANGE PASSES.  k_exist above sends every k-mer's probe wherever its hash points: two tables of
// 256 MiB .. 4 GiB, far beyond the L2, so every probe is one 128-byte DRAM fetch (B200 fetches 128 B per random access) and
// the kernel sits at the random-access roofline of the DRAM (49 G probes/s).  Random gathers from a table that FITS the L2
// run at 176 G/s (<= 64 MiB; 110 G/s at 128 MiB: profiles/l2_fit_r02.txt).  So:
//   k_exist_hash : every k-mer once, rolling (one base in, one base out): low-complexity mask, both hashes, stored 32 bits
//                  each; the exist byte becomes "pending" (plus the hash bits above 32)
//   k_exist_pass : one launch per table slice (stage 0: first table, stage 1: second table, for the k-mers whose first
//                  bit was set): streams the exist bytes, probes the k-mers whose hash falls into THIS slice -- the slice
//                  stays in the L2 for the whole launch -- and writes the outcome back into the exist byte
// After the last pass an exist byte is 0 or 1 exactly as k_exist would have written it (a probe is a pure function of the
// k-mer; the order of the probes is free).
//   exist byte while in flight: bit 7 pending, bit 6 stage, bits 5..3 / 2..0 = bits 32.. of the first / second hash
// ------------------------------------------------------------------------------------------
#define DSB_EX_PENDING 0x80u
#define DSB_EX_STAGE1 0x40u
#define DSB_EX_RUN 16            // consecutive k-mers per thread in k_exist_hash
#ifndef DSB_EX_ILP
#define DSB_EX_ILP 8             // k-mers per thread and step in k_exist_pass
#endif

// where the hashes of read r's strand live: 16-entry aligned, so that a thread's run is one 64-byte store target
static __device__ __forceinline__ size_t dsb_hash_base(uint64_t seq_off, uint32_t r, uint32_t L, uint32_t strand)
{
    const size_t b0 = (2 * (size_t)seq_off + 32 * (size_t)r + 15) & ~(size_t)15;
    return strand ? ((b0 + L + 15) & ~(size_t)15) : b0;
}

__global__ void __launch_bounds__(256) k_exist_hash(const __grid_constant__ DsbIndexView ix,
                                                    const dsb_read_desc *__restrict__ desc, const uint8_t *__restrict__ bin,
                                                    uint8_t *__restrict__ exist, uint32_t *__restrict__ h1v,
                                                    uint32_t *__restrict__ h2v, uint32_t n_reads)
{
    const int k = ix.ek_len;
    const uint64_t kmask = (k >= 32) ? ~0ull : ((1ull << (2 * k)) - 1ull);
    // counts of the four bases in four byte lanes; a lane reaches 0x80 when its count reaches single_base_max
    const uint32_t bias = (uint32_t)(128 - ix.single_base_max) * 0x01010101u;
    for (uint32_t r = blockIdx.x; r < n_reads; r += gridDim.x) {
        const dsb_read_desc d = desc[r];
        const uint32_t L = d.len;
        if (L < 40) continue;                                   // never classified (src/cly.c:3084)
        const uint8_t *F = bin + dsb_bin_off(d.seq_off, r) + DSB_HEAD_PAD;
        const uint32_t n = L - (uint32_t)k + 1;
        const uint32_t runs = (n + DSB_EX_RUN - 1) / DSB_EX_RUN;             // per strand
        for (uint32_t u = threadIdx.x; u < 2 * runs; u += blockDim.x) {
            const uint32_t strand = (u >= runs) ? 1u : 0u;
            const uint32_t base = (u - strand * runs) * DSB_EX_RUN;
            const uint32_t m = DSB_MIN((uint32_t)DSB_EX_RUN, n - base);
            const uint8_t *sq = F + strand * L + base;
            uint8_t *ex = exist + 2 * d.seq_off + strand * L + base;
            const size_t hb = dsb_hash_base(d.seq_off, r, L, strand) + base;
            uint64_t kmer = 0;
            uint32_t cnt = 0;
            for (int i = 0; i < k - 1; i++) {
                const uint32_t b = sq[i] & 3u;
                kmer = (kmer << 2) | b;
                cnt += 1u << (8 * b);
            }
            uint32_t a1[4], a2[4];
#pragma unroll 1
            for (uint32_t j0 = 0; j0 < m; j0 += 4) {
#pragma unroll
                for (uint32_t jj = 0; jj < 4; jj++) {
                    const uint32_t j = j0 + jj;
                    uint32_t e = 0, v1 = 0, v2 = 0;
                    if (j < m) {
                        const uint32_t b = sq[j + k - 1] & 3u;
                        kmer = ((kmer << 2) | b) & kmask;
                        cnt += 1u << (8 * b);
                        if ((((cnt + bias) & 0x80808080u) == 0) && kmer != 0) {
                            const uint64_t g1 = dsb_hash64_1(kmer) & ix.ek_mask, g2 = dsb_hash64_2(kmer) & ix.ek_mask;
                            v1 = (uint32_t)g1; v2 = (uint32_t)g2;
                            e = DSB_EX_PENDING | ((uint32_t)(g1 >> 32) << 3) | (uint32_t)(g2 >> 32);
                        }
                        __stcs(ex + j, (uint8_t)e);
                        cnt -= 1u << (8 * (uint32_t)((kmer >> (2 * k - 2)) & 3u));       // the base that leaves next
                    }
                    a1[jj] = v1; a2[jj] = v2;
                }
                __stcs(reinterpret_cast<uint4 *>(h1v + hb + j0), make_uint4(a1[0], a1[1], a1[2], a1[3]));
                __stcs(reinterpret_cast<uint4 *>(h2v + hb + j0), make_uint4(a2[0], a2[1], a2[2], a2[3]));
            }
        }
    }
}

// one slice of one table: stage 0 = first table (pending k-mers), stage 1 = second table (k-mers whose first bit was set)
__global__ void __launch_bounds__(256) k_exist_pass(const __grid_constant__ DsbIndexView ix,
                                                    const dsb_read_desc *__restrict__ desc, uint8_t *__restrict__ exist,
                                                    const uint32_t *__restrict__ hv, uint32_t n_reads, uint32_t stage,
                                                    uint32_t slice, uint32_t slice_shift, DsbCtrl *ctrl)
{
    uint32_t probes = 0;
    const uint32_t k = (uint32_t)ix.ek_len;
    const uint8_t *__restrict__ table = stage ? ix.ek1 : ix.ek0;
    const uint32_t want = DSB_EX_PENDING | (stage ? DSB_EX_STAGE1 : 0u);
    for (uint32_t r = blockIdx.x; r < n_reads; r += gridDim.x) {
        const dsb_read_desc d = desc[r];
        const uint32_t L = d.len;
        if (L < 40) continue;
        const uint32_t n = L - k + 1;
        uint8_t *ex0 = exist + 2 * d.seq_off;
        const size_t hb0 = dsb_hash_base(d.seq_off, r, L, 0), hb1 = dsb_hash_base(d.seq_off, r, L, 1);
        // DSB_EX_ILP k-mers per thread and step, in phases (flags, hashes, table bytes, write-back): the loads of a phase are
        // independent, so a thread keeps DSB_EX_ILP of them in flight (one byte per thread and step left the launch
        // latency-bound: 7-8 ms per pass over 2 G k-mers instead of the 1.5 ms its 10 GB of streaming take)
        for (uint32_t t0 = threadIdx.x; t0 < 2 * n; t0 += blockDim.x * DSB_EX_ILP) {
            uint8_t *ex[DSB_EX_ILP];
            uint32_t e[DSB_EX_ILP], lo[DSB_EX_ILP];
            size_t hx[DSB_EX_ILP];
            bool go[DSB_EX_ILP];
#pragma unroll
            for (int u = 0; u < DSB_EX_ILP; u++) {
                const uint32_t t = t0 + (uint32_t)u * blockDim.x;
                const uint32_t strand = (t >= n) ? 1u : 0u;
                const uint32_t p = t - strand * n;
                ex[u] = ex0 + strand * L + p;
                hx[u] = (strand ? hb1 : hb0) + p;
                e[u] = (t < 2 * n) ? (uint32_t)__ldcs(ex[u]) : 0u;
            }
#pragma unroll
            for (int u = 0; u < DSB_EX_ILP; u++) {
                go[u] = (e[u] & (DSB_EX_PENDING | DSB_EX_STAGE1)) == want;
                lo[u] = go[u] ? __ldcs(hv + hx[u]) : 0u;
            }
            uint64_t h[DSB_EX_ILP];
            uint32_t byte[DSB_EX_ILP];
#pragma unroll
            for (int u = 0; u < DSB_EX_ILP; u++) {
                const uint64_t hi = stage ? (e[u] & 7u) : ((e[u] >> 3) & 7u);
                h[u] = (hi << 32) | lo[u];
                go[u] = go[u] && (uint32_t)(h[u] >> slice_shift) == slice;
                byte[u] = go[u] ? (uint32_t)table[h[u] >> 3] : 0u;
            }
#pragma unroll
            for (int u = 0; u < DSB_EX_ILP; u++) {
                if (!go[u]) continue;
                probes++;
                const uint32_t bit = (byte[u] >> (7 - (uint32_t)(h[u] & 7))) & 1u;
                __stcs(ex[u], (uint8_t)(stage ? bit : (bit ? (e[u] | DSB_EX_STAGE1) : 0u)));
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) probes += __shfl_down_sync(0xffffffffu, probes, o);
    if ((threadIdx.x & 31) == 0 && probes)
        atomicAdd((unsigned long long *)(stage ? &ctrl->counters.exist_probes1 : &ctrl->counters.exist_probes0), (unsigned long long)probes);
}

// ------------------------------------------------------------------------------------------
// split pipeline: k_islands -> k_seed -> k_classify (seeded)
// ------------------------------------------------------------------------------------------
// Lock-step forms of the FM seeding primitives for k_seed.  Every lane of `mask` calls them together with its own
// job and an `act` flag; loops run `while (any lane still has work)` with the work predicated, so the lanes meet again
// at every iteration instead of finishing their loops one after the other (in the per-lane forms the single-row walk
// ran with ONE active thread per instruction: ncu, profiles/).  Results per lane are those of dsb_sp_set_insert /
// dsb_single_search / dsb_mem_search (dsb_core.cuh), which they restate.
// start fetching the FM block of row r (one 64-byte line) into L1 without waiting for it
static __device__ __forceinline__ void dsb_fm_prefetch(const DsbIndexView *ix, uint64_t r)
{
#if !defined(DSB_NO_FM_PREFETCH)
    asm volatile("prefetch.global.L1 [%0];" ::"l"(ix->fm + (r >> 6)));
#endif
}

static __device__ __forceinline__ int dsb_sp_set_insert_ls(uint32_t mask, bool act, uint64_t node, DsbSpSet *s)
{
    int l = 0;
    if (act) {
        if (s->l == s->m) s->l = 0;
        l = s->l;
    }
    const uint64_t *set = s->set;
    bool found = false;
    for (int i = 0; __any_sync(mask, act && !found && i < l); i += 2) {
        if (act && !found && i < l) {
            const uint64_t a = set[i];
            const uint64_t b = set[(i + 1 < l) ? i + 1 : i];
            found = (a == node) || (b == node);
        }
    }
    if (act && !found) {
        s->set[l] = node;
        s->l = l + 1;
    }
    return (act && !found) ? 1 : 0;
}

static __device__ __noinline__ void dsb_single_search_ls(DsbWork *w, uint32_t mask, bool act, uint64_t sp, const uint8_t *str,
                                                         int max_match_len, DsbSpSet *sp_set, DsbMem *m)
{
    const uint64_t *rank = w->ix->rank;
    uint64_t sa_sp = 0xffffffffffffffffull;
    int match_len = 0, sa_sp_l = 0;
    bool run = act, dup = false;
    while (__any_sync(mask, run)) {
        bool step = false;
        uint64_t new_sp = 0;
        if (run) {
            if (match_len >= max_match_len) run = false;
            else {
                if ((sp & DSB_SA_MASK) == 0) { sa_sp = sp; sa_sp_l = 0; }
                else sa_sp_l--;
                uint8_t c = 0xff;
                new_sp = dsb_occ(w, sp, &c);
                new_sp += rank[c];
                if (c != *str) run = false;
                else {
                    match_len++; str--; step = true;
                    dsb_fm_prefetch(w->ix, new_sp);      // the next step's block travels while the SP_SET is scanned
                }
            }
        }
        const int ins = dsb_sp_set_insert_ls(mask, step, new_sp, sp_set);
        if (step) {
            if (!ins) { dup = true; run = false; }
            else sp = new_sp;
        }
    }
    if (act) {
        if (dup) m->match_len = -1000;
        else { m->sp = sp; m->match_len = match_len; m->sa_sp = sa_sp; m->sa_sp_l = sa_sp_l; }
    }
}

static __device__ __noinline__ int dsb_mem_search_ls(DsbWork *w, uint32_t mask, bool act, const uint8_t *str, uint64_t pre_v,
                                                     int max_rst, int l_min_mth, int l_max_mth, DsbSpSet *sp_set, DsbMem *mem)
{
    const uint64_t *rank = w->ix->rank;
    int n_rst = 0;
    uint64_t sp = 0, ep = 0, new_sp = 0, new_ep = 0;
    int match_len = DSB_L_PRE_IDX;
    bool run = act, ret0 = false;
    if (act) {
        w->st.mem_search++;
        w->st.lut++;
        sp = w->ix->lut[pre_v];
        ep = w->ix->lut[pre_v + 1];
        str -= DSB_L_PRE_IDX;
    }
    while (__any_sync(mask, run)) {
        if (run) {
            uint8_t c = *str;
            str--;
            dsb_fm_prefetch(w->ix, ep);                  // both ends of the interval in flight together
            new_sp = rank[c] + dsb_occ(w, sp, &c);
            new_ep = rank[c] + dsb_occ(w, ep, &c);
            bool brk = false;
            if (match_len >= l_min_mth - 1) {
                if (new_sp + (uint64_t)(int64_t)max_rst >= new_ep) brk = true;
                else if (match_len >= l_max_mth) { ret0 = true; brk = true; }
            }
            if (!brk && new_sp + 1 >= new_ep) brk = true;
            if (brk) run = false;
            else { match_len++; sp = new_sp; ep = new_ep; }
        }
    }
    // the rows left (at most max_rst): insert, walk on alone, keep what is long enough.  One row (new_sp + 1 ==
    // new_ep) and several rows read the same in the reference (:1419-1440)
    const bool go = act && !ret0 && new_sp < new_ep;
    const uint64_t n_rows = go ? (new_ep - new_sp) : 0;
    for (uint64_t k = 0; __any_sync(mask, k < n_rows); k++) {
        const bool a = k < n_rows;
        const uint64_t c_sp = new_sp + k;
        const int ins = dsb_sp_set_insert_ls(mask, a, c_sp, sp_set);
        const bool walk = a && ins;
        dsb_single_search_ls(w, mask, walk, c_sp, str, DSB_MAX(0, l_max_mth - match_len), sp_set, mem + n_rst);
        if (walk) {
            mem[n_rst].match_len += match_len + 1;
            if (mem[n_rst].match_len >= l_min_mth) n_rst++;
        }
    }
    return n_rst;
}

// dsb_lv_extd (dsb_core.cuh; reference lv_extd src/cly.c:505-604) in lock-step form.  On its own a lane leaves the
// 5 x 10 cell loops at a data-dependent `return`, so 32 lanes ran them one after the other (2.8 active threads per
// instruction, 39 % of k_seed's instructions).  Here all lanes of `mask` walk the cells together, a finished lane just
// stops updating; the exact-run loop of a cell is one XOR + find-first-set on the strings packed 4 bits per symbol
// (bases 0..3; the callers' fillers 0xF1..0xF3 -> 5..7; '#' -> 8; '$' -> 9; DSB_LV_NO_BYTE -> 11; bytes before ref[0]
// never match and are handled on use: a one-to-one map, which is all equality needs).
static __device__ __forceinline__ uint32_t dsb_lv_sym(uint8_t b) { return (uint32_t)(b & 3u) | ((uint32_t)(b >> 5) & 4u); }
#define DSB_LV_QOFF 3     // query index k sits in nibble k + 3 (k = -3 .. 12)
static __device__ __noinline__ int32_t dsb_lv_extd_ls(uint32_t mask, const uint8_t *ref_in, const uint8_t *query_in, int32_t len,
                                                      uint8_t query_m1)
{
    bool done = (len <= 0);                   // returns 0 whatever the bytes are
    int32_t result = 0;
    const int32_t ref_length = len, query_length = len;
    uint64_t R = 0, Q = 0xbbull | ((uint64_t)((query_m1 == DSB_LV_NO_BYTE) ? 11u : dsb_lv_sym(query_m1)) << 8);
    for (int i = 0; i < len; i++) {
        R |= (uint64_t)dsb_lv_sym(ref_in[i]) << (4 * i);
        Q |= (uint64_t)dsb_lv_sym(query_in[i]) << (4 * (i + DSB_LV_QOFF));
    }
    if (!done) {
        R |= 8ull << (4 * len);
        Q |= 9ull << (4 * (len + DSB_LV_QOFF));
    }
    int32_t mn_d[14], ed_d[14];
    int32_t *mn = mn_d + DSB_LV_ERROR + 1, *ed = ed_d + DSB_LV_ERROR + 1;
    for (int i = -DSB_LV_ERROR - 1; i <= DSB_LV_ERROR + 1; i++) {
        mn[i] = -1;
        ed[i] = (i > 0) ? i : -i;
    }
    mn[DSB_LV_ERROR + 2] = -1; ed[DSB_LV_ERROR + 2] = DSB_LV_ERROR + 2;
    int32_t best_score = query_length;
#pragma unroll 1
    for (int i = 0; i <= DSB_LV_ERROR; i++) {
        int32_t prev_mn = -1, cur_mn = i - 1, next_mn = mn[-i + 1];
        int32_t prev_ed = i + 1, cur_ed = i, next_ed = ed[-i + 1];
#pragma unroll 1
        for (int j = -i; j <= DSB_LV_ERROR; j++) {
            if (!__any_sync(mask, !done)) return result;
            if (!done) {
                int mn_v, ed_v;
                if (cur_mn + j < ref_length - 1) {
                    int m = cur_mn + 1 - cur_ed;
                    mn_v = cur_mn + 1;
                    ed_v = cur_ed + 1;
                    if (m < next_mn + 1 - next_ed) { mn_v = next_mn + 1; ed_v = next_ed + 1; m = next_mn - next_ed; }
                    if (m < prev_mn - prev_ed) { mn_v = prev_mn + 1; ed_v = prev_ed + 1; }
                } else {
                    int m = cur_mn - cur_ed;
                    mn_v = cur_mn;
                    ed_v = cur_ed + 1;
                    if (m < prev_mn - prev_ed) { mn_v = prev_mn; ed_v = prev_ed + 1; m = prev_mn - prev_ed; }
                    if (m < next_mn + 1 - next_ed) { mn_v = next_mn + 1; ed_v = next_ed + 1; }
                }
                int mn_j = DSB_MIN(mn_v, query_length);
                mn_j = DSB_MIN(mn_j, ref_length - j);
                // the exact run from (ref[mn_j + j], query[mn_j]): a sentinel ends it at the latest
                if (mn_j + j >= 0) {
                    const uint64_t x = (R >> (4 * (mn_j + j))) ^ (Q >> (4 * (mn_j + DSB_LV_QOFF)));
                    const uint64_t nz = (x | (x >> 1) | (x >> 2) | (x >> 3)) & 0x1111111111111111ull;
                    mn_j += (__ffsll((long long)nz) - 1) >> 2;
                }
                mn[j] = mn_j;
                ed[j] = ed_v;
                const uint32_t q_sym = (uint32_t)(Q >> (4 * (mn_j + DSB_LV_QOFF))) & 0xfu;
                const uint32_t r_sym = (mn_j + j >= 0) ? ((uint32_t)(R >> (4 * (mn_j + j))) & 0xfu) : 10u;
                if (q_sym == 9u || r_sym == 8u) {
                    best_score = DSB_MIN(ed_v - 1, best_score);
                    if (j <= i + 1) { result = best_score; done = true; }
                }
                prev_mn = cur_mn; cur_mn = next_mn; next_mn = mn[j + 2];
                prev_ed = cur_ed; cur_ed = next_ed; next_ed = ed[j + 2];
            }
        }
    }
    return done ? result : best_score;
}

// dsb_get_new_ed (dsb_core.cuh) in lock-step form: is_fwd is the same for every lane of a call site
static __device__ __forceinline__ void dsb_get_new_ed_ls(DsbWork *w, uint32_t mask, bool act, uint32_t *e_d, uint32_t *len_,
                                                         uint32_t *l_mem_ext, int32_t q_off, uint64_t t_off, uint32_t l_read,
                                                         const uint8_t *q_b, bool is_fwd)
{
    uint8_t q_buff[13], t_buff[13];
    const uint8_t *q = q_buff;
    uint8_t *t = t_buff;
    uint32_t len = 0, max_len = 0;
    if (act) {
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
    }
    bool run = act && len > 0 && t[0] == q[0];
    while (__any_sync(mask, run)) {
        if (run) {
            uint32_t mtc;
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
            } else
                run = false;
        }
    }
    const int32_t d = dsb_lv_extd_ls(mask, t, q, act ? (int32_t)len : 0, (act && !is_fwd) ? q[-1] : (uint8_t)DSB_LV_NO_BYTE);
    if (act) { *e_d = (uint32_t)d; *len_ = len; }
}

// dsb_map_seed (dsb_core.cuh; reference map_seed src/cly.c:701-934) in lock-step form: the `break`s of its
// do { } while (0) become an `alive` flag, its loops run while any lane of `mask` still iterates, and the Landau-
// Vishkin calls are made by all lanes together (length 0 for the lanes with nothing to extend).
static __device__ __noinline__ int32_t dsb_map_seed_ls(DsbWork *w, uint32_t mask, bool act, DsbMem *m_r, const uint8_t *q_b,
                                                       uint32_t read_L, uint8_t direction, int *smax /* 32 ints of shared memory */)
{
#if defined(DSB_SEED_DIAG)
    const long long td_in = clock64();
#endif
    const DsbIndexView *ix = w->ix;
    const int32_t *Q_MEM = ix->q_mem;
    const int32_t *Q_LV = ix->q_lv;
    uint64_t b_p = 0, sa_sp = 0;
    int32_t q_off = 0, sa_sp_l = 0;
    uint32_t l_m = 0;
    if (act) {
        b_p = m_r->sp; q_off = m_r->read_offset; l_m = (uint32_t)m_r->match_len;
        sa_sp = m_r->sa_sp; sa_sp_l = m_r->sa_sp_l;
        w->st.map_seed++;
    }
    bool have_uni = false;
    uint64_t uni = 0;
    uint32_t uni_len = 0, u_off = 0;
    uint64_t t_off = 0;
    uint32_t l_pre = 0, l_suf = 0, d_pre = 0, d_suf = 0;
    int32_t s = 0, max_s = 0;
    uint8_t q_pre[DSB_LV_L + 1], t_pre[DSB_LV_L + 1], t_suf[DSB_LV_L + 1];
    const uint8_t *q_suf = q_pre;
    for (int k = 0; k <= DSB_LV_L; k++) { q_pre[k] = 0xF1; t_pre[k] = 0xF2; t_suf[k] = 0xF3; }
    bool alive = act;
    int s_l = 0;
    if (act) {
        l_pre = (uint32_t)DSB_MIN(q_off + 1, DSB_LV_L);
        for (uint8_t k = 0; k < l_pre; k++) q_pre[k] = q_b[q_off - k];
    }
    // (1) to the left until a sampled SA row (or l_pre bases), collecting the BWT symbols
    const bool walk = act && sa_sp == 0xffffffffffffffffull;
    {
        bool run = walk;
        while (__any_sync(mask, run)) {
            if (run) {
                if ((b_p & DSB_SA_MASK) == 0) run = false;
                else {
                    uint8_t c = 0xff;
                    uint64_t new_sp = dsb_occ(w, b_p, &c);
                    new_sp += ix->rank[c];
                    if (c == 4) run = false;
                    else {
                        t_pre[s_l++] = c;
                        b_p = new_sp;
                        if (s_l >= (int)l_pre) run = false;
                    }
                }
            }
        }
    }
    if (act && (!walk || (b_p & DSB_SA_MASK) == 0)) {
        uni = dsb_get_uni(w, walk ? b_p : sa_sp, walk ? s_l : sa_sp_l, &t_off, &u_off);
        have_uni = true;
    } else if (walk)
        l_pre = (uint32_t)s_l;
    if (alive && have_uni) {
        uni_len = (uint32_t)(ix->unitig[uni] >> 32);
        if (uni_len < DSB_MIN_UNI_L) alive = false;
        else {
            l_pre = DSB_MIN(l_pre, u_off);
            dsb_get_ref(w, t_pre, t_off - 1, l_pre, false);
        }
    }
    {
        const int32_t d = dsb_lv_extd_ls(mask, t_pre, q_pre, alive ? (int32_t)l_pre : 0, DSB_LV_NO_BYTE);
        if (alive) {
            d_pre = (uint32_t)d;
            s = Q_MEM[DSB_QMEM_IX(l_m)] + Q_LV[d_pre * 20 + l_pre];
            if (s < DSB_MIN_S_1 && l_pre == DSB_LV_L && !have_uni) { s = 0; alive = false; }
        }
    }
    // (2) the rest of the way to a sampled row
    const bool walk2 = alive && !have_uni;
    {
        bool run = walk2;
        while (__any_sync(mask, run)) {
            if (run) {
                if ((b_p & DSB_SA_MASK) == 0) run = false;
                else {
                    uint8_t c = 0xff;
                    b_p = dsb_occ(w, b_p, &c);
                    b_p += ix->rank[c];
                    s_l++;
                }
            }
        }
    }
    if (walk2) {
        uni = dsb_get_uni(w, b_p, s_l, &t_off, &u_off);
        have_uni = true;
        uni_len = (uint32_t)(ix->unitig[uni] >> 32);
        if (uni_len < DSB_MIN_UNI_L) { s = 0; alive = false; }
    }
    // (3) right side
    uint32_t l_max_suf = 0;
    bool suf = false;
    if (alive) {
        const int32_t q_off_r = q_off + (int32_t)l_m + 1;
        l_max_suf = DSB_MIN(uni_len - u_off - l_m, read_L - (uint32_t)q_off_r);
        suf = (l_max_suf != 0);
        if (suf) {
            l_suf = DSB_MIN(l_max_suf, (uint32_t)DSB_LV_L);
            q_suf = q_b + q_off_r;
            dsb_get_ref(w, t_suf, t_off + l_m, l_suf, true);
        } else
            l_suf = d_suf = 0;
    }
    {
        bool run = suf && t_suf[0] == q_suf[0];
        while (__any_sync(mask, run)) {
            if (run) {
                uint32_t mtc;
                for (mtc = 0; mtc < l_suf; mtc++)
                    if (t_suf[mtc] != q_suf[mtc]) break;
                if (mtc > 0) {
                    l_m += mtc;
                    s = Q_MEM[DSB_QMEM_IX(l_m)] + Q_LV[d_pre * 20 + l_pre];
                    l_max_suf -= mtc;
                    l_suf = DSB_MIN(l_max_suf, (uint32_t)DSB_LV_L);
                    q_suf += mtc;
                    dsb_get_ref(w, t_suf, t_off + l_m, l_suf, true);
                } else
                    run = false;
            }
        }
    }
    {
        const int32_t d = dsb_lv_extd_ls(mask, t_suf, q_suf, suf ? (int32_t)l_suf : 0, suf ? q_suf[-1] : (uint8_t)DSB_LV_NO_BYTE);
        if (suf) {
            d_suf = (uint32_t)d;
            s += Q_LV[d_suf * 20 + l_suf];
        }
    }
    if (alive && s <= DSB_MIN_S_2 && l_suf == DSB_LV_L) { s = 0; alive = false; }
    if (!alive) s = 0;

#if defined(DSB_SEED_DIAG)
    const long long td1 = clock64();
#endif
    // (4) one anchor per occurrence of the unitig.  The occurrence counts differ wildly between lanes (1 .. 999), so
    // the (seed, occurrence) pairs of the whole warp are dealt out evenly: pair g goes to lane g % 32, which fetches the
    // owning lane's seed by shuffle, extends both sides and writes the anchor into the OWNER's staging area at the
    // position the owner's serial loop would have used (pairs of one owner sit in consecutive lanes in occurrence order).
    uint32_t rp_s = 0, n_rp = 0;
    bool ret50 = false;
    if (act && s > 0) {
        rp_s = (uint32_t)ix->unitig[uni];
        const uint32_t rp_e = (uint32_t)ix->unitig[uni + 1];
        w->st.unitig++;
        if ((int64_t)rp_e - (int64_t)rp_s > 50 && !((int64_t)rp_e - (int64_t)rp_s < 1000)) ret50 = true;
        else n_rp = (rp_e > rp_s) ? rp_e - rp_s : 0;
    }
    const uint32_t lane = threadIdx.x & 31;
    uint32_t incl = n_rp;
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane >= o) incl += v;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return ret50 ? 50 : 0;
    smax[lane] = 0;
    __syncwarp();
    // the owner's side of a pair, packed for the shuffles
    const uint32_t my_flags = ((l_pre < DSB_LV_L || d_pre == 0) ? 1u : 0u) | ((l_suf < DSB_LV_L || d_suf == 0) ? 2u : 0u) |
                              ((uint32_t)direction << 8);
    const uint32_t my_lens = (l_pre & 0xffu) | ((d_pre & 0xffu) << 8) | ((l_suf & 0xffu) << 16) | ((d_suf & 0xffu) << 24);
    for (uint32_t g0 = 0; g0 < total; g0 += 32) {
        const uint32_t g = g0 + lane;
        const bool a = g < total;
        uint32_t owner = 0;                                   // number of lanes whose inclusive count is <= g
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            const uint32_t v = __shfl_sync(0xffffffffu, incl, (int)(owner + step - 1));
            if (v <= g) owner += step;
        }
        owner &= 31u;
        const uint32_t o_incl = __shfl_sync(0xffffffffu, incl, (int)owner);
        const uint32_t o_n = __shfl_sync(0xffffffffu, n_rp, (int)owner);
        const uint32_t o_rp_s = __shfl_sync(0xffffffffu, rp_s, (int)owner);
        const int32_t o_q_off = __shfl_sync(0xffffffffu, q_off, (int)owner);
        const uint32_t o_l_m = __shfl_sync(0xffffffffu, l_m, (int)owner);
        const uint32_t o_u_off = __shfl_sync(0xffffffffu, u_off, (int)owner);
        const uint32_t o_flags = __shfl_sync(0xffffffffu, my_flags, (int)owner);
        const uint32_t o_lens = __shfl_sync(0xffffffffu, my_lens, (int)owner);
        const int32_t o_s = __shfl_sync(0xffffffffu, s, (int)owner);
        const uint32_t o_read_L = __shfl_sync(0xffffffffu, read_L, (int)owner);
        const uint8_t *o_q_b = (const uint8_t *)__shfl_sync(0xffffffffu, (unsigned long long)q_b, (int)owner);
        const uint32_t o_base = __shfl_sync(0xffffffffu, w->n_anchor, (int)owner);
        const uint32_t o_excl = o_incl - o_n;
        const uint32_t k = g - o_excl;
        uint64_t r_global = 0;
        uint32_t r_ref_ID = 0;
        if (a) {
            const uint64_t rpv = ix->ref_pos[o_rp_s + k];
            w->st.refpos++;
            r_global = rpv & 0xFFFFFFFFFFull;
            r_ref_ID = (uint32_t)((rpv >> 40) & 0x7FFFFF);
        }
        uint16_t am_mtch = (uint16_t)o_l_m;
        int16_t am_score = (int16_t)o_s;
        uint8_t am_left_len = (uint8_t)(o_lens & 0xffu), am_left_ED = (uint8_t)((o_lens >> 8) & 0xffu);
        uint8_t am_rigt_len = (uint8_t)((o_lens >> 16) & 0xffu), am_rigt_ED = (uint8_t)(o_lens >> 24);
        uint32_t ed_l = 0, ed_r = 0, len_l = 0, len_r = 0;
        uint32_t l_m_ext_l = 0, l_m_ext_r = 0;
        const bool do_l = a && (o_flags & 1u), do_r = a && (o_flags & 2u);
        dsb_get_new_ed_ls(w, mask, do_l, &ed_l, &len_l, &l_m_ext_l, o_q_off, r_global + o_u_off - 1, o_read_L, o_q_b, true);
        if (do_l) { am_left_len = (uint8_t)len_l; am_left_ED = (uint8_t)ed_l; }
        if (do_l || do_r) am_mtch = (uint16_t)(o_l_m + l_m_ext_l);
        dsb_get_new_ed_ls(w, mask, do_r, &ed_r, &len_r, &l_m_ext_r, o_q_off + (int32_t)o_l_m + 1, r_global + o_u_off + o_l_m,
                          o_read_L, o_q_b, false);
        if (do_r) {
            am_rigt_len = (uint8_t)len_r;
            am_rigt_ED = (uint8_t)ed_r;
            am_mtch = (uint16_t)(am_mtch + l_m_ext_r);
        }
        bool keep = a;
        if (do_l || do_r) {
            am_score = (int16_t)(Q_MEM[DSB_QMEM_IX(am_mtch)] + Q_LV[am_left_ED * 20 + am_left_len] + Q_LV[am_rigt_ED * 20 + am_rigt_len]);
            if (am_score < DSB_MIN_S_2) keep = false;
        }
        // where the anchor goes: after the owner's earlier anchors and the kept pairs of this owner in lower lanes
        const uint32_t kept = __ballot_sync(0xffffffffu, keep);
        const uint32_t o_lo = (o_excl > g0) ? (o_excl - g0) : 0u;                 // the owner's first lane of this wave
        const uint32_t below = kept & ((1u << lane) - 1u) & ~((1u << o_lo) - 1u);
        if (keep) {
            atomicMax(&smax[owner], (int)am_score);
            const uint32_t pos = o_base + (uint32_t)__popc(below);
            if (pos >= w->cap_stage) w->overflow = 1;
            else {
                DsbAnchor an;
                an.direction = (uint8_t)(o_flags >> 8);
                an.index_in_read = (uint32_t)(o_q_off + 1) - l_m_ext_l;
                const uint64_t gpos = r_global + o_u_off - l_m_ext_l;
                an.ref_ID = r_ref_ID;
                an.ref_offset = (uint32_t)(gpos - ix->ref_off[r_ref_ID]);
                an.mtch_len = am_mtch;
                an.score = am_score;
                an.pre = -1;
                an.useless = 0;
                an.duplicate = 0;
                an.pad = 0;
                w->stage[(size_t)owner * w->cap_stage + pos] = an;
            }
        }
        // every owner takes note of what was appended to its area in this wave
        {
            const uint32_t my_excl = incl - n_rp;
            const uint32_t lo = (my_excl > g0) ? (my_excl - g0) : 0u;
            const uint32_t hi = (incl > g0) ? DSB_MIN(incl - g0, 32u) : 0u;
            if (n_rp && hi > lo) {
                const uint32_t range = ((hi >= 32u) ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
                w->n_anchor = DSB_MIN(w->n_anchor + (uint32_t)__popc(kept & range), w->cap_stage);
            }
        }
        __syncwarp();
    }
    max_s = smax[lane];
#if defined(DSB_SEED_DIAG)
    __syncwarp(mask);
    w->st.cyc[7] += (unsigned long long)(clock64() - td1);
    w->st.cyc[6] += (unsigned long long)(td1 - td_in);
#endif
    return ret50 ? 50 : max_s;
}

// Split pipeline, middle kernel: the body of one persistent warp of k_seed.  Same lock-step rounds as the device
// branch of dsb_fast_seeding, but the 32 lanes pull islands from the BATCH's job list, so a warp stays full whatever
// the number of top islands of a single read is (5 kb reads have a handful, which left 3-4 lanes of 32 busy).
// A finished island's anchors move from the lane's staging area to the batch's anchor pool.
// w: bound to the warp's scratch slice; bin_base / desc: the batch.  Returns through jobs[], pool, *n_pool, w->st.
// slow_classify (dsb_core.cuh; reference src/cly.c:1545-1606) with ONE LANE PER ISLAND, 32 islands of the strand at a time, the
// lanes in lock step through the same primitives as k_seed.  An island's work (its k-mer loop of bwt_MEM_search calls with up
// to 8 rows each, the sort of its MEMs, map_seed on the best 8, the useless marks) touches nothing outside the island: SP_SET
// restarts empty and the anchors are only appended -- so the islands of a wave run side by side on private SP_SET / MEM list /
// anchor staging and their anchors are appended in island order afterwards, which is the serial result.  The warp-serial form
// spent ~11 ms on a 10 kb read without a usable chain (600 islands, every search a chain of dependent rank queries with 32
// lanes doing the same thing): 11 % of k_classify for 5 % of the reads, and the kernel's tail.  Private MEM lists live in the
// match-block arena (not in use before scoring); a lane that would run out of MEM slots or of anchor staging makes the whole
// strand fall back to the serial form (returns false, nothing changed).
__device__ __noinline__ bool dsb_slow_seeding_lanes(DsbWork *w, int strand)
{
    if ((size_t)w->cap_sms * sizeof(DsbSms) < (size_t)32 * DSB_SLOW_LANE_MEMS * sizeof(DsbMem) || !w->smax) return false;
    const uint32_t lane = threadIdx.x & 31;
    const int l_ek = w->ix->ek_len;
    const uint8_t *bin_read = w->bin + (strand ? w->L : 0);
    const uint8_t direction = strand ? DSB_REVERSE : DSB_FORWARD;
    const DsbSeed *sv = w->seeds[strand];
    const uint32_t nsv = w->n_seeds[strand];
    if (nsv == 0) return true;
    const uint32_t first_top = sv[0].top;
    const int min_match_len = DSB_MIN(20 - 1, l_ek + 1);
    DsbAnchor *const main_anchors = w->anchors;
    const uint32_t main_n = w->n_anchor, main_cap = w->cap_anchor;
    const DsbStats snap = w->st;
    __syncwarp();
    uint8_t *priv = w->lane_priv + (size_t)lane * DSB_LANE_PRIV_BYTES;
    DsbSpSet sp_set = {(uint64_t *)priv, 0, DSB_SP_SET_MAX, 1};
    DsbMem *m_r = (DsbMem *)w->sms + (size_t)lane * DSB_SLOW_LANE_MEMS;
    DsbAnchor *const my_stage = w->stage + (size_t)lane * w->cap_stage;
    const int lane_mems = w->slow_lane_mems ? (int)w->slow_lane_mems : DSB_SLOW_LANE_MEMS;
    uint32_t merged = main_n;                   // anchors of the finished waves, in the main array
    bool bail = false;
    uint32_t cur = 0;
    while (cur < nsv && !bail) {
        // the next (up to) 32 islands that pass the filter (it looks at the FIRST seed's top), in order
        int mine = -1;
        uint32_t filled = 0;
        while (filled < 32 && cur < nsv) {
            const uint32_t c = cur + lane;
            // an island of one k-mer has no k-mer loop (j = len - 1 down to 1): no search, no MEM, no anchor -- it need not take a lane
            const bool q = c < nsv && sv[c].len >= 2 && !((int)sv[c].len < 3 && first_top == 0);
            const uint32_t m = __ballot_sync(0xffffffffu, q);
            const uint32_t cnt = (uint32_t)__popc(m), take = DSB_MIN(cnt, 32u - filled);
            const int t = (int)lane - (int)filled;
            if (t >= 0 && (uint32_t)t < take) mine = (int)(cur + __fns(m, 0, t + 1));
            if (take < cnt) { cur += __fns(m, 0, (int)take) + 1; filled = 32; }
            else { cur += 32; filled += take; }
        }
        if (cur > nsv) cur = nsv;
        // the wave: every lane its island, private state
        w->anchors = my_stage;
        w->cap_anchor = w->cap_stage;
        w->n_anchor = 0;
        w->solo = 1;
        bool have = mine >= 0;
        DsbSeed sd = {0, 0, 0};
        if (have) sd = sv[mine];
        int j = (int)sd.len - 1, n_mem = 0, t_map = -1;          // t_map >= 0: the island is in its map_seed phase
        sp_set.l = 0;
        while (__any_sync(0xffffffffu, have)) {
            bool search = false;
            int k_idx = 0, s_idx = 0;
            uint64_t pre_v = 0;
            if (have && t_map < 0) {
                if (j < 1) {
                    if (n_mem > 1) dsb_sort_mems(m_r, n_mem);
                    t_map = 0;
                } else if (n_mem + 8 > lane_mems) {
                    bail = true; have = false;
                } else {
                    k_idx = (int)sd.offset + j;
                    s_idx = k_idx + l_ek - 1;
                    pre_v = dsb_prefix13(bin_read + s_idx);
                    search = true;
                }
            }
            if (__any_sync(0xffffffffu, search)) {
                const int got = dsb_mem_search_ls(w, 0xffffffffu, search, bin_read + s_idx, pre_v, 8, min_match_len, s_idx, &sp_set,
                                                  m_r + n_mem);
                if (search) {
                    for (int t = n_mem; t < n_mem + got; t++) m_r[t].read_offset = k_idx + l_ek - 1 - m_r[t].match_len;
                    n_mem += got;
                    j -= 2;
                }
                __syncwarp();
            }
            const bool mp = have && t_map >= 0 && t_map < DSB_MIN(n_mem, 8);
            if (__any_sync(0xffffffffu, mp)) {
                dsb_map_seed_ls(w, 0xffffffffu, mp, m_r + (mp ? t_map : 0), bin_read, w->L, direction, w->smax);
                if (mp) t_map++;
                __syncwarp();
            }
            if (have && t_map >= 0 && t_map >= DSB_MIN(n_mem, 8)) {
                dsb_mark_useless(w, 0, w->n_anchor);
                have = false;
            }
        }
        w->solo = 0;
        const uint32_t my_cnt = w->n_anchor;
        // a lane's staging (an eighth of the read's anchor arena) ran full: the serial form has the whole arena
        bail = __any_sync(0xffffffffu, bail || w->overflow != 0);
        w->overflow = 0;
        w->anchors = main_anchors;
        w->cap_anchor = main_cap;
        if (bail) break;
        // append the wave's anchors in island (= lane) order
        uint32_t pre = my_cnt;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, pre, o);
            if ((int)lane >= o) pre += v;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
        if (merged + total > main_cap) { w->overflow = 1; w->n_anchor = main_n; break; }
        for (uint32_t k = 0; k < my_cnt; k++) main_anchors[merged + (pre - my_cnt) + k] = my_stage[k];
        merged += total;
        __syncwarp();
    }
    w->anchors = main_anchors;
    w->cap_anchor = main_cap;
    w->solo = 0;
    __syncwarp();
    if (bail) { w->n_anchor = main_n; w->st = snap; return false; }
#define DSB_FOLD(f) w->st.f = snap.f + __reduce_add_sync(0xffffffffu, (unsigned)(w->st.f - snap.f))
    DSB_FOLD(occ); DSB_FOLD(lut); DSB_FOLD(sa); DSB_FOLD(unitig); DSB_FOLD(refpos); DSB_FOLD(ref_bytes);
    DSB_FOLD(mem_search); DSB_FOLD(map_seed);
#undef DSB_FOLD
    if (!w->overflow) w->n_anchor = merged;
    return true;
}

static __device__ __noinline__ void dsb_seed_jobs(DsbWork *w, const dsb_read_desc *desc, uint8_t *bin_base, DsbSeedJob *jobs, uint32_t n_jobs,
                           uint32_t *next_job, DsbAnchor *pool, uint32_t cap_pool, uint32_t *n_pool, int *smax)
{
    const uint32_t lane = threadIdx.x & 31;
    uint8_t *priv = w->lane_priv + (size_t)lane * DSB_LANE_PRIV_BYTES;
    DsbSpSet sp_set = {(uint64_t *)priv, 0, DSB_SP_SET_MAX, 1};
    DsbMem *m_r = (DsbMem *)(priv + 512 * 8);
    w->anchors = w->stage + (size_t)lane * w->cap_stage;
    w->cap_anchor = w->cap_stage;
    w->n_anchor = 0;
    w->solo = 1;
    const int l_ek = w->ix->ek_len;
    const int min_index = 21 - l_ek;
    bool have = false, dry = false;
    uint32_t jb = 0, j_strand = 0, skip_next = 0;
    int j = 0;
    DsbSeed sd = {0, 0, 0};
    for (;;) {
        const uint32_t need = __ballot_sync(0xffffffffu, !have && !dry);
        if (need) {
            const int leader = __ffs((int)need) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(next_job, (uint32_t)__popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!have && !dry) {
                jb = base + (uint32_t)__popc(need & ((1u << lane) - 1u));
                if (jb < n_jobs) {
                    const DsbSeedJob job = jobs[jb];
                    const dsb_read_desc d = desc[job.read];
                    w->bin = bin_base + dsb_bin_off(d.seq_off, job.read) + DSB_HEAD_PAD;
                    w->L = d.len;
                    j_strand = job.strand_ci >> 31;
                    sd.offset = job.offset; sd.len = job.len;
                    j = (int)sd.len - 1;
                    w->n_anchor = 0;
                    skip_next = 0;
                    sp_set.l = 0;
                    have = true;
                } else
                    dry = true;
            }
        }
        if (!__any_sync(0xffffffffu, have)) break;
        bool finish = false;
        int n_mem = 0, string_index = 0;
        const uint8_t *bin_read = w->bin + (j_strand ? w->L : 0);
        bool search = false;
        uint64_t pre_v = 0;
        if (have) {
            if (j < min_index) finish = true;
            else {
                string_index = (int)sd.offset + j + l_ek - 1;
                pre_v = dsb_prefix13(bin_read + string_index);
                search = true;
            }
        }
#if defined(DSB_SEED_DIAG)
        const long long td0 = clock64();
#endif
        n_mem = dsb_mem_search_ls(w, 0xffffffffu, search, bin_read + string_index, pre_v, 2, 21 - 1, string_index, &sp_set, m_r);
        __syncwarp();
#if defined(DSB_SEED_DIAG)
        w->st.cyc[5] += (unsigned long long)(clock64() - td0);
#endif
        int max_score = 0;
        for (int mi = 0; mi < 2; mi++) {
            const bool ms = have && mi < n_mem;
            if (ms) m_r[mi].read_offset = string_index - m_r[mi].match_len;
            const int c_score = dsb_map_seed_ls(w, 0xffffffffu, ms, m_r + mi, bin_read, w->L, j_strand ? DSB_REVERSE : DSB_FORWARD, smax);
            if (ms) max_score = DSB_MAX(c_score, max_score);
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
            dsb_mark_useless(w, 0, w->n_anchor);
            const uint32_t cnt = w->n_anchor;
            uint32_t off = 0;
            if (cnt) {
                off = atomicAdd(n_pool, cnt);
                if (off + cnt > cap_pool) w->overflow |= 8u;
                else
                    for (uint32_t k = 0; k < cnt; k++) pool[off + k] = w->anchors[k];
            }
            // pool full: the batch is re-run with a larger pool; until then nobody may follow a_off past the pool's end
            const bool lost = (w->overflow & 8u) != 0;
            jobs[jb].a_off = lost ? 0u : off;
            jobs[jb].count = lost ? 0u : cnt;
            jobs[jb].skip = skip_next;
            have = false;
        }
    }
    w->solo = 0;
    __syncwarp();
}

#ifndef DSB_MIN_BLOCKS
#define DSB_MIN_BLOCKS 6          // 80 registers -> 24 resident warps per SM.  Same-box A/Bs (profiles/): as ONE 10 000-read launch 64 registers / 32 warps is 1.7 % faster (classify stage 10.08 vs 10.26 ms), but behind read_classify(), where three to five smaller batches overlap and a batch is about one read per warp, the faster 80-register code wins: 18.6-18.9 -> 17.8-18.2 ms per 50 Mbase call, 65.4-66.0 -> 61.9-62.2 ms per 198 Mbase call (56 registers / 7 blocks in between, 96 / 5 blocks slower in both)
#endif

// warp per read: the exist-k-mer islands of both strands, then the read's top islands as jobs of the batch-wide list
__global__ void __launch_bounds__(DSB_CLASSIFY_THREADS, DSB_MIN_BLOCKS) k_islands(
    const __grid_constant__ DsbIndexView ix, const dsb_read_desc *__restrict__ desc, const uint32_t *__restrict__ order,
    uint8_t *__restrict__ bin, const uint8_t *__restrict__ exist, uint8_t *__restrict__ scratch,
    const __grid_constant__ DsbScratchDims dims, DsbReadSeeds *__restrict__ rinfo, DsbSeedJob *__restrict__ jobs,
    uint32_t cap_jobs, uint32_t n_reads, DsbCtrl *ctrl)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    DsbWork w;
    memset(&w, 0, sizeof w);
    w.ix = &ix;
    dsb_bind_scratch(&w, scratch + (size_t)warp * dims.bytes, &dims);
    unsigned long long cyc = 0;
    for (;;) {
        uint32_t slot = 0;
        if (lane == 0) slot = atomicAdd(&ctrl->next_island_read, 1u);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= n_reads) break;
        const uint32_t r = order[slot];
        const dsb_read_desc d = desc[r];
        DsbReadSeeds rs = {0, 0, {0, 0}};
        if (d.len >= 40) {
            const long long t0 = clock64();
            w.bin = bin + dsb_bin_off(d.seq_off, r) + DSB_HEAD_PAD;
            w.L = d.len;
            w.exist[0] = exist ? exist + 2 * d.seq_off : nullptr;     // no exist bytes: the scan probes the tables itself
            w.exist[1] = exist ? w.exist[0] + d.len : nullptr;
            dsb_islands_both(&w);
            rs.total_score[0] = w.total_score[0];
            rs.total_score[1] = w.total_score[1];
            int dir0 = 0, dir1 = 1;
            if (w.total_score[0] < w.total_score[1]) { dir0 = 1; dir1 = 0; }
            const uint32_t s0 = w.total_score[dir0], s1 = w.total_score[dir1];
            const bool both_direction = ((s0 - s1) <= (s0 >> 3));
            // count the top islands, reserve their jobs, write them in the serial order
            uint32_t cnt = 0;
            for (int pass = 0; pass < (both_direction ? 2 : 1); pass++) {
                const int strand = pass ? dir1 : dir0;
                const DsbSeed *sv = w.seeds[strand];
                const uint32_t nsv = w.n_seeds[strand];
                for (uint32_t c0 = 0; c0 < nsv; c0 += 32) {
                    const uint32_t ci = c0 + lane;
                    cnt += (uint32_t)__popc(__ballot_sync(0xffffffffu, ci < nsv && sv[ci].top != 0));
                }
            }
            uint32_t base = 0;
            if (lane == 0 && cnt) base = atomicAdd(&ctrl->n_jobs, cnt);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + cnt > cap_jobs) {
                if (lane == 0) atomicOr(&ctrl->overflow, 8u);
            } else {
                uint32_t k = base;
                for (int pass = 0; pass < (both_direction ? 2 : 1); pass++) {
                    const int strand = pass ? dir1 : dir0;
                    const DsbSeed *sv = w.seeds[strand];
                    const uint32_t nsv = w.n_seeds[strand];
                    for (uint32_t c0 = 0; c0 < nsv; c0 += 32) {
                        const uint32_t ci = c0 + lane;
                        const bool top = ci < nsv && sv[ci].top != 0;
                        const uint32_t m = __ballot_sync(0xffffffffu, top);
                        if (top) {
                            DsbSeedJob jb;
                            jb.read = r; jb.strand_ci = ((uint32_t)strand << 31) | ci;
                            jb.offset = sv[ci].offset; jb.len = sv[ci].len;
                            jb.a_off = 0; jb.count = 0; jb.skip = 0; jb.pad = 0;
                            jobs[k + (uint32_t)__popc(m & ((1u << lane) - 1u))] = jb;
                        }
                        k += (uint32_t)__popc(m);
                    }
                }
                rs.job_base = base;
                rs.n_jobs = cnt;
            }
            cyc += (unsigned long long)(clock64() - t0);
        }
        if (lane == 0) rinfo[r] = rs;
        __syncwarp();
    }
    if (lane == 0 && cyc) atomicAdd((unsigned long long *)&ctrl->counters.phase_cycles[0], cyc);
    if (lane == 0 && w.st.ex0) {
        atomicAdd((unsigned long long *)&ctrl->counters.exist_probes0, (unsigned long long)w.st.ex0);
        atomicAdd((unsigned long long *)&ctrl->counters.exist_probes1, (unsigned long long)w.st.ex1);
    }
}

// persistent warps, ONE LANE PER ISLAND of the whole batch
#ifndef DSB_SEED_MIN_BLOCKS
#define DSB_SEED_MIN_BLOCKS 8        // 64 registers, 32 warps per SM: 2.66 -> 2.49 ms (10 blocks / 48 registers: 2.59)
#endif
__global__ void __launch_bounds__(DSB_CLASSIFY_THREADS, DSB_SEED_MIN_BLOCKS) k_seed(
    const __grid_constant__ DsbIndexView ix, const dsb_read_desc *__restrict__ desc, uint8_t *__restrict__ bin,
    uint8_t *__restrict__ scratch, const __grid_constant__ DsbScratchDims dims, DsbSeedJob *__restrict__ jobs,
    DsbAnchor *__restrict__ pool, uint32_t cap_pool, DsbCtrl *ctrl)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (ctrl->overflow & 8u) return;            // k_islands ran out of job slots: n_jobs counts jobs that were never written
    const uint32_t n_jobs = ctrl->n_jobs;
    DsbWork w;
    memset(&w, 0, sizeof w);
    w.ix = &ix;
    dsb_bind_scratch(&w, scratch + (size_t)warp * dims.bytes, &dims);
    const long long t0 = clock64();
    __shared__ int s_seed_max[DSB_CLASSIFY_THREADS / 32][32];
    dsb_seed_jobs(&w, desc, bin, jobs, n_jobs, &ctrl->next_job, pool, cap_pool, &ctrl->n_pool, s_seed_max[threadIdx.x >> 5]);
    const unsigned long long cyc = (unsigned long long)(clock64() - t0);
    // every lane counted its own work
    if (__any_sync(0xffffffffu, w.overflow != 0)) {
        const uint32_t of = __reduce_or_sync(0xffffffffu, w.overflow);
        if (lane == 0) atomicOr(&ctrl->overflow, of);
    }
    dsb_counters *cn = &ctrl->counters;
#define DSB_FOLD_ADD(field, val) do { const unsigned long long v_ = (val); unsigned long long s_ = v_; \
        for (int o_ = 16; o_ > 0; o_ >>= 1) s_ += __shfl_xor_sync(0xffffffffu, s_, o_); \
        if (lane == 0 && s_) atomicAdd((unsigned long long *)&cn->field, s_); } while (0)
    DSB_FOLD_ADD(occ_queries, w.st.occ); DSB_FOLD_ADD(lut_lookups, w.st.lut); DSB_FOLD_ADD(sa_lookups, w.st.sa);
    DSB_FOLD_ADD(unitig_reads, w.st.unitig); DSB_FOLD_ADD(refpos_reads, w.st.refpos);
    DSB_FOLD_ADD(ref_window_bytes, w.st.ref_bytes); DSB_FOLD_ADD(mem_searches, w.st.mem_search);
    DSB_FOLD_ADD(map_seeds, w.st.map_seed);
    DSB_FOLD_ADD(seed_occ_queries, w.st.occ);
    DSB_FOLD_ADD(seed_alg_bytes, (unsigned long long)w.st.occ * 64ull + (unsigned long long)w.st.lut * 16ull +
                                 ((unsigned long long)w.st.sa + w.st.unitig + w.st.refpos) * 8ull + w.st.ref_bytes);
#undef DSB_FOLD_ADD
    if (lane == 0) atomicAdd((unsigned long long *)&cn->phase_cycles[1], cyc);
#if defined(DSB_SEED_DIAG)
    if (lane == 0) for (int ph = 5; ph < 8; ph++) atomicAdd((unsigned long long *)&cn->phase_cycles[ph], w.st.cyc[ph]);
#endif
}

// persistent: every warp pulls the next read of the (longest-first) order until none is left
#ifndef DSB_MIN_BLOCKS
#define DSB_MIN_BLOCKS 6          // 80 registers -> 24 resident warps per SM.  Same-box A/Bs (profiles/): as ONE 10 000-read launch 64 registers / 32 warps is 1.7 % faster (classify stage 10.08 vs 10.26 ms), but behind read_classify(), where three to five smaller batches overlap and a batch is about one read per warp, the faster 80-register code wins: 18.6-18.9 -> 17.8-18.2 ms per 50 Mbase call, 65.4-66.0 -> 61.9-62.2 ms per 198 Mbase call (56 registers / 7 blocks in between, 96 / 5 blocks slower in both)
#endif
__global__ void __launch_bounds__(DSB_CLASSIFY_THREADS, DSB_MIN_BLOCKS) k_classify(
    const __grid_constant__ DsbIndexView ix, const dsb_read_desc *__restrict__ desc, const uint32_t *__restrict__ order,
    uint8_t *__restrict__ bin, const uint8_t *__restrict__ exist, uint8_t *__restrict__ scratch,
    const __grid_constant__ DsbScratchDims dims, dsb_read_result *__restrict__ results, dsb_hit *__restrict__ hits,
    uint32_t cap_hits, uint32_t n_reads, DsbCtrl *ctrl,
    const DsbReadSeeds *__restrict__ rinfo, const DsbSeedJob *__restrict__ jobs, const DsbAnchor *__restrict__ pool)
{
    __shared__ uint32_t s_job_ctr[DSB_CLASSIFY_THREADS / 32];
    __shared__ int s_seed_max[DSB_CLASSIFY_THREADS / 32][32];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (rinfo && (ctrl->overflow & 8u)) return; // job list or anchor pool overflowed upstream: the host regrows and re-runs the batch
    DsbWork w;
    memset(&w, 0, sizeof w);
    w.ix = &ix;
    dsb_bind_scratch(&w, scratch + (size_t)warp * dims.bytes, &dims);
    w.warp_ctr = &s_job_ctr[threadIdx.x >> 5];
    w.smax = s_seed_max[threadIdx.x >> 5];
    __syncwarp();
    for (;;) {
        uint32_t slot = 0;
        if (lane == 0) slot = atomicAdd(&ctrl->next_read, 1u);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= n_reads) break;
        const uint32_t r = order[slot];
        const dsb_read_desc d = desc[r];
        w.bin = bin + dsb_bin_off(d.seq_off, r) + DSB_HEAD_PAD;
        w.L = d.len;
        w.is_2g = (d.flags & DSB_READ_2G) ? 1u : 0u;
        w.exist[0] = exist ? exist + 2 * d.seq_off : nullptr;
        w.exist[1] = exist ? w.exist[0] + d.len : nullptr;
        w.overflow = 0;
        memset(&w.st, 0, sizeof w.st);
#if defined(DSB_READ_CYCLES)
        unsigned long long t_begin;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
#endif
        if (rinfo) {                                   // split pipeline: islands and fast seeding were done by k_islands / k_seed
            const DsbReadSeeds rs = rinfo[r];
            dsb_classify_seeded(&w, &rs, jobs, pool);
        } else
            dsb_classify_read(&w);
        __syncwarp();
        uint32_t hit_off = 0;
        if (lane == 0) {
            if (w.overflow) atomicOr(&ctrl->overflow, w.overflow);
            else if (w.n_chain) hit_off = atomicAdd(&ctrl->n_hits, w.n_chain);
        }
        hit_off = __shfl_sync(0xffffffffu, hit_off, 0);
        if (w.overflow) continue;
        if (hit_off + w.n_chain > cap_hits) {
            if (lane == 0) atomicOr(&ctrl->overflow, 4u);
            continue;
        }
        for (uint32_t c = lane; c < w.n_chain; c += 32) {
            const DsbChain &ch = w.chains[c];
            dsb_hit h;
            h.ref_ID = ch.ref_ID; h.t_st = ch.t_st; h.t_ed = ch.t_ed; h.q_st = ch.q_st; h.q_ed = ch.q_ed;
            h.sum_score = ch.sum_score; h.indel = ch.indel; h.anchor_number = ch.anchor_number;
            h.direction = ch.direction; h.primary = ch.primary; h.pri_index = ch.pri_index;
            h.with_top_anchor = ch.with_top;
            hits[hit_off + c] = h;
        }
        if (lane == 0) {
            dsb_read_result rr;
            rr.hit_off = hit_off; rr.n_hit = w.n_chain; rr.fast_classify = (w.fast_classify ? DSB_RES_FAST : 0u) | (w.had_chain ? DSB_RES_HAD_CHAIN : 0u); rr.n_anchor = w.n_anchor;
#if defined(DSB_READ_CYCLES)
            // profiling build (tools/read_times.py): when this read started and how long it took, in microseconds
            unsigned long long t_end;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
            rr.hit_off = (uint32_t)(t_begin / 1000);
            rr.n_anchor = (uint32_t)((t_end - t_begin) / 1000);
#endif
            results[r] = rr;
            dsb_counters *cn = &ctrl->counters;
#define DSB_ADD(field, val) if (val) atomicAdd((unsigned long long *)&cn->field, (unsigned long long)(val))
            DSB_ADD(occ_queries, w.st.occ); DSB_ADD(lut_lookups, w.st.lut); DSB_ADD(sa_lookups, w.st.sa);
            DSB_ADD(unitig_reads, w.st.unitig); DSB_ADD(refpos_reads, w.st.refpos);
            DSB_ADD(ref_window_bytes, w.st.ref_bytes); DSB_ADD(mem_searches, w.st.mem_search);
            DSB_ADD(map_seeds, w.st.map_seed); DSB_ADD(sdp_windows, w.st.sdp_win);
            DSB_ADD(anchors, w.n_anchor); DSB_ADD(slow_reads, w.st.slow); DSB_ADD(slow_serial_strands, w.st.slow_serial);
            DSB_ADD(exist_probes0, w.st.ex0); DSB_ADD(exist_probes1, w.st.ex1);
            for (int ph = 0; ph < 8; ph++) DSB_ADD(phase_cycles[ph], w.st.cyc[ph]);
#undef DSB_ADD
        }
        __syncwarp();
    }
}


// bulk primitives (parity known-answer entry + roofline probes)
// Four independent queries per thread per step: all their block loads are issued before the first
// popcount, so a thread keeps 4 x 64 bytes of random HBM traffic in flight.
#define DSB_FM_ILP 4
__global__ void __launch_bounds__(256) k_fm_rank(const __grid_constant__ DsbIndexView ix, const uint64_t *__restrict__ row,
                                                 const uint8_t *__restrict__ c_in, uint64_t n, uint64_t *__restrict__ out,
                                                 uint8_t *__restrict__ c_out)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i0 = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i0 < n; i0 += stride * DSB_FM_ILP) {
        uint64_t r[DSB_FM_ILP], o[DSB_FM_ILP];
        uint8_t c[DSB_FM_ILP];
#pragma unroll
        for (int u = 0; u < DSB_FM_ILP; u++) {
            const uint64_t i = i0 + (uint64_t)u * stride;
            r[u] = (i < n) ? row[i] : 0;
            c[u] = (i < n) ? c_in[i] : (uint8_t)0;
        }
#pragma unroll
        for (int u = 0; u < DSB_FM_ILP; u++) o[u] = dsb_occ_ix(&ix, r[u], &c[u]);
#pragma unroll
        for (int u = 0; u < DSB_FM_ILP; u++) {
            const uint64_t i = i0 + (uint64_t)u * stride;
            if (i < n) { out[i] = o[u]; c_out[i] = c[u]; }
        }
    }
}

// Random-access roofline probe: every thread gathers `chunk`-byte aligned pieces (32, 64 or 128 bytes) from
// pseudo-random places of a table far larger than L2, DSB_FM_ILP of them in flight, and folds them into one
// word.  Its GB/s is the ceiling a rank-query kernel (one 64-byte block per query) can approach.
template <int CHUNK>
__global__ void __launch_bounds__(256) k_random_gather(const uint4 *__restrict__ table, uint64_t n_chunks, uint64_t n_access,
                                                       uint32_t *__restrict__ sink)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    for (uint64_t i0 = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i0 < n_access; i0 += stride * DSB_FM_ILP) {
        uint4 v[DSB_FM_ILP][CHUNK / 16];
#pragma unroll
        for (int u = 0; u < DSB_FM_ILP; u++) {
            uint64_t x = (i0 + (uint64_t)u * stride) * 0x9E3779B97F4A7C15ull;       // splitmix-style scramble
            x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
            const uint4 *p = table + (x % n_chunks) * (CHUNK / 16);
#pragma unroll
            for (int k = 0; k < CHUNK / 16; k++) v[u][k] = __ldg(p + k);
        }
#pragma unroll
        for (int u = 0; u < DSB_FM_ILP; u++)
#pragma unroll
            for (int k = 0; k < CHUNK / 16; k++) acc ^= v[u][k].x ^ v[u][k].y ^ v[u][k].z ^ v[u][k].w;
    }
    if (acc == 0x12345679u) sink[0] = acc;                  // keeps the loads alive
}

__global__ void __launch_bounds__(256) k_exist_flat(const __grid_constant__ DsbIndexView ix, const uint8_t *__restrict__ codes,
                                                    uint64_t n_kmer, uint8_t *__restrict__ exist)
{
    uint32_t p0 = 0, p1 = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_kmer; i += (uint64_t)gridDim.x * blockDim.x)
        exist[i] = dsb_exist_kmer(&ix, codes + i, &p0, &p1);
}

// ------------------------------------------------------------------------------------------
// C-ABI: index
// ------------------------------------------------------------------------------------------

// ---------------------------------------------------------------------------------------
// FM re-blocking on the device (the index upload's only transform): one thread per 64-symbol block.
// Input = the reference's occ array as it is on disk: blocks of 168 B = 5 x u64 counts of A,C,G,T,# before
// symbol 256*b, then 256 symbols as nibbles, low nibble first (reference src/bwt.c:32-41,52,214-220).
// Output = DsbFmBlock (dsb_layout.cuh): cumulative A/C/G/T counts + four bit planes.  Device twin of the
// host function dsb_fm_reblock (kept for the test-only host simulator and as the checker of this kernel).
// ---------------------------------------------------------------------------------------
static __device__ __forceinline__ uint64_t dsb_gather_every_4th_bit(uint64_t y)   // bits 4k of y -> bits k
{
    y = (y | (y >> 3)) & 0x0303030303030303ull;
    y = (y | (y >> 6)) & 0x000f000f000f000full;
    y = (y | (y >> 12)) & 0x000000ff000000ffull;
    y = (y | (y >> 24)) & 0xffffull;
    return y;
}

static __device__ __forceinline__ void dsb_count_nibbles(uint64_t x, uint64_t cnt[4])
{
#pragma unroll
    for (int c = 0; c < 4; c++) {
        uint64_t z = x ^ (0x1111111111111111ull * (uint64_t)c);
        z |= z >> 1;
        z |= z >> 2;                                     // bit 0 of every nibble = the nibble differs from c
        cnt[c] += 16u - (uint32_t)__popcll(z & 0x1111111111111111ull);
    }
}

__global__ void __launch_bounds__(256) k_fm_reblock(const uint64_t *__restrict__ occ, uint64_t nb, DsbFmBlock *__restrict__ out,
                                                    unsigned long long *__restrict__ dollar_row)
{
    const uint64_t t = (uint64_t)blockIdx.x * 256u + threadIdx.x;
    if (t >= nb * 4) return;
    const uint64_t b = t >> 2;
    const int s = (int)(t & 3);
    const uint64_t *blk = occ + b * 21;                  // 168 B = 21 words
    uint64_t cnt[4] = {blk[0], blk[1], blk[2], blk[3]};
    const uint64_t *data = blk + 5;                      // 16 words x 16 nibbles
    for (int w = 0; w < 4 * s; w++) dsb_count_nibbles(data[w], cnt);
    DsbFmBlock o;
#pragma unroll
    for (int c = 0; c < 4; c++) { o.cnt[c] = cnt[c]; o.plane[c] = 0; }
#pragma unroll
    for (int w = 0; w < 4; w++) {
        const uint64_t x = data[4 * s + w];
#pragma unroll
        for (int pl = 0; pl < 4; pl++)
            o.plane[pl] |= dsb_gather_every_4th_bit((x >> pl) & 0x1111111111111111ull) << (16 * w);
    }
    out[t] = o;
    {   // the one row that holds '$' (symbol 5 = planes 0 and 2 set, 1 and 3 clear)
        const uint64_t is5 = o.plane[0] & ~o.plane[1] & o.plane[2] & ~o.plane[3];
        if (is5) atomicMin(dollar_row, (unsigned long long)(t * 64 + (uint64_t)(__ffsll((long long)is5) - 1)));
    }
    if (t == nb * 4 - 1) {                               // terminal block: counts after the last symbol, so that r == n works
        for (int w = 0; w < 4; w++) dsb_count_nibbles(data[4 * s + w], cnt);
        DsbFmBlock e;
        for (int c = 0; c < 4; c++) { e.cnt[c] = cnt[c]; e.plane[c] = 0; }
        out[nb * 4] = e;
    }
}

// L2 fetch granularity of the device (a hint): the path's index accesses are single 32-byte sectors (exist probes) and
// 64-byte FM blocks at random places.  DSB_L2_FETCH=32|64|128 asks for a value (A/B runs); DSB_VERBOSE prints what the device
// reports.  Measured on B200 (profiles/l2_fetch_r02.txt): no setting changes the time or the DRAM bytes of the probes.
static void dsb_apply_l2_fetch_env(void)
{
    const char *e = getenv("DSB_L2_FETCH");
    size_t before = 0, got = 0;
    cudaDeviceGetLimit(&before, cudaLimitMaxL2FetchGranularity);
    cudaError_t rc = cudaSuccess;
    if (e && atoi(e) > 0) rc = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(e));
    cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
    if (getenv("DSB_VERBOSE"))
        fprintf(stderr, "[dsb] L2 fetch granularity: device default %zu, asked %s (%s), now %zu\n", before, e ? e : "-",
                cudaGetErrorString(rc), got);
    cudaGetLastError();
}

// Host -> device copy of an index array through two pinned staging buffers: the memcpy out of the (mapped, pageable) source
// into one buffer overlaps the DMA of the other.  A plain cudaMemcpy from pageable memory does the same inside the driver
// with one small buffer and no overlap (index load of the 1 Gbp index: 2.7 s -> see profiles/).
struct DsbStager {
    uint8_t *buf[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaStream_t st = nullptr;
    size_t cap = (size_t)32 << 20;
    int n = 0;
    bool ok = false;
    DsbStager()
    {
        ok = cudaHostAlloc((void **)&buf[0], cap, cudaHostAllocDefault) == cudaSuccess &&
             cudaHostAlloc((void **)&buf[1], cap, cudaHostAllocDefault) == cudaSuccess &&
             cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming) == cudaSuccess &&
             cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess;
        if (!ok) cudaGetLastError();
    }
    ~DsbStager()
    {
        if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
        for (int i = 0; i < 2; i++) { if (ev[i]) cudaEventDestroy(ev[i]); if (buf[i]) cudaFreeHost(buf[i]); }
    }
    cudaError_t copy(void *dst, const void *src, size_t bytes)
    {
        if (!ok) return cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice);
        for (size_t off = 0; off < bytes; off += cap, n++) {
            const int b = n & 1;
            const size_t m = bytes - off < cap ? bytes - off : cap;
            if (n >= 2) { cudaError_t e = cudaEventSynchronize(ev[b]); if (e != cudaSuccess) return e; }
            memcpy(buf[b], (const uint8_t *)src + off, m);
            cudaError_t e = cudaMemcpyAsync((uint8_t *)dst + off, buf[b], m, cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess) return e;
            cudaEventRecord(ev[b], st);
        }
        return cudaSuccess;
    }
    cudaError_t finish() { return ok ? cudaStreamSynchronize(st) : cudaSuccess; }
};
static thread_local DsbStager *g_stager = nullptr;      // alive during one dsb_index_upload

template <typename T>
static int upload(dsb_index *ix, const T *src, uint64_t count, uint64_t extra_zero, const T **out)
{
    void *d = nullptr;
    const size_t bytes = (size_t)(count + extra_zero) * sizeof(T);
    CU(cudaMalloc(&d, bytes ? bytes : 16));
    ix->allocs.push_back(d);
    ix->alloc_bytes.push_back(bytes ? bytes : 16);
    if (count) {
        if (g_stager) CU(g_stager->copy(d, src, (size_t)count * sizeof(T)));
        else CU(cudaMemcpy(d, src, (size_t)count * sizeof(T), cudaMemcpyHostToDevice));
    }
    if (extra_zero) CU(cudaMemset((char *)d + (size_t)count * sizeof(T), 0, (size_t)extra_zero * sizeof(T)));
    ix->bytes += bytes;
    *out = (const T *)d;
    return 0;
}

extern "C" {

const char *dsb_last_error(void) { return g_err.c_str(); }
int dsb_kernel_launch_count(void) { return g_launches.load(); }

int dsb_gpu_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { fail("cudaGetDeviceCount", e); return 0; }
    return n;
}

void dsb_index_free(dsb_index *ix)
{
    if (!ix) return;
    cudaSetDevice(ix->dev);
    for (void *p : ix->allocs) cudaFree(p);
    delete ix;
}

int dsb_index_upload(const dsb_host_index *h, int dev, dsb_index **out)
{
    *out = nullptr;
    CU(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) {
        g_err = "device is not sm_100-class; this library carries sm_100a code only";
        return -1;
    }
    dsb_apply_l2_fetch_env();
    dsb_index *ix = new dsb_index();
    ix->dev = dev;
    ix->bytes = 0;
    ix->sm_count = prop.multiProcessorCount;
    DsbIndexView &v = ix->v;
    memset(&v, 0, sizeof v);
    int rc = 0;
    DsbStager stager;
    g_stager = &stager;
    struct StagerScope { ~StagerScope() { g_stager = nullptr; } } stager_scope;
    {   // FM index: the occ array goes up as it is on disk and is re-blocked into 64-symbol / 64-byte blocks by k_fm_reblock
        const uint64_t nb = h->bwt_bytes / 168;
        const uint64_t nblk = dsb_fm_blocks_needed(h->bwt_bytes);
        void *d_fm = nullptr, *d_raw = nullptr;
        CU(cudaMalloc(&d_fm, nblk * sizeof(DsbFmBlock)));
        ix->allocs.push_back(d_fm);
        ix->alloc_bytes.push_back(nblk * sizeof(DsbFmBlock));
        ix->bytes += nblk * sizeof(DsbFmBlock);
        v.fm = (const DsbFmBlock *)d_fm;
        v.n_fm_blocks = nblk;
        v.dollar_row = ~0ull;
        if (nb == 0) {
            CU(cudaMemset(d_fm, 0, nblk * sizeof(DsbFmBlock)));
        } else {
            unsigned long long *d_row = nullptr, h_row = ~0ull;
            CU(cudaMalloc(&d_raw, nb * 168));
            CU(cudaMalloc((void **)&d_row, 8));
            CU(cudaMemcpy(d_row, &h_row, 8, cudaMemcpyHostToDevice));
            CU(stager.copy(d_raw, h->bwt_occ, nb * 168));
            CU(stager.finish());
            k_fm_reblock<<<(unsigned)((nb * 4 + 255) / 256), 256>>>((const uint64_t *)d_raw, nb, (DsbFmBlock *)d_fm, d_row);
            g_launches++;
            cudaError_t e = cudaMemcpy(&h_row, d_row, 8, cudaMemcpyDeviceToHost);
            cudaFree(d_raw);
            cudaFree(d_row);
            if (e != cudaSuccess) { fail("k_fm_reblock", e); dsb_index_free(ix); return -1; }
            v.dollar_row = h_row;
        }
        if (getenv("DSB_VERIFY_REBLOCK")) {              // tests: the device blocks equal the host function's, byte for byte
            std::vector<DsbFmBlock> want(nblk), got(nblk);
            uint64_t bad = 0, want_row = ~0ull;
            dsb_fm_reblock(h->bwt_occ, h->bwt_bytes, &want_row, want.data(), &bad);
            CU(cudaMemcpy(got.data(), d_fm, nblk * sizeof(DsbFmBlock), cudaMemcpyDeviceToHost));
            if (bad) {
                g_err = "FM index: the stored '#' checkpoints differ from the derived ones in " + std::to_string(bad) + " blocks";
                dsb_index_free(ix);
                return -1;
            }
            if (want_row != v.dollar_row || memcmp(want.data(), got.data(), nblk * sizeof(DsbFmBlock)) != 0) {
                g_err = "k_fm_reblock differs from the host re-blocking";
                dsb_index_free(ix);
                return -1;
            }
            fprintf(stderr, "[dsb] k_fm_reblock verified against the host re-blocking: %llu blocks identical\n", (unsigned long long)nblk);
        }
    }
    memcpy(v.rank, h->rank, sizeof v.rank);
    v.dollar_pos = h->dollar_pos;
    rc |= upload(ix, h->hash_index, (1ull << 26) + 1, 1, &v.lut);
    rc |= upload(ix, h->sa_taxon, h->sa_n, 1, &v.sa);
    v.sa_n = h->sa_n;
    rc |= upload(ix, h->e_kmer0, h->e_kmer_size, 0, &v.ek0);
    rc |= upload(ix, h->e_kmer1, h->e_kmer_size, 0, &v.ek1);
    int k;
    uint64_t mask;
    dsb_ekmer_params(h->e_kmer_size, &k, &mask);
    v.ek_mask = mask;
    v.ek_len = k;
    v.single_base_max = (int)(0.8 * k);              // src/idx.c:981
    rc |= upload(ix, h->unitig, h->n_unitig, 2, &v.unitig);
    v.n_unitig = h->n_unitig;
    rc |= upload(ix, h->ref_pos, h->n_ref_pos, 2, &v.ref_pos);
    rc |= upload(ix, h->ref_bin, h->ref_bin_n, 64, &v.ref_bin);
    v.ref_bin_n = h->ref_bin_n;
    rc |= upload(ix, h->ref_seq_offset, h->n_ref, 1, &v.ref_off);
    rc |= upload(ix, h->ref_seq_len, h->n_ref, 1, &v.ref_len);
    v.n_ref = h->n_ref;
    rc |= upload(ix, h->q_mem, 2000, 0, &v.q_mem);
    rc |= upload(ix, h->q_lv, 400, 0, &v.q_lv);
    if (!rc && stager.finish() != cudaSuccess) { g_err = "index upload: staged copy failed"; rc = -1; }
    v.filter_min_length = h->filter_min_length;
    v.filter_min_score = h->filter_min_score;
    v.filter_min_score_lv3 = h->filter_min_score_lv3;
    if (rc) { dsb_index_free(ix); return -1; }
    // the classifier's frames are deep (LV tables, merge-sort stacks): give them room
    cudaDeviceSetLimit(cudaLimitStackSize, 8192);
    *out = ix;
    return 0;
}

// A replica of a resident index on another GPU of the box, copied array by array from the first GPU's HBM
// (cudaMemcpyPeer: NVLink / NVSwitch when peer access is possible, staged by the driver otherwise) instead of being
// uploaded and re-blocked from host memory again.
int dsb_index_clone(const dsb_index *src, int dev, dsb_index **out)
{
    *out = nullptr;
    if (!src) { g_err = "dsb_index_clone: no source index"; return -1; }
    CU(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) {
        g_err = "device is not sm_100-class; this library carries sm_100a code only";
        return -1;
    }
    if (dev != src->dev) {
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, dev, src->dev) == cudaSuccess && can) {
            cudaError_t e = cudaDeviceEnablePeerAccess(src->dev, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { fail("cudaDeviceEnablePeerAccess", e); return -1; }
            cudaGetLastError();
        }
    }
    dsb_index *ix = new dsb_index();
    ix->dev = dev;
    ix->bytes = 0;
    ix->sm_count = prop.multiProcessorCount;
    ix->v = src->v;
    for (size_t i = 0; i < src->allocs.size(); i++) {
        void *d = nullptr;
        const size_t bytes = src->alloc_bytes[i];
        cudaError_t e = cudaMalloc(&d, bytes);
        if (e == cudaSuccess) {
            ix->allocs.push_back(d);
            ix->alloc_bytes.push_back(bytes);
            ix->bytes += bytes;
            e = cudaMemcpyPeer(d, dev, src->allocs[i], src->dev, bytes);
        }
        if (e != cudaSuccess) { fail("dsb_index_clone", e); dsb_index_free(ix); return -1; }
    }
    CU(cudaDeviceSynchronize());
    // every device pointer of the view moves to the same offset of the corresponding new array
    bool ok = true;
    auto remap = [&](const void *p) -> const void * {
        if (!p) return nullptr;
        for (size_t i = 0; i < src->allocs.size(); i++) {
            const char *b = (const char *)src->allocs[i];
            if ((const char *)p >= b && (const char *)p < b + src->alloc_bytes[i]) return (const char *)ix->allocs[i] + ((const char *)p - b);
        }
        ok = false;
        return nullptr;
    };
    DsbIndexView &v = ix->v;
    v.fm = (const DsbFmBlock *)remap(v.fm);
    v.lut = (const uint64_t *)remap(v.lut);
    v.sa = (const uint64_t *)remap(v.sa);
    v.ek0 = (const uint8_t *)remap(v.ek0);
    v.ek1 = (const uint8_t *)remap(v.ek1);
    v.unitig = (const uint64_t *)remap(v.unitig);
    v.ref_pos = (const uint64_t *)remap(v.ref_pos);
    v.ref_bin = (const uint8_t *)remap(v.ref_bin);
    v.ref_off = (const uint64_t *)remap(v.ref_off);
    v.ref_len = (const uint64_t *)remap(v.ref_len);
    v.q_mem = (const int32_t *)remap(v.q_mem);
    v.q_lv = (const int32_t *)remap(v.q_lv);
    if (!ok) { g_err = "dsb_index_clone: a view pointer lies outside the index's arrays"; dsb_index_free(ix); return -1; }
    cudaDeviceSetLimit(cudaLimitStackSize, 8192);
    *out = ix;
    return 0;
}

uint64_t dsb_index_device_bytes(const dsb_index *ix) { return ix ? ix->bytes : 0; }

// ------------------------------------------------------------------------------------------
// C-ABI: slots (one CUDA stream + pinned staging + device buffers)
// ------------------------------------------------------------------------------------------
static int ensure_scratch(dsb_slot *s, uint32_t Lmax)
{
    DsbScratchDims d;
    dsb_scratch_dims(&d, Lmax, s->cap_anchor, s->cap_sms);
    const size_t need = d.bytes * (size_t)s->n_warps;
    if (need > s->scratch_bytes) {
        if (s->d_scratch) CU(cudaFree(s->d_scratch));
        s->d_scratch = nullptr;
        s->scratch_bytes = 0;
        CU(cudaMalloc((void **)&s->d_scratch, need));
        s->scratch_bytes = need;
    }
    {
        // tests: a lane of the slow path gets so few MEM slots that strands fall back to the serial form (tests/test_gpu.py)
        const char *e = getenv("DSB_TEST_SLOW_LANE_MEMS");
        if (e && atoi(e) > 0) d.slow_lane_mems = (uint32_t)atoi(e);
    }
    s->dims = d;
    return 0;
}

static int ensure_hits(dsb_slot *s, uint32_t cap)
{
    if (cap > s->cap_hits) {
        if (s->d_hits) CU(cudaFree(s->d_hits));
        s->d_hits = nullptr;
        s->cap_hits = 0;
        CU(cudaMalloc((void **)&s->d_hits, (size_t)cap * sizeof(dsb_hit)));
        s->cap_hits = cap;
    }
    return 0;
}

void dsb_slot_free(dsb_slot *s)
{
    if (!s) return;
    cudaSetDevice(s->ix->dev);
    if (s->stream) cudaStreamSynchronize(s->stream);
    cudaFreeHost(s->h_seq); cudaFreeHost(s->h_desc); cudaFreeHost(s->h_edge); cudaFreeHost(s->h_results);
    cudaFreeHost(s->h_hits); cudaFreeHost(s->h_ctrl); cudaFreeHost(s->h_order);
    cudaFree(s->d_seq); cudaFree(s->d_desc); cudaFree(s->d_edge); cudaFree(s->d_bin); cudaFree(s->d_exist); cudaFree(s->d_h1); cudaFree(s->d_h2);
    cudaFree(s->d_results); cudaFree(s->d_hits); cudaFree(s->d_ctrl); cudaFree(s->d_order); cudaFree(s->d_scratch);
    cudaFree(s->d_rinfo); cudaFree(s->d_jobs); cudaFree(s->d_pool);
    for (int i = 0; i < 8; i++) if (s->ev[i]) cudaEventDestroy(s->ev[i]);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

int dsb_slot_create(dsb_index *ix, uint64_t max_bases, uint32_t max_reads, dsb_slot **out)
{
    *out = nullptr;
    CU(cudaSetDevice(ix->dev));
    dsb_slot *s = new dsb_slot();
    memset((void *)s, 0, sizeof *s);
    s->ix = ix;
    s->max_bases = max_bases;
    s->max_reads = max_reads;
    const size_t edge_bytes = (size_t)max_reads * (DSB_HEAD_BYTES + DSB_TAIL_BYTES);
    int rc = 0;
#define TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { rc = fail(#call, e_); goto done; } } while (0)
    TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    for (int i = 0; i < 8; i++) TRY(cudaEventCreate(&s->ev[i]));
    TRY(cudaHostAlloc((void **)&s->h_seq, max_bases + 64, cudaHostAllocDefault));
    TRY(cudaHostAlloc((void **)&s->h_desc, (size_t)max_reads * sizeof(dsb_read_desc), cudaHostAllocDefault));
    TRY(cudaHostAlloc((void **)&s->h_edge, edge_bytes, cudaHostAllocDefault));
    TRY(cudaHostAlloc((void **)&s->h_results, (size_t)max_reads * sizeof(dsb_read_result), cudaHostAllocDefault));
    TRY(cudaHostAlloc((void **)&s->h_ctrl, sizeof(DsbCtrl), cudaHostAllocDefault));
    TRY(cudaHostAlloc((void **)&s->h_order, (size_t)max_reads * 4, cudaHostAllocDefault));
    TRY(cudaMalloc((void **)&s->d_seq, max_bases + 64));
    TRY(cudaMalloc((void **)&s->d_desc, (size_t)max_reads * sizeof(dsb_read_desc)));
    TRY(cudaMalloc((void **)&s->d_edge, edge_bytes));
    TRY(cudaMalloc((void **)&s->d_bin, dsb_bin_bytes(max_bases, max_reads)));
    {
        // A/B on cfg3 (profiles/ab_r02.txt): probing on demand saves only 16 % of the first-table probes -- on a 1 Gbp index
        // one 17-mer in ten passes both tables, so the scan meets an island every ~30 positions and looks at most of the row
        // anyway -- and turns a bandwidth-bound pass into a latency-bound one: 72.7 ms against 49.8 + 9.7 ms
        const char *e = getenv("DSB_EXIST_ON_DEMAND");
        s->exist_bytes = !(e && atoi(e) != 0);
    }
    if (s->exist_bytes) TRY(cudaMalloc((void **)&s->d_exist, 2 * max_bases + 64));
    if (s->exist_bytes) {
        // probes in hash-range passes (see k_exist_hash): DSB_EXIST_SLICE_MB = slice size (a power of two); default 0 = the
        // one-pass k_exist, which is still the faster one (profiles/ab_r02.txt)
        const uint64_t table_bits = ix->v.ek_mask + 1;
        uint64_t slice_mb = 0;
        const char *e = getenv("DSB_EXIST_SLICE_MB");
        if (e) slice_mb = (uint64_t)atoll(e);
        s->exist_slices = 0;
        if (slice_mb > 0 && (slice_mb & (slice_mb - 1)) == 0 && table_bits > (slice_mb << 23) && (table_bits >> 35) <= 1) {
            uint32_t sh = 23;
            while ((1ull << sh) < (slice_mb << 23)) sh++;
            s->exist_slice_shift = sh;
            s->exist_slices = (int)(table_bits >> sh);
            const size_t n_h = 2 * (size_t)max_bases + 32 * (size_t)max_reads + 64;
            TRY(cudaMalloc((void **)&s->d_h1, n_h * 4));
            TRY(cudaMalloc((void **)&s->d_h2, n_h * 4));
        }
    }
    TRY(cudaMalloc((void **)&s->d_results, (size_t)max_reads * sizeof(dsb_read_result)));
    TRY(cudaMalloc((void **)&s->d_ctrl, sizeof(DsbCtrl)));
    TRY(cudaMalloc((void **)&s->d_order, (size_t)max_reads * 4));
    {
        // L1 / shared-memory split of the classify kernels: the driver's own choice is the best one measured (classify stage
        // 10.06 ms; preferring the largest L1: 10.8 ms; 25 % shared memory: 13.2 ms).  DSB_CARVEOUT=<percent> overrides.
        const char *ce = getenv("DSB_CARVEOUT");
        const int carve = ce ? atoi(ce) : -1;
        if (carve >= 0) {
            cudaFuncSetAttribute(k_islands, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_seed, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_classify, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaGetLastError();
        }
        // persistent grid: as many warps as the device keeps resident
        int per_sm = 0;
        TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_classify, DSB_CLASSIFY_THREADS, 0));
        if (per_sm < 1) per_sm = 1;
        const char *e = getenv("DSB_BLOCKS_PER_SM");
        if (e && atoi(e) > 0 && atoi(e) < per_sm) per_sm = atoi(e);
        s->threads = DSB_CLASSIFY_THREADS;
        s->blocks = per_sm * ix->sm_count;
        s->n_warps = s->blocks * (DSB_CLASSIFY_THREADS / 32);
        int per_sm_seed = 0;
        TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_seed, k_seed, DSB_CLASSIFY_THREADS, 0));
        if (per_sm_seed < 1) per_sm_seed = 1;
        const char *es = getenv("DSB_SEED_BLOCKS_PER_SM");
        if (es && atoi(es) > 0 && atoi(es) < per_sm_seed) per_sm_seed = atoi(es);
        s->seed_blocks = per_sm_seed * ix->sm_count;
        if (s->seed_blocks * (DSB_CLASSIFY_THREADS / 32) > s->n_warps) s->n_warps = s->seed_blocks * (DSB_CLASSIFY_THREADS / 32);
    }
    {
        // North-star item N1: pin part of an index array in L2 with an access-policy window on the slot's stream.
        // DSB_L2_PERSIST=lut|ek0|fm|sa picks the array (the window covers its first accessPolicyMaxWindowSize bytes, hit ratio =
        // persisting capacity / window).  Off by default: the path's index accesses are uniformly random over arrays 4 - 30x
        // the L2, there is no hot subset to keep (A/B on the 1 Gbp index: profiles/ab_r02.txt).
        const char *e = getenv("DSB_L2_PERSIST");
        if (e && *e) {
            cudaDeviceProp prop;
            TRY(cudaGetDeviceProperties(&prop, ix->dev));
            const void *base = nullptr;
            size_t bytes = 0;
            if (!strcmp(e, "lut")) { base = ix->v.lut; bytes = ((size_t)1 << 26) * 8; }
            else if (!strcmp(e, "ek0")) { base = ix->v.ek0; bytes = (size_t)(ix->v.ek_mask + 1) / 8; }
            else if (!strcmp(e, "fm")) { base = ix->v.fm; bytes = (size_t)ix->v.n_fm_blocks * sizeof(DsbFmBlock); }
            else if (!strcmp(e, "sa")) { base = ix->v.sa; bytes = (size_t)ix->v.sa_n * 8; }
            if (base && prop.persistingL2CacheMaxSize > 0) {
                cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)prop.persistingL2CacheMaxSize);
                cudaStreamAttrValue a;
                memset(&a, 0, sizeof a);
                a.accessPolicyWindow.base_ptr = const_cast<void *>(base);
                a.accessPolicyWindow.num_bytes = bytes < (size_t)prop.accessPolicyMaxWindowSize ? bytes : (size_t)prop.accessPolicyMaxWindowSize;
                const double r = (double)prop.persistingL2CacheMaxSize / (double)a.accessPolicyWindow.num_bytes;
                a.accessPolicyWindow.hitRatio = (float)(r < 1.0 ? r : 1.0);
                a.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                a.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                cudaError_t pe = cudaStreamSetAttribute(s->stream, cudaStreamAttributeAccessPolicyWindow, &a);
                if (getenv("DSB_VERBOSE"))
                    fprintf(stderr, "[dsb] L2 access-policy window on %s: %zu MB window, %d MB persisting, hit ratio %.2f (%s)\n", e,
                            a.accessPolicyWindow.num_bytes >> 20, prop.persistingL2CacheMaxSize >> 20, a.accessPolicyWindow.hitRatio,
                            cudaGetErrorString(pe));
                cudaGetLastError();
            }
        }
    }
    s->cap_anchor = 2048;
    s->cap_sms = 8192;
    {
        const char *e = getenv("DSB_SPLIT");
        s->split = !(e && atoi(e) == 0);
        if (s->split) {
            // every read lists at most 2 * (L / 100 + 2) top islands
            s->cap_jobs = (uint32_t)(max_bases / 50) + 12u * max_reads + 64u;
            TRY(cudaMalloc((void **)&s->d_rinfo, (size_t)max_reads * sizeof(DsbReadSeeds)));
            TRY(cudaMalloc((void **)&s->d_jobs, (size_t)s->cap_jobs * sizeof(DsbSeedJob)));
            s->cap_pool = (uint32_t)(max_bases / 8) + 64u * max_reads + 1024u;
            // tests: start with arenas that are far too small, so that the regrow-and-rerun path runs (tests/test_gpu.py)
            if (getenv("DSB_TEST_TINY_ARENAS")) { s->cap_jobs = 64; s->cap_pool = 256; }
            TRY(cudaMalloc((void **)&s->d_pool, (size_t)s->cap_pool * sizeof(DsbAnchor)));
        }
    }
    if (ensure_hits(s, max_reads * 8u + 4096u)) { rc = -1; goto done; }
    s->h_cap_hits = s->cap_hits;
    TRY(cudaHostAlloc((void **)&s->h_hits, (size_t)s->h_cap_hits * sizeof(dsb_hit), cudaHostAllocDefault));
#undef TRY
done:
    if (rc) { dsb_slot_free(s); return rc; }
    *out = s;
    return 0;
}

int dsb_slot_staging(dsb_slot *s, char **seq, dsb_read_desc **desc, uint8_t **edge)
{
    *seq = s->h_seq;
    *desc = s->h_desc;
    *edge = s->h_edge;
    return 0;
}

// kernels of one batch, inputs already on the device
static int launch_kernels(dsb_slot *s)
{
    const uint32_t n = s->n_reads;
    CU(cudaMemsetAsync(s->d_ctrl, 0, sizeof(DsbCtrl), s->stream));
    const int sm = s->ix->sm_count;
    const uint32_t grid_r = n < (uint32_t)(sm * 8) ? n : (uint32_t)(sm * 8);
    k_encode<<<grid_r, 256, 0, s->stream>>>(s->d_seq, s->d_desc, s->d_edge, s->d_bin, n);
    CU(cudaEventRecord(s->ev[1], s->stream));
    if (s->exist_bytes && s->exist_slices > 0) {
        k_exist_hash<<<grid_r, 256, 0, s->stream>>>(s->ix->v, s->d_desc, s->d_bin, s->d_exist, s->d_h1, s->d_h2, n);
        for (uint32_t stage = 0; stage < 2; stage++)
            for (int sl = 0; sl < s->exist_slices; sl++)
                k_exist_pass<<<grid_r, 256, 0, s->stream>>>(s->ix->v, s->d_desc, s->d_exist, stage ? s->d_h2 : s->d_h1, n, stage,
                                                            (uint32_t)sl, s->exist_slice_shift, s->d_ctrl);
        g_launches += 2 * s->exist_slices;
    } else if (s->exist_bytes)
        k_exist<<<grid_r, 256, 0, s->stream>>>(s->ix->v, s->d_desc, s->d_bin, s->d_exist, n, s->d_ctrl);
    CU(cudaEventRecord(s->ev[2], s->stream));
    // small batches: no more warps than reads (the other slots' kernels fill the rest of the device)
    const int warps_per_block = s->threads / 32;
    int blocks = (int)((n + warps_per_block - 1) / warps_per_block);
    if (blocks > s->blocks) blocks = s->blocks;
    if (s->split) {
        k_islands<<<blocks, s->threads, 0, s->stream>>>(s->ix->v, s->d_desc, s->d_order, s->d_bin, s->d_exist, s->d_scratch,
                                                          s->dims, s->d_rinfo, s->d_jobs, s->cap_jobs, n, s->d_ctrl);
        CU(cudaEventRecord(s->ev[6], s->stream));
        // the island jobs of the whole batch, one lane each: always the full persistent grid
        k_seed<<<s->seed_blocks, s->threads, 0, s->stream>>>(s->ix->v, s->d_desc, s->d_bin, s->d_scratch, s->dims, s->d_jobs,
                                                         s->d_pool, s->cap_pool, s->d_ctrl);
        CU(cudaEventRecord(s->ev[7], s->stream));
        k_classify<<<blocks, s->threads, 0, s->stream>>>(s->ix->v, s->d_desc, s->d_order, s->d_bin, s->d_exist, s->d_scratch,
                                                           s->dims, s->d_results, s->d_hits, s->cap_hits, n, s->d_ctrl,
                                                           s->d_rinfo, s->d_jobs, s->d_pool);
        g_launches += 2;
    } else {
        k_classify<<<blocks, s->threads, 0, s->stream>>>(s->ix->v, s->d_desc, s->d_order, s->d_bin, s->d_exist, s->d_scratch,
                                                           s->dims, s->d_results, s->d_hits, s->cap_hits, n, s->d_ctrl,
                                                           nullptr, nullptr, nullptr);
    }
    CU(cudaEventRecord(s->ev[3], s->stream));
    CU(cudaGetLastError());
    g_launches += s->exist_bytes ? 3 : 2;
    CU(cudaMemcpyAsync(s->h_ctrl, s->d_ctrl, sizeof(DsbCtrl), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaMemcpyAsync(s->h_results, s->d_results, (size_t)n * sizeof(dsb_read_result), cudaMemcpyDeviceToHost, s->stream));
    return 0;
}

int dsb_slot_submit(dsb_slot *s, uint32_t n_reads, uint64_t n_bases)
{
    if (n_reads == 0 || n_reads > s->max_reads || n_bases > s->max_bases) {
        g_err = "dsb_slot_submit: batch exceeds the slot's capacity";
        return -1;
    }
    CU(cudaSetDevice(s->ix->dev));
    s->n_reads = n_reads;
    s->n_bases = n_bases;
    // longest reads first: the persistent grid's tail is then made of short reads
    uint32_t Lmax = 0;
    {
        std::vector<uint64_t> key(n_reads);
        for (uint32_t i = 0; i < n_reads; i++) {
            const uint32_t L = s->h_desc[i].len;
            if (L > Lmax) Lmax = L;
            key[i] = ((uint64_t)(0xffffffffu - L) << 32) | i;
        }
        std::sort(key.begin(), key.end());
        for (uint32_t i = 0; i < n_reads; i++) s->h_order[i] = (uint32_t)key[i];
    }
    s->Lmax = Lmax;
    if (ensure_scratch(s, Lmax)) return -1;
    CU(cudaEventRecord(s->ev[0], s->stream));
    CU(cudaMemcpyAsync(s->d_seq, s->h_seq, n_bases, cudaMemcpyHostToDevice, s->stream));
    CU(cudaMemcpyAsync(s->d_desc, s->h_desc, (size_t)n_reads * sizeof(dsb_read_desc), cudaMemcpyHostToDevice, s->stream));
    CU(cudaMemcpyAsync(s->d_edge, s->h_edge, (size_t)n_reads * (DSB_HEAD_BYTES + DSB_TAIL_BYTES), cudaMemcpyHostToDevice,
                       s->stream));
    CU(cudaMemcpyAsync(s->d_order, s->h_order, (size_t)n_reads * 4, cudaMemcpyHostToDevice, s->stream));
    if (launch_kernels(s)) return -1;
    s->in_flight = true;
    return 0;
}

int dsb_slot_rerun_device(dsb_slot *s)
{
    if (s->n_reads == 0) { g_err = "dsb_slot_rerun_device: nothing was submitted"; return -1; }
    CU(cudaSetDevice(s->ix->dev));
    CU(cudaEventRecord(s->ev[0], s->stream));
    if (launch_kernels(s)) return -1;
    s->in_flight = true;
    return 0;
}

int dsb_slot_wait(dsb_slot *s, const dsb_read_result **results, const dsb_hit **hits, uint32_t *n_hits,
                  dsb_counters *counters)
{
    if (!s->in_flight) { g_err = "dsb_slot_wait: no batch in flight"; return -1; }
    CU(cudaSetDevice(s->ix->dev));
    for (int attempt = 0;; attempt++) {
        CU(cudaStreamSynchronize(s->stream));
        const uint32_t of = s->h_ctrl->overflow;
        if (of == 0) break;
        if (of & 2u) {
            s->in_flight = false;
            g_err = "a chain gap of >= 2000 reference bases reached sdp_middle: the reference aborts here (src/cly.c:2468)";
            return -2;
        }
        if (of & 16u) {
            s->in_flight = false;
            g_err = "a read of 2^24 bases or more reached the device: the read's 9-mer table holds 24-bit positions";
            return -3;
        }
        if (attempt >= 8) { s->in_flight = false; g_err = "scratch arenas still too small after 8 regrows"; return -1; }
        // grow whatever overflowed and run the batch again (inputs are still on the device)
        if (of & 1u) {
            s->cap_anchor *= 4;
            s->cap_sms *= 4;
            if (ensure_scratch(s, s->Lmax)) return -1;
        }
        if (of & 4u) {
            if (ensure_hits(s, s->cap_hits * 4u)) return -1;
        }
        if (of & 8u) {                                 // anchor pool (or job list) of the split pipeline
            cudaFree(s->d_pool); cudaFree(s->d_jobs);
            s->d_pool = nullptr; s->d_jobs = nullptr;
            // both allocators keep counting past their capacity, so they say how much this batch needs (the pool's count is
            // only complete when k_seed ran, i.e. when the job list fitted)
            s->cap_pool *= 2; s->cap_jobs *= 2;
            if (s->h_ctrl->n_jobs > s->cap_jobs) s->cap_jobs = s->h_ctrl->n_jobs + 64u;
            if (s->h_ctrl->n_pool > s->cap_pool) s->cap_pool = s->h_ctrl->n_pool + s->h_ctrl->n_pool / 4u + 1024u;
            CU(cudaMalloc((void **)&s->d_pool, (size_t)s->cap_pool * sizeof(DsbAnchor)));
            CU(cudaMalloc((void **)&s->d_jobs, (size_t)s->cap_jobs * sizeof(DsbSeedJob)));
        }
        if (launch_kernels(s)) return -1;
    }
    const uint32_t nh = s->h_ctrl->n_hits;
    if (nh > s->h_cap_hits) {
        cudaFreeHost(s->h_hits);
        s->h_hits = nullptr;
        s->h_cap_hits = 0;
        CU(cudaHostAlloc((void **)&s->h_hits, (size_t)s->cap_hits * sizeof(dsb_hit), cudaHostAllocDefault));
        s->h_cap_hits = s->cap_hits;
    }
    if (nh) CU(cudaMemcpyAsync(s->h_hits, s->d_hits, (size_t)nh * sizeof(dsb_hit), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaEventRecord(s->ev[4], s->stream));
    CU(cudaStreamSynchronize(s->stream));
    cudaEventElapsedTime(&s->ms_total, s->ev[0], s->ev[4]);
    cudaEventElapsedTime(&s->ms_exist, s->ev[1], s->ev[2]);
    cudaEventElapsedTime(&s->ms_classify, s->ev[2], s->ev[3]);
    s->ms_islands = s->ms_seed = 0.f;
    if (s->split) {
        cudaEventElapsedTime(&s->ms_islands, s->ev[2], s->ev[6]);
        cudaEventElapsedTime(&s->ms_seed, s->ev[6], s->ev[7]);
    }
    s->in_flight = false;
    *results = s->h_results;
    *hits = s->h_hits;
    *n_hits = nh;
    if (counters) *counters = s->h_ctrl->counters;
    return 0;
}

int dsb_slot_timing(dsb_slot *s, float *ms_total, float *ms_exist, float *ms_classify)
{
    *ms_total = s->ms_total;
    *ms_exist = s->ms_exist;
    *ms_classify = s->ms_classify;
    return 0;
}

int dsb_slot_timing_stages(dsb_slot *s, float *ms_islands, float *ms_seed, float *ms_score)
{
    *ms_islands = s->ms_islands;
    *ms_seed = s->ms_seed;
    *ms_score = s->ms_classify - s->ms_islands - s->ms_seed;
    return s->split ? 1 : 0;
}

int dsb_fm_rank_batch(dsb_index *ix, const uint64_t *row, const uint8_t *c, uint64_t n, uint64_t *out, uint8_t *c_out,
                      int repeat, float *ms)
{
    CU(cudaSetDevice(ix->dev));
    uint64_t *d_row = nullptr, *d_out = nullptr;
    uint8_t *d_c = nullptr, *d_co = nullptr;
    cudaEvent_t e0, e1;
    CU(cudaMalloc((void **)&d_row, n * 8 + 8));
    CU(cudaMalloc((void **)&d_out, n * 8 + 8));
    CU(cudaMalloc((void **)&d_c, n + 8));
    CU(cudaMalloc((void **)&d_co, n + 8));
    CU(cudaMemcpy(d_row, row, n * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_c, c, n, cudaMemcpyHostToDevice));
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int grid = ix->sm_count * 8;
    if (repeat < 1) repeat = 1;
    k_fm_rank<<<grid, 256>>>(ix->v, d_row, d_c, n, d_out, d_co);        // warm-up
    CU(cudaEventRecord(e0));
    for (int r = 0; r < repeat; r++) k_fm_rank<<<grid, 256>>>(ix->v, d_row, d_c, n, d_out, d_co);
    CU(cudaEventRecord(e1));
    CU(cudaDeviceSynchronize());
    CU(cudaGetLastError());
    g_launches += repeat + 1;
    float t = 0;
    cudaEventElapsedTime(&t, e0, e1);
    if (ms) *ms = t / repeat;
    CU(cudaMemcpy(out, d_out, n * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(c_out, d_co, n, cudaMemcpyDeviceToHost));
    cudaFree(d_row); cudaFree(d_out); cudaFree(d_c); cudaFree(d_co);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

int dsb_random_access_probe(int dev, uint64_t table_bytes, uint32_t chunk_bytes, uint64_t n_access, int repeat, float *ms)
{
    CU(cudaSetDevice(dev));
    if (chunk_bytes != 32 && chunk_bytes != 64 && chunk_bytes != 128) { g_err = "chunk_bytes must be 32, 64 or 128"; return -1; }
    dsb_apply_l2_fetch_env();
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    uint4 *table = nullptr;
    uint32_t *sink = nullptr;
    cudaEvent_t e0, e1;
    CU(cudaMalloc((void **)&table, table_bytes));
    CU(cudaMalloc((void **)&sink, 64));
    CU(cudaMemset(table, 0x5a, table_bytes));
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int grid = prop.multiProcessorCount * 8;
    const uint64_t n_chunks = table_bytes / chunk_bytes;
    if (repeat < 1) repeat = 1;
    for (int r = -1; r < repeat; r++) {
        if (r == 0) CU(cudaEventRecord(e0));
        if (chunk_bytes == 32) k_random_gather<32><<<grid, 256>>>(table, n_chunks, n_access, sink);
        else if (chunk_bytes == 64) k_random_gather<64><<<grid, 256>>>(table, n_chunks, n_access, sink);
        else k_random_gather<128><<<grid, 256>>>(table, n_chunks, n_access, sink);
    }
    CU(cudaEventRecord(e1));
    CU(cudaDeviceSynchronize());
    CU(cudaGetLastError());
    g_launches += repeat + 1;
    float t = 0;
    cudaEventElapsedTime(&t, e0, e1);
    if (ms) *ms = t / repeat;
    cudaFree(table); cudaFree(sink);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

int dsb_exist_batch(dsb_index *ix, const uint8_t *codes, uint64_t n, uint8_t *exist, int repeat, float *ms)
{
    CU(cudaSetDevice(ix->dev));
    const uint64_t k = (uint64_t)ix->v.ek_len;
    if (n < k) { g_err = "dsb_exist_batch: string shorter than k"; return -1; }
    const uint64_t nk = n - k + 1;
    uint8_t *d_codes = nullptr, *d_ex = nullptr;
    cudaEvent_t e0, e1;
    CU(cudaMalloc((void **)&d_codes, n + 64));
    CU(cudaMalloc((void **)&d_ex, nk + 64));
    CU(cudaMemcpy(d_codes, codes, n, cudaMemcpyHostToDevice));
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int grid = ix->sm_count * 8;
    if (repeat < 1) repeat = 1;
    k_exist_flat<<<grid, 256>>>(ix->v, d_codes, nk, d_ex);
    CU(cudaEventRecord(e0));
    for (int r = 0; r < repeat; r++) k_exist_flat<<<grid, 256>>>(ix->v, d_codes, nk, d_ex);
    CU(cudaEventRecord(e1));
    CU(cudaDeviceSynchronize());
    CU(cudaGetLastError());
    g_launches += repeat + 1;
    float t = 0;
    cudaEventElapsedTime(&t, e0, e1);
    if (ms) *ms = t / repeat;
    CU(cudaMemcpy(exist, d_ex, nk, cudaMemcpyDeviceToHost));
    cudaFree(d_codes); cudaFree(d_ex);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

// ------------------------------------------------------------------------------------------
// Exist-k-mer probes in bulk WITHOUT any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  The device-side working buffers of dsb_exist_batch -- the 2-bit codes the kernel reads and the exist bytes
// it writes -- live in ONE file-scope device array of DSB_EXIST_STATIC_MAX codes + as many exist bytes (8 MiB + padding,
// held by every GPU that loads this library until the process exits).  In host code `g_exist_static` is only a stand-in
// for the symbol, never a device address: the address is fetched once per device, with that device current, by
// cudaGetSymbolAddress and handed to the kernel.  All calls share the array and it keeps the last call's contents: the
// part a call uses is cleared first, and two calls that overlap on one GPU race (callers serialise them; within a
// process a host mutex does).
// ------------------------------------------------------------------------------------------
#define DSB_EXIST_STATIC_PAD 256
static __device__ __align__(256) uint8_t g_exist_static[2 * (DSB_EXIST_STATIC_MAX + DSB_EXIST_STATIC_PAD)];

// same work as k_exist_flat, on the static scratch: codes at scratch[0 ..), exist bytes at scratch[half ..)
__global__ void __launch_bounds__(256) k_exist_flat_static(const __grid_constant__ DsbIndexView ix, uint8_t *__restrict__ scratch,
                                                           uint64_t n_kmer)
{
    const uint8_t *codes = scratch;
    uint8_t *exist = scratch + (DSB_EXIST_STATIC_MAX + DSB_EXIST_STATIC_PAD);
    uint32_t p0 = 0, p1 = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_kmer; i += (uint64_t)gridDim.x * blockDim.x)
        exist[i] = dsb_exist_kmer(&ix, codes + i, &p0, &p1);
}

#define DSB_MAX_DEVICES 64
static std::mutex g_exist_static_mu;
static uint8_t *g_exist_static_ptr[DSB_MAX_DEVICES];      // device address of g_exist_static on each GPU (set up on first use)

int dsb_exist_batch_static(dsb_index *ix, const uint8_t *codes, uint64_t n, uint8_t *exist)
{
    if (!ix || !codes || !exist) { g_err = "dsb_exist_batch_static: null argument"; return -1; }
    const uint64_t k = (uint64_t)ix->v.ek_len;
    if (n < k) { g_err = "dsb_exist_batch_static: string shorter than k"; return -1; }
    if (n > (uint64_t)DSB_EXIST_STATIC_MAX) {
        g_err = "dsb_exist_batch_static: " + std::to_string(n) + " codes exceed the static scratch of " +
                std::to_string((uint64_t)DSB_EXIST_STATIC_MAX) + " (DSB_EXIST_STATIC_MAX); split the string or use dsb_exist_batch";
        return -1;
    }
    if (ix->dev < 0 || ix->dev >= DSB_MAX_DEVICES) { g_err = "dsb_exist_batch_static: device ordinal out of range"; return -1; }
    std::lock_guard<std::mutex> lock(g_exist_static_mu);
    CU(cudaSetDevice(ix->dev));
    uint8_t *scratch = g_exist_static_ptr[ix->dev];
    if (!scratch) {                                     // once per device, with that device current
        void *p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_exist_static));
        if (!p) { g_err = "dsb_exist_batch_static: cudaGetSymbolAddress returned no address"; return -1; }
        scratch = g_exist_static_ptr[ix->dev] = (uint8_t *)p;
    }
    const uint64_t nk = n - k + 1;
    const uint64_t half = DSB_EXIST_STATIC_MAX + DSB_EXIST_STATIC_PAD;
    // the array keeps the last call's contents: clear what this call reads beyond its codes (the k-mer of the last start
    // reads up to 16 bytes) and the exist bytes it is about to produce
    CU(cudaMemset(scratch + n, 0, DSB_EXIST_STATIC_PAD));
    CU(cudaMemset(scratch + half, 0, nk));
    CU(cudaMemcpy(scratch, codes, n, cudaMemcpyHostToDevice));
    k_exist_flat_static<<<ix->sm_count * 8, 256>>>(ix->v, scratch, nk);
    CU(cudaGetLastError());
    g_launches++;
    CU(cudaMemcpy(exist, scratch + half, nk, cudaMemcpyDeviceToHost));     // synchronises with the kernel (default stream)
    return 0;
}

}  // extern "C"
