/* dsb_cuda.h -- thin C-ABI in front of the sm_100a classify kernels (libdsb_cuda.so).
 *
 * This is the FFI boundary the reference's C host code would bind for its hot path: the
 * per-read classifier `classify_seq` (reference src/cly.c:3059-3127) as it is driven by
 * `worker_for` / `classify_pipeline` step 1 (reference src/cly_mt.c:352-376), plus the one-time
 * index residency that replaces `load_idx` keeping host arrays (reference src/idx.c:1103-1160,
 * src/bwt.c:68-104).  Plain pointers and sizes only; no C++/torch types cross this line.
 * Every function returns 0 on success or a negative code; dsb_last_error() describes it.
 */
#ifndef DSB_CUDA_H
#define DSB_CUDA_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- index ------------------------------------------------------------------------------- */

/* Host view of a loaded index, exactly the arrays of the reference's DA_IDX that classify_seq
 * touches (reference src/idx.h:83-110, src/bwt.h:15-42).  All pointers are HOST pointers that
 * only need to stay valid during dsb_index_upload(). */
typedef struct dsb_host_index {
    const uint8_t  *bwt_occ;      /* reference FM blocks: 168 B = 5 x u64 counts + 128 B of 4-bit symbols (src/bwt.c:32-41) */
    uint64_t        bwt_bytes;    /* bt.byteLen */
    uint64_t        rank[6];      /* bt.rank, rank[5] = rank[0]-1 (src/bwt.c:81) */
    uint64_t        dollar_pos;   /* bt.DOLLOR_POS (src/idx.c:1128) */
    const uint64_t *hash_index;   /* 13-mer prefix intervals, 2^26+1 entries (src/bwt.c:83-85) */
    const uint64_t *sa_taxon;     /* SA_taxon {u32 unitig_ID; u32 offset} viewed as u64 (src/bwt.h:10-13) */
    uint64_t        sa_n;
    const uint8_t  *e_kmer0;      /* exist-k-mer bit tables (src/idx.h:71-79) */
    const uint8_t  *e_kmer1;
    uint64_t        e_kmer_size;  /* bytes per table; fixes k and the hash mask (src/idx.c:966-982) */
    const uint64_t *unitig;       /* UNITIG {u32 ref_list; u32 length} as u64, n_unitig entries INCLUDING the sentinel (src/idx.c:1127) */
    uint64_t        n_unitig;
    const uint64_t *ref_pos;      /* REF_POS bit-field {global_offset:40, ref_ID:23, direction:1} (src/idx.h:42-48) */
    uint64_t        n_ref_pos;
    const uint8_t  *ref_bin;      /* 2-bit packed references, 4 bases / byte, first base in the top bits (src/cly.c:434-461) */
    uint64_t        ref_bin_n;
    const uint64_t *ref_seq_offset; /* REF_INFO.seq_offset per reference (src/idx.h:22-26) */
    const uint64_t *ref_seq_len;    /* REF_INFO.seq_l */
    uint64_t        n_ref;
    const int32_t  *q_mem;        /* MAPQ.Q_MEM[2000]   (src/cly_mt.c:396-420) */
    const int32_t  *q_lv;         /* MAPQ.Q_LV[20][20] row-major */
    int32_t         filter_min_length;     /* idx.filter_min_length     (src/cly_mt.c:518-520) */
    int32_t         filter_min_score;      /* idx.filter_min_score */
    int32_t         filter_min_score_lv3;  /* idx.filter_min_score_LV3 */
} dsb_host_index;

typedef struct dsb_index dsb_index;   /* device-resident, immutable after upload */

int  dsb_gpu_count(void);
/* Copy every array to device `dev`; the FM occ array (reference src/bwt.c:68-104 layout) is re-blocked for
 * 128-bit loads by a kernel on the device.  Replaces the per-process load of reference src/idx.c:1103-1160. */
int  dsb_index_upload(const dsb_host_index *h, int dev, dsb_index **out);
/* A replica of `src` on device `dev`, copied GPU to GPU (peer copy over NVLink when available): how the second
 * and later GPUs of a box get the index. */
int  dsb_index_clone(const dsb_index *src, int dev, dsb_index **out);
void dsb_index_free(dsb_index *ix);
/* bytes resident on the device for this index */
uint64_t dsb_index_device_bytes(const dsb_index *ix);

/* ---- read batches ------------------------------------------------------------------------ */

#define DSB_TAIL_BYTES 64   /* bytes after the reverse strand the reference may read (src/cly.c:2365-2367) */
#define DSB_HEAD_BYTES 64   /* bytes before the forward strand the reference may read (src/cly.c:2414-2416) */

/* One read of a batch.  `seq_off` indexes the batch's ASCII base buffer. */
typedef struct dsb_read_desc {
    uint64_t seq_off;
    uint32_t len;
    uint32_t flags;         /* bit0: the reference worker's max_read_l, after this read, is < 510 ("2G reads" filter,
                             * src/cly.c:2953-2963): a running maximum over the EARLIER READS THAT HAD A CHAIN and this one;
                             * evaluated on the host in read order from the DSB_RES_HAD_CHAIN bits of the reads before */
} dsb_read_desc;
#define DSB_READ_2G 1u

/* One reported chain = the fields of the reference's chain_item that output uses
 * (reference src/cly.h:69-89, src/cly_mt.c:229-327). */
typedef struct dsb_hit {
    uint32_t ref_ID;
    uint32_t t_st, t_ed;
    uint32_t q_st, q_ed;
    uint32_t sum_score;
    uint32_t indel;
    uint32_t anchor_number;
    uint8_t  direction;     /* 1 forward, 0 reverse (src/lib/utils.h:66-67) */
    uint8_t  primary;       /* 1 PRI, 2 SEC, 3 SUP (src/cly.h:65-67) */
    uint8_t  pri_index;
    uint8_t  with_top_anchor;
} dsb_hit;

/* Per-read result header. */
typedef struct dsb_read_result {
    uint32_t hit_off;       /* first hit of this read in the batch's hit array */
    uint32_t n_hit;
    uint32_t fast_classify; /* bit 0 (DSB_RES_FAST): cly_r.fast_classify (src/cly.h:97); bit 1 (DSB_RES_HAD_CHAIN): the read
                             * entered delete_small_score_rst with >= 1 chain, i.e. it raised the reference worker's
                             * max_read_l (src/cly.c:2884-2885,2953) */
    uint32_t n_anchor;      /* cly_r.anchor_v.n */
} dsb_read_result;

#define DSB_RES_FAST      1u
#define DSB_RES_HAD_CHAIN 2u

/* Work counters, summed over the batch: the algorithmic units of SURVEY.md 8(d). */
typedef struct dsb_counters {
    uint64_t exist_probes0, exist_probes1; /* first / second exist-table probes */
    uint64_t occ_queries;                  /* FM rank queries */
    uint64_t lut_lookups;                  /* 13-mer LUT pairs */
    uint64_t sa_lookups;
    uint64_t unitig_reads, refpos_reads;
    uint64_t ref_window_bytes;             /* bytes of 2-bit reference fetched */
    uint64_t mem_searches, map_seeds, sdp_windows, anchors, slow_reads;
    /* SM clock cycles the classify kernel's warps spent per phase (profiling aid; lane 0 of each warp):
     * 0 islands, 1 fast seeding, 2 chaining, 3 slow seeding, 4 read 9-mer table, 5 chain scoring,
     * 6 filters + primary, 7 result write-back */
    uint64_t phase_cycles[8];
    /* k_seed's own share of the counts above (the FM-search kernel of the split pipeline): its rank queries, and its
     * algorithmic bytes = 64 x rank queries + 16 x LUT pairs + 8 x (SA + unitig + REF_POS reads) + reference-window bytes */
    uint64_t seed_occ_queries, seed_alg_bytes;
    /* strands of slow-path reads that went through the warp-serial seeding form because a lane of the lane-per-island form
     * would have run out of MEM slots or anchor staging (normally 0) */
    uint64_t slow_serial_strands;
} dsb_counters;

typedef struct dsb_slot dsb_slot;  /* one CUDA stream + pinned staging + device scratch */

/* Create a slot on the index's device.  max_bases / max_reads bound one batch. */
int  dsb_slot_create(dsb_index *ix, uint64_t max_bases, uint32_t max_reads, dsb_slot **out);
void dsb_slot_free(dsb_slot *s);
/* Pinned host staging the caller fills before submit (double-buffer by using two slots):
 * seq: max_bases bytes of ASCII; desc: max_reads descriptors;
 * edge: max_reads x (DSB_HEAD_BYTES + DSB_TAIL_BYTES) bytes = what the reference's worker buffer
 * holds before F[0] and after R[L-1] for that read (history model). */
int  dsb_slot_staging(dsb_slot *s, char **seq, dsb_read_desc **desc, uint8_t **edge);
/* Async: H2D of the staged batch, all kernels, D2H of results, on the slot's stream. */
int  dsb_slot_submit(dsb_slot *s, uint32_t n_reads, uint64_t n_bases);
/* Block until the slot's batch is done.  Returned pointers are pinned host memory owned by the
 * slot, valid until the next submit on it. */
int  dsb_slot_wait(dsb_slot *s, const dsb_read_result **results, const dsb_hit **hits,
                   uint32_t *n_hits, dsb_counters *counters);
/* Device time of the slot's last batch in ms: whole batch and the classify kernel alone. */
int  dsb_slot_timing(dsb_slot *s, float *ms_total, float *ms_exist, float *ms_classify);
/* The classify stage is three launches (k_islands, k_seed, k_classify) unless DSB_SPLIT=0 selected the single-kernel
 * form when the slot was created: their device times of the last batch.  Returns 1 in the split form, else 0
 * (then ms_score = the whole stage). */
int  dsb_slot_timing_stages(dsb_slot *s, float *ms_islands, float *ms_seed, float *ms_score);

/* Device-resident variants used by bench.py's `value` leg (inputs already in HBM):
 * re-run the kernels of the last submitted batch without any host<->device copy. */
int  dsb_slot_rerun_device(dsb_slot *s);

/* FM rank queries in bulk (one thread per query): out[i] = occ(row[i], c[i]) exactly as the reference's
 * occ() (src/bwt.c:43-65) returns it, c_out[i] = the symbol used (c[i] == 0xff asks for the symbol at
 * the row).  Host pointers.  `repeat` >= 1 re-runs the kernel that many times on the device-resident
 * queries and *ms returns the mean kernel time: this is the FM-search roofline probe of bench.py and the
 * known-answer entry of the parity tests. */
int  dsb_fm_rank_batch(dsb_index *ix, const uint64_t *row, const uint8_t *c, uint64_t n, uint64_t *out,
                       uint8_t *c_out, int repeat, float *ms);
/* Random-access roofline probe (measurement aid, no index needed): gather n_access aligned pieces of
 * chunk_bytes (32, 64 or 128) from pseudo-random places of a table_bytes device buffer; *ms = mean kernel time.
 * n_access * chunk_bytes / time is the ceiling the FM-search kernel is reported against (bench.py). */
int  dsb_random_access_probe(int dev, uint64_t table_bytes, uint32_t chunk_bytes, uint64_t n_access, int repeat, float *ms);
/* Exist-k-mer probes in bulk for 2-bit coded strings (test / roofline entry): exist[i] for every
 * k-mer start of codes[0..n), same rule as k_exist.  Host pointers. */
int  dsb_exist_batch(dsb_index *ix, const uint8_t *codes, uint64_t n, uint8_t *exist, int repeat, float *ms);

/* ---- index build --------------------------------------------------------------- */

#define DSB_BUILD_MAX_PHASES 32
typedef struct dsb_build_stats {
    uint64_t n_ref, n_bases;         /* reference sequences, bases */
    uint64_t n_kmer;                 /* distinct 31-mers (what kmer.srt holds, reference src/idx_sort.c:196-200) */
    uint64_t n_unitig;               /* unitigs, without the sentinel */
    uint64_t n_rows;                 /* BWT rows = bases of unitig text incl. terminators */
    uint64_t n_ref_pos;              /* REF_POS entries */
    uint64_t exist_table_bytes;      /* bytes of each of the two exist-k-mer tables */
    int32_t  exist_k;
    int32_t  kernel_launches;
    uint64_t peak_device_bytes;
    double   seconds;
    int32_t  n_phases;
    char     phase_name[DSB_BUILD_MAX_PHASES][40];
    double   phase_s[DSB_BUILD_MAX_PHASES];
} dsb_build_stats;

/* `deSAMBA index` on device `dev`: reference FASTA (plain or gzip) -> the ten deSAMBA.* files in out_dir, byte for byte
 * what the reference builder writes (reference src/idx.c:1255-1282 build_index_main and everything below it; the
 * unspecified bytes -- heap garbage after each reference name in .ref_i -- are zero).  kmer_srt may be NULL: the 31-mer
 * set is derived from the FASTA on the device; when given, it must be that same set (reference src/idx_sort.c:196-200).
 * Returns 0, or -1 with dsb_build_last_error(). */
int dsb_index_build(const char *ref_fasta, const char *kmer_srt, const char *out_dir, int dev, dsb_build_stats *stats);
const char *dsb_build_last_error(void);

/* Exist-k-mer probes in bulk with NO allocation of any kind (no cudaMalloc / cudaFree / allocating library call): the
 * same results as dsb_exist_batch for codes[0..n) -> exist[0..n-k+1) (host pointers), computed by a variant of its kernel
 * whose device-side working buffers are one file-scope `static __device__` array inside libdsb_cuda.so.
 * This entry point has NO COUNTERPART IN THE REFERENCE: it was added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Limit: n <= DSB_EXIST_STATIC_MAX codes per call (larger n is rejected with -1 and a dsb_last_error() that says so);
 * the array (2 x (DSB_EXIST_STATIC_MAX + 256) bytes = 8 MiB) is held by every GPU that loads the library until exit.
 * All calls share the array: calls that overlap on one GPU race (calls of one process are serialised by a host mutex;
 * processes sharing a GPU each have their own copy of the array). */
#define DSB_EXIST_STATIC_MAX 4194304u   /* 4 Mi codes */
int  dsb_exist_batch_static(dsb_index *ix, const uint8_t *codes, uint64_t n, uint8_t *exist);

const char *dsb_last_error(void);
int  dsb_kernel_launch_count(void);   /* kernels launched by this library so far */

#ifdef __cplusplus
}
#endif
#endif
