/* desamba.h -- public C API of libdesamba.so (B200 build).
 *
 * Drop-in for the reference header of the same name: identical symbols, signatures and
 * ownership rules (reference desamba.h:10, :23, :41-43; implementations at reference
 * src/cly_mt.c:1238, :1309, :1329).  A caller that dlopen()s the reference library and
 * dlsym()s these three names (reference main_test.c:29-32) works unchanged.
 *
 * Differences that a caller can observe (documented in INTEGRATION.md):
 *   - the classification itself runs on the B200(s) visible to the process; there is no CPU
 *     fallback: if no CUDA device or the CUDA layer is missing the process aborts with a
 *     message, exactly like every other fatal error of the reference (xassert -> abort);
 *   - `thread_num` is accepted and ignored for results: the reference is only deterministic
 *     for thread_num == 1; this library always returns the thread_num == 1
 *     answer.
 */
#ifndef DESAMBA_H
#define DESAMBA_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Load the index in `dirPath` (deSAMBA.* files, nodes.dmp, names.dmp), upload it to every
 * visible GPU, return an opaque handle through *idx.  (reference desamba.h:10) */
void load_index(void **idx, const char *dirPath);

/* Classify FASTQ/FASTA text (`input_n` bytes) or, when input_n == (uint64_t)-1, the file named by
 * `input` (plain or gzip).  *output is a malloc()ed NUL-terminated SAM_FULL text of *output_n
 * bytes that the caller free()s.  input_n == 0 sets *output_n = 0 and leaves *output alone.
 * (reference desamba.h:23, src/cly_mt.c:1309-1316) */
void read_classify(void *idx, char *input, uint64_t input_n, char **output, uint64_t *output_n,
                   int thread_id, int thread_num);

#define META_USE_READ_NUM 0
#define META_USE_BASE_NUM 1
/* Per-taxon tally of a read_classify() result.  (reference desamba.h:43, src/cly_mt.c:1329) */
void meta_analysis(void *idx, char *input, uint64_t input_n, char **output, uint64_t *output_n,
                   int thread_id, int flag, uint64_t max_snapshot_len, char **human_snapshot,
                   uint64_t *human_snapshot_n);

#ifdef __cplusplus
}
#endif
#endif
