/* dsb_meta.c -- meta_analysis (reference src/cly_mt.c:1083-1235, :1329-1414): SURVEY.md 8(f)
 * rank 1, a "next" row.  Host-side text processing, no GPU involvement.
 * PLACEHOLDER until the hot path meets its parity bar: loading is a no-op and the call aborts
 * loudly instead of returning something that looks like an answer.
 */
#include "dsb_host.h"
#include <stdlib.h>

typedef struct dsb_taxonomy { int unused; } dsb_taxonomy;

dsb_taxonomy *dsb_taxonomy_load(const char *dir)
{
    (void)dir;
    return NULL;
}

void dsb_meta_analysis_impl(dsb_taxonomy *tax, const char *input, uint64_t input_n, char **output, uint64_t *output_n,
                            int flag, uint64_t max_snapshot_len, char **human_snapshot, uint64_t *human_snapshot_n)
{
    (void)tax; (void)input; (void)input_n; (void)output; (void)output_n; (void)flag; (void)max_snapshot_len;
    (void)human_snapshot; (void)human_snapshot_n;
    dsb_fatal("meta_analysis is not implemented in this build yet");
}
