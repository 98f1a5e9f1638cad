/* dbg_anchors.c -- TEST INFRASTRUCTURE ONLY (debug aid).  Our own file; #includes the reference's
 * src/cly.c in place (like kat_harness.c) and prints, for every read of a FASTQ, the anchors and the hits the
 * reference's classify_seq leaves behind, so that a parity difference can be located by stage.
 *   dbg_anchors <index_dir> <reads.fq>
 */
#include REF_CLY_C
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
extern void calculate_MAPQ_TABLE(int *Q_MEM, int (*Q_LV)[MAX_LV_R_LEN], double P_E, uint64_t L_REF);

int main(int argc, char **argv)
{
    if (argc < 3) return 1;
    DA_IDX *idx = (DA_IDX *)calloc(1, sizeof(DA_IDX));
    load_idx(idx, argv[1]);
    idx->filter_min_length = 170; idx->filter_min_score = 64; idx->filter_min_score_LV3 = 74;
    idx->mapQ.Q_MEM = malloc(sizeof(int) * Q_MEM_MAX);
    idx->mapQ.Q_LV = (int(*)[MAX_LV_R_LEN])malloc(MAX_LV_WRONG * MAX_LV_R_LEN * sizeof(int));
    calculate_MAPQ_TABLE(idx->mapQ.Q_MEM, idx->mapQ.Q_LV, 0.15, idx->ref_bin.n * 4);
    Classify_buff_pool buff;
    memset(&buff, 0, sizeof buff);
    buff.sa_hash[0] = malloc(sizeof(sparse_align_HASH) * 0x100000);
    buff.sa_hash[1] = malloc(sizeof(sparse_align_HASH) * 0x100000);
    FILE *fq = fopen(argv[2], "r");
    char *name = NULL, *seq = NULL, *tmp = NULL;
    size_t c1 = 0, c2 = 0, c3 = 0;
    while (getline(&name, &c1, fq) > 0 && getline(&seq, &c2, fq) > 0) {
        getline(&tmp, &c3, fq); getline(&tmp, &c3, fq);
        uint32_t L = (uint32_t)strlen(seq);
        while (L && (seq[L - 1] == '\n' || seq[L - 1] == '\r')) seq[--L] = 0;
        kseq_t ks; memset(&ks, 0, sizeof ks);
        ks.seq.s = seq; ks.seq.l = L; ks.name.s = name; ks.name.l = strlen(name);
        cly_r res; memset(&res, 0, sizeof res);
        classify_seq(&ks, idx, &res, &buff);
        printf("READ %s L=%u anchors=%d hits=%d fast=%d\n", name, L, (int)res.anchor_v.n, (int)res.hit.n, (int)res.fast_classify);
        for (size_t i = 0; i < res.anchor_v.n; i++) {
            Anchor *a = res.anchor_v.a + i;
            printf("  A %zu ref=%u dir=%d q=%u t=%u len=%u score=%d useless=%d lL=%d lE=%d rL=%d rE=%d\n", i, a->ref_ID, a->direction,
                   a->index_in_read, a->ref_offset, a->a_m.mtch_len, a->a_m.score, a->anchor_useless, a->a_m.left_len, a->a_m.left_ED,
                   a->a_m.rigt_len, a->a_m.rigt_ED);
        }
        for (size_t i = 0; i < res.hit.n; i++) {
            chain_item *c = res.hit.a + i;
            printf("  H %zu ref=%u dir=%d q=%u-%u t=%u-%u score=%u an=%u\n", i, c->ref_ID, c->direction, c->q_st, c->q_ed, c->t_st, c->t_ed,
                   c->sum_score, c->anchor_number);
        }
    }
    return 0;
}
