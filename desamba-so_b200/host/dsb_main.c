/* dsb_main.c -- `deSAMBA classify` and `deSAMBA index` command lines, flag-compatible with the reference
 * (reference src/main.c:35-53 dispatcher, src/cly_mt.c:431-562 classify_main / usage, src/idx.c:1239-1282
 * build_index_main).  `index` runs on the GPU and needs no Jellyfish / kmersort step: the 31-mer set is derived from the
 * FASTA itself; a kmer.srt given in the reference's argument position is checked against it.
 * The analysis tools are outside the hot path and remain the reference binary's job.
 */
#include "dsb_host.h"
#include <getopt.h>
#include <stdlib.h>
#include <string.h>

static void classify_usage(void)
{
    fprintf(stderr, "\n");
    fprintf(stderr, "Program:   deSAMBA (B200 build)\n");
    fprintf(stderr, "  Usage:     deSAMBA  classify  [Options] <IndexDir> [ReadFiles.fa][...]>\n");
    fprintf(stderr, "  Basic:   \n");
    fprintf(stderr, "    <IndexDir>      FOLDER   the directory contains deSAMBA index\n");
    fprintf(stderr, "    [ReadFiles.fa]  FILES    reads files, FASTQ(A) format, separated by space\n");
    fprintf(stderr, "  Options:\n");
    fprintf(stderr, "    -h,             help\n");
    fprintf(stderr, "    -t, INT         number of threads[4] (accepted for compatibility; the GPUs do the work)\n");
    fprintf(stderr, "    -l, INT         minimum matching length, ignored for NGS reads [170]\n");
    fprintf(stderr, "    -r, INT         max Output number of secondary alignments[5]\n");
    fprintf(stderr, "    -o, FILE        output results into file [stdout]\n");
    fprintf(stderr, "    -s, INT         MIN score[64]\n");
    fprintf(stderr, "    -f, STR         output format, one of:\n");
    fprintf(stderr, "                    - SAM: SAM-like results without SEQ and QUAL and header, default\n");
    fprintf(stderr, "                    - SAM_FULL: SAM-like results with SEQ and QUAL\n");
    fprintf(stderr, "                    - DES: smallest format\n");
    fprintf(stderr, "                    - DES_FULL: all results are showed, ignore '-r' opinion\n");
    fprintf(stderr, "    -g, INT         number of GPUs to use [all visible]\n");
    fprintf(stderr, "\n");
}

static void sink_file(void *ctx, const char *text, size_t n)
{
    if (fwrite(text, 1, n, (FILE *)ctx) != n) dsb_fatal("short write on output");
}

static int classify_main(int argc, char *argv[])
{
    int c, L_min_matching = 170, max_sec_N = 5, out_format = DSB_OUT_SAM, min_score = 64, n_gpu = 0;
    FILE *outfile = stdout;
    while ((c = getopt(argc, argv, "ht:l:r:f:o:s:g:")) >= 0) {
        if (c == 'h') { classify_usage(); return 0; }
        else if (c == 't') { /* accepted and ignored: results are those of -t 1 */ }
        else if (c == 'l') L_min_matching = atoi(optarg);
        else if (c == 'r') max_sec_N = atoi(optarg);
        else if (c == 'o') { outfile = fopen(optarg, "w"); if (!outfile) dsb_fatal("cannot open %s", optarg); }
        else if (c == 's') min_score = atoi(optarg);
        else if (c == 'g') n_gpu = atoi(optarg);
        else if (c == 'f') {
            if (strcmp(optarg, "SAM") == 0) out_format = DSB_OUT_SAM;
            else if (strcmp(optarg, "SAM_FULL") == 0) out_format = DSB_OUT_SAM_FULL;
            else if (strcmp(optarg, "DES") == 0) out_format = DSB_OUT_DES;
            else if (strcmp(optarg, "DES_FULL") == 0) out_format = DSB_OUT_DES_FULL;
        }
    }
    if (optind + 2 > argc) { classify_usage(); return 0; }
    const char *index_dir = argv[optind++];
    fprintf(stderr, "loading index\t");
    double t0 = dsb_now();
    dsb_loaded_index ix;
    dsb_index_load(&ix, index_dir, L_min_matching, min_score);
    dsb_engine *e = dsb_engine_create(&ix, n_gpu);
    dsb_index_release_bulk(&ix);
    fprintf(stderr, "index resident on %d GPU(s) in %.3fs\n", dsb_engine_n_gpu(e), dsb_now() - t0);
    fprintf(stderr, "Start classify\n");
    dsb_session *s = dsb_session_create(e);
    dsb_out_opt o = {out_format, max_sec_N};
    uint64_t total_reads = 0, total_bases = 0;
    double t1 = dsb_now();
    for (; optind < argc; optind++) {
        fprintf(stderr, "Processing file: [%s].\n", argv[optind]);
        dsb_fq *f = dsb_fq_open_path(argv[optind]);
        dsb_run_stats st;
        dsb_session_run(s, f, &o, sink_file, outfile, &st);
        dsb_fq_close(f);
        total_reads += st.n_reads;
        total_bases += st.n_bases;
        fprintf(stderr, "  batches %llu  parse %.3fs  gpu-wait %.3fs  format %.3fs  gpu-time %.1f ms (exist %.1f, classify %.1f)\n",
                (unsigned long long)st.n_batches, st.t_parse, st.t_gpu_wait, st.t_format, st.gpu_ms_total, st.gpu_ms_exist,
                st.gpu_ms_classify);
        fprintf(stderr, "  slow-path reads %llu (strands through the serial form: %llu)\n", (unsigned long long)st.counters.slow_reads,
                (unsigned long long)st.counters.slow_serial_strands);
    }
    double sec = dsb_now() - t1;
    fflush(outfile);
    fprintf(stderr, "%llu sequences processed in %.3fs (%.1f Kseq/m), %.2f Mbp/s.\n", (unsigned long long)total_reads, sec,
            total_reads / 1.0e3 / (sec / 60), total_bases / 1.0e6 / sec);
    dsb_session_destroy(s);
    dsb_engine_destroy(e);
    dsb_index_destroy(&ix);
    if (outfile != stdout) fclose(outfile);
    return 0;
}

static int index_usage(void)
{
    fprintf(stderr, "\nProgram:   deSAMBA (B200 build)\n");
    fprintf(stderr, "  Usage:     deSAMBA  index  [Options] [SortedKmer] <Reference> <IndexDir>\n");
    fprintf(stderr, "  Basic:     \n");
    fprintf(stderr, "    [SortedKmer]  FILE   optional: \"kmer.srt\" of the reference's kmersort; checked against the FASTA's own 31-mers\n");
    fprintf(stderr, "    <Reference>   FILE   one fasta REF file (plain or gzip), multiple files need to be combined\n");
    fprintf(stderr, "    <IndexDir>    FOLDER the directory to store deSAMBA index\n");
    fprintf(stderr, "  Options:\n    -h            help\n    -g INT        GPU to build on [0]\n\n");
    return 0;
}

/* reference src/idx.c:1255-1282: `index [SortedKmer] [Reference] [IndexDir]` */
static int index_main(int argc, char *argv[])
{
    int c, dev = 0;
    while ((c = getopt(argc, argv, "k:hg:")) >= 0) {
        if (c == 'h') return index_usage();
        else if (c == 'g') dev = atoi(optarg);
    }
    if (optind + 2 > argc) {
        fprintf(stderr, "Warning: [Reference] or [IndexDir] missed.\n");
        return index_usage();
    }
    const char *kmer = NULL;
    if (optind + 3 <= argc) kmer = argv[optind++];
    const char *ref = argv[optind++], *dir = argv[optind++];
    dsb_build_stats st;
    if (dsb_index_build(ref, kmer, dir, dev, &st) != 0) dsb_fatal("index: %s", dsb_build_last_error());
    fprintf(stderr, "index built in %.2fs: %llu references, %llu bases, %llu distinct 31-mers, %llu unitigs, %llu BWT rows, "
            "exist k = %d (2 x %llu MiB); %d kernel launches, %.2f GB peak on the device\n", st.seconds,
            (unsigned long long)st.n_ref, (unsigned long long)st.n_bases, (unsigned long long)st.n_kmer,
            (unsigned long long)st.n_unitig, (unsigned long long)st.n_rows, st.exist_k,
            (unsigned long long)(st.exist_table_bytes >> 20), st.kernel_launches, st.peak_device_bytes / 1e9);
    for (int i = 0; i < st.n_phases; i++) fprintf(stderr, "  %-34s %8.3f s\n", st.phase_name[i], st.phase_s[i]);
    return 0;
}

static int usage(void)
{
    fprintf(stderr, "\nProgram: deSAMBA (B200 build)\n\n  Usage  : deSAMBA <command> \n  Command: \n");
    fprintf(stderr, "		classify:	classify meta-genomic reads (runs on the GPUs)\n");
    fprintf(stderr, "		index:		build the index from a FASTA (runs on the GPU; no kmersort step needed)\n");
    fprintf(stderr, "		analysis:	not part of the B200 hot path; use the reference binary\n\n");
    return 1;
}

int main(int argc, char *argv[])
{
    if (argc <= 1) return usage();
    if (strcmp(argv[1], "classify") == 0) return classify_main(argc - 1, argv + 1);
    if (strcmp(argv[1], "index") == 0) return index_main(argc - 1, argv + 1);
    return usage();
}
