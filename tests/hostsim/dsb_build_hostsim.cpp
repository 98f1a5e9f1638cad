/* TEST INFRASTRUCTURE ONLY: the index builder's source (desamba-so_b200/csrc/dsb_build.cu) compiled for the host, with
 * the data-parallel primitives of dsb_px.cuh as plain loops, so that its output can be compared byte for byte with the
 * reference's `deSAMBA index` in a container without a GPU.  Never linked into the product. */
#define DSB_BUILD_HOSTSIM 1
#include "../../desamba-so_b200/csrc/dsb_build.cu"

#ifdef DSB_BUILD_MAIN
int main(int argc, char **argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s ref.fa out_dir [kmer.srt]\n", argv[0]); return 2; }
    dsb_build_stats st;
    if (dsb_index_build(argv[1], argc > 3 ? argv[3] : NULL, argv[2], 0, &st) != 0) {
        fprintf(stderr, "build failed: %s\n", dsb_build_last_error());
        return 1;
    }
    fprintf(stderr, "refs %llu bases %llu kmers %llu unitigs %llu rows %llu ref_pos %llu exist k=%d %.2fs\n",
            (unsigned long long)st.n_ref, (unsigned long long)st.n_bases, (unsigned long long)st.n_kmer,
            (unsigned long long)st.n_unitig, (unsigned long long)st.n_rows, (unsigned long long)st.n_ref_pos, st.exist_k, st.seconds);
    return 0;
}
#endif
