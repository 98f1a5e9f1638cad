/* dsb_build.cu -- `deSAMBA index` on the B200: reference FASTA -> the ten deSAMBA.* index files, byte for byte what the
 * reference's builder writes (reference src/idx.c:884-963 build_UNITIG, :125-239 build_deb, :386-540 setLabel,
 * :768-854 get_uni_v, :554-706 set_ref_lists, :345-389 merge_kmer, :986-1026 get_EXIST_kmer, :1046-1101 write_idx,
 * :1163-1237 bwt_cal_SA; src/bwt.c:106-276 build_BWT / write_bwt).  SURVEY 8(f)-4.
 *
 * NOT a port: the reference walks the de Bruijn graph k-mer by k-mer with binary searches, merge-sorts 30 "special"
 * suffixes per unitig on the host and finds the suffix-array samples by one serial LF walk over the whole BWT.  Here
 * every step is a data-parallel pass over REFERENCE POSITIONS or over sorted K-MERS, because every edge of the graph
 * is a (k+1)-mer of the reference (jellyfish is run without -C on the same FASTA, reference build-index:88):
 *
 *   1. 31-mer of every position (rolling, one thread per 64 positions) -> radix sort -> unique       = kmer.srt
 *   2. position -> k-mer index (13-mer bucket table + binary search), edge bits OR-ed per k-mer      = build_deb
 *   3. Start / End labels per k-mer from its own edge counts + one look-up per edge                  = setLabel
 *   4. unitig ids = prefix sum of Start flags over the sorted k-mers (the reference numbers unitigs by start k-mer)
 *   5. a unitig's occurrence in a reference is the stretch from a Start position to the next End position: one
 *      max-scan over positions gives every position its unitig and offset -> lengths, end k-mers, REF_POS lists
 *   6. the 30 short suffixes of each unitig end: radix sort on (A-padded suffix) with (length, unitig) as the stable
 *      tie order == the reference's comparator (src/idx.c:857-881); merged with the k-mers by counting
 *   7. BWT symbol per row, 13-mer prefix table, rank checkpoints, 4-bit packing
 *   8. suffix-array samples scattered directly from (unitig, offset) of each row -- no LF walk
 *   9. exist-k-mer tables from the positions (every k-mer window inside a run of >= 31 ACGT)
 *
 * Limits (checked, reported through dsb_last_error): < 2^32 distinct 31-mers, < 2^32 / 30 unitigs.
 * `kmer.srt` (reference src/idx_sort.c:196-200) is optional: when given it must equal the FASTA's own 31-mer set.
 */
#include "dsb_px.cuh"
#include "dsb_cuda.h"
#include <errno.h>
#include <fcntl.h>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>
#include <vector>
#include <zlib.h>

namespace px {
int g_launches = 0;
char g_error[512] = "";
size_t g_live_bytes = 0, g_peak_bytes = 0;
}

namespace dsbb {

static const int KB = 31;                              /* B_KMER, reference src/idx.c:8 */
static const uint64_t KMASK = ((uint64_t)1 << 62) - 1;
static const uint64_t SENT = (uint64_t)1 << 62;       /* "no k-mer here": sorts after every k-mer */
static const uint32_t NONE32 = 0xffffffffu;
static const uint64_t NONE64 = ~(uint64_t)0;
/* per-k-mer info word: bits 0-3 out edges, 4-7 in edges, 8 End, 9 Start (reference src/idx.c:51-58); ours: 13 head of a
 * run, 14 tail of a run (the reference keeps those as the `heads` / `tails` lists, src/idx.c:163,208) */
static const uint32_t I_END = 1u << 8, I_START = 1u << 9, I_HEAD = 1u << 13, I_TAIL = 1u << 14;
static const int LUT_BITS = 26;                        /* L_PRE_IDX = 13 bases, reference src/idx.h:58 */

static double now()
{
    struct timeval tv;
    gettimeofday(&tv, NULL);
    return tv.tv_sec + 1e-6 * tv.tv_usec;
}

/* ------------------------------------------------------------------ FASTA (host) */
struct RefSet {
    std::vector<std::string> name;
    std::vector<uint64_t> len, off;
    std::vector<char> seq;        /* all sequences back to back, ASCII */
};

/* Same record grammar as kseq (reference src/lib/utils.c:939-977): a record opens at a line starting with '>' or '@',
 * the name ends at the first white space, sequence lines run until a line starting with '>', '+' or '@'; after '+' as
 * many quality characters as there were bases are skipped. */
static void parse_fasta(const char *buf, size_t n, RefSet &rs)
{
    size_t p = 0;
    while (p < n && buf[p] != '>' && buf[p] != '@') {   /* skip to the first header */
        const char *nl = (const char *)memchr(buf + p, '\n', n - p);
        if (!nl) return;
        p = (size_t)(nl - buf) + 1;
    }
    while (p < n) {
        /* header line */
        size_t e = p + 1;
        while (e < n && buf[e] != '\n' && buf[e] != ' ' && buf[e] != '\t' && buf[e] != '\r') e++;
        rs.name.emplace_back(buf + p + 1, e - p - 1);
        const char *nl = (const char *)memchr(buf + e, '\n', n - e);
        p = nl ? (size_t)(nl - buf) + 1 : n;
        uint64_t start = rs.seq.size();
        int c = -1;
        while (p < n) {
            c = (unsigned char)buf[p];
            if (c == '>' || c == '+' || c == '@') break;
            nl = (const char *)memchr(buf + p, '\n', n - p);
            size_t le = nl ? (size_t)(nl - buf) : n;
            size_t ll = le - p;
            if (ll && buf[le - 1] == '\r') ll--;
            rs.seq.insert(rs.seq.end(), buf + p, buf + p + ll);
            p = nl ? le + 1 : n;
            c = -1;
        }
        uint64_t L = rs.seq.size() - start;
        rs.off.push_back(start);
        rs.len.push_back(L);
        if (c == '+') {                                   /* FASTQ: skip the '+' line and L quality characters */
            nl = (const char *)memchr(buf + p, '\n', n - p);
            p = nl ? (size_t)(nl - buf) + 1 : n;
            uint64_t q = 0;
            while (p < n && q < L) {
                nl = (const char *)memchr(buf + p, '\n', n - p);
                size_t le = nl ? (size_t)(nl - buf) : n;
                size_t ll = le - p;
                if (ll && buf[le - 1] == '\r') ll--;
                q += ll;
                p = nl ? le + 1 : n;
            }
            while (p < n && buf[p] != '>' && buf[p] != '@') {
                nl = (const char *)memchr(buf + p, '\n', n - p);
                p = nl ? (size_t)(nl - buf) + 1 : n;
            }
        }
    }
}

static bool load_refset(const char *path, RefSet &rs)
{
    int fd = open(path, O_RDONLY);
    if (fd < 0) { snprintf(px::g_error, sizeof px::g_error, "cannot open %s: %s", path, strerror(errno)); return false; }
    struct stat st;
    fstat(fd, &st);
    unsigned char magic[2] = {0, 0};
    if (pread(fd, magic, 2, 0) == 2 && magic[0] == 0x1f && magic[1] == 0x8b) {
        close(fd);
        gzFile gz = gzopen(path, "rb");
        if (!gz) { snprintf(px::g_error, sizeof px::g_error, "cannot gzopen %s", path); return false; }
        std::vector<char> all;
        std::vector<char> piece(1 << 24);
        int got;
        while ((got = gzread(gz, piece.data(), (unsigned)piece.size())) > 0) all.insert(all.end(), piece.begin(), piece.begin() + got);
        gzclose(gz);
        rs.seq.reserve(all.size());
        parse_fasta(all.data(), all.size(), rs);
        return true;
    }
    if (st.st_size == 0) { close(fd); return true; }
    void *m = mmap(NULL, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (m == MAP_FAILED) { snprintf(px::g_error, sizeof px::g_error, "cannot map %s", path); return false; }
    madvise(m, (size_t)st.st_size, MADV_SEQUENTIAL);
    rs.seq.reserve((size_t)st.st_size);
    parse_fasta((const char *)m, (size_t)st.st_size, rs);
    munmap(m, (size_t)st.st_size);
    return true;
}

/* ------------------------------------------------------------------ small device helpers */
/* `Bit` of the reference (src/idx.c:9-27): ACGT in either case -> 0..3, anything else 4 */
PX_HD uint32_t base_code(uint8_t c)
{
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 4;
    }
}
PX_HD int popc4(uint32_t v) { return (int)((v & 1) + ((v >> 1) & 1) + ((v >> 2) & 1) + ((v >> 3) & 1)); }
PX_HD int low_bit4(uint32_t v) { return (v & 1) ? 0 : (v & 2) ? 1 : (v & 4) ? 2 : 3; }

/* first index in kv[counter[p] .. counter[p+1]) (p = 13-mer prefix of key) whose k-mer is >= key: the reference's
 * bucketed binary search (src/idx.c:74-88, getCounter :98-115) as a lower bound */
PX_HD uint64_t kmer_lower_bound(const uint64_t *kv, const uint64_t *counter, uint64_t key)
{
    uint64_t p = key >> (2 * (KB - 13));
    uint64_t lo = counter[p], hi = counter[p + 1];
    while (lo < hi) {
        uint64_t mid = lo + ((hi - lo) >> 1);
        if (kv[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

/* hash64_1 / hash64_2 of the reference (src/lib/utils.c:1067-1091) */
PX_HD uint64_t hash_a(uint64_t key)
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
PX_HD uint64_t hash_b(uint64_t key)
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

template <class P> static uint64_t count_if(uint64_t n, P pred)
{
    px::Buf<uint64_t> c(1);
    px::zero(c.p, 8);
    uint64_t *cp = c.p;
    const uint64_t chunk = 256;
    px::for_each((n + chunk - 1) / chunk, PX_LAMBDA(uint64_t t) {
        uint64_t s = t * chunk, e = s + chunk < n ? s + chunk : n, m = 0;
        for (uint64_t i = s; i < e; i++) m += pred(i) ? 1 : 0;
        if (m) px::atomic_add64(cp, m);
    });
    uint64_t h;
    px::d2h(&h, cp, 8);
    return h;
}

/* ------------------------------------------------------------------ file output */
struct Out {
    std::string dir;
    FILE *open_part(const char *suffix)
    {
        std::string p = dir + "/deSAMBA" + suffix;
        FILE *fp = fopen(p.c_str(), "wb");
        if (!fp) px::fail("cannot create", p.c_str());
        return fp;
    }
};
static void put(FILE *fp, const void *p, size_t n)
{
    const char *c = (const char *)p;
    while (n) {
        size_t w = n > ((size_t)1 << 30) ? ((size_t)1 << 30) : n;
        if (fwrite(c, 1, w, fp) != w) px::fail("write", strerror(errno));
        c += w;
        n -= w;
    }
}
/* device array -> file in pieces (no full-size host copy) */
static void put_device(FILE *fp, const void *d, size_t n)
{
    const size_t piece = (size_t)256 << 20;
    std::vector<char> h(n < piece ? n : piece);
    for (size_t o = 0; o < n; o += piece) {
        size_t w = n - o < piece ? n - o : piece;
        px::d2h(h.data(), (const char *)d + o, w);
        put(fp, h.data(), w);
    }
}

/* exist-table class from the number of distinct 31-mers (reference src/idx.c:986-998, set_ekmer_par :966-982) */
static void ekmer_class(uint64_t n_kmer, uint64_t *bytes, int *k, int *mask_bits)
{
    static const uint64_t size[8] = {0x8000000ull, 0x10000000ull, 0x20000000ull, 0x40000000ull,
                                     0x80000000ull, 0x100000000ull, 0x200000000ull, 0x400000000ull};
    static const int klen[8] = {16, 17, 17, 18, 18, 19, 19, 20};
    int c = 7;
    for (int i = 0; i < 8; i++)
        if (n_kmer < ((uint64_t)1 << (31 + i)) / 9) { c = i; break; }
    *bytes = size[c];
    *k = klen[c];
    *mask_bits = 30 + c;
}

/* ------------------------------------------------------------------ the build */
static void build(const char *ref_path, const char *kmer_path, const char *out_dir, dsb_build_stats *st)
{
    double t0 = now(), t_prev = t0;
    int n_phase = 0;
    auto lap = [&](const char *name) {
        px::sync();
        double t = now();
        if (n_phase < DSB_BUILD_MAX_PHASES) {
            snprintf(st->phase_name[n_phase], sizeof st->phase_name[0], "%s", name);
            st->phase_s[n_phase++] = t - t_prev;
            st->n_phases = n_phase;
        }
        if (getenv("DSB_VERBOSE")) fprintf(stderr, "[dsb-build] %-28s %7.3f s  (device live %.2f GB)\n", name, t - t_prev, px::g_live_bytes / 1e9);
        t_prev = t;
    };

    /* ---- 0. reference on the host, then on the device ---- */
    RefSet rs;
    if (!load_refset(ref_path, rs)) throw px::Fail();
    const uint64_t G = rs.seq.size(), n_ref = rs.name.size();
    if (G < (uint64_t)KB) px::fail("reference", "shorter than one 31-mer");
    lap("read FASTA");

    px::Buf<uint8_t> code(G + 64);           /* bits 0-2: base code (4 = not ACGT), bit 3: first base of a sequence */
    {
        px::Buf<uint8_t> ascii(G);
        px::h2d(ascii.p, rs.seq.data(), G);
        uint8_t *cd = code.p;
        const uint8_t *as = ascii.p;
        px::for_each(G + 64, PX_LAMBDA(uint64_t i) { cd[i] = i < G ? (uint8_t)base_code(as[i]) : (uint8_t)4; });
        px::Buf<uint64_t> offs(n_ref);
        px::h2d(offs.p, rs.off.data(), n_ref * 8);
        const uint64_t *op = offs.p;
        px::for_each(n_ref, PX_LAMBDA(uint64_t r) { if (op[r] < G) cd[op[r]] |= 8; });
        px::sync();
    }
    std::vector<char>().swap(rs.seq);
    const uint8_t *cd = code.p;

    /* 2-bit packed reference, 4 bases per byte, first base in the top bits; everything that is not ACGT packs as A
     * (`bin_Bit`, reference src/idx.c:29-47; set_ref_lists :596-606).  Straight from the codes: 4 -> 0. */
    const uint64_t ref_bin_n = (G + 3) >> 2;
    px::Buf<uint8_t> ref_bin(ref_bin_n);
    {
        uint8_t *rb = ref_bin.p;
        px::for_each(ref_bin_n, PX_LAMBDA(uint64_t b) {
            uint32_t v = 0;
            for (int k = 0; k < 4; k++) {
                uint64_t i = b * 4 + k;
                uint32_t c = i < G ? (cd[i] & 7) : 0;
                v = (v << 2) | (c < 4 ? c : 0);
            }
            rb[b] = (uint8_t)v;
        });
    }
    lap("codes + packed reference");

    /* ---- 1. the 31-mer set ---- */
    const uint64_t CH = 64;                                  /* positions per thread in the rolling passes */
    const uint64_t n_chunk = (G + CH - 1) / CH;
    uint64_t K = 0;
    px::Buf<uint64_t> kmer_v;
    {
        px::Buf<uint64_t> key(G), alt(G);
        uint64_t *kp = key.p;
        px::for_each(n_chunk, PX_LAMBDA(uint64_t c) {
            uint64_t s = c * CH, e = s + CH < G ? s + CH : G;
            uint64_t kmer = 0;
            uint32_t clean = 0;
            for (uint64_t j = s; j < e + (KB - 1) && j < G; j++) {
                uint8_t x = cd[j];
                if (x & 8) clean = 0;                          /* a window may start at a sequence start, not span it */
                if ((x & 7) >= 4) { clean = 0; kmer = 0; }
                else { kmer = ((kmer << 2) | (x & 3)) & KMASK; clean++; }
                if (j >= s + (KB - 1)) kp[j - (KB - 1)] = clean >= (uint32_t)KB ? kmer : SENT;
            }
            uint64_t tail = G >= (uint64_t)(KB - 1) ? G - (KB - 1) : 0;   /* windows that would pass the end */
            for (uint64_t i = s > tail ? s : tail; i < e; i++) kp[i] = SENT;
        });
        uint64_t *k0 = key.p, *k1 = alt.p;
        px::sort_keys(k0, k1, G, 63);
        lap("31-mers of every position, sort");
        uint64_t m = px::unique(k0, G, k1);
        /* the sentinel, if any position was not a k-mer, is the last distinct value */
        uint64_t last;
        px::d2h(&last, k1 + (m - 1), 8);
        K = last == SENT ? m - 1 : m;
        if (K == 0) px::fail("reference", "no 31-mer of A/C/G/T");
        if (K >= (uint64_t)0xfffffff0u) px::fail("index too large for this builder", ">= 2^32 distinct 31-mers");
        kmer_v.alloc(K);
        px::d2d(kmer_v.p, k1, K * 8);
    }
    const uint64_t *kv = kmer_v.p;
    lap("distinct 31-mers");

    if (kmer_path && kmer_path[0]) {     /* optional kmer.srt: must be this very set (reference src/idx_sort.c:196-200) */
        FILE *fp = fopen(kmer_path, "rb");
        if (!fp) px::fail("cannot open", kmer_path);
        uint64_t n = 0;
        if (fread(&n, 8, 1, fp) != 1) px::fail("kmer.srt", "truncated");
        if (n != K) {
            char m[160];
            snprintf(m, sizeof m, "%llu k-mers in the file, %llu distinct 31-mers in the reference (jellyfish must be run on the "
                     "same FASTA without -C)", (unsigned long long)n, (unsigned long long)K);
            px::fail("kmer.srt does not match the reference", m);
        }
        const uint64_t piece = (uint64_t)1 << 24;
        std::vector<uint64_t> h(piece);
        px::Buf<uint64_t> d(piece), bad(1);
        px::zero(bad.p, 8);
        uint64_t *dp = d.p, *bp = bad.p;
        for (uint64_t o = 0; o < K; o += piece) {
            uint64_t w = K - o < piece ? K - o : piece;
            if (fread(h.data(), 8, w, fp) != w) px::fail("kmer.srt", "truncated");
            px::h2d(dp, h.data(), w * 8);
            px::for_each(w, PX_LAMBDA(uint64_t i) { if (dp[i] != kv[o + i]) px::atomic_add64(bp, 1); });
        }
        fclose(fp);
        uint64_t nb;
        px::d2h(&nb, bp, 8);
        if (nb) px::fail("kmer.srt does not match the reference", "same count, different 31-mers");
        lap("kmer.srt verified");
    }

    /* 13-mer bucket table over the sorted k-mers (`getCounter`, reference src/idx.c:98-115) */
    const uint64_t n_lut = ((uint64_t)1 << LUT_BITS) + 1;
    px::Buf<uint64_t> counter(n_lut);
    {
        uint64_t *cp = counter.p;
        px::for_each(n_lut, PX_LAMBDA(uint64_t p) {
            uint64_t key = p << (2 * (KB - 13));
            uint64_t lo = 0, hi = K;
            while (lo < hi) { uint64_t mid = lo + ((hi - lo) >> 1); if (kv[mid] < key) lo = mid + 1; else hi = mid; }
            cp[p] = lo;
        });
    }
    const uint64_t *ct = counter.p;

    /* ---- 2. position -> k-mer index; edges (build_deb, reference src/idx.c:125-239) ---- */
    px::Buf<uint32_t> loc(G + 1);
    px::Buf<uint32_t> kinfo(K);
    px::zero(kinfo.p, K * 4);
    {
        uint32_t *lp = loc.p;
        px::for_each(n_chunk, PX_LAMBDA(uint64_t c) {
            uint64_t s = c * CH, e = s + CH < G ? s + CH : G;
            uint64_t kmer = 0;
            uint32_t clean = 0;
            for (uint64_t j = s; j < e + (KB - 1) && j < G; j++) {
                uint8_t x = cd[j];
                if (x & 8) clean = 0;
                if ((x & 7) >= 4) { clean = 0; kmer = 0; }
                else { kmer = ((kmer << 2) | (x & 3)) & KMASK; clean++; }
                if (j >= s + (KB - 1)) lp[j - (KB - 1)] = clean >= (uint32_t)KB ? (uint32_t)kmer_lower_bound(kv, ct, kmer) : NONE32;
            }
            uint64_t tail = G >= (uint64_t)(KB - 1) ? G - (KB - 1) : 0;
            for (uint64_t i = s > tail ? s : tail; i < e; i++) lp[i] = NONE32;
            if (c == 0) lp[G] = NONE32;
        });
    }
    const uint32_t *lc = loc.p;
    lap("position -> k-mer index");
    {
        uint32_t *ki = kinfo.p;
        px::for_each(G, PX_LAMBDA(uint64_t i) {
            uint32_t j = lc[i];
            if (j == NONE32) return;
            uint32_t bits = 0;
            if (i >= 1 && lc[i - 1] != NONE32) bits |= 1u << (4 + (cd[i - 1] & 3)); else bits |= I_HEAD;
            if (lc[i + 1] != NONE32) bits |= 1u << (cd[i + KB] & 3); else bits |= I_TAIL;
            if ((ki[j] & bits) != bits) px::atomic_or32(ki + j, bits);
        });
    }
    lap("edges");

    /* ---- 3. Start / End labels (setLabel, reference src/idx.c:386-484) ---- */
    {
        uint32_t *ki = kinfo.p;
        px::for_each(K, PX_LAMBDA(uint64_t j) {
            uint32_t info = ki[j];
            uint32_t in = (info >> 4) & 15, out = info & 15;
            uint64_t v = kv[j];
            if (popc4(in) != 1 || (info & I_HEAD)) {
                px::atomic_or32(ki + j, I_START);
                for (uint32_t c = 0; c < 4; c++)
                    if (in & (1u << c)) {
                        uint64_t key = (v >> 2) | ((uint64_t)c << (2 * (KB - 1)));
                        px::atomic_or32(ki + kmer_lower_bound(kv, ct, key), I_END);
                    }
            }
            if (popc4(out) != 1 || (info & I_TAIL)) {
                px::atomic_or32(ki + j, I_END);
                for (uint32_t c = 0; c < 4; c++)
                    if (out & (1u << c)) {
                        uint64_t key = ((v << 2) | c) & KMASK;
                        px::atomic_or32(ki + kmer_lower_bound(kv, ct, key), I_START);
                    }
            }
        });
    }
    const uint32_t *ki = kinfo.p;
    lap("unitig start / end labels");

    /* ---- 4. unitig ids: the reference numbers unitigs by ascending start k-mer (get_uni_v, src/idx.c:768-854) ---- */
    px::Buf<uint32_t> uid(K);
    const uint64_t U = px::exclusive_sum<uint32_t>(K, PX_LAMBDA(uint64_t j) { return (uint64_t)((ki[j] >> 9) & 1); }, uid.p);
    const uint64_t n_end = count_if(K, PX_LAMBDA(uint64_t j) { return (ki[j] & I_END) != 0; });
    if (U == 0 || U != n_end) px::fail("de Bruijn graph", "start and end counts differ (a perfectly circular sequence?)");
    if (U * 30 >= (uint64_t)0xfffffff0u) px::fail("index too large for this builder", ">= 2^32 / 30 unitigs");
    const uint32_t *ud = uid.p;
    {   /* the reference gives '$' to the first start k-mer of its FIRST worker's range [0, K/16) (src/idx.c:737,792-805);
         * without one its own SA walk aborts (src/idx.c:1208,1235) */
        px::Buf<uint64_t> first(1);
        px::fill_ff(first.p, 8);
        uint64_t *fp = first.p;
        px::for_each(K, PX_LAMBDA(uint64_t j) { if ((ki[j] & I_START) && ud[j] == 0) *fp = j; });
        uint64_t j0;
        px::d2h(&j0, fp, 8);
        if (j0 >= K / 16) px::fail("reference too small", "the smallest unitig start k-mer is not in the first 1/16 of the k-mer list "
                                   "(the reference builder aborts on such input too)");
    }
    lap("unitig ids");

    /* ---- 5. unitigs along the references ---- */
    px::Buf<uint32_t> ulen(U);
    px::Buf<uint64_t> uendk(U);
    px::Buf<uint64_t> ref_pos;
    px::Buf<uint64_t> unv(U + 1);
    uint64_t R = 0;
    px::Buf<uint64_t> seg(G);       /* 1 + the last Start position at or before i */
    px::inclusive_max(G, PX_LAMBDA(uint64_t i) {
        uint32_t j = lc[i];
        return (j != NONE32 && (ki[j] & I_START)) ? i + 1 : (uint64_t)0;
    }, seg.p);
    const uint64_t *sg = seg.p;
    {
        uint32_t *ul = ulen.p;
        uint64_t *ue = uendk.p;
        px::for_each(G, PX_LAMBDA(uint64_t i) {
            uint32_t j = lc[i];
            if (j == NONE32 || !(ki[j] & I_END)) return;
            uint64_t s = sg[i] - 1;
            uint32_t u = ud[lc[s]];
            ul[u] = (uint32_t)(i - s) + KB;
            ue[u] = kv[j];
        });
    }
    const uint32_t *ul = ulen.p;
    lap("unitig lengths");
    {   /* REF_UNI list in reference order, then stably by unitig (set_ref_lists, reference src/idx.c:608-667, sort :669-677) */
        px::Buf<uint64_t> rp(G / 4 + 1024), rp_alt;
        /* worst case one entry per 5 positions (a unitig of >= 35 bases advances by >= 5) */
        R = px::select_index(G, PX_LAMBDA(uint64_t i) {
            uint32_t j = lc[i];
            return j != NONE32 && (ki[j] & I_START) && ul[ud[j]] >= 35;       /* MIN_UNI_L, reference src/idx.h:40 */
        }, rp.p);
        if (R >= (uint64_t)0xfffffff0u) px::fail("index too large for this builder", ">= 2^32 unitig occurrences");
        px::Buf<uint32_t> rk(R), rk_alt(R);
        rp_alt.alloc(R);
        uint32_t *rkp = rk.p;
        uint64_t *rpp = rp.p;
        px::for_each(R, PX_LAMBDA(uint64_t r) { rkp[r] = ud[lc[rpp[r]]]; });
        int ubits = 1;
        while (ubits < 32 && ((uint64_t)1 << ubits) < U) ubits++;
        uint32_t *k0 = rk.p, *k1 = rk_alt.p;
        uint64_t *v0 = rp.p, *v1 = rp_alt.p;
        px::sort_pairs(k0, k1, v0, v1, R, ubits);
        /* REF_POS {global_offset:40, ref_ID:23, direction:1 = FORWARD} (reference src/idx.h:42-48, write_idx :1086-1093) */
        ref_pos.alloc(R);
        px::Buf<uint64_t> offs(n_ref);
        px::h2d(offs.p, rs.off.data(), n_ref * 8);
        const uint64_t *op = offs.p;
        uint64_t *out = ref_pos.p;
        /* per unitig: first entry and one past the last (ref_list, reference src/idx.c:679-700) */
        px::Buf<uint32_t> ufirst(U), ulast(U);
        px::fill_ff(ufirst.p, U * 4);
        px::zero(ulast.p, U * 4);
        uint32_t *uf = ufirst.p, *ulst = ulast.p;
        px::for_each(R, PX_LAMBDA(uint64_t r) {
            uint64_t pos = v0[r];
            uint64_t lo = 0, hi = n_ref;                     /* last reference starting at or before pos */
            while (lo + 1 < hi) { uint64_t mid = (lo + hi) >> 1; if (op[mid] <= pos) lo = mid; else hi = mid; }
            out[r] = (pos & (((uint64_t)1 << 40) - 1)) | ((lo & 0x7fffff) << 40) | ((uint64_t)1 << 63);
            uint32_t u = k0[r];
            if (r == 0 || k0[r - 1] != u) uf[u] = (uint32_t)r;
            if (r + 1 == R || k0[r + 1] != u) ulst[u] = (uint32_t)(r + 1);
        });
        uint64_t *uv = unv.p;
        px::for_each(U + 1, PX_LAMBDA(uint64_t u) {
            if (u == U) { uv[u] = R & 0xffffffffu; return; }            /* sentinel {ref_list = n, length = 0} (:702-705) */
            uint32_t rl = 0;
            if (uf[u] != NONE32) rl = uf[u];
            else if (u > 0 && uf[u - 1] != NONE32) rl = ulst[u - 1];
            uv[u] = (uint64_t)rl | ((uint64_t)ul[u] << 32);
        });
        px::sync();
    }
    lap("reference positions of unitigs");

    /* ---- 6. the 30 short suffixes of every unitig end (genSpKmers, reference src/idx.c:514-526; order :857-881) ---- */
    const uint64_t n_sp = U * 30;
    const uint64_t N = U + n_sp + K;                          /* rows of the BWT = length of the unitig text */
    px::Buf<uint64_t> sp_key(n_sp), sp_val(n_sp);
    {
        px::Buf<uint64_t> ka(n_sp), va(n_sp);
        uint64_t *kp = sp_key.p, *vp = sp_val.p;
        const uint64_t *ue = uendk.p;
        px::for_each(n_sp, PX_LAMBDA(uint64_t s) {
            uint64_t l = s / U + 1, u = s % U;                /* generated in (length, unitig) order: the stable tie order */
            uint64_t ek = ue[u];
            uint64_t val = ek & (((uint64_t)1 << (2 * l)) - 1);
            kp[s] = val << (2 * (30 - l));                    /* the suffix padded with A to 30 bases */
            vp[s] = (u << 8) | (l << 2) | ((ek >> (2 * l)) & 3);
        });
        uint64_t *k0 = sp_key.p, *k1 = ka.p, *v0 = sp_val.p, *v1 = va.p;
        px::sort_pairs(k0, k1, v0, v1, n_sp, 60);
        if (k0 != sp_key.p) {                                 /* result landed in the alternates: keep those */
            std::swap(sp_key.p, ka.p);
            std::swap(sp_val.p, va.p);
        }
    }
    lap("unitig-end suffixes sorted");
    /* merge with the k-mers by counting (merge_kmer, reference src/idx.c:345-389): a suffix S# goes before every k-mer
     * >= S padded with A */
    px::Buf<uint32_t> sp_lb(n_sp);
    px::Buf<uint32_t> below(K + 2);      /* below[j + 1] = number of short suffixes placed before k-mer j */
    {
        px::Buf<uint32_t> cnt(K + 2);
        px::zero(cnt.p, (K + 2) * 4);
        uint32_t *cp = cnt.p, *lb = sp_lb.p;
        const uint64_t *kp = sp_key.p;
        px::for_each(n_sp, PX_LAMBDA(uint64_t s) {
            uint64_t l = kmer_lower_bound(kv, ct, kp[s] << 2);
            lb[s] = (uint32_t)l;
            px::atomic_add32(cp + l, 1);
        });
        const uint32_t *ccp = cnt.p;
        /* below[m] = sum_{t < m} cnt[t], so below[j + 1] = suffixes whose lower bound is <= j */
        px::exclusive_sum<uint32_t>(K + 2, PX_LAMBDA(uint64_t j) { return (uint64_t)ccp[j]; }, below.p);
    }
    const uint32_t *bl = below.p;    /* bl[j + 1] = suffixes with lower bound <= j */
    lap("suffixes merged with k-mers");

    /* ---- 7. BWT symbols ---- */
    px::Buf<uint8_t> bwt8(N + 512);
    {
        uint8_t *bw = bwt8.p;
        const uint64_t *ue = uendk.p, *vp = sp_val.p;
        const uint32_t *lb = sp_lb.p;
        px::for_each(512, PX_LAMBDA(uint64_t i) { bw[N + i] = 0xf; });
        px::for_each(U, PX_LAMBDA(uint64_t u) { bw[u] = (uint8_t)(ue[u] & 3); });            /* rows of '#'/'$' suffixes (:920-921) */
        px::for_each(n_sp, PX_LAMBDA(uint64_t s) { bw[U + s + lb[s]] = (uint8_t)(vp[s] & 3); });
        px::for_each(K, PX_LAMBDA(uint64_t j) {
            uint32_t info = ki[j];
            uint8_t c = (info & I_START) ? (ud[j] == 0 ? 5 : 4) : (uint8_t)low_bit4((info >> 4) & 15);
            bw[U + j + bl[j + 1]] = c;
        });
    }
    lap("BWT symbols");

    /* 13-mer prefix intervals (store_hash + compression, reference src/idx.c:333-343,935-957) */
    px::Buf<uint64_t> lut(n_lut);
    {
        px::Buf<uint64_t> first(n_lut - 1), end(n_lut - 1), exm(n_lut - 1);
        px::fill_ff(first.p, (n_lut - 1) * 8);
        px::zero(end.p, (n_lut - 1) * 8);
        uint64_t *fp = first.p, *ep = end.p;
        const uint64_t *kp = sp_key.p, *vp = sp_val.p;
        const uint32_t *lb = sp_lb.p;
        px::for_each(K, PX_LAMBDA(uint64_t j) {
            uint64_t p = kv[j] >> (2 * (KB - 13));
            uint64_t row = U + j + bl[j + 1];
            if (j == ct[p]) px::atomic_min64(fp + p, row);
            if (j + 1 == ct[p + 1]) px::atomic_max64(ep + p, row + 1);
        });
        px::for_each(n_sp, PX_LAMBDA(uint64_t s) {
            uint64_t l = (vp[s] >> 2) & 63;
            if (l < 13) return;
            uint64_t p = kp[s] >> (2 * (30 - 13));
            uint64_t row = U + s + lb[s];
            px::atomic_min64(fp + p, row);
            px::atomic_max64(ep + p, row + 1);
        });
        uint64_t last = px::exclusive_max(n_lut - 1, end.p, exm.p);
        uint64_t *lp = lut.p;
        const uint64_t *xm = exm.p;
        px::for_each(n_lut, PX_LAMBDA(uint64_t p) {
            if (p == n_lut - 1) { lp[p] = last; return; }
            lp[p] = fp[p] != NONE64 ? fp[p] : xm[p];
        });
        px::sync();
    }
    lap("13-mer prefix table");

    /* ---- 8. suffix-array samples: the (unitig, offset) of the symbol BEFORE each sampled row's suffix
     *         (bwt_cal_SA, reference src/idx.c:1163-1237, walks the whole BWT by LF for this) ---- */
    const uint64_t n_sa = (N + 7) / 8;
    px::Buf<uint64_t> sa(n_sa);
    {
        uint64_t *sp = sa.p;
        const uint64_t *vp = sp_val.p;
        const uint32_t *lb = sp_lb.p;
        px::zero(sp, n_sa * 8);
        px::for_each((U + 7) / 8, PX_LAMBDA(uint64_t q) { uint64_t u = q * 8; sp[q] = u | ((uint64_t)(ul[u] - 1) << 32); });
        px::for_each(n_sp, PX_LAMBDA(uint64_t s) {
            uint64_t row = U + s + lb[s];
            if (row & 7) return;
            uint64_t u = vp[s] >> 8, l = (vp[s] >> 2) & 63;
            sp[row >> 3] = u | ((uint64_t)(ul[u] - l - 1) << 32);
        });
        px::for_each(G, PX_LAMBDA(uint64_t i) {
            uint32_t j = lc[i];
            if (j == NONE32) return;
            uint64_t row = U + j + bl[j + 1];
            if (row & 7) return;
            uint64_t s = sg[i] - 1;
            uint64_t u = ud[lc[s]], off = i - s;
            uint64_t e;
            if (off > 0) e = u | ((off - 1) << 32);
            else if (u > 0) e = (u - 1) | ((uint64_t)ul[u - 1] << 32);
            else e = (U - 1) | ((uint64_t)ul[U - 1] << 32);
            sp[row >> 3] = e;
        });
    }
    seg.release();
    lap("suffix-array samples");

    /* ---- rank checkpoints every 256 rows + 4-bit packing (build_BWT / write_bwt, reference src/bwt.c:106-162,198-263) ---- */
    const uint64_t n_blk = (N + 255) / 256;
    const uint64_t occ_bytes = n_blk * 168;
    px::Buf<uint8_t> occ(occ_bytes);
    uint64_t rank[5];
    {
        px::Buf<uint32_t> bc(n_blk * 5);
        px::Buf<uint64_t> base(n_blk * 5);
        uint32_t *bcp = bc.p;
        const uint8_t *bw = bwt8.p;
        px::for_each(n_blk, PX_LAMBDA(uint64_t b) {
            uint32_t c[6] = {0, 0, 0, 0, 0, 0};
            uint64_t s = b * 256, e = s + 256 < N ? s + 256 : N;
            for (uint64_t r = s; r < e; r++) c[bw[r]]++;
            for (int k = 0; k < 5; k++) bcp[(uint64_t)k * n_blk + b] = c[k];
        });
        uint64_t tot[5];
        for (int k = 0; k < 5; k++) {
            const uint32_t *src = bc.p + (uint64_t)k * n_blk;
            tot[k] = px::exclusive_sum<uint64_t>(n_blk, PX_LAMBDA(uint64_t b) { return (uint64_t)src[b]; }, base.p + (uint64_t)k * n_blk);
        }
        rank[0] = tot[4] + 1;            /* '#' and the one '$' sort before A (src/bwt.c:133) */
        rank[1] = rank[0] + tot[0];
        rank[2] = rank[1] + tot[1];
        rank[3] = rank[2] + tot[2];
        rank[4] = 0;
        if (tot[4] + 1 != U) px::fail("BWT", "terminator count differs from the unitig count");
        uint64_t *op = (uint64_t *)occ.p;
        const uint64_t *bp = base.p;
        const uint64_t Lbin = (N + 1) >> 1;                   /* bytes of packed symbols (src/bwt.c:144) */
        px::for_each(n_blk * 21, PX_LAMBDA(uint64_t w) {
            uint64_t b = w / 21, q = w % 21;
            if (q < 5) { op[w] = bp[q * n_blk + b]; return; }
            uint64_t v = 0;
            for (int k = 0; k < 8; k++) {
                uint64_t g = b * 128 + (q - 5) * 8 + k;      /* byte of the packed string */
                uint64_t src = g;
                /* write_bwt copies only the bytes that exist into the last block and writes the block whole: the rest
                 * is what its 256-block staging buffer held, i.e. the same bytes of block b-256 (src/bwt.c:236-241) */
                if (g >= Lbin) { if (b < 256) continue; src = g - 256 * 128; }
                uint8_t lo = bw[2 * src], hi = bw[2 * src + 1];     /* low nibble first; a missing last symbol packs as 0xF (:150-153) */
                v |= (uint64_t)(((hi & 15) << 4) | (lo & 15)) << (8 * k);
            }
            op[w] = v;
        });
        px::sync();
    }
    bwt8.release();
    lap("rank checkpoints + packing");

    /* ---- 9. exist-k-mer tables (get_EXIST_kmer, reference src/idx.c:986-1026): every k-mer window inside a unitig
     *         = every window inside a run of >= 31 ACGT of a reference ---- */
    uint64_t ek_bytes;
    int ek_k, ek_bits;
    ekmer_class(K, &ek_bytes, &ek_k, &ek_bits);
    px::Buf<uint32_t> ek0(ek_bytes / 4), ek1(ek_bytes / 4);
    px::zero(ek0.p, ek_bytes);
    px::zero(ek1.p, ek_bytes);
    {
        uint32_t *t0 = ek0.p, *t1 = ek1.p;
        const uint64_t hmask = ((uint64_t)1 << ek_bits) - 1, emask = ((uint64_t)1 << (2 * ek_k)) - 1;
        const int k = ek_k;
        px::for_each(G, PX_LAMBDA(uint64_t q) {
            if (q + k > G) return;
            /* some valid 31-mer window [i, i+31) with i <= q and i + 31 >= q + k must hold the k-mer window */
            uint64_t a = q + k >= (uint64_t)KB ? q + k - KB : 0;
            bool ok = false;
            for (uint64_t i = q;; i--) {
                if (lc[i] != NONE32) { ok = true; break; }
                if (i == a) break;
            }
            if (!ok) return;
            uint64_t km = 0;
            for (int t = 0; t < k; t++) km = (km << 2) | (cd[q + t] & 3);
            km &= emask;
            uint64_t h1 = hash_a(km) & hmask, h2 = hash_b(km) & hmask;
            uint32_t m1 = 1u << (8 * ((h1 >> 3) & 3) + (7 - (h1 & 7)));   /* byte h>>3, bit 0x80 >> (h & 7) (:1016-1019) */
            uint32_t m2 = 1u << (8 * ((h2 >> 3) & 3) + (7 - (h2 & 7)));
            if (!(t0[h1 >> 5] & m1)) px::atomic_or32(t0 + (h1 >> 5), m1);
            if (!(t1[h2 >> 5] & m2)) px::atomic_or32(t1 + (h2 >> 5), m2);
        });
    }
    lap("exist-k-mer tables");

    /* ---- write the ten files (write_idx / write_bwt, reference src/idx.c:1046-1101, src/bwt.c:198-263) ---- */
    mkdir(out_dir, 0777);
    Out o{out_dir};
    FILE *fp = o.open_part(".bwt");
    put(fp, &occ_bytes, 8);
    put_device(fp, occ.p, occ_bytes);
    put(fp, rank, 40);
    put_device(fp, lut.p, n_lut * 8);
    fclose(fp);
    fp = o.open_part(".sa");
    put(fp, &n_sa, 8);
    put_device(fp, sa.p, n_sa * 8);
    fclose(fp);
    fp = o.open_part(".acg");                                 /* bwt_cal_AGCTCounter, reference src/bwt.c:169-182 */
    {
        uint64_t sz = 1 << 16;
        put(fp, &sz, 8);
        std::vector<uint8_t> t(sz);
        static const uint32_t mask[5] = {0x0000, 0x1111, 0x2222, 0x3333, 0x4444};
        for (int c = 0; c < 5; c++) {
            for (uint32_t i = 0; i < sz; i++) {
                uint32_t v = i ^ mask[c];
                t[i] = (uint8_t)(((v & 0xf) == 0) + ((v & 0xf0) == 0) + ((v & 0xf00) == 0) + ((v & 0xf000) == 0));
            }
            put(fp, t.data(), sz);
        }
    }
    fclose(fp);
    fp = o.open_part(".exki");
    put(fp, &ek_bytes, 8);
    fclose(fp);
    fp = o.open_part(".exk0");
    put_device(fp, ek0.p, ek_bytes);
    fclose(fp);
    fp = o.open_part(".exk1");
    put_device(fp, ek1.p, ek_bytes);
    fclose(fp);
    fp = o.open_part(".unv");
    {
        uint64_t n = U + 1;
        put(fp, &n, 8);
        put_device(fp, unv.p, n * 8);
    }
    fclose(fp);
    fp = o.open_part(".ref_b");
    put(fp, &ref_bin_n, 8);
    put_device(fp, ref_bin.p, ref_bin_n);
    fclose(fp);
    fp = o.open_part(".ref_i");
    {
        put(fp, &n_ref, 8);
        for (uint64_t r = 0; r < n_ref; r++) {
            /* REF_INFO {char ref_name[128]; u64 seq_l; u64 seq_offset} (reference src/idx.h:22-26).  The reference leaves
             * the bytes after the name's NUL as its heap had them; they are zero here. */
            char rec[144];
            memset(rec, 0, sizeof rec);
            snprintf(rec, 128, "%s", rs.name[r].c_str());
            memcpy(rec + 128, &rs.len[r], 8);
            memcpy(rec + 136, &rs.off[r], 8);
            put(fp, rec, 144);
        }
    }
    fclose(fp);
    fp = o.open_part(".ref_p");
    put(fp, &R, 8);
    put_device(fp, ref_pos.p, R * 8);
    fclose(fp);
    lap("files written");

    st->n_ref = n_ref;
    st->n_bases = G;
    st->n_kmer = K;
    st->n_unitig = U;
    st->n_rows = N;
    st->n_ref_pos = R;
    st->exist_table_bytes = ek_bytes;
    st->exist_k = ek_k;
    st->seconds = now() - t0;
    st->peak_device_bytes = px::g_peak_bytes;
    st->kernel_launches = px::g_launches;
}

}  // namespace dsbb

extern "C" int dsb_index_build(const char *ref_fasta, const char *kmer_srt, const char *out_dir, int dev, dsb_build_stats *stats)
{
    dsb_build_stats local;
    if (!stats) stats = &local;
    memset(stats, 0, sizeof *stats);
    px::g_error[0] = 0;
    px::g_launches = 0;
    px::g_live_bytes = px::g_peak_bytes = 0;
#ifndef DSB_BUILD_HOSTSIM
    if (cudaSetDevice(dev) != cudaSuccess) {
        snprintf(px::g_error, sizeof px::g_error, "cudaSetDevice(%d) failed: no usable GPU (there is no CPU path)", dev);
        return -1;
    }
#else
    (void)dev;
#endif
    try {
        dsbb::build(ref_fasta, kmer_srt, out_dir, stats);
    } catch (px::Fail &) {
        return -1;
    } catch (std::bad_alloc &) {
        snprintf(px::g_error, sizeof px::g_error, "out of host memory");
        return -1;
    }
    return 0;
}

extern "C" const char *dsb_build_last_error(void) { return px::g_error; }
