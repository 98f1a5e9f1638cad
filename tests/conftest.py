"""Shared fixtures.  `-m "not gpu"` runs everywhere; `-m gpu` needs a B200 and calls through the C-ABI."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "desamba-so_b200")
sys.path.insert(0, PKG)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests", "hostsim"))
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); runs the sm_100a kernels through the C-ABI")


def _make(args):
    r = subprocess.run(["make"] + args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    assert r.returncode == 0, r.stdout.decode()[-3000:]


@pytest.fixture(scope="session")
def built():
    """Everything the tests load: product libraries, host simulator, oracle (checker)."""
    if not os.path.exists(os.path.join(PKG, "lib", "libdesamba.so")):
        _make(["-C", PKG, "all"])
    if not os.path.exists(os.path.join(ROOT, "tests", "hostsim", "libdesamba_hostsim.so")):
        _make(["-C", PKG, "hostsim"])
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "liboracle_port.so")):
        _make(["-C", os.path.join(ROOT, "oracle"), "port"])
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "deSAMBA")) and os.path.isdir("/root/reference/src"):
        _make(["-C", os.path.join(ROOT, "oracle"), "ref"])
    return True


@pytest.fixture(scope="session")
def have_ref_binary(built):
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "deSAMBA"))


@pytest.fixture(scope="session")
def tiny(built, have_ref_binary):
    if not have_ref_binary:
        pytest.skip("oracle/_ref/deSAMBA (needed to build a test index) is absent")
    import dataset
    return dataset.build("tiny")


@pytest.fixture(scope="session")
def mixed(built, have_ref_binary):
    if not have_ref_binary:
        pytest.skip("oracle/_ref/deSAMBA (needed to build a test index) is absent")
    import dataset
    return dataset.build("mixed")


def sam_records(text):
    """{read name: [classification fields of each record]} -- QNAME FLAG RNAME POS MAPQ CIGAR AS."""
    d = {}
    if isinstance(text, bytes):
        text = text.decode()
    for line in text.split("\n"):
        if not line:
            continue
        f = line.split("\t")
        d.setdefault(f[0], []).append(tuple(f[:6] + f[11:12]))
    return d


def ref_resident_session(idx_dir, slices, flag, snap):
    """The UNMODIFIED reference library (oracle/_ref/libdesamba.so, the checker) driven the way BASELINE.json
    configs[4] describes: one load_index, then read_classify + meta_analysis per slice on one thread_id.  Runs in
    a child process (the reference prints to stderr and aborts on errors)."""
    code = r'''
import ctypes as C, sys, pickle
lib = C.CDLL(sys.argv[1]); libc = C.CDLL(None); libc.free.argtypes = [C.c_void_p]
idx = C.c_void_p(); lib.load_index(C.byref(idx), sys.argv[2].encode())
lib.read_classify.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int, C.c_int]
lib.meta_analysis.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int, C.c_int,
                              C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
slices, flag, snap = pickle.load(open(sys.argv[3], "rb"))
res = []
for fq in slices:
    o, n = C.c_void_p(), C.c_uint64(0)
    lib.read_classify(idx, fq, len(fq), C.byref(o), C.byref(n), 7, 1)
    sam = C.string_at(o.value, n.value)
    out, m, hs, hn = C.c_void_p(), C.c_uint64(0), C.c_void_p(), C.c_uint64(0)
    lib.meta_analysis(idx, sam, len(sam), C.byref(out), C.byref(m), 7, flag, snap, C.byref(hs), C.byref(hn))
    res.append((sam, C.string_at(out.value, m.value), C.string_at(hs.value, hn.value) if hs.value else None))
pickle.dump(res, open(sys.argv[4], "wb"))
'''
    import pickle
    import sys
    import tempfile
    d = tempfile.mkdtemp()
    pickle.dump((slices, flag, snap), open(os.path.join(d, "in.pkl"), "wb"))
    r = subprocess.run([sys.executable, "-c", code, os.path.join(ROOT, "oracle", "_ref", "libdesamba.so"), idx_dir + "/",
                        os.path.join(d, "in.pkl"), os.path.join(d, "out.pkl")], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-1500:]
    return pickle.load(open(os.path.join(d, "out.pkl"), "rb"))


def fastq_slices(path, n_slices):
    recs = open(path, "rb").read().split(b"\n")
    n_rec = len(recs) // 4
    per = (n_rec + n_slices - 1) // n_slices
    return [b"\n".join(recs[4 * i:4 * min(i + per, n_rec)]) + b"\n" for i in range(0, n_rec, per)]


def ref_cli(args):
    """stdout of the UNMODIFIED reference binary (oracle/_ref/deSAMBA, the checker): `classify -t 1 <args>`."""
    r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "deSAMBA"), "classify", "-t", "1"] + args, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-1500:]
    return r.stdout


# the classify command line's options and formats (reference src/cly_mt.c:431-499), as argument lists without index / reads
CLI_VARIANTS = [["-f", "DES"], ["-f", "DES_FULL"], ["-f", "SAM_FULL"], ["-f", "SAM", "-r", "1"], ["-f", "DES", "-r", "0"],
                ["-f", "SAM", "-s", "200"], ["-f", "SAM", "-l", "400"], ["-f", "SAM", "-l", "60", "-s", "20", "-r", "9"]]


def odd_base_reads(src_fastq, dst_fastq, seed=5):
    """Reads with N runs, lower case, IUPAC / junk characters (all encode as 'C') and lengths around the 40-base
    minimum (src/cly.c:3084), derived from an existing FASTQ."""
    import random
    rng = random.Random(seed)
    lines = open(src_fastq).read().split("\n")
    out = []
    for i in range(0, len(lines) - 1, 4):
        name, seq, plus, qual = lines[i:i + 4]
        k = (i // 4) % 8
        if k == 0:
            s = list(seq)
            for _ in range(3):
                p = rng.randrange(0, max(1, len(s) - 30))
                for j in range(p, min(len(s), p + rng.randrange(1, 30))):
                    s[j] = "N"
            seq = "".join(s)
        elif k == 1:
            seq = seq.lower()
        elif k == 2:
            seq = "".join(c if rng.random() > 0.02 else rng.choice("RYKMSWnx-") for c in seq)
        elif k == 3:
            seq = seq[:rng.choice([1, 10, 39, 40, 41, 60])]
        elif k == 4:
            seq = "".join(c.lower() if rng.random() < 0.5 else c for c in seq)
        out += [name, seq, plus, qual[:len(seq)]]
    open(dst_fastq, "w").write("\n".join(out) + "\n")


def ref_call_sequence(idx_dir, calls):
    """read_classify() of the UNMODIFIED reference library for a list of (fastq text, thread_id, thread_num) calls on one
    load_index, in a child process; returns the SAM_FULL text of every call."""
    code = r'''
import ctypes as C, sys, pickle
lib = C.CDLL(sys.argv[1]); idx = C.c_void_p(); lib.load_index(C.byref(idx), sys.argv[2].encode())
lib.read_classify.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int, C.c_int]
res = []
for fq, tid, tn in pickle.load(open(sys.argv[3], "rb")):
    o, n = C.c_void_p(), C.c_uint64(0)
    lib.read_classify(idx, fq, len(fq), C.byref(o), C.byref(n), tid, tn)
    res.append(C.string_at(o.value, n.value))
pickle.dump(res, open(sys.argv[4], "wb"))
'''
    import pickle
    import sys
    import tempfile
    d = tempfile.mkdtemp()
    pickle.dump(calls, open(os.path.join(d, "in.pkl"), "wb"))
    r = subprocess.run([sys.executable, "-c", code, os.path.join(ROOT, "oracle", "_ref", "libdesamba.so"), idx_dir + "/",
                        os.path.join(d, "in.pkl"), os.path.join(d, "out.pkl")], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-1500:]
    return pickle.load(open(os.path.join(d, "out.pkl"), "rb"))


def max_read_l_stream(src_fastq, dst_fastq, junk_first=True, seed=17):
    """A read stream for the reference's max_read_l rule (src/cly.c:2884-2885,2953-2981): three long reads that cannot have
    a chain (random bases; with junk_first=False, three long REAL reads instead), then 60 reads cut down to 100..509 bases,
    one long real read, then 40 more short reads.  Until the first long read WITH a chain has been classified the short reads
    take the "2G" filter; afterwards the long-read filters."""
    import random
    rng = random.Random(seed)
    lines = open(src_fastq).read().split("\n")
    recs = [lines[i:i + 4] for i in range(0, len(lines) - 1, 4)]
    real = [r for r in recs if "junk" not in r[0]]
    longs = [r for r in real if len(r[1]) >= 1500]
    out = []
    for i in range(3):
        if junk_first:
            s = "".join(rng.choice("ACGT") for _ in range(3000))
            out += ["@lead%d_junk" % i, s, "+", "I" * len(s)]
        else:
            out += ["@lead%d" % i] + longs[i][1:]

    def short(r, k):
        n = rng.choice([100, 150, 250, 309, 310, 400, 509])
        n = min(n, len(r[1]))
        st = rng.randrange(0, len(r[1]) - n + 1)
        return ["@s%d_%s" % (k, r[0][1:]), r[1][st:st + n], "+", r[3][st:st + n]]
    pool = [r for r in real if len(r[1]) >= 520]
    for k in range(60):
        out += short(pool[k % len(pool)], k)
    out += ["@mid_long"] + longs[5][1:]
    for k in range(60, 100):
        out += short(pool[k % len(pool)], k)
    open(dst_fastq, "w").write("\n".join(out) + "\n")


def exact_copy_reads(meta, tmp_path, n=3, length=3000):
    """FASTQ of error-free copies of the reference (an exact match of 2000 bases or more makes the reference
    index past its Q_MEM table -- reference-undefined; this build scores such a seed like a 1999-base one)."""
    ref = os.path.join(os.path.dirname(meta["idx"].rstrip("/")), "ref.fa")
    seqs, cur = [], []
    for line in open(ref):
        if line.startswith(">"):
            if cur:
                seqs.append("".join(cur))
            cur = []
        else:
            cur.append(line.strip())
    seqs.append("".join(cur))
    path = os.path.join(str(tmp_path), "exact.fq")
    comp = str.maketrans("ACGT", "TGCA")
    with open(path, "w") as f:
        for i in range(n):
            s = seqs[i % len(seqs)][5000 + 1000 * i:5000 + 1000 * i + length]
            if i % 2:
                s = s.translate(comp)[::-1]
            f.write("@exact%d\n%s\n+\n%s\n" % (i, s, "I" * len(s)))
    return path
