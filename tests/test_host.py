"""Host-side logic without a GPU: the C-ABI libraries load and export what include/*.h declares, and
the C host code (FASTQ parser, history model, batcher, SAM formatter) reproduces the reference's SAM
when the kernel source runs on the host simulator (tests/hostsim, test infrastructure)."""
import ctypes as C
import gzip
import os
import re
import subprocess
import sys

import pytest

import pydsb_hostsim

from conftest import GOLD, PKG, ROOT, sam_records

SIM = os.path.join(ROOT, "tests", "hostsim", "deSAMBA_hostsim")


def declared_functions(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(\w+)\s*\([^;{]*\)\s*;", text)))


def test_cuda_abi_exports_every_declared_symbol(built):
    names = declared_functions("dsb_cuda.h")
    assert "dsb_slot_submit" in names and "dsb_fm_rank_batch" in names and len(names) >= 15
    lib = C.CDLL(os.path.join(PKG, "lib", "libdsb_cuda.so"))
    for n in names:
        assert hasattr(lib, n), n


def test_api_exports_exactly_the_reference_symbols(built):
    names = declared_functions("desamba.h")
    assert names == ["load_index", "meta_analysis", "read_classify"]
    lib = C.CDLL(os.path.join(PKG, "lib", "libdesamba.so"))
    for n in names:
        assert hasattr(lib, n), n


def test_product_library_has_no_oracle_or_simulator_dependency(built):
    for so in ("libdesamba.so", "libdsb_cuda.so"):
        out = subprocess.run(["ldd", os.path.join(PKG, "lib", so)], stdout=subprocess.PIPE).stdout.decode()
        assert "hostsim" not in out and "oracle" not in out
    # the product CLI aborts loudly without a CUDA device instead of falling back
    if not os.path.exists("/dev/nvidia0"):
        r = subprocess.run([os.path.join(PKG, "bin", "deSAMBA"), "classify", os.path.join(GOLD, "nonexistent"), "x.fq"],
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r.returncode != 0


def run_sim(idx, reads, fmt="SAM", env=None, extra=()):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([SIM, "classify", "-f", fmt] + list(extra) + [idx] + list(reads), stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, env=e)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    return r.stdout


@pytest.mark.parametrize("name", ["tiny", "mixed"])
def test_hostsim_cli_equals_reference_golden(name, request):
    meta = request.getfixturevalue(name)
    out = run_sim(meta["idx"], [meta["reads"]])
    assert out == open(os.path.join(GOLD, name + ".sam"), "rb").read()


def test_batch_boundaries_and_device_count_do_not_change_results(mixed):
    want = open(os.path.join(GOLD, "mixed.sam"), "rb").read()
    assert run_sim(mixed["idx"], [mixed["reads"]], env={"DSB_BATCH_READS": "37"}) == want
    assert run_sim(mixed["idx"], [mixed["reads"]], env={"DSB_BATCH_BASES": "20000"}) == want
    assert run_sim(mixed["idx"], [mixed["reads"]], env={"DSB_BATCH_READS": "11", "DSB_HOSTSIM_GPUS": "2"}) == want


def test_gzip_and_fasta_inputs(tiny, tmp_path):
    recs = open(tiny["reads"], "rb").read().split(b"\n")
    n = 40
    fq = b"\n".join(recs[:4 * n]) + b"\n"
    gz = tmp_path / "r.fq.gz"
    gz.write_bytes(gzip.compress(fq))
    fa = tmp_path / "r.fa"
    with open(fa, "wb") as fh:
        for i in range(n):
            seq = recs[4 * i + 1]
            fh.write(b">" + recs[4 * i][1:] + b" some comment\n")
            for j in range(0, len(seq), 61):          # multi-line FASTA
                fh.write(seq[j:j + 61] + b"\n")
    plain = tmp_path / "r.fq"
    plain.write_bytes(fq)
    a = run_sim(tiny["idx"], [str(plain)])
    assert run_sim(tiny["idx"], [str(gz)]) == a
    assert run_sim(tiny["idx"], [str(fa)]) == a
    want = sam_records(open(os.path.join(GOLD, "tiny.sam"), "rb").read())
    got = sam_records(a)
    assert len(got) == n
    for k, v in got.items():
        assert want[k] == v


def test_api_read_classify_through_hostsim_library(tiny):
    """read_classify(): text input, path input, empty input, history carried across calls per thread_id."""
    import pydsb
    d = pydsb_hostsim.Desamba()
    d.load_index(tiny["idx"])
    text = open(tiny["reads"], "rb").read()
    recs = text.split(b"\n")
    half = len(recs) // 8 * 4
    a, b = b"\n".join(recs[:half]) + b"\n", b"\n".join(recs[half:])
    want = sam_records(open(os.path.join(GOLD, "tiny.sam"), "rb").read())
    # two consecutive calls on one thread_id == one reference run over the concatenation
    got = sam_records(d.read_classify(a + b"\0", thread_id=3) + d.read_classify(b, thread_id=3))
    assert got == want
    # path input on a fresh thread_id
    assert sam_records(d.read_classify(tiny["reads"], thread_id=4)) == want
    # input_n == 0 leaves the outputs alone
    out, out_n = C.c_void_p(0x1234), C.c_uint64(77)
    d.lib.read_classify(d.idx, b"", 0, C.byref(out), C.byref(out_n), 0, 1)
    assert out.value == 0x1234 and out_n.value == 0


def test_sam_full_format_carries_seq_and_qual(tiny):
    out = run_sim(tiny["idx"], [tiny["reads"]], fmt="SAM_FULL").decode().split("\n")
    prim = [ln.split("\t") for ln in out if ln and not (int(ln.split("\t")[1]) & 0x900)]
    assert prim and all(len(f[9]) == len(f[10]) and (f[9] == "*" or set(f[10]) == {"I"}) for f in prim)


def _ref_meta(idx_dir, sams, flag, snap):
    """meta_analysis of the UNMODIFIED reference library (oracle/_ref/libdesamba.so, the checker), in a child
    process: the reference prints to stderr and aborts on errors."""
    code = r'''
import ctypes as C, sys, pickle
lib = C.CDLL(sys.argv[1]); libc = C.CDLL(None); libc.free.argtypes = [C.c_void_p]
idx = C.c_void_p(); lib.load_index(C.byref(idx), sys.argv[2].encode())
lib.meta_analysis.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int, C.c_int,
                              C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
sams, flag, snap = pickle.load(open(sys.argv[3], "rb"))
# the reference creates a thread_id's buffers in read_classify (meta_analysis only looks them up): one short read first
lib.read_classify.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int, C.c_int]
fq = b"@w\n" + b"ACGT" * 20 + b"\n+\n" + b"I" * 80 + b"\n"
o, n = C.c_void_p(), C.c_uint64(0)
lib.read_classify(idx, fq, len(fq), C.byref(o), C.byref(n), 0, 1)
res = []
for s in sams:
    out, n, hs, hn = C.c_void_p(), C.c_uint64(0), C.c_void_p(), C.c_uint64(0)
    lib.meta_analysis(idx, s, len(s), C.byref(out), C.byref(n), 0, flag, snap, C.byref(hs), C.byref(hn))
    res.append((C.string_at(out.value, n.value), C.string_at(hs.value, hn.value) if hs.value else None))
pickle.dump(res, open(sys.argv[4], "wb"))
'''
    import pickle
    import sys
    import tempfile
    d = tempfile.mkdtemp()
    pickle.dump((sams, flag, snap), open(os.path.join(d, "in.pkl"), "wb"))
    r = subprocess.run([sys.executable, "-c", code, os.path.join(ROOT, "oracle", "_ref", "libdesamba.so"), idx_dir + "/",
                        os.path.join(d, "in.pkl"), os.path.join(d, "out.pkl")], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-1500:]
    return pickle.load(open(os.path.join(d, "out.pkl"), "rb"))


def test_meta_analysis_equals_reference(tiny, mixed, have_ref_binary):
    """meta_analysis() on the SAM text of classified reads, both weighting flags, against the
    reference's own libdesamba.so: SAM and SAM_FULL input, a human (tid 9606) subset for the snapshot, an
    all-unclassified input (the 'no_match' line), headers, and a truncated snapshot."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdesamba.so")):
        pytest.skip("reference library absent")
    import pydsb
    sam = open(os.path.join(GOLD, "tiny.sam"), "rb").read()
    full = run_sim(tiny["idx"], [tiny["reads"]], fmt="SAM_FULL")
    lines = full.split(b"\n")
    human = b"\n".join(l.replace(b"tid|1003|", b"tid|9606|") for l in lines[:400])
    unmapped = b"\n".join(l for l in lines if b"\t4\t*\t" in l) + b"\n"
    header = b"@SQ\tSN:x\tLN:5\n@PG\tID:deSAMBA\n" + sam
    mixed_sam = open(os.path.join(GOLD, "mixed.sam"), "rb").read()
    cases = [sam, full, human, unmapped, header, mixed_sam[: mixed_sam.rfind(b"\n", 0, 30000) + 1]]
    # lines the non-modifying fast parser must hand to the strtok-faithful one (or parse to the same record): RNAME
    # without '|' (the reference then takes POS for the taxid), signs and doubled / leading '|', MAPQ with a leading blank,
    # RNAME ending right after the taxid, and a text whose last line has no newline
    recs = [l for l in sam.split(b"\n") if l][:120]

    def mutated(kind):
        out = []
        for i, l in enumerate(recs):
            f = l.split(b"\t")
            if f[2] != b"*" and i % 3 == 0:
                if kind == 0:
                    f[2] = b"chr1"
                elif kind == 1:
                    f[2] = f[2].replace(b"tid|", b"tid|+")
                elif kind == 2:
                    f[2] = f[2].replace(b"tid|", b"tid||")
                elif kind == 3:
                    f[2] = b"|" + f[2]
                elif kind == 4:
                    f[4] = b" 7"
                elif kind == 5:
                    f[2] = b"tid|1003"
            out.append(b"\t".join(f))
        return b"\n".join(out) + b"\n"
    cases += [mutated(k) for k in range(6)] + [sam.rstrip(b"\n")]
    cases.append(full * 3)              # > 4 MB: read by several threads (pieces cut at line ends, joined in order)
    d = pydsb_hostsim.Desamba()
    d.load_index(tiny["idx"])
    for flag in (0, 1):
        for snap in (65536, 1000, 0):          # 0: some human bases -> an empty block, not NULL
            want = _ref_meta(tiny["idx"], cases, flag, snap)
            for s, (w_text, w_snap) in zip(cases, want):
                got_text, got_snap = d.meta_analysis(s, thread_id=0, flag=flag, max_snapshot_len=snap)
                assert got_text == w_text, (flag, snap, got_text, w_text)
                assert got_snap == w_snap
    assert any(w[1] is not None for w in want)                       # the human case produced a snapshot


class _RecView(C.Structure):
    _fields_ = [("name", C.c_void_p), ("seq", C.c_void_p), ("qual", C.c_void_p), ("name_l", C.c_size_t), ("seq_l", C.c_size_t),
                ("is_fastq", C.c_int), ("stable", C.c_int)]


def _records(lib, text, threads):
    lib.dsb_fq_open_mem.restype = C.c_void_p
    lib.dsb_fq_open_mem.argtypes = [C.c_char_p, C.c_uint64]
    lib.dsb_fq_next_view.restype = C.c_int64
    lib.dsb_fq_next_view.argtypes = [C.c_void_p, C.POINTER(_RecView)]
    lib.dsb_fq_prescan.argtypes = [C.c_void_p, C.c_int]
    lib.dsb_fq_close.argtypes = [C.c_void_p]
    f = lib.dsb_fq_open_mem(text, len(text))
    if threads:
        lib.dsb_fq_prescan(f, threads)
    out, v = [], _RecView()
    while True:
        r = lib.dsb_fq_next_view(f, C.byref(v))
        if r < 0:
            out.append(r)
            break
        out.append((C.string_at(v.name, v.name_l), C.string_at(v.seq, v.seq_l),
                    C.string_at(v.qual, v.seq_l) if v.qual else None, v.is_fastq))
    _records.pieces = lib.dsb_fq_scan_pieces(C.c_void_p(f))
    lib.dsb_fq_close(f)
    return out


def test_parallel_fastq_prescan_equals_sequential_reader(built):
    """The threaded pre-scan of in-memory input must return exactly the sequential (kseq-semantics) reader's
    records: plain 4-line FASTQ, qualities that start with '@' or '+', wrapped records, FASTA records, comments,
    blank lines, a truncated tail."""
    import random
    lib = C.CDLL(os.path.join(ROOT, "tests", "hostsim", "libdesamba_hostsim.so"))
    rng = random.Random(5)
    parts = []
    for i in range(9000):
        L = rng.randint(40, 600)
        seq = "".join(rng.choice("ACGTN") for _ in range(L))
        qual = "".join(rng.choice("@+I5>!~") for _ in range(L))
        kind = rng.random()
        if kind < 0.90:
            parts.append("@r%d some comment\n%s\n+\n%s\n" % (i, seq, qual))
        elif kind < 0.94:
            h = L // 2
            parts.append("@w%d\n%s\n%s\n+w%d\n%s\n%s\n" % (i, seq[:h], seq[h:], i, qual[:h], qual[h:]))
        elif kind < 0.97:
            parts.append(">f%d\n%s\n" % (i, seq))
        else:
            parts.append("\n\n@b%d\n%s\n+\n%s\n\n" % (i, seq, qual))
    parts.append("@s\nACGTACGT\n+\nIIIII\n")                                  # short quality at the end: kseq returns -2
    text = "".join(parts).encode()
    assert len(text) > (2 << 20)
    want = _records(lib, text, 0)
    assert len(want) > 8000
    for threads in (2, 5, 8):
        assert _records(lib, text, threads) == want
    clean = "".join(p for p in parts if p.startswith("@r")).encode()
    assert _records(lib, clean, 8) == _records(lib, clean, 0)
    assert _records(lib, clean[:-7], 8) == _records(lib, clean[:-7], 0)
    # the block is scanned piece by piece (the first batch starts after the first piece): several pieces, then a rest too
    # small for the threads, then -- in the mixed text -- a piece that ends at a record the fast recogniser refuses
    os.environ["DSB_SCAN_PIECE"] = str(1 << 20)
    try:
        big = clean * 3
        want_big = _records(lib, big, 0)
        for threads in (1, 3, 8):
            assert _records(lib, big, threads) == want_big
            assert _records.pieces >= 5
        assert _records(lib, clean * 2 + text, 4) == _records(lib, clean * 2 + text, 0)
    finally:
        del os.environ["DSB_SCAN_PIECE"]


def test_resident_index_call_sequence_through_hostsim(mixed, have_ref_binary):
    """BASELINE.json configs[4], host logic on CPU: one load_index, consecutive read_classify slices on one thread_id,
    meta_analysis after each; the whole call sequence equals the unmodified reference library driven the same way
    (the buffer history carries across calls).  The GPU twin is tests/test_gpu.py::test_resident_index_repeated_batches_with_meta_analysis."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdesamba.so")):
        pytest.skip("reference library absent")
    import pydsb
    from conftest import fastq_slices, ref_resident_session
    slices = fastq_slices(mixed["reads"], 5)
    slices = slices + slices[:2]
    want = ref_resident_session(mixed["idx"], slices, 1, 65536)
    d = pydsb_hostsim.Desamba()
    d.load_index(mixed["idx"])
    for i, (fq, (w_sam, w_meta, w_snap)) in enumerate(zip(slices, want)):
        sam = d.read_classify(fq, thread_id=7, thread_num=1)
        assert sam == w_sam, i
        meta, snap = d.meta_analysis(sam, thread_id=7, flag=1, max_snapshot_len=65536)
        assert (meta, snap) == (w_meta, w_snap), i


def test_api_splice_formatter_equals_cli_formatter(mixed, tmp_path):
    """read_classify() writes SAM_FULL by splicing SEQ / QUAL straight into the output block (dsb_batch.c splice_part);
    the CLI formats whole records through the sink.  Same bytes, for FASTQ and for FASTA (no qualities) input."""
    import pydsb
    fq = open(mixed["reads"], "rb").read()
    lines = fq.split(b"\n")
    fa = b"".join(b">" + lines[i][1:] + b"\n" + lines[i + 1] + b"\n" for i in range(0, len(lines) - 1, 4))
    fa_path = str(tmp_path / "reads.fa")
    open(fa_path, "wb").write(fa)
    for data, path in ((fq, mixed["reads"]), (fa, fa_path)):
        d = pydsb_hostsim.Desamba()
        d.load_index(mixed["idx"])
        assert d.read_classify(data, thread_id=3) == run_sim(mixed["idx"], [path], fmt="SAM_FULL")


def test_mapped_file_input_equals_stream_reader(mixed, tmp_path):
    """A plain reads file is mapped and parsed in memory mode (threaded pre-scan, zero-copy views); DSB_NO_MMAP=1 keeps
    the kseq-style stream reader the reference uses.  Same records, same output: FASTQ, FASTA with wrapped lines, and a
    file without a final newline."""
    want = open(os.path.join(GOLD, "mixed.sam"), "rb").read()
    assert run_sim(mixed["idx"], [mixed["reads"]]) == want
    assert run_sim(mixed["idx"], [mixed["reads"]], env={"DSB_NO_MMAP": "1"}) == want
    fq = open(mixed["reads"], "rb").read()
    lines = fq.split(b"\n")
    fa = b"".join(b">" + lines[i][1:] + b" some comment\n" + b"\n".join(lines[i + 1][k:k + 70] for k in range(0, len(lines[i + 1]), 70)) + b"\n"
                  for i in range(0, len(lines) - 1, 4))
    p1, p2 = str(tmp_path / "wrapped.fa"), str(tmp_path / "nonl.fq")
    open(p1, "wb").write(fa)
    open(p2, "wb").write(fq.rstrip(b"\n"))
    for p in (p1, p2):
        a = run_sim(mixed["idx"], [p])
        assert a == run_sim(mixed["idx"], [p], env={"DSB_NO_MMAP": "1"})
        assert sam_records(a) == sam_records(want)


def test_reblocked_fm_rank_equals_reference_occ_for_every_symbol(tiny):
    """The re-blocked FM layout (bit planes + A/C/G/T counts; '#' derived from the true '$' row) answers the reference's
    occ known answers for every symbol, '#' and '$' included: map_seed's second LF walk does consume a '#' rank.  Host twin
    of tests/test_gpu.py::test_fm_rank_kernel_equals_reference_occ (same layout code, compiled for the host)."""
    import struct
    import numpy as np
    import pydsb
    raw = open(os.path.join(GOLD, "kat_tiny", "occ.bin"), "rb").read()
    n = struct.unpack_from("<Q", raw, 0)[0]
    o = 8
    r = np.frombuffer(raw, dtype=np.uint64, count=n, offset=o); o += 8 * n
    ci = np.frombuffer(raw, dtype=np.uint8, count=n, offset=o); o += n
    co = np.frombuffer(raw, dtype=np.uint8, count=n, offset=o); o += n
    res = np.frombuffer(raw, dtype=np.uint64, count=n, offset=o)
    ix = pydsb_hostsim.DeviceIndex(tiny["idx"])
    got, c_out, _ms = ix.fm_rank(r, ci)
    ix.close()
    assert (c_out == co).all() and (got == res).all()
    assert (co == 4).any()


def test_cli_options_and_formats_equal_reference_hostsim(mixed, have_ref_binary, tmp_path):
    """Every output format and the -r / -s / -l options of `deSAMBA classify` (reference src/cly_mt.c:431-499), and -o FILE:
    byte-identical to the reference binary run here on the same input (host side + host simulator of the kernels)."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    from conftest import CLI_VARIANTS, ref_cli
    for opts in CLI_VARIANTS:
        want = ref_cli(opts + [mixed["idx"] + "/", mixed["reads"]])
        r = subprocess.run([SIM, "classify"] + opts + [mixed["idx"], mixed["reads"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r.returncode == 0, r.stderr.decode()[-1500:]
        assert r.stdout == want, opts
    out = str(tmp_path / "o.sam")
    r = subprocess.run([SIM, "classify", "-o", out, mixed["idx"], mixed["reads"], mixed["reads"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0 and r.stdout == b""
    assert open(out, "rb").read() == ref_cli([mixed["idx"] + "/", mixed["reads"], mixed["reads"]])      # two files: the history carries over


def _concurrent_calls(lib_kind, tiny, mixed_reads=None):
    """two caller threads, distinct thread_ids, the same resident index (reference contract: src/cly_mt.c:1279-1307)"""
    import threading
    import pydsb
    d = (pydsb_hostsim.Desamba() if lib_kind == "hostsim" else pydsb.Desamba())
    d.load_index(tiny["idx"])
    text = open(tiny["reads"], "rb").read()
    res = {}

    def work(tid):
        res[tid] = [d.read_classify(text, thread_id=tid, thread_num=1) for _ in range(2)]
    th = [threading.Thread(target=work, args=(t,)) for t in (11, 12, 13)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    return res


def test_concurrent_callers_with_distinct_thread_ids(tiny):
    """Each thread_id owns its session (slots, history): three concurrent callers get what a lone caller gets -- the first
    call equals the golden output, the second call (same reads again, history of the first) is the same for all three."""
    res = _concurrent_calls("hostsim", tiny)
    want = sam_records(open(os.path.join(GOLD, "tiny.sam"), "rb").read())
    for tid, (first, second) in res.items():
        assert sam_records(first) == want, tid
    assert len({second for _first, second in res.values()}) == 1


def test_non_acgt_lower_case_and_short_reads_equal_reference_hostsim(mixed, have_ref_binary, tmp_path):
    """N runs, lower case, IUPAC codes, junk characters and reads of 1 .. 60 bases: byte-identical to the reference binary."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    from conftest import odd_base_reads, ref_cli
    fq = str(tmp_path / "odd.fq")
    odd_base_reads(mixed["reads"], fq)
    for fmt in ("SAM", "SAM_FULL"):
        assert run_sim(mixed["idx"], [fq], fmt=fmt) == ref_cli(["-f", fmt, mixed["idx"] + "/", fq])


def test_text_shapes_of_the_api_input_equal_reference_library(mixed, have_ref_binary):
    """read_classify() on in-memory text of several shapes -- plain, with the trailing NUL callers pass (main_test_2.c:73),
    without a final newline, CRLF line ends, multi-line FASTQ with comments, blank lines -- equals the reference library's
    answer byte for byte.  FASTA is the one documented deviation: the reference gives every record slot its own kseq_t
    (src/cly_mt.c:29-43), forgets the '>' it has already consumed and loses every second FASTA record; here all records
    come out, and the ones the reference does return are identical."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdesamba.so")):
        pytest.skip("reference library absent")
    import pydsb
    from conftest import ref_resident_session
    lines = open(mixed["reads"], "rb").read().split(b"\n")
    recs = [lines[i:i + 4] for i in range(0, len(lines) - 1, 4)][:60]
    plain = b"".join(b"\n".join(r) + b"\n" for r in recs)

    def wrap(s, n):
        return b"\n".join(s[k:k + n] for k in range(0, len(s), n))
    multi = b"".join(r[0] + b" extra words\n" + wrap(r[1], 61) + b"\n+\n" + wrap(r[3], 61) + b"\n" for r in recs)
    fasta = b"".join(b">" + r[0][1:] + b"\n" + wrap(r[1], 70) + b"\n" for r in recs)
    cases = [plain, plain + b"\0", plain.rstrip(b"\n"), plain.replace(b"\n", b"\r\n"), multi, plain.replace(b"\n@", b"\n\n@", 5)]
    for text in cases:
        want = ref_resident_session(mixed["idx"], [text], 1, 65536)[0][0]
        d = pydsb_hostsim.Desamba()
        d.load_index(mixed["idx"])
        assert d.read_classify(text, thread_id=7) == want
    want = sam_records(ref_resident_session(mixed["idx"], [fasta], 1, 65536)[0][0])
    d = pydsb_hostsim.Desamba()
    d.load_index(mixed["idx"])
    got = sam_records(d.read_classify(fasta, thread_id=7))
    assert len(got) == 60 and len(want) == 30
    assert all(got[k] == v for k, v in want.items())


def test_rejected_fastq_records_follow_the_reference_pipeline(mixed, have_ref_binary):
    """A record kseq rejects (quality string of another length) ends the reference's read batch; reading carries on behind
    it, and only the third rejected record that OPENS a batch ends the run (one pipeline worker ends per empty batch,
    src/cly_mt.c:29-43,367,461, src/lib/kthread.c:136-141).  Same output as the reference library for rejected records in
    the middle, adjacent ones, at the start, and enough of them to end the run early."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdesamba.so")):
        pytest.skip("reference library absent")
    import pydsb
    from conftest import ref_resident_session
    lines = open(mixed["reads"], "rb").read().split(b"\n")
    recs = [lines[i:i + 4] for i in range(0, len(lines) - 1, 4)][:30]

    def text(bad, cut=()):
        out = []
        for i, r in enumerate(recs):
            q = r[3] + b"II" if i in bad else (r[3][:40] if i in cut else r[3])
            out.append(b"\n".join([r[0], r[1], r[2], q]) + b"\n")
        return b"".join(out)
    cases = [text({10}), text({10, 11}), text({0}), text({0, 1}), text({0, 1, 2}), text({10, 11, 12}), text({10, 11, 12, 13}),
             text({0, 5, 6, 20, 21}), text((), cut={7}), text(()) + b"@x\nACGTACGT\n+\n"]
    n_lines = []
    for t in cases:
        want = ref_resident_session(mixed["idx"], [t], 1, 65536)[0][0]
        d = pydsb_hostsim.Desamba()
        d.load_index(mixed["idx"])
        assert d.read_classify(t, thread_id=7) == want
        n_lines.append(want.count(b"\n"))
    assert n_lines[4] == 0 and n_lines[6] < n_lines[5] < n_lines[0]      # the early ends really happened


def test_thread_num_change_restarts_the_buffer_history(mixed, have_ref_binary):
    """The reference throws a thread_id's buffers away when the caller changes thread_num (src/cly_mt.c:1286-1295), so the
    buffer history of earlier calls ends there.  A call sequence that switches 1 -> 3 -> 1 on one thread_id: every
    thread_num == 1 call (the reproducible ones) equals the reference library's answer, the one after the switch included."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdesamba.so")):
        pytest.skip("reference library absent")
    import pydsb
    from conftest import fastq_slices, ref_call_sequence
    sl = fastq_slices(mixed["reads"], 5)
    calls = [(sl[0], 7, 1), (sl[1], 7, 1), (sl[2], 7, 3), (sl[3], 7, 1), (sl[4], 7, 1), (sl[1], 7, 1)]
    want = ref_call_sequence(mixed["idx"], calls)
    d = pydsb_hostsim.Desamba()
    d.load_index(mixed["idx"])
    for i, (fq, tid, tn) in enumerate(calls):
        got = d.read_classify(fq, thread_id=tid, thread_num=tn)
        if tn == 1:
            assert got == want[i], i


def test_max_read_l_counts_only_reads_that_had_a_chain(mixed, have_ref_binary, tmp_path):
    """The reference worker's max_read_l is raised only by reads that reach the filter WITH a chain (src/cly.c:2884-2885
    returns first, :2953 updates): after long reads without any hit, short reads still take the 2G filter.  A stream
    that opens with three 3 kb junk reads, then 100..509-base reads, a long real read, more short reads -- byte-identical
    to the reference binary, at several batch sizes (the class of a short read may hang on a batch still in flight)."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    from conftest import max_read_l_stream, ref_cli
    fq, fq2 = str(tmp_path / "a.fq"), str(tmp_path / "b.fq")
    max_read_l_stream(mixed["reads"], fq, junk_first=True)
    max_read_l_stream(mixed["reads"], fq2, junk_first=False)
    want = ref_cli([mixed["idx"] + "/", fq])
    other = ref_cli([mixed["idx"] + "/", fq2])
    # teeth: in the reference itself the short reads' records depend on whether the leading long reads had chains
    tail = lambda t: [l for l in t.split(b"\n") if l.startswith(b"s")]
    assert tail(want) != tail(other)
    for env in ({}, {"DSB_BATCH_READS": "7"}, {"DSB_BATCH_READS": "1"}, {"DSB_BATCH_READS": "64", "DSB_HOSTSIM_GPUS": "2"}):
        assert run_sim(mixed["idx"], [fq], env=env) == want, env
        assert run_sim(mixed["idx"], [fq2], env=env) == other, env


def test_exact_matches_of_2000_bases_do_not_leave_the_score_table(tiny, tmp_path):
    """SURVEY.md Q12 on the host twin of the kernel source: error-free 3 kb copies of the reference give seeds with exact
    matches >= 2000 bases; Q_MEM has 2000 entries and the reference reads past it (undefined).  Here the look-up is clamped:
    the run finishes, every read is reported on its own genome, and the bytes do not depend on the batching."""
    from conftest import exact_copy_reads
    fq = exact_copy_reads(tiny, tmp_path)
    out = run_sim(tiny["idx"], [fq])
    recs = [ln.split(b"\t") for ln in out.split(b"\n") if ln.startswith(b"exact")]
    assert {r[0] for r in recs} == {b"exact0", b"exact1", b"exact2"}
    assert all(r[2] != b"*" for r in recs if int(r[1]) in (0, 16))
    assert run_sim(tiny["idx"], [fq], env={"DSB_BATCH_READS": "1"}) == out


def test_meta_analysis_from_hit_records_equals_text_parser_and_reference(mixed, tiny, have_ref_binary):
    """meta_analysis() on the very block read_classify() returned answers from the hit records the batcher
    kept (no text parsed).  Same bytes as the text parser on a copy of the text, for both weights and several snapshot
    sizes, over a call sequence on one thread_id -- and the same as the unmodified reference library driven the same way."""
    from conftest import fastq_slices, ref_resident_session
    for meta_ds, n_slices in ((mixed, 3), (tiny, 2)):
        slices = fastq_slices(meta_ds["reads"], n_slices)
        for flag, snap in ((1, 65536), (0, 100), (1, 0)):
            d = pydsb_hostsim.Desamba()
            d.load_index(meta_ds["idx"])
            want = ref_resident_session(meta_ds["idx"], slices, flag, snap) if have_ref_binary else None
            for i, fq in enumerate(slices):
                sam, meta, hs, _tc, _tm = d.classify_then_analyse(fq, thread_id=5, flag=flag, max_snapshot_len=snap)
                assert d.meta_path == 1, "the hit-record path was not taken"
                assert (meta, hs) == d.meta_analysis(sam, thread_id=5, flag=flag, max_snapshot_len=snap)     # a copy: text parser
                if want:
                    assert (sam, meta, hs) == want[i], (flag, snap, i)


def test_meta_analysis_falls_back_to_the_text_when_the_records_do_not_describe_it(mixed):
    """The records are used only for a text whose every line has the usual shape and only for that very text: a read name
    that starts with '@' (the text parser would skip it as a header line), an empty name, a NUL inside a sequence, a text
    the caller changed -- all go through the text parser, and give what it gives on a copy."""
    recs = open(mixed["reads"], "rb").read().split(b"\n")
    first = b"\n".join(recs[:40]) + b"\n"
    cases = {
        "plain": (first, 1, False),
        "header-like name": (b"@" + first, 0, False),
        "text changed by the caller": (first, 0, True),
        "FASTA": (b"".join(b">" + recs[i][1:] + b"\n" + recs[i + 1] + b"\n" for i in range(0, 40, 4)), 1, False),
    }
    for what, (fq, path, tamper) in cases.items():
        d = pydsb_hostsim.Desamba()
        d.load_index(mixed["idx"])
        sam, meta, hs, _tc, _tm = d.classify_then_analyse(fq, thread_id=2, flag=1, max_snapshot_len=1000, tamper=tamper)
        assert d.meta_path == path, what
        if tamper:
            sam = b"X" + sam[1:]
        assert (meta, hs) == d.meta_analysis(sam, thread_id=2, flag=1, max_snapshot_len=1000), what


def test_meta_analysis_on_text_the_reference_cannot_parse_still_stops_like_the_text_parser(mixed, tmp_path):
    """An empty read name or a NUL inside a sequence give SAM lines the reference's getOneSAM crashes on; the text parser
    here stops with a message.  The hit records must not paper over that (they are not kept for such calls)."""
    recs = open(mixed["reads"], "rb").read().split(b"\n")
    cases = {"empty name": b"@\n" + b"\n".join(recs[1:40]) + b"\n",
             "NUL in a sequence": b"\n".join(recs[:5] + [recs[5][:100] + b"\0" + recs[5][101:]] + recs[6:40]) + b"\n"}
    code = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r); import pydsb_hostsim\n"
            "d = pydsb_hostsim.Desamba(); d.load_index(sys.argv[1])\n"
            "d.classify_then_analyse(open(sys.argv[2], 'rb').read(), thread_id=2)\n"
            "print('answered, path', d.meta_path)\n") % (os.path.join(ROOT, "desamba-so_b200"), os.path.join(ROOT, "tests", "hostsim"))
    for what, fq in cases.items():
        p = tmp_path / "in.fq"
        p.write_bytes(fq)
        r = subprocess.run([sys.executable, "-c", code, mixed["idx"], str(p)], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r.returncode != 0 and b"malformed SAM record" in r.stderr, (what, r.stdout, r.stderr[-300:])
