"""Parity tests proper: the sm_100a kernels, called through the C-ABI (libdesamba.so / libdsb_cuda.so),
against the unmodified reference's output and against function-level known answers."""
import os
import struct
import subprocess

import numpy as np
import pytest

from conftest import GOLD, PKG, ROOT, sam_records, ref_resident_session, fastq_slices

pytestmark = pytest.mark.gpu
CLI = os.path.join(PKG, "bin", "deSAMBA")


@pytest.fixture(scope="module")
def api(built):
    import pydsb
    return pydsb


def run_cli(idx, reads, fmt="SAM", env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([CLI, "classify", "-g", "1", "-f", fmt, idx, reads], stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=e)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    return r.stdout


def parity_against_reference(name, max_undefined, env=None):
    """Classify the data set's parity reads on the GPU and compare with the unmodified reference at -t 1, reads in file
    order.  (1) same reads in the same order; (2) every read whose reference answer is DEFINED is identical; (3) on the
    reference-undefined reads (some instrumented build of the unmodified sources -- zero / pattern initialised automatic
    variables, -O1 frame layout: oracle/Makefile `twins` -- answers differently from the stock binary)
    our answer is counted against the answers of the four reference builds; (4) a twin that crashed gives no verdict on
    the reads it did not reach, and its coverage is part of the summary.  One JSON line per data set goes to
    gpurun_out/parity_summary.jsonl (copied to profiles/ per round)."""
    import json
    import dataset
    meta = dataset.build(name)
    reads = meta.get("parity_reads") or meta["reads"]
    out = run_cli(meta["idx"], reads, env=env)
    got, want = sam_records(out), sam_records(open(meta["golden"], "rb").read())
    assert list(got) == list(want), name                  # same reads, same order
    bad = {k for k in want if got[k] != want[k]}
    undefined = dataset.undefined_reads(meta)
    info = json.load(open(os.path.join(meta["dir"], "undefined.json")))
    builds = [want] + [sam_records(open(os.path.join(meta["dir"], t + ".sam"), "rb").read()) for t in dataset.TWINS]
    in_family = sum(1 for k in undefined if any(b.get(k) == got.get(k) for b in builds))   # we give one of the reference builds' answers
    line = dict(data=name, reads=len(want), undefined=len(undefined), differ_from_stock=len(bad),
                differ_outside_undefined=len(bad - undefined), undefined_matching_a_reference_build=in_family,
                twins={t: dict(reached=v["reached"], differ=v["differ"], rc=v["rc"]) for t, v in info["twins"].items()})
    print("PARITY " + json.dumps(line))
    od = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(od):
        with open(os.path.join(od, "parity_summary.jsonl"), "a") as fh:
            fh.write(json.dumps(line) + "\n")
    assert bad <= undefined, (name, sorted(bad - undefined)[:5])
    # on a reference-undefined read our answer must still be the answer of ONE of the four builds of the reference
    assert in_family == len(undefined), (name, in_family, len(undefined))
    assert len(undefined) <= max_undefined * len(want), (name, len(undefined))
    for t, v in info["twins"].items():                    # a twin that stopped early must still have judged most of the set
        assert v["reached"] >= 0.9 * len(want), (name, t, v)
    return meta, out


@pytest.mark.parametrize("name", ["tiny", "cfg2"])
def test_index_builder_equals_reference(name, have_ref_binary, tmp_path):
    """`deSAMBA index` on the GPU (csrc/dsb_build.cu) writes the ten index files byte for byte as the unmodified reference
    builder does from the same FASTA (reference src/idx.c:1255-1282) -- on the smoke set and on BASELINE.json configs[1]'s
    50 Mbp reference; the only bytes masked are the heap garbage after each name in .ref_i (tools/idxcmp.py)."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    import dataset
    import idxcmp
    p = dict(dataset.CONFIGS[name])
    rd, idx = dataset.build_index(p)
    out = str(tmp_path / "idx")
    r = subprocess.run([CLI, "index", os.path.join(rd, "ref.fa"), out], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    print(r.stderr.decode()[-1500:])
    assert idxcmp.compare(idx, out) == []


@pytest.mark.parametrize("name", ["tiny", "mixed"])
def test_cli_is_byte_identical_to_reference(name, request):
    meta = request.getfixturevalue(name)
    assert run_cli(meta["idx"], meta["reads"]) == open(os.path.join(GOLD, name + ".sam"), "rb").read()


def test_read_classify_api_equals_reference_and_keeps_history_across_calls(api, tiny):
    d = api.Desamba()
    d.load_index(tiny["idx"])
    text = open(tiny["reads"], "rb").read()
    recs = text.split(b"\n")
    half = len(recs) // 8 * 4
    a, b = b"\n".join(recs[:half]) + b"\n", b"\n".join(recs[half:])
    want = sam_records(open(os.path.join(GOLD, "tiny.sam"), "rb").read())
    assert sam_records(d.read_classify(a, thread_id=1) + d.read_classify(b, thread_id=1)) == want
    assert sam_records(d.read_classify(tiny["reads"], thread_id=2, thread_num=4)) == want
    assert api.load_cuda_layer().dsb_kernel_launch_count() >= 9


def test_fm_rank_kernel_equals_reference_occ(api, tiny):
    raw = open(os.path.join(GOLD, "kat_tiny", "occ.bin"), "rb").read()
    n = struct.unpack_from("<Q", raw, 0)[0]
    o = 8
    r = np.frombuffer(raw, dtype=np.uint64, count=n, offset=o); o += 8 * n
    ci = np.frombuffer(raw, dtype=np.uint8, count=n, offset=o); o += n
    co = np.frombuffer(raw, dtype=np.uint8, count=n, offset=o); o += n
    res = np.frombuffer(raw, dtype=np.uint64, count=n, offset=o)
    ix = api.DeviceIndex(tiny["idx"])
    got, c_out, _ms = ix.fm_rank(r, ci)
    ix.close()
    # every symbol, '#' included: map_seed's second LF walk (src/cly.c:776-781) does not stop at '#', so its rank is
    # consumed (rarely) and must equal the reference's stored checkpoint; '$' returns dollar_pos
    assert (c_out == co).all()
    assert (got == res).all()
    assert (co == 4).any() and (co < 4).any()


def test_exist_kernel_equals_reference_probes(api, tiny):
    from test_oracle import parse_reads_kat
    ix = api.DeviceIndex(tiny["idx"])
    k = ix.ek_len()
    for rd in parse_reads_kat(k):
        L = rd["L"]
        for s in range(2):
            got, _ms = ix.exist(rd["bin"][s * L:(s + 1) * L])
            assert (got == rd["exist"][s]).all()
    ix.close()


def test_batch_size_and_rerun_do_not_change_results(mixed):
    want = open(os.path.join(GOLD, "mixed.sam"), "rb").read()
    assert run_cli(mixed["idx"], mixed["reads"], env={"DSB_BATCH_READS": "37"}) == want
    assert run_cli(mixed["idx"], mixed["reads"], env={"DSB_BATCH_BASES": "30000"}) == want
    assert run_cli(mixed["idx"], mixed["reads"], env={"DSB_BLOCKS_PER_SM": "1"}) == want


def test_max_read_l_counts_only_reads_that_had_a_chain(mixed, have_ref_binary, tmp_path):
    """GPU form of tests/test_host.py's test of the same name (reference src/cly.c:2884-2885,2953-2981)."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    from conftest import max_read_l_stream, ref_cli
    for junk_first in (True, False):
        fq = str(tmp_path / ("m%d.fq" % junk_first))
        max_read_l_stream(mixed["reads"], fq, junk_first=junk_first)
        want = ref_cli([mixed["idx"] + "/", fq])
        for env in ({}, {"DSB_BATCH_READS": "7"}, {"DSB_BATCH_READS": "1"}):
            assert run_cli(mixed["idx"], fq, env=env) == want, (junk_first, env)


def test_job_list_and_anchor_pool_regrow_and_rerun(mixed):
    """The split pipeline's batch-wide job list and anchor pool start far too small (DSB_TEST_TINY_ARENAS: 64 jobs, 256
    anchors): k_islands / k_seed flag the overflow, k_seed / k_classify stand down, the host doubles both until the batch
    fits and re-runs it; the bytes are the golden's."""
    want = open(os.path.join(GOLD, "mixed.sam"), "rb").read()
    assert run_cli(mixed["idx"], mixed["reads"], env={"DSB_TEST_TINY_ARENAS": "1"}) == want
    assert run_cli(mixed["idx"], mixed["reads"], env={"DSB_TEST_TINY_ARENAS": "1", "DSB_BATCH_READS": "50"}) == want


@pytest.mark.parametrize("name", ["tiny", "mixed"])
def test_single_kernel_form_equals_split_pipeline(name, request):
    """DSB_SPLIT=0 (everything in k_classify) and the default k_islands -> k_seed -> k_classify give the same bytes."""
    meta = request.getfixturevalue(name)
    want = open(os.path.join(GOLD, name + ".sam"), "rb").read()
    assert run_cli(meta["idx"], meta["reads"], env={"DSB_SPLIT": "0"}) == want
    assert run_cli(meta["idx"], meta["reads"], env={"DSB_SPLIT": "1", "DSB_BATCH_BASES": "300000"}) == want


def test_cfg2_full_size_against_reference(have_ref_binary):
    """BASELINE.json configs[1] at full size: 10 000 x 5 kb reads on the 50 Mbp index (rule: parity_against_reference)."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    meta, out = parity_against_reference("cfg2", 0.01)
    # idempotence / batch independence at full size
    assert run_cli(meta["idx"], meta["reads"], env={"DSB_BATCH_READS": "1500"}) == out


def test_cfg3_headline_index_against_reference(have_ref_binary):
    """BASELINE.json configs[2] on the bench's own index: 1 Gbp reference (200 families x 5 strains x 1 Mbp, BWT ~0.95 G
    rows, exist tables 2 x 256 MiB / k = 17), built on this box by OUR index builder, ONT-like 10 kb reads; the first
    20 000 reads of the bench set against the unmodified reference's `classify -t 1` on the same index."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    meta, out = parity_against_reference("cfg3", 0.05)
    placed_where_simulated(meta, out, 0.9)


def placed_where_simulated(meta, sam_text, min_share):
    """The C simulator names a read r<i>_g<genome>_<start>_<strand>: the PRIMARY record of a classified read must name that
    genome's reference (or a strain of the same family: strains differ by 1 % SNPs, ties between them are legitimate) at the
    simulated position (+-10 % of the read length for indels).  A check of the INDEX's content that parity with the reference
    cannot give, because both sides read the same index files."""
    import re
    ok = n = 0
    for line in sam_text.decode().split("\n"):
        f = line.split("\t")
        if len(f) < 6 or (int(f[1]) & 0x900):
            continue
        m = re.match(r"r\d+_g(\d+)_(\d+)_([+-])$", f[0])
        if not m or f[2] == "*":
            continue
        g, st = int(m.group(1)), int(m.group(2))
        r = re.search(r"SYN_(\d+)", f[2])
        n += 1
        clip = re.match(r"(-?\d+)S", f[5])
        lead = int(clip.group(1)) if clip else 0
        if r and int(r.group(1)) // 5 == g // 5 and abs(int(f[3]) - lead * (m.group(3) == "+") - st) <= 2000:
            ok += 1
    assert n > 0
    print("placed where simulated: %d of %d primary records (%.1f %%)" % (ok, n, 100.0 * ok / n))
    assert ok >= min_share * n, (ok, n)


def test_cfg3x_full_scale_index_against_reference(have_ref_binary):
    """BASELINE.json configs[2] at its own index size: 4 Gbp reference (800 families x 5 strains x 1 Mbp), BWT 3.8 G rows (the
    reference's FM array is 2.5 GB, the SA samples 3.8 GB, the exist tables 2 x 1 GiB with k = 18), built on this box by OUR
    index builder in seconds; 5 000 ONT-like 10 kb reads against the unmodified reference's `classify -t 1` on that index, and
    the reads must land where they were simulated from."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    import shutil
    if shutil.disk_usage("/tmp").free < (40 << 30):
        pytest.skip("less than 40 GB free under /tmp")
    meta, out = parity_against_reference("cfg3x", 0.05)
    placed_where_simulated(meta, out, 0.9)


def test_length_error_sweep_against_reference(have_ref_binary):
    """BASELINE.json configs[3] at test scale, all 24 cells: reads of 1 / 2 / 5 / 10 / 20 / 50 kb at 5 / 10 / 15 / 20 % error
    (indel-heavy mix, +-30 % length jitter, 5 % junk reads) on the cfg2 index, ~3 Mbases per cell.  The per-cell table
    (reads / undefined / differing / matching a reference build) goes to gpurun_out/parity_summary.jsonl."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    import dataset
    dataset.build_many(dataset.SWEEP)
    for name in dataset.SWEEP:
        # long noisy reads cross more windows whose reference answer reads uninitialised stack
        L = dataset.CONFIGS[name]["read_len"]
        parity_against_reference(name, 0.02 if L <= 2000 else 0.10 if L <= 10000 else 0.45)


def test_resident_index_repeated_batches_with_meta_analysis(api, mixed, have_ref_binary):
    """BASELINE.json configs[4] at test scale: ONE load_index, then many read_classify() calls on consecutive slices
    of the read set on one thread_id, each followed by meta_analysis(META_USE_BASE_NUM, 65536).  The reference's
    buffer history carries across calls, so the whole call sequence is compared with the unmodified
    reference library driven the same way: SAM_FULL text, the analysis lines and the human snapshot of every call.
    Then the same session keeps going (300 more calls): outputs repeat the history-free steady state and neither
    device memory nor the host's resident set grows."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdesamba.so")):
        pytest.skip("reference library absent")
    import psutil
    import torch
    slices = fastq_slices(mixed["reads"], 10) * 2            # 20 calls: the second pass starts from the first pass's history
    want = ref_resident_session(mixed["idx"], slices, 1, 65536)
    d = api.Desamba()
    d.load_index(mixed["idx"])
    for i, (fq, (w_sam, w_meta, w_snap)) in enumerate(zip(slices, want)):
        sam = d.read_classify(fq, thread_id=7, thread_num=1)
        assert sam_records(sam) == sam_records(w_sam), i
        assert sam == w_sam, i                                # SAM_FULL, byte for byte (SEQ / QUAL on primaries included)
        meta, snap = d.meta_analysis(sam, thread_id=7, flag=1, max_snapshot_len=65536)
        assert meta == w_meta, (i, meta, w_meta)
        assert snap == w_snap, i
    # the same call sequence as a C caller writes it: meta_analysis() on the very block read_classify() returned, which is
    # answered from the hit records the batcher kept -- same bytes as the reference library again
    d2 = api.Desamba()
    d2.load_index(mixed["idx"])
    for i, (fq, (w_sam, w_meta, w_snap)) in enumerate(zip(slices, want)):
        sam, meta, snap, _tc, _tm = d2.classify_then_analyse(fq, thread_id=9, flag=1, max_snapshot_len=65536)
        assert d2.meta_path == 1, i
        assert (sam, meta, snap) == (w_sam, w_meta, w_snap), i
    proc = psutil.Process()
    free0, rss0 = None, None
    period = slices[10:]
    ref_out = None
    for rep in range(30):
        outs = []
        for fq in period:
            sam = d.read_classify(fq, thread_id=7, thread_num=1)
            outs.append((sam, d.meta_analysis(sam, thread_id=7, flag=1, max_snapshot_len=65536)))
        if rep == 0:
            ref_out = outs                                     # history is periodic from the second pass on
            free0, rss0 = torch.cuda.mem_get_info(0)[0], proc.memory_info().rss
        else:
            assert outs == ref_out, rep
    free1, rss1 = torch.cuda.mem_get_info(0)[0], proc.memory_info().rss
    assert free0 - free1 <= (8 << 20), (free0, free1)          # device memory: no growth over 290 calls
    assert rss1 - rss0 <= (64 << 20), (rss0, rss1)             # host: allocator noise only


@pytest.mark.parametrize("name", ["tiny", "mixed", "cfg2"])
def test_exist_on_demand_equals_exist_bytes(name, request):
    """DSB_EXIST_ON_DEMAND=1: no k_exist pass, the island scan of k_islands (and of the slow path in k_classify) probes the
    exist tables itself, half a warp per strand.  A probe is a pure function of the k-mer, so the output must not change."""
    request.getfixturevalue("tiny")                 # the build and the reference binary
    import dataset
    ds = dataset.build(name)
    assert run_cli(ds["idx"], ds["reads"], env={"DSB_EXIST_ON_DEMAND": "1"}) == run_cli(ds["idx"], ds["reads"])


@pytest.mark.parametrize("name", ["mixed", "cfg2"])
def test_slow_path_serial_fallback_equals_lane_per_island(name, request):
    """The slow path (reads without a usable chain) runs one lane per island; a lane that would run out of MEM slots sends
    its strand through the warp-serial form instead.  DSB_TEST_SLOW_LANE_MEMS=8 makes that happen for every island with a
    second search that follows a hit; the output must not change."""
    import re
    request.getfixturevalue("tiny")                 # the build and the reference binary
    import dataset
    ds = dataset.build(name)
    e = dict(os.environ, DSB_TEST_SLOW_LANE_MEMS="8")
    r = subprocess.run([CLI, "classify", "-g", "1", "-f", "SAM", ds["idx"], ds["reads"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=e)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    m = re.search(r"slow-path reads (\d+) \(strands through the serial form: (\d+)\)", r.stderr.decode())
    assert m and int(m.group(1)) > 0 and int(m.group(2)) > 0, r.stderr.decode()[-600:]      # the fallback did run
    assert r.stdout == run_cli(ds["idx"], ds["reads"])


def test_device_reblocking_equals_host_reblocking(tiny):
    """Index upload: k_fm_reblock (device) writes the same FM blocks, byte for byte, as the host re-blocking function
    that the host simulator uses (DSB_VERIFY_REBLOCK makes dsb_index_upload compare them and fail otherwise)."""
    e = dict(os.environ, DSB_VERIFY_REBLOCK="1")
    r = subprocess.run([CLI, "classify", "-g", "1", "-f", "SAM", tiny["idx"], tiny["reads"]], stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, env=e)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    assert b"k_fm_reblock verified" in r.stderr
    assert r.stdout == open(os.path.join(GOLD, "tiny.sam"), "rb").read()


def test_two_gpus_of_one_process_equal_one_gpu(mixed):
    """index replicated (GPU 1 gets it by peer copy from GPU 0), batches alternate between the devices,
    output order and bytes unchanged.  Needs a box with two GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU on this box")
    want = open(os.path.join(GOLD, "mixed.sam"), "rb").read()
    for env in ({"DSB_BATCH_READS": "23"}, {"DSB_BATCH_READS": "23", "DSB_INDEX_CLONE": "0"}):
        e = dict(os.environ, **env)
        r = subprocess.run([CLI, "classify", "-g", "2", "-f", "SAM", mixed["idx"], mixed["reads"]], stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, env=e)
        assert r.returncode == 0, r.stderr.decode()[-2000:]
        assert r.stdout == want


def test_cfg1_demo_standin_against_reference(have_ref_binary):
    """BASELINE.json configs[0] (`classify demo_index/human demo/human.fq`): the demo files are not in the reference tree
    (.gitignore:3-4), so the stand-in is a synthetic 5 x 1 Mbp "human" (tid 9606) contig set with 2 000 x 5 kb reads
    (SURVEY 8c/8d).  Same rule as cfg2; the SAM_FULL text feeds meta_analysis, whose human snapshot must be non-empty."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    import dataset
    import pydsb
    meta, _out = parity_against_reference("cfg1", 0.02)
    d = pydsb.Desamba()
    d.load_index(meta["idx"])
    sam = d.read_classify(meta["reads"], thread_id=0, thread_num=1)
    text, snap = d.meta_analysis(sam, thread_id=0, flag=1, max_snapshot_len=65536)
    assert text and snap


def test_cli_options_and_formats_equal_reference(mixed, have_ref_binary, tmp_path):
    """Every output format and the -r / -s / -l options of `deSAMBA classify`, and -o FILE with two input files (the buffer
    history carries from one file into the next): byte-identical to the reference binary run on the box."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    from conftest import CLI_VARIANTS, ref_cli
    for opts in CLI_VARIANTS:
        want = ref_cli(opts + [mixed["idx"] + "/", mixed["reads"]])
        r = subprocess.run([CLI, "classify", "-g", "1"] + opts + [mixed["idx"], mixed["reads"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r.returncode == 0, r.stderr.decode()[-1500:]
        assert r.stdout == want, opts
    out = str(tmp_path / "o.sam")
    r = subprocess.run([CLI, "classify", "-g", "1", "-o", out, mixed["idx"], mixed["reads"], mixed["reads"]], stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE)
    assert r.returncode == 0 and r.stdout == b""
    assert open(out, "rb").read() == ref_cli([mixed["idx"] + "/", mixed["reads"], mixed["reads"]])


def test_concurrent_callers_with_distinct_thread_ids(tiny):
    """Three caller threads with their own thread_ids on one resident index (reference contract: src/cly_mt.c:1279-1307):
    each gets what a lone caller gets."""
    from test_host import _concurrent_calls
    res = _concurrent_calls("cuda", tiny)
    want = sam_records(open(os.path.join(GOLD, "tiny.sam"), "rb").read())
    for tid, (first, second) in res.items():
        assert sam_records(first) == want, tid
    assert len({second for _first, second in res.values()}) == 1


@pytest.mark.parametrize("table_bytes,k,bits", [(0x10000000, 17, 31), (0x40000000, 18, 33), (0x100000000, 19, 35)])
def test_exist_kernel_on_larger_table_classes(api, built, table_bytes, k, bits):
    """The exist-k-mer tables of larger indexes (reference src/idx.c:966-982: 256 MiB -> k 17 / 31-bit mask, 1 GiB -> k 18 /
    33 bits, 4 GiB -> k 19 / 35 bits, i.e. bit offsets beyond 2^32): k_exist against the oracle's probe on a synthetic
    table (random 64 MiB tile repeated; the real tables of such indexes cannot be built on the box)."""
    import ctypes
    orc = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_port.so"))
    orc.orc_exist_bits.restype = ctypes.c_int64
    orc.orc_exist_bits.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                   ctypes.c_void_p]
    rng = np.random.default_rng(table_bytes & 0xffff)
    tile = rng.integers(0, 256, size=1 << 26, dtype=np.uint8)
    tile &= rng.integers(0, 256, size=1 << 26, dtype=np.uint8)          # a quarter of the bits set
    table = np.tile(tile, table_bytes >> 26)
    for j in range(0, table_bytes, 1 << 26):                             # make the tiles differ
        table[j + (j >> 26) % 4096] ^= 0xff
    ix = api.synthetic_fm_index(1 << 16, ek_table=table)
    assert ix.ek_len() == k
    codes = rng.integers(0, 4, size=300000, dtype=np.uint8)
    codes[1000:1040] = 2                                                 # a low-complexity stretch
    got, _ms = ix.exist(codes)
    want = np.zeros(len(codes) - k + 1, dtype=np.uint8)
    n = orc.orc_exist_bits(codes.ctypes.data, len(codes), k, table.ctypes.data, table.ctypes.data, (1 << bits) - 1, want.ctypes.data)
    ix.close()
    assert n == len(want)
    assert (got == want).all()
    assert 0.005 < want.mean() < 0.5


def test_fm_rank_on_rows_beyond_2_to_32(api, built):
    """A 4 Gbp index has more than 2^32 BWT rows: k_fm_reblock + k_fm_rank on a synthetic 2^33-symbol BWT (the reference's
    168-byte block layout, random symbols) against the oracle's occ on the same blocks, rows drawn from the whole range."""
    import ctypes
    orc = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_port.so"))
    orc.orc_occ.restype = ctypes.c_uint64
    orc.orc_occ.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint8)]
    n_sym = 1 << 33
    ix = api.synthetic_fm_index(n_sym, seed=11)
    rng = np.random.default_rng(12)
    rows = np.concatenate([rng.integers(0, n_sym - 64, size=20000, dtype=np.uint64),
                           rng.integers((1 << 32) - 300, (1 << 32) + 300, size=2000, dtype=np.uint64)])
    c = rng.integers(0, 4, size=len(rows), dtype=np.uint8)
    c[::7] = 0xff                                                        # "the symbol at this row"
    got, c_out, _ms = ix.fm_rank(rows, c)
    blocks = ix.keep["blocks"]
    for i in range(len(rows)):
        cc = ctypes.c_uint8(int(c[i]))
        want = orc.orc_occ(blocks.ctypes.data, n_sym + 5, int(rows[i]), ctypes.byref(cc))
        assert cc.value == int(c_out[i]) and want == int(got[i]), (i, int(rows[i]))
    ix.close()
    assert (rows > (1 << 32)).sum() > 5000


def test_non_acgt_lower_case_and_short_reads_equal_reference(mixed, have_ref_binary, tmp_path):
    """N runs, lower case, IUPAC codes, junk characters and reads of 1 .. 60 bases: byte-identical to the reference binary."""
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    from conftest import odd_base_reads, ref_cli
    fq = str(tmp_path / "odd.fq")
    odd_base_reads(mixed["reads"], fq)
    for fmt in ("SAM", "SAM_FULL"):
        assert run_cli(mixed["idx"], fq, fmt=fmt) == ref_cli(["-f", fmt, mixed["idx"] + "/", fq])


def test_exact_matches_of_2000_bases_do_not_leave_the_score_table(tiny, tmp_path):
    """SURVEY.md Q12 on the device (the host twin: tests/test_host.py): exact matches >= 2000 bases index Q_MEM[1999]
    instead of running past the table; the run finishes, reports every read, and does not depend on the batching."""
    from conftest import exact_copy_reads
    fq = exact_copy_reads(tiny, tmp_path)
    out = run_cli(tiny["idx"], fq)
    recs = [ln.split(b"\t") for ln in out.split(b"\n") if ln.startswith(b"exact")]
    assert {r[0] for r in recs} == {b"exact0", b"exact1", b"exact2"}
    assert all(r[2] != b"*" for r in recs if int(r[1]) in (0, 16))
    assert run_cli(tiny["idx"], fq, env={"DSB_BATCH_READS": "1"}) == out
