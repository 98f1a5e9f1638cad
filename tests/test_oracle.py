"""The CPU oracle (oracle/oracle_*.c) against known answers produced by the unmodified reference
(tests/golden/kat_tiny, written by oracle/kat_harness.c which includes the reference's cly.c), and the
committed golden SAM against the reference binary when it is present.  No GPU."""
import ctypes as C
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from conftest import GOLD, ROOT

KAT = os.path.join(GOLD, "kat_tiny")


@pytest.fixture(scope="module")
def orc(built):
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_port.so"))
    lib.orc_occ.restype = C.c_uint64
    lib.orc_occ.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint8)]
    lib.orc_hash64_1.restype = C.c_uint64
    lib.orc_hash64_1.argtypes = [C.c_uint64]
    lib.orc_hash64_2.restype = C.c_uint64
    lib.orc_hash64_2.argtypes = [C.c_uint64]
    lib.orc_lv_extd.restype = C.c_int32
    lib.orc_lv_extd.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
    lib.orc_exist_bits.restype = C.c_int64
    lib.orc_exist_bits.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.orc_islands.restype = C.c_int64
    lib.orc_islands.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.POINTER(C.c_uint32)]
    lib.orc_get_ref.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_int]
    lib.orc_mapq_tables.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_uint64]
    return lib


def load_kat(name):
    return open(os.path.join(KAT, name), "rb").read()


def parse_reads_kat(k):
    """reads.bin -> list of dict(L, bin, exist[2], strands[2] = (total_score, seeds[n,3]))"""
    raw = load_kat("reads.bin")
    n = struct.unpack_from("<Q", raw, 0)[0]
    o = 8
    out = []
    for _ in range(n):
        L = struct.unpack_from("<I", raw, o)[0]
        o += 4
        b = np.frombuffer(raw, dtype=np.uint8, count=2 * L, offset=o)
        o += 2 * L
        nk = L - k + 1
        ex = [np.frombuffer(raw, dtype=np.uint8, count=nk, offset=o + s * nk) for s in range(2)]
        o += 2 * nk
        strands = []
        for _s in range(2):
            total, ns = struct.unpack_from("<II", raw, o)
            o += 8
            seeds = np.frombuffer(raw, dtype=np.uint32, count=3 * ns, offset=o).reshape(ns, 3)
            o += 12 * ns
            strands.append((total, seeds))
        out.append(dict(L=L, bin=b, exist=ex, strands=strands))
    assert o == len(raw)
    return out


def test_hash_kat(orc):
    raw = load_kat("hash.bin")
    n = struct.unpack_from("<Q", raw, 0)[0]
    a = np.frombuffer(raw, dtype=np.uint64, offset=8).reshape(3, n)
    for i in range(n):
        assert orc.orc_hash64_1(int(a[0, i])) == int(a[1, i])
        assert orc.orc_hash64_2(int(a[0, i])) == int(a[2, i])


def test_lv_kat(orc):
    raw = load_kat("lv.bin")
    n = struct.unpack_from("<Q", raw, 0)[0]
    o = 8
    ln = np.frombuffer(raw, dtype=np.int32, count=n, offset=o); o += 4 * n
    rf = np.frombuffer(raw, dtype=np.uint8, count=12 * n, offset=o).reshape(n, 12).copy(); o += 12 * n
    qr = np.frombuffer(raw, dtype=np.uint8, count=12 * n, offset=o).reshape(n, 12).copy(); o += 12 * n
    res = np.frombuffer(raw, dtype=np.int32, count=n, offset=o)
    assert set(ln.tolist()) == set(range(13))
    for i in range(n):
        got = orc.orc_lv_extd(rf[i].ctypes.data, qr[i].ctypes.data, int(ln[i]))
        assert got == int(res[i]), (i, int(ln[i]), rf[i], qr[i])


def test_mapq_kat(orc):
    raw = load_kat("mapq.bin")
    l_ref = struct.unpack_from("<Q", raw, 0)[0]
    want = np.frombuffer(raw, dtype=np.int32, offset=8)
    q_mem = np.zeros(2000, dtype=np.int32)
    q_lv = np.zeros(400, dtype=np.int32)
    orc.orc_mapq_tables(q_mem.ctypes.data, q_lv.ctypes.data, 0.15, l_ref)
    assert (q_mem == want[:2000]).all()
    assert (q_lv == want[2000:]).all()


def test_occ_kat(orc, tiny):
    import pydsb
    ix = pydsb.read_index_files(tiny["idx"])
    raw = load_kat("occ.bin")
    n = struct.unpack_from("<Q", raw, 0)[0]
    o = 8
    r = np.frombuffer(raw, dtype=np.uint64, count=n, offset=o); o += 8 * n
    ci = np.frombuffer(raw, dtype=np.uint8, count=n, offset=o); o += n
    co = np.frombuffer(raw, dtype=np.uint8, count=n, offset=o); o += n
    res = np.frombuffer(raw, dtype=np.uint64, count=n, offset=o)
    bw = np.ascontiguousarray(ix["bwt_occ"])
    for i in range(n):
        c = C.c_uint8(int(ci[i]))
        got = orc.orc_occ(bw.ctypes.data, ix["dollar_pos"], int(r[i]), C.byref(c))
        assert c.value == int(co[i]) and got == int(res[i]), i


def test_getref_kat(orc, tiny):
    import pydsb
    ix = pydsb.read_index_files(tiny["idx"])
    raw = load_kat("getref.bin")
    n = struct.unpack_from("<Q", raw, 0)[0]
    o = 8
    off = np.frombuffer(raw, dtype=np.uint64, count=n, offset=o); o += 8 * n
    ln = np.frombuffer(raw, dtype=np.uint32, count=n, offset=o); o += 4 * n
    fw = np.frombuffer(raw, dtype=np.uint8, count=n, offset=o); o += n
    want = np.frombuffer(raw, dtype=np.uint8, count=16 * n, offset=o).reshape(n, 16)
    rb = np.ascontiguousarray(ix["ref_bin"])
    buf = np.zeros(16, dtype=np.uint8)
    for i in range(n):
        buf[:] = 0
        orc.orc_get_ref(rb.ctypes.data, buf.ctypes.data, int(off[i]), int(ln[i]), int(fw[i]))
        assert (buf == want[i]).all(), i


def test_exist_and_islands_kat(orc, tiny):
    import pydsb
    ix = pydsb.read_index_files(tiny["idx"])
    k = ix["k"]
    reads = parse_reads_kat(k)
    assert len(reads) >= 10
    ek0, ek1 = np.ascontiguousarray(ix["ek0"]), np.ascontiguousarray(ix["ek1"])
    for rd in reads:
        L = rd["L"]
        for s in range(2):
            codes = np.ascontiguousarray(rd["bin"][s * L:(s + 1) * L])
            ex = np.zeros(L - k + 1, dtype=np.uint8)
            nk = orc.orc_exist_bits(codes.ctypes.data, L, k, ek0.ctypes.data, ek1.ctypes.data, ix["ek_mask"], ex.ctypes.data)
            assert nk == L - k + 1
            assert (ex == rd["exist"][s]).all()
            seeds = np.zeros((L // 4 + 8, 3), dtype=np.uint32)
            total = C.c_uint32(0)
            ns = orc.orc_islands(ex.ctypes.data, nk, s, seeds.ctypes.data, C.byref(total))
            want_total, want_seeds = rd["strands"][s]
            assert ns == len(want_seeds) and total.value == want_total
            assert (seeds[:ns] == want_seeds).all()


def test_msort_model_matches_libc_qsort(orc):
    """glibc qsort here is the stable merge sort; the comparator of the final hit
    ordering (reference src/cly.c:53-63) is inconsistent (returns sum_score % 2 on ties)."""
    libc = C.CDLL(None)
    CMP = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p)
    CMPI = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint32, C.c_uint32)
    # without argtypes ctypes passes a Python int as a C int: a buffer above 4 GB (a grown heap) arrives truncated
    libc.qsort.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, CMP]
    libc.qsort.restype = None
    orc.orc_msort_idx.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, CMPI, C.c_void_p]
    orc.orc_msort_idx.restype = None
    rng = np.random.default_rng(5)
    for trial in range(300):
        n = int(rng.integers(1, 60))
        score = rng.integers(0, 6, size=n).astype(np.uint32)
        rec = np.zeros((n, 2), dtype=np.uint32)
        rec[:, 0] = score
        rec[:, 1] = np.arange(n)

        def cmp_ptr(a, b):
            sa = C.cast(a, C.POINTER(C.c_uint32))[0]
            sb = C.cast(b, C.POINTER(C.c_uint32))[0]
            if sa < sb:
                return 1
            if sa > sb:
                return -1
            return int(sa % 2)

        def cmp_idx(_ctx, i, j):
            if score[i] < score[j]:
                return 1
            if score[i] > score[j]:
                return -1
            return int(score[i] % 2)

        libc.qsort(rec.ctypes.data, n, 8, CMP(cmp_ptr))
        idx = np.zeros(n, dtype=np.uint32)
        tmp = np.zeros(n, dtype=np.uint32)
        orc.orc_msort_idx(idx.ctypes.data, tmp.ctypes.data, n, CMPI(cmp_idx), None)
        assert (rec[:, 1] == idx).all(), trial


@pytest.mark.parametrize("name", ["tiny", "mixed"])
def test_committed_golden_is_what_the_reference_prints(name, have_ref_binary, request):
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    meta = request.getfixturevalue(name)
    want = json.load(open(os.path.join(GOLD, name + ".meta.json")))
    got = hashlib.md5(open(meta["golden"], "rb").read()).hexdigest()
    assert got == want["golden_md5"]
    assert open(meta["golden"], "rb").read() == open(os.path.join(GOLD, name + ".sam"), "rb").read()
    assert meta["n_reads"] == want["n_reads"]
