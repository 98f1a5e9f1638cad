"""ctypes bindings of the two C-ABI layers, for tests/, bench.py and __graft_entry__.py only.

  * libdesamba.so  -- the reference-facing API (include/desamba.h == reference desamba.h:10,23,43)
  * libdsb_cuda.so -- the thin CUDA C-ABI under it (include/dsb_cuda.h)

Python is a driver here, not part of the product: the host side of the product is C
(desamba-so_b200/host/*.c), as in the reference.  Loading fails loudly when the libraries are not
built; there is no fallback of any kind.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.environ.get("DSB_LIB_DIR") or os.path.join(HERE, "lib")     # DSB_LIB_DIR: A/B builds of the CUDA layer

API_SYMBOLS = ["load_index", "read_classify", "meta_analysis"]
CUDA_SYMBOLS = ["dsb_gpu_count", "dsb_index_upload", "dsb_index_clone", "dsb_index_free", "dsb_index_device_bytes", "dsb_slot_create",
                "dsb_slot_free", "dsb_slot_staging", "dsb_slot_submit", "dsb_slot_wait", "dsb_slot_timing",
                "dsb_slot_rerun_device", "dsb_last_error", "dsb_kernel_launch_count", "dsb_fm_rank_batch",
                "dsb_exist_batch", "dsb_exist_batch_static", "dsb_index_build", "dsb_build_last_error"]
HOST_SYMBOLS = ["dsb_index_load", "dsb_index_release_bulk", "dsb_index_destroy", "dsb_engine_create",
                "dsb_engine_destroy", "dsb_session_create", "dsb_session_destroy", "dsb_session_run",
                "dsb_fq_open_path", "dsb_fq_open_mem", "dsb_fq_close", "dsb_history_new", "dsb_history_next",
                "dsb_history_free", "dsb_mapq_tables", "dsb_format_read"]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "exist_probes0", "exist_probes1", "occ_queries", "lut_lookups", "sa_lookups", "unitig_reads", "refpos_reads",
        "ref_window_bytes", "mem_searches", "map_seeds", "sdp_windows", "anchors", "slow_reads")] + [
        ("phase_cycles", C.c_uint64 * 8), ("seed_occ_queries", C.c_uint64), ("seed_alg_bytes", C.c_uint64), ("slow_serial_strands", C.c_uint64)]

    def as_dict(self):
        d = {n: int(getattr(self, n)) for n, _ in self._fields_ if n != "phase_cycles"}
        d["phase_cycles"] = [int(x) for x in self.phase_cycles]
        return d


class ReadDesc(C.Structure):
    _fields_ = [("seq_off", C.c_uint64), ("len", C.c_uint32), ("flags", C.c_uint32)]


class Hit(C.Structure):
    _fields_ = [("ref_ID", C.c_uint32), ("t_st", C.c_uint32), ("t_ed", C.c_uint32), ("q_st", C.c_uint32),
                ("q_ed", C.c_uint32), ("sum_score", C.c_uint32), ("indel", C.c_uint32), ("anchor_number", C.c_uint32),
                ("direction", C.c_uint8), ("primary", C.c_uint8), ("pri_index", C.c_uint8), ("with_top_anchor", C.c_uint8)]


class ReadResult(C.Structure):
    _fields_ = [("hit_off", C.c_uint32), ("n_hit", C.c_uint32), ("fast_classify", C.c_uint32), ("n_anchor", C.c_uint32)]


class RunStats(C.Structure):
    _fields_ = [("n_reads", C.c_uint64), ("n_bases", C.c_uint64), ("n_batches", C.c_uint64), ("t_parse", C.c_double),
                ("t_gpu_wait", C.c_double), ("t_format", C.c_double), ("t_total", C.c_double), ("counters", Counters),
                ("gpu_ms_total", C.c_double), ("gpu_ms_exist", C.c_double), ("gpu_ms_classify", C.c_double)]


def lib_path(name):
    return os.path.join(LIB_DIR, name)


def _load(path):
    if not os.path.exists(path):
        raise RuntimeError("%s is not built (run __graft_entry__.build())" % path)
    return C.CDLL(path)


def load_cuda_layer(path=None):
    """libdsb_cuda.so, the sm_100a kernels behind include/dsb_cuda.h.  There is no other implementation in this package
    (the tests' host simulator binds its own libraries from tests/hostsim/pydsb_hostsim.py)."""
    lib = _load(path or lib_path("libdsb_cuda.so"))
    lib.dsb_last_error.restype = C.c_char_p
    lib.dsb_index_device_bytes.restype = C.c_uint64
    lib.dsb_build_last_error.restype = C.c_char_p
    return lib


def build_index(ref_fasta, out_dir, kmer_srt=None, dev=0, _cuda_lib=None):
    """`deSAMBA index` through the C-ABI (dsb_index_build): returns the stats as a dict."""
    class Stats(C.Structure):
        _fields_ = [("n_ref", C.c_uint64), ("n_bases", C.c_uint64), ("n_kmer", C.c_uint64), ("n_unitig", C.c_uint64),
                    ("n_rows", C.c_uint64), ("n_ref_pos", C.c_uint64), ("exist_table_bytes", C.c_uint64),
                    ("exist_k", C.c_int32), ("kernel_launches", C.c_int32), ("peak_device_bytes", C.c_uint64),
                    ("seconds", C.c_double), ("n_phases", C.c_int32), ("phase_name", (C.c_char * 40) * 32),
                    ("phase_s", C.c_double * 32)]
    cuda = load_cuda_layer(_cuda_lib)
    st = Stats()
    rc = cuda.dsb_index_build(ref_fasta.encode(), kmer_srt.encode() if kmer_srt else None, out_dir.encode(), dev, C.byref(st))
    if rc != 0:
        raise RuntimeError("dsb_index_build: " + cuda.dsb_build_last_error().decode())
    d = {n: getattr(st, n) for n, _ in Stats._fields_[:11]}
    d["phases"] = {st.phase_name[i].value.decode(): st.phase_s[i] for i in range(st.n_phases)}
    return d


class Desamba:
    """The three-symbol API, exactly as a C caller would use it (reference main_test.c:29-57)."""

    API_LIB = None          # overridden only by tests/hostsim/pydsb_hostsim.py
    CUDA_LIB = None

    def __init__(self):
        load_cuda_layer(self.CUDA_LIB)
        self.lib = _load(self.API_LIB or lib_path("libdesamba.so"))
        self.libc = C.CDLL(None)
        self.libc.free.argtypes = [C.c_void_p]
        self.lib.load_index.argtypes = [C.POINTER(C.c_void_p), C.c_char_p]
        self.lib.read_classify.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_uint64), C.c_int, C.c_int]
        self.lib.meta_analysis.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_uint64), C.c_int, C.c_int, C.c_uint64, C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_uint64)]
        self.idx = C.c_void_p()

    def load_index(self, path):
        self.lib.load_index(C.byref(self.idx), path.encode())

    def read_classify(self, data, thread_id=0, thread_num=1):
        """data: bytes of FASTQ/FASTA text, or a str path (input_n = (uint64)-1)."""
        out, out_n = C.c_void_p(), C.c_uint64(0)
        if isinstance(data, str):
            self.lib.read_classify(self.idx, data.encode(), C.c_uint64(2 ** 64 - 1), C.byref(out), C.byref(out_n),
                                   thread_id, thread_num)
        else:
            self.lib.read_classify(self.idx, data, len(data), C.byref(out), C.byref(out_n), thread_id, thread_num)
        if not out.value:
            return b""
        text = C.string_at(out.value, out_n.value)
        self.libc.free(out)
        return text

    def read_classify_timed(self, data, thread_id=0, thread_num=1):
        """What a C caller does and nothing else: read_classify() on host text, then free() of the result.
        Returns (seconds, output bytes); the text is not copied into Python."""
        import time
        out, out_n = C.c_void_p(), C.c_uint64(0)
        t0 = time.perf_counter()
        self.lib.read_classify(self.idx, data, len(data), C.byref(out), C.byref(out_n), thread_id, thread_num)
        if out.value:
            self.libc.free(out)
        return time.perf_counter() - t0, out_n.value

    def classify_then_analyse(self, data, thread_id=0, thread_num=1, flag=1, max_snapshot_len=65536, calls=1, tamper=False):
        """The resident loop of BASELINE.json configs[4] as a C caller writes it: read_classify(), then meta_analysis()
        on the block it returned (the same pointer), then free().  meta_analysis() then answers from the hit records
        read_classify() kept.  Returns (sam text, meta text, snapshot, seconds classify, seconds meta)."""
        import time
        out, out_n = C.c_void_p(), C.c_uint64(0)
        t0 = time.perf_counter()
        self.lib.read_classify(self.idx, data, len(data), C.byref(out), C.byref(out_n), thread_id, thread_num)
        t1 = time.perf_counter()
        if not out.value:
            return b"", b"", None, t1 - t0, 0.0
        if tamper:                                       # tests: the caller changed the text before handing it back
            C.memmove(out.value, b"X", 1)
        meta = snap = None
        for _ in range(calls):
            mo, mo_n, hs, hs_n = C.c_void_p(), C.c_uint64(0), C.c_void_p(), C.c_uint64(0)
            self.lib.meta_analysis(self.idx, C.cast(out, C.c_char_p), out_n, C.byref(mo), C.byref(mo_n), thread_id, flag,
                                   max_snapshot_len, C.byref(hs), C.byref(hs_n))
            meta = C.string_at(mo.value, mo_n.value) if mo.value else b""
            snap = C.string_at(hs.value, hs_n.value) if hs.value else None
            if mo.value:
                self.libc.free(mo)
            if hs.value:
                self.libc.free(hs)
        t2 = time.perf_counter()
        self.meta_path = int(self.lib.dsb_meta_last_path())      # 1: from the hit records, 0: text parser
        sam = C.string_at(out.value, out_n.value)
        self.libc.free(out)
        return sam, meta, snap, t1 - t0, (t2 - t1) / calls

    def meta_analysis(self, sam, thread_id=0, flag=1, max_snapshot_len=65536):
        out, out_n, hs, hs_n = C.c_void_p(), C.c_uint64(0), C.c_void_p(), C.c_uint64(0)
        self.lib.meta_analysis(self.idx, sam, len(sam), C.byref(out), C.byref(out_n), thread_id, flag, max_snapshot_len,
                               C.byref(hs), C.byref(hs_n))
        text = C.string_at(out.value, out_n.value) if out.value else b""
        snap = C.string_at(hs.value, hs_n.value) if hs.value else None
        if out.value:
            self.libc.free(out)
        if hs.value:
            self.libc.free(hs)
        return text, snap


class DeviceIndex:
    """Index loaded by the C host loader and made resident through the CUDA C-ABI (or the host simulator)."""

    API_LIB = None
    CUDA_LIB = None

    def __init__(self, index_dir, dev=0):
        import numpy as np
        self.np = np
        self.cuda = load_cuda_layer(self.CUDA_LIB)
        self.host = _load(self.API_LIB or lib_path("libdesamba.so"))
        self.buf = C.create_string_buffer(1 << 16)          # dsb_loaded_index (its first member is dsb_host_index)
        self.host.dsb_index_load(self.buf, index_dir.encode(), 170, 64)
        self.handle = C.c_void_p()
        rc = self.cuda.dsb_index_upload(self.buf, dev, C.byref(self.handle))
        if rc != 0:
            raise RuntimeError("dsb_index_upload: " + self.cuda.dsb_last_error().decode())

    def close(self):
        if self.handle:
            self.cuda.dsb_index_free(self.handle)
            self.handle = C.c_void_p()
            self.host.dsb_index_destroy(self.buf)

    def fm_rank(self, rows, c, repeat=1):
        np = self.np
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        c = np.ascontiguousarray(c, dtype=np.uint8)
        out = np.empty(len(rows), dtype=np.uint64)
        c_out = np.empty(len(rows), dtype=np.uint8)
        ms = C.c_float(0)
        rc = self.cuda.dsb_fm_rank_batch(self.handle, rows.ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p),
                                         C.c_uint64(len(rows)), out.ctypes.data_as(C.c_void_p),
                                         c_out.ctypes.data_as(C.c_void_p), repeat, C.byref(ms))
        if rc != 0:
            raise RuntimeError("dsb_fm_rank_batch: " + self.cuda.dsb_last_error().decode())
        return out, c_out, ms.value

    def exist(self, codes, repeat=1):
        np = self.np
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        k = self.ek_len()
        out = np.empty(len(codes) - k + 1, dtype=np.uint8)
        ms = C.c_float(0)
        rc = self.cuda.dsb_exist_batch(self.handle, codes.ctypes.data_as(C.c_void_p), C.c_uint64(len(codes)),
                                       out.ctypes.data_as(C.c_void_p), repeat, C.byref(ms))
        if rc != 0:
            raise RuntimeError("dsb_exist_batch: " + self.cuda.dsb_last_error().decode())
        return out, ms.value

    def exist_static(self, codes):
        """dsb_exist_batch_static: the same probes through the entry point that allocates nothing."""
        np = self.np
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        k = self.ek_len()
        out = np.full(max(0, len(codes) - k + 1), 0xEE, dtype=np.uint8)
        rc = self.cuda.dsb_exist_batch_static(self.handle, codes.ctypes.data_as(C.c_void_p), C.c_uint64(len(codes)),
                                              out.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise RuntimeError("dsb_exist_batch_static: " + self.cuda.dsb_last_error().decode())
        return out

    def ek_len(self):
        self.host.dsb_ekmer_len.argtypes = [C.c_uint64]
        return self.host.dsb_ekmer_len(C.c_uint64(self.ek_bytes()))

    def ek_bytes(self):
        # dsb_host_index.e_kmer_size: see include/dsb_cuda.h (field order is part of the C-ABI)
        off = 8 + 8 + 6 * 8 + 8 + 8 + 8 + 8 + 8 + 8
        return C.c_uint64.from_buffer(self.buf, off).value


def read_index_files(index_dir):
    """Raw views of the reference's on-disk index (formats: reference src/idx.c:1103-1160, src/bwt.c:68-104)."""
    import numpy as np
    d = {}
    raw = np.fromfile(os.path.join(index_dir, "deSAMBA.bwt"), dtype=np.uint8)
    n = int(raw[:8].view(np.uint64)[0])
    d["bwt_occ"] = raw[8:8 + n]
    d["rank"] = raw[8 + n:8 + n + 40].view(np.uint64).copy()
    d["lut"] = raw[8 + n + 40:].view(np.uint64)
    d["ek0"] = np.fromfile(os.path.join(index_dir, "deSAMBA.exk0"), dtype=np.uint8)
    d["ek1"] = np.fromfile(os.path.join(index_dir, "deSAMBA.exk1"), dtype=np.uint8)
    d["ref_bin"] = np.fromfile(os.path.join(index_dir, "deSAMBA.ref_b"), dtype=np.uint8)[8:]
    unv = np.fromfile(os.path.join(index_dir, "deSAMBA.unv"), dtype=np.uint64)
    d["n_unitig_file"] = int(unv[0])
    d["dollar_pos"] = int(unv[0]) - 2              # DOLLOR_POS = unitig_v.n - 1 - 1 (src/idx.c:1128)
    bits = {0x8000000: (30, 16), 0x10000000: (31, 17), 0x20000000: (32, 17), 0x40000000: (33, 18),
            0x80000000: (34, 18), 0x100000000: (35, 19), 0x200000000: (36, 19), 0x400000000: (37, 20)}
    b, k = bits[len(d["ek0"])]
    d["ek_mask"] = (1 << b) - 1
    d["k"] = k
    return d


class HostIndex(C.Structure):
    """dsb_host_index of include/dsb_cuda.h, field for field."""
    _fields_ = [("bwt_occ", C.c_void_p), ("bwt_bytes", C.c_uint64), ("rank", C.c_uint64 * 6), ("dollar_pos", C.c_uint64),
                ("hash_index", C.c_void_p), ("sa_taxon", C.c_void_p), ("sa_n", C.c_uint64), ("e_kmer0", C.c_void_p),
                ("e_kmer1", C.c_void_p), ("e_kmer_size", C.c_uint64), ("unitig", C.c_void_p), ("n_unitig", C.c_uint64),
                ("ref_pos", C.c_void_p), ("n_ref_pos", C.c_uint64), ("ref_bin", C.c_void_p), ("ref_bin_n", C.c_uint64),
                ("ref_seq_offset", C.c_void_p), ("ref_seq_len", C.c_void_p), ("n_ref", C.c_uint64), ("q_mem", C.c_void_p),
                ("q_lv", C.c_void_p), ("filter_min_length", C.c_int32), ("filter_min_score", C.c_int32),
                ("filter_min_score_lv3", C.c_int32)]


class _SyntheticIndex(DeviceIndex):
    def __init__(self, handle, cuda, keep):
        import numpy as np
        self.np = np
        self.cuda = cuda
        self.handle = handle
        self.keep = keep

    def close(self):
        if self.handle:
            self.cuda.dsb_index_free(self.handle)
            self.handle = C.c_void_p()

    def ek_bytes(self):
        return self._ek_bytes


def synthetic_fm_index(n_sym, dev=0, seed=3, ek_table=None, _libs=(None, None)):
    """A device index whose FM part is a random BWT of n_sym symbols in the reference's 168-byte block layout
    (counts are consistent inside a block, which is all a rank query reads).  Only dsb_fm_rank_batch is
    meaningful on it: it exists to measure rank-query throughput on a table far larger than L2."""
    import numpy as np
    cuda = load_cuda_layer(_libs[1])
    rng = np.random.default_rng(seed)
    nb = n_sym // 256
    blocks = np.zeros((nb, 168), dtype=np.uint8)
    blocks[:, 40:] = rng.integers(0, 256, size=(nb, 128), dtype=np.uint8) & 0x33   # nibbles in 0..3
    keep = dict(blocks=blocks, lut=np.zeros((1 << 26) + 1, dtype=np.uint64), sa=np.zeros(8, dtype=np.uint64),
                ek=np.zeros(0x8000000, dtype=np.uint8), small=np.zeros(64, dtype=np.uint64),
                q=np.zeros(2400, dtype=np.int32))
    h = HostIndex()
    h.bwt_occ = blocks.ctypes.data
    h.bwt_bytes = nb * 168
    h.dollar_pos = n_sym + 5
    h.hash_index = keep["lut"].ctypes.data
    h.sa_taxon = keep["sa"].ctypes.data
    h.sa_n = 8
    if ek_table is not None:                 # exist-k-mer tables of another size class (both tables = this array)
        keep["ek"] = ek_table
    h.e_kmer0 = h.e_kmer1 = keep["ek"].ctypes.data
    h.e_kmer_size = len(keep["ek"])
    h.unitig = h.ref_pos = h.ref_seq_offset = h.ref_seq_len = keep["small"].ctypes.data
    h.n_unitig = h.n_ref_pos = h.n_ref = 8
    h.ref_bin = keep["small"].ctypes.data
    h.ref_bin_n = 64
    h.q_mem = keep["q"].ctypes.data
    h.q_lv = keep["q"].ctypes.data + 8000
    handle = C.c_void_p()
    rc = cuda.dsb_index_upload(C.byref(h), dev, C.byref(handle))
    if rc != 0:
        raise RuntimeError("dsb_index_upload: " + cuda.dsb_last_error().decode())
    ix = _SyntheticIndex(handle, cuda, keep)
    ix.host = _load(_libs[0] or lib_path("libdesamba.so"))
    ix._ek_bytes = len(keep["ek"])
    return ix
