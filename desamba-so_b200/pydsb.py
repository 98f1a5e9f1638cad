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
LIB_DIR = os.path.join(HERE, "lib")

API_SYMBOLS = ["load_index", "read_classify", "meta_analysis"]
CUDA_SYMBOLS = ["dsb_gpu_count", "dsb_index_upload", "dsb_index_free", "dsb_index_device_bytes", "dsb_slot_create",
                "dsb_slot_free", "dsb_slot_staging", "dsb_slot_submit", "dsb_slot_wait", "dsb_slot_timing",
                "dsb_slot_rerun_device", "dsb_last_error", "dsb_kernel_launch_count"]
HOST_SYMBOLS = ["dsb_index_load", "dsb_index_release_bulk", "dsb_index_destroy", "dsb_engine_create",
                "dsb_engine_destroy", "dsb_session_create", "dsb_session_destroy", "dsb_session_run",
                "dsb_fq_open_path", "dsb_fq_open_mem", "dsb_fq_close", "dsb_history_new", "dsb_history_next",
                "dsb_history_free", "dsb_mapq_tables", "dsb_format_read"]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "exist_probes0", "exist_probes1", "occ_queries", "lut_lookups", "sa_lookups", "unitig_reads", "refpos_reads",
        "ref_window_bytes", "mem_searches", "map_seeds", "sdp_windows", "anchors", "slow_reads")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


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


def lib_path(name, hostsim=False):
    if hostsim:
        return os.path.join(os.path.dirname(HERE), "tests", "hostsim", name.replace(".so", "_hostsim.so"))
    return os.path.join(LIB_DIR, name)


def load_cuda_layer(hostsim=False):
    p = lib_path("libdsb_cuda.so", hostsim) if not hostsim else os.path.join(os.path.dirname(HERE), "tests", "hostsim",
                                                                          "libdsb_hostsim.so")
    if not os.path.exists(p):
        raise RuntimeError("%s is not built (run __graft_entry__.build())" % p)
    lib = C.CDLL(p, mode=C.RTLD_GLOBAL)
    lib.dsb_last_error.restype = C.c_char_p
    lib.dsb_index_device_bytes.restype = C.c_uint64
    return lib


class Desamba:
    """The three-symbol API, exactly as a C caller would use it (reference main_test.c:29-57)."""

    def __init__(self, hostsim=False):
        load_cuda_layer(hostsim)
        p = lib_path("libdesamba.so", hostsim)
        if not os.path.exists(p):
            raise RuntimeError("%s is not built (run __graft_entry__.build())" % p)
        self.lib = C.CDLL(p)
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
