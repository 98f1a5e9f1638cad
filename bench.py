#!/usr/bin/env python3
"""bench.py -- Mbp/s classified on the BASELINE.json headline workload (configs[2]: RefSeq-scale unitig index, simulated
10 kb ONT-like reads, read-sharded over 1/2/4/8 B200 with the index replicated).  Default = `cfg3`: a 1 Gbp reference
(200 families x 5 strains x 1 Mbp: BWT 0.95 G rows, FM blocks 0.95 GB, prefix table 512 MiB, exist tables 2 x 256 MiB /
k = 17 -- everything far larger than the 126 MB L2), built on the box by this repository's own `deSAMBA index`, and
100 000 reads (~1 Gbase) PER GPU.  DSB_WORKLOAD=cfg3x is the same at 4 Gbp (configs[2]'s own size), cfg2 is configs[1].

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the classify hot path over the rank's batch of reads.
  value        whole-job Mbp/s with the batch already resident in HBM (the path's kernels + D2H of the results; CUDA events
               on the library's stream, max over ranks); `h2d_kernels_d2h` is the same with the H2D of the staged batch
               inside the events (SURVEY 8d's definition)
  e2e          same metric through read_classify() of libdesamba.so: FASTQ text in host memory in, SAM text
               out (parse, pinned H2D, kernels, D2H, SAM formatting all inside the timed region)
  roofline     dominant kernel (k_classify): algorithmic bytes (counted by the kernels) / its
               CUDA-event time vs the measured HBM peak; `fm_search` = bulk FM rank probe on a BWT larger
               than L2 vs the same peak (the random-access roofline the north star names)
  cpu_baseline the UNMODIFIED reference (oracle/_ref/deSAMBA classify -t <cores>) on a bounded sample of the same reads
               (the first REF_SAMPLE_READS reads), same index, same box
  e2e_inproc   (N > 1, rank 0, outside the timed region) ONE process driving all N GPUs through read_classify() on N x
               the per-GPU reads -- the north star's own multi-GPU design -- and `inproc_parity`: md5 of a read set
               classified with 1 and with N devices in one process
`--impl reference` times that reference binary alone (rank 0 only) on the same bounded sample.
The oracle is only the checker / the CPU baseline here; nothing under oracle/ is on the measured GPU path.
"""
import argparse
import ctypes as C
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "desamba-so_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))

WORKLOAD = os.environ.get("DSB_WORKLOAD", "cfg3")
WORKLOAD_TEXT = {
    "cfg2": "cfg2 = BASELINE configs[1]: 50 Mbp 10x5x1Mbp index, 10000 x 5 kb reads @10% error per GPU",
    "cfg3": "cfg3 = BASELINE configs[2] at 1 Gbp: 200x5x1Mbp reference, BWT 0.95 G rows, exist tables 2x256 MiB (k=17), index "
            "built by this repo's `deSAMBA index`; 100000 ONT-like 10 kb reads (~1 Gbase, 10% error 30/30/40 sub/ins/del, "
            "+-50% length) per GPU",
    "cfg3x": "cfg3x = BASELINE configs[2] at 4 Gbp: 800x5x1Mbp reference, BWT 3.8 G rows (> 2^32 with the reference's "
             "block bytes), exist tables 2x1 GiB (k=18), index built by this repo's `deSAMBA index`; 100000 ONT-like 10 kb "
             "reads (~1 Gbase) per GPU",
}.get(WORKLOAD, WORKLOAD) + " (tools/dataset.py)"
REF_SAMPLE_READS = int(os.environ.get("DSB_REF_SAMPLE_READS", "20000"))     # bounded sample of the reference arm
METRIC = "Mbp/s classified"
UNIT = "Mbp/s"


def env_int(name, dflt):
    try:
        return int(os.environ.get(name, dflt))
    except ValueError:
        return dflt


RANK = env_int("RANK", 0)
LOCAL_RANK = env_int("LOCAL_RANK", 0)
WORLD = env_int("WORLD_SIZE", 1)


# Only the JSON line may reach stdout: native libraries (NCCL prints its version banner there) get stderr.
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


def log(*a):
    print("[bench r%d]" % RANK, *a, file=sys.stderr, flush=True)


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of this rank's GPU, sampled through NVML every 10 ms during the timed regions."""

    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu = gpu
        self.samples = []
        self.reasons = set()
        self.sm_max = None
        self.stop_flag = False
        self.active = False          # samples are kept only while a timed region is open

    def _nvml_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        ids = [x for x in vis.split(",") if x.strip()]
        if ids and all(x.strip().isdigit() for x in ids) and self.gpu < len(ids):
            return int(ids[self.gpu])
        return self.gpu

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self._nvml_index())
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        except Exception as e:
            log("clock sampler: NVML unavailable (%s)" % e)
            return
        while not self.stop_flag:
            if self.active:
                try:
                    self.samples.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                    r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for n, bit in self.REASONS:
                        if r & bit:
                            self.reasons.add(n)
                except Exception:
                    pass
            time.sleep(0.01)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons),
                "samples": len(s), "how": "NVML, 10 ms period, during the timed kernel loop and the timed e2e calls"}


def dist_setup():
    """torch.distributed is plumbing only: a barrier and a max over ranks (no data-path collective)."""
    if WORLD <= 1:
        return None
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(LOCAL_RANK)
    dist.init_process_group("nccl", device_id=torch.device("cuda", LOCAL_RANK))
    global CPU_GROUP
    CPU_GROUP = dist.new_group(backend="gloo")      # for waits that must not occupy a GPU (see cpu_barrier)
    return dist


CPU_GROUP = None


def cpu_barrier(dist):
    """A barrier on the host only.  An NCCL barrier is a kernel that spins on every waiting GPU; while rank 0's single-process
    leg drives ALL the GPUs from a child process, those spinning kernels would share each GPU with it by time slicing
    (measured: 590 instead of 334 ms per step at N = 2)."""
    if dist is not None:
        dist.barrier(group=CPU_GROUP)


def barrier(dist):
    if dist is not None:
        import torch
        dist.barrier()
        torch.cuda.synchronize()


def max_over_ranks(dist, v):
    if dist is None:
        return v
    import torch
    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(dist, v):
    if dist is None:
        return v
    import torch
    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


# ------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the unmodified reference binary on the box's host cores
# ------------------------------------------------------------------------------------------------
def reference_sample(meta):
    """The bounded sample the reference is timed on: the first REF_SAMPLE_READS reads of the workload's read set
    -> (path, n_reads, n_bases)."""
    path = os.path.join(meta["dir"], "ref_sample_%d.fq" % REF_SAMPLE_READS)
    n = min(REF_SAMPLE_READS, meta["n_reads"])
    if n == meta["n_reads"]:
        return meta["reads"], meta["n_reads"], meta["n_bases"]
    info = path + ".json"
    if not (os.path.exists(path) and os.path.exists(info)):
        bases = 0
        with open(meta["reads"], "rb") as fi, open(path + ".tmp", "wb") as fo:
            for i in range(4 * n):
                line = fi.readline()
                if i % 4 == 1:
                    bases += len(line) - 1
                fo.write(line)
        os.replace(path + ".tmp", path)
        json.dump({"n_reads": n, "n_bases": bases}, open(info, "w"))
    d = json.load(open(info))
    return path, d["n_reads"], d["n_bases"]


def reference_run(meta, threads):
    """One `deSAMBA classify -t threads` pass over the bounded sample; returns (seconds of classify excluding index
    load -- the reference's own 'sequences processed in' line --, bases, reads)."""
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "deSAMBA")
    if not os.path.exists(ref_bin):
        raise RuntimeError("oracle/_ref/deSAMBA missing (built by __graft_entry__.build() where /root/reference exists)")
    fq, n_reads, n_bases = reference_sample(meta)
    r = subprocess.run([ref_bin, "classify", "-t", str(threads), "-f", "SAM", meta["idx"] + "/", fq],
                       stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    err = r.stderr.decode(errors="replace")
    m = re.search(r"sequences processed in ([0-9.]+)s", err)
    if r.returncode != 0 or not m:
        raise RuntimeError("reference classify failed: " + err[-500:])
    return float(m.group(1)), n_bases, n_reads


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def run_reference_arm(args):
    import dataset
    if RANK != 0:
        return
    meta = dataset.build(WORKLOAD, log=sys.stderr, golden=False)
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        reference_run(meta, cores)
    t = 0.0
    n_bases = n_reads = 0
    for _ in range(args.steps):
        s, n_bases, n_reads = reference_run(meta, cores)
        t += s
    v = n_bases * args.steps / t / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/int64", "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT,
                   "sample_reads_per_step": n_reads, "sample_bases_per_step": n_bases,
                   "timing": "reference's own 'sequences processed in' line, index load excluded"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "the first %d reads of the workload (%d bases) per step, deSAMBA classify -t %d, %s" %
                                   (n_reads, n_bases, cores, cpu_model())},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def parse_fastq(path):
    """-> (names, concatenated sequence bytes, offsets, lengths); plain 4-line FASTQ written by tools/synth.py"""
    import numpy as np
    lines = open(path, "rb").read().split(b"\n")
    seqs = lines[1::4]
    if seqs and seqs[-1] == b"" and len(lines) % 4 == 1:
        pass
    seqs = [s for s in seqs[:len(lines) // 4]]
    lens = np.array([len(s) for s in seqs], dtype=np.uint32)
    offs = np.zeros(len(seqs), dtype=np.uint64)
    if len(seqs) > 1:
        offs[1:] = np.cumsum(lens[:-1], dtype=np.uint64)
    return b"".join(seqs), offs, lens


class DeviceBatch:
    """One slot of the CUDA C-ABI holding the rank's whole batch (inputs staged once, then device-resident)."""

    def __init__(self, index_dir, fq_path, dev):
        import numpy as np
        import pydsb
        self.pydsb = pydsb
        self.ix = pydsb.DeviceIndex(index_dir, dev=dev)
        self.cuda = self.ix.cuda
        host = self.ix.host
        seq, offs, lens = parse_fastq(fq_path)
        self.n_reads = len(lens)
        self.n_bases = int(lens.sum())
        self.slot = C.c_void_p()
        rc = self.cuda.dsb_slot_create(self.ix.handle, C.c_uint64(self.n_bases + 64), C.c_uint32(self.n_reads), C.byref(self.slot))
        if rc != 0:
            raise RuntimeError("dsb_slot_create: " + self.cuda.dsb_last_error().decode())
        p_seq, p_desc, p_edge = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self.cuda.dsb_slot_staging(self.slot, C.byref(p_seq), C.byref(p_desc), C.byref(p_edge))
        C.memmove(p_seq, seq, len(seq))
        # history model of the reference's worker buffer, in read order (host C code of the product)
        host.dsb_history_new.restype = C.c_void_p
        host.dsb_history_next.restype = C.c_uint32
        host.dsb_history_next.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        host.dsb_history_free.argtypes = [C.c_void_p]
        # the 2G / 3G filter class of a read shorter than 510 bases depends on earlier results (dsb_history.c); the bench
        # workloads have none, so every read is in the long-read class and one launch can hold the whole set
        if int(lens.min()) < 510:
            raise RuntimeError("bench workload holds reads shorter than 510 bases")
        h = host.dsb_history_new()
        desc = (pydsb.ReadDesc * self.n_reads).from_address(p_desc.value)
        edge_stride = 128
        for i in range(self.n_reads):
            o, L = int(offs[i]), int(lens[i])
            desc[i].seq_off = o
            desc[i].len = L
            desc[i].flags = host.dsb_history_next(h, p_seq.value + o, L, p_edge.value + i * edge_stride)
        host.dsb_history_free(h)
        self.h2d_bytes = self.n_bases + self.n_reads * (16 + edge_stride + 4)
        self.counters = None
        self.n_hits = 0

    def _wait(self):
        res, hits, n_hits = C.c_void_p(), C.c_void_p(), C.c_uint32(0)
        cnt = self.pydsb.Counters()
        rc = self.cuda.dsb_slot_wait(self.slot, C.byref(res), C.byref(hits), C.byref(n_hits), C.byref(cnt))
        if rc != 0:
            raise RuntimeError("dsb_slot_wait: " + self.cuda.dsb_last_error().decode())
        self.counters = cnt.as_dict()
        self.n_hits = n_hits.value
        a, b, c = C.c_float(0), C.c_float(0), C.c_float(0)
        self.cuda.dsb_slot_timing(self.slot, C.byref(a), C.byref(b), C.byref(c))
        i, sd, sc = C.c_float(0), C.c_float(0), C.c_float(0)
        self.split = bool(self.cuda.dsb_slot_timing_stages(self.slot, C.byref(i), C.byref(sd), C.byref(sc)))
        self.stage_ms = {"k_islands": i.value, "k_seed": sd.value, "k_classify": sc.value}
        return a.value, b.value, c.value

    def submit(self):
        rc = self.cuda.dsb_slot_submit(self.slot, C.c_uint32(self.n_reads), C.c_uint64(self.n_bases))
        if rc != 0:
            raise RuntimeError("dsb_slot_submit: " + self.cuda.dsb_last_error().decode())
        return self._wait()

    def rerun(self):
        rc = self.cuda.dsb_slot_rerun_device(self.slot)
        if rc != 0:
            raise RuntimeError("dsb_slot_rerun_device: " + self.cuda.dsb_last_error().decode())
        return self._wait()

    def d2h_bytes(self):
        return self.n_reads * 16 + self.n_hits * 36 + 128

    def close(self):
        self.cuda.dsb_slot_free(self.slot)
        self.ix.close()


def algorithmic_bytes(c, n_bases):
    """DESIGN.md 'algorithmic bytes': what one pass of the classify stage must fetch at minimum."""
    return (c["occ_queries"] * 64 + c["lut_lookups"] * 16 + c["sa_lookups"] * 8 + c["unitig_reads"] * 8 +
            c["refpos_reads"] * 8 + c["ref_window_bytes"] + 4 * n_bases)


def random_access_peak(dev):
    """The random-access ceiling of this GPU, measured in this run: plain gathers of 32 / 64 / 128-byte pieces from an
    8 GiB table (k_random_gather, no arithmetic) -> GB/s per piece size."""
    import pydsb
    cuda = pydsb.load_cuda_layer()
    ra = {}
    for chunk in (32, 64, 128):
        t = C.c_float(0)
        rc = cuda.dsb_random_access_probe(dev, C.c_uint64(8 << 30), C.c_uint32(chunk), C.c_uint64(1 << 26), 3, C.byref(t))
        if rc != 0:
            raise RuntimeError("dsb_random_access_probe: " + cuda.dsb_last_error().decode())
        ra["%dB" % chunk] = (1 << 26) * chunk / (t.value * 1e-3) / 1e9
    return ra


def fm_search_product(counters, k_seed_ms, ra, index_bytes):
    """The north star's FM-search figure for the PRODUCT kernel: rank queries issued by k_seed (one 64-byte FM block
    each; the kernel counts them) / k_seed's CUDA-event time, against the random-64 B ceiling measured in this run.
    k_seed also does the LUT, SA, unitig and reference-window reads and the Landau-Vishkin extensions in that time, so
    this is a lower bound on its rank-query rate."""
    q = counters.get("seed_occ_queries") or counters["occ_queries"]
    gbs = q * 64 / (k_seed_ms * 1e-3) / 1e9
    return {"kernel": "k_seed", "occ_queries": q, "counted_by": "seed_occ_queries" if counters.get("seed_occ_queries") else
            "occ_queries (k_seed + the slow-seeding path of k_classify)", "bytes_per_query": 64, "ms": k_seed_ms,
            "achieved": gbs, "unit": "GB/s", "random_64B_peak": ra["64B"], "frac_of_random_64B": gbs / ra["64B"],
            "alg_bytes_all": counters.get("seed_alg_bytes"),
            "achieved_all_bytes": (counters.get("seed_alg_bytes") or 0) / (k_seed_ms * 1e-3) / 1e9,
            "index_device_bytes": index_bytes}


def fm_probe(ix_dir_unused, dev, hbm_peak):
    """Bulk FM rank queries on a synthetic BWT much larger than L2: the random-access roofline probe."""
    import numpy as np
    import pydsb
    n_sym = 1 << 31                                       # 2 Gi symbols -> 2 GiB of 64-byte FM blocks on the device
    ix = pydsb.synthetic_fm_index(n_sym, dev)
    rng = np.random.default_rng(1)
    n = 1 << 24
    rows = rng.integers(0, n_sym - 64, size=n, dtype=np.uint64)
    c = rng.integers(0, 4, size=n, dtype=np.uint8)
    _o, _c, ms = ix.fm_rank(rows, c, repeat=5)
    ix.close()
    gbs = n * 64 / (ms * 1e-3) / 1e9
    ra = random_access_peak(dev)
    return {"kernel": "k_fm_rank (synthetic micro-kernel: validates the 64-byte FM block, NOT the product's FM search)", "queries": n, "ms": ms, "bytes_per_query": 64, "achieved": gbs, "peak": hbm_peak,
            "unit": "GB/s", "frac": gbs / hbm_peak, "table_bytes": (n_sym // 64) * 64,
            "random_access_peak": ra, "frac_of_random_64B": gbs / ra["64B"],
            "note": "one 64 B block (2 sectors) per rank query, random rows, table 16x larger than L2; "
                    "random_access_peak = k_random_gather (pure gathers, 8 GiB table) measured in the same run"}


def ncu_traffic(n_reads, n_bases):
    """dram bytes read + written by one k_classify launch, from the committed `ncu --set full` capture of this
    same workload (profiles/ncu_k_classify_traffic.json; a profiler number cannot be taken inside a timed run)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_k_classify_traffic.json")))
        if t.get("n_reads") == n_reads and t.get("n_bases") == n_bases:
            return float(t["dram_bytes_read"]) + float(t["dram_bytes_write"]), t.get("source")
    except Exception:
        pass
    return None, None


def e2e_leg(meta, dev, steps, warmup, sampler=None):
    """read_classify() of libdesamba.so on FASTQ text held in host memory; wall clock around each call."""
    import pydsb
    os.environ["DSB_GPUS"] = "1"
    os.environ["DSB_GPU_FIRST"] = str(dev)
    d = pydsb.Desamba()
    d.load_index(meta["idx"])
    text = open(meta["reads"], "rb").read()
    n_sam = 0
    for _ in range(max(1, warmup)):
        _t, n_sam = d.read_classify_timed(text, thread_id=0, thread_num=1)
    dt = 0.0
    if sampler is not None:
        sampler.active = True
    for _ in range(steps):
        t, n = d.read_classify_timed(text, thread_id=0, thread_num=1)   # the C call + free(), as a C caller would
        if n != n_sam:
            raise RuntimeError("read_classify returned %d bytes, expected %d" % (n, n_sam))
        dt += t
    if sampler is not None:
        sampler.active = False
    return dt / steps, n_sam


def run_ours(args):
    import dataset
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6500.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6500 GB/s (B200_PROFILING.md)"
    dist = dist_setup()
    # rank 0 builds the shared index + its reads first, the others then add their own read sets
    name = WORKLOAD if RANK == 0 else "%s@%d" % (WORKLOAD, RANK)
    if RANK == 0:
        meta = dataset.build(name, golden=False)
    barrier(dist)
    if RANK != 0:
        meta = dataset.build(name, golden=False)
    t_load = time.perf_counter()
    batch = DeviceBatch(meta["idx"], meta["reads"], LOCAL_RANK)
    batch.cuda.dsb_index_device_bytes.restype = C.c_uint64
    index_bytes = int(batch.cuda.dsb_index_device_bytes(batch.ix.handle))
    log("index (%.2f GB on the device) + batch resident in %.1fs, %d reads %d bases" %
        (index_bytes / 1e9, time.perf_counter() - t_load, batch.n_reads, batch.n_bases))
    first = batch.submit()                      # H2D + first pass (also sizes the scratch arenas)
    log("first pass: total %.1f ms (exist %.2f, classify %.1f)" % first)
    for _ in range(args.warmup):
        batch.rerun()
    sampler = ClockSampler(LOCAL_RANK)
    sampler.start()
    barrier(dist)
    sampler.active = True
    launches0 = batch.cuda.dsb_kernel_launch_count()
    ms_total = ms_exist = ms_classify = 0.0
    stage = {"k_islands": 0.0, "k_seed": 0.0, "k_classify": 0.0}
    t0 = time.perf_counter()
    for _ in range(args.steps):
        a, b, c = batch.rerun()
        ms_total += a
        ms_exist += b
        ms_classify += c
        for k in stage:
            stage[k] += batch.stage_ms[k]
    wall = time.perf_counter() - t0
    sampler.active = False
    barrier(dist)
    launches = batch.cuda.dsb_kernel_launch_count() - launches0
    t_dev = max_over_ranks(dist, ms_total * 1e-3)
    bases_all = sum_over_ranks(dist, float(batch.n_bases))
    value = bases_all * args.steps / t_dev / 1e6
    counters = batch.counters
    d2h = batch.d2h_bytes()
    # SURVEY 8(d)'s own definition: first H2D of the staged batch -> last D2H of the results, per step
    n_sub = max(3, min(args.steps, 5))
    ms_sub = 0.0
    sampler.active = True
    for _ in range(n_sub):
        ms_sub += batch.submit()[0]
    sampler.active = False
    t_sub = max_over_ranks(dist, ms_sub * 1e-3 / n_sub)
    # end to end through the public API (every rank runs it; max over ranks)
    n_reads, n_bases = batch.n_reads, batch.n_bases
    h2d_bytes, split = batch.h2d_bytes, batch.split
    batch.close()                               # the e2e leg brings its own index replica and slots
    e2e_s, n_sam = e2e_leg(meta, LOCAL_RANK, args.steps, args.warmup, sampler)
    barrier(dist)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    e2e_t = max_over_ranks(dist, e2e_s)
    e2e_value = bases_all / e2e_t / 1e6
    if RANK != 0:
        import torch
        torch.cuda.synchronize()
        cpu_barrier(dist)                       # rank 0 runs the single-process multi-GPU leg meanwhile: wait off the GPU
        return
    alg = algorithmic_bytes(counters, n_bases)
    k_ms = ms_classify / args.steps
    stage = {k: v / args.steps for k, v in stage.items()}
    achieved = alg / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic(n_reads, n_bases)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": WORLD, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/int64", "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT,
                   "n_reads_per_gpu": n_reads, "n_bases_per_gpu": n_bases, "index_device_bytes": index_bytes,
                   "parallelism": "reads sharded, index replicated, no collective",
                   "value_scope": "inputs resident in HBM: k_encode, k_exist, k_islands, k_seed, k_classify + D2H of the results",
                   "l2": "inputs (%d MB ASCII + %d MB derived per step) and the index (%.1f GB) exceed the 126 MB L2; no flush needed" %
                         (n_bases // 1000000, 4 * n_bases // 1000000, index_bytes / 1e9),
                   "wall_ms_per_step": 1e3 * wall / args.steps},
        "h2d_kernels_d2h": {"value": bases_all / t_sub / 1e6, "unit": UNIT, "ms_per_step": 1e3 * t_sub, "steps": n_sub,
                            "scope": "SURVEY 8(d): pinned staging -> H2D -> kernels -> D2H of the results, CUDA events on the slot's stream"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h,
                "api": "read_classify() of libdesamba.so: FASTQ text -> SAM_FULL text (%d bytes)" % n_sam,
                "ms_per_step": 1e3 * e2e_t},
        "gpu_launches": launches,
        "kernel_ms": {"exist": ms_exist / args.steps, "classify": k_ms, "total": ms_total / args.steps,
                      "classify_stages": stage if split else None},
        "roofline": {"bound": "hbm", "kernel": "classify stage = k_islands + k_seed + k_classify (the path's work counters span the three launches)" if split else "k_classify", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes": alg, "peak_source": peak_src},
        "counters": counters,
        "clocks": sampler.result(),
    }
    # k_exist: one random bit per probe, and B200 moves 128 B of DRAM per random access: probes/s against
    # hbm peak / 128 B, the rate at which the DRAM can serve distinct lines at all
    try:
        probes = counters["exist_probes0"] + counters["exist_probes1"]
        ex_ms = ms_exist / args.steps
        if ex_ms > 0:
            line["exist_probe"] = {"kernel": "k_exist", "probes": probes, "ms": ex_ms, "achieved": probes / (ex_ms * 1e-3) / 1e9,
                                   "unit": "G probes/s", "peak": hbm_peak / 128.0, "frac": probes / (ex_ms * 1e-3) / 1e9 / (hbm_peak / 128.0),
                                   "peak_is": "MEASURED_PEAKS hbm GB/s / 128 B per random DRAM access"}
    except Exception as e:
        line["exist_probe"] = {"error": str(e)}
    if not args.no_fm_probe:
        try:
            ra = random_access_peak(LOCAL_RANK)
            line["roofline"]["random_access_peak"] = ra
            line["roofline"]["frac_of_random_64B"] = achieved / ra["64B"]
            if split:
                line["fm_search_product"] = fm_search_product(counters, stage["k_seed"], ra, index_bytes)
            line["fm_search"] = fm_probe(None, LOCAL_RANK, hbm_peak)
        except Exception as e:                                   # the probes are auxiliary: report, do not hide
            line["fm_search"] = {"error": str(e)}
    if not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        try:
            s, b, nr = reference_run(meta, cores)
            line["cpu_baseline"] = {"value": b / s / 1e6, "unit": UNIT, "cores": cores, "kind": "reference",
                                    "sample": "the first %d reads of the workload (%d bases), oracle/_ref/deSAMBA classify -t %d, %.2fs, %s" %
                                              (nr, b, cores, s, cpu_model())}
        except Exception as e:
            line["cpu_baseline"] = {"error": str(e)}
    if WORLD > 1 and not args.no_inproc:
        try:
            line.update(inproc_leg(meta, WORLD, e2e_value / WORLD))
        except Exception as e:
            line["e2e_inproc"] = {"error": str(e)}
    cpu_barrier(dist)
    emit(line)


INPROC_CHILD = r'''
import hashlib, json, os, sys, time
sys.path.insert(0, os.path.join(sys.argv[1], "desamba-so_b200"))
import pydsb
idx, reads, sample, n_gpu, steps = sys.argv[2], sys.argv[3], sys.argv[4], int(sys.argv[5]), int(sys.argv[6])
out = {}
text = open(sample, "rb").read()
md5 = {}
for g in (1, n_gpu):
    os.environ["DSB_GPUS"] = str(g)
    os.environ["DSB_GPU_FIRST"] = "0"
    d = pydsb.Desamba()
    t0 = time.perf_counter()
    d.load_index(idx)
    out["load_index_s_%d_gpu" % g] = time.perf_counter() - t0
    md5[g] = hashlib.md5(d.read_classify(text, thread_id=0, thread_num=1)).hexdigest()
    if g == n_gpu:
        one = open(reads, "rb").read()
        n_bases = sum(len(l) for l in one.split(b"\n")[1::4])
        big = one * n_gpu
        del one
        d.read_classify_timed(big, thread_id=1, thread_num=1)
        dt = 0.0
        for _ in range(steps):
            t, n = d.read_classify_timed(big, thread_id=1, thread_num=1)
            dt += t
        out["ms_per_step"] = 1e3 * dt / steps
        out["value"] = n_bases * n_gpu / (dt / steps) / 1e6
        out["bases_per_step"] = n_bases * n_gpu
        out["sam_bytes"] = n
out["md5"] = {str(k): v for k, v in md5.items()}
print("INPROC " + json.dumps(out))
'''


def inproc_leg(meta, n_gpu, e2e_per_gpu):
    """One process, DSB_GPUS=N (SURVEY 8e / the north star's design: one host process shards the reads of a call over
    the box's GPUs, index replicated by peer copy, results merged in read order on the host).  Run by rank 0 in a child
    process while the other ranks wait at a barrier; not part of the timed region of `value` / `e2e`."""
    sample, _nr, _nb = reference_sample(meta)
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    r = subprocess.run([sys.executable, "-c", INPROC_CHILD, ROOT, meta["idx"], meta["reads"], sample, str(n_gpu), "3"],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=900)
    m = re.search(r"^INPROC (.*)$", r.stdout.decode(errors="replace"), re.M)
    if r.returncode != 0 or not m:
        raise RuntimeError("in-process leg failed: " + r.stderr.decode(errors="replace")[-600:])
    o = json.loads(m.group(1))
    return {"e2e_inproc": {"value": o["value"], "unit": UNIT, "n_gpus": n_gpu, "ms_per_step": o["ms_per_step"],
                           "bases_per_step": o["bases_per_step"], "steps": 3,
                           "vs_one_gpu_e2e": o["value"] / e2e_per_gpu if e2e_per_gpu else None,
                           "load_index_s": {"1_gpu": o["load_index_s_1_gpu"], "%d_gpus" % n_gpu: o["load_index_s_%d_gpu" % n_gpu]},
                           "api": "ONE process, read_classify() with DSB_GPUS=%d on %d x the per-GPU reads" % (n_gpu, n_gpu)},
            "inproc_parity": o["md5"]["1"] == o["md5"][str(n_gpu)],
            "inproc_md5": o["md5"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-fm-probe", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-inproc", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    if WORLD != args.gpus and args.gpus > 1:
        log("warning: --gpus %d but WORLD_SIZE %d (launch with torch.distributed.run)" % (args.gpus, WORLD))
    run_ours(args)
    if WORLD > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
