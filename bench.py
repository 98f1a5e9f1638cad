#!/usr/bin/env python3
"""bench.py -- Mbp/s classified on the BASELINE.json workload (configs[1]: 50 Mbp synthetic multi-genome
index, 10 000 simulated 5 kb reads at 10 % error), 1..8 B200, index replicated, reads sharded.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the classify hot path over the rank's batch of reads.
  value        whole-job Mbp/s with the batch already resident in HBM (kernels only, CUDA events on the
               library's stream, max over ranks)
  e2e          same metric through read_classify() of libdesamba.so: FASTQ text in host memory in, SAM text
               out (parse, pinned H2D, kernels, D2H, SAM formatting all inside the timed region)
  roofline     dominant kernel (k_classify): algorithmic bytes (counted by the kernels) / its
               CUDA-event time vs the measured HBM peak; `fm_search` = bulk FM rank probe on a BWT larger
               than L2 vs the same peak (the random-access roofline the north star names)
  cpu_baseline the UNMODIFIED reference (oracle/_ref/deSAMBA classify -t <cores>) on the same reads, same box
`--impl reference` times that reference binary alone (rank 0 only).
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

# The bench workload is BASELINE.json configs[1] (cfg2).  DSB_WORKLOAD=cfg3s measures configs[2] at the scale the
# reference's index builder allows on the box (tools/dataset.py); that line is evidence for profiles/, not the headline.
WORKLOAD = os.environ.get("DSB_WORKLOAD", "cfg2")
WORKLOAD_TEXT = {
    "cfg2": "cfg2: 50 Mbp 10x5x1Mbp index, 10000 x 5 kb reads @10% error",
    "cfg3s": "cfg3s (configs[2] scaled): 200 Mbp 40x5x1Mbp index, 20000 x 10 kb ONT-like reads @10% error",
}.get(WORKLOAD, WORKLOAD)
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
    return dist


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
def reference_run(meta, threads):
    """One `deSAMBA classify -t threads` pass; returns (seconds of classify excluding index load, bases)."""
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "deSAMBA")
    if not os.path.exists(ref_bin):
        raise RuntimeError("oracle/_ref/deSAMBA missing (built by __graft_entry__.build() where /root/reference exists)")
    r = subprocess.run([ref_bin, "classify", "-t", str(threads), "-f", "SAM", meta["idx"] + "/", meta["reads"]],
                       stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    err = r.stderr.decode(errors="replace")
    m = re.search(r"sequences processed in ([0-9.]+)s", err)
    if r.returncode != 0 or not m:
        raise RuntimeError("reference classify failed: " + err[-500:])
    return float(m.group(1)), meta["n_bases"]


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
    meta = dataset.build(WORKLOAD, log=sys.stderr)
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        reference_run(meta, cores)
    t = 0.0
    for _ in range(args.steps):
        s, _b = reference_run(meta, cores)
        t += s
    v = meta["n_bases"] * args.steps / t / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/int64", "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT + " (tools/dataset.py)",
                   "n_reads": meta["n_reads"], "n_bases": meta["n_bases"], "timing": "reference's own 'sequences processed in' line, index load excluded"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "all %d reads, deSAMBA classify -t %d, %s" % (meta["n_reads"], cores, cpu_model())},
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
        host.dsb_history_note_max.argtypes = [C.c_void_p, C.c_uint32]
        host.dsb_history_free.argtypes = [C.c_void_p]
        h = host.dsb_history_new()
        desc = (pydsb.ReadDesc * self.n_reads).from_address(p_desc.value)
        edge_stride = 128
        for i in range(self.n_reads):
            o, L = int(offs[i]), int(lens[i])
            desc[i].seq_off = o
            desc[i].len = L
            desc[i].flags = host.dsb_history_next(h, p_seq.value + o, L, p_edge.value + i * edge_stride)
            if L >= 40:
                host.dsb_history_note_max(h, L)
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
    """DESIGN.md 'algorithmic bytes': what one pass of k_classify must fetch at minimum."""
    return (c["occ_queries"] * 64 + c["lut_lookups"] * 16 + c["sa_lookups"] * 8 + c["unitig_reads"] * 8 +
            c["refpos_reads"] * 8 + c["ref_window_bytes"] + 4 * n_bases)


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
    # the ceiling: plain random gathers of 32 / 64 / 128-byte pieces from an 8 GiB table (no arithmetic)
    cuda = pydsb.load_cuda_layer()
    ra = {}
    for chunk in (32, 64, 128):
        t = C.c_float(0)
        rc = cuda.dsb_random_access_probe(dev, C.c_uint64(8 << 30), C.c_uint32(chunk), C.c_uint64(1 << 26), 3, C.byref(t))
        if rc != 0:
            raise RuntimeError("dsb_random_access_probe: " + cuda.dsb_last_error().decode())
        ra["%dB" % chunk] = (1 << 26) * chunk / (t.value * 1e-3) / 1e9
    return {"kernel": "k_fm_rank", "queries": n, "ms": ms, "bytes_per_query": 64, "achieved": gbs, "peak": hbm_peak,
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
        n_sam = len(d.read_classify(text, thread_id=0, thread_num=1))
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
    log("index + batch resident in %.1fs, %d reads %d bases" % (time.perf_counter() - t_load, batch.n_reads, batch.n_bases))
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
    t0 = time.perf_counter()
    for _ in range(args.steps):
        a, b, c = batch.rerun()
        ms_total += a
        ms_exist += b
        ms_classify += c
    wall = time.perf_counter() - t0
    sampler.active = False
    barrier(dist)
    launches = batch.cuda.dsb_kernel_launch_count() - launches0
    t_dev = max_over_ranks(dist, ms_total * 1e-3)
    bases_all = sum_over_ranks(dist, float(batch.n_bases))
    value = bases_all * args.steps / t_dev / 1e6
    counters = batch.counters
    d2h = batch.d2h_bytes()
    # end to end through the public API (every rank runs it; max over ranks)
    e2e_s, n_sam = e2e_leg(meta, LOCAL_RANK, args.steps, args.warmup, sampler)
    barrier(dist)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    e2e_t = max_over_ranks(dist, e2e_s)
    e2e_value = bases_all / e2e_t / 1e6
    if RANK != 0:
        batch.close()
        return
    alg = algorithmic_bytes(counters, batch.n_bases)
    k_ms = ms_classify / args.steps
    achieved = alg / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic(batch.n_reads, batch.n_bases)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": WORLD, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/int64", "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT + " per GPU (tools/dataset.py)",
                   "n_reads_per_gpu": batch.n_reads, "n_bases_per_gpu": batch.n_bases, "parallelism": "reads sharded, index replicated, no collective",
                   "l2": "inputs (%d MB ASCII + %d MB derived per step) and the index (>= 0.9 GB) exceed the 126 MB L2; no flush needed" %
                         (batch.n_bases // 1000000, 4 * batch.n_bases // 1000000),
                   "wall_ms_per_step": 1e3 * wall / args.steps},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": batch.h2d_bytes, "d2h_bytes_per_step": d2h,
                "api": "read_classify() of libdesamba.so: FASTQ text -> SAM_FULL text (%d bytes)" % n_sam,
                "ms_per_step": 1e3 * e2e_t},
        "gpu_launches": launches,
        "kernel_ms": {"exist": ms_exist / args.steps, "classify": k_ms, "total": ms_total / args.steps,
                      "classify_stages_last_step": batch.stage_ms if batch.split else None},
        "roofline": {"bound": "hbm", "kernel": "classify stage = k_islands + k_seed + k_classify (the path's work counters span the three launches)" if batch.split else "k_classify", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes": alg, "peak_source": peak_src},
        "counters": counters,
        "clocks": sampler.result(),
    }
    batch.close()
    if not args.no_fm_probe:
        try:
            line["fm_search"] = fm_probe(None, LOCAL_RANK, hbm_peak)
        except Exception as e:                                   # the probe is auxiliary: report, do not hide
            line["fm_search"] = {"error": str(e)}
    if not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        try:
            gm = dataset.build(WORKLOAD, golden=True)
            s, b = reference_run(gm, cores)
            line["cpu_baseline"] = {"value": b / s / 1e6, "unit": UNIT, "cores": cores, "kind": "reference",
                                    "sample": "all %d reads of the workload, oracle/_ref/deSAMBA classify -t %d, %.2fs, %s" %
                                              (gm["n_reads"], cores, s, cpu_model())}
        except Exception as e:
            line["cpu_baseline"] = {"error": str(e)}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-fm-probe", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
