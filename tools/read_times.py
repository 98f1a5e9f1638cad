#!/usr/bin/env python3
"""Per-read device time of k_classify on the bench workload (needs the -DDSB_READ_CYCLES profiling build:
tools/variant.sh rc -DDSB_READ_CYCLES; DSB_LIB_DIR=desamba-so_b200/lib_rc python tools/read_times.py).

Prints the distribution of per-read times, the occupancy of the warp slots over the launch, and what list
scheduling would give for other read orders (the kernel hands reads to warps in `order`, first come first served)."""
import ctypes as C
import heapq
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "desamba-so_b200"))
import bench      # noqa: E402
import dataset    # noqa: E402


def makespan(durs, n_workers):
    h = [0.0] * n_workers
    heapq.heapify(h)
    end = 0.0
    for d in durs:
        t = heapq.heappop(h) + d
        end = max(end, t)
        heapq.heappush(h, t)
    return end


def main():
    meta = dataset.build(sys.argv[1] if len(sys.argv) > 1 else "cfg2", golden=False)
    b = bench.DeviceBatch(meta["idx"], meta["reads"], 0)
    b.submit()
    b.rerun()
    res, hits, n_hits = C.c_void_p(), C.c_void_p(), C.c_uint32(0)
    b.rerun_nowait = None
    rc = b.cuda.dsb_slot_rerun_device(b.slot)
    cnt = b.pydsb.Counters()
    rc = b.cuda.dsb_slot_wait(b.slot, C.byref(res), C.byref(hits), C.byref(n_hits), C.byref(cnt))
    assert rc == 0
    a, e, c = C.c_float(0), C.c_float(0), C.c_float(0)
    b.cuda.dsb_slot_timing(b.slot, C.byref(a), C.byref(e), C.byref(c))
    r = np.ctypeslib.as_array((C.c_uint32 * (4 * b.n_reads)).from_address(res.value)).reshape(-1, 4).copy()
    start = r[:, 0].astype(np.int64)
    dur = r[:, 3].astype(np.float64)
    fast = r[:, 2]
    start -= start.min()
    end = start + dur
    span = end.max()
    n_workers = int(os.environ.get("DSB_WARPS", "3552"))
    out = {"k_classify_ms": c.value, "span_us": float(span), "n_reads": int(b.n_reads),
           "dur_us": {k: float(np.percentile(dur, q)) for k, q in (("p10", 10), ("p50", 50), ("p90", 90), ("p99", 99), ("max", 100))},
           "dur_mean_us": float(dur.mean()), "sum_dur_ms": float(dur.sum() / 1e3),
           "busy_fraction": float(dur.sum() / (span * n_workers)),
           "fast_classify_share": float((fast != 0).mean()),
           "dur_mean_fast_us": float(dur[fast != 0].mean()) if (fast != 0).any() else None,
           "dur_mean_slow_us": float(dur[fast == 0].mean()) if (fast == 0).any() else None}
    # active warps over time (20 bins)
    bins = np.linspace(0, span, 21)
    act = []
    for lo, hi in zip(bins[:-1], bins[1:]):
        act.append(float((np.minimum(end, hi) - np.maximum(start, lo)).clip(0).sum() / (hi - lo)))
    out["active_warps_by_twentieth"] = [round(x) for x in act]
    # list scheduling with the measured durations (assumes a read's time does not depend on what runs beside it)
    order = np.argsort(start, kind="stable")
    out["sim_makespan_us"] = {"as_run": makespan(dur[order], n_workers),
                              "longest_first": makespan(np.sort(dur)[::-1], n_workers),
                              "ideal": float(dur.sum() / n_workers)}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    open(os.path.join(ROOT, "gpurun_out", "read_times.json"), "w").write(json.dumps(out) + "\n")
    np.save(os.path.join(ROOT, "gpurun_out", "read_times.npy"), np.stack([start, dur, fast.astype(np.int64)], 1))
    print(json.dumps(out), file=sys.stderr)


if __name__ == "__main__":
    main()
