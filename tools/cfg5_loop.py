#!/usr/bin/env python3
"""BASELINE.json configs[4] at bench scale: ONE load_index, then N x (read_classify of one batch of reads + meta_analysis of
its SAM_FULL text) on one thread_id.  Prints per-call times, steady-state Mbp/s, device / host memory before and after.
usage: tools/cfg5_loop.py [n_calls=100] [workload=cfg2]      (parity of this call sequence: tests/test_gpu.py)"""
import ctypes as C
import json
import os
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "desamba-so_b200"))
import dataset   # noqa: E402
import pydsb     # noqa: E402
import psutil    # noqa: E402
import torch     # noqa: E402

n_calls = int(sys.argv[1]) if len(sys.argv) > 1 else 100
m = dataset.build(sys.argv[2] if len(sys.argv) > 2 else "cfg2", golden=False)
text = open(m["reads"], "rb").read()
d = pydsb.Desamba()
t0 = time.perf_counter()
d.load_index(m["idx"])
t_load = time.perf_counter() - t0
proc = psutil.Process()
t_cls, t_meta = [], []
first = None
free0 = rss0 = None
for i in range(n_calls):
    out, out_n = C.c_void_p(), C.c_uint64(0)
    a = time.perf_counter()
    d.lib.read_classify(d.idx, text, len(text), C.byref(out), C.byref(out_n), 0, 1)
    b = time.perf_counter()
    mo, mn, hs, hn = C.c_void_p(), C.c_uint64(0), C.c_void_p(), C.c_uint64(0)
    d.lib.meta_analysis(d.idx, C.cast(out, C.c_char_p), out_n, C.byref(mo), C.byref(mn), 0, 1, 65536, C.byref(hs), C.byref(hn))
    c = time.perf_counter()
    rep = C.string_at(mo.value, mn.value)
    for p in (out, mo, hs):
        if p.value:
            d.libc.free(p)
    if i == 0:
        first = rep
    elif i == 4:
        free0, rss0 = torch.cuda.mem_get_info(0)[0], proc.memory_info().rss
    if i >= 2 and rep != ref_rep:
        raise SystemExit("call %d: the analysis text changed" % i)
    if i == 1:
        ref_rep = rep          # from the second call on the buffer history is periodic
    t_cls.append(b - a)
    t_meta.append(c - b)
free1, rss1 = torch.cuda.mem_get_info(0)[0], proc.memory_info().rss
k = max(1, n_calls // 10)
cls = sorted(t_cls[k:])
met = sorted(t_meta[k:])
tot = sum(t_cls[k:]) + sum(t_meta[k:])
print(json.dumps({
    "workload": m["name"] if "name" in m else "cfg2", "calls": n_calls, "reads_per_call": m["n_reads"], "bases_per_call": m["n_bases"],
    "load_index_s": t_load,
    "read_classify_ms": {"median": 1e3 * cls[len(cls) // 2], "max": 1e3 * cls[-1]},
    "meta_analysis_ms": {"median": 1e3 * met[len(met) // 2], "max": 1e3 * met[-1]},
    "steady_state_Mbp_s_classify_plus_analysis": m["n_bases"] * (n_calls - k) / tot / 1e6,
    "steady_state_Mbp_s_classify_only": m["n_bases"] * (n_calls - k) / sum(t_cls[k:]) / 1e6,
    "device_free_bytes_call5_vs_last": [free0, free1], "host_rss_bytes_call5_vs_last": [rss0, rss1],
    "analysis": first.decode().strip().split("\n")}))
