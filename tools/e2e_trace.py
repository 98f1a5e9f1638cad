#!/usr/bin/env python3
"""read_classify() on the bench workload with DSB_VERBOSE=1: the host pipeline's own stage timings (stderr)."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "desamba-so_b200"))
import dataset   # noqa: E402
import pydsb     # noqa: E402

m = dataset.build(os.environ.get("DSB_WORKLOAD", "cfg3"), golden=False)
d = pydsb.Desamba()
d.load_index(m["idx"])
text = open(m["reads"], "rb").read() * int(os.environ.get("DSB_TRACE_COPIES", "1"))
for i in range(2):
    d.read_classify_timed(text, thread_id=0, thread_num=1)
os.environ["DSB_VERBOSE"] = "1"
for i in range(2):
    t, n = d.read_classify_timed(text, thread_id=0, thread_num=1)
    print("call %d: %.2f ms, %d bytes" % (i, 1e3 * t, n), file=sys.stderr)
