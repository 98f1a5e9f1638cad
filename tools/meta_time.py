#!/usr/bin/env python3
"""BASELINE.json configs[4] step at bench scale: read_classify() of the bench reads, then meta_analysis(META_USE_BASE_NUM,
65536) on the returned block -- from the hit records the batcher kept (default) and through the text parser (DSB_META_TEXT=1)."""
import hashlib
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "desamba-so_b200"))
import dataset   # noqa: E402
import pydsb     # noqa: E402

m = dataset.build(os.environ.get("DSB_WORKLOAD", "cfg3"), golden=False)
text = open(m["reads"], "rb").read()
res = {}
for mode in ("records", "text"):
    if mode == "text":
        os.environ["DSB_META_TEXT"] = "1"
    d = pydsb.Desamba()
    d.load_index(m["idx"])
    for i in range(3):
        sam, meta, snap, tc, tm = d.classify_then_analyse(text, thread_id=0, flag=1, max_snapshot_len=65536, calls=3)
    res[mode] = (hashlib.md5(sam).hexdigest(), meta, snap)
    print("%-8s path %d: read_classify %.1f ms, meta_analysis %.2f ms per call (%d bytes of SAM_FULL)" %
          (mode, d.meta_path, 1e3 * tc, 1e3 * tm, len(sam)), flush=True)
    del d
print("identical:", res["records"] == res["text"])
print(res["records"][1].decode())
