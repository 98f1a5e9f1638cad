#!/usr/bin/env python3
"""Same-box comparison of host-pipeline settings: read_classify() on the bench workload, one library instance per setting.
usage: tools/e2e_sweep.py ["VAR=val VAR=val" ...]   (an empty string = defaults)"""
import hashlib
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "desamba-so_b200"))
import dataset   # noqa: E402
import pydsb     # noqa: E402

import subprocess   # noqa: E402

if len(sys.argv) > 1 and sys.argv[1] == "--one":
    m = dataset.build(os.environ.get("DSB_WORKLOAD", "cfg2"), golden=False)
    text = open(m["reads"], "rb").read()
    d = pydsb.Desamba()
    d.load_index(m["idx"])
    md5 = hashlib.md5(d.read_classify(text, thread_id=0, thread_num=1)).hexdigest()
    for _ in range(3):
        d.read_classify_timed(text, thread_id=0, thread_num=1)
    ts = [d.read_classify_timed(text, thread_id=0, thread_num=1)[0] for _ in range(int(os.environ.get("SWEEP_CALLS", "12")))]
    print("mean %.2f ms  min %.2f ms  (%.0f Mbp/s)  md5 %s" % (1e3 * sum(ts) / len(ts), 1e3 * min(ts), m["n_bases"] / (sum(ts) / len(ts)) / 1e6, md5[:8]))
    sys.exit(0)

settings = sys.argv[1:] or [""]
for rnd in range(2):
    for st in settings:
        env = dict(os.environ)
        env.update(dict(x.split("=", 1) for x in st.split()))
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        out = r.stdout.decode().strip() if r.returncode == 0 else "FAILED: " + r.stderr.decode()[-300:]
        print("round %d  %-58s %s" % (rnd, st or "(defaults)", out), flush=True)
