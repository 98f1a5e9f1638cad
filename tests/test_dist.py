"""N>1 path on CPU: the bench's rank plumbing (barrier, max / sum over ranks) over gloo, world_size 2.
The data path itself has no collective: ranks classify disjoint read sets against replicated indexes."""
import os
import subprocess
import sys
import textwrap

from conftest import ROOT


def test_two_rank_gloo_reduction(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent("""
        import os, sys, json
        import torch, torch.distributed as dist
        dist.init_process_group("gloo")
        r = dist.get_rank()
        sys.path.insert(0, %r)
        sys.path.insert(0, os.path.join(%r, "tools"))
        import dataset
        # every rank derives its own read-set name and parameters deterministically
        name = "cfg2" if r == 0 else "cfg2@%%d" %% r
        p = dict(dataset.CONFIGS["cfg2"])
        seed = p["read_seed"] + 1000 * r
        t = torch.tensor([float(10 + r)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        s = torch.tensor([float(100 * (r + 1))], dtype=torch.float64)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        dist.barrier()
        if r == 0:
            print(json.dumps({"max": t.item(), "sum": s.item(), "seed": seed, "name": name}))
        dist.destroy_process_group()
    """ % (ROOT, ROOT)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=300)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    line = [ln for ln in r.stdout.decode().splitlines() if ln.startswith("{")][-1]
    import json
    d = json.loads(line)
    assert d["max"] == 11.0 and d["sum"] == 300.0 and d["name"] == "cfg2"
