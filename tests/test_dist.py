"""N>1 path on CPU, world_size 2 over gloo: what `bench.py --gpus N` does under torchrun, with the host simulator in place
of the GPUs.  Every rank is its own stream of reads against its own index replica (the reference analogue: one worker
per read set, `worker_for`, reference src/cly_mt.c:352-359): the ranks classify DISJOINT shards of a read set through the
product's host pipeline (libdesamba_hostsim.so = the real host C code over the kernel source compiled for the host), the
texts are gathered on rank 0 and compared with the unmodified reference's answer for each shard; the bench's plumbing
(barrier, max and sum over ranks) runs over the same process group.  The data path has no collective."""
import json
import os
import subprocess
import sys
import textwrap

import pytest

from conftest import ROOT, fastq_slices, ref_cli

WORKER = """
    import hashlib, json, os, sys, time
    import torch, torch.distributed as dist
    dist.init_process_group("gloo")
    r, n = dist.get_rank(), dist.get_world_size()
    root, idx, shard_dir = sys.argv[1], sys.argv[2], sys.argv[3]
    sys.path.insert(0, os.path.join(root, "desamba-so_b200"))
    sys.path.insert(0, os.path.join(root, "tests", "hostsim"))
    import pydsb_hostsim
    d = pydsb_hostsim.Desamba()
    d.load_index(idx)
    text = open(os.path.join(shard_dir, "shard%d.fq" % r), "rb").read()
    dist.barrier()
    t0 = time.perf_counter()
    sam = d.read_classify(text, thread_id=0, thread_num=1)
    dt = time.perf_counter() - t0
    bases = sum(len(l) for l in text.split(b"\\n")[1::4])
    # the bench's reductions: time = max over ranks, work = sum over ranks
    t = torch.tensor([dt], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    b = torch.tensor([float(bases)], dtype=torch.float64)
    dist.all_reduce(b, op=dist.ReduceOp.SUM)
    out = [None] * n
    dist.all_gather_object(out, sam)          # results gathered on the host, in rank order = read order
    dist.barrier()
    if r == 0:
        open(os.path.join(shard_dir, "merged.sam"), "wb").write(b"".join(out))
        print(json.dumps({"t_max": t.item(), "bases": b.item(), "mine": dt, "sizes": [len(x) for x in out]}))
    dist.destroy_process_group()
"""


def test_two_ranks_classify_disjoint_shards_and_gather(mixed, have_ref_binary, tmp_path):
    if not have_ref_binary:
        pytest.skip("reference binary absent")
    shards = fastq_slices(mixed["reads"], 2)
    assert len(shards) == 2
    want = b""
    n_bases = 0
    for r, fq in enumerate(shards):
        p = tmp_path / ("shard%d.fq" % r)
        p.write_bytes(fq)
        n_bases += sum(len(x) for x in fq.split(b"\n")[1::4])
        # every rank is a worker of its own: the reference's answer for the shard as a stream of its own, SAM_FULL like the API
        want += ref_cli(["-f", "SAM_FULL", mixed["idx"] + "/", str(p)])
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent(WORKER))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script), ROOT, mixed["idx"], str(tmp_path)],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    d = json.loads([ln for ln in r.stdout.decode().splitlines() if ln.startswith("{")][-1])
    assert d["bases"] == float(n_bases)
    assert d["t_max"] >= d["mine"] > 0
    assert (tmp_path / "merged.sam").read_bytes() == want
