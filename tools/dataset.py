#!/usr/bin/env python3
"""Build a named synthetic data set for the classify hot path (test / bench infrastructure).

A data set = reference FASTA + kmer.srt + index directory (built by the UNMODIFIED reference
binary oracle/_ref/deSAMBA) + reads FASTQ + the reference's own `classify -t 1`
SAM ("golden").  Everything is derived from seeds, so the GPU box rebuilds byte-identical inputs
from this script alone; nothing here reads /root/reference.

Reads on which the reference binary itself crashes are dropped (they are "reference-undefined":
e.g. sdp_left_M2 reads ref_bin at a negative offset for a chain near the start of reference 0,
reference src/cly.c:2719-2722).  The crashers are found by bisection and recorded in
<dir>/dropped.txt so the same filtered FASTQ is used by both sides.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "deSAMBA")
sys.path.insert(0, HERE)
import synth  # noqa: E402

# name -> parameters.  `cfg1` is the stand-in for BASELINE.json configs[0] (the demo data are
# not in the reference tree), `cfg2` is configs[1] (the bench workload), `tiny` is for smoke.
CONFIGS = {
    "tiny": dict(families=4, strains=3, length=80000, snp=0.01, ref_seed=7, human_first=False,
                 n_reads=300, read_len=5000, err=0.10, read_seed=9, junk=0.05, mix="1,1,1", jitter=0.0),
    "mixed": dict(families=4, strains=3, length=80000, snp=0.01, ref_seed=7, human_first=False,
                  n_reads=400, read_len=2000, err=0.12, read_seed=11, junk=0.10, mix="1,1,1", jitter=0.95),
    "cfg1": dict(families=1, strains=5, length=1000000, snp=0.01, ref_seed=1, human_first=True,
                 n_reads=2000, read_len=5000, err=0.10, read_seed=2, junk=0.05, mix="1,1,1", jitter=0.0),
    "cfg2": dict(families=10, strains=5, length=1000000, snp=0.01, ref_seed=42, human_first=False,
                 n_reads=10000, read_len=5000, err=0.10, read_seed=43, junk=0.05, mix="1,1,1", jitter=0.0),
    "ont10k": dict(families=10, strains=5, length=1000000, snp=0.01, ref_seed=42, human_first=False,
                   n_reads=4000, read_len=10000, err=0.10, read_seed=45, junk=0.05, mix="3,3,4", jitter=0.5),
    # BASELINE.json configs[2] scaled to what the reference's index builder finishes in minutes on the GPU box
    # (4 Gbp needs > 60 GB of host memory and hours): 200 Mbp = 40 families x 5 strains x 1 Mbp,
    # 20 000 ONT-like 10 kb reads (10 % error, 30/30/40 sub/ins/del)
    "cfg3s": dict(families=40, strains=5, length=1000000, snp=0.01, ref_seed=44, human_first=False,
                  n_reads=20000, read_len=10000, err=0.10, read_seed=45, junk=0.05, mix="3,3,4", jitter=0.5),
}


# BASELINE.json configs[2] at the size this box builds in seconds with OUR index builder (`deSAMBA index` on the
# GPU, byte-identical to the reference builder on tiny / cfg2): 1 Gbp = 200 families x 5 strains x 1 Mbp
# (BWT ~0.95 G rows, exist tables 2 x 256 MiB, k = 17) and 100 000 ONT-like 10 kb reads (1 Gbase) per read set;
# the reference's own `-t 1` answer (the golden) is computed for the first `golden_reads` reads.
CONFIGS["cfg3"] = dict(families=200, strains=5, length=1000000, snp=0.01, ref_seed=44, human_first=False,
                       n_reads=100000, read_len=10000, err=0.10, read_seed=45, junk=0.05, mix="3,3,4", jitter=0.5,
                       builder="gpu", sim="c", golden_reads=20000)
# configs[2] at its full index size: 4 Gbp = 800 x 5 x 1 Mbp (BWT ~3.8 G rows, exist tables 2 x 1 GiB, k = 18)
CONFIGS["cfg3x"] = dict(families=800, strains=5, length=1000000, snp=0.01, ref_seed=44, human_first=False,
                        n_reads=100000, read_len=10000, err=0.10, read_seed=45, junk=0.05, mix="3,3,4", jitter=0.5,
                        builder="gpu", sim="c", golden_reads=5000)
# small twin of cfg3 for CPU-side checks of the tooling (reference builder, C simulator)
CONFIGS["cfg3t"] = dict(families=4, strains=3, length=80000, snp=0.01, ref_seed=7, human_first=False,
                        n_reads=300, read_len=4000, err=0.10, read_seed=45, junk=0.05, mix="3,3,4", jitter=0.5,
                        sim="c", golden_reads=200)

# BASELINE.json configs[3] (read-length / error sweep) at test scale, on the cfg2 genomes: ~3 Mbp per cell, all 24 cells
for _L in (1000, 2000, 5000, 10000, 20000, 50000):
    for _e in (5, 10, 15, 20):
        CONFIGS["sw_L%d_e%d" % (_L, _e)] = dict(families=10, strains=5, length=1000000, snp=0.01, ref_seed=42, human_first=False,
                                                n_reads=max(40, 3000000 // _L), read_len=_L, err=_e / 100.0,
                                                read_seed=100 + _L // 1000 + _e, junk=0.05, mix="3,3,4", jitter=0.3)
# reads longer than 65 535 bases: the 32-bit form of the read's 9-mer table heads (dsb_core.cuh DSB_HEAD_NARROW)
CONFIGS["sw_L80000_e10"] = dict(families=10, strains=5, length=1000000, snp=0.01, ref_seed=42, human_first=False,
                                n_reads=40, read_len=80000, err=0.10, read_seed=190, junk=0.05, mix="3,3,4", jitter=0.1)
SWEEP = [k for k in CONFIGS if k.startswith("sw_")]


def data_root():
    return os.environ.get("DSB_DATA", "/tmp/dsb_data")


def ref_key(p):
    return "ref_f%d_s%d_l%d_snp%g_seed%d_h%d" % (p["families"], p["strains"], p["length"], p["snp"], p["ref_seed"],
                                                 int(p["human_first"]))


def run(cmd, **kw):
    return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, **kw)


OUR_BIN = os.path.join(ROOT, "desamba-so_b200", "bin", "deSAMBA")


class _Lock:
    """Cross-process lock on a data directory (the ranks of one bench run share /tmp/dsb_data)."""

    def __init__(self, path):
        self.path = path

    def __enter__(self):
        import fcntl
        os.makedirs(os.path.dirname(self.path), exist_ok=True)
        self.fh = open(self.path, "w")
        fcntl.flock(self.fh, fcntl.LOCK_EX)

    def __exit__(self, *a):
        import fcntl
        fcntl.flock(self.fh, fcntl.LOCK_UN)
        self.fh.close()


def build_index(p, log=sys.stderr):
    """Reference FASTA + index directory shared by every read set on the same genomes.

    builder "ref" (default): the UNMODIFIED reference binary (`deSAMBA index` of oracle/_ref) from a kmer.srt written
    here; builder "gpu": this repository's own `deSAMBA index` (GPU; byte-identical to the reference builder where both
    run, tests/test_gpu.py::test_index_builder_equals_reference) -- the only one that finishes a >= 1 Gbp index in
    minutes.  DSB_INDEX_BUILDER=hostsim (local experiments without a GPU) runs the builder's host twin instead."""
    d = os.path.join(data_root(), ref_key(p))
    idx = os.path.join(d, "idx")
    with _Lock(os.path.join(data_root(), ref_key(p) + ".lock")):
        if os.path.exists(os.path.join(d, "index.ok")):
            return d, idx
        os.makedirs(idx, exist_ok=True)
        t0 = time.time()
        g = synth.make_genomes(p["families"], p["strains"], p["length"], p["snp"], p["ref_seed"])
        synth.write_fasta(os.path.join(d, "ref.fa"), g, p["human_first"])
        synth.write_taxonomy(idx, g, p["strains"], p["human_first"])
        t1 = time.time()
        builder = p.get("builder", "ref")
        if builder == "gpu":
            del g
            exe = OUR_BIN
            if os.environ.get("DSB_INDEX_BUILDER") == "hostsim":
                exe = os.path.join(ROOT, "tests", "hostsim", "deSAMBA_hostsim")
            r = run([exe, "index", "ref.fa", "idx"], cwd=d)
            if r.returncode != 0 or not os.path.exists(os.path.join(idx, "deSAMBA.bwt")):
                raise RuntimeError("index build failed:\n" + r.stderr.decode()[-2000:])
            open(os.path.join(d, "build.log"), "wb").write(r.stderr)
        else:
            if not os.path.exists(REF_BIN):
                raise RuntimeError("oracle/_ref/deSAMBA is missing: run `make -C oracle ref` where /root/reference exists")
            synth.write_kmer_srt(os.path.join(d, "kmer.srt"), g)
            r = run([REF_BIN, "index", "kmer.srt", "ref.fa", "idx"], cwd=d)
            if r.returncode != 0 or not os.path.exists(os.path.join(idx, "deSAMBA.bwt")):
                raise RuntimeError("reference index build failed:\n" + r.stderr.decode()[-2000:])
            os.remove(os.path.join(d, "kmer.srt"))
        open(os.path.join(d, "index.ok"), "w").write("ok\n")
        print("[dataset] index %s: genomes + FASTA %.1fs, %s builder %.1fs" % (ref_key(p), t1 - t0, builder, time.time() - t1),
              file=log)
    return d, idx


def readsim_exe():
    """tools/readsim.c compiled on first use (gcc is on both machines)."""
    exe = os.path.join(data_root(), "readsim")
    src = os.path.join(HERE, "readsim.c")
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
        os.makedirs(data_root(), exist_ok=True)
        tmp = exe + ".%d" % os.getpid()
        r = run(["gcc", "-O2", "-fopenmp", "-o", tmp, src])
        if r.returncode != 0:
            raise RuntimeError("readsim build failed:\n" + r.stderr.decode())
        os.replace(tmp, exe)
    return exe


def read_fastq_records(path):
    with open(path, "rb") as fh:
        lines = fh.read().split(b"\n")
    return [b"\n".join(lines[i:i + 4]) + b"\n" for i in range(0, len(lines) - 1, 4)]


def ref_classify(idx, fq, out, threads=1, fmt="SAM"):
    with open(out, "wb") as fo:
        r = subprocess.run([REF_BIN, "classify", "-t", str(threads), "-f", fmt, idx + "/", fq], stdout=fo,
                           stderr=subprocess.PIPE)
    return r.returncode


def find_crashers(idx, recs, tmp):
    """Indices of records on which the reference crashes when classified alone-in-context."""
    bad = []

    def crashes(ids):
        with open(tmp, "wb") as fh:
            for i in ids:
                fh.write(recs[i])
        return ref_classify(idx, tmp, tmp + ".sam") != 0

    def rec(ids):
        if not crashes(ids):
            return
        if len(ids) == 1:
            bad.append(ids[0])
            return
        h = len(ids) // 2
        rec(ids[:h])
        rec(ids[h:])

    chunk = 256
    for s in range(0, len(recs), chunk):
        rec(list(range(s, min(s + chunk, len(recs)))))
    return bad


def build(name, log=sys.stderr, golden=True):
    """Returns dict(dir, idx, reads, golden, n_reads, n_bases)."""
    p = dict(CONFIGS[name.split("@")[0]])
    if "@" in name:                      # "<cfg>@<k>": same genomes, k-th independent read set (bench shards)
        p["read_seed"] += 1000 * int(name.split("@")[1])
    rd, idx = build_index(p, log)
    d = os.path.join(data_root(), name)
    meta_path = os.path.join(d, "meta.json")
    if os.path.exists(meta_path):
        meta = json.load(open(meta_path))
        if not golden or meta.get("golden"):
            return meta
    os.makedirs(d, exist_ok=True)
    t0 = time.time()
    if p.get("sim") == "c":
        return _build_large(name, p, rd, idx, d, golden, log)
    g = synth.make_genomes(p["families"], p["strains"], p["length"], p["snp"], p["ref_seed"])
    w = [float(x) for x in p["mix"].split(",")]
    raw = os.path.join(d, "raw.fq")
    synth.write_reads(raw, g, p["n_reads"], p["read_len"], p["err"], p["read_seed"], p["junk"],
                      tuple(x / sum(w) for x in w), p["jitter"])
    recs = read_fastq_records(raw)
    fq = os.path.join(d, "reads.fq")
    sam = os.path.join(d, "golden.sam")
    dropped = []
    committed = os.path.join(ROOT, "tests", "golden", name + ".dropped.txt")
    if os.path.exists(committed):
        names = set(open(committed).read().split())
        dropped = [i for i, r in enumerate(recs) if r.split(b"\n", 1)[0][1:].decode() in names]
    for _ in range(4):
        keep = [r for i, r in enumerate(recs) if i not in set(dropped)]
        with open(fq, "wb") as fh:
            fh.write(b"".join(keep))
        if not golden:
            break
        if ref_classify(idx, fq, sam) == 0:
            break
        print("[dataset] %s: reference crashed, bisecting" % name, file=log)
        live = [i for i in range(len(recs)) if i not in set(dropped)]
        sub = [recs[i] for i in live]
        dropped += [live[j] for j in find_crashers(idx, sub, os.path.join(d, "bisect.fq"))]
    else:
        raise RuntimeError("reference keeps crashing on " + name)
    os.remove(raw)
    with open(os.path.join(d, "dropped.txt"), "w") as fh:
        for i in dropped:
            fh.write(recs[i].split(b"\n", 1)[0][1:].decode() + "\n")
    n_bases = sum(len(r.split(b"\n")[1]) for i, r in enumerate(recs) if i not in set(dropped))
    meta = dict(name=name, dir=d, idx=idx, reads=fq, golden=sam if golden else None,
                n_reads=len(recs) - len(dropped), n_bases=n_bases, dropped=len(dropped))
    if golden:
        meta["golden_md5"] = hashlib.md5(open(sam, "rb").read()).hexdigest()
    json.dump(meta, open(meta_path, "w"))
    print("[dataset] %s: %d reads, %d bases, %d dropped, %.1fs" % (name, meta["n_reads"], n_bases, len(dropped),
                                                                 time.time() - t0), file=log)
    return meta


def _build_large(name, p, rd, idx, d, golden, log):
    """Read sets of the C simulator: `reads` = the whole set (bench), `parity_reads` = its first golden_reads reads (what
    the reference is run on: golden, twins, and the reference arm of bench.py).  Reads the reference crashes on are
    dropped from the parity subset only."""
    t0 = time.time()
    exe = readsim_exe()
    fq = os.path.join(d, "reads.fq")
    args = [str(p["read_len"]), str(p["err"]), str(p["read_seed"]), str(p["junk"]), p["mix"], str(p["jitter"]), "r"]
    r = run([exe, os.path.join(rd, "ref.fa"), fq, str(p["n_reads"])] + args)
    if r.returncode != 0:
        raise RuntimeError("readsim failed: " + r.stderr.decode())
    n_bases = int(r.stdout.decode().split("bases=")[1])
    meta = dict(name=name, dir=d, idx=idx, reads=fq, golden=None, n_reads=p["n_reads"], n_bases=n_bases, dropped=0)
    if golden:
        ng = min(p.get("golden_reads", p["n_reads"]), p["n_reads"])
        raw = os.path.join(d, "parity_raw.fq")
        r = run([exe, os.path.join(rd, "ref.fa"), raw, str(ng)] + args)   # per-read seeding: the first ng reads of the set
        if r.returncode != 0:
            raise RuntimeError("readsim failed: " + r.stderr.decode())
        recs = read_fastq_records(raw)
        pfq, sam = os.path.join(d, "parity.fq"), os.path.join(d, "golden.sam")
        dropped = []
        for _ in range(4):
            keep = [x for i, x in enumerate(recs) if i not in set(dropped)]
            with open(pfq, "wb") as fh:
                fh.write(b"".join(keep))
            if ref_classify(idx, pfq, sam) == 0:
                break
            print("[dataset] %s: reference crashed, bisecting" % name, file=log)
            live = [i for i in range(len(recs)) if i not in set(dropped)]
            dropped += [live[j] for j in find_crashers(idx, [recs[i] for i in live], os.path.join(d, "bisect.fq"))]
        else:
            raise RuntimeError("reference keeps crashing on " + name)
        os.remove(raw)
        meta.update(parity_reads=pfq, golden=sam, n_parity_reads=len(recs) - len(dropped),
                    n_parity_bases=sum(len(x.split(b"\n")[1]) for i, x in enumerate(recs) if i not in set(dropped)),
                    dropped=len(dropped), golden_md5=hashlib.md5(open(sam, "rb").read()).hexdigest())
    json.dump(meta, open(os.path.join(d, "meta.json"), "w"))
    print("[dataset] %s: %d reads, %d bases, %.1fs" % (name, meta["n_reads"], n_bases, time.time() - t0), file=log)
    return meta


TWINS = ("deSAMBA_zinit", "deSAMBA_pinit", "deSAMBA_o1")


def _records_by_read(path):
    out = {}
    for line in open(path, "rb"):
        f = line.rstrip(b"\n").split(b"\t")
        out.setdefault(f[0], []).append((f[1:6], f[11:12]))
    return out


def undefined_reads(meta, log=sys.stderr):
    """Names of the reads whose REFERENCE answer is not defined by the reference's source: any instrumented
    build of the unmodified sources (oracle/Makefile `twins`: zero- / pattern-initialised automatic variables,
    -O1 frame layout) classifies them differently from the stock -O3 binary at -t 1.  Cached per data set.

    A twin that crashes gives no verdict on the reads it did not reach (they are NOT counted as undefined); the read it
    stopped at is undefined (a build of the same sources crashing on it is undefined behaviour by definition).  The
    per-twin counts and crashes are kept in <dir>/undefined.json; the twins' own answers stay in <dir>/<twin>.sam so
    that tests can require our answer on an undefined read to be ONE of the reference builds' answers."""
    path = os.path.join(meta["dir"], "undefined.txt")
    if os.path.exists(path) and os.path.exists(os.path.join(meta["dir"], "undefined.json")):
        return set(open(path).read().split())
    reads = meta.get("parity_reads") or meta["reads"]
    stock = _records_by_read(meta["golden"])
    order = []
    for line in open(meta["golden"], "rb"):
        k = line.split(b"\t", 1)[0]
        if not order or order[-1] != k:
            order.append(k)
    bad = set()
    info = {"n_reads": len(order), "twins": {}}
    procs = []
    t0 = time.time()
    for tw in TWINS:
        b = os.path.join(ROOT, "oracle", "_ref", tw)
        if not os.path.exists(b):
            raise RuntimeError("%s is missing: run `make -C oracle twins` where /root/reference exists" % b)
        out = os.path.join(meta["dir"], tw + ".sam")
        fo = open(out, "wb")
        procs.append((tw, out, fo, subprocess.Popen([b, "classify", "-t", "1", "-f", "SAM", meta["idx"] + "/", reads],
                                                    stdout=fo, stderr=subprocess.DEVNULL)))
    for tw, out, fo, pr in procs:
        rc = pr.wait()
        fo.close()
        other = _records_by_read(out)
        reached = [k for k in order if k in other]
        d = {k.decode() for k in reached if other[k] != stock[k]}
        ti = {"rc": rc, "reached": len(reached), "differ": len(d)}
        if rc != 0 or len(reached) < len(order):
            nxt = next((k for k in order if k not in other), None)
            if nxt is not None:
                d.add(nxt.decode())
                ti["stopped_at"] = nxt.decode()
            print("[dataset] twin %s stopped on %s after %d of %d reads (rc %d): no verdict on the rest" %
                  (tw, meta["name"], len(reached), len(order), rc), file=log)
        info["twins"][tw] = ti
        print("[dataset] %s vs stock on %s: %d reads differ" % (tw, meta["name"], len(d)), file=log)
        bad |= d
    info["undefined"] = len(bad)
    info["seconds"] = round(time.time() - t0, 1)
    json.dump(info, open(os.path.join(meta["dir"], "undefined.json"), "w"), indent=1)
    with open(path, "w") as fh:
        fh.write("\n".join(sorted(bad)) + "\n")
    return bad


def build_many(names, golden=True, workers=None, log=sys.stderr):
    """Build several data sets (and their undefined-read lists) with a process pool: the reference's `-t 1` runs are
    serial and there are many small sets (the 24 sweep cells)."""
    import concurrent.futures as cf
    metas = {}
    for n in names:                       # indexes first, one at a time (shared by the cells)
        build_index(dict(CONFIGS[n.split("@")[0]]), log)
    with cf.ProcessPoolExecutor(max_workers=workers or max(1, (os.cpu_count() or 4) // 2)) as ex:
        futs = {ex.submit(_build_one, n, golden): n for n in names}
        for f in cf.as_completed(futs):
            metas[futs[f]] = f.result()
    return metas


def _build_one(name, golden):
    m = build(name, log=open(os.devnull, "w"), golden=golden)
    if golden:
        undefined_reads(m, log=open(os.devnull, "w"))
    return m


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("name")
    ap.add_argument("--no-golden", action="store_true")
    a = ap.parse_args()
    print(json.dumps(build(a.name, golden=not a.no_golden), indent=1))
