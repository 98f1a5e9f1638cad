#!/usr/bin/env python3
"""Synthetic inputs for the deSAMBA classify hot path (test/bench infrastructure).

Nothing ships with the reference (demo/, demo_index/ are git-ignored there), and
Jellyfish is absent, so every index used by tests and bench.py is built from data made here:

  * multi-genome reference FASTA  (families x strains, strains = family base + SNPs),
    headers ``>tid|<taxid>|ref|SYN_<i>`` as the reference's SAM parser expects
    (reference src/cly_mt.c:779-785, build-index:22-24)
  * ``kmer.srt``: u64 n + n ascending distinct u64 31-mers, 2 bit/base A0 C1 G2 T3, first base
    most significant -- the file ``deSAMBA kmersort`` would have produced
    (reference src/idx_sort.c:196-200, src/idx.c:887-892)
  * simulated reads (FASTQ): uniform start, 50 % reverse strand, sub/ins/del errors, junk reads
  * nodes.dmp / names.dmp for load_index (reference src/cly_mt.c:590-670)

Everything is seeded; the same arguments give byte-identical files.
"""
import argparse
import os
import sys

import numpy as np

BASES = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGT", b"TGCA"):
    COMP[a] = b


def make_genomes(n_fam, n_strain, length, snp, seed):
    """Return list of (taxid, name, uint8 ASCII array)."""
    rng = np.random.default_rng(seed)
    out = []
    for f in range(n_fam):
        base = rng.integers(0, 4, size=length, dtype=np.uint8)
        for s in range(n_strain):
            g = base.copy()
            if s > 0 or n_strain == 1:
                pass
            if s > 0:
                m = rng.random(length) < snp
                g[m] = (g[m] + rng.integers(1, 4, size=int(m.sum()), dtype=np.uint8)) & 3
            i = f * n_strain + s
            out.append((1000 + i, "SYN_%d" % i, BASES[g]))
    return out


def write_fasta(path, genomes, human_first=False):
    with open(path, "wb") as fh:
        for k, (tid, name, seq) in enumerate(genomes):
            if human_first and k == 0:
                tid = 9606
            fh.write((">tid|%d|ref|%s\n" % (tid, name)).encode())
            n = len(seq)
            full = (n // 70) * 70
            if full:
                body = np.empty((full // 70, 71), dtype=np.uint8)
                body[:, :70] = seq[:full].reshape(-1, 70)
                body[:, 70] = 10
                fh.write(body.tobytes())
            if full < n:
                fh.write(seq[full:].tobytes() + b"\n")


def write_kmer_srt(path, genomes, k=31):
    code = np.zeros(256, dtype=np.uint64)
    for i, c in enumerate(b"ACGT"):
        code[c] = i
    parts = []
    for _, _, seq in genomes:
        c = code[seq]
        n = len(c) - k + 1
        if n <= 0:
            continue
        v = np.zeros(n, dtype=np.uint64)
        for j in range(k):
            v <<= np.uint64(2)
            v |= c[j:j + n]
        parts.append(v)
    allk = np.unique(np.concatenate(parts))
    with open(path, "wb") as fh:
        fh.write(np.array([len(allk)], dtype=np.uint64).tobytes())
        fh.write(allk.tobytes())
    return len(allk)


def write_taxonomy(dirpath, genomes, n_strain, human_first=False):
    """root 1 -> family genus nodes (tid 500+f) -> strain species tids."""
    nodes, names = [(1, 1, "no rank")], [(1, "root")]
    fams = sorted({(k // n_strain) for k in range(len(genomes))})
    for f in fams:
        nodes.append((500 + f, 1, "genus"))
        names.append((500 + f, "SynGenus_%d" % f))
    for k, (tid, name, _) in enumerate(genomes):
        if human_first and k == 0:
            tid, name = 9606, "Homo sapiens"
        nodes.append((tid, 500 + k // n_strain, "species"))
        names.append((tid, name))
    with open(os.path.join(dirpath, "nodes.dmp"), "w") as fh:
        for t, p, r in nodes:
            fh.write("%d\t|\t%d\t|\t%s\t|\n" % (t, p, r))
    with open(os.path.join(dirpath, "names.dmp"), "w") as fh:
        for t, n in names:
            fh.write("%d\t|\t%s\t|\t\t|\tscientific name\t|\n" % (t, n))


def mutate(seq, err, mix, rng):
    """Apply per-base errors; mix = (p_sub, p_ins, p_del) normalised."""
    n = len(seq)
    r = rng.random(n)
    kind = np.zeros(n, dtype=np.uint8)  # 0 ok 1 sub 2 ins 3 del
    ps, pi, pd = [err * m for m in mix]
    kind[r < ps] = 1
    kind[(r >= ps) & (r < ps + pi)] = 2
    kind[(r >= ps + pi) & (r < ps + pi + pd)] = 3
    out = []
    rb = rng.integers(0, 4, size=n, dtype=np.uint8)
    code = {65: 0, 67: 1, 71: 2, 84: 3}
    # vectorised assembly: each position emits 0, 1 or 2 bases
    emit1 = seq.copy()
    sub = kind == 1
    if sub.any():
        cur = np.array([code[x] for x in seq[sub]], dtype=np.uint8)
        emit1[sub] = BASES[(cur + 1 + rb[sub] % 3) & 3]
    cnt = np.ones(n, dtype=np.int64)
    cnt[kind == 3] = 0
    cnt[kind == 2] = 2
    total = int(cnt.sum())
    res = np.empty(total, dtype=np.uint8)
    pos = np.cumsum(cnt) - cnt
    keep = cnt >= 1
    res[pos[keep]] = emit1[keep]
    ins = kind == 2
    res[pos[ins] + 1] = BASES[rb[ins]]
    return res


def write_reads(path, genomes, n_reads, length, err, seed, junk=0.05, mix=(1 / 3, 1 / 3, 1 / 3),
                len_jitter=0.0, avoid_edge=1000, name_prefix="r"):
    rng = np.random.default_rng(seed)
    glen = np.array([len(g[2]) for g in genomes], dtype=np.int64)
    total_bases = 0
    with open(path, "wb") as fh:
        for i in range(n_reads):
            L = length
            if len_jitter > 0:
                L = max(50, int(length * (1 + len_jitter * (rng.random() * 2 - 1))))
            if rng.random() < junk:
                seq = BASES[rng.integers(0, 4, size=L, dtype=np.uint8)]
                tag = "junk"
            else:
                gi = int(rng.integers(0, len(genomes)))
                span = min(L, int(glen[gi]) - 2 * avoid_edge)
                st = int(rng.integers(avoid_edge, glen[gi] - avoid_edge - span + 1))
                seq = genomes[gi][2][st:st + span]
                rev = rng.random() < 0.5
                if rev:
                    seq = COMP[seq[::-1]]
                seq = mutate(seq, err, mix, rng)
                tag = "g%d_%d_%c" % (gi, st, "-" if rev else "+")
            total_bases += len(seq)
            fh.write(b"@%s%d_%s\n" % (name_prefix.encode(), i, tag.encode()))
            fh.write(seq.tobytes())
            fh.write(b"\n+\n")
            fh.write(b"I" * len(seq))
            fh.write(b"\n")
    return total_bases


def main():
    ap = argparse.ArgumentParser()
    sub = ap.add_subparsers(dest="cmd", required=True)
    r = sub.add_parser("ref")
    r.add_argument("--out", required=True, help="output directory (ref.fa, kmer.srt, nodes.dmp, names.dmp)")
    r.add_argument("--families", type=int, default=10)
    r.add_argument("--strains", type=int, default=5)
    r.add_argument("--length", type=int, default=1000000)
    r.add_argument("--snp", type=float, default=0.01)
    r.add_argument("--seed", type=int, default=42)
    r.add_argument("--human-first", action="store_true")
    q = sub.add_parser("reads")
    q.add_argument("--out", required=True)
    q.add_argument("--families", type=int, default=10)
    q.add_argument("--strains", type=int, default=5)
    q.add_argument("--length", type=int, default=1000000)
    q.add_argument("--snp", type=float, default=0.01)
    q.add_argument("--ref-seed", type=int, default=42)
    q.add_argument("-n", type=int, default=10000)
    q.add_argument("--read-len", type=int, default=5000)
    q.add_argument("--err", type=float, default=0.10)
    q.add_argument("--seed", type=int, default=43)
    q.add_argument("--junk", type=float, default=0.05)
    q.add_argument("--mix", default="1,1,1", help="sub,ins,del weights")
    q.add_argument("--len-jitter", type=float, default=0.0)
    q.add_argument("--prefix", default="r")
    a = ap.parse_args()
    if a.cmd == "ref":
        os.makedirs(a.out, exist_ok=True)
        g = make_genomes(a.families, a.strains, a.length, a.snp, a.seed)
        write_fasta(os.path.join(a.out, "ref.fa"), g, a.human_first)
        n = write_kmer_srt(os.path.join(a.out, "kmer.srt"), g)
        write_taxonomy(a.out, g, a.strains, a.human_first)
        print("genomes=%d bases=%d kmers=%d" % (len(g), sum(len(x[2]) for x in g), n))
    else:
        g = make_genomes(a.families, a.strains, a.length, a.snp, a.ref_seed)
        w = [float(x) for x in a.mix.split(",")]
        s = sum(w)
        tb = write_reads(a.out, g, a.n, a.read_len, a.err, a.seed, a.junk, tuple(x / s for x in w),
                         a.len_jitter, name_prefix=a.prefix)
        print("reads=%d bases=%d" % (a.n, tb))


if __name__ == "__main__":
    sys.exit(main())
