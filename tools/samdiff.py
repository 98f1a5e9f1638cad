#!/usr/bin/env python3
"""Read-level comparison of two SAM files: prints the read names whose record sets differ."""
import sys
from collections import defaultdict


def load(path):
    d = defaultdict(list)
    for line in open(path, "rb"):
        f = line.rstrip(b"\n").split(b"\t")
        d[f[0]].append(b"\t".join(f[:6] + f[11:12]))
    return d


def differing(a, b):
    A, B = load(a), load(b)
    return sorted(k.decode() for k in set(A) | set(B) if A.get(k) != B.get(k))


if __name__ == "__main__":
    for n in differing(sys.argv[1], sys.argv[2]):
        print(n)
