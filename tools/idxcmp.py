#!/usr/bin/env python3
"""Compare two deSAMBA index directories file by file (test infrastructure).

Byte-identical except where the reference's own output is unspecified: in .ref_i the bytes after each reference
name's NUL are whatever the reference's heap held (reference src/idx.c:587-591 strcpy into realloc'd memory) and are
masked on both sides."""
import os
import sys

FILES = [".bwt", ".sa", ".acg", ".exki", ".exk0", ".exk1", ".unv", ".ref_b", ".ref_i", ".ref_p"]


def norm_ref_i(b):
    n = int.from_bytes(b[:8], "little")
    out = bytearray(b)
    for r in range(n):
        o = 8 + 144 * r
        z = b.index(0, o, o + 128) if 0 in b[o:o + 128] else o + 128
        for i in range(z, o + 128):
            out[i] = 0
    return bytes(out)


def compare(a, b):
    bad = []
    for s in FILES:
        pa, pb = os.path.join(a, "deSAMBA" + s), os.path.join(b, "deSAMBA" + s)
        if not (os.path.exists(pa) and os.path.exists(pb)):
            bad.append((s, "missing"))
            continue
        if os.path.getsize(pa) != os.path.getsize(pb):
            bad.append((s, "size %d vs %d" % (os.path.getsize(pa), os.path.getsize(pb))))
            continue
        if s == ".ref_i":
            da, db = norm_ref_i(open(pa, "rb").read()), norm_ref_i(open(pb, "rb").read())
            if da != db:
                bad.append((s, "content"))
            continue
        with open(pa, "rb") as fa, open(pb, "rb") as fb:
            off = 0
            while True:
                ca, cb = fa.read(1 << 24), fb.read(1 << 24)
                if ca != cb:
                    i = next(i for i in range(min(len(ca), len(cb))) if ca[i] != cb[i])
                    bad.append((s, "first difference at byte %d" % (off + i)))
                    break
                if not ca:
                    break
                off += len(ca)
    return bad


if __name__ == "__main__":
    bad = compare(sys.argv[1], sys.argv[2])
    for s, why in bad:
        print("DIFF deSAMBA%s: %s" % (s, why))
    print("identical" if not bad else "%d files differ" % len(bad))
    sys.exit(1 if bad else 0)
