#!/usr/bin/env python3
"""Per-device-function breakdown of an ncu capture of k_classify.

usage: tools/ncu_by_function.py <report.ncu-rep> [libdsb_cuda.so] [--kernel k_classify] [--lines N]

ncu's SASS source page lists every instruction of the kernel (the out-of-line device functions are part of
the kernel's section) with stall samples and executed-instruction counts but without function names;
nvdisasm of the cubin gives the function labels in the same instruction order.  This joins the two by
instruction index and prints, per function: stall samples (share), long-scoreboard / no-instruction /
wait samples, warp instructions executed, average active threads.  With --lines N it also prints the N
hottest SASS instructions with their function.
"""
import csv
import os
import re
import subprocess
import sys
import tempfile


def function_ranges(so_path, kernel="k_classify"):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so_path)], cwd=d, stdout=subprocess.DEVNULL, check=True)
    txt = ""
    for cubin in sorted(f for f in os.listdir(d) if f.endswith(".cubin")):      # the library holds one cubin per .cu file
        t = subprocess.run(["nvdisasm", "-c", os.path.join(d, cubin)], stdout=subprocess.PIPE, check=True).stdout.decode()
        if re.search(r"^\.text\.\S*" + re.escape(kernel), t, re.M):
            txt = t
            break
    names, in_kernel, idx = [], False, 0
    for line in txt.splitlines():
        m = re.match(r"^\.text\.(\S+):", line)
        if m:
            in_kernel = kernel in m.group(1)
            if in_kernel:
                names.append((0, kernel))
                idx = 0
            continue
        if not in_kernel:
            continue
        m = re.match(r"^\$\S+\$(_Z\w+):", line)
        if m:
            sym = m.group(1)
            mi = re.match(r"_ZN\d+_INTERNAL_[0-9a-f]{8}_\d+_\w+?_cu_[0-9a-f]{8}(\d+)", sym)     # file-local (static) functions
            if mi:
                n = int(mi.group(1))
                names.append((idx, sym[mi.end():mi.end() + n]))
                continue
            mm = re.match(r"_ZL?(\d+)", sym)
            if not mm:
                names.append((idx, sym))
                continue
            n = int(mm.group(1))
            names.append((idx, sym[mm.end():mm.end() + n]))
            continue
        if re.match(r"^\s+/\*[0-9a-f]{4,}\*/", line):
            idx += 1
    return names, idx


def main():
    rep = sys.argv[1]
    so = "desamba-so_b200/lib/libdsb_cuda.so"
    n_lines = 0
    kernel = "k_classify"
    args = sys.argv[2:]
    while args:
        a = args.pop(0)
        if a == "--lines":
            n_lines = int(args.pop(0))
        elif a == "--kernel":
            kernel = args.pop(0)
        else:
            so = a
    out = subprocess.run(["ncu", "-i", rep, "--kernel-name", "regex:^" + kernel, "--page", "source", "--csv"],
                         stdout=subprocess.PIPE, check=True).stdout.decode()
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    col = {n: i for i, n in enumerate(hdr)}
    body = []
    for r in rows[2:]:                      # one section per captured launch of the kernel: keep the first
        if r and r[0] == "Kernel Name":
            break
        body.append(r)
    ranges, n_inst = function_ranges(so, kernel)
    if n_inst != len(body):
        print("warning: nvdisasm has %d instructions, the report %d (different build?)" % (n_inst, len(body)), file=sys.stderr)
    starts = [r[0] for r in ranges]

    def func_of(i):
        lo = 0
        for k, s in enumerate(starts):
            if s <= i:
                lo = k
        return ranges[lo][1]

    agg = {}
    keys = ["# Samples", "stall_long_sb", "stall_no_inst", "stall_wait", "stall_short_sb", "stall_branch_resolving",
            "Instructions Executed", "Thread Instructions Executed", "L2 Theoretical Sectors Global",
            "L2 Theoretical Sectors Local"]
    hot = []
    for i, r in enumerate(body):
        f = func_of(i)
        a = agg.setdefault(f, [0] * len(keys))
        for k, name in enumerate(keys):
            try:
                a[k] += int(float(r[col[name]]))
            except (ValueError, KeyError):
                pass
        hot.append((int(float(r[col["# Samples"]] or 0)), f, r[col["Source"]].strip(), int(float(r[col["stall_long_sb"]] or 0))))
    tot = sum(a[0] for a in agg.values()) or 1
    tot_i = sum(a[6] for a in agg.values()) or 1
    print("%-28s %7s %6s %8s %8s %8s %8s %6s %6s %9s %9s" % ("function", "samples", "%", "long_sb", "no_inst", "wait", "inst(M)",
                                                             "inst%", "thr", "gsect(M)", "lsect(M)"))
    for f, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print("%-28s %7d %6.1f %8d %8d %8d %8.1f %6.1f %6.1f %9.1f %9.1f" % (
            f, a[0], 100.0 * a[0] / tot, a[1], a[2], a[3], a[6] / 1e6, 100.0 * a[6] / tot_i, a[7] / a[6] if a[6] else 0,
            a[8] / 1e6, a[9] / 1e6))
    print("total samples %d, warp instructions %.2f G; gsect/lsect = sectors the global / local accesses would move with no "
          "L1 (L2 Theoretical Sectors), %.0f M / %.0f M in total" % (tot, tot_i / 1e9, sum(a[8] for a in agg.values()) / 1e6,
                                                                      sum(a[9] for a in agg.values()) / 1e6))
    if n_lines:
        print()
        for s, f, src, lsb in sorted(hot, key=lambda x: -x[0])[:n_lines]:
            print("%7d %5.2f%% lsb=%-6d %-22s %s" % (s, 100.0 * s / tot, lsb, f, src))


if __name__ == "__main__":
    main()
