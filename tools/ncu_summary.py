#!/usr/bin/env python3
"""Per-kernel summary of an ncu --set full capture: duration, DRAM bytes, L1 / L2 hit rates, issue activity, active threads
per instruction, local vs global sectors.   usage: tools/ncu_summary.py report.ncu-rep [--md]"""
import csv
import io
import subprocess
import sys

WANT = [("gpu__time_duration.sum", "ms", 1e-6), ("dram__bytes_read.sum", "dram_rd_GB", 1e-9), ("dram__bytes_write.sum", "dram_wr_GB", 1e-9),
        ("lts__t_sector_hit_rate.pct", "l2_hit_%", 1), ("l1tex__t_sector_hit_rate.pct", "l1_hit_%", 1),
        ("sm__inst_executed.sum", "warp_inst_G", 1e-9), ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_%", 1), ("smsp__issue_active.avg.pct", "issue_%", 1),
        ("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "lsect_ld_M", 1e-6), ("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum", "lsect_st_M", 1e-6),
        ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "gsect_ld_M", 1e-6), ("l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "gsect_st_M", 1e-6),
        ("launch__registers_per_thread", "regs", 1)]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, check=True).stdout.decode()
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in rows[2:]:
        d = {"kernel": r[col["Kernel Name"]].split("(")[0]}
        for m, name, scale in WANT:
            if m in col and r[col[m]] not in ("", "n/a"):
                v = float(r[col[m]].replace(",", ""))
                u = units[col[m]]
                if m.startswith("gpu__time") and u in ("us", "usecond"):
                    v *= 1e3
                elif m.startswith("gpu__time") and u in ("ms", "msecond"):
                    v *= 1e6
                elif m.startswith("gpu__time") and u in ("s", "second"):
                    v *= 1e9
                elif m.startswith("dram__bytes"):
                    v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
                d[name] = v * scale
        out.append(d)
    names = ["kernel"] + [n for _m, n, _s in WANT]
    print("| " + " | ".join(names) + " |")
    print("|" + "---|" * len(names))
    for d in out:
        print("| " + " | ".join(("%.3f" % d[n] if isinstance(d.get(n), float) else str(d.get(n, ""))) for n in names) + " |")


if __name__ == "__main__":
    main()
