#!/usr/bin/env python3
"""Per-kernel summary of an ncu --set full capture: duration, DRAM bytes, L1 / L2 hit rates, issue activity, active threads
per instruction, local vs global sectors.   usage: tools/ncu_summary.py report.ncu-rep [--traffic-json bench.json out.json]
--traffic-json also writes the DRAM bytes of the classify stage's launches (k_islands + k_seed + k_classify of one step) in the
form bench.py's roofline.traffic reads, with the workload size taken from a bench line of the same workload."""
import json
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
    if "--traffic-json" in sys.argv:
        i = sys.argv.index("--traffic-json")
        traffic_json(out, sys.argv[i + 1], sys.argv[i + 2], rep)
    names = ["kernel"] + [n for _m, n, _s in WANT]
    print("| " + " | ".join(names) + " |")
    print("|" + "---|" * len(names))
    for d in out:
        print("| " + " | ".join(("%.3f" % d[n] if isinstance(d.get(n), float) else str(d.get(n, ""))) for n in names) + " |")


def traffic_json(out, bench_json, dst, rep):
    line = json.loads(open(bench_json).read().strip().splitlines()[-1])
    per = {}
    for d in out:
        if d["kernel"] in ("k_islands", "k_seed", "k_classify") and d["kernel"] not in per:
            per[d["kernel"]] = {"read": d["dram_rd_GB"] * 1e9, "write": d["dram_wr_GB"] * 1e9, "ms_under_ncu": d["ms"]}
    t = {"kernels": "k_islands + k_seed + k_classify (one step)", "workload": line["config"]["workload"][:60],
         "n_reads": line["config"]["n_reads_per_gpu"], "n_bases": line["config"]["n_bases_per_gpu"],
         "dram_bytes_read": sum(v["read"] for v in per.values()), "dram_bytes_write": sum(v["write"] for v in per.values()),
         "per_kernel": per, "source": "ncu --set full --clock-control none, one launch of each kernel of one step (%s)" % rep}
    json.dump(t, open(dst, "w"), indent=1)


if __name__ == "__main__":
    main()
