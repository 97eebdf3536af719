#!/usr/bin/env python3
"""Text summary of an .ncu-rep (run here, no GPU needed): headline metrics per captured launch, stall-reason and
opcode histograms from the source page.   python tools/ncu_summary.py <rep> [--json out.json] [--top N]"""
import csv
import json
import subprocess
import sys
from collections import Counter

WANT = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__thread_inst_executed_per_inst_executed.ratio"]


def ncu(*args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True).stdout


def to_bytes(v, unit):
    m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(v) * m.get(unit, 1)


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 12
    rows = list(csv.reader(ncu("-i", rep, "--page", "raw", "--csv").splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print("== %s: %d captured launch(es)" % (rep, len(data)))
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print("%-62s %-16s %s" % (w, units[i], [r[i] for r in data]))
    if "--json" in sys.argv:
        ir, iw, it = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
        rd = [to_bytes(r[ir], units[ir]) for r in data]
        wr = [to_bytes(r[iw], units[iw]) for r in data]
        out = {"kernel": data[0][hdr.index("Kernel Name")], "launches": len(data), "dram_bytes_read_per_launch": sum(rd) / len(rd),
               "dram_bytes_write_per_launch": sum(wr) / len(wr), "dram_bytes_per_launch": (sum(rd) + sum(wr)) / len(rd),
               "ncu_duration_us": [float(r[it]) for r in data], "source": rep}
        json.dump(out, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
    rows = list(csv.reader(ncu("-i", rep, "--page", "source", "--csv", "--print-source", "sass").splitlines()))
    k, out, h = 0, [], None
    for r in rows:
        if r and r[0] == "Address":
            h = r; k += 1; continue
        if r and r[0] == "Kernel Name":
            continue
        if h and k == 1 and len(r) == len(h):
            out.append(r)
    if not out:
        return
    iS, iI, iSrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source")
    tot_s, tot_i = sum(int(r[iS]) for r in out), sum(int(r[iI]) for r in out)
    print("-- first launch: %d SASS instructions, %d warp-instructions executed, %d stall samples" % (len(out), tot_i, tot_s))
    stalls = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
    agg = sorted(((s, sum(int(r[h.index(s)]) for r in out)) for s in stalls), key=lambda x: -x[1])[:8]
    print("-- stall samples:", ", ".join("%s %.1f%%" % (s[6:], 100.0 * n / max(1, tot_s)) for s, n in agg))
    c, cs = Counter(), Counter()
    for r in out:
        t = r[iSrc].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        c[op] += int(r[iI]); cs[op] += int(r[iS])
    print("-- opcode mix (executed %, stall-sample %):")
    for op, n in c.most_common(top):
        print("   %-8s %5.1f%%  %5.1f%%" % (op, 100.0 * n / tot_i, 100.0 * cs[op] / max(1, tot_s)))
    tma = sum(n for op, n in c.items() if op in ("UBLKCP", "UTMALDG", "UTMASTG"))
    print("-- TMA instructions executed (UBLKCP/UTMALDG/UTMASTG): %d" % tma)


if __name__ == "__main__":
    main()
