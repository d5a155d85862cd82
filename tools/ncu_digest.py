#!/usr/bin/env python
"""Digest of an .ncu-rep for one kernel: headline metrics, stall reasons, opcode mix, hottest source
lines.  Usage: python tools/ncu_digest.py report.ncu-rep [--lines N] [--regions file:lo-hi:name,...]"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "launch__registers_per_thread ",
        "launch__occupancy_limit", "launch__grid_size", "launch__waves", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct",
        "smsp__inst_executed.sum ", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct", "l1tex__throughput.avg.pct",
        "lts__throughput.avg.pct", "gpu__dram_throughput.avg.pct", "sm__inst_executed_pipe_lsu", "smsp__inst_executed_pipe_fp64",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum ", "smsp__average_warps_issue_stalled"]


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    nlines = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 25
    raw = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units, vals = raw[0], raw[1], raw[2]
    print("== kernel:", vals[hdr.index("Kernel Name")][:100])
    for h, u, v in zip(hdr, units, vals):
        if any(k in h + " " for k in KEYS):
            if "stalled" in h and float(v or 0) < 0.05:
                continue
            print("%-90s %-14s %s" % (h, u, v))
    rows = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]))))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
    h = rows[hi]
    ie, ss = h.index("Instructions Executed"), h.index("# Samples")
    cur = ""
    lines, ops, tot, tots = [], {}, 0, 0
    for r in rows[hi + 1:]:
        if r and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if len(r) <= ie:
            continue
        try:
            n, s = int(r[ie]), int(r[ss] or 0)
        except ValueError:
            continue
        if r[0] != "":
            lines.append((cur, int(r[0]), r[1].strip()[:100], n, s))
            tot += n
            tots += s
        else:
            t = r[3].split()
            op = (t[1] if t and t[0].startswith("@") else t[0]).split(".")[0] if t else "?"
            o = ops.setdefault(op, [0, 0])
            o[0] += n
            o[1] += s
    print("== total warp-instructions (line-attributed) %d, samples %d" % (tot, tots))
    print("== opcode mix (warp-inst, %, stall samples)")
    ot = sum(v[0] for v in ops.values())
    for k, v in sorted(ops.items(), key=lambda kv: -kv[1][0])[:28]:
        print("  %-10s %12d %5.1f%% %8d" % (k, v[0], 100.0 * v[0] / max(ot, 1), v[1]))
    print("== hottest lines by instructions")
    for l in sorted(lines, key=lambda a: -a[3])[:nlines]:
        print("  %-18s %4d %11d %5.1f%% smp %6d | %s" % (l[0], l[1], l[3], 100.0 * l[3] / max(tot, 1), l[4], l[2]))
    print("== hottest lines by stall samples")
    for l in sorted(lines, key=lambda a: -a[4])[:nlines]:
        print("  %-18s %4d %11d smp %6d %5.1f%% | %s" % (l[0], l[1], l[3], l[4], 100.0 * l[4] / max(tots, 1), l[2]))


if __name__ == "__main__":
    main()
