#!/usr/bin/env python
"""Derives the two ncu-only constants bench.py reports next to its live measurements from `ncu --set full`
reports of the level-0 kernels (one report per window radius):

    python tools/ncu_constants.py --config 1080p --pairs 16 r7=gpurun_out/x_r7.ncu-rep r1=gpurun_out/x_r1.ncu-rep

  profiles/traffic.json   dram__bytes_read.sum + dram__bytes_write.sum of one level-0 launch  (roofline.traffic)
  profiles/fp64_ops.json  FP64 instructions (DADD + DMUL + DFMA, per thread) per level-0 pixel  (roofline.fp64)
"""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SIZES = {"1080p": 1920 * 1080, "4k": 3840 * 2160}


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    cfg = sys.argv[sys.argv.index("--config") + 1]
    pairs = int(sys.argv[sys.argv.index("--pairs") + 1])
    traffic = {"config": cfg, "pairs": pairs}
    ops = {}
    src = []
    for a in sys.argv[1:]:
        if "=" not in a:
            continue
        key, rep = a.split("=", 1)
        raw = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
        hdr, units, vals = raw[0], raw[1], raw[2]
        d = {h: (u, v) for h, u, v in zip(hdr, units, vals)}

        def mb(name):
            u, v = d[name]
            return float(v) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[u]
        traffic["level0_%s_bytes_per_launch" % key] = int(mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum"))
        rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "sass"]))))
        h = rows[1]
        ie, isrc = h.index("Instructions Executed"), h.index("Source")
        n = 0
        for r in rows[2:]:
            try:
                t = r[isrc].split()
                op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
                if op in ("DADD", "DMUL", "DFMA"):
                    n += int(r[ie])
            except (ValueError, IndexError):
                pass
        ops[key] = round(n * 32.0 / (SIZES[cfg] * pairs), 1)
        src.append("%s: %s (%s)" % (key, os.path.basename(rep), d["Kernel Name"][1][:60]))
    traffic["source"] = "ncu --set full --clock-control none, one level-0 launch of a %d-pair %s batch; %s" % (pairs, cfg, "; ".join(src))
    ops["source"] = "ncu source page, DADD + DMUL + DFMA warp instructions x 32 / (pixels x pairs); " + "; ".join(src)
    json.dump(traffic, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    json.dump(ops, open(os.path.join(ROOT, "profiles", "fp64_ops.json"), "w"), indent=1)
    print(json.dumps(traffic, indent=1)); print(json.dumps(ops, indent=1))


if __name__ == "__main__":
    main()
