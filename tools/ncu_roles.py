#!/usr/bin/env python
"""Per-role instruction / stall-sample breakdown of the strip kernel from an ncu report.
Roles are contiguous SASS address regions; an instruction is classified by a majority vote of the
role-specific source lines in a window around it."""
import csv, io, subprocess, sys, re
rep = sys.argv[1]
srcfile = sys.argv[2] if len(sys.argv) > 2 else "lk_level_strip.cu"
src = open('/root/repo/optical-flow_b200/csrc/' + srcfile).read().split('\n')
def find(pat):
    hits = [i + 1 for i, l in enumerate(src) if re.search(pat, l)]
    return hits[0] if hits else 10 ** 9
if len(sys.argv) > 3:      # role = function name pattern
    b = [(p, find(r"void " + p)) for p in sys.argv[3:]]
    ROLES = tuple(sys.argv[3:])
else:
    b = [("producer", find(r"void producer_role")), ("helpers", find(r"// -+ consumer")),
         ("consumer", min(find(r"void consumer_role"), find(r"void consumer_xx_role"))), ("epilogue", min(find(r"void epilogue_role"), find(r"void consumer_xt_role"))), ("kernel", find(r"lk_level_strip_kernel\("))]
    ROLES = ("producer", "consumer", "epilogue")
b.sort(key=lambda x: x[1])
def role_of(line):
    r = None
    for name, lo in b:
        if line >= lo and lo < 10 ** 9: r = name
    return r
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
h = rows[hi]; ie = h.index("Instructions Executed"); ss = h.index("# Samples")
stall_cols = [(i, x) for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
cur = ""; tag = None
ins = []   # (addr, tag, op, n, s)
for r in rows[hi + 1:]:
    if r and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) <= ie: continue
    if r[0] == "Line No": continue
    if r[0] != "":
        tag = None
        if cur == "" or cur.endswith(srcfile):
            rr = role_of(int(r[0]))
            if rr in ROLES: tag = rr
        continue
    t = r[3].split()
    op = (t[1] if t and t[0].startswith("@") else t[0]).split(".")[0] if t else "?"
    try: n, s = int(r[ie]), int(r[ss] or 0)
    except ValueError: continue
    st = {}
    for i, x in stall_cols:
        try: st[x] = int(r[i] or 0)
        except ValueError: st[x] = 0
    ins.append([int(r[2], 16), tag, op, n, s, st])
ins.sort(key=lambda x: x[0])
Wd = int(__import__('os').environ.get('ROLE_WIN', '60'))
agg = {}
for i, x in enumerate(ins):
    votes = {}
    for y in ins[max(0, i - Wd): i + Wd]:
        if y[1]: votes[y[1]] = votes.get(y[1], 0) + 1
    role = max(votes, key=votes.get) if votes else "?"
    a = agg.setdefault(role, {"n": 0, "s": 0, "ops": {}, "count": 0, "st": {}})
    for k2, v2 in x[5].items(): a["st"][k2] = a["st"].get(k2, 0) + v2
    a["n"] += x[3]; a["s"] += x[4]; a["count"] += 1
    o = a["ops"].setdefault(x[2], [0, 0]); o[0] += x[3]; o[1] += x[4]
tot = sum(a["n"] for a in agg.values())
for k, a in agg.items():
    fp = sum(a["ops"].get(x, [0, 0])[0] for x in ("DADD", "DMUL", "DFMA"))
    print("== %-9s static %5d  executed %12d (%4.1f%%)  fp64 %11d (%4.1f%% of role)  samples %7d" % (k, a["count"], a["n"], 100.0 * a["n"] / tot, fp, 100.0 * fp / max(a["n"], 1), a["s"]))
    print("     " + "  ".join("%s %.1fM/%d" % (op, v[0] / 1e6, v[1]) for op, v in sorted(a["ops"].items(), key=lambda kv: -kv[1][0])[:16]))
    print("     stalls: " + "  ".join("%s %d" % (k2[6:], v2) for k2, v2 in sorted(a["st"].items(), key=lambda kv: -kv[1])[:9]))
