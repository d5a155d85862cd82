#!/usr/bin/env python
"""Splits the SASS of a profiled kernel at its BAR/EXIT instructions (= the phases of the
phase-synchronous kernel) and prints executed warp instructions, stall samples and the top
opcodes per segment.  usage: ncu_segments.py report.ncu-rep [hot]   (hot: dump SASS of segment N)"""
import csv, io, subprocess, sys
rep = sys.argv[1]
dump = int(sys.argv[2]) if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]
ie = h.index('Instructions Executed'); ss = h.index('# Samples'); isrc = h.index('Source')
stall_cols = [(i, x) for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
segs = []; cur = {'n': 0, 's': 0, 'cnt': 0, 'ops': {}, 'st': {}, 'txt': []}
for r in rows[2:]:
    try: n = int(r[ie]); s = int(r[ss] or 0)
    except (ValueError, IndexError): continue
    txt = r[isrc]; t = txt.split(); op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    cur['n'] += n; cur['s'] += s; cur['cnt'] += 1
    o = cur['ops'].setdefault(op, [0, 0]); o[0] += n; o[1] += s
    for i, x in stall_cols:
        try: cur['st'][x] = cur['st'].get(x, 0) + int(r[i] or 0)
        except ValueError: pass
    cur['txt'].append((n, s, txt))
    if op in ('BAR', 'EXIT'):
        segs.append(cur); cur = {'n': 0, 's': 0, 'cnt': 0, 'ops': {}, 'st': {}, 'txt': []}
segs.append(cur)
T = sum(x['n'] for x in segs); S = sum(x['s'] for x in segs)
print('total warp instr', T, 'samples', S)
for i, x in enumerate(segs):
    if x['n'] < T / 1000: continue
    print('seg %d static %d  exec %.1fM (%.1f%%)  samples %d (%.1f%%)' % (i, x['cnt'], x['n'] / 1e6, 100 * x['n'] / T, x['s'], 100 * x['s'] / S))
    print('    ' + '  '.join('%s %.1fM/%d' % (k, v[0] / 1e6, v[1]) for k, v in sorted(x['ops'].items(), key=lambda kv: -kv[1][0])[:14]))
    print('    stalls: ' + '  '.join('%s %d' % (k[6:], v) for k, v in sorted(x['st'].items(), key=lambda kv: -kv[1])[:8]))
if dump is not None:
    mx = max(n for n, s, t in segs[dump]['txt'])
    for n, s, t in segs[dump]['txt']:
        if n > 0.3 * mx: print('%5.2f %5d  %s' % (n / mx, s, t[:100]))
