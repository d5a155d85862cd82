"""A/B timing of library variants (development aid): python tools/ab_time.py libA.so libB.so ...
Each variant runs in its own process (OF_B200_LIB), twice, 1080p L=3 r=7, 16 pairs."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys
sys.path.insert(0, %r)
import torch, importlib
from of_b200_loader import ofb
synth = importlib.import_module("optical_flow_b200.synth")
h, w, L, r, n = 1080, 1920, 3, int(os.environ.get("AB_R", "7")), 16
fr = torch.from_numpy(synth.synth_frames(n + 1, h, w, seed=1234)).cuda()
U = torch.empty((n, h, w), dtype=torch.float64, device="cuda"); V = torch.empty_like(U)
S = torch.zeros((n, 4), dtype=torch.float64, device="cuda") if os.environ.get("AB_STATS", "1") == "1" else None
with ofb.Context(0, win_radius=r) as ctx:
    for _ in range(3): ctx.flow_batch_dev(fr, L, U, V, S)
    torch.cuda.synchronize()
    ctx.set_timing(True)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): ctx.flow_batch_dev(fr, L, U, V, S)
    e1.record(); torch.cuda.synchronize()
    t = ctx.get_timing()
    l0 = [ms for k, l, ms in t if k == 1 and l == 0]; l1 = [ms for k, l, ms in t if k == 1 and l == 1]; l2 = [ms for k, l, ms in t if k == 1 and l == 2]
    ms = e0.elapsed_time(e1) / 10
    other = {}
    for k, l, m_ in t:
        if k != 1: other[(k, l)] = other.get((k, l), 0.0) + m_ / 10
    print("   other kernels (kind, level): ms/batch", {k: round(v, 4) for k, v in sorted(other.items())})
    print("%%-28s %%.3f ms/batch  %%8.1f pairs/s   L0 %%.3f  L1 %%.3f  L2 %%.3f  sum(u) %%.6f" %% (os.path.basename(os.environ.get("OF_B200_LIB", "default")), ms, n / ms * 1e3,
          sum(l0) / len(l0), sum(l1) / len(l1), sum(l2) / len(l2), float(U.sum())))
''' % ROOT
for rep in range(int(os.environ.get("AB_REPS", "2"))):
    for lib in sys.argv[1:]:
        env = dict(os.environ)
        env["OF_B200_LIB"] = os.path.join(ROOT, "optical-flow_b200", lib)
        subprocess.run([sys.executable, "-c", CHILD], env=env)
