"""Development aid: per-warp cycle breakdown of the phase kernel (library built with -DPH_PROF).
usage: OF_B200_LIB=.../libof_prof.so OF_B200_VARIANT=phase python tools/phase_prof.py"""
import ctypes, os, sys, importlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from of_b200_loader import ofb
synth = importlib.import_module("optical_flow_b200.synth")
h, w, L, r, n = 1080, 1920, 3, 7, 16
fr = torch.from_numpy(synth.synth_frames(n + 1, h, w, seed=1234)).cuda()
U = torch.empty((n, h, w), dtype=torch.float64, device="cuda"); V = torch.empty_like(U)
lib = ctypes.CDLL(os.environ["OF_B200_LIB"])
buf = (ctypes.c_ulonglong * 256)()
with ofb.Context(0, win_radius=r) as ctx:
    for _ in range(3): ctx.flow_batch_dev(fr, L, U, V)
    torch.cuda.synchronize()
    lib.of_b200_debug_phase_prof(buf, 1)
    reps = 4
    for _ in range(reps): ctx.flow_batch_dev(fr, L, U, V)
    torch.cuda.synchronize()
    lib.of_b200_debug_phase_prof(buf, 0)
names = ["beta", "bar0", "P3", "P1", "bar12", "P4", "bar_end", "misc"]
print("warp " + " ".join("%8s" % x for x in names) + "    total   (cycles per step, level 0, CTA (1,0,0))")
steps = 74 * reps
for wp in range(32):
    v = [buf[wp * 8 + i] / steps for i in range(8)]
    if sum(v) == 0: continue
    print("%4d " % wp + " ".join("%8.0f" % x for x in v) + " %8.0f" % sum(v))
if hasattr(lib, "of_b200_debug_p1_prof"):
    # the counters were reset with the others only if reset was called on them too: report totals / all calls
    lib.of_b200_debug_p1_prof(buf, 0)
    names = ["setup", "vert", "park", "tapsA", "blendA", "tapsB", "blendB", "store"]
    print("P1 breakdown (cycles per step, includes the warm-up calls: scaled by %d/%d)" % (reps, reps + 3))
    print("warp " + " ".join("%8s" % x for x in names) + "    total")
    for wp in range(32):
        v = [buf[wp * 8 + i] / (74 * (reps + 3)) for i in range(8)]
        if sum(v) == 0: continue
        print("%4d " % wp + " ".join("%8.0f" % x for x in v) + " %8.0f" % sum(v))
