"""Quick device-resident timing of the batch entry point (development aid, not the bench)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from of_b200_loader import ofb
import importlib
synth = importlib.import_module("optical_flow_b200.synth")

def run(h, w, L, r, npairs, iters=5):
    fr = synth.synth_frames(npairs + 1, h, w, seed=1234)
    d = torch.from_numpy(fr).cuda()
    U = torch.empty((npairs, h, w), dtype=torch.float64, device="cuda"); V = torch.empty_like(U)
    with ofb.Context(0, win_radius=r) as ctx:
        for _ in range(2): ctx.flow_batch_dev(d, L, U, V)
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): ctx.flow_batch_dev(d, L, U, V)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
    alg = 0
    N = [h * w // 4 ** l for l in range(L)]
    alg = 16 * (sum(N[:-1]) + sum(N[1:])) + sum(32 * N[l] + (16 * N[l + 1] if l < L - 1 else 0) for l in range(L))
    pps = npairs / (ms * 1e-3)
    print("%dx%d L=%d r=%d pairs=%d: %.3f ms/batch  %.1f pairs/s  %.1f us/pair  roofline %.3f (variant %s)" % (
        w, h, L, r, npairs, ms, pps, 1e3 * ms / npairs, alg * pps / 6539.5e9, os.environ.get("OF_B200_VARIANT", "auto")))

if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    for r in (1, 7):
        run(1080, 1920, 3, r, 16)
    run(1080, 1080, 4, 1, 16)
    run(2160, 3840, 4, 7, 4)
