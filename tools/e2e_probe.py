"""End-to-end (host-pointer) stream call: time per 16-pair call against the chunk size."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from of_b200_loader import ofb
import importlib
synth = importlib.import_module("optical_flow_b200.synth")
H, W, P = 1080, 1920, 16
fr = synth.synth_frames(P + 1, H, W, seed=1234)
hin = torch.empty((P + 1, H, W), dtype=torch.float64).pin_memory(); hin.copy_(torch.from_numpy(fr))
hu = torch.empty((P, H, W), dtype=torch.float64).pin_memory()
hv = torch.empty((P, H, W), dtype=torch.float64).pin_memory()
fin, u, v = hin.numpy(), hu.numpy(), hv.numpy()
for chunk in (16, 8, 4, 2, 1):
    os.environ["OF_B200_STREAM_CHUNK"] = str(chunk)
    with ofb.Context(0, win_radius=7) as ctx:
        for flow in (True, False):
            kw = dict(out_u=u, out_v=v) if flow else dict(want_flow=False)
            ctx.flow_stream(fin, 3, **kw)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                ctx.flow_stream(fin, 3, **kw)
            dt = (time.perf_counter() - t0) / 5
            print("chunk %2d flow=%d: %.2f ms per call, %.0f pairs/s" % (chunk, flow, dt * 1e3, P / dt), flush=True)
