"""End-to-end throughput of the single-process multi-GPU entry points (of_mgpu_ctf_*_stream: one host thread
and one context per GPU inside the C library) on 1, 2, 4, ... GPUs of this box: pinned host frames in; full
f64 flow fields out, statistics only, and 8-bit frames in / statistics out.  One JSON object on stdout.

    python tools/mgpu_e2e.py [--config 1080p|4k]"""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from of_b200_loader import ofb

synth = importlib.import_module("optical_flow_b200.synth")
cfg = sys.argv[sys.argv.index("--config") + 1] if "--config" in sys.argv else "1080p"
W, H, L, PER = (1920, 1080, 3, 16) if cfg == "1080p" else (3840, 2160, 4, 8)
n_dev = torch.cuda.device_count()
base = synth.synth_frames(PER + 1, H, W, seed=1234)
out = {"config": cfg, "pairs_per_gpu_per_call": PER, "win_radius": 7, "results": []}
N = 1
while N <= n_dev:
    n_pairs = PER * N
    idx = [k % (2 * PER) if k % (2 * PER) <= PER else 2 * PER - k % (2 * PER) for k in range(n_pairs + 1)]   # forth and back
    fin = torch.empty((n_pairs + 1, H, W), dtype=torch.float64).pin_memory()
    fin.copy_(torch.from_numpy(base[idx]))
    hu = torch.empty((n_pairs, H, W), dtype=torch.float64).pin_memory()
    hv = torch.empty((n_pairs, H, W), dtype=torch.float64).pin_memory()
    q8 = torch.empty((n_pairs + 1, H, W), dtype=torch.uint8).pin_memory()
    q8.copy_(torch.from_numpy(np.clip(np.rint(base[idx] * 255.0), 0, 255).astype(np.uint8)))
    row = {"n_gpus": N}
    with ofb.MultiGpu(list(range(N)), win_radius=7) as m:
        for name, kw in (("f64_in_flow_out", dict(frames=fin.numpy(), out_u=hu.numpy(), out_v=hv.numpy())),
                         ("f64_in_stats_out", dict(frames=fin.numpy(), want_flow=False)),
                         ("u8_in_stats_out", dict(frames=q8.numpy(), want_flow=False))):
            f = kw.pop("frames")
            m.flow_stream(f, L, **kw)
            reps = 3
            t0 = time.perf_counter()
            for _ in range(reps):
                m.flow_stream(f, L, **kw)
            row[name] = n_pairs * reps / (time.perf_counter() - t0)
    out["results"].append(row)
    del fin, hu, hv, q8
    N *= 2
print(json.dumps(out))
