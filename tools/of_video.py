#!/usr/bin/env python
"""Headless replacement of the reference's live loop (ParallelVideoPyr.cpp:769-905) without camera
and window: a frame stream -> coarse-to-fine flow of every consecutive pair -> per-pair gesture
statistics (centroid of the pixels whose flow magnitude is >= 20, count, "detected" iff count >
500, :845-888) as JSON lines; optionally the flow fields as .npy.

  python tools/of_video.py frames.npy|frames.npz|synthetic[:N[:HxW]] [--levels L] [--radius R]
                           [--chunk PAIRS] [--flow-out stem] [--mag-thresh 20] [--count-thresh 500]

frames.npy: (n, H, W) float64 gray frames (already scaled as the caller wants them: the live loop
feeds unnormalised Sobel sums, the harnesses gray / 255).  frames.npz: first array inside;
uint8 frames are divided by 255.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from of_b200_loader import ofb  # noqa: E402
import importlib  # noqa: E402


def load_frames(spec):
    if spec.startswith("synthetic"):
        parts = spec.split(":")
        n = int(parts[1]) if len(parts) > 1 else 17
        h, w = (int(x) for x in parts[2].split("x")) if len(parts) > 2 else (1080, 1920)
        synth = importlib.import_module("optical_flow_b200.synth")
        return synth.synth_frames(n, h, w, seed=1234, gesture=True)
    d = np.load(spec)
    if hasattr(d, "files"):
        d = d[d.files[0]]
    if d.dtype == np.uint8:
        d = d.astype(np.float64) / 255.0
    return np.ascontiguousarray(d, dtype=np.float64)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("frames")
    ap.add_argument("--levels", type=int, default=0, help="0 = getMaxLayer of the frame size")
    ap.add_argument("--radius", type=int, default=1)
    ap.add_argument("--chunk", type=int, default=64, help="pairs per library call")
    ap.add_argument("--flow-out", help="also save the flow fields: <stem>_u.npy, <stem>_v.npy")
    ap.add_argument("--mag-thresh", type=float, default=20.0)
    ap.add_argument("--count-thresh", type=int, default=500)
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args(argv)

    fr = load_frames(a.frames)
    n, h, w = fr.shape
    levels = a.levels or ofb.getMaxLayer(w, h)
    us, vs = [], []
    t0 = time.perf_counter()
    with ofb.Context(a.device, win_radius=a.radius, mag_thresh=a.mag_thresh, count_thresh=a.count_thresh) as ctx:
        for p0 in range(0, n - 1, a.chunk):
            p1 = min(p0 + a.chunk, n - 1)
            u, v, st = ctx.flow_stream(fr[p0:p1 + 1], levels, want_flow=bool(a.flow_out))
            for k in range(p1 - p0):
                s = st[k]
                c = float(s["count"])
                # centroid like :884-885 (sum / count); None when no pixel passed the threshold
                print(json.dumps({"pair": p0 + k, "count": int(c), "cx": float(s["sum_x"]) / c if c else None,
                                  "cy": float(s["sum_y"]) / c if c else None, "detected": bool(s["detected"])}))
            if a.flow_out:
                us.append(u)
                vs.append(v)
    dt = time.perf_counter() - t0
    if a.flow_out:
        np.save(a.flow_out + "_u.npy", np.concatenate(us))
        np.save(a.flow_out + "_v.npy", np.concatenate(vs))
    print(json.dumps({"pairs": n - 1, "width": w, "height": h, "levels": levels, "radius": a.radius,
                      "seconds": dt, "pairs_per_s": (n - 1) / dt}), file=sys.stderr)
    return 0


if __name__ == "__main__":
    sys.exit(main())
