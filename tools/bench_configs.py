#!/usr/bin/env python
"""Throughput of every configuration BASELINE.json lists (bench.py itself times configs[3] only),
device-resident through of_ctf_f64_batch_dev, with the C + OpenMP port of the reference timed beside
each one on a bounded sample.  One JSON object on stdout; run on a B200:

    python tools/bench_configs.py > profiles/<round>_configs.json

Configs: hand pair (3264x2448, L = getMaxLayer = 5, 3x3), table pair KAT-1 / KAT-7 (640x448),
pacman.mp4 (194 pairs, 240x160, L = 5, 3x3), synthetic 1080p L = 3 (15x15 and 3x3), the live loop's
1080x1080 L = 4 3x3 (ParallelVideoPyr.cpp:782,831), synthetic 4K L = 4 (15x15 and 3x3).
"""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["OMP_NUM_THREADS"] = str(len(os.sched_getaffinity(0)))

import torch  # noqa: E402
from of_b200_loader import ofb  # noqa: E402
from oracle import c_oracle as CO  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
synth = importlib.import_module("optical_flow_b200.synth")
PEAK = 6549.1
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def alg_bytes(w, h, L):
    n = [(w >> l) * (h >> l) for l in range(L)]
    return 16 * (sum(n[:-1]) + sum(n[1:])) + sum(32 * n[l] + (16 * n[l + 1] if l < L - 1 else 0) for l in range(L))


def gray_png(name):
    import cv2
    return cv2.imread(os.path.join(GOLDEN, name), cv2.IMREAD_UNCHANGED).astype(np.float64) / 255.0


def run(name, frames, L, r, pairs_mode, cpu_pairs, reps):
    """frames: (n, H, W) float64.  pairs_mode 'stream' = consecutive pairs, 'split' = (k, k + n/2)."""
    h, w = frames.shape[1:]
    d = torch.from_numpy(np.ascontiguousarray(frames)).cuda()
    n_pairs = len(frames) - 1 if pairs_mode == "stream" else len(frames) // 2
    U = torch.empty((n_pairs, h, w), dtype=torch.float64, device="cuda")
    V = torch.empty_like(U)
    kw = {} if pairs_mode == "stream" else {"pairs": "split"}
    with ofb.Context(0, win_radius=r) as ctx:
        for _ in range(3):
            ctx.flow_batch_dev(d, L, U, V, **kw)
        torch.cuda.synchronize()
        l0 = ctx.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            ctx.flow_batch_dev(d, L, U, V, **kw)
        e1.record()
        torch.cuda.synchronize()
        launches = (ctx.launch_count - l0) // reps
    ms = e0.elapsed_time(e1) / reps
    gpu = n_pairs / ms * 1e3
    # CPU port on the first cpu_pairs pairs, all host threads
    cores = CO.set_threads(len(os.sched_getaffinity(0)))
    t0 = time.perf_counter()
    for k in range(cpu_pairs):
        a, b = (frames[k], frames[k + 1]) if pairs_mode == "stream" else (frames[k], frames[k + len(frames) // 2])
        uo, vo = CO.coarse_to_fine(a, b, L, r)
    cpu = cpu_pairs / (time.perf_counter() - t0)
    err = max(np.abs(U[cpu_pairs - 1].cpu().numpy() - uo).max(), np.abs(V[cpu_pairs - 1].cpu().numpy() - vo).max())
    return {"config": name, "width": w, "height": h, "levels": L, "win_radius": r, "pairs_per_call": n_pairs,
            "gpu_pairs_per_s": gpu, "ms_per_call": ms, "kernel_launches_per_call": int(launches),
            "whole_path_roofline_frac": alg_bytes(w, h, L) * gpu / (PEAK * 1e9),
            "cpu_port_pairs_per_s": cpu, "cpu_threads": cores, "cpu_sample_pairs": cpu_pairs,
            "max_abs_diff_vs_port_last_sampled_pair": float(err)}


def latency(name, frames, L, r, reps=200):
    """One pair per call (what the reference's live loop does, ParallelVideoPyr.cpp:831-832), device-resident."""
    h, w = frames.shape[1:]
    d = torch.from_numpy(np.ascontiguousarray(frames[:2])).cuda()
    u = torch.empty((h, w), dtype=torch.float64, device="cuda")
    v = torch.empty_like(u)
    with ofb.Context(0, win_radius=r) as ctx:
        for _ in range(5):
            ctx.coarseToFineEstimation_dev(d[0], d[1], L, u, v)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            ctx.coarseToFineEstimation_dev(d[0], d[1], L, u, v)
        e1.record()
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    return {"config": name, "width": w, "height": h, "levels": L, "win_radius": r, "pairs_per_call": 1,
            "us_per_call_back_to_back": us, "gpu_pairs_per_s": 1e6 / us}


def main():
    out = []
    hand = np.stack([gray_png("hand1_gray.png"), gray_png("hand2_gray.png")])
    out.append(run("hand1 -> hand2 (configs[0])", hand, ofb.getMaxLayer(hand[0]), 1, "split", 1, 20))
    t = np.load(os.path.join(GOLDEN, "table_pair.npz"))
    table = np.stack([t["img1"], t["img2"]]).astype(np.float64) / 255.0
    out.append(run("table1 -> table2, single level KAT-1 (configs[1])", table, 1, 1, "split", 1, 50))
    out.append(run("table1 -> table2, 7 levels KAT-7 (configs[1])", table, 7, 1, "split", 1, 50))
    pac = np.load(os.path.join(GOLDEN, "pacman_gray.npz"))["frames"].astype(np.float64) / 255.0
    out.append(run("pacman.mp4, all 194 consecutive pairs (configs[2])", pac, 5, 1, "stream", 16, 10))
    f = synth.synth_frames(17, 1080, 1920, seed=1234)
    out.append(run("synthetic 1080p, 15x15 (configs[3])", f, 3, 7, "stream", 4, 20))
    out.append(run("synthetic 1080p, 3x3 (configs[3], the reference's window)", f, 3, 1, "stream", 4, 20))
    f = synth.synth_frames(17, 1080, 1920, seed=1234, gesture=True)
    out.append(run("synthetic 1080p gesture-like stream (moving disc, exact zeros elsewhere), 15x15", f, 3, 7, "stream", 4, 20))
    out.append(run("synthetic 1080p gesture-like stream, 3x3", f, 3, 1, "stream", 4, 20))
    f = synth.synth_frames(17, 1080, 1080, seed=1234)
    out.append(run("synthetic 1080x1080 L=4 3x3 (the live loop's size, ParallelVideoPyr.cpp:782)", f, 4, 1, "stream", 4, 20))
    f = synth.synth_frames(9, 2160, 3840, seed=1234)
    out.append(run("synthetic 4K L=4, 15x15 (configs[4], one GPU's share)", f, 4, 7, "stream", 2, 10))
    out.append(run("synthetic 4K L=4, 3x3 (configs[4])", f, 4, 1, "stream", 2, 10))
    f = synth.synth_frames(2, 1080, 1080, seed=1234)
    out.append(latency("one pair per call, 1080x1080 L=4 3x3 (the live loop's call)", f, 4, 1))
    f = synth.synth_frames(2, 1080, 1920, seed=1234)
    out.append(latency("one pair per call, 1080p L=3 15x15", f, 3, 7))
    json.dump({"hbm_peak_gbs": PEAK, "gpu": torch.cuda.get_device_name(0), "results": out}, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
