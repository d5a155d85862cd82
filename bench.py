#!/usr/bin/env python
"""Benchmark of the dense pyramidal Lucas-Kanade path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # CPU baseline arm
    python bench.py --config 4k ...                          # BASELINE.json configs[4] (3840x2160, L = 4)

Workload (BASELINE.json configs[3], the configuration the metric is quoted on): synthetic
1920x1080 grayscale video, 3-level pyramid, 15x15 window (r = 7), FP64.  One "step" = one batch
of PAIRS consecutive frame pairs (PAIRS+1 frames) through of_ctf_f64_batch_dev with inputs and
outputs resident in HBM.  Each step touches (PAIRS+1)*16.6 MB of frames and PAIRS*33.2 MB of
flow, far more than the 126 MB L2, so no explicit L2 flush is needed between steps.

value  = frame pairs per second, whole job (all ranks), device-resident, CUDA-event timed,
         max over ranks.
e2e    = the same metric through the host-pointer stream call (of_ctf_f64_stream): frames in
         pinned host memory, flow fields read back to pinned host memory, copies inside the
         timed region.
roofline = the level-0 fused kernel (dominant): algorithmic bytes per launch (36 B per level-0
         pixel per pair: read A,B 16 + write u,v 16 + coarse flow 4) / its average launch
         duration measured with CUDA events on the launching stream inside the timed region.
cpu_baseline = the C+OpenMP oracle port of the reference's path on the host cores, bounded
         sample, rank 0, N = 1 only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs[3] (the configuration the metric is quoted on) and configs[4]
CONFIGS = {
    "1080p": {"w": 1920, "h": 1080, "levels": 3, "radius": 7, "pairs": 16, "ref_pairs": 2,
              "metric": "frame_pairs_per_sec_1080p_3level_lk",
              "workload": "synthetic 1920x1080 grayscale video, 3-level pyramid, 15x15 window (r=7), consecutive frame pairs"},
    "4k": {"w": 3840, "h": 2160, "levels": 4, "radius": 7, "pairs": 8, "ref_pairs": 1,
           "metric": "frame_pairs_per_sec_4k_4level_lk",
           "workload": "synthetic 3840x2160 grayscale video, 4-level pyramid, 15x15 window (r=7), consecutive frame pairs"},
}
W, H, LEVELS, RADIUS = 1920, 1080, 3, 7      # set from --config in main()
PAIRS = 16                      # frame pairs per step per GPU
METRIC = CONFIGS["1080p"]["metric"]
WORKLOAD = CONFIGS["1080p"]["workload"]
REF_PAIRS = 2
UNIT = "pairs/s"
MIN_TIMED_S = 0.5               # the timed region repeats the K-step loop until this much has elapsed ...
MIN_REPS = 5                    # ... and at least this many times; the median repetition is reported


def level_sizes(w, h, L):
    return [(w >> l) * (h >> l) for l in range(L)]


def algorithmic_bytes_per_pair(w, h, L):
    """pyramid 16*(sum_{l<L-1} N_l + sum_{l>=1} N_l) + per level 32*N_l (+16*N_{l+1})."""
    n = level_sizes(w, h, L)
    pyr = 16 * (sum(n[:-1]) + sum(n[1:]))
    lev = sum(32 * n[l] + (16 * n[l + 1] if l < L - 1 else 0) for l in range(L))
    return pyr + lev


def level0_bytes_per_pair(w, h, L):
    n = level_sizes(w, h, L)
    return 32 * n[0] + (16 * n[1] if L > 1 else 0)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def config_block(args, world):
    """The `config` object of the JSON line: the workload BASELINE.json names -- identical in both arms (ours and
    --impl reference); how much of it an arm's step covers is said in that arm's own fields (the reference arm's
    bounded sample: cpu_baseline.sample)."""
    frame_bytes = H * W * 8
    return {"workload": WORKLOAD, "name": args.config, "width": W, "height": H, "levels": LEVELS,
            "win_radius": args.radius, "pairs_per_step": args.pairs,
            "l2": "inputs+outputs per step (%.0f MB) exceed the 126 MB L2; no flush" %
                  (((args.pairs + 1) + 2 * args.pairs) * frame_bytes / 1e6),
            "exchange": "per-step gather of gesture statistics (32 B/pair) to rank 0" if world > 1 else "none"}


def recorded(name):
    """A committed measurement under profiles/ (ncu-derived constants the live run cannot take itself)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", name)))
    except Exception:
        return None


def fp64_roofline(radius, n_px, launch_ms, sm_mhz, n_sm):
    """Second roofline of the level-0 kernel: FP64 issue.  ops_per_px = FP64 instructions (DADD + DMUL +
    DFMA, per thread) the kernel executes per level-0 output pixel, counted by ncu on this kernel
    (profiles/fp64_ops.json); peak = 64 FP64 lanes / clk / SM (tools/ubench.cu, profiles/r01_ubench.txt:
    18.36 T instructions/s at 1.94 GHz on 148 SMs) at the SM clock sampled during the run."""
    rec = recorded("fp64_ops.json")
    if not rec or ("r%d" % radius) not in rec or not launch_ms or not sm_mhz:
        return None
    ops = float(rec["r%d" % radius])
    achieved = ops * n_px / (launch_ms * 1e-3) / 1e12
    peak = 64.0 * n_sm * sm_mhz * 1e6 / 1e12
    return {"ops_per_px": ops, "achieved": achieved, "peak": peak, "unit": "T FP64 instructions/s",
            "frac": achieved / peak, "source": rec.get("source")}


class ClockSampler:
    """Samples nvidia-smi clocks and throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            t0 = time.time()                       # nvidia-smi needs ~0.1-0.5 s to deliver its first row
            while not self.rows and time.time() - t0 < 3.0:
                time.sleep(0.01)
        except Exception:
            self.proc = None

    def mark(self):
        """Start of the timed region: drop what was sampled before it."""
        self.rows = []
        self.n_region = None

    def mark_end(self):
        """End of the timed region; sampling goes on (the end-to-end legs run the same kernels) so that a
        short --steps run still gets a reading."""
        self.n_region = len(self.rows)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        n_region = getattr(self, "n_region", None)
        window = "timed region"
        rows = self.rows
        if n_region is not None and n_region >= 3:
            rows = self.rows[:n_region]
        elif n_region is not None:
            window = "timed region (%d samples) + the end-to-end legs after it" % n_region
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": window}


_emit = print


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def synth(n_frames, seed=1234):
    import importlib
    from of_b200_loader import ofb  # noqa: F401
    return importlib.import_module("optical_flow_b200.synth").synth_frames(n_frames, H, W, seed=seed)


def cpu_baseline_run(frames, n_pairs, repeats=1, radius=None):
    """Times the C+OpenMP oracle port (oracle/lk_oracle.c) on n_pairs pairs of the workload,
    all host threads.  Returns (pairs_per_s, cores, seconds)."""
    from oracle import c_oracle as CO
    CO.build()
    cores = CO.set_threads(host_threads())
    CO.coarse_to_fine(frames[0][:64, :64].copy(), frames[1][:64, :64].copy(), 1, 1)   # thread-pool warm-up
    t0 = time.perf_counter()
    done = 0
    for _ in range(repeats):
        for k in range(n_pairs):
            CO.coarse_to_fine(frames[k % (len(frames) - 1)], frames[k % (len(frames) - 1) + 1], LEVELS,
                              RADIUS if radius is None else radius)
            done += 1
    dt = time.perf_counter() - t0
    return done / dt, cores, dt


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The reference itself
    (OpenCV 2.4 C++) cannot be compiled in this image, so this times the oracle port of it
    (oracle/lk_oracle.c, C + OpenMP, every host thread) on the same workload; one step = a
    bounded sample of REF_PAIRS pairs."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = REF_PAIRS
    frames = synth(sample + 1)
    from oracle import c_oracle as CO
    CO.build()
    cores = CO.set_threads(host_threads())      # torchrun exports OMP_NUM_THREADS=1: use every host thread anyway
    radius = args.radius
    for _ in range(max(args.warmup, 1)):
        CO.coarse_to_fine(frames[0], frames[1], LEVELS, radius)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for k in range(sample):
            CO.coarse_to_fine(frames[k], frames[k + 1], LEVELS, radius)
    dt = time.perf_counter() - t0
    val = args.steps * sample / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": config_block(args, args.gpus),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d steps x %d consecutive pairs of the same synthetic stream, "
                                   "oracle/lk_oracle.c (gcc -O3 -fopenmp), all host threads" % (args.steps, sample)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(json.dumps(line))


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from of_b200_loader import ofb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pairs = args.pairs
    radius = args.radius

    frames_h = synth(pairs + 1, seed=1234 + rank)             # each rank: its own stream segment (weak scaling)
    frames = torch.from_numpy(frames_h).to(dev)
    U = torch.empty((pairs, H, W), dtype=torch.float64, device=dev)
    V = torch.empty_like(U)
    S = torch.zeros((pairs, 4), dtype=torch.float64, device=dev)
    ctx = ofb.Context(local_rank, win_radius=radius)
    import importlib
    shard = importlib.import_module("optical_flow_b200.shard")
    counts = [pairs] * world

    def step():
        ctx.flow_batch_dev(frames, LEVELS, U, V, S)
        if world > 1:                                         # the path's only exchange: stats to rank 0
            shard.gather_to_rank0(S, counts)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    # ---- device-resident timed region ------------------------------------------------------
    # One repetition = args.steps steps between two CUDA events; repetitions go on until MIN_TIMED_S
    # have elapsed (and at least MIN_REPS times), so that a short --steps still gives the clocks
    # sampler and the per-launch statistics a real window.  The median repetition is reported.
    ctx.set_timing(True)
    lvl0_ms, lvl_ms, pyr_ms = [], [], []
    sampler = ClockSampler(torch.cuda.current_device() if os.environ.get("CUDA_VISIBLE_DEVICES") is None else local_rank)
    launches0 = ctx.launch_count
    sampler.start()
    barrier()
    sampler.mark()
    rep_ms = []
    t_wall = time.perf_counter()
    while True:
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for i in range(args.steps):
            step()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rep_ms.append(float(t.item()))
        go = torch.tensor([1.0 if (time.perf_counter() - t_wall < MIN_TIMED_S or len(rep_ms) < MIN_REPS) else 0.0],
                          dtype=torch.float64, device=dev)
        if world > 1:
            dist.broadcast(go, 0)                              # every rank runs the same number of repetitions
        if go.item() == 0.0 or len(rep_ms) >= 400:
            break
    sampler.mark_end()
    n_reps = len(rep_ms)
    launches = (ctx.launch_count - launches0) // n_reps
    ms_total = sorted(rep_ms)[n_reps // 2]
    # per-launch durations of every kernel launched inside the timed region (events on the
    # launching stream, recorded by the library around each launch; the most recent 8192)
    per_kernel = {}
    for kind, level, ms in ctx.get_timing():
        per_kernel.setdefault((kind, level), []).append(ms)
        if kind == 1 and level == 0:
            lvl0_ms.append(ms)
        elif kind == 1:
            lvl_ms.append(ms)
        else:
            pyr_ms.append(ms)
    ctx.set_timing(False)
    n_timed_steps = max(len(lvl0_ms), 1)                      # one level-0 launch per step
    value = world * pairs * args.steps / (ms_total * 1e-3)

    # ---- end-to-end through the host-pointer stream call --------------------------------------
    host_in = torch.empty((pairs + 1, H, W), dtype=torch.float64).pin_memory()
    host_in.copy_(torch.from_numpy(frames_h))
    host_u = torch.empty((pairs, H, W), dtype=torch.float64).pin_memory()
    host_v = torch.empty((pairs, H, W), dtype=torch.float64).pin_memory()
    e2e_steps = max(2, min(args.steps, 10))
    fin, hu, hv = host_in.numpy(), host_u.numpy(), host_v.numpy()
    ctx.flow_stream(fin, LEVELS, out_u=hu, out_v=hv)          # warm-up (allocates staging)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx.flow_stream(fin, LEVELS, out_u=hu, out_v=hv)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = world * pairs * e2e_steps / float(t.item())
    frame_bytes = H * W * 8
    e2e_stats_val = e2e_u8_val = None
    stats_bytes, n8 = 32 * pairs, 0
    if not args.lean:
        # the same call when only the per-pair gesture statistics come back (what the reference's
        # live loop consumes, ParallelVideoPyr.cpp:845-905): the read-back shrinks to bytes per pair
        ctx.flow_stream(fin, LEVELS, want_flow=False)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            _, _, stats_h = ctx.flow_stream(fin, LEVELS, want_flow=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_stats_val = world * pairs * e2e_steps / float(t.item())
        stats_bytes = int(stats_h.nbytes)
        # 8-bit frames in (what imread / VideoCapture deliver; converted on the device like the harness
        # mains' convertTo(CV_64FC1, 1.0/255)), statistics out: the call the live loop would make.
        # A longer stream (the 17 frames played forth and back) so that chunks overlap.
        idx = list(range(pairs + 1)) + list(range(pairs - 1, -1, -1)) + list(range(1, pairs + 1)) + list(range(pairs - 1, -1, -1))
        q8 = np.clip(np.rint(frames_h * 255.0), 0, 255).astype(np.uint8)[idx]
        host_q8 = torch.empty(q8.shape, dtype=torch.uint8).pin_memory()
        host_q8.copy_(torch.from_numpy(q8))
        f8 = host_q8.numpy()
        n8 = len(idx) - 1
        ctx.flow_stream(f8, LEVELS, want_flow=False)
        barrier()
        u8_steps = max(2, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(u8_steps):
            ctx.flow_stream(f8, LEVELS, want_flow=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_u8_val = world * n8 * u8_steps / float(t.item())
    clocks = sampler.stop()

    # ---- the reference's own window (3x3, r = 1; bit-exact path), device-resident like `value` ---
    r1 = None
    if radius != 1 and not args.lean:
        with ofb.Context(local_rank, win_radius=1) as c1:
            for _ in range(3):
                c1.flow_batch_dev(frames, LEVELS, U, V, S)
            barrier()
            c1.set_timing(True)
            r1_steps = max(2, min(args.steps, 20))
            e0.record()
            for _ in range(r1_steps):
                c1.flow_batch_dev(frames, LEVELS, U, V, S)
            e1.record()
            barrier()
            l0s = [ms for kind, level, ms in c1.get_timing() if kind == 1 and level == 0]
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            r1 = {"value": world * pairs * r1_steps / (float(t.item()) * 1e-3), "unit": UNIT, "steps": r1_steps,
                  "level0_launch_ms": sum(l0s) / len(l0s) if l0s else None,
                  "note": "same workload with the reference's own 3x3 window (get_Sum9), bit-exact kernel"}
            c1.set_timing(False)
            # ... and end to end through the host-pointer stream call, like `e2e`
            c1.flow_stream(fin, LEVELS, out_u=hu, out_v=hv)
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                c1.flow_stream(fin, LEVELS, out_u=hu, out_v=hv)
            torch.cuda.synchronize()
            t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            r1["e2e"] = {"value": world * pairs * e2e_steps / float(t.item()), "unit": UNIT,
                         "h2d_bytes_per_step": (pairs + 1) * frame_bytes, "d2h_bytes_per_step": 2 * pairs * frame_bytes,
                         "api": "of_ctf_f64_stream, win_radius = 1"}

    if rank == 0:
        peak, peak_src = measured_peaks()
        if r1 and r1["level0_launch_ms"]:
            r1["level0_frac"] = level0_bytes_per_pair(W, H, LEVELS) * pairs / (r1["level0_launch_ms"] * 1e-3) / 1e9 / peak
            r1["whole_path_frac"] = algorithmic_bytes_per_pair(W, H, LEVELS) * (r1["value"] / world) / (peak * 1e9)
        l0 = sum(lvl0_ms) / len(lvl0_ms) if lvl0_ms else None
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        roof = None
        if l0:
            achieved = level0_bytes_per_pair(W, H, LEVELS) * pairs / (l0 * 1e-3) / 1e9
            # DRAM bytes of one level-0 launch: ncu cannot run inside the bench, so the figure is the
            # committed ncu reading of this kernel at this batch size (profiles/traffic.json); null when
            # the batch size or configuration differs from the recorded one
            traffic, traffic_src = None, None
            rec = recorded("traffic.json")
            if rec and rec.get("config") == args.config and rec.get("pairs") == pairs:
                traffic = rec.get("level0_r%d_bytes_per_launch" % radius)
                traffic_src = rec.get("source")
            f64 = fp64_roofline(radius, W * H * pairs, l0, clocks.get("sm_mhz"), n_sm)
            hbm_frac = achieved / peak
            roof = {"bound": "fp64" if f64 and f64["frac"] > hbm_frac else "hbm",
                    "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": hbm_frac,
                    "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "fp64": f64, "kernel": "lk_level (level 0, fused)",
                    "launch_ms": l0, "algorithmic_bytes_per_launch": level0_bytes_per_pair(W, H, LEVELS) * pairs,
                    "whole_path_frac": algorithmic_bytes_per_pair(W, H, LEVELS) * (value / world) / (peak * 1e9),
                    "kernel_ms_per_step": {"level0": l0, "coarser_levels": sum(lvl_ms) / n_timed_steps,
                                           "pyr_down": sum(pyr_ms) / n_timed_steps}}
        if roof:
            # bytes-per-pixel roofline of every kernel of the step: fused level kernel of level l
            # 32 B per level-l pixel (+ 16 B per coarser pixel), pyrDown producing level l 8 B read per source
            # pixel + 8 B written per destination pixel, per frame
            n = level_sizes(W, H, LEVELS)
            table = []
            for (kind, level), v in sorted(per_kernel.items()):
                ms_launch = sum(v) / len(v)
                launches_per_step = len(v) / n_timed_steps
                if kind == 1:
                    b = (32 * n[level] + (16 * n[level + 1] if level < LEVELS - 1 else 0)) * pairs
                    name = "lk_level L%d" % level
                else:
                    b = 8 * (n[level - 1] + n[level]) * (pairs + 1) / launches_per_step
                    name = "pyrDown -> L%d" % level
                gbs = b / (ms_launch * 1e-3) / 1e9
                table.append({"kernel": name, "launch_ms": ms_launch, "launches_per_step": launches_per_step,
                              "algorithmic_bytes_per_launch": int(b), "achieved_gbs": gbs, "frac": gbs / peak})
            roof["per_kernel"] = table
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            n_cpu = 64 if W * H <= 1920 * 1080 else 16           # about 10-15 s of CPU work
            v, cores, secs = cpu_baseline_run(frames_h, n_cpu, radius=radius)
            from oracle import c_oracle as CO
            CO.set_threads(8)                  # the reference's hard-coded team size (ParallelVideoPyr.cpp:164)
            n8t = 8 if W * H <= 1920 * 1080 else 4
            t0 = time.perf_counter()
            for k in range(n8t):
                CO.coarse_to_fine(frames_h[k], frames_h[k + 1], LEVELS, radius)
            v8 = n8t / (time.perf_counter() - t0)
            CO.set_threads(host_threads())
            n31 = 16 if W * H <= 1920 * 1080 else 4
            t0 = time.perf_counter()
            for k in range(n31):
                CO.coarse_to_fine(frames_h[k % pairs], frames_h[k % pairs + 1], LEVELS, 1)
            v31 = n31 / (time.perf_counter() - t0)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "value_8_threads": v8,
                   "value_3x3_window": v31,
                   "sample": "%d pairs (the first %d consecutive pairs of the same synthetic stream, L=%d, r=%d, cycled), "
                             "oracle/lk_oracle.c C+OpenMP port of the reference path, all host threads, %.1f s"
                             % (n_cpu, pairs, LEVELS, radius, secs)}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "timed": {"repetitions": n_reps, "steps_per_repetition": args.steps, "reported": "median repetition",
                      "rep_ms_min": min(rep_ms), "rep_ms_max": max(rep_ms), "min_timed_s": MIN_TIMED_S},
            "config": config_block(args, world),
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": (pairs + 1) * frame_bytes,
                    "d2h_bytes_per_step": 2 * pairs * frame_bytes, "steps": e2e_steps,
                    "api": "of_ctf_f64_stream (pinned host frames in, full u/v flow fields out)"},
            "e2e_stats_only": None if args.lean else {"value": e2e_stats_val, "unit": UNIT, "h2d_bytes_per_step": (pairs + 1) * frame_bytes,
                               "d2h_bytes_per_step": stats_bytes,
                               "api": "of_ctf_f64_stream, u = v = NULL: gesture statistics per pair only"},
            "e2e_u8_stats_only": None if args.lean else {"value": e2e_u8_val, "unit": UNIT, "h2d_bytes_per_step": (n8 + 1) * H * W,
                                  "d2h_bytes_per_step": 32 * n8, "pairs_per_call": n8,
                                  "api": "of_ctf_u8_stream: 8-bit frames in (the synthetic stream quantised to 8 bits), "
                                         "converted on the device, gesture statistics per pair out"},
            "window_3x3": r1,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roof,
            "cpu_baseline": cpu,
        }
        _emit(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="1080p", choices=sorted(CONFIGS),
                    help="1080p = BASELINE.json configs[3] (the metric's configuration), 4k = configs[4]")
    ap.add_argument("--pairs", type=int, default=None, help="frame pairs per step per GPU (default: the configuration's)")
    ap.add_argument("--radius", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lean", action="store_true",
                    help="only the device-resident measurement and the e2e leg (scaling sweeps): no statistics-only, "
                         "8-bit, 3x3 and CPU legs")
    args = ap.parse_args()
    if args.lean:
        args.no_cpu_baseline = True
    global W, H, LEVELS, RADIUS, PAIRS, METRIC, WORKLOAD, REF_PAIRS
    c = CONFIGS[args.config]
    W, H, LEVELS, RADIUS, PAIRS, METRIC, REF_PAIRS = c["w"], c["h"], c["levels"], c["radius"], c["pairs"], c["metric"], c["ref_pairs"]
    if args.pairs is None:
        args.pairs = PAIRS
    if args.radius is None:
        args.radius = RADIUS
    WORKLOAD = c["workload"].replace("15x15 window (r=7)", "%dx%d window (r=%d)" % (2 * args.radius + 1, 2 * args.radius + 1, args.radius))
    # torchrun exports OMP_NUM_THREADS=1; the CPU legs (reference arm, cpu_baseline) are to use every
    # host thread, and libgomp reads the variable when it is loaded, so set it before anything imports it
    os.environ["OMP_NUM_THREADS"] = str(host_threads())
    # stdout carries exactly one JSON line: anything a library prints there (NCCL's version banner, ...)
    # goes to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    global _emit
    _emit = lambda text: os.write(real_stdout, (text + "\n").encode())
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
