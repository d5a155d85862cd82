"""Host<->device copy bandwidth of this box (pinned memory): H2D alone, D2H alone and both at once, per
GPU and summed over the ranks when launched under torchrun (one rank per GPU, all ranks copying at the
same time).  It bounds the end-to-end (host-pointer) numbers of bench.py: a 1080p pair moves 16.6 MB in
and 33.2 MB out.

    python tools/pcie_probe.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py

Rank 0 prints one JSON line."""
import json
import os
import time

import torch
import torch.distributed as dist

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

n = 256 << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def run(up, down, reps=8):
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)           # all ranks copy concurrently: the slowest one counts
    return n * reps / float(dt.item()) / 1e9


run(True, True, 2)
res = {"n_gpus": world, "per_gpu_gbs": {"h2d_alone": run(True, False), "d2h_alone": run(False, True),
                                        "each_direction_when_both": run(True, True)}}
res["aggregate_gbs"] = {k: v * world for k, v in res["per_gpu_gbs"].items()}
res["host_threads"] = len(os.sched_getaffinity(0))
res["note"] = "256 MiB pinned buffers, 8 repetitions, every rank copying at the same time; per-GPU = slowest rank"
if rank == 0:
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
