"""Worker of test_torchrun_two_ranks_reproduce_single_gpu_rows (launched under torchrun, one rank per GPU):
each rank computes its contiguous block of pairs of the same 33-frame stream on its own GPU and the
per-pair statistics are gathered to rank 0 (shard.gather_to_rank0, NCCL)."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from of_b200_loader import ofb

shard = importlib.import_module("optical_flow_b200.shard")
synth = importlib.import_module("optical_flow_b200.synth")

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
fr = synth.synth_frames(33, 272, 480, seed=1234, gesture=True)
n_pairs = fr.shape[0] - 1
a, b = shard.partition_pairs(n_pairs, world, rank)
counts = [shard.partition_pairs(n_pairs, world, r)[1] - shard.partition_pairs(n_pairs, world, r)[0] for r in range(world)]
d = torch.from_numpy(fr[a:b + 1]).cuda()
S = torch.zeros((b - a, 4), dtype=torch.float64, device="cuda")
with ofb.Context(local, win_radius=7, mag_thresh=0.5, count_thresh=100) as ctx:
    ctx.flow_batch_dev(d, 3, None, None, S)
    torch.cuda.synchronize()
rows = shard.gather_to_rank0(S, counts)
if rank == 0:
    np.save(sys.argv[1], rows.cpu().numpy())
dist.destroy_process_group()
