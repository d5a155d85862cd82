"""CPU suite: the multi-GPU host logic (pair partitioning + gather to rank 0) with
world_size 2 over gloo.  The per-rank "flow" is replaced by a deterministic stand-in so that
no GPU is needed; the GPU path itself is covered by -m gpu tests."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _shard():
    from of_b200_loader import ofb  # noqa: F401
    return importlib.import_module("optical_flow_b200.shard")


def test_partition_covers_all_pairs_once():
    sh = _shard()
    for n in (0, 1, 7, 63, 194):
        for world in (1, 2, 3, 4, 8):
            got = []
            for r in range(world):
                a, b = sh.partition_pairs(n, world, r)
                assert 0 <= a <= b <= n
                got += list(range(a, b))
            assert got == list(range(n))
            sizes = [sh.partition_pairs(n, world, r)[1] - sh.partition_pairs(n, world, r)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.partition_pairs(4, 2, 2)


def _worker(rank, world, port, n_pairs, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = _shard()
        a, b = sh.partition_pairs(n_pairs, world, rank)
        counts = [sh.partition_pairs(n_pairs, world, r)[1] - sh.partition_pairs(n_pairs, world, r)[0]
                  for r in range(world)]
        # stand-in per-pair statistics: row k = (k, 2k, 3k, k % 2)
        k = torch.arange(a, b, dtype=torch.float64)
        local = torch.stack([k, 2 * k, 3 * k, k % 2], dim=1)
        res = sh.gather_to_rank0(local, counts)
        if rank == 0:
            np.save(out, res.numpy())
        else:
            assert res is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_pairs", [7, 8])
def test_gather_stats_world2_gloo(tmp_path, n_pairs):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "g.npy")
    mp.spawn(_worker, args=(2, port, n_pairs, out), nprocs=2, join=True)
    g = np.load(out)
    k = np.arange(n_pairs, dtype=np.float64)
    assert np.array_equal(g, np.stack([k, 2 * k, 3 * k, k % 2], axis=1))
