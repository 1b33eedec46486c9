"""World-size-2 and -3 tests (gloo, CPU) of the multi-GPU host logic: row sharding y % G == rank and the
framebuffer gather + de-interleave of ray-tracer_b200/dist.py.  The oracle stands in as the shard renderer (it is the
checker of what each rank must produce); on the GPU box the same gather runs over NCCL with the CUDA kernel."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, height, width, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    for p in (ROOT, os.path.join(ROOT, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rtdist = importlib.import_module("ray-tracer_b200.dist")
        import oracle_py as orc

        scene = orc.Scene("cornell-box", 1)
        spp, depth, seed = 2, 6, 9
        rows = rtdist.shard_rows(height, rank, world)
        shard, _, _ = scene.render(width, height, spp, depth, seed=seed, row_begin=rank, row_stride=world)
        assert shard.shape[0] == rows
        local = torch.zeros((rtdist.padded_rows(height, world), width, 3), dtype=torch.float32)
        local[:rows] = torch.from_numpy(shard.astype(np.float32))
        full = rtdist.gather_rows(local, height, world)
        assert full.shape == (height, width, 3)
        if rank == 0:
            ref, _, _ = scene.render(width, height, spp, depth, seed=seed)
            np.save(os.path.join(out_dir, "full.npy"), full.numpy())
            np.save(os.path.join(out_dir, "ref.npy"), ref.astype(np.float32))
        # every rank must hold the same image
        chk = torch.tensor([float(full.double().sum())], dtype=torch.float64)
        lst = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(lst, chk)
        assert all(float(v) == float(lst[0]) for v in lst)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,height", [(2, 12), (3, 10), (2, 7)])  # 10 % 3 != 0 and 7 % 2 != 0: ragged shards are padded
def test_row_shards_gather_to_the_full_image(orc, tmp_path, world, height):
    width = 9
    mp.spawn(_worker, args=(world, _free_port(), height, width, str(tmp_path)), nprocs=world, join=True)
    full, ref = np.load(tmp_path / "full.npy"), np.load(tmp_path / "ref.npy")
    assert np.array_equal(full, ref)


def test_shard_arithmetic():
    rtdist = importlib.import_module("ray-tracer_b200.dist")
    for h in (1, 7, 600, 2160):
        for g in (1, 2, 3, 4, 8):
            assert sum(rtdist.shard_rows(h, r, g) for r in range(g)) == h
            assert rtdist.padded_rows(h, g) == max(rtdist.shard_rows(h, r, g) for r in range(g))
