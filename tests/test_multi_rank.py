"""World-size-2 checks (gloo, CPU) of the host-side logic of the ray-sharded mapping step
(step.py): batch sharding, the flat gradient bucket and its single SUM
all-reduce, and the MAX-reduced batch depth scalar.  The kernels themselves are covered by the
-m gpu tests; here each rank fabricates its gradients."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nice_slam_annotation_b200.step import GradBucket, global_depth_max, shard_slice


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                                   # identical replicas on every rank
        grid = torch.randn(1, 32, 3, 4, 5).contiguous(memory_format=torch.channels_last_3d)
        weight = torch.randn(32, 93)
        bias = torch.randn(32)
        bucket = GradBucket([grid, weight, bias])
        # views share the leaves' shape and memory layout
        vg = bucket.view_of(grid)
        assert vg.shape == grid.shape and vg.stride() == grid.stride()
        # a rank's "backward": gradients of ITS rays only
        n = 101
        sl = shard_slice(n, world, rank)
        rays = torch.arange(n, dtype=torch.float32)
        bucket.zero()
        for r in rays[sl]:
            vg.add_(r * 1e-3)
            bucket.view_of(weight).add_(r)
            bucket.view_of(bias)[int(r) % 32] += 1.0
        bucket.all_reduce()
        bucket.assign()
        total = rays.sum().item()
        ok = torch.allclose(grid.grad, torch.full_like(grid, total * 1e-3), rtol=1e-5)
        ok &= torch.allclose(weight.grad, torch.full_like(weight, total))
        ok &= abs(bias.grad.sum().item() - n) < 1e-6
        ok &= grid.grad.data_ptr() == vg.data_ptr()               # no copy between bucket and .grad
        # batch-global depth scalar: max over the kept rays of all ranks
        gd = torch.tensor([1.0, 2.0 + rank, 9.0], dtype=torch.float32)
        keep = torch.tensor([1, 1, 0], dtype=torch.uint8)
        m = global_depth_max(gd, keep)
        ok &= float(m) == 2.0 + (world - 1)
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_shard_slice_partitions_the_batch():
    for n in (0, 1, 7, 1000, 1001):
        for world in (1, 2, 3, 8):
            idx = []
            for r in range(world):
                s = shard_slice(n, world, r)
                idx += list(range(n))[s]
            assert idx == list(range(n))
            sizes = [len(range(n)[shard_slice(n, world, r)]) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_bucket_layout_single_process():
    grid = torch.randn(1, 32, 2, 3, 4).contiguous(memory_format=torch.channels_last_3d)
    w = torch.randn(4, 32)
    b = GradBucket([grid, w])
    assert b.flat.numel() % 64 == 0 and b.offsets[1] % 64 == 0
    b.view_of(grid)[0, 5, 1, 2, 3] = 7.0
    # channels-last: element (c, z, y, x) lives at ((z*Y + y)*X + x)*32 + c of the segment
    assert b.flat[((1 * 3 + 2) * 4 + 3) * 32 + 5] == 7.0
    b.zero()
    assert float(b.flat.abs().sum()) == 0.0
    with pytest.raises(KeyError):
        b.view_of(torch.zeros(1))


def test_sharded_gradients_sum_over_two_ranks():
    world = 2
    mgr = mp.get_context("spawn").Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert dict(ret) == {0: True, 1: True}
