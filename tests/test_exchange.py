"""Gradient exchange over peer memory (csrc/nsb_exchange.cu): slice arithmetic on the CPU, the kernels
on >= 2 GPUs of one box (scripts/exchange_check.py under torchrun: all-reduce and fused reduce + Adam
against NCCL + nsb_adam, both the NVLS multicast and the peer load/store paths, CUDA-graph replay)."""
import ctypes as C
import json
import os
import subprocess
import sys

import pytest
import torch

from nice_slam_annotation_b200 import _lib

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exchange_slices_partition_the_range():
    L = _lib.lib()
    for world in (1, 2, 3, 4, 8):
        for n in (0, 4, 64, 4 * 1001, 32 * 4099, 12120192):
            nxt, lo = 0, C.c_int64()
            for r in range(world):
                cnt = L.nsb_exchange_slice(r, world, n, C.byref(lo))
                assert cnt % 4 == 0 and lo.value % 4 == 0
                if cnt:
                    assert lo.value == nxt
                    nxt += cnt
            assert nxt == n
    assert L.nsb_exchange_slice(0, 2, 6, None) == -1          # not a multiple of 4 floats
    assert L.nsb_exchange_signal_words() >= 2 * 8


def test_exchange_rejects_bad_groups():
    L = _lib.lib()
    p = _lib.NsbPeers()
    p.rank, p.world = 0, 9
    assert L.nsb_bucket_allreduce(C.byref(p), 0, 64, None) == _lib.NSB_EINVAL
    p.world = 2                                                # no mappings filled in
    assert L.nsb_bucket_allreduce(C.byref(p), 0, 64, None) == _lib.NSB_EINVAL
    assert b"rank" in L.nsb_last_error()


@pytest.mark.gpu
def test_exchange_two_gpus(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on one box")
    out = tmp_path / "x.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(REPO, "scripts", "exchange_check.py"), "--json", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=REPO)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    res = json.load(open(out))
    assert res["checks"]["graph_replay"] is True
    assert any(k.startswith("reduce_adam_") for k in res["checks"])


@pytest.mark.gpu
def test_sharded_mapper_two_gpus(tmp_path):
    """Mapper.optimize_map after enable_sharding (rays of every iteration split over 2 ranks, gradients
    summed through the peer-memory exchange inside the iteration's CUDA graph, identical Adam step on
    every rank) reproduces the single-GPU call, with frustum masks and with local BA, and the replicas
    stay bit-identical (scripts/sharded_map_check.py)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on one box")
    out = tmp_path / "s.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29534", os.path.join(REPO, "scripts", "sharded_map_check.py"), "--iters", "6",
           "--json", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=REPO)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    res = json.load(open(out))
    assert set(res["checks"]) == {"no_BA", "BA"}
