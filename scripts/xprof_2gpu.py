#!/usr/bin/env python
"""Developer probe (torchrun, 2+ GPUs): a few un-graphed sharded engine steps with NSB_X_PROFILE_PRINT=1 so that the
sparse exchange prints its phase times (scan / marks, opening barrier incl. the wait for the slowest rank, reduce +
store, closing barrier)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nice_slam_annotation_b200.peers import PeerGroup  # noqa: E402
from nice_slam_annotation_b200.scene import build_scene  # noqa: E402
from nice_slam_annotation_b200.step import FusedStep  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = f"cuda:{local}"
dist.init_process_group("nccl", device_id=torch.device(dev))
ns = bench.desaturate(build_scene("replica_room0", dev, seed=0, n_frames=bench.FRAMES))
bench.train_only_color_decoder(ns.shared_decoders)
peers = PeerGroup(device=torch.device(dev))
step = FusedStep(ns.renderer, ns.shared_decoders, bench.grids_of(ns), "color", 1000, peers=peers, **bench.STEP_KW)
batches = bench.make_batches(ns, 1000, 4, seed=100 + rank)
flush = torch.empty(bench.L2_FLUSH_BYTES // 4, device=dev)
for i in range(6):
    flush.fill_(float(i))
    dist.barrier()
    step.load(*batches[i % 4])
    step.run()
torch.cuda.synchronize()
peers.check_error()
touched = float((step.bucket.flat.view(-1, 32) != 0).any(1).float().mean())
print(f"rank {rank}: lines non-zero after the exchange {touched:.3f}", flush=True)
dist.barrier()
os._exit(0)
