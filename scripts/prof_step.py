#!/usr/bin/env python
"""Developer probe: a few un-graphed engine steps (bench.py's step) of one shape / ray count, for ncu captures and
for the in-kernel phase timers of an NSB_TC_PROFILE build (NSB_LIB=<that .so> NSB_TC_PROFILE_PRINT=1).

    python scripts/prof_step.py [shape] [rays] [iters]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nice_slam_annotation_b200.scene import build_scene  # noqa: E402
from nice_slam_annotation_b200.step import FusedStep  # noqa: E402

shape = sys.argv[1] if len(sys.argv) > 1 else "replica_room0"
rays = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dev = "cuda:0"
ns = bench.desaturate(build_scene(shape, dev, seed=0, n_frames=bench.FRAMES))
bench.train_only_color_decoder(ns.shared_decoders)
step = FusedStep(ns.renderer, ns.shared_decoders, bench.grids_of(ns), "color", rays, **bench.STEP_KW)
batches = bench.make_batches(ns, rays, 2, seed=100)
flush = torch.empty(bench.L2_FLUSH_BYTES // 4, device=dev)
for i in range(iters):
    flush.fill_(float(i))
    step.load(*batches[i % 2])
    step.run()
torch.cuda.synchronize()
print("done", float(step.loss))
