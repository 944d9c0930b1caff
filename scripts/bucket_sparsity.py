"""How much of the dense gradient bucket does one mapping step touch?  (decides whether a sparse exchange
pays: SURVEY.md 8e "compact masked gradients").  Replica room0 bench scene, N rays from 5 frames."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as B                                                     # noqa: E402
from nice_slam_annotation_b200.scene import build_scene               # noqa: E402
from nice_slam_annotation_b200.step import FusedStep                  # noqa: E402

dev = "cuda:0"
ns = build_scene(B.SHAPE, dev, seed=0, n_frames=B.FRAMES)
out = []
for rays in (1000, 2000, 4000, 8000):
    grids = {k: v.detach() for k, v in ns.shared_c.items()}
    batches = B.make_batches(ns, rays, 1, seed=100)
    step = FusedStep(ns.renderer, ns.shared_decoders, grids, "color", rays, loss="mapping", w_color=0.2, use_color=True,
                     grid_grads=("grid_middle", "grid_fine", "grid_color"), decoder_grads=("color_decoder",))
    step.load(*batches[0])
    step.run()
    torch.cuda.synchronize()
    flat = step.bucket.flat
    row = {"rays": rays, "bucket_mb": flat.numel() * 4 / 1e6}
    for name, chunk_floats in (("128B", 32), ("4KB", 1024), ("16KB", 4096), ("64KB", 16384)):
        n = flat.numel() // chunk_floats * chunk_floats
        nz = (flat[:n].view(-1, chunk_floats) != 0).any(1).float().mean().item()
        row["touched_" + name] = round(nz, 4)
    out.append(row)
    print(json.dumps(row), flush=True)
