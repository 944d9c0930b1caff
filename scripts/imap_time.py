#!/usr/bin/env python
"""Timing of the iMAP* path on one GPU (BASELINE configs[4], the tensor-core MLP stress case):
MLP forward / forward+backward over P points, and Renderer.render_batch_ray forward + backward at
the iMAP config (N rays x (32 + 12) samples, density compositing).  Prints one JSON object.

    python scripts/imap_time.py [--rays 5000,20000] [--steps 10]
"""
import argparse
import json
import os
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nice_slam_annotation_b200 import synth  # noqa: E402
from nice_slam_annotation_b200.common import get_samples  # noqa: E402
from nice_slam_annotation_b200.conv_onet.models.decoder import MLP  # noqa: E402
from nice_slam_annotation_b200.scene import load_bound  # noqa: E402
from nice_slam_annotation_b200.utils.Renderer import Renderer  # noqa: E402

FLOP_FWD = 2 * (3 * 93 + 93 * 256 + 3 * 256 * 256 + 256 * 4)      # per point: embedding + 4 layers + head


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", default="5000,20000")
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    dev = "cuda:0"
    torch.manual_seed(0)
    H, W, fx, fy, cx, cy, bnd = synth.SHAPES["replica_room0"]
    scale = 0.1                                                     # configs/imap.yaml: scale 0.1
    bound = load_bound(bnd, scale)
    dec = MLP(name="", dim=3, c_dim=0, color=True, hidden_size=256, skips=[], n_blocks=4,
              pos_embedding_method="fourier").to(dev)
    with torch.no_grad():
        dec.output_linear.bias.add_(torch.tensor([0.3, 0.5, 0.4, 2.0], device=dev))
    dec.bound = bound
    cfg = {"rendering": {"lindisp": False, "perturb": 0.0, "N_samples": 32, "N_surface": 0, "N_importance": 12},
           "scale": scale, "occupancy": False, "model": {"coarse_bound_enlarge": 2}}
    slam = types.SimpleNamespace(nice=False, bound=bound, H=H, W=W, fx=fx, fy=fy, cx=cx, cy=cy)
    r = Renderer(cfg, None, slam)
    room = synth.room_of(bound, margin=0.05)
    c2w = synth.orbit_poses(room, 1, seed=0)[0].to(dev)
    depth, color = synth.make_frame(H, W, fx, fy, cx, cy, room, c2w, zero_frac=0.05, seed=0, device=dev)
    rows = []
    for n in [int(x) for x in a.rays.split(",")]:
        idx = torch.randint(H * W, (n,), generator=torch.Generator().manual_seed(n))
        ro, rd, gd, gc = get_samples(0, H, 0, W, n, H, W, fx, fy, cx, cy, c2w, depth, color, dev, indices=idx)
        P = n * 44
        pts = (torch.rand(P, 3, device=dev) - 0.5) * 0.5

        def fwd_only():
            with torch.no_grad():
                dec(pts)

        def fwd_bwd():
            out = dec(pts)
            out.sum().backward()
            for q in dec.parameters():
                q.grad = None

        def render_step():
            d, _, c = r.render_batch_ray({}, dec, rd, ro, dev, "color", gt_depth=gd)
            loss = (d - gd).abs().sum() + 0.05 * (c - gc).abs().sum()
            loss.backward()
            for q in dec.parameters():
                q.grad = None

        ms_f, ms_fb, ms_r = timed(fwd_only, a.steps), timed(fwd_bwd, a.steps), timed(render_step, a.steps)
        rows.append({"rays": n, "points": P, "mlp_forward_ms": ms_f, "mlp_forward_tflops": FLOP_FWD * P / ms_f / 1e9,
                     "mlp_fwd_bwd_ms": ms_fb, "mlp_fwd_bwd_tflops": 3 * FLOP_FWD * P / ms_fb / 1e9,
                     "render_fwd_bwd_ms": ms_r, "render_rays_per_s": n / (ms_r * 1e-3),
                     "render_tflops": FLOP_FWD * n * (32 + 3 * 44) / ms_r / 1e9})
        print(json.dumps(rows[-1]), file=sys.stderr, flush=True)
    print(json.dumps({"what": "iMAP* path, 1 B200: MLP 93-256-256-256-256-4 as chained 3xTF32 tcgen05 GEMMs; TFLOP/s are "
                              "algorithmic fp32 FLOPs (3xTF32 issues 3x as many tensor-core FLOPs)", "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
