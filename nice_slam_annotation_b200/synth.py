"""Synthetic RGB-D frames of an axis-aligned box room.

The benchmark, the GPU tests and the golden-vector generator all need RGB-D input of the
reference's dataset shapes but there is no dataset on the box (no network), so frames are
rendered analytically: the camera sits inside a box room (the scene bound shrunk by
``margin`` per side), depth is the exact ray/box z-depth (the reference's rays are not
normalised, their parameter IS z-depth: src/common.py:82-87) and colour is
``0.5 + 0.5 sin(3 x_world)``.  A fraction of pixels is set to depth 0 to exercise the
zero-depth branch of the sampler (src/utils/Renderer.py:193-200).
Pure torch, CPU or CUDA; no dependency on the CUDA extension.
"""
import math

import torch

# dataset shapes used by the reference configs (after crop_edge), SURVEY.md section 8
SHAPES = {
    # name: (H, W, fx, fy, cx, cy, bound)
    "demo": (460, 620, 577.590698, 578.729797, 308.905426, 232.683609,
             [[0.0, 6.5], [0.0, 4.0], [0.0, 3.5]]),
    "replica_room0": (680, 1200, 600.0, 600.0, 599.5, 339.5,
                      [[-2.9, 8.9], [-3.2, 5.5], [-3.5, 3.3]]),
    "scannet_scene0000": (460, 620, 577.590698, 578.729797, 308.905426, 232.683609,
                          [[-2.0, 11.0], [-2.0, 11.5], [-2.0, 5.5]]),
    "apartment": (720, 1280, 607.4694213867188, 607.4534912109375, 636.9967041015625, 369.2689514160156,
                  [[-5.8, 11.3], [-4.0, 4.5], [-7.9, 4.9]]),
    # configs/TUM_RGBD/freiburg3_office.yaml: 480x640 resized to 384x512, crop_edge 8; 127^3 fine/colour grids
    "tum_fr3_office": (368, 496, 428.32, 431.36, 248.08, 190.08,
                       [[-5.0, 5.0], [-5.0, 5.0], [-5.0, 5.0]]),
}
# per-shape grid_len overrides of the dataset configs (configs/TUM_RGBD/tum.yaml:12-16)
GRID_LEN = {"tum_fr3_office": {"coarse": 2, "middle": 0.16, "fine": 0.08, "color": 0.08}}


def look_at(eye, target, up=(0.0, 0.0, 1.0)):
    """c2w (4x4, fp32) of a camera at ``eye`` whose -z axis points at ``target``."""
    eye = torch.as_tensor(eye, dtype=torch.float64)
    fwd = torch.as_tensor(target, dtype=torch.float64) - eye
    fwd = fwd / fwd.norm()
    upv = torch.as_tensor(up, dtype=torch.float64)
    right = torch.linalg.cross(fwd, upv)
    right = right / right.norm()
    up2 = torch.linalg.cross(right, fwd)
    c2w = torch.eye(4, dtype=torch.float64)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, up2, -fwd, eye
    return c2w.float()


def room_of(bound, margin=0.5):
    b = torch.as_tensor(bound, dtype=torch.float64)
    return torch.stack([b[:, 0] + margin, b[:, 1] - margin], 1)


def make_frame(H, W, fx, fy, cx, cy, room, c2w, zero_frac=0.05, seed=0, device="cpu"):
    """depth [H,W] fp32, colour [H,W,3] fp32 of the box room seen from c2w."""
    room = torch.as_tensor(room, dtype=torch.float64, device=device)
    c2w = c2w.to(device=device, dtype=torch.float64)
    jj, ii = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float64),
                            torch.arange(W, device=device, dtype=torch.float64), indexing="ij")
    dirs = torch.stack([(ii - cx) / fx, -(jj - cy) / fy, -torch.ones_like(ii)], -1)
    d = dirs @ c2w[:3, :3].T
    o = c2w[:3, 3]
    t = (room.T[:, None, None, :] - o) / d            # [2,H,W,3]
    depth = t.max(0)[0].min(-1)[0]
    hit = o + d * depth[..., None]
    color = 0.5 + 0.5 * torch.sin(3.0 * hit)
    if zero_frac > 0:
        g = torch.Generator().manual_seed(seed)
        drop = (torch.rand(H, W, generator=g) < zero_frac).to(device)
        depth = torch.where(drop, torch.zeros_like(depth), depth)
    return depth.float(), color.float()


def orbit_poses(room, n, seed=0):
    """n camera poses inside the room looking roughly across it (deterministic)."""
    room = torch.as_tensor(room, dtype=torch.float64)
    ctr = room.mean(1)
    ext = room[:, 1] - room[:, 0]
    g = torch.Generator().manual_seed(seed)
    poses = []
    for k in range(n):
        a = 2 * math.pi * k / max(n, 1) + 0.3
        jit = (torch.rand(3, generator=g, dtype=torch.float64) - 0.5) * 0.2
        eye = ctr + torch.tensor([math.cos(a), math.sin(a), 0.0], dtype=torch.float64) * ext * 0.15 + jit * ext
        tgt = ctr + torch.tensor([math.cos(a + 2.5), math.sin(a + 2.5), -0.1], dtype=torch.float64) * ext * 0.45
        poses.append(look_at(eye, tgt))
    return poses
