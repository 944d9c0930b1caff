"""Scene set-up around the hot path: what ``NICE_SLAM.__init__`` does before the first render
(src/NICE_SLAM.py:188-320), restated for the product so that benchmarks, smoke tests and
integrations can build decoders + grids without the rest of the SLAM system.

``grid_init`` allocates the grids CHANNELS-LAST (torch.channels_last_3d): the logical shape
stays the reference's [1, c_dim, Z, Y, X] (checkpoints, Mesher, boolean masks keep working),
while memory holds one 128-byte line per voxel, which is what the gather/scatter kernels read.
"""
import types

import torch

from . import synth
from .conv_onet import config as conv_config

DEFAULT_GRID_LEN = {"coarse": 2, "middle": 0.32, "fine": 0.16, "color": 0.16}


def load_bound(bound, scale=1.0, divisible=0.32):
    """Scene bound as an fp64 [3,2] tensor whose upper ends are pushed to lo + k*0.32 so that
    every level divides it (NICE_SLAM.py:188-201)."""
    b = torch.as_tensor(bound, dtype=torch.float64).clone() * scale
    b[:, 1] = (((b[:, 1] - b[:, 0]) / divisible).int() + 1) * divisible + b[:, 0]
    return b


def assign_bound(decoders, bound, coarse_bound_enlarge=2):
    """NICE_SLAM.py:202-208: every decoder carries the bound; the coarse one an enlarged copy."""
    decoders.middle_decoder.bound = bound
    decoders.fine_decoder.bound = bound
    decoders.color_decoder.bound = bound
    if hasattr(decoders, "coarse_decoder"):
        decoders.coarse_decoder.bound = bound * coarse_bound_enlarge
    return decoders


def grid_init(bound, grid_len=None, c_dim=32, coarse=True, coarse_bound_enlarge=2, device="cpu", generator=None):
    """Feature grids {'grid_coarse', 'grid_middle', 'grid_fine', 'grid_color'} initialised like
    NICE_SLAM.py:243-320 (N(0, 0.01); fine N(0, 1e-4)), channels-last in memory."""
    grid_len = dict(DEFAULT_GRID_LEN, **(grid_len or {}))
    xyz_len = bound[:, 1] - bound[:, 0]
    c = {}
    for name, std in (("coarse", 0.01), ("middle", 0.01), ("fine", 0.0001), ("color", 0.01)):
        if name == "coarse" and not coarse:
            continue
        ext = xyz_len * coarse_bound_enlarge if name == "coarse" else xyz_len
        shape = list(map(int, (ext / grid_len[name]).tolist()))
        shape[0], shape[2] = shape[2], shape[0]
        val = torch.zeros([1, c_dim, *shape]).normal_(mean=0, std=std, generator=generator)
        c["grid_" + name] = val.to(device).contiguous(memory_format=torch.channels_last_3d)
    return c


def base_cfg(device="cuda:0"):
    """The hot-path keys of configs/nice_slam.yaml (rendering :121-126, model, grid_len :9-14,
    tracking :36-52, mapping :53-107) with their reference defaults."""
    stage = {
        "coarse": {"decoders_lr": 0.0, "coarse_lr": 0.001, "middle_lr": 0.0, "fine_lr": 0.0, "color_lr": 0.0},
        "middle": {"decoders_lr": 0.0, "coarse_lr": 0.0, "middle_lr": 0.1, "fine_lr": 0.0, "color_lr": 0.0},
        "fine": {"decoders_lr": 0.0, "coarse_lr": 0.0, "middle_lr": 0.005, "fine_lr": 0.005, "color_lr": 0.0},
        "color": {"decoders_lr": 0.005, "coarse_lr": 0.0, "middle_lr": 0.005, "fine_lr": 0.005, "color_lr": 0.005},
    }
    return {
        "scale": 1, "coarse": True, "occupancy": True, "sync_method": "strict",
        "data": {"dim": 3},
        "grid_len": dict(DEFAULT_GRID_LEN),
        "model": {"c_dim": 32, "coarse_bound_enlarge": 2, "pos_embedding_method": "fourier"},
        "rendering": {"N_samples": 32, "N_surface": 16, "N_importance": 0, "lindisp": False, "perturb": 0.0},
        "tracking": {"lr": 1e-3, "device": device, "iters": 10, "gt_camera": False, "pixels": 200,
                     "seperate_LR": False, "w_color_loss": 0.5, "ignore_edge_W": 20, "ignore_edge_H": 20,
                     "handle_dynamic": True, "use_color_in_tracking": True, "const_speed_assumption": True},
        "mapping": {"device": device, "every_frame": 5, "no_vis_on_first_frame": True, "fix_fine": True,
                    "BA_cam_lr": 0.001, "fix_color": False, "pixels": 1000, "iters": 60, "color_refine": True,
                    "w_color_loss": 0.2, "keyframe_every": 50, "fine_iter_ratio": 0.6, "middle_iter_ratio": 0.4,
                    "mapping_window_size": 5, "frustum_feature_selection": True,
                    "keyframe_selection_method": "overlap", "save_selected_keyframes_info": False,
                    "stage": stage},
    }


def build_scene(shape="replica_room0", device="cuda:0", seed=0, n_frames=5, cfg=None):
    """Decoders, grids, bound, camera and ``n_frames`` synthetic RGB-D frames of a dataset shape
    (synth.SHAPES).  Returns a SimpleNamespace with the attributes Renderer / Tracker / Mapper
    read from the reference's NICE_SLAM object, plus ``frames`` = [(c2w, depth, color)]."""
    from .utils.Renderer import Renderer
    H, W, fx, fy, cx, cy, bnd = synth.SHAPES[shape]
    if cfg is None:
        cfg = base_cfg(device)
        cfg["grid_len"].update(synth.GRID_LEN.get(shape, {}))
    torch.manual_seed(seed)
    bound = load_bound(bnd, cfg["scale"])
    decoders = conv_config.get_model(cfg, nice=True)
    assign_bound(decoders, bound, cfg["model"]["coarse_bound_enlarge"])
    decoders = decoders.to(device)
    gen = torch.Generator().manual_seed(seed + 1)
    c = grid_init(bound, cfg["grid_len"], cfg["model"]["c_dim"], cfg["coarse"],
                  cfg["model"]["coarse_bound_enlarge"], device, gen)
    room = synth.room_of(bound)
    frames = []
    for c2w in synth.orbit_poses(room, n_frames, seed=seed):
        depth, color = synth.make_frame(H, W, fx, fy, cx, cy, room, c2w, zero_frac=0.05, seed=seed, device=device)
        frames.append((c2w.to(device), depth, color))
    ns = types.SimpleNamespace(nice=True, bound=bound, H=H, W=W, fx=fx, fy=fy, cx=cx, cy=cy, cfg=cfg,
                               shared_c=c, shared_decoders=decoders, frames=frames, shape=shape)
    ns.renderer = Renderer(cfg, None, ns)
    return ns
