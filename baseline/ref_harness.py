"""Harness that runs the UNMODIFIED reference (WohimLee/NICE-SLAM-annotation) from ``baseline/_ref`` (see
make_ref.py) -- or, in the build container, from /root/reference -- as a baseline and as the other half of the
drop-in seam test.  Test / bench infrastructure only: the product package never imports this file.

What is harness code and what is reference code
* reference: every value-producing call -- ``src.utils.Renderer.Renderer.render_batch_ray``,
  ``src.conv_onet`` decoders built by ``src.config.get_model``, ``src.Tracker.Tracker.optimize_cam_in_batch``,
  ``src.Mapper.Mapper.optimize_map``, ``src.common`` -- runs from the copied files, untouched.
* harness: (1) the two shims SURVEY.md 8c documents (``Tensor.to('cuda:-1')`` -> cpu for CPU runs; stub
  modules for the cosmetic / offline dependencies colorama, matplotlib, open3d, trimesh, skimage, mathutils that
  ``src.Tracker`` / ``src.Mapper`` import at module level); (2) ``Tracker`` / ``Mapper`` objects are made with
  ``object.__new__`` + attribute injection because their ``__init__`` opens a dataset;
  (3) the mapping loss of ``src/Mapper.py:604-612`` is restated in ``mapping_step`` (8 lines that live inline
  in ``optimize_map``).
"""
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
_REF = None


def reference_root():
    """baseline/_ref when it was made, else $NICE_SLAM_REFERENCE / /root/reference, else None."""
    for p in (os.path.join(HERE, "_ref"), os.environ.get("NICE_SLAM_REFERENCE", "/root/reference")):
        if p and os.path.exists(os.path.join(p, "src", "utils", "Renderer.py")):
            return p
    return None


class _Blank:
    def __getattr__(self, name):
        return ""


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    try:
        __import__(name)
        return sys.modules[name]
    except Exception:                                    # noqa: BLE001 - absent or broken: stub it
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m


def load():
    """Import the reference; returns a namespace of its modules (cached).  Raises if it is not available."""
    global _REF
    if _REF is not None:
        return _REF
    root = reference_root()
    if root is None:
        raise FileNotFoundError("the reference is not available: run baseline/make_ref.py in the build container")
    if root not in sys.path:
        sys.path.insert(0, root)
    _to = torch.Tensor.to
    if not getattr(torch.Tensor.to, "_nsb_cpu_safe", False):
        def _to_cpu_safe(self, *a, **k):                 # decoder.py:340-358, common.py:150 on CPU tensors
            if a and (isinstance(a[0], str) and a[0] == "cuda:-1" or (type(a[0]) is int and a[0] == -1)):
                a = ("cpu",) + a[1:]
            return _to(self, *a, **k)
        _to_cpu_safe._nsb_cpu_safe = True
        torch.Tensor.to = _to_cpu_safe
    _stub("colorama", Fore=_Blank(), Style=_Blank())
    plt = _stub("matplotlib.pyplot")
    _stub("matplotlib", pyplot=plt)
    for n in ("open3d", "trimesh", "skimage", "mathutils"):
        _stub(n)
    from src import common, config                       # noqa: E402
    from src.Mapper import Mapper
    from src.NICE_SLAM import NICE_SLAM
    from src.Tracker import Tracker
    from src.utils.Renderer import Renderer
    _REF = types.SimpleNamespace(root=root, common=common, config=config, Mapper=Mapper, Tracker=Tracker,
                                 NICE_SLAM=NICE_SLAM, Renderer=Renderer)
    return _REF


def available():
    return reference_root() is not None


def reference_scene(ns, device, renderer=None, decoders=None):
    """The reference's scene objects holding the SAME state as a product scene ``ns`` (scene.build_scene):
    decoders from ``src.config.get_model`` loaded with the product's state_dict (the keys are identical), the
    grids as plain contiguous [1,32,Z,Y,X] tensors (``NICE_SLAM.grid_init`` layout), the reference Renderer.
    ``renderer`` / ``decoders``: pass the product's objects instead to build the drop-in seam."""
    R = load()
    cfg = ns.cfg
    slam = types.SimpleNamespace(nice=ns.nice, bound=ns.bound.clone(), H=ns.H, W=ns.W, fx=ns.fx, fy=ns.fy, cx=ns.cx,
                                 cy=ns.cy, cfg=cfg)
    if decoders is None:
        cfg_m = dict(cfg)
        cfg_m.setdefault("data", {"dim": 3})
        decoders = R.config.get_model(cfg_m, nice=ns.nice)
        decoders.load_state_dict({k: v.detach().cpu().clone() for k, v in ns.shared_decoders.state_dict().items()})
        decoders = decoders.to(device)
        if ns.nice:
            decoders.bound = slam.bound
            for n in ("middle_decoder", "fine_decoder", "color_decoder"):
                getattr(decoders, n).bound = slam.bound
            if cfg.get("coarse", True):
                decoders.coarse_decoder.bound = slam.bound * cfg["model"]["coarse_bound_enlarge"]
        else:
            decoders.bound = slam.bound
    slam.shared_decoders = decoders
    slam.shared_c = {k: v.detach().to(device).contiguous().clone() for k, v in ns.shared_c.items()}
    slam.renderer = renderer if renderer is not None else R.Renderer(cfg, None, slam)
    return slam


def _worker(cls, slam, cfg, device, **extra):
    w = object.__new__(cls)
    w.cfg, w.args = cfg, None
    for k in ("nice", "bound", "renderer", "H", "W", "fx", "fy", "cx", "cy"):
        setattr(w, k, getattr(slam, k))
    w.scale, w.coarse, w.occupancy = cfg["scale"], cfg["coarse"], cfg["occupancy"]
    w.device, w.verbose, w.output = device, False, "Demo"
    w.__dict__.update(extra)
    return w


def make_tracker(slam, cfg, device):
    """``src.Tracker.Tracker`` without its dataset plumbing (attribute list = what Tracker.py:78-135 reads)."""
    t = cfg["tracking"]
    return _worker(load().Tracker, slam, cfg, device, c=slam.shared_c, decoders=slam.shared_decoders,
                   ignore_edge_W=t["ignore_edge_W"], ignore_edge_H=t["ignore_edge_H"], handle_dynamic=t["handle_dynamic"],
                   use_color_in_tracking=t["use_color_in_tracking"], w_color_loss=t["w_color_loss"])


def make_mapper(slam, cfg, device, frustum=False, ba=False, coarse_mapper=False):
    """``src.Mapper.Mapper`` without its dataset / visualiser plumbing (what Mapper.py:295-669 reads)."""
    m = cfg["mapping"]
    return _worker(load().Mapper, slam, cfg, device, coarse_mapper=coarse_mapper, c=slam.shared_c,
                   decoders=slam.shared_decoders, fix_fine=m["fix_fine"], fix_color=m["fix_color"], BA=ba,
                   BA_cam_lr=m["BA_cam_lr"], mapping_pixels=m["pixels"], w_color_loss=m["w_color_loss"],
                   fine_iter_ratio=m["fine_iter_ratio"], middle_iter_ratio=m["middle_iter_ratio"],
                   mapping_window_size=m["mapping_window_size"], frustum_feature_selection=frustum,
                   keyframe_selection_method=m.get("keyframe_selection_method", "overlap"),
                   save_selected_keyframes_info=False, keyframe_dict=[], keyframe_list=[],
                   no_vis_on_first_frame=True, color_refine=m["color_refine"])


def mapping_step(slam, device, stage="color", w_color=0.2):
    """fn(rays_o, rays_d, gt_depth, gt_color) -> loss: one iteration body of ``Mapper.optimize_map``
    (src/Mapper.py:598-624 without the optimiser): reference ``render_batch_ray`` forward, the mapping loss,
    backward to the three grids and the colour decoder."""
    dec, c = slam.shared_decoders, slam.shared_c
    for p in dec.parameters():
        p.requires_grad_(False)
    for p in dec.color_decoder.parameters():
        p.requires_grad_(True)
    for k, v in c.items():
        v.requires_grad_(k != "grid_coarse")
    leaves = [v for k, v in c.items() if k != "grid_coarse"] + list(dec.color_decoder.parameters())

    def step(ro, rd, gd, gc):
        depth, _, color = slam.renderer.render_batch_ray(c, dec, rd, ro, device, stage, gt_depth=gd)
        depth_mask = gd > 0                                            # Mapper.py:604-612
        loss = torch.abs(gd[depth_mask] - depth[depth_mask]).sum()
        if stage == "color":
            loss = loss + w_color * torch.abs(gc - color).sum()
        loss.backward()
        for t in leaves:
            t.grad = None
        return loss
    return step
