#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by RUNNING THE REFERENCE ITSELF.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python tests/golden/make_golden.py

The reference source is imported unmodified.  Two harness-side shims are applied, both
documented in SURVEY.md section 8c: (1) ``Tensor.to`` maps the ``'cuda:-1'`` / ``-1``
device the reference builds for CPU tensors (decoder.py:340-358, common.py:150) to
``'cpu'``; (2) cosmetic / offline-tool modules that are not installed here (colorama,
matplotlib, open3d, trimesh, skimage, mathutils) are stubbed so ``src.Tracker`` and
``src.Mapper`` import.  Nothing from the oracle or the product package computes a value
stored here, except ``synth.make_frame`` which only fabricates the INPUT RGB-D frame.
"""
import os
import sys
import types

import numpy as np
import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NICE_SLAM_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
sys.path.insert(1, REPO)

# ---- shim 1: CPU device strings -----------------------------------------------------
_to = torch.Tensor.to


def _to_cpu_safe(self, *a, **k):
    if a and (isinstance(a[0], str) and a[0] == "cuda:-1" or (type(a[0]) is int and a[0] == -1)):
        a = ("cpu",) + a[1:]
    return _to(self, *a, **k)


torch.Tensor.to = _to_cpu_safe


# ---- shim 2: stub modules -----------------------------------------------------------
class _Blank:
    def __getattr__(self, name):
        return ""


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules.setdefault(name, m)
    return sys.modules[name]


_stub("colorama", Fore=_Blank(), Style=_Blank())
_plt = _stub("matplotlib.pyplot")
_stub("matplotlib", pyplot=_plt)
for _n in ("open3d", "trimesh", "skimage", "mathutils"):
    _stub(_n)


class _Matrix:
    """Stand-in for ``mathutils.Matrix`` (not installed): the one method ``get_tensor_from_camera`` uses
    (common.py:199-202), rotation matrix -> quaternion (w, x, y, z), through scipy.  Only the BA golden
    (``mapper_variants``) reaches it; q and -q describe the same rotation, so the sign convention is immaterial."""

    def __init__(self, R):
        self.R = np.asarray(R, dtype=np.float64)

    def to_quaternion(self):
        from scipy.spatial.transform import Rotation
        x, y, z, w = Rotation.from_matrix(self.R).as_quat()
        return np.array([w, x, y, z])


sys.modules["mathutils"].__dict__.setdefault("Matrix", _Matrix)

from src import config as ref_config                      # noqa: E402
from src.common import (get_camera_from_tensor, get_rays, get_samples,   # noqa: E402
                        raw2outputs_nerf_color, sample_pdf)
from src.conv_onet.models.decoder import MLP as RefMLP    # noqa: E402
from src.Mapper import Mapper                             # noqa: E402
from src.NICE_SLAM import NICE_SLAM                       # noqa: E402
from src.Tracker import Tracker                           # noqa: E402
from src.utils.Renderer import Renderer                   # noqa: E402

from nice_slam_annotation_b200 import synth               # noqa: E402


def load_cfg(nice=True):
    os.chdir(REF)
    with open(os.path.join(REF, "configs/nice_slam.yaml" if nice else "configs/imap.yaml")) as f:
        cfg = yaml.full_load(f)
    os.chdir(REPO)
    return cfg


def build_scene(cfg, cam, bound, nice=True, seed=0, decoders=None):
    """The hot-path part of NICE_SLAM.__init__ (src/NICE_SLAM.py:26-148) without datasets,
    processes, pretrained weights (absent) or mesher."""
    torch.manual_seed(seed)
    cfg["mapping"]["bound"] = bound
    slam = object.__new__(NICE_SLAM)
    slam.cfg, slam.nice = cfg, nice
    slam.coarse, slam.scale = cfg["coarse"], cfg["scale"]
    slam.coarse_bound_enlarge = cfg["model"]["coarse_bound_enlarge"]
    slam.H, slam.W, slam.fx, slam.fy, slam.cx, slam.cy = cam
    slam.shared_decoders = decoders if decoders is not None else ref_config.get_model(cfg, nice=nice)
    NICE_SLAM.load_bound(slam, cfg)
    if nice:
        NICE_SLAM.grid_init(slam, cfg)
    else:
        slam.shared_c = {}
    slam.renderer = Renderer(cfg, None, slam)
    return slam


def np_state(sd):
    return {k: v.detach().cpu().numpy().copy() for k, v in sd.items()}


def flat(prefix, d):
    return {prefix + k: v for k, v in d.items()}


def room_frame(cam, bound_t, margin, pose_idx=0, n_poses=4, seed=0):
    H, W, fx, fy, cx, cy = cam
    room = torch.stack([bound_t[:, 0] + margin, bound_t[:, 1] - margin], 1)
    c2w = synth.orbit_poses(room, n_poses, seed=seed)[pose_idx]
    depth, color = synth.make_frame(H, W, fx, fy, cx, cy, room, c2w, zero_frac=0.05, seed=seed)
    return depth, color, c2w


def render_case(slam, stage, rays_o, rays_d, gt_depth, coef, want_param_grads=True):
    """Reference render_batch_ray forward + backward of a fixed linear functional."""
    c = {k: v.clone().requires_grad_(True) for k, v in slam.shared_c.items()}
    dec = slam.shared_decoders
    dec.zero_grad()
    ro = rays_o.clone().requires_grad_(True)
    rd = rays_d.clone().requires_grad_(True)
    depth, var, color = slam.renderer.render_batch_ray(c, dec, rd, ro, "cpu", stage, gt_depth=gt_depth)
    a, b, w = coef
    loss = (depth * a).sum() + (var * b).sum() + (color * w).sum()
    loss.backward()
    out = {"depth": depth, "var": var, "color": color, "g_rays_o": ro.grad, "g_rays_d": rd.grad}
    for k, v in c.items():
        if v.grad is not None:
            out["g_" + k] = v.grad
    if want_param_grads:
        for k, p in dec.named_parameters():
            if p.grad is not None:
                out["gp_" + k] = p.grad
    return {k: v.detach().numpy().copy() for k, v in out.items()}


# =====================================================================================
def gen_small_nice():
    """F1 on a tiny scene: reference-constructor decoders + grid_init grids, 4 stages."""
    cfg = load_cfg(True)
    cfg["grid_len"].update({"coarse": 0.8, "middle": 0.32, "fine": 0.16, "color": 0.16})
    cam = (48, 64, 60.0, 60.0, 31.5, 23.5)
    bound = [[-0.6, 0.9], [-0.5, 0.7], [-0.4, 0.5]]
    slam = build_scene(cfg, cam, bound, seed=0)
    # grids larger in magnitude than grid_init's so alpha is not saturated everywhere
    g = torch.Generator().manual_seed(5)
    for k, v in slam.shared_c.items():
        v.add_(torch.randn(v.shape, generator=g) * 0.3)
    depth, color, c2w = room_frame(cam, slam.bound, 0.15, seed=1)
    torch.manual_seed(11)
    N = 48
    ro, rd, gd, gc = get_samples(0, cam[0], 0, cam[1], N, *cam, c2w, depth, color, "cpu")
    ro, rd = ro.contiguous(), rd.contiguous()
    # edge cases in the batch: two rays pointing out of the box early, one origin near a wall
    rd[0] = torch.tensor([0.0, 0.0, 1.0])
    rd[1] = torch.tensor([1.0, 0.0, 0.0])
    gd[2] = 0.0
    gg = torch.Generator().manual_seed(3)
    coef = (torch.randn(N, generator=gg).double(), torch.randn(N, generator=gg).double(),
            torch.randn(N, 3, generator=gg))
    out = {"bound": slam.bound.numpy(), "cam": np.array(cam), "c2w": c2w.numpy(),
           "rays_o": ro.numpy(), "rays_d": rd.numpy(), "gt_depth": gd.numpy(), "gt_color": gc.numpy(),
           "coef_a": coef[0].numpy(), "coef_b": coef[1].numpy(), "coef_w": coef[2].numpy()}
    out.update(flat("c.", np_state(slam.shared_c)))
    out.update(flat("dec.", np_state(slam.shared_decoders.state_dict())))
    for stage in ("coarse", "middle", "fine", "color"):
        r = render_case(slam, stage, ro, rd, gd, coef)
        out.update(flat(stage + ".", r))
    # no-gt-depth variant (gt_depth=None at a non-coarse stage)
    r = render_case(slam, "middle", ro, rd, None, coef, want_param_grads=False)
    out.update(flat("middle_nogt.", r))
    # eval_points on explicit points, some outside the bound
    gp = torch.Generator().manual_seed(4)
    pts = (torch.rand(40, 3, generator=gp).double() * 1.3 - 0.15) * (slam.bound[:, 1] - slam.bound[:, 0]) + slam.bound[:, 0]
    pts[0] = slam.bound[:, 0]          # exactly on the lower corner -> out of bound (strict)
    pts[1] = slam.bound[:, 1]
    with torch.no_grad():
        for stage in ("coarse", "middle", "fine", "color"):
            out[f"evalpts.{stage}"] = slam.renderer.eval_points(
                pts, slam.shared_decoders, slam.shared_c, stage, "cpu").numpy()
    out["evalpts.p"] = pts.numpy()
    np.savez_compressed(os.path.join(HERE, "small_nice.npz"), **out)
    print("small_nice", {k: v.shape for k, v in out.items() if k.startswith("color.")})


def gen_sampling_variants():
    """lindisp / perturb / odd sample counts on the sampling + compositing front end."""
    cfg = load_cfg(True)
    cfg["grid_len"].update({"coarse": 0.8})
    cam = (48, 64, 60.0, 60.0, 31.5, 23.5)
    bound = [[-0.6, 0.9], [-0.5, 0.7], [-0.4, 0.5]]
    out = {}
    for name, rend in (("lindisp", dict(lindisp=True)), ("perturb", dict(perturb=1.0)),
                       ("odd", dict(N_samples=20, N_surface=7))):
        cfg2 = load_cfg(True)
        cfg2["grid_len"].update({"coarse": 0.8})
        cfg2["rendering"].update(rend)
        slam = build_scene(cfg2, cam, bound, seed=0)
        g = torch.Generator().manual_seed(5)
        for k, v in slam.shared_c.items():
            v.add_(torch.randn(v.shape, generator=g) * 0.3)
        depth, color, c2w = room_frame(cam, slam.bound, 0.15, seed=1)
        torch.manual_seed(11)
        N = 24
        ro, rd, gd, gc = get_samples(0, cam[0], 0, cam[1], N, *cam, c2w, depth, color, "cpu")
        gd[3] = 0.0
        if name == "lindisp":
            # lindisp with a zero sensor depth is NaN in the reference itself (1/near = inf,
            # inf * 0 at t = 1): not a meaningful case, keep every depth positive here
            gd = torch.where(gd > 0, gd, torch.ones_like(gd))
        if name == "perturb":
            torch.manual_seed(77)
            u = torch.rand(N, cfg2["rendering"]["N_samples"])    # the draw render_batch_ray makes
            out["perturb.u"] = u.numpy()
            torch.manual_seed(77)
        with torch.no_grad():
            d, v, c = slam.renderer.render_batch_ray(slam.shared_c, slam.shared_decoders, rd, ro,
                                                     "cpu", "color", gt_depth=gd)
        out.update({f"{name}.depth": d.numpy(), f"{name}.var": v.numpy(), f"{name}.color": c.numpy(),
                    f"{name}.rays_o": ro.contiguous().numpy(), f"{name}.rays_d": rd.numpy(),
                    f"{name}.gt_depth": gd.numpy()})
    # state is identical for the three (same seeds): store once
    out["bound"] = slam.bound.numpy()
    out.update(flat("c.", np_state(slam.shared_c)))
    out.update(flat("dec.", np_state(slam.shared_decoders.state_dict())))
    np.savez_compressed(os.path.join(HERE, "sampling_variants.npz"), **out)
    print("sampling_variants ok")


def gen_units():
    out = {}
    g = torch.Generator().manual_seed(0)
    # sample_pdf, deterministic and with supplied uniforms (common.py:19-63)
    z = torch.sort(torch.rand(9, 32, generator=g).double() * 3 + 0.1, -1)[0]
    w = torch.rand(9, 32, generator=g)
    w[0] = 0.0                                    # all-zero weights -> uniform pdf
    w[1, 5:] = 0.0
    zm = .5 * (z[..., 1:] + z[..., :-1])
    out["pdf.bins"], out["pdf.weights"] = zm.numpy(), w[..., 1:-1].numpy()
    out["pdf.det"] = sample_pdf(zm, w[..., 1:-1], 12, det=True, device="cpu").numpy()
    torch.manual_seed(5)
    u = torch.rand(9, 12)
    torch.manual_seed(5)
    out["pdf.u"] = u.numpy()
    out["pdf.rand"] = sample_pdf(zm, w[..., 1:-1], 12, det=False, device="cpu").numpy()
    # raw2outputs, both modes (common.py:215-256)
    raw = torch.randn(9, 32, 4, generator=g)
    raw[..., 3] *= 0.3
    raw[2, 10:, 3] = 100.0
    rd = torch.randn(9, 3, generator=g)
    for occ in (True, False):
        d, v, c, ww = raw2outputs_nerf_color(raw.clone(), z, rd, occupancy=occ, device="cpu")
        tag = "occ" if occ else "den"
        out.update({f"comp.{tag}.depth": d.numpy(), f"comp.{tag}.var": v.numpy(),
                    f"comp.{tag}.rgb": c.numpy(), f"comp.{tag}.weights": ww.numpy()})
    out["comp.raw"], out["comp.z"], out["comp.rays_d"] = raw.numpy(), z.numpy(), rd.numpy()
    # camera tensor -> c2w and its gradient (common.py:137-176)
    cam = torch.tensor([0.9, 0.1, -0.3, 0.2, 0.5, -1.0, 2.0], requires_grad=True)
    c2w = get_camera_from_tensor(cam)
    cw = torch.randn(3, 4, generator=g)
    (c2w * cw).sum().backward()
    out.update({"cam.tensor": cam.detach().numpy(), "cam.c2w": c2w.detach().numpy(),
                "cam.cw": cw.numpy(), "cam.grad": cam.grad.numpy()})
    # get_samples with its own randint draw (common.py:92-134) and get_rays (common.py:259-277)
    H, W, fx, fy, cx, cy = 30, 40, 35.0, 36.0, 19.5, 14.5
    depth = torch.rand(H, W, generator=g) * 3
    color = torch.rand(H, W, 3, generator=g)
    c2w0 = synth.look_at([0.1, 0.2, 0.3], [2.0, 1.0, 0.0])
    torch.manual_seed(9)
    idx = torch.randint((H - 8) * (W - 10), (17,))
    torch.manual_seed(9)
    ro, rdd, sd, sc = get_samples(4, H - 4, 5, W - 5, 17, H, W, fx, fy, cx, cy, c2w0, depth, color, "cpu")
    out.update({"gs.depth": depth.numpy(), "gs.color": color.numpy(), "gs.c2w": c2w0.numpy(),
                "gs.cam": np.array([H, W, fx, fy, cx, cy]), "gs.crop": np.array([4, H - 4, 5, W - 5]),
                "gs.indices": idx.numpy(), "gs.rays_o": ro.contiguous().numpy(), "gs.rays_d": rdd.numpy(),
                "gs.sdepth": sd.numpy(), "gs.scolor": sc.numpy()})
    ro_all, rd_all = get_rays(H, W, fx, fy, cx, cy, c2w0, "cpu")
    out["gr.rays_o"], out["gr.rays_d"] = ro_all.contiguous().numpy(), rd_all.numpy()
    np.savez_compressed(os.path.join(HERE, "units.npz"), **out)
    print("units ok")


def gen_imap(hidden=64, name="imap"):
    """iMAP*: single MLP without grids, density compositing, importance resampling,
    regulariser (configs/imap.yaml; Renderer.py:244-259,343-409).  ``imap``: the reference MLP
    class at hidden 64 (small fixture for the oracle); ``imap256``: the factory's own width 256
    (src/conv_onet/config.py:33-37), the case the tensor-core kernels are built for."""
    cfg = load_cfg(False)
    cam = (48, 64, 60.0, 60.0, 31.5, 23.5)
    bound = [[-6.0, 9.0], [-5.0, 7.0], [-4.0, 5.0]]          # metres; cfg scale 0.1 applies
    torch.manual_seed(0)
    dec = RefMLP(name="", dim=3, c_dim=0, color=True, hidden_size=hidden, skips=[], n_blocks=4,
                 pos_embedding_method="fourier")
    with torch.no_grad():
        dec.embedder._B.mul_(0.2)                            # keep densities moderate
        dec.output_linear.bias.add_(torch.tensor([0.3, 0.5, 0.4, 2.0]))
    slam = build_scene(cfg, cam, bound, nice=False, seed=0, decoders=dec)
    depth, color, c2w = room_frame(cam, slam.bound, 0.15, seed=1)
    torch.manual_seed(11)
    N = 32
    ro, rd, gd, gc = get_samples(0, cam[0], 0, cam[1], N, *cam, c2w, depth, color, "cpu")
    ro = ro.contiguous()
    gd[4] = 0.0
    out = {"bound": slam.bound.numpy(), "rays_o": ro.numpy(), "rays_d": rd.numpy(), "gt_depth": gd.numpy()}
    out.update(flat("dec.", np_state(dec.state_dict())))
    gg = torch.Generator().manual_seed(3)
    coef = (torch.randn(N, generator=gg).double(), torch.zeros(N).double(), torch.randn(N, 3, generator=gg))
    out.update({"coef_a": coef[0].numpy(), "coef_w": coef[2].numpy()})
    r = render_case(slam, "color", ro, rd, gd, coef)
    out.update(flat("det.", r))
    # stochastic variant: perturb > 0 draws rand(z.shape) then rand([N, N_importance])
    slam.renderer.perturb = 1.0
    torch.manual_seed(21)
    u1 = torch.rand(N, 32)
    u2 = torch.rand(N, 12)
    out["rand.u"], out["rand.u_importance"] = u1.numpy(), u2.numpy()
    torch.manual_seed(21)
    r = render_case(slam, "color", ro, rd, gd, coef, want_param_grads=False)
    out.update(flat("rand.", r))
    slam.renderer.perturb = 0.0
    # regulariser (always jittered: one rand(z.shape) draw)
    torch.manual_seed(31)
    ur = torch.rand(N, 32)
    torch.manual_seed(31)
    sig = slam.renderer.regulation({}, dec, rd, ro, gd, "cpu", "color")
    out["reg.u"], out["reg.sigma"] = ur.numpy(), sig.detach().numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "ok", float(r["depth"].mean()))


def _make_worker(cls, slam, cfg, **extra):
    w = object.__new__(cls)
    w.cfg, w.args = cfg, None
    for k in ("nice", "bound", "renderer", "H", "W", "fx", "fy", "cx", "cy"):
        setattr(w, k, getattr(slam, k))
    w.scale, w.coarse, w.occupancy = cfg["scale"], cfg["coarse"], cfg["occupancy"]
    w.device, w.verbose, w.output = "cpu", False, "Demo"
    w.__dict__.update(extra)
    return w


def make_mapper(slam, cfg, coarse_mapper=False, frustum=False):
    m = cfg["mapping"]
    return _make_worker(
        Mapper, slam, cfg, coarse_mapper=coarse_mapper, c=slam.shared_c, decoders=slam.shared_decoders,
        fix_fine=m["fix_fine"], fix_color=m["fix_color"], BA=False, BA_cam_lr=m["BA_cam_lr"],
        mapping_pixels=m["pixels"], w_color_loss=m["w_color_loss"], fine_iter_ratio=m["fine_iter_ratio"],
        middle_iter_ratio=m["middle_iter_ratio"], mapping_window_size=m["mapping_window_size"],
        frustum_feature_selection=frustum, keyframe_selection_method="overlap",
        save_selected_keyframes_info=False, keyframe_dict=[], keyframe_list=[],
        no_vis_on_first_frame=True, color_refine=m["color_refine"])


def make_tracker(slam, cfg):
    t = cfg["tracking"]
    return _make_worker(
        Tracker, slam, cfg, c=slam.shared_c, decoders=slam.shared_decoders,
        ignore_edge_W=t["ignore_edge_W"], ignore_edge_H=t["ignore_edge_H"],
        handle_dynamic=t["handle_dynamic"], use_color_in_tracking=t["use_color_in_tracking"],
        w_color_loss=t["w_color_loss"])


def gen_trained():
    """F2 'trained-like' fixture: reference init, then the reference's OWN Mapper.optimize_map
    (first-frame schedule: lr x5) on a synthetic box-room frame; then one reference tracking
    iteration and a short mapper run recorded for step-level parity."""
    cfg = load_cfg(True)
    cfg["grid_len"].update({"coarse": 1.0})
    cfg["mapping"].update({"pixels": 400})
    cfg["tracking"].update({"ignore_edge_W": 4, "ignore_edge_H": 4})
    cam = (60, 80, 70.0, 70.0, 39.5, 29.5)
    bound = [[-1.4, 1.6], [-1.1, 1.3], [-1.0, 1.1]]
    slam = build_scene(cfg, cam, bound, seed=0)
    depth, color, c2w = room_frame(cam, slam.bound, 0.3, seed=2)
    c2w4 = c2w.clone()
    mapper = make_mapper(slam, cfg, frustum=False)
    torch.manual_seed(123)
    iters = int(os.environ.get("GOLDEN_TRAIN_ITERS", "200"))
    mapper.optimize_map(iters, cfg["mapping"]["lr_first_factor"], 0, color, depth, c2w4, [], [], c2w4)
    for k in slam.shared_c:
        slam.shared_c[k] = slam.shared_c[k].detach()
    out = {"bound": slam.bound.numpy(), "cam": np.array(cam), "c2w": c2w4.numpy(),
           "depth_img": depth.numpy(), "color_img": color.numpy()}
    out.update(flat("c.", np_state(slam.shared_c)))
    out.update(flat("dec.", np_state(slam.shared_decoders.state_dict())))

    # (a) render fwd+bwd at stage color on the trained state
    torch.manual_seed(11)
    N = 96
    ro, rd, gd, gc = get_samples(0, cam[0], 0, cam[1], N, *cam, c2w4, depth, color, "cpu")
    ro = ro.contiguous()
    gg = torch.Generator().manual_seed(3)
    coef = (torch.randn(N, generator=gg).double(), torch.randn(N, generator=gg).double() * 0.0,
            torch.randn(N, 3, generator=gg))
    out.update({"r.rays_o": ro.numpy(), "r.rays_d": rd.numpy(), "r.gt_depth": gd.numpy(),
                "r.coef_a": coef[0].numpy(), "r.coef_w": coef[2].numpy()})
    for stage in ("middle", "fine", "color"):
        out.update(flat(f"r.{stage}.", render_case(slam, stage, ro, rd, gd, coef)))

    # (b) one tracking iteration from a perturbed pose (Tracker.optimize_cam_in_batch)
    tracker = make_tracker(slam, cfg)
    # quaternion of c2w via a closed form (mathutils is not installed): use scipy
    from scipy.spatial.transform import Rotation
    q = Rotation.from_matrix(c2w4[:3, :3].double().numpy()).as_quat()       # x,y,z,w
    cam_t = torch.tensor([q[3], q[0], q[1], q[2], *c2w4[:3, 3].tolist()], dtype=torch.float32)
    cam_t = cam_t + torch.tensor([0.01, -0.008, 0.006, 0.004, 0.02, -0.015, 0.01])
    out["t.cam0"] = cam_t.numpy()
    cam_p = cam_t.clone().requires_grad_(True)
    opt = torch.optim.Adam([cam_p], lr=1e-3)
    Hc, Wc = cam[0] - 8, cam[1] - 8
    torch.manual_seed(41)
    out["t.indices"] = torch.randint(Hc * Wc, (150,)).numpy()
    torch.manual_seed(41)

    class _Spy(torch.optim.Adam):
        def step(self_inner, *a, **k):
            out["t.grad"] = cam_p.grad.detach().numpy().copy()
            return super().step(*a, **k)
    opt.__class__ = _Spy
    loss = tracker.optimize_cam_in_batch(cam_p, color, depth, 150, opt)
    out["t.loss"], out["t.cam1"] = np.array(loss), cam_p.detach().numpy().copy()

    # (c) three more mapper iterations from the trained state, dense (no frustum mask)
    torch.manual_seed(51)
    Hf, Wf = cam[0], cam[1]
    out["m.indices"] = np.stack([torch.randint(Hf * Wf, (400,)).numpy() for _ in range(3)])
    torch.manual_seed(51)
    # schedule: with 3 iterations, joint_iter 0,1 -> middle (<= int(3*.4)=1), 2 -> color? (<= int(3*.6)=1 is fine)
    mapper.optimize_map(3, 1.0, 5, color, depth, c2w4, [], [], c2w4)
    out.update(flat("m.c_after.", np_state({k: v.detach() for k, v in slam.shared_c.items()})))
    out.update(flat("m.dec_after.", np_state(slam.shared_decoders.state_dict())))
    np.savez_compressed(os.path.join(HERE, "trained_nice.npz"), **out)
    d = out["r.color.depth"]
    print("trained ok: depth range", d.min(), d.max(), "gt", gd.min().item(), gd.max().item(),
          "L1", np.abs(d - gd.numpy())[gd.numpy() > 0].mean())


def gen_frustum():
    """f2: Mapper.get_mask_from_c2w (cv2.remap inside) and the keyframe scores of
    Mapper.keyframe_selection_overlap, from the reference itself."""
    import src.Mapper as ref_mapper
    cfg = load_cfg(True)
    cfg["grid_len"].update({"coarse": 1.0, "middle": 0.16, "fine": 0.08, "color": 0.08})
    cam = (60, 80, 70.0, 70.0, 39.5, 29.5)
    bound = [[-1.4, 1.6], [-1.1, 1.3], [-1.0, 1.1]]
    slam = build_scene(cfg, cam, bound, seed=0)
    mapper = make_mapper(slam, cfg, frustum=True)
    out = {"bound": slam.bound.numpy(), "cam": np.array(cam)}
    room = torch.stack([slam.bound[:, 0] + 0.3, slam.bound[:, 1] - 0.3], 1)
    poses = synth.orbit_poses(room, 6, seed=4)
    n_mask = 0
    for i in (0, 2, 5):
        c2w = poses[i].float()
        depth, color = synth.make_frame(*cam, room, poses[i], zero_frac=0.05, seed=10 + i)
        out[f"f{i}.c2w"], out[f"f{i}.depth"] = c2w.numpy(), depth.numpy()
        for key in ("grid_coarse", "grid_middle", "grid_fine"):
            val_shape = slam.shared_c[key].shape[2:]
            m = mapper.get_mask_from_c2w(c2w, key, val_shape, depth.numpy())
            out[f"f{i}.{key}.shape"] = np.array(val_shape)
            out[f"f{i}.{key}.mask"] = np.packbits(np.asarray(m).reshape(-1))
            n_mask += 1
            print("frustum", i, key, tuple(val_shape), float(np.asarray(m).mean()))
    # keyframe scores: spy on the sort of list_keyframe (Mapper.py:254)
    seen = {}

    def spy_sorted(lst, **kw):
        seen["scores"] = [d["percent_inside"] for d in lst]
        return sorted(lst, **kw)
    ref_mapper.sorted = spy_sorted
    c2w = poses[0].float()
    depth, color = synth.make_frame(*cam, room, poses[0], zero_frac=0.0, seed=3)
    kfs = [{"est_c2w": p.float()} for p in poses]
    torch.manual_seed(77)
    out["k.indices"] = torch.randint(cam[0] * cam[1], (100,)).numpy()
    torch.manual_seed(77)
    np.random.seed(0)
    sel = mapper.keyframe_selection_overlap(color, depth, c2w, kfs, 3)
    del ref_mapper.sorted
    out["k.c2w"], out["k.depth"], out["k.color"] = c2w.numpy(), depth.numpy(), color.numpy()
    out["k.kf_c2w"] = np.stack([p.float().numpy() for p in poses])
    out["k.scores"] = np.array(seen["scores"], dtype=np.float64)
    out["k.selected"] = np.array(sel)
    print("keyframe scores", out["k.scores"], "selected", sel)
    np.savez_compressed(os.path.join(HERE, "frustum.npz"), **out)


def gen_mapper_variants():
    """f1 pinned to the reference: Mapper.optimize_map from the trained state (trained_nice.npz) with
    ``frustum_feature_selection=True`` (compact val[mask] Adam + write-back, Mapper.py:393-404,491-500,633-642) and
    with local BA (camera tensors in the optimiser, Mapper.py:430-458,644-667), two keyframes + the current frame,
    10 iterations (middle x5, fine x2, color x3).  Stores the pixel draws, the updated grids / decoders and poses."""
    d = dict(np.load(os.path.join(HERE, "trained_nice.npz")))
    cam = tuple(float(v) for v in d["cam"])
    cam = (int(cam[0]), int(cam[1])) + cam[2:]
    bound = [[-1.4, 1.6], [-1.1, 1.3], [-1.0, 1.1]]
    depth, color = torch.from_numpy(d["depth_img"]), torch.from_numpy(d["color_img"])
    c2w = torch.from_numpy(d["c2w"]).float()
    if c2w.shape[0] == 3:
        c2w = torch.cat([c2w, torch.tensor([[0, 0, 0, 1.]])], 0)
    iters, pix = 10, 399
    out = {"iters": np.array(iters), "pixels": np.array(pix)}
    for name, frustum, ba in (("frustum", True, False), ("ba", False, True), ("frustum_ba", True, True)):
        cfg = load_cfg(True)
        cfg["grid_len"].update({"coarse": 1.0})
        cfg["mapping"].update({"pixels": pix})
        slam = build_scene(cfg, cam, bound, seed=0)
        assert np.allclose(slam.bound.numpy(), d["bound"])
        slam.shared_c = {k[2:]: torch.from_numpy(v.copy()) for k, v in d.items() if k.startswith("c.")}
        slam.shared_decoders.load_state_dict({k[4:]: torch.from_numpy(v.copy()) for k, v in d.items() if k.startswith("dec.")})
        mapper = make_mapper(slam, cfg, frustum=frustum)
        mapper.BA = ba
        mapper.keyframe_selection_method = "global"
        kf = []
        for i in range(2):
            pose = c2w.clone()
            pose[:3, 3] += 0.02 * (i + 1)
            kf.append({"gt_c2w": pose.clone(), "idx": i, "color": color, "depth": depth, "est_c2w": pose.clone()})
        mapper.keyframe_dict = kf
        np.random.seed(0)
        torch.manual_seed(61)
        draws = np.stack([torch.randint(cam[0] * cam[1], (pix // 3,)).numpy() for _ in range(3 * iters)])
        torch.manual_seed(61)
        ret = mapper.optimize_map(iters, 1.0, 7, color, depth, c2w, kf, [0, 1], c2w.clone())
        out[f"{name}.indices"] = draws
        for k, v in slam.shared_c.items():
            if k != "grid_coarse":
                out[f"{name}.c_after.{k}"] = v.detach().numpy().copy()
        for k, v in slam.shared_decoders.state_dict().items():
            if k.startswith("color_decoder."):
                out[f"{name}.dec_after.{k}"] = v.detach().numpy().copy()
        out[f"{name}.kf_c2w"] = np.stack([np.asarray(k_["est_c2w"].detach().numpy()) for k_ in kf])
        out[f"{name}.kf_c2w0"] = np.stack([np.asarray(k_["gt_c2w"].detach().numpy()) for k_ in kf])
        if ba:
            out[f"{name}.cur_c2w"] = ret.detach().numpy().copy()
        else:
            assert ret is None
        moved = {k: float((v.detach() != torch.from_numpy(d["c." + k])).double().mean()) for k, v in slam.shared_c.items()}
        print("mapper_variants", name, "moved fraction", moved,
              "pose delta", float(np.abs(out[f"{name}.kf_c2w"] - out[f"{name}.kf_c2w0"]).max()))
    np.savez_compressed(os.path.join(HERE, "mapper_variants.npz"), **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["small", "variants", "units", "imap", "imap256", "trained", "frustum", "mapper_variants"]
    torch.set_num_threads(1)
    if "small" in which:
        gen_small_nice()
    if "variants" in which:
        gen_sampling_variants()
    if "units" in which:
        gen_units()
    if "imap" in which:
        gen_imap()
    if "imap256" in which:
        gen_imap(256, "imap256")
    if "trained" in which:
        gen_trained()
    if "frustum" in which:
        gen_frustum()
    if "mapper_variants" in which:
        gen_mapper_variants()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)) // 1024, "KiB")
