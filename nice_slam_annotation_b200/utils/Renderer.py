"""Renderer with the reference's interface (src/utils/Renderer.py), backed by the fused
sm_100a kernels.  Drop-in for ``src.utils.Renderer.Renderer``: same constructor, same four
methods, same return shapes and dtypes (depth / uncertainty fp64, colour fp32)."""
import torch

from .. import fused, imap
from ..common import get_rays


class Renderer(object):
    def __init__(self, cfg, args, slam, points_batch_size=500000, ray_batch_size=100000):
        # Renderer.py:6-34 (kept verbatim: the object is pickled into the worker processes, so
        # it holds plain Python values only)
        self.ray_batch_size = ray_batch_size
        self.points_batch_size = points_batch_size
        self.lindisp = cfg['rendering']['lindisp']
        self.perturb = cfg['rendering']['perturb']
        self.N_samples = cfg['rendering']['N_samples']
        self.N_surface = cfg['rendering']['N_surface']
        self.N_importance = cfg['rendering']['N_importance']
        self.scale = cfg['scale']
        self.occupancy = cfg['occupancy']
        self.nice = slam.nice
        self.bound = slam.bound
        self.coarse_bound_enlarge = cfg.get('model', {}).get('coarse_bound_enlarge', 2)
        self.H, self.W, self.fx, self.fy, self.cx, self.cy = slam.H, slam.W, slam.fx, slam.fy, slam.cx, slam.cy

    def eval_points(self, p, decoders, c=None, stage='color', device='cuda:0'):
        """Occupancy (and colour) of points p [N,3] -> raw [N,4] fp32 (Renderer.py:36-93).
        Differentiable when any input requires grad."""
        p = p.to(device)
        rets = []
        for pi in torch.split(p, self.points_batch_size):      # Renderer.py:51
            if not self.nice:                                  # Renderer.py:73-74: decoders(pi, c_grid=None)
                rets.append(imap.eval_points(pi, decoders, self.bound))
                continue
            needs_grad = torch.is_grad_enabled() and (
                pi.requires_grad or any(g.requires_grad for g in c.values())
                or any(q.requires_grad for q in decoders.parameters()))
            if needs_grad:
                rets.append(fused.eval_points_autograd(pi, decoders, c, stage, self.bound, self.coarse_bound_enlarge))
            else:
                rets.append(fused.eval_points(pi, decoders, c, stage, self.bound, self.coarse_bound_enlarge))
        return torch.cat(rets, dim=0) if len(rets) != 1 else rets[0]

    def fixed_shape_ok(self):
        """True when a pre-filter mask can replace batch compaction without changing any value:
        no jitter draws (their count depends on the batch size) and a single sampling pass."""
        return bool(self.nice) and self.perturb == 0. and self.N_importance == 0

    def render_batch_ray(self, c, decoders, rays_d, rays_o, device, stage, gt_depth=None, keep=None):
        """Depth, uncertainty and colour of a batch of rays (Renderer.py:95-262).  ``keep``
        (extension, [N] uint8): rays a caller would have dropped by the box pre-filter; they are
        rendered but excluded from the batch depth max (see fixed_shape_ok)."""
        N = rays_o.shape[0]
        u = u_imp = None
        if self.perturb > 0.:
            # same draws, in the same order, from the same (CPU) generator as Renderer.py:218
            # and common.py:36
            n_first = self.N_samples
            u = torch.rand(N, n_first)
            if self.N_importance > 0:
                u_imp = torch.rand(N, self.N_importance)
        if not self.nice:
            if keep is not None:
                raise NotImplementedError("keep masks are supported on the single-pass (NICE) render only")
            return imap.render_batch_ray(
                decoders, rays_d.to(device), rays_o.to(device), None if gt_depth is None else gt_depth.to(device),
                self.bound, N_samples=self.N_samples, N_surface=self.N_surface, N_importance=self.N_importance,
                lindisp=self.lindisp, occupancy=self.occupancy, u=u, u_importance=u_imp)
        return fused.render_batch_ray(
            c, decoders, rays_d.to(device), rays_o.to(device), stage,
            None if gt_depth is None else gt_depth.to(device), self.bound,
            N_samples=self.N_samples, N_surface=self.N_surface, N_importance=self.N_importance,
            lindisp=self.lindisp, occupancy=self.occupancy, u=u, u_importance=u_imp,
            coarse_bound_enlarge=self.coarse_bound_enlarge, keep=keep)

    def render_img(self, c, decoders, c2w, device, stage, gt_depth=None):
        """Full-image rendering in ray batches (Renderer.py:264-340): depth [H,W] f64,
        uncertainty [H,W] f64, colour [H,W,3]."""
        with torch.no_grad():
            H, W = self.H, self.W
            rays_o, rays_d = get_rays(H, W, self.fx, self.fy, self.cx, self.cy, c2w, device)
            rays_o = rays_o.reshape(-1, 3)
            rays_d = rays_d.reshape(-1, 3)
            depth_list, uncertainty_list, color_list = [], [], []
            ray_batch_size = self.ray_batch_size
            if gt_depth is not None:
                gt_depth = gt_depth.reshape(-1)
            for i in range(0, rays_d.shape[0], ray_batch_size):
                rays_d_batch = rays_d[i:i + ray_batch_size]
                rays_o_batch = rays_o[i:i + ray_batch_size]
                gt_depth_batch = None if gt_depth is None else gt_depth[i:i + ray_batch_size]
                depth, uncertainty, color = self.render_batch_ray(
                    c, decoders, rays_d_batch, rays_o_batch, device, stage, gt_depth=gt_depth_batch)
                depth_list.append(depth.double())
                uncertainty_list.append(uncertainty.double())
                color_list.append(color)
            depth = torch.cat(depth_list, dim=0).reshape(H, W)
            uncertainty = torch.cat(uncertainty_list, dim=0).reshape(H, W)
            color = torch.cat(color_list, dim=0).reshape(H, W, 3)
            return depth, uncertainty, color

    def regulation(self, c, decoders, rays_d, rays_o, gt_depth, device, stage='color'):
        """iMAP free-space regulariser (Renderer.py:343-409): densities at N_samples jittered samples in
        [0, 0.85 d] of every ray -> [N * N_samples].  Depths and points come from one kernel
        (csrc/nsb_sample.cu:k_regulation_points); the uniforms are drawn where the reference draws them
        (CPU generator, Renderer.py:385), so a seeded run sees the same samples."""
        n = rays_o.shape[0]
        u = torch.rand(n, self.N_samples).to(device, non_blocking=True)
        pts = fused.regulation_points(rays_o.to(device), rays_d.to(device), gt_depth.to(device).reshape(-1),
                                      self.N_samples, u)
        return self.eval_points(pts, decoders, c, stage, device)[:, -1]
