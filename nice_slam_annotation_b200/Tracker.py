"""Tracker step with the reference's interface (src/Tracker.py): ``optimize_cam_in_batch``
is the per-iteration hot path (Tracker.py:78-135) and runs on the fused sm_100a kernels.

The frame loop ``run()`` (Tracker.py:151-299: dataset reader, process synchronisation with
the mapper, visualiser) is a caller of this path and stays with the reference; only the
state it hands to the step is mirrored here (same attribute names, read from cfg / slam).
"""
import copy
import ctypes as C

import torch

from . import fused
from ._lib import check, lib
from .fused import _bound6, _ptr, _stream


class _CameraRaysFn(torch.autograd.Function):
    """get_camera_from_tensor + get_samples + the box pre-filter (common.py:163-176,92-134,
    Tracker.py:100-107) as ONE kernel; backward is the fused pose gradient
    (d(rays_o), d(rays_d)) -> d(camera_tensor[7]) (csrc/nsb_composite.cu:k_pose_grad)."""

    @staticmethod
    def forward(ctx, cam7, indices, crop, intr, depth, color, bound):
        H0, H1, W0, W1 = crop
        fx, fy, cx, cy = intr
        dev = depth.device
        n = indices.shape[0]
        cam = cam7.detach().float().contiguous()
        ro = torch.empty(n, 3, dtype=torch.float32, device=dev)
        rd = torch.empty(n, 3, dtype=torch.float32, device=dev)
        gd = torch.empty(n, dtype=torch.float32, device=dev)
        gc = torch.empty(n, 3, dtype=torch.float32, device=dev)
        keep = torch.empty(n, dtype=torch.uint8, device=dev)
        check(lib().nsb_sample_rays(_ptr(indices), n, H0, H1, W0, W1, depth.shape[1], fx, fy, cx, cy, None,
                                    _ptr(cam), _ptr(depth), _ptr(color), int(color.dtype == torch.float64),
                                    None if bound is None else _bound6(bound), _ptr(ro), _ptr(rd), _ptr(gd),
                                    _ptr(gc), _ptr(keep), _stream()), "sample_rays")
        ctx.save_for_backward(indices, cam)
        ctx.geom = (crop, intr)
        ctx.mark_non_differentiable(gd, gc, keep)
        return ro, rd, gd, gc, keep

    @staticmethod
    def backward(ctx, g_o, g_d, _a, _b, _c):
        indices, cam = ctx.saved_tensors
        (H0, H1, W0, W1), (fx, fy, cx, cy) = ctx.geom
        n = indices.shape[0]
        dev = cam.device
        g_o = torch.zeros(n, 3, device=dev) if g_o is None else g_o.float().contiguous()
        g_d = torch.zeros(n, 3, device=dev) if g_d is None else g_d.float().contiguous()
        d_cam = torch.empty(7, dtype=torch.float32, device=dev)
        check(lib().nsb_pose_grad(_ptr(g_o), _ptr(g_d), _ptr(indices), None, n, H0, W0, W1 - W0, fx, fy, cx, cy,
                                  _ptr(cam), _ptr(d_cam), _stream()), "pose_grad")
        return d_cam, None, None, None, None, None, None


class Tracker(object):
    def __init__(self, cfg, args, slam):
        # Tracker.py:19-76 without the dataset reader / visualiser (callers of the step)
        self.cfg = cfg
        self.args = args
        self.scale = cfg['scale']
        self.coarse = cfg['coarse']
        self.occupancy = cfg['occupancy']
        self.sync_method = cfg.get('sync_method', 'strict')
        for name in ('idx', 'mesher', 'output', 'verbose', 'shared_c', 'gt_c2w_list', 'low_gpu_mem',
                     'mapping_idx', 'mapping_cnt', 'shared_decoders', 'estimate_c2w_list'):
            setattr(self, name, getattr(slam, name, None))
        self.nice = slam.nice
        self.bound = slam.bound
        self.renderer = slam.renderer
        t = cfg['tracking']
        self.cam_lr = t['lr']
        self.device = t['device']
        self.num_cam_iters = t['iters']
        self.gt_camera = t['gt_camera']
        self.tracking_pixels = t['pixels']
        self.seperate_LR = t['seperate_LR']
        self.w_color_loss = t['w_color_loss']
        self.ignore_edge_W = t['ignore_edge_W']
        self.ignore_edge_H = t['ignore_edge_H']
        self.handle_dynamic = t['handle_dynamic']
        self.use_color_in_tracking = t['use_color_in_tracking']
        self.const_speed_assumption = t['const_speed_assumption']
        self.every_frame = cfg['mapping']['every_frame']
        self.no_vis_on_first_frame = cfg['mapping']['no_vis_on_first_frame']
        self.prev_mapping_idx = -1
        self.c = {}
        self.decoders = None
        self._fused_loop = None
        self._api_step = None            # engine behind optimize_cam_in_batch (graph without the camera optimiser)
        self._api_frame = None           # (storage, version) of the frame loaded into it
        self.H, self.W, self.fx, self.fy, self.cx, self.cy = slam.H, slam.W, slam.fx, slam.fy, slam.cx, slam.cy

    def optimize_cam_in_batch(self, camera_tensor, gt_color, gt_depth, batch_size, optimizer, indices=None):
        """One camera iteration: sample pixels, render depth/colour, loss, backward to the 7-dof
        camera tensor, optimiser step (Tracker.py:78-135).  Returns the loss value.

        ``indices`` optionally replaces the torch.randint pixel draw (common.py:99)."""
        device = self.device
        H, W, fx, fy, cx, cy = self.H, self.W, self.fx, self.fy, self.cx, self.cy
        optimizer.zero_grad()
        if self.nice and self.renderer.fixed_shape_ok() and camera_tensor.is_cuda and camera_tensor.requires_grad:
            return self._optimize_cam_in_batch_graph(camera_tensor, gt_color, gt_depth, batch_size, optimizer, indices)
        Wedge, Hedge = self.ignore_edge_W, self.ignore_edge_H
        H0, H1, W0, W1 = Hedge, H - Hedge, Wedge, W - Wedge
        gt_depth = gt_depth.to(device)
        gt_color = gt_color.to(device)
        if gt_depth.dtype != torch.float32:
            gt_depth = gt_depth.float()
        if gt_color.dtype not in (torch.float32, torch.float64):
            gt_color = gt_color.float()
        gt_depth, gt_color = gt_depth.contiguous(), gt_color.contiguous()
        if indices is None:
            indices = torch.randint((H1 - H0) * (W1 - W0), (batch_size,), device=gt_depth.device)
        indices = indices.to(gt_depth.device).contiguous()
        rays_o, rays_d, batch_gt_depth, batch_gt_color, keep = _CameraRaysFn.apply(
            camera_tensor, indices, (H0, H1, W0, W1), (float(fx), float(fy), float(cx), float(cy)),
            gt_depth, gt_color, self.bound if self.nice else None)
        # Tracker.py:100-111 drops the rays that leave the scene box before their sensor depth by
        # boolean indexing (a device->host sync per iteration); the fused path keeps the batch
        # shape and hands the mask to the kernels instead: dropped rays are excluded from the
        # batch depth max, the median and the loss, so every value is the same.
        if not self.nice:
            keep = None
        elif not self.renderer.fixed_shape_ok():
            inside_mask = keep.bool()
            rays_d, rays_o = rays_d[inside_mask], rays_o[inside_mask]
            batch_gt_depth, batch_gt_color = batch_gt_depth[inside_mask], batch_gt_color[inside_mask]
            keep = None
        depth, uncertainty, color = self.renderer.render_batch_ray(
            self.c, self.decoders, rays_d, rays_o, self.device, stage='color', gt_depth=batch_gt_depth, keep=keep)
        loss = fused.tracking_loss(depth, uncertainty, color, batch_gt_depth, batch_gt_color,
                                   self.w_color_loss, self.use_color_in_tracking, self.handle_dynamic, keep=keep)
        loss.backward()
        optimizer.step()
        optimizer.zero_grad()
        return loss.item()

    def _optimize_cam_in_batch_graph(self, camera_tensor, gt_color, gt_depth, batch_size, optimizer, indices):
        """optimize_cam_in_batch on the engine: the iteration up to d(loss)/d(camera tensor) -- pixel draw, camera
        tensor to rays + box pre-filter, render, tracking loss, backward to the rays, pose gradient -- is ONE replayed
        CUDA graph (step.TrackingStep without its own Adam).  What stays on the host is what the reference's caller
        owns: the hand-over of the gradient to autograd (``camera_tensor`` may be ``cat([quad, T])`` of two leaves,
        Tracker.py:226-231), ``optimizer.step()`` and the ``loss.item()`` read.  Same values as the autograd path."""
        from .step import TrackingStep
        ts = self._api_step
        if ts is None or ts.N != batch_size:
            ts = self._api_step = TrackingStep(self, batch_size, adam=False)
            self._api_frame = None
        # the reference passes the same frame for all its iterations (Tracker.py:244-258): load it once.  The key holds
        # the tensors themselves, so their storage cannot be recycled for another frame while the key is compared.
        frame = (gt_color.data_ptr(), gt_color._version, gt_depth.data_ptr(), gt_depth._version,
                 tuple(gt_color.shape), tuple(gt_depth.shape))
        if self._api_frame is None or frame != self._api_frame[0]:
            ts.load_frame(gt_color.to(self.device), gt_depth.to(self.device))
            self._api_frame = (frame, gt_color, gt_depth)
        ts.cam.copy_(camera_tensor.detach().reshape(7), non_blocking=True)
        draw = indices is None
        if not draw:
            ts.indices.copy_(indices.reshape(-1), non_blocking=True)
        ts.iterate_api(draw)
        camera_tensor.backward(ts.d_cam)                # autograd carries it to the leaves the optimiser owns
        optimizer.step()
        optimizer.zero_grad()
        return ts.step.loss.item()

    def optimize_cam_fused(self, camera_tensor, gt_color, gt_depth, batch_size, num_iters=None, graph=True):
        """All camera iterations of one frame (the loop of Tracker.py:244-275 around
        optimize_cam_in_batch) without a host round trip per iteration: pixel draw, render,
        loss, pose gradient, Adam and the best-candidate rule run on the device
        (step.TrackingStep, one CUDA graph replayed ``num_iters`` times).  Returns
        (candidate_cam_tensor [7], its loss)."""
        from .step import TrackingStep
        if not (self.nice and self.renderer.fixed_shape_ok()):
            raise NotImplementedError("the fused tracking loop needs the NICE decoders, perturb == 0, N_importance == 0")
        ts = self._fused_loop
        if ts is None or ts.N != batch_size:
            ts = self._fused_loop = TrackingStep(self, batch_size)
        ts.load_frame(gt_color.to(self.device), gt_depth.to(self.device))
        ts.reset(camera_tensor)
        if graph and ts.graph is None:
            ts.iterate(1)                       # warm-up launch outside the capture
            ts.reset(camera_tensor)
            ts.capture()
        ts.iterate(self.num_cam_iters if num_iters is None else num_iters)
        return ts.result()

    def update_para_from_mapping(self):
        """Copy decoders / grids from the mapper's shared state (Tracker.py:137-149).  The copies
        are detached and parameter gradients are switched off: tracking only needs d(camera)
        (the reference computes, and discards, the decoder gradients here)."""
        if self.mapping_idx is None or self.mapping_idx[0] != self.prev_mapping_idx:
            self.decoders = copy.deepcopy(self.shared_decoders).to(self.device)
            self._fused_loop = None             # its packed weights / grid pointers are stale
            self._api_step, self._api_frame = None, None
            for p in self.decoders.parameters():
                p.requires_grad_(False)
            for key, val in self.shared_c.items():
                self.c[key] = val.detach().clone().to(self.device)
            if self.mapping_idx is not None:
                self.prev_mapping_idx = self.mapping_idx[0].clone()

    def run(self):
        raise NotImplementedError(
            "Tracker.run (dataset loop, mapper synchronisation, visualiser: src/Tracker.py:151-299) is "
            "outside the render hot path; drive optimize_cam_in_batch from the reference's loop")
