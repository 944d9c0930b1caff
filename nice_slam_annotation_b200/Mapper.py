"""Mapper step with the reference's interface (src/Mapper.py): ``optimize_map`` is the
per-mapped-frame optimisation whose iteration body (Mapper.py:488-642) is the render hot path
and runs on the fused sm_100a kernels.

Grids are kept channels-last in memory (one 128-byte line per voxel, the kernels' layout)
while keeping the reference's logical shape [1,32,Z,Y,X]; everything the reference does with
them (boolean-mask indexing, Adam, checkpointing) works unchanged on such tensors.

``get_mask_from_c2w`` and ``keyframe_selection_overlap`` (Mapper.py:100-293) are callers of
the path that run once per mapped frame: both are kernels
(csrc/nsb_frustum.cu), bit-exact with the reference's numpy + cv2.remap pass.  The frame loop
``run()`` stays with the reference.
"""
import numpy as np
import torch

from . import fused
from ._lib import check, lib
from .common import get_camera_from_tensor, get_samples, get_tensor_from_camera, random_select


class Mapper(object):
    def __init__(self, cfg, args, slam, coarse_mapper=False):
        # Mapper.py:23-98 without dataset reader / visualiser / logger plumbing
        self.cfg = cfg
        self.args = args
        self.coarse_mapper = coarse_mapper
        for name in ('idx', 'logger', 'mesher', 'output', 'verbose', 'low_gpu_mem', 'mapping_idx',
                     'mapping_cnt', 'estimate_c2w_list', 'mapping_first_frame'):
            setattr(self, name, getattr(slam, name, None))
        self.nice = slam.nice
        self.c = slam.shared_c
        self.bound = slam.bound
        self.renderer = slam.renderer
        self.decoders = slam.shared_decoders
        self.scale = cfg['scale']
        self.coarse = cfg['coarse']
        self.occupancy = cfg['occupancy']
        self.sync_method = cfg.get('sync_method', 'strict')
        m = cfg['mapping']
        self.device = m['device']
        self.fix_fine = m['fix_fine']
        self.BA = False
        self.BA_cam_lr = m['BA_cam_lr']
        self.fix_color = m['fix_color']
        self.mapping_pixels = m['pixels']
        self.num_joint_iters = m['iters']
        self.every_frame = m['every_frame']
        self.color_refine = m['color_refine']
        self.w_color_loss = m['w_color_loss']
        self.keyframe_every = m['keyframe_every']
        self.fine_iter_ratio = m['fine_iter_ratio']
        self.middle_iter_ratio = m['middle_iter_ratio']
        self.mapping_window_size = m['mapping_window_size']
        self.no_vis_on_first_frame = m['no_vis_on_first_frame']
        self.frustum_feature_selection = m['frustum_feature_selection']
        self.keyframe_selection_method = m['keyframe_selection_method']
        self.save_selected_keyframes_info = m['save_selected_keyframes_info']
        if self.save_selected_keyframes_info:
            self.selected_keyframes = {}
        if self.nice and coarse_mapper:
            self.keyframe_selection_method = 'global'
        self.keyframe_dict = []
        self.keyframe_list = []
        # optional iterator of pixel-index tensors replacing the torch.randint draws of
        # get_samples (common.py:99), one per frame per iteration: lets a test share the draws
        self.index_source = None
        # True: the iterations of optimize_map run as the device-resident loop of map_loop.py
        # whenever its preconditions hold; False: the autograd path below (same values)
        self.fused_loop = True
        self._loops = {}
        # shared.prepare_shared(slam) ran before the objects were shared with the other processes: the loop
        # must then never re-allocate a grid or a decoder parameter (the tracker would keep reading the old storage)
        self.shared_prepared = bool(getattr(slam, 'shared_layout_prepared', False))
        # ray-sharded mapping over the GPUs of one box: see enable_sharding
        self.shard_group, self.shard_peers, self.shard_world, self.shard_rank = None, None, 1, 0
        self.H, self.W, self.fx, self.fy, self.cx, self.cy = slam.H, slam.W, slam.fx, slam.fy, slam.cx, slam.cy

    # ---- callers of the hot path, once per mapped frame (kernels: csrc/nsb_frustum.cu) -------
    @staticmethod
    def _w2c_host(c2w):
        """np.linalg.inv of the float32 pose on the host, as the reference does (Mapper.py:121-122,
        225-226): 16 floats of scalar prep, so the kernels see bit-identical matrices."""
        m = np.asarray(c2w.detach().cpu().numpy() if torch.is_tensor(c2w) else c2w, dtype=np.float32)
        if m.shape[-2] == 3:
            pad = np.broadcast_to(np.array([[0, 0, 0, 1]], dtype=np.float32), m.shape[:-2] + (1, 4))
            m = np.concatenate([m, pad], -2)
        return m, np.linalg.inv(m)

    def get_mask_from_c2w(self, c2w, key, val_shape, depth_np):
        """Frustum feature selection (Mapper.py:100-187): boolean numpy mask [X,Y,Z] of the voxels
        of grid ``key`` seen by the current frame.  ``depth_np`` may be a numpy array or a tensor."""
        return self._frustum_mask(c2w, key, val_shape, depth_np, order='xyz').cpu().numpy()

    def _frustum_mask(self, c2w, key, val_shape, depth_np, order='zyx'):
        """get_mask_from_c2w as a device tensor (nsb_frustum_mask, bit-exact with the reference's
        numpy + cv2.remap pass): bool [X,Y,Z] (order='xyz', the reference's array) or [Z,Y,X]
        (order='zyx', one flag per voxel line of the channels-last grid)."""
        dev = self.device
        nz, ny, nx = (int(v) for v in val_shape)
        if key == 'grid_coarse':
            return torch.ones((nx, ny, nz) if order == 'xyz' else (nz, ny, nx), dtype=torch.bool, device=dev)
        b = self.bound
        xs = torch.linspace(float(b[0][0]), float(b[0][1]), nx).to(dev)
        ys = torch.linspace(float(b[1][0]), float(b[1][1]), ny).to(dev)
        zs = torch.linspace(float(b[2][0]), float(b[2][1]), nz).to(dev)
        depth = torch.as_tensor(depth_np, dtype=torch.float32).to(dev).contiguous()
        m, w2c = self._w2c_host(c2w)
        w2c = np.ascontiguousarray(w2c, dtype=np.float32)
        cam = np.ascontiguousarray(m[:3, 3], dtype=np.float32)
        out = torch.empty((nx, ny, nz) if order == 'xyz' else (nz, ny, nx), dtype=torch.uint8, device=dev)
        scratch = torch.empty(1, dtype=torch.int32, device=dev)
        check(lib().nsb_frustum_mask(w2c.ctypes.data, cam.ctypes.data, fused._ptr(xs), fused._ptr(ys), fused._ptr(zs),
                                     nx, ny, nz, fused._ptr(depth), int(self.H), int(self.W), float(self.fx),
                                     float(self.fy), float(self.cx), float(self.cy), fused._ptr(scratch),
                                     fused._ptr(out) if order == 'xyz' else None,
                                     fused._ptr(out) if order != 'xyz' else None, fused._stream()), "frustum_mask")
        return out.bool()

    def keyframe_overlap_scores(self, rays_o, rays_d, gt_depth, keyframe_c2ws, N_samples=16, edge=20):
        """percent_inside of every keyframe (Mapper.py:224-252) as one kernel launch
        (nsb_keyframe_overlap): float64 numpy [K]."""
        dev = self.device
        K = len(keyframe_c2ws)
        if K == 0:
            return np.zeros(0)
        _, w2c = self._w2c_host(torch.stack([torch.as_tensor(c).detach().cpu().float() for c in keyframe_c2ws]))
        w2c = torch.from_numpy(np.ascontiguousarray(w2c, dtype=np.float32)).to(dev)
        ro, rd = rays_o.detach().float().contiguous(), rays_d.detach().float().contiguous()
        gd = gt_depth.detach().float().reshape(-1).contiguous()
        t_vals = torch.linspace(0., 1., steps=N_samples).to(dev)
        counts = torch.zeros(K, dtype=torch.int32, device=dev)
        check(lib().nsb_keyframe_overlap(fused._ptr(ro), fused._ptr(rd), fused._ptr(gd), ro.shape[0], N_samples,
                                         fused._ptr(t_vals), fused._ptr(w2c), K, int(self.H), int(self.W),
                                         float(self.fx), float(self.fy), float(self.cx), float(self.cy), int(edge),
                                         fused._ptr(counts), fused._stream()), "keyframe_overlap")
        return counts.cpu().numpy() / float(ro.shape[0] * N_samples)

    def keyframe_selection_overlap(self, gt_color, gt_depth, c2w, keyframe_dict, k, N_samples=16, pixels=100):
        """Keyframes whose view overlaps the current frame (Mapper.py:189-293)."""
        device = self.device
        H, W, fx, fy, cx, cy = self.H, self.W, self.fx, self.fy, self.cx, self.cy
        rays_o, rays_d, gt_depth, gt_color = get_samples(
            0, H, 0, W, pixels, H, W, fx, fy, cx, cy, c2w, gt_depth.to(device), gt_color.to(device), device)
        scores = self.keyframe_overlap_scores(rays_o, rays_d, gt_depth, [kf['est_c2w'] for kf in keyframe_dict], N_samples)
        list_keyframe = [{'id': i, 'percent_inside': s} for i, s in enumerate(scores)]
        list_keyframe = sorted(list_keyframe, key=lambda i: i['percent_inside'], reverse=True)
        selected = [dic['id'] for dic in list_keyframe if dic['percent_inside'] > 0.00]
        return list(np.random.permutation(np.array(selected))[:k])

    # ---- ray-sharded mapping ----------------------------------------------------------------
    def enable_sharding(self, group=None, peer_memory=True):
        """Shard the rays of every optimize_map iteration over the ranks of ``group`` (default: the
        world group; one process per GPU, all ranks call optimize_map with the same arguments and hold
        identical grids / decoders).  Every rank samples mapping_pixels / world pixels per iteration,
        the gradients are summed over ranks (peer-memory exchange kernels inside the iteration's CUDA
        graph when ``peer_memory``, else one NCCL all-reduce), and every rank applies the identical
        optimiser step, so the replicas stay bit-identical.  Give every rank its own RNG stream
        (``torch.cuda.manual_seed(seed + rank)``) so that the ranks draw different pixels."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("enable_sharding needs an initialised torch.distributed process group")
        self.shard_group = group
        self.shard_world, self.shard_rank = dist.get_world_size(group), dist.get_rank(group)
        self.shard_peers = None
        if peer_memory and self.shard_world > 1:
            from .peers import PeerGroup
            self.shard_peers = PeerGroup(group, device=torch.device(self.device))
        self._loops = {}

        # ---- the mapping optimisation ----------------------------------------------------------
    def optimize_map(self, num_joint_iters, lr_factor, idx, cur_gt_color, cur_gt_depth, gt_cur_c2w,
                     keyframe_dict, keyframe_list, cur_c2w):
        """Mapping iterations: sample pixels from the selected keyframes, optimise the scene
        representation (and the camera poses under local BA) (Mapper.py:295-669).
        Returns the updated cur_c2w when BA ran, else None."""
        H, W, fx, fy, cx, cy = self.H, self.W, self.fx, self.fy, self.cx, self.cy
        c = self.c
        cfg = self.cfg
        device = self.device
        bottom = torch.tensor([[0, 0, 0, 1.]], dtype=torch.float32, device=device)
        if len(keyframe_dict) == 0:
            optimize_frame = []
        else:
            num = self.mapping_window_size - 2
            if self.keyframe_selection_method == 'global':
                optimize_frame = random_select(len(self.keyframe_dict) - 1, num)
            else:
                optimize_frame = self.keyframe_selection_overlap(
                    cur_gt_color, cur_gt_depth, cur_c2w, keyframe_dict[:-1], num)
        if self.shard_world > 1:
            # the selection draws random numbers: every rank optimises the frames rank 0 picked
            import torch.distributed as dist
            box = [[int(f) for f in optimize_frame]]
            dist.broadcast_object_list(box, src=dist.get_global_rank(self.shard_group, 0) if self.shard_group is not None else 0,
                                       group=self.shard_group)
            optimize_frame = list(box[0])
        oldest_frame = None
        if len(keyframe_list) > 0:
            optimize_frame = optimize_frame + [len(keyframe_list) - 1]
            oldest_frame = min(optimize_frame)
        optimize_frame += [-1]
        if self.save_selected_keyframes_info:
            info = []
            for frame in optimize_frame:
                if frame != -1:
                    info.append({'idx': keyframe_list[frame], 'gt_c2w': keyframe_dict[frame]['gt_c2w'],
                                 'est_c2w': keyframe_dict[frame]['est_c2w']})
                else:
                    info.append({'idx': idx, 'gt_c2w': gt_cur_c2w, 'est_c2w': cur_c2w})
            self.selected_keyframes[idx] = info
        pixs_per_image = self.mapping_pixels // len(optimize_frame)
        if self.fused_loop and self.nice and self.occupancy and self.renderer.fixed_shape_ok():
            return self._optimize_map_fused(num_joint_iters, lr_factor, optimize_frame, oldest_frame, pixs_per_image,
                                            cur_gt_color, cur_gt_depth, keyframe_dict, cur_c2w)

        if self.shard_world > 1:
            raise NotImplementedError("ray-sharded mapping runs on the fused loop only (NICE decoders, occupancy, "
                                      "no jitter / importance sampling, fused_loop = True)")
        decoders_para_list = []
        grid_para = {'grid_coarse': [], 'grid_middle': [], 'grid_fine': [], 'grid_color': []}
        if self.nice:
            if self.frustum_feature_selection:
                masked_c_grad = {}
                mask_c2w = cur_c2w
            for key, val in c.items():
                val = val.to(device)
                if val.dim() == 5:     # the kernels' layout; logical shape unchanged
                    val = val.contiguous(memory_format=torch.channels_last_3d)
                if not self.frustum_feature_selection:
                    val = val.detach().requires_grad_(True)
                    c[key] = val
                    grid_para[key].append(val)
                else:
                    mask = self.get_mask_from_c2w(mask_c2w, key, val.shape[2:], cur_gt_depth)
                    mask = torch.from_numpy(mask).permute(2, 1, 0).unsqueeze(0).unsqueeze(0).repeat(
                        1, val.shape[1], 1, 1, 1).to(device)
                    c[key] = val
                    val_grad = val[mask].clone().detach().requires_grad_(True)
                    masked_c_grad[key] = val_grad
                    masked_c_grad[key + 'mask'] = mask
                    grid_para[key].append(val_grad)
            if not self.fix_fine:
                decoders_para_list += list(self.decoders.fine_decoder.parameters())
            if not self.fix_color:
                decoders_para_list += list(self.decoders.color_decoder.parameters())
        else:
            decoders_para_list += list(self.decoders.parameters())
        # only parameters an optimiser group owns need a gradient (the reference lets autograd
        # compute the others and drops them)
        wanted = {id(p) for p in decoders_para_list}
        for p in self.decoders.parameters():
            p.requires_grad_(id(p) in wanted)

        camera_tensor_list, gt_camera_tensor_list = [], []
        if self.BA:
            for frame in optimize_frame:
                if frame != oldest_frame:
                    if frame != -1:
                        c2w, gt_c2w = keyframe_dict[frame]['est_c2w'], keyframe_dict[frame]['gt_c2w']
                    else:
                        c2w, gt_c2w = cur_c2w, gt_cur_c2w
                    camera_tensor = get_tensor_from_camera(c2w).to(device).requires_grad_(True)
                    camera_tensor_list.append(camera_tensor)
                    gt_camera_tensor_list.append(get_tensor_from_camera(gt_c2w))
        if self.nice:
            groups = [{'params': decoders_para_list, 'lr': 0},
                      {'params': grid_para['grid_coarse'], 'lr': 0},
                      {'params': grid_para['grid_middle'], 'lr': 0},
                      {'params': grid_para['grid_fine'], 'lr': 0},
                      {'params': grid_para['grid_color'], 'lr': 0}]
            if self.BA:
                groups.append({'params': camera_tensor_list, 'lr': 0})
            optimizer = torch.optim.Adam(groups)
        else:
            groups = [{'params': decoders_para_list, 'lr': 0}]
            if self.BA:
                groups.append({'params': camera_tensor_list, 'lr': 0})
            optimizer = torch.optim.Adam(groups)
            from torch.optim.lr_scheduler import StepLR
            scheduler = StepLR(optimizer, step_size=200, gamma=0.8)

        for joint_iter in range(num_joint_iters):
            if self.nice:
                if self.frustum_feature_selection:
                    for key, val in c.items():
                        if (self.coarse_mapper and 'coarse' in key) or \
                                ((not self.coarse_mapper) and ('coarse' not in key)):
                            val_grad = masked_c_grad[key]
                            mask = masked_c_grad[key + 'mask']
                            val = val.to(device)
                            val[mask] = val_grad
                            c[key] = val
                if self.coarse_mapper:
                    self.stage = 'coarse'
                elif joint_iter <= int(num_joint_iters * self.middle_iter_ratio):
                    self.stage = 'middle'
                elif joint_iter <= int(num_joint_iters * self.fine_iter_ratio):
                    self.stage = 'fine'
                else:
                    self.stage = 'color'
                st = cfg['mapping']['stage'][self.stage]
                optimizer.param_groups[0]['lr'] = st['decoders_lr'] * lr_factor
                optimizer.param_groups[1]['lr'] = st['coarse_lr'] * lr_factor
                optimizer.param_groups[2]['lr'] = st['middle_lr'] * lr_factor
                optimizer.param_groups[3]['lr'] = st['fine_lr'] * lr_factor
                optimizer.param_groups[4]['lr'] = st['color_lr'] * lr_factor
                if self.BA and self.stage == 'color':
                    optimizer.param_groups[5]['lr'] = self.BA_cam_lr
            else:
                self.stage = 'color'
                optimizer.param_groups[0]['lr'] = cfg['mapping']['imap_decoders_lr']
                if self.BA:
                    optimizer.param_groups[1]['lr'] = self.BA_cam_lr

            optimizer.zero_grad()
            ro_l, rd_l, gd_l, gc_l = [], [], [], []
            camera_tensor_id = 0
            for frame in optimize_frame:
                if frame != -1:
                    gt_depth = keyframe_dict[frame]['depth'].to(device)
                    gt_color = keyframe_dict[frame]['color'].to(device)
                    if self.BA and frame != oldest_frame:
                        camera_tensor = camera_tensor_list[camera_tensor_id]
                        camera_tensor_id += 1
                        c2w = get_camera_from_tensor(camera_tensor)
                    else:
                        c2w = keyframe_dict[frame]['est_c2w']
                else:
                    gt_depth = cur_gt_depth.to(device)
                    gt_color = cur_gt_color.to(device)
                    if self.BA:
                        camera_tensor = camera_tensor_list[camera_tensor_id]
                        c2w = get_camera_from_tensor(camera_tensor)
                    else:
                        c2w = cur_c2w
                ro, rd, gd, gc = get_samples(0, H, 0, W, pixs_per_image, H, W, fx, fy, cx, cy, c2w,
                                             gt_depth, gt_color, self.device,
                                             indices=None if self.index_source is None else next(self.index_source))
                ro_l.append(ro.float()), rd_l.append(rd.float()), gd_l.append(gd.float()), gc_l.append(gc.float())
            batch_rays_d, batch_rays_o = torch.cat(rd_l), torch.cat(ro_l)
            batch_gt_depth, batch_gt_color = torch.cat(gd_l), torch.cat(gc_l)

            keep = None
            if self.nice:                                     # Mapper.py:580-596
                # the reference compacts the batch with boolean indexing (a host sync per
                # iteration); the fused path keeps the shape and passes the mask to the kernels
                keep = fused.box_prefilter(batch_rays_o, batch_rays_d, batch_gt_depth, self.bound)
            if keep is not None and not self.renderer.fixed_shape_ok():
                m = keep.bool()
                batch_rays_d, batch_rays_o = batch_rays_d[m], batch_rays_o[m]
                batch_gt_depth, batch_gt_color = batch_gt_depth[m], batch_gt_color[m]
                keep = None
            depth, uncertainty, color = self.renderer.render_batch_ray(
                c, self.decoders, batch_rays_d, batch_rays_o, device, self.stage,
                gt_depth=None if self.coarse_mapper else batch_gt_depth, keep=keep)
            loss = fused.mapping_loss(depth, color, batch_gt_depth, batch_gt_color, self.w_color_loss,
                                      use_color=(not self.nice) or (self.stage == 'color'), keep=keep)
            if not self.occupancy:                            # iMAP regulariser, Mapper.py:614-621
                point_sigma = self.renderer.regulation(
                    c, self.decoders, batch_rays_d, batch_rays_o, batch_gt_depth, device, self.stage)
                loss += 0.0005 * torch.abs(point_sigma).sum()
            loss.backward(retain_graph=False)
            optimizer.step()
            if not self.nice:
                scheduler.step()
            optimizer.zero_grad()

            if self.nice and self.frustum_feature_selection:  # Mapper.py:633-642
                for key, val in c.items():
                    if (self.coarse_mapper and 'coarse' in key) or \
                            ((not self.coarse_mapper) and ('coarse' not in key)):
                        val_grad = masked_c_grad[key]
                        mask = masked_c_grad[key + 'mask']
                        val = val.detach()
                        val[mask] = val_grad.clone().detach()
                        c[key] = val

        if self.BA:                                           # Mapper.py:644-667
            camera_tensor_id = 0
            for frame in optimize_frame:
                if frame != -1:
                    if frame != oldest_frame:
                        c2w = get_camera_from_tensor(camera_tensor_list[camera_tensor_id].detach())
                        c2w = torch.cat([c2w, bottom], dim=0)
                        camera_tensor_id += 1
                        keyframe_dict[frame]['est_c2w'] = c2w.clone()
                else:
                    c2w = get_camera_from_tensor(camera_tensor_list[-1].detach())
                    cur_c2w = torch.cat([c2w, bottom], dim=0).clone()
            return cur_c2w
        return None

    def _stage_of(self, joint_iter, num_joint_iters):
        if self.coarse_mapper:                                # Mapper.py:506-515
            return 'coarse'
        if joint_iter <= int(num_joint_iters * self.middle_iter_ratio):
            return 'middle'
        if joint_iter <= int(num_joint_iters * self.fine_iter_ratio):
            return 'fine'
        return 'color'

    def _optimize_map_fused(self, num_joint_iters, lr_factor, optimize_frame, oldest_frame, pixs_per_image,
                            cur_gt_color, cur_gt_depth, keyframe_dict, cur_c2w):
        """The iteration loop of optimize_map (Mapper.py:488-667) through map_loop.MappingLoop: same
        schedule, same sampled frames, same optimiser arithmetic, no autograd graph and no host
        round trip per iteration."""
        from .map_loop import MappingLoop
        device = self.device
        bottom = torch.tensor([[0, 0, 0, 1.]], dtype=torch.float32, device=device)
        frames, ba_slots = [], []
        for frame in optimize_frame:
            if frame != -1:
                depth, color = keyframe_dict[frame]['depth'], keyframe_dict[frame]['color']
                c2w = keyframe_dict[frame]['est_c2w']
            else:
                depth, color, c2w = cur_gt_depth, cur_gt_color, cur_c2w
            ba = bool(self.BA and frame != oldest_frame)
            ba_slots.append(ba)
            frames.append((depth.to(device), color.to(device),
                           get_tensor_from_camera(c2w).to(device) if ba else c2w.to(device).float()))
        masked = bool(self.frustum_feature_selection)
        f64 = False      # the reference casts the sampled colours to fp32 (Mapper.py:575); casting the frame first is the same
        if self.shard_world > 1:                              # every rank renders its share of the pixels
            pixs_per_image = max(1, pixs_per_image // self.shard_world)
        key = (len(frames), pixs_per_image, tuple(ba_slots), masked, f64, self.shard_world)
        loop = self._loops.get(key)
        if loop is None or not loop.valid_for(self):
            self._loops = {k: v for k, v in self._loops.items() if v.valid_for(self)}
            loop = self._loops[key] = MappingLoop(self, len(frames), pixs_per_image, ba_slots, masked, f64,
                                                  group=self.shard_group if self.shard_world > 1 else None,
                                                  peers=self.shard_peers, sharded=self.shard_world > 1)
        masks = None
        if masked:
            masks = {}
            for gk in loop.grid_keys:
                masks[gk] = self._frustum_mask(cur_c2w, gk, self.c[gk].shape[2:], cur_gt_depth)
        loop.begin(frames, masks)
        prev = None
        for joint_iter in range(num_joint_iters):
            self.stage = self._stage_of(joint_iter, num_joint_iters)
            if self.stage != prev:
                loop.set_lr(self.stage, lr_factor)
                prev = self.stage
            loop.draw = self.index_source is None            # set on EVERY call: the loops (and their graphs) are cached
            if self.index_source is not None:
                for i in range(loop.K):
                    loop.indices[i * loop.pix:(i + 1) * loop.pix].copy_(next(self.index_source))
            loop.iterate(self.stage)
        loop.finish()
        if self.shard_peers is not None:
            # a rank that missed a peer barrier leaves partial sums behind and the replicas diverge: one host
            # read per optimize_map call, and raise
            self.shard_peers.check_error()
        if self.BA:                                           # Mapper.py:644-667
            for i, frame in enumerate(optimize_frame):
                if not ba_slots[i]:
                    continue
                c2w = torch.cat([get_camera_from_tensor(loop.camera(i)), bottom], dim=0)
                if frame != -1:
                    keyframe_dict[frame]['est_c2w'] = c2w.clone()
                else:
                    cur_c2w = c2w.clone()
            return cur_c2w
        return None

    def run(self):
        raise NotImplementedError(
            "Mapper.run (dataset loop, tracker synchronisation, keyframe bookkeeping, meshing, logging: "
            "src/Mapper.py:671-806) is outside the render hot path; drive optimize_map from the reference's loop")
