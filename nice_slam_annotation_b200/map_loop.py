"""The iterations of one ``Mapper.optimize_map`` call on the device (src/Mapper.py:488-642).

Reference iteration: 6 boolean-mask ``index_put``s over full grids (+ ``nonzero`` syncs), K
``get_samples`` calls, ~620 ATen launches for render + loss + backward, ``torch.optim.Adam`` over
the compact ``val[mask]`` copies, write-back copies.  Here one iteration is

    pixel draw -> K x nsb_sample_rays (rays + gt gather + box pre-filter mask)
    -> FusedStep (zero the gradient bucket, re-pack trained decoders, render forward, mapping loss,
       render backward: grid scatter, decoder dW, ray gradients under BA)
    -> [sharded: ONE all-reduce of the bucket]
    -> nsb_adam per optimised tensor: grids IN PLACE under the frustum voxel mask (SURVEY.md 8f-1:
       the elementwise update of val[mask] + its write-back, without the compact copy), the trained
       decoders as one flat tensor each, camera tensors with nsb_pose_grad + nsb_cam_adam under BA

and, on one GPU, the whole list is replayed as ONE CUDA graph per stage.  Optimiser state is fresh
per call (the reference builds a new Adam per optimize_map, Mapper.py:462-482); learning rates live
in device scalars so the graphs follow the stage schedule and ``lr_factor``.
"""
import torch

from . import fused
from ._lib import check, lib
from .fused import DEC_NAMES, GRID_KEYS, STAGE_DECODERS, _ptr, _stream, canonical_params
from .step import FusedStep, graph_capture

BETAS, EPS = (0.9, 0.999), 1e-8          # torch.optim.Adam defaults (the reference passes none)
GROUP_LR = {"grid_coarse": "coarse_lr", "grid_middle": "middle_lr", "grid_fine": "fine_lr", "grid_color": "color_lr"}


def flatten_decoder(dec, kind, originals=None):
    """Re-point the parameters of one decoder at views of ONE flat buffer in canonical order (the
    order of its flat gradient), so that its Adam step is a single kernel.  Parameter objects,
    names and values are unchanged (state_dict / load_state_dict / deepcopy keep working).

    The storages the parameters had before are NOT the ones trained afterwards.  When they are shared with
    another process (the reference passes ``shared_decoders`` to the tracker through CUDA IPC at spawn,
    NICE_SLAM.py:134,357-376) flatten BEFORE sharing (``shared.prepare_shared``); a caller that flattens later
    passes ``originals`` (a list) to receive (old storage, new view) pairs and copies the trained values back
    (``MappingLoop.finish``)."""
    params = canonical_params(dec, kind)
    flat = getattr(dec, "_nsb_flat", None)
    n = sum(p.numel() for p in params)
    if flat is not None and flat.numel() == n and all(
            p.data_ptr() == flat.data_ptr() + 4 * o for p, o in zip(params, _offsets(params))):
        return flat
    flat = torch.empty(n, dtype=torch.float32, device=params[0].device)
    for p, o in zip(params, _offsets(params)):
        seg = flat[o:o + p.numel()].view(p.shape)
        seg.copy_(p.data)
        if originals is not None:
            originals.append((p.data, seg))
        p.data = seg
    dec._nsb_flat = flat
    return flat


def _offsets(params):
    out, o = [], 0
    for p in params:
        out.append(o)
        o += p.numel()
    return out


class _Opt:
    """Adam state of one flat tensor: m, v, [step, ticket], lr scalar (all on the device)."""

    def __init__(self, p_flat, mask=None):
        self.p = p_flat
        self.m = torch.zeros_like(p_flat)
        self.v = torch.zeros_like(p_flat)
        self.state2 = torch.zeros(2, dtype=torch.int32, device=p_flat.device)
        self.lr = torch.zeros(1, device=p_flat.device)
        self.mask = mask

    def reset(self):
        self.m.zero_()
        self.v.zero_()
        self.state2.zero_()

    def step(self, g_flat):
        check(lib().nsb_adam(_ptr(self.p), _ptr(g_flat), _ptr(self.m), _ptr(self.v), self.p.numel(),
                             _ptr(self.mask), _ptr(self.state2), _ptr(self.lr), BETAS[0], BETAS[1], EPS, _stream()),
              "adam")


def _flat_of(grid):
    """The channels-last memory of a [1,32,Z,Y,X] grid as a flat tensor (no copy)."""
    return grid.permute(0, 2, 3, 4, 1).reshape(-1)


class MappingLoop:
    """mapper : nice_slam_annotation_b200.Mapper.Mapper (cfg, renderer, decoders, c, bound, intrinsics)
    n_frames, pixs : frames per iteration and pixels per frame (Mapper.py:489-490)
    ba_slots : per frame slot, True if its pose is optimised (local BA, Mapper.py:430-458)
    masked : frustum_feature_selection (a voxel mask per grid is loaded per call)
    """

    def __init__(self, mapper, n_frames, pixs, ba_slots=None, masked=True, color_f64=False, group=None, peers=None,
                 sharded=None):
        mp = self.mp = mapper
        self.K, self.pix = int(n_frames), int(pixs)
        self.N = self.K * self.pix
        self.ba = [bool(x) for x in (ba_slots or [False] * self.K)]
        self.masked, self.group = masked, group
        dev = self.dev = next(mp.decoders.parameters()).device
        H, W = mp.H, mp.W
        self.stages = ["coarse"] if mp.coarse_mapper else ["middle", "fine", "color"]
        # static frame slots
        self.depth = torch.zeros(self.K, H, W, device=dev)
        self.color = torch.zeros(self.K, H, W, 3, dtype=torch.float64 if color_f64 else torch.float32, device=dev)
        self.c2w = torch.zeros(self.K, 12, device=dev)
        self.cam = torch.zeros(self.K, 7, device=dev)
        # camera gradients: rows of 8 floats; in symmetric memory when ranks exchange through peer memory
        self.peers = peers
        if peers is not None:
            self._d_cam_flat = peers.alloc(self.K * 8)
            self._d_cam_px = peers.struct(self._d_cam_flat, use_multicast=False)
        else:
            self._d_cam_flat = torch.zeros(self.K * 8, device=dev)
        self.d_cam = self._d_cam_flat.view(self.K, 8)[:, :7]
        self.cam_m = torch.zeros(self.K, 7, device=dev)
        self.cam_v = torch.zeros(self.K, 7, device=dev)
        self.cam_t = torch.zeros(self.K, 1, dtype=torch.int32, device=dev)
        self.indices = torch.zeros(self.N, dtype=torch.int64, device=dev)
        self.draw = True
        self.bound6 = fused._bound6(mp.bound)
        self.intr = tuple(float(x) for x in (mp.fx, mp.fy, mp.cx, mp.cy))
        # grids: optimised in place, in the kernels' layout.  A grid that is not in that layout yet is re-allocated;
        # its old storage may be the one another process reads (shared_c travels to the tracker by CUDA IPC), so it
        # is kept and refreshed by finish() -- unless the caller promised prepared shared objects, then it is an error
        self._write_back = []
        strict = bool(getattr(mp, "shared_prepared", False))
        for key in list(mp.c.keys()):
            g = mp.c[key]
            if not (g.is_cuda and g.is_contiguous(memory_format=torch.channels_last_3d)):
                if strict:
                    raise RuntimeError(f"{key} is not channels_last_3d on the device although shared.prepare_shared ran: "
                                       "re-allocating it here would detach it from the other processes")
                new = g.detach().to(dev).contiguous(memory_format=torch.channels_last_3d)
                if g.is_cuda:
                    self._write_back.append((g.detach(), new))
                mp.c[key] = new
            elif g.requires_grad:
                mp.c[key] = g.detach()
        self.grid_keys = ["grid_coarse"] if mp.coarse_mapper else ["grid_middle", "grid_fine", "grid_color"]
        self.grid_ptrs = {k: mp.c[k].data_ptr() for k in self.grid_keys}
        self.masks = {k: (torch.ones(mp.c[k].shape[2:], dtype=torch.uint8, device=dev) if masked else None)
                      for k in self.grid_keys}
        self.opt = {k: _Opt(_flat_of(mp.c[k]), self.masks[k]) for k in self.grid_keys}
        # trained decoders (Mapper.py:419-428), each one flat tensor
        self.train_dec = []
        if not mp.coarse_mapper:
            if not mp.fix_fine:
                self.train_dec.append("fine_decoder")
            if not mp.fix_color:
                self.train_dec.append("color_decoder")
        self.dec_kind = {name: k for k, name in enumerate(DEC_NAMES)}
        for name in self.train_dec:
            n_before = len(self._write_back)
            self.opt[name] = _Opt(flatten_decoder(getattr(mp.decoders, name), self.dec_kind[name], self._write_back))
            if strict and len(self._write_back) != n_before:
                raise RuntimeError(f"{name} had to be re-flattened although shared.prepare_shared ran: its old "
                                   "storages are the ones the other processes read")
        want = {id(p) for name in self.train_dec for p in getattr(mp.decoders, name).parameters()}
        for p in mp.decoders.parameters():
            p.requires_grad_(id(p) in want)
        # one engine step per stage
        self.steps, self.graphs = {}, {}
        for st in self.stages:
            kinds = STAGE_DECODERS[st]
            self.steps[st] = FusedStep(
                mp.renderer, mp.decoders, mp.c, st, self.N, loss="mapping", w_color=mp.w_color_loss,
                use_color=(st == "color"), grid_grads=tuple(GRID_KEYS[k] for k in kinds),
                decoder_grads=tuple(n for n in self.train_dec if self.dec_kind[n] in kinds),
                ray_grads=any(self.ba), with_gt_depth=not mp.coarse_mapper, group=group, peers=peers, sharded=sharded)

    # -- per call ---------------------------------------------------------------------------
    def valid_for(self, mapper):
        """Grid storages and the flat decoder buffers are still the ones the kernels point at."""
        if not all(mapper.c[k].data_ptr() == p for k, p in self.grid_ptrs.items()):
            return False
        for name in self.train_dec:
            dec = getattr(mapper.decoders, name)
            flat = getattr(dec, "_nsb_flat", None)
            if flat is None or flat is not self.opt[name].p or \
                    canonical_params(dec, self.dec_kind[name])[0].data_ptr() != flat.data_ptr():
                return False
        return True

    def begin(self, frames, masks=None):
        """frames: K x (depth [H,W], color [H,W,3], c2w [3|4,4] or camera tensor [7] for BA slots).
        masks: {grid key: [Z,Y,X] bool / uint8} when the loop is masked."""
        for i, (d, c, pose) in enumerate(frames):
            self.depth[i].copy_(d, non_blocking=True)
            self.color[i].copy_(c, non_blocking=True)
            if self.ba[i]:
                self.cam[i].copy_(pose.detach().reshape(7), non_blocking=True)
            else:
                self.c2w[i].copy_(pose.detach().reshape(-1)[:12], non_blocking=True)
        if self.masked:
            for k in self.grid_keys:
                self.masks[k].copy_(masks[k].to(torch.uint8), non_blocking=True)
        for st in self.steps.values():                           # decoders may have been loaded / updated since
            for k in st.kinds:
                st._pack(k)
        for o in self.opt.values():
            o.reset()
        for x in (self.cam_m, self.cam_v, self.cam_t):
            x.zero_()

    def set_lr(self, stage, lr_factor):
        st = self.mp.cfg["mapping"]["stage"][stage]
        for k in self.grid_keys:
            self.opt[k].lr.fill_(st[GROUP_LR[k]] * lr_factor)
        for name in self.train_dec:
            self.opt[name].lr.fill_(st["decoders_lr"] * lr_factor)

    # -- one iteration ------------------------------------------------------------------------
    def _sample(self, st):
        s, L, mp = _stream(), lib(), self.mp
        fx, fy, cx, cy = self.intr
        if self.draw:                                            # common.py:99, once for the K frames
            torch.randint(0, mp.H * mp.W, (self.N,), device=self.dev, out=self.indices)
        f64 = int(self.color.dtype == torch.float64)
        for i in range(self.K):
            a, b = i * self.pix, (i + 1) * self.pix
            check(L.nsb_sample_rays(_ptr(self.indices[a:b]), self.pix, 0, mp.H, 0, mp.W, mp.W, fx, fy, cx, cy,
                                    None if self.ba[i] else _ptr(self.c2w[i]), _ptr(self.cam[i]) if self.ba[i] else None,
                                    _ptr(self.depth[i]), _ptr(self.color[i]), f64, self.bound6,
                                    _ptr(st.rays_o[a:b]), _ptr(st.rays_d[a:b]), _ptr(st.gt_depth[a:b]),
                                    _ptr(st.gt_color[a:b]), _ptr(st.keep[a:b]), s), "sample_rays")

    def _update(self, stage, st):
        s, L, mp = _stream(), lib(), self.mp
        fx, fy, cx, cy = self.intr
        for k in st.grid_grad_kinds:
            key = GRID_KEYS[k]
            self.opt[key].step(_flat_of(st.bucket.view_of(st.grids[k])))
        for k, holder in zip(st.dec_grad_kinds, st.dw_holders):
            self.opt[DEC_NAMES[k]].step(st.bucket.view_of(holder))
        lr_cam = float(mp.BA_cam_lr) if stage == "color" else 0.0       # Mapper.py:527-528
        for i in range(self.K):
            if not self.ba[i]:
                continue
            a, b = i * self.pix, (i + 1) * self.pix
            check(L.nsb_pose_grad(_ptr(st.g_rays_o[a:b]), _ptr(st.g_rays_d[a:b]), _ptr(self.indices[a:b]),
                                  _ptr(st.keep[a:b]), self.pix, 0, 0, mp.W, fx, fy, cx, cy, _ptr(self.cam[i]),
                                  _ptr(self.d_cam[i]), s), "pose_grad")
        if any(self.ba) and st._sharded():
            # the pose gradient is linear in the ray gradients: partial d(camera) of the ranks add
            if self.peers is not None:
                import ctypes as C
                check(L.nsb_bucket_allreduce(C.byref(self._d_cam_px), 0, self.K * 8, s), "cam_allreduce")
            else:
                import torch.distributed as dist
                dist.all_reduce(self._d_cam_flat, group=self.group)
        for i in range(self.K):
            if not self.ba[i]:
                continue
            check(L.nsb_cam_adam(_ptr(self.cam[i]), _ptr(self.d_cam[i]), _ptr(self.cam_m[i]), _ptr(self.cam_v[i]),
                                 _ptr(self.cam_t[i]), lr_cam, lr_cam, BETAS[0], BETAS[1], EPS, None, None, None, s),
                  "cam_adam")

    def _launch(self, stage):
        st = self.steps[stage]
        self._sample(st)
        if st._sharded() and st.with_gt_depth and not st._peer_max:       # batch depth max over all ranks (NCCL)
            import torch.distributed as dist
            check(lib().nsb_depth_max(_ptr(st.gt_depth), _ptr(st.keep), st.N, _ptr(st.depth_max), _stream()), "depth_max")
            dist.all_reduce(st.depth_max, op=dist.ReduceOp.MAX, group=self.group)
        st._launch()
        if st._sharded() and st.bucket.peers is None:
            st.bucket.all_reduce(self.group)                     # NCCL; the peer exchange ran inside st._launch
        self._update(stage, st)

    def iterate(self, stage, graph=True):
        """One mapping iteration of ``stage`` (call ``set_lr`` when the stage changes)."""
        st = self.steps[stage]
        if not graph or (st._sharded() and st.bucket.peers is None):    # NCCL calls are not captured
            self._launch(stage)
            return
        key = (stage, bool(self.draw))           # the pixel draw is a node of the graph: one graph per draw flag
        g = self.graphs.get(key)
        if g is None:
            # warm-up outside the capture must not advance the optimiser (render only) nor the RNG stream
            # (a seeded run must see the same draws with and without graphs)
            rng = torch.cuda.get_rng_state(self.dev) if self.draw else None
            keep_idx = self.indices.clone()
            self._sample(st)
            st._launch(exchange=False)
            if rng is not None:
                torch.cuda.set_rng_state(rng, self.dev)
            self.indices.copy_(keep_idx)
            g = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            with graph_capture(g):
                self._launch(stage)
            self.graphs[key] = g
        g.replay()

    def finish(self):
        """The kernels updated the trained parameters in place: bump their autograd version so
        that version-keyed caches (fused.packed_decoder) re-pack them."""
        for name in self.train_dec:
            for p in getattr(self.mp.decoders, name).parameters():
                torch.autograd.graph.increment_version(p)
        for old, new in self._write_back:        # storages this loop replaced (see __init__): keep them current
            old.copy_(new)

    def camera(self, slot):
        return self.cam[slot].clone()
