"""One SLAM optimisation iteration as three library calls: render forward, loss, render
backward -- no autograd graph, fixed shapes, every buffer allocated once.

This is the engine under ``Mapper.optimize_map`` / ``Tracker.optimize_cam_in_batch`` when the
iteration has the static shape the fused path wants (NICE decoders, no jitter, pre-filter as a
mask): the reference's iteration body (src/Mapper.py:598-624, src/Tracker.py:113-132) is ~620
ATen launches + 3 host syncs, here it is ~10 launches that can be replayed as one CUDA graph.

Mapping shards by rays: every rank renders its slice of the batch against
replicated grids / decoders, the gradients land in ONE flat bucket (``GradBucket``) that is
sum-all-reduced in a single NCCL call, after which every rank applies the identical optimiser
step.  The batch-global scalar ``max(gt_depth)`` (src/utils/Renderer.py:157,194) is MAX-reduced
so the sharded render equals the unsharded one.
"""
import ctypes as C
import contextlib
import gc
import os

import torch

from . import fused
from ._lib import GRAD_GRID, GRAD_POINT, GRAD_WEIGHT, NsbRenderCall, STAGES, check, lib
from .fused import DEC_NAMES, GRID_KEYS, STAGE_DECODERS, _linspace, _ptr, _stream, canonical_params, pack_table


# ---------------------------------------------------------------------------------------
# gradient bucket (pure torch: also exercised on CPU / gloo by the tests)
# ---------------------------------------------------------------------------------------
@contextlib.contextmanager
def graph_capture(g):
    """``torch.cuda.graph(g)`` with Python's cyclic garbage collector off for the duration of the capture.  A collection
    that happens to run while a stream is capturing can free CUDA objects of unrelated, already dead owners (steps,
    graphs, side streams of earlier work) and the allocator's bookkeeping for them invalidates the capture
    (cudaErrorStreamCaptureInvalidated at the next launch) -- seen when merely the number of collected test items moved
    the collector's trigger into a capture.  torch collects once before the capture starts; nothing needs collecting
    during it."""
    was = gc.isenabled()
    gc.collect()
    gc.disable()
    try:
        with torch.cuda.graph(g):
            yield
    finally:
        if was:
            gc.enable()


class GradBucket:
    """Gradients of a list of leaves as views of one flat fp32 buffer.

    A view has the shape AND the memory layout of its leaf (channels-last grids stay
    channels-last), so kernels scatter straight into it and ``leaf.grad = view`` costs nothing.
    ``all_reduce`` is then one collective over the flat buffer, with no flatten / unflatten copies.
    """

    def __init__(self, leaves, device=None, peers=None):
        self.leaves = list(leaves)
        device = device if device is not None else self.leaves[0].device
        self.offsets, n = [], 0
        for p in self.leaves:
            self.offsets.append(n)
            n += (p.numel() + 63) // 64 * 64            # 256-byte aligned segments
        # peers (peers.PeerGroup): the bucket lives in symmetric memory and all_reduce is ONE kernel over
        # NVLink peer / multicast mappings (csrc/nsb_exchange.cu) instead of a NCCL call
        self.peers = peers
        if peers is not None:
            self.flat = peers.alloc(n)
            # sparse exchange (only the voxel lines some rank touched travel); NSB_X_DENSE=1 keeps the dense one
            self.sparse = os.environ.get("NSB_X_DENSE", "0") != "1"
            self.touched = peers.alloc_touched(n) if self.sparse else None
            self._px = peers.struct(self.flat, touched=self.touched)
            # touched-line marks (local): the tcgen05 backward writes one byte per 128-byte line it scatters into, the
            # sparse exchange reads them instead of scanning the bucket (first word: "marks are valid"); they cover the
            # leading 5-d leaves (the grids), the decoder segments behind them are scanned
            self.marks = None
            n_grid = 0
            for p_, o_ in zip(self.leaves, self.offsets):
                if p_.dim() != 5:
                    break
                n_grid = o_ + (p_.numel() + 63) // 64 * 64
            self.n_grid = n_grid                     # floats of the leading grid segments
            if self.sparse and os.environ.get("NSB_X_MARKS", "1") != "0":
                self.marks = torch.zeros(128 + (n + 31) // 32, dtype=torch.uint8, device=self.flat.device)
                self._px.line_marks = self.marks.data_ptr()
                self._px.line_marks_floats = n_grid
        else:
            self.flat = torch.zeros(n, dtype=torch.float32, device=device)
            self.marks = None
            if os.environ.get("NSB_FORCE_MARKS") == "1":          # timing experiments: mark without an exchange
                self.marks = torch.zeros(128 + (n + 31) // 32, dtype=torch.uint8, device=device)
        self.views = [self._view(p, o) for p, o in zip(self.leaves, self.offsets)]

    def _view(self, p, off):
        seg = self.flat[off:off + p.numel()]
        if p.dim() == 5 and p.is_contiguous(memory_format=torch.channels_last_3d) and not p.is_contiguous():
            n, c, z, y, x = p.shape
            return seg.view(n, z, y, x, c).permute(0, 4, 1, 2, 3)
        if not p.is_contiguous():
            raise RuntimeError("GradBucket: leaves must be contiguous or channels_last_3d")
        return seg.view(p.shape)

    def view_of(self, leaf):
        for p, v in zip(self.leaves, self.views):
            if p is leaf:
                return v
        raise KeyError("leaf is not in the bucket")

    def zero(self):
        self.flat.zero_()

    def all_reduce(self, group=None):
        """SUM over ranks (the mapping loss is a plain sum over rays, src/Mapper.py:605-612)."""
        if self.peers is not None:
            fn = lib().nsb_bucket_allreduce_sparse if self.sparse else lib().nsb_bucket_allreduce
            check(fn(C.byref(self._px), 0, self.flat.numel(), _stream()), "bucket_allreduce")
            return
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)

    def all_reduce_range(self, offset, n, max_ctas=0):
        """Peer-memory SUM of flat[offset, offset + n) (sparse exchange), optionally with few CTAs so that it can run
        beside another kernel.  Every rank must make the same sequence of calls."""
        if n <= 0:
            return
        check(lib().nsb_bucket_allreduce_sparse_ctas(C.byref(self._px), int(offset), int(n), int(max_ctas), _stream()),
              "bucket_allreduce_sparse")

    def assign(self):
        """leaf.grad = its view (what optimizer.step() reads)."""
        for p, v in zip(self.leaves, self.views):
            p.grad = v


def shard_slice(n, world, rank):
    """Contiguous slice of an n-ray batch owned by ``rank`` (sizes differ by at most one)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return slice(lo, lo + base + (1 if rank < rem else 0))


def global_depth_max(gt_depth, keep=None, group=None):
    """max(gt_depth) over the kept rays of ALL ranks, as a [1] fp32 tensor (Renderer.py:157)."""
    import torch.distributed as dist
    g = gt_depth.float()
    if keep is not None:
        g = torch.where(keep.bool(), g, torch.full_like(g, float("-inf")))
    m = g.max().reshape(1) if g.numel() else torch.full((1,), float("-inf"), device=gt_depth.device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(m, op=dist.ReduceOp.MAX, group=group)
    return m


# ---------------------------------------------------------------------------------------
# the fused iteration
# ---------------------------------------------------------------------------------------
class FusedStep:
    """Forward + loss + backward of one iteration at a fixed batch size.

    renderer : nice_slam_annotation_b200.utils.Renderer.Renderer (sampling parameters, bound)
    decoders : NICE module; c : dict of grids ([1,32,Z,Y,X], channels_last_3d)
    n_rays   : batch size of every call (mask dropped rays with ``keep``)
    loss     : 'mapping' (Mapper.py:604-612) or 'tracking' (Tracker.py:113-130)
    grid_grads / decoder_grads : names whose gradient is produced (into ``self.bucket``)
    ray_grads : also produce d(rays_o), d(rays_d) (tracking, BA)
    """

    def __init__(self, renderer, decoders, c, stage, n_rays, loss="mapping", w_color=0.2, use_color=True,
                 handle_dynamic=False, grid_grads=(), decoder_grads=(), ray_grads=False, with_gt_depth=True,
                 group=None, peers=None, sharded=None):
        if not renderer.fixed_shape_ok():
            raise NotImplementedError("FusedStep needs the NICE decoders, perturb == 0 and N_importance == 0")
        self.r, self.decoders, self.stage, self.N = renderer, decoders, stage, int(n_rays)
        self.kinds = STAGE_DECODERS[stage]
        self.loss_kind, self.w_color, self.use_color, self.handle_dynamic = loss, w_color, use_color, handle_dynamic
        self.group, self.ray_grads = group, ray_grads
        # sharded: this step renders a slice of a batch that other ranks complete (gradients are summed over
        # ``group``).  None = yes whenever a process group with more than one rank is initialised; pass
        # False for a step that is whole on this rank although torch.distributed is up.
        self._sharded_flag = sharded
        self.with_gt_depth = with_gt_depth and stage != "coarse"
        dev = next(decoders.parameters()).device
        self.dev = dev
        self.grids = {k: fused.grid_to_kernel_layout(c[GRID_KEYS[k]]).detach() for k in self.kinds}
        N = self.N
        self.S = renderer.N_samples + (renderer.N_surface if self.with_gt_depth else 0)
        # static inputs
        self.rays_o = torch.zeros(N, 3, device=dev)
        self.rays_d = torch.zeros(N, 3, device=dev)
        self.gt_depth = torch.zeros(N, device=dev)
        self.gt_color = torch.zeros(N, 3, device=dev)
        self.keep = torch.ones(N, dtype=torch.uint8, device=dev)
        self.depth_max = torch.zeros(1, device=dev)
        # outputs
        self.depth = torch.zeros(N, dtype=torch.float64, device=dev)
        self.var = torch.zeros(N, dtype=torch.float64, device=dev)
        self.color = torch.zeros(N, 3, device=dev)
        self.loss = torch.zeros(1, dtype=torch.float64, device=dev)
        self.g_depth = torch.zeros(N, dtype=torch.float64, device=dev)
        self.g_color = torch.zeros(N, 3, device=dev)
        self.g_rays_o = torch.zeros(N, 3, device=dev) if ray_grads else None
        self.g_rays_d = torch.zeros(N, 3, device=dev) if ray_grads else None
        # Large mapping batches run as ray chunks (forward + loss + backward per chunk, gradients accumulate in the bucket):
        # rays are independent once the batch depth max is known, so the result is the same, the per-call buffers stay
        # small and every chunk keeps the deposit + tcgen05 backward path (whole-batch records, nsb_layout.h).
        chunk = int(os.environ.get("NSB_STEP_CHUNK_RAYS", "16384"))
        can_chunk = (loss == "mapping" and self.with_gt_depth and bool(renderer.occupancy))      # = fused_loss below
        self.chunk = min(N, max(chunk, 1)) if can_chunk else N
        self.nchunk = (N + self.chunk - 1) // self.chunk
        self.ws = torch.empty(max(lib().nsb_render_workspace_bytes(self.chunk, self.S), 1), dtype=torch.uint8, device=dev)
        # gradient bucket: grids first, then one flat canonical segment per trainable decoder
        self.grid_grad_kinds = [k for k in self.kinds if GRID_KEYS[k] in grid_grads]
        self.dec_grad_kinds = [k for k in self.kinds if DEC_NAMES[k] in decoder_grads]
        leaves = [self.grids[k] for k in self.grid_grad_kinds]
        self.dw_holders = [torch.empty(lib().nsb_flat_numel(k), device=dev) for k in self.dec_grad_kinds]
        self.bucket = GradBucket(leaves + self.dw_holders, dev, peers) if (leaves or self.dw_holders) else None
        self.scratch = None
        if self.dec_grad_kinds:
            nbytes = len(self.dec_grad_kinds) * lib().nsb_dweight_scratch_bytes(self.dec_grad_kinds[0], self.chunk * self.S)
            self.scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        # static packed decoder blobs, re-packed in place from the live parameters
        self.packed, self._pack_args = {}, {}
        for k in self.kinds:
            params = [p.detach() for p in canonical_params(getattr(decoders, DEC_NAMES[k]), k)]
            if any(not p.is_contiguous() for p in params):
                raise RuntimeError("decoder parameters must be contiguous")
            ptrs = torch.tensor([p.data_ptr() for p in params], dtype=torch.int64).to(dev)
            self.packed[k] = torch.empty(lib().nsb_packed_numel(k), device=dev)
            self._pack_args[k] = (ptrs, params, pack_table(k, dev))
            self._pack(k)
        self.loss_parts = torch.zeros(self.nchunk, dtype=torch.float64, device=dev) if self.nchunk > 1 else None
        self._calls = [self._build_call(i * self.chunk, min(self.chunk, N - i * self.chunk), i) for i in range(self.nchunk)]
        self._call = self._calls[0]
        self.graph = None
        self._side = None
        self._xside = None
        self._packed_dev = None
        self.graph_host = None           # capture_host(): H2D of the staged batch + iteration + D2H of the loss
        self._host_stage = self._host_keep = self._loss_host = self._keep_dev = None
        self._host_n = -1
        self._peer_max = bool(self.with_gt_depth and self._sharded() and peers is not None and self.bucket is not None)

    # -- plumbing -----------------------------------------------------------------------
    def _sharded(self):
        import torch.distributed as dist
        if self._sharded_flag is False:
            return False
        return dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1

    def _pack(self, k):
        ptrs, _, table = self._pack_args[k]
        check(lib().nsb_pack_weights(k, _ptr(ptrs), _ptr(table), _ptr(self.packed[k]), _stream()), "pack_weights")

    def _build_call(self, off=0, n=None, part=0):
        """The render call of rays [off, off + n) (one chunk; the whole batch by default)."""
        r, dev = self.r, self.dev
        n = self.N if n is None else n
        call = NsbRenderCall()
        call.rays_o, call.rays_d = self.rays_o[off:].data_ptr(), self.rays_d[off:].data_ptr()
        call.rays_o_stride = call.rays_d_stride = 3
        call.N = n
        call.gt_depth = self.gt_depth[off:].data_ptr() if self.with_gt_depth else 0
        call.keep = self.keep[off:].data_ptr()
        # sharded: the batch depth max is MAX-reduced over ranks by load(); chunked: _launch computes it over the whole
        # batch first; otherwise the render call computes it itself (inside the captured graph)
        call.depth_max = self.depth_max.data_ptr() if (self.with_gt_depth and (self._sharded() or self.nchunk > 1)) else 0
        call.N_samples, call.N_surface = r.N_samples, r.N_surface
        call.lindisp, call.stage, call.occupancy = int(r.lindisp), STAGES[self.stage], int(r.occupancy)
        call.t_uniform = _linspace(r.N_samples, torch.float32, dev).data_ptr()
        call.t_surface = _linspace(r.N_surface, torch.float64, dev).data_ptr() if r.N_surface > 0 else 0
        grads = {k: self.bucket.view_of(self.grids[k]) for k in self.grid_grad_kinds}
        call.scene = fused.make_scene(r.bound, self.grids, self.packed, grads, r.coarse_bound_enlarge)
        call.workspace, call.workspace_bytes = self.ws.data_ptr(), self.ws.numel()
        call.depth, call.var, call.color = self.depth[off:].data_ptr(), self.var[off:].data_ptr(), self.color[off:].data_ptr()
        call.g_depth, call.g_color = self.g_depth[off:].data_ptr(), self.g_color[off:].data_ptr()
        flags, mask = (GRAD_GRID if self.grid_grad_kinds else 0), 0
        for k, h in zip(self.dec_grad_kinds, self.dw_holders):
            call.d_weight[k] = self.bucket.view_of(h).data_ptr()
            mask |= 1 << k
        if mask:
            flags |= GRAD_WEIGHT
            call.scratch, call.scratch_bytes = self.scratch.data_ptr(), self.scratch.numel()
        if self.ray_grads:
            flags |= GRAD_POINT
            call.g_rays_o, call.g_rays_d = self.g_rays_o[off:].data_ptr(), self.g_rays_d[off:].data_ptr()
        call.flags, call.dweight_mask = flags, mask
        if self.bucket is not None and getattr(self.bucket, "marks", None) is not None:
            call.line_marks, call.line_marks_base = self.bucket.marks.data_ptr(), self.bucket.flat.data_ptr()
        # one struct serves the forward and the backward call: the forward may deposit the trained decoders'
        # layer inputs into the weight-gradient records (the backward then skips their recomputation)
        call.deposit_dweight_inputs = 1 if (mask and os.environ.get("NSB_DEPOSIT", "1") != "0") else 0
        # mapping: compositing + loss + their adjoint run as one kernel inside nsb_render_backward
        self.fused_loss = self.loss_kind == "mapping" and self.with_gt_depth and bool(r.occupancy)
        if self.fused_loss:
            call.gt_color = self.gt_color[off:].data_ptr()
            call.loss = self.loss.data_ptr() if self.nchunk == 1 else self.loss_parts[part:].data_ptr()
            call.w_color, call.use_color = float(self.w_color), int(self.use_color)
        return call

    def load(self, rays_o, rays_d, gt_depth, gt_color, keep=None, depth_max=None):
        """Copy one batch into the static input buffers (n <= N rays; the rest is masked off)."""
        n = rays_o.shape[0]
        if n > self.N:
            raise ValueError(f"batch of {n} rays exceeds the step's capacity {self.N}")
        if rays_o.is_cuda:
            src = [rays_o.float().contiguous(), rays_d.float().contiguous(), gt_depth.reshape(-1).float().contiguous(),
                   gt_color.reshape(-1, 3).float().contiguous(), None if keep is None else keep.to(torch.uint8).contiguous()]
        else:
            # host batch: asynchronous H2D copies straight into the static buffers, then ONE kernel that
            # masks the tail off (no source pointers: the live entries are already in place)
            self.rays_o[:n].copy_(rays_o, non_blocking=True)
            self.rays_d[:n].copy_(rays_d, non_blocking=True)
            self.gt_depth[:n].copy_(gt_depth.reshape(-1), non_blocking=True)
            self.gt_color[:n].copy_(gt_color.reshape(-1, 3), non_blocking=True)
            if keep is not None:
                self.keep[:n].copy_(keep, non_blocking=True)
            src = [None, None, None, None, None if keep is None else self.keep]
        check(lib().nsb_load_batch(_ptr(src[0]), _ptr(src[1]), _ptr(src[2]), _ptr(src[3]), _ptr(src[4]), n, self.N,
                                   _ptr(self.rays_o), _ptr(self.rays_d), _ptr(self.gt_depth), _ptr(self.gt_color),
                                   _ptr(self.keep), _stream()), "load_batch")
        self._peer_max = False
        if self.with_gt_depth and self._sharded():
            if depth_max is not None:
                self.depth_max.copy_(depth_max.reshape(1), non_blocking=True)
            elif self.bucket is not None and self.bucket.peers is not None:
                self._peer_max = True        # nsb_depth_max_peers at the head of _launch (inside the graph)
            else:
                check(lib().nsb_depth_max(_ptr(self.gt_depth), _ptr(self.keep), self.N, _ptr(self.depth_max), _stream()),
                      "depth_max")
                import torch.distributed as dist
                dist.all_reduce(self.depth_max, op=dist.ReduceOp.MAX, group=self.group)

    @staticmethod
    def pack_batch(rays_o, rays_d, gt_depth, gt_color, pin=True):
        """One [n,10] fp32 host array (rays_o 3, rays_d 3, gt_depth 1, gt_color 3 per ray) for
        ``load_packed``: a producer on the host hands a batch over with a single H2D copy."""
        p = torch.cat([rays_o.float().reshape(-1, 3).cpu(), rays_d.float().reshape(-1, 3).cpu(),
                       gt_depth.float().reshape(-1, 1).cpu(), gt_color.float().reshape(-1, 3).cpu()], 1).contiguous()
        return p.pin_memory() if pin and torch.cuda.is_available() else p

    def load_packed(self, packed, keep=None):
        """``load`` from one packed [n,10] tensor (host, ideally pinned, or device): one copy + one kernel."""
        n = packed.shape[0]
        if n > self.N or packed.shape[1] != 10 or packed.dtype != torch.float32:
            raise ValueError("load_packed wants a float32 [n<=N, 10] tensor")
        if self._sharded() and not self._peer_max:
            raise NotImplementedError("load_packed on a sharded step needs the peer-memory depth max")
        if packed.is_cuda:
            dev = packed.contiguous()
        else:
            if self._packed_dev is None:
                self._packed_dev = torch.empty(self.N, 10, device=self.dev)
            dev = self._packed_dev[:n]
            dev.copy_(packed, non_blocking=True)
        kp = None if keep is None else keep.to(self.dev, torch.uint8, non_blocking=True).contiguous()
        check(lib().nsb_load_batch_packed(_ptr(dev), _ptr(kp), n, self.N, _ptr(self.rays_o), _ptr(self.rays_d),
                                          _ptr(self.gt_depth), _ptr(self.gt_color), _ptr(self.keep), _stream()),
              "load_batch_packed")

    def _launch(self, exchange=True):
        """memset(bucket) -> re-pack trainable decoders -> forward -> loss -> backward
        [-> peer-memory all-reduce of the bucket when it lives in symmetric memory]."""
        s = _stream()
        # the bucket memset (48 MB at Replica scale) has no consumer before the backward: it runs on a
        # side branch, under the forward
        joined = None
        if self.bucket is not None:
            main = torch.cuda.current_stream()
            if self._side is None:
                self._side = torch.cuda.Stream()
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side):
                self.bucket.zero()
            joined = self._side
        for k in self.dec_grad_kinds:
            self._pack(k)
        call = self._call
        if self._peer_max:                       # batch depth max over all ranks, through peer memory
            check(lib().nsb_depth_max_peers(C.byref(self.bucket._px), _ptr(self.gt_depth), _ptr(self.keep), self.N,
                                            _ptr(self.depth_max), s), "depth_max_peers")
        if self.nchunk > 1:
            if not self._sharded():              # the batch-global depth max, once for all chunks
                check(lib().nsb_depth_max(_ptr(self.gt_depth), _ptr(self.keep), self.N, _ptr(self.depth_max), s), "depth_max")
            for i, c in enumerate(self._calls):
                check(lib().nsb_render_forward(C.byref(c), s), "render_forward")
                if i == 0 and joined is not None:
                    torch.cuda.current_stream().wait_stream(joined)
                check(lib().nsb_render_backward(C.byref(c), s), "render_backward")
            torch.sum(self.loss_parts, dim=0, keepdim=True, out=self.loss)
            if exchange and self.bucket is not None and self.bucket.peers is not None:
                self.bucket.all_reduce()
            return
        check(lib().nsb_render_forward(C.byref(call), s), "render_forward")
        if joined is not None:
            torch.cuda.current_stream().wait_stream(joined)
        gd = _ptr(self.gt_depth)
        if self.fused_loss:
            pass
        elif self.loss_kind == "mapping":
            check(lib().nsb_mapping_loss(_ptr(self.depth), _ptr(self.color), gd, _ptr(self.gt_color), _ptr(self.keep),
                                         self.N, float(self.w_color), int(self.use_color), _ptr(self.loss),
                                         _ptr(self.g_depth), _ptr(self.g_color), s), "mapping_loss")
        else:
            check(lib().nsb_tracking_loss(_ptr(self.depth), _ptr(self.var), _ptr(self.color), gd, _ptr(self.gt_color),
                                          _ptr(self.keep), self.N, float(self.w_color), int(self.use_color),
                                          int(self.handle_dynamic), _ptr(self.loss), _ptr(self.g_depth),
                                          _ptr(self.g_color), s), "tracking_loss")
        # Sharded with weight gradients: the grid gradients are complete when the decoder backward retires, the
        # weight-gradient kernels (SM-bound) follow.  The exchange is bound by NVLink round trips, not by SMs: its grid
        # segment starts right away on a side stream with a few CTAs, beside the weight-gradient kernels on the other
        # SMs; only the small decoder segment is exchanged after the join.
        overlap = (exchange and self.bucket is not None and self.bucket.peers is not None and self.bucket.sparse
                   and bool(self.dec_grad_kinds) and call.deposit_dweight_inputs and self.bucket.n_grid > 0
                   and os.environ.get("NSB_X_OVERLAP", "0") == "1")
        if not overlap:
            check(lib().nsb_render_backward(C.byref(call), s), "render_backward")
            if exchange and self.bucket is not None and self.bucket.peers is not None:
                self.bucket.all_reduce()
            return
        side_ctas = int(os.environ.get("NSB_X_SIDE_CTAS", "20"))
        sms = torch.cuda.get_device_properties(self.dev).multi_processor_count
        main = torch.cuda.current_stream()
        if self._xside is None:
            self._xside = torch.cuda.Stream()
        call.dweight_phase = 1
        check(lib().nsb_render_backward(C.byref(call), s), "render_backward")
        self._xside.wait_stream(main)
        with torch.cuda.stream(self._xside):
            self.bucket.all_reduce_range(0, self.bucket.n_grid, side_ctas)
        call.dweight_phase, call.dweight_max_ctas = 2, max(sms - side_ctas, 1)
        check(lib().nsb_render_backward(C.byref(call), s), "render_backward (weight gradients)")
        call.dweight_phase, call.dweight_max_ctas = 0, 0
        main.wait_stream(self._xside)
        n = self.bucket.flat.numel()
        self.bucket.all_reduce_range(self.bucket.n_grid, n - self.bucket.n_grid)

    def capture(self):
        """Record the iteration into a CUDA graph (call once after a warm-up ``run``)."""
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        with graph_capture(g):
            self._launch()
        self.graph = g

    def capture_host(self):
        """Record a whole host-to-host step into ONE graph: copy of the staged batch (pinned host memory) to the
        device, ``nsb_load_batch_packed``, the iteration, copy of the loss back to pinned host memory.  Call after
        a warm-up ``run``; then ``step_from_host`` costs the host one memcpy into the staging buffer, one graph
        launch and one stream synchronisation per iteration."""
        if self._sharded() and (self.bucket is None or self.bucket.peers is None or not self._peer_max):
            raise NotImplementedError("capture_host on a sharded step needs the peer-memory exchange")
        if self._host_stage is None:
            self._host_stage = torch.zeros(self.N, 10, dtype=torch.float32).pin_memory()
            self._host_keep = torch.zeros(self.N, dtype=torch.uint8).pin_memory()
            self._loss_host = torch.zeros(1, dtype=torch.float64).pin_memory()
            self._keep_dev = torch.zeros(self.N, dtype=torch.uint8, device=self.dev)
            if self._packed_dev is None:
                self._packed_dev = torch.empty(self.N, 10, device=self.dev)
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        with graph_capture(g):
            self._packed_dev.copy_(self._host_stage, non_blocking=True)
            self._keep_dev.copy_(self._host_keep, non_blocking=True)
            check(lib().nsb_load_batch_packed(_ptr(self._packed_dev), _ptr(self._keep_dev), self.N, self.N,
                                              _ptr(self.rays_o), _ptr(self.rays_d), _ptr(self.gt_depth),
                                              _ptr(self.gt_color), _ptr(self.keep), _stream()), "load_batch_packed")
            self._launch()
            self._loss_host.copy_(self.loss, non_blocking=True)
        self.graph_host = g

    def step_from_host(self, packed, keep=None):
        """One iteration on a packed [n,10] float32 HOST batch (``pack_batch``), loss returned as a Python float;
        gradients are in ``self.bucket``.  Needs ``capture_host``."""
        if self.graph_host is None:
            raise RuntimeError("step_from_host needs capture_host() first")
        n = packed.shape[0]
        if n > self.N or packed.shape[1] != 10 or packed.dtype != torch.float32 or packed.is_cuda:
            raise ValueError("step_from_host wants a float32 [n<=N, 10] host tensor")
        self._host_stage[:n].copy_(packed)
        if keep is not None:
            self._host_keep[:n].copy_(keep.to(torch.uint8))
            self._host_keep[n:] = 0
            self._host_n = -1
        elif n != self._host_n:
            self._host_keep[:n] = 1
            self._host_keep[n:] = 0
            self._host_n = n
        self.graph_host.replay()
        torch.cuda.current_stream().synchronize()
        return float(self._loss_host[0])

    def run(self, all_reduce=True):
        """One iteration on the loaded batch.  Returns the loss as a [1] fp64 device tensor (this
        rank's partial sum when sharded); gradients are in ``self.bucket`` (summed over ranks)."""
        if self.graph is not None:
            self.graph.replay()
        else:
            self._launch()
        if all_reduce and self.bucket is not None and self.bucket.peers is None and self._sharded():
            self.bucket.all_reduce(self.group)         # NCCL; the peer-memory exchange is part of _launch
        return self.loss

    def live_tiles(self, tile_points=16):
        """Diagnostic (after ``run``): share of the ``tile_points``-point tiles whose incoming gradient is not exactly
        zero, as {'occupancy': f, 'rgb': f}: what the decoder backward cannot skip on this batch / fixture."""
        counts = torch.zeros(3, dtype=torch.int64, device=self.dev)
        check(lib().nsb_render_live_tiles(C.byref(self._call), int(tile_points), _ptr(counts), _stream()), "live_tiles")
        n, occ, rgb = (int(x) for x in counts.tolist())
        return {"tile_points": int(tile_points), "occupancy": occ / max(n, 1), "rgb": rgb / max(n, 1)}

    def decoder_grads(self, kind):
        """Per-parameter gradient views of a trainable decoder, in state_dict order of
        ``canonical_params``."""
        flat = self.bucket.view_of(self.dw_holders[self.dec_grad_kinds.index(kind)])
        out, off = [], 0
        for p in self._pack_args[kind][1]:
            out.append(flat[off:off + p.numel()].view(p.shape))
            off += p.numel()
        return out

    def assign_grads(self, c):
        """Set ``.grad`` of the grid leaves in ``c`` and of the trainable decoder parameters."""
        for k in self.grid_grad_kinds:
            c[GRID_KEYS[k]].grad = self.bucket.view_of(self.grids[k])
        for k in self.dec_grad_kinds:
            params = canonical_params(getattr(self.decoders, DEC_NAMES[k]), k)
            for p, g in zip(params, self.decoder_grads(k)):
                p.grad = g


# ---------------------------------------------------------------------------------------
# the fused tracking iteration
# ---------------------------------------------------------------------------------------
class TrackingStep:
    """One whole camera-tracking iteration on the device (src/Tracker.py:78-135 + the optimiser
    and candidate bookkeeping of Tracker.run, :232-275): draw pixels -> camera tensor to rays +
    box pre-filter -> render -> tracking loss -> backward to the rays -> pose gradient -> Adam on
    the 7 camera parameters -> keep the best candidate.  No host round trip per iteration (the
    reference reads ``loss.item()`` every iteration), so ``iterate(n)`` replays one CUDA graph n
    times and the host reads the result once per frame.

    tracker : object with the attributes of nice_slam_annotation_b200.Tracker.Tracker
              (renderer, decoders, c, bound, H, W, fx, fy, cx, cy, ignore_edge_*, w_color_loss, ...)
    n_rays  : tracking pixels per iteration (cfg['tracking']['pixels'])
    """

    def __init__(self, tracker, n_rays, lr=None, separate_lr=None, betas=(0.9, 0.999), eps=1e-8, adam=True):
        t = tracker
        self.t, self.N = t, int(n_rays)
        self.adam = bool(adam)          # False: stop at d(camera tensor); the caller's optimiser takes the step
        self._api_graphs = {}
        dev = next(t.decoders.parameters()).device
        self.dev = dev
        self.step = FusedStep(t.renderer, t.decoders, t.c, "color", n_rays, loss="tracking",
                              w_color=t.w_color_loss, use_color=t.use_color_in_tracking,
                              handle_dynamic=t.handle_dynamic, ray_grads=True, sharded=False)   # tracking: replicas only
        lr = t.cam_lr if lr is None else lr
        separate_lr = t.seperate_LR if separate_lr is None else separate_lr
        # Tracker.py:232-239: translation at cam_lr, quaternion at 0.2 cam_lr when seperate_LR
        self.lr_t, self.lr_q = float(lr), float(lr) * (0.2 if separate_lr else 1.0)
        self.betas, self.eps = betas, float(eps)
        H0, W0 = t.ignore_edge_H, t.ignore_edge_W
        self.crop = (H0, t.H - H0, W0, t.W - W0)
        self.intr = tuple(float(x) for x in (t.fx, t.fy, t.cx, t.cy))
        self.bound6 = fused._bound6(t.bound)
        self.depth_img = torch.zeros(t.H, t.W, device=dev)
        self.color_img = torch.zeros(t.H, t.W, 3, device=dev)
        self.indices = torch.zeros(self.N, dtype=torch.int64, device=dev)
        self.cam = torch.zeros(7, device=dev)
        self.d_cam = torch.zeros(7, device=dev)
        self.m = torch.zeros(7, device=dev)
        self.v = torch.zeros(7, device=dev)
        self.count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.best_loss = torch.full((1,), float("inf"), dtype=torch.float64, device=dev)
        self.best_cam = torch.zeros(7, device=dev)
        self.first_loss = torch.zeros(1, dtype=torch.float64, device=dev)
        self.draw = True            # False: iterate on the indices the caller wrote into self.indices
        self.graph = None

    def load_frame(self, gt_color, gt_depth):
        self.depth_img.copy_(gt_depth, non_blocking=True)
        self.color_img.copy_(gt_color, non_blocking=True)       # fp64 dataset colour is cast here

    def reset(self, camera_tensor):
        """New frame / new initial pose: camera tensor [quat(4), t(3)], fresh optimiser state
        (the reference builds a new Adam per frame, Tracker.py:232-242)."""
        self.cam.copy_(camera_tensor.detach().reshape(7), non_blocking=True)
        for x in (self.m, self.v, self.count):
            x.zero_()
        self.best_loss.fill_(float("inf"))
        self.best_cam.copy_(self.cam)

    def _launch(self):
        s, st, L = _stream(), self.step, lib()
        H0, H1, W0, W1 = self.crop
        fx, fy, cx, cy = self.intr
        if self.draw:                                            # common.py:99
            torch.randint(0, (H1 - H0) * (W1 - W0), (self.N,), device=self.dev, out=self.indices)
        check(L.nsb_sample_rays(_ptr(self.indices), self.N, H0, H1, W0, W1, self.t.W, fx, fy, cx, cy, None,
                                _ptr(self.cam), _ptr(self.depth_img), _ptr(self.color_img), 0, self.bound6,
                                _ptr(st.rays_o), _ptr(st.rays_d), _ptr(st.gt_depth), _ptr(st.gt_color),
                                _ptr(st.keep), s), "sample_rays")
        st._launch()
        if self.adam:       # pose gradient + Adam on the camera tensor + best-candidate rule: one launch
            check(L.nsb_pose_grad_adam(_ptr(st.g_rays_o), _ptr(st.g_rays_d), _ptr(self.indices), _ptr(st.keep), self.N, H0, W0,
                                       W1 - W0, fx, fy, cx, cy, _ptr(self.cam), _ptr(self.d_cam), _ptr(self.m), _ptr(self.v),
                                       _ptr(self.count), self.lr_q, self.lr_t, self.betas[0], self.betas[1], self.eps,
                                       _ptr(st.loss), _ptr(self.best_loss), _ptr(self.best_cam), s), "pose_grad_adam")
        else:
            check(L.nsb_pose_grad(_ptr(st.g_rays_o), _ptr(st.g_rays_d), _ptr(self.indices), _ptr(st.keep), self.N, H0, W0,
                                  W1 - W0, fx, fy, cx, cy, _ptr(self.cam), _ptr(self.d_cam), s), "pose_grad")

    def capture(self):
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        with graph_capture(g):
            self._launch()
        self.graph = g

    def iterate(self, n=1):
        for _ in range(n):
            if self.graph is not None:
                self.graph.replay()
            else:
                self._launch()

    def iterate_api(self, draw):
        """One iteration up to ``d_cam`` for Tracker.optimize_cam_in_batch: a graph per draw flag (the pixel draw is a
        node of the graph), captured on first use after one plain launch."""
        g = self._api_graphs.get(bool(draw))
        self.draw = bool(draw)
        if g is None:
            keep = self.indices.clone()
            rng = torch.cuda.get_rng_state(self.dev) if draw else None
            self._launch()                              # warm-up outside the capture, RNG- and index-neutral
            if rng is not None:
                torch.cuda.set_rng_state(rng, self.dev)
            self.indices.copy_(keep)
            g = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            with graph_capture(g):
                self._launch()
            self._api_graphs[bool(draw)] = g
        g.replay()

    def result(self):
        """(candidate camera tensor [7], its loss) -- the one host read of the frame."""
        return self.best_cam.clone(), float(self.best_loss.item())
