"""Host side of the fused render path: tensors in, C-ABI calls, tensors out.

Everything here is plumbing (device memory, streams, autograd bookkeeping); every value
is produced by the kernels in ``csrc/`` through ``include/nice_slam_b200.h``.  The single
autograd node ``_RenderFn`` replaces the ~620 ATen kernels the reference launches for one
``Renderer.render_batch_ray`` forward + backward.
"""
import ctypes as C
import os
import weakref

import numpy as np
import torch

from . import _lib
from ._lib import (DEC_COARSE, DEC_COLOR, DEC_FINE, DEC_MIDDLE, GRAD_GRID, GRAD_POINT, GRAD_WEIGHT,
                   STAGES, NsbGrid, NsbPoints, NsbRenderCall, NsbScene, check, lib)

GRID_KEYS = ("grid_coarse", "grid_middle", "grid_fine", "grid_color")
DEC_NAMES = ("coarse_decoder", "middle_decoder", "fine_decoder", "color_decoder")
# decoders a stage runs (src/conv_onet/models/decoder.py:336-370), in kernel task order
STAGE_DECODERS = {"coarse": (DEC_COARSE,), "middle": (DEC_MIDDLE,), "fine": (DEC_MIDDLE, DEC_FINE),
                  "color": (DEC_MIDDLE, DEC_FINE, DEC_COLOR)}


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor: the fused render path has no CPU fallback")


# ---------------------------------------------------------------------------------------
# decoder parameters
# ---------------------------------------------------------------------------------------
def canonical_params(dec, kind):
    """Parameters of one decoder in the canonical order of include/nice_slam_b200.h."""
    ps = [dec.pts_linears[i].weight for i in range(5)] + [dec.pts_linears[i].bias for i in range(5)]
    if kind != DEC_COARSE:
        ps += [dec.fc_c[i].weight for i in range(5)] + [dec.fc_c[i].bias for i in range(5)]
    ps += [dec.output_linear.weight, dec.output_linear.bias]
    if kind != DEC_COARSE:
        ps.append(dec.embedder._B)
    return ps


_tables = {}          # (kind, device) -> int32 device tensor
_pack_cache = weakref.WeakKeyDictionary()   # decoder module -> (key, packed, ptr tensor)


def pack_table(kind, device):
    key = (kind, str(device))
    if key not in _tables:
        n = lib().nsb_packed_numel(kind)
        tab = np.empty(n, dtype=np.int32)
        check(lib().nsb_pack_table_host(kind, tab.ctypes.data_as(C.c_void_p)), "pack_table")
        _tables[key] = torch.from_numpy(tab).to(device)
    return _tables[key]


def pack_params(kind, params):
    """Packed blob (fp32 device tensor) for a list of parameter tensors in canonical order."""
    dev = params[0].device
    _require_cuda(params[0], "decoder parameters")
    n_expected = lib().nsb_param_count(kind)
    if len(params) != n_expected:
        raise RuntimeError(f"decoder kind {kind}: expected {n_expected} parameters, got {len(params)}")
    keep = []
    for i, p in enumerate(params):
        if p.dtype != torch.float32 or p.numel() != lib().nsb_param_numel(kind, i):
            raise RuntimeError(f"decoder kind {kind}: parameter {i} has shape {tuple(p.shape)} / {p.dtype}")
        keep.append(p.detach() if p.is_contiguous() else p.detach().contiguous())
    ptrs = torch.tensor([p.data_ptr() for p in keep], dtype=torch.int64).to(dev, non_blocking=False)
    packed = torch.empty(lib().nsb_packed_numel(kind), dtype=torch.float32, device=dev)
    check(lib().nsb_pack_weights(kind, _ptr(ptrs), _ptr(pack_table(kind, dev)), _ptr(packed), _stream()),
          "pack_weights")
    return packed, (ptrs, keep)


def packed_decoder(dec, kind):
    """Packed blob of a decoder module, re-packed only when a parameter changed."""
    params = canonical_params(dec, kind)
    key = tuple((p.data_ptr(), p._version) for p in params)
    hit = _pack_cache.get(dec)
    if hit is not None and hit[0] == key:
        return hit[1], params
    packed, hold = pack_params(kind, params)
    _pack_cache[dec] = (key, packed, hold)
    return packed, params


# ---------------------------------------------------------------------------------------
# structs
# ---------------------------------------------------------------------------------------
_layout_cache = {}        # (data_ptr, version, shape, device) -> channels-last copy; small LRU
_LAYOUT_CACHE_MAX = 8


def grid_to_kernel_layout(g):
    """[1,32,Z,Y,X] fp32 -> same logical tensor, channels-last in memory.

    No copy when the grid already has the kernels' layout (scene.grid_init, shared.prepare_shared, the
    package's Mapper).  A channels-FIRST grid -- what the reference's own ``NICE_SLAM.grid_init`` and its
    ``val[mask] = val_grad`` non-leaf produce (NICE_SLAM.py:243-320, Mapper.py:499) -- has to be re-laid out:
    * without autograd history (the tracker's detached copies, forward-only callers) the converted copy is
      cached by (storage address, version counter), so the copy is paid once per ``update_para_from_mapping``
      and not once per render call;
    * with autograd history (the reference's mapper driving this Renderer) the conversion must stay a
      differentiable op: one layout copy per call forward and one for its gradient (0.1 ms per 48 MB grid on
      B200) -- use ``shared.prepare_shared`` / this package's Mapper to avoid it."""
    if g.dim() != 5 or g.shape[0] != 1 or g.shape[1] != 32:
        raise RuntimeError(f"feature grid must be [1,32,Z,Y,X], got {tuple(g.shape)}")
    if g.dtype != torch.float32:
        raise RuntimeError("feature grid must be float32")
    _require_cuda(g, "feature grid")
    if g.is_contiguous(memory_format=torch.channels_last_3d):
        return g
    if g.requires_grad and torch.is_grad_enabled():
        return g.contiguous(memory_format=torch.channels_last_3d)
    key = (g.data_ptr(), g._version, tuple(g.shape), tuple(g.stride()), str(g.device))
    hit = _layout_cache.get(key)
    if hit is not None:
        return hit
    out = g.detach().contiguous(memory_format=torch.channels_last_3d)
    if len(_layout_cache) >= _LAYOUT_CACHE_MAX:
        _layout_cache.pop(next(iter(_layout_cache)))
    _layout_cache[key] = out
    return out


def _bound6(bound):
    b = torch.as_tensor(bound, dtype=torch.float64, device="cpu").reshape(3, 2)
    return (C.c_double * 6)(*b.reshape(-1).tolist())


def make_scene(bound, grids, packed, grads=None, coarse_bound_enlarge=2.0, coarse_bound=None):
    """NsbScene from {kind: grid tensor}, {kind: packed tensor}, {kind: grad tensor}."""
    sc = NsbScene()
    b = torch.as_tensor(bound, dtype=torch.float64, device="cpu").reshape(3, 2)
    cb = b * coarse_bound_enlarge if coarse_bound is None else \
        torch.as_tensor(coarse_bound, dtype=torch.float64, device="cpu").reshape(3, 2)
    for a in range(3):
        for k in range(2):
            sc.bound[a][k] = float(b[a, k])
            sc.coarse_bound[a][k] = float(cb[a, k])
    for kind in range(4):
        g = grids.get(kind)
        if g is not None:
            sc.grid[kind].data = g.data_ptr()
            gr = None if grads is None else grads.get(kind)
            sc.grid[kind].grad = 0 if gr is None else gr.data_ptr()
            sc.grid[kind].Z, sc.grid[kind].Y, sc.grid[kind].X = g.shape[2], g.shape[3], g.shape[4]
        pk = packed.get(kind)
        sc.packed[kind] = 0 if pk is None else pk.data_ptr()
    return sc


def ray_points(rays_o, rays_d, z):
    p = NsbPoints()
    p.rays_o, p.rays_d = rays_o.data_ptr(), rays_d.data_ptr()
    p.rays_o_stride, p.rays_d_stride = rays_o.stride(0), rays_d.stride(0)
    p.z = 0 if z is None else z.data_ptr()
    p.N = rays_o.shape[0]
    p.S = 0 if z is None else z.shape[1]
    p.P = p.N * p.S
    return p


def explicit_points(pts):
    p = NsbPoints()
    if pts.dtype == torch.float64:
        p.p64 = pts.data_ptr()
    elif pts.dtype == torch.float32:
        p.p32 = pts.data_ptr()
    else:
        raise RuntimeError(f"points must be float32 or float64, got {pts.dtype}")
    p.P = pts.shape[0]
    return p


def _row_major_rays(t):
    """[N,3] fp32 with unit inner stride; a stride-0 expand (common.py:88) is kept as is."""
    if t.dtype != torch.float32:
        t = t.float()
    if t.dim() != 2 or t.shape[1] != 3:
        raise RuntimeError(f"rays must be [N,3], got {tuple(t.shape)}")
    if t.shape[0] > 0 and (t.stride(1) != 1 or t.stride(0) not in (0, 3)):
        t = t.contiguous()
    return t


_linspace_cache = {}


def _linspace(n, dtype, device):
    key = (n, dtype, str(device))
    if key not in _linspace_cache:
        v = torch.linspace(0., 1., steps=n)          # fp32 on the CPU, as the reference draws it
        if dtype == torch.float64:
            v = v.double()
        _linspace_cache[key] = v.to(device)
    return _linspace_cache[key]


# ---------------------------------------------------------------------------------------
# kernels as functions of tensors
# ---------------------------------------------------------------------------------------
def sample_z(rays_o, rays_d, gt_depth, bound, N_samples, N_surface, lindisp=False, u=None):
    """Sorted fp64 sample depths [N,S] (src/utils/Renderer.py:116-224)."""
    dev = rays_o.device
    N = rays_o.shape[0]
    if gt_depth is None:
        N_surface = 0
    S = N_samples + N_surface
    z = torch.empty(N, S, dtype=torch.float64, device=dev)
    if N == 0:
        return z
    dmax = None
    if gt_depth is not None:
        gt_depth = gt_depth.reshape(-1).float().contiguous()
        dmax = torch.empty(1, dtype=torch.float32, device=dev)
        check(lib().nsb_depth_max(_ptr(gt_depth), None, N, _ptr(dmax), _stream()), "depth_max")
    if u is not None:
        u = u.to(device=dev, dtype=torch.float32).contiguous()
    rays = ray_points(rays_o, rays_d, None)
    ts = _linspace(N_surface, torch.float64, dev) if N_surface > 0 else None
    check(lib().nsb_sample_z(C.byref(rays), _ptr(gt_depth), _bound6(bound), N_samples, N_surface,
                             int(bool(lindisp)), _ptr(_linspace(N_samples, torch.float32, dev)), _ptr(ts),
                             _ptr(u), _ptr(dmax), _ptr(z), _stream()), "sample_z")
    return z


def box_prefilter(rays_o, rays_d, gt_depth, bound):
    """[N] uint8: 1 for the rays the reference keeps (src/Tracker.py:100-107, src/Mapper.py:587-594)."""
    _require_cuda(rays_o, "rays_o")
    N = rays_o.shape[0]
    keep = torch.empty(N, dtype=torch.uint8, device=rays_o.device)
    check(lib().nsb_box_prefilter(_ptr(rays_o.detach().float().contiguous()), _ptr(rays_d.detach().float().contiguous()),
                                  _ptr(gt_depth.detach().reshape(-1).float().contiguous()), N, _bound6(bound),
                                  _ptr(keep), _stream()), "box_prefilter")
    return keep


class _RegulationPointsFn(torch.autograd.Function):
    """Renderer.py:360-394 as one kernel; backward carries d(points) to the rays (bundle adjustment)."""

    @staticmethod
    def forward(ctx, rays_o, rays_d, gt_depth, N_samples, u):
        dev = rays_o.device
        N = rays_o.shape[0]
        ro, rd = rays_o.detach().float().contiguous(), rays_d.detach().float().contiguous()
        gd = gt_depth.detach().reshape(-1).float().contiguous()
        u = u.to(device=dev, dtype=torch.float32).contiguous()
        pts = torch.empty(N * N_samples, 3, dtype=torch.float32, device=dev)
        z = torch.empty(N, N_samples, dtype=torch.float32, device=dev)
        check(lib().nsb_regulation_points(_ptr(ro), _ptr(rd), _ptr(gd), N, N_samples,
                                          _ptr(_linspace(N_samples, torch.float32, dev)), _ptr(u), _ptr(pts), _ptr(z),
                                          _stream()), "regulation_points")
        ctx.save_for_backward(z)
        ctx.mark_non_differentiable(z)
        return pts, z

    @staticmethod
    def backward(ctx, g_pts, _gz):
        (z,) = ctx.saved_tensors
        N, S = z.shape
        g = g_pts.float().contiguous()
        d_o = torch.empty(N, 3, dtype=torch.float32, device=z.device) if ctx.needs_input_grad[0] else None
        d_d = torch.empty(N, 3, dtype=torch.float32, device=z.device) if ctx.needs_input_grad[1] else None
        if N and (d_o is not None or d_d is not None):
            check(lib().nsb_regulation_points_backward(_ptr(g), _ptr(z), N, S, _ptr(d_o), _ptr(d_d), _stream()),
                  "regulation_points_backward")
        return d_o, d_d, None, None, None


def regulation_points(rays_o, rays_d, gt_depth, N_samples, u, return_z=False):
    """Points [N*N_samples,3] of the iMAP free-space regulariser (src/utils/Renderer.py:360-394): jittered depths in
    [0, 0.85 gt_depth]; u [N,N_samples] uniforms.  Differentiable towards rays_o / rays_d."""
    _require_cuda(rays_o, "rays_o")
    pts, z = _RegulationPointsFn.apply(rays_o, rays_d, gt_depth, int(N_samples), u)
    return (pts, z) if return_z else pts


class _Planes:
    """Raw decoder outputs of one stage."""

    def __init__(self, stage, P, device):
        self.occ0 = self.occ1 = self.rgb = None
        kinds = STAGE_DECODERS[stage]
        if DEC_COARSE in kinds or DEC_MIDDLE in kinds:
            self.occ0 = torch.empty(P, dtype=torch.float32, device=device)
        if DEC_FINE in kinds:
            self.occ1 = torch.empty(P, dtype=torch.float32, device=device)
        if DEC_COLOR in kinds:
            self.rgb = torch.empty(P, 4, dtype=torch.float32, device=device)


def decoders_forward(points, scene, stage, P, device):
    planes = _Planes(stage, P, device)
    if P > 0:
        check(lib().nsb_decoders_forward(C.byref(points), C.byref(scene), STAGES[stage], _ptr(planes.occ0),
                                         _ptr(planes.occ1), _ptr(planes.rgb), _stream()), "decoders_forward")
    return planes


def _gather_scene(stage, bound, c, decoders, enlarge):
    """Grids (kernel layout) and packed decoders a stage needs."""
    kinds = STAGE_DECODERS[stage]
    grids, packed, params = {}, {}, {}
    for kind in kinds:
        grids[kind] = grid_to_kernel_layout(c[GRID_KEYS[kind]])
        packed[kind], params[kind] = packed_decoder(getattr(decoders, DEC_NAMES[kind]), kind)
    return grids, packed, params


def _render_call(meta, rays_o, rays_d, grids, grads, ws, depth, var, color):
    """NsbRenderCall of one render_batch_ray (forward fields; the backward adds its own)."""
    dev = rays_o.device
    call = NsbRenderCall()
    call.rays_o, call.rays_d = rays_o.data_ptr(), rays_d.data_ptr()
    call.rays_o_stride, call.rays_d_stride = rays_o.stride(0), rays_d.stride(0)
    call.N = rays_o.shape[0]
    gd, keep, u = meta["gt_depth"], meta.get("keep"), meta.get("u")
    call.gt_depth = 0 if gd is None else gd.data_ptr()
    call.keep = 0 if keep is None else keep.data_ptr()
    dmax = meta.get("depth_max")
    call.depth_max = 0 if dmax is None else dmax.data_ptr()
    call.N_samples, call.N_surface = meta["N_samples"], meta["N_surface"]
    call.lindisp, call.stage, call.occupancy = int(meta["lindisp"]), STAGES[meta["stage"]], int(meta["occupancy"])
    call.t_uniform = _linspace(meta["N_samples"], torch.float32, dev).data_ptr()
    call.t_surface = _linspace(meta["N_surface"], torch.float64, dev).data_ptr() if meta["N_surface"] > 0 else 0
    call.u = 0 if u is None else u.data_ptr()
    call.scene = make_scene(meta["bound"], grids, meta["packed"], grads, meta["coarse_bound_enlarge"])
    call.workspace, call.workspace_bytes = ws.data_ptr(), ws.numel()
    call.depth, call.var, call.color = depth.data_ptr(), var.data_ptr(), color.data_ptr()
    return call


class _RenderFn(torch.autograd.Function):
    """render_batch_ray (src/utils/Renderer.py:95-262) as one autograd node.  Without
    importance resampling (the NICE configs) forward and backward are ONE library call each
    (nsb_render_forward / nsb_render_backward)."""

    @staticmethod
    def forward(ctx, meta, rays_o, rays_d, *tensors):
        stage, bound = meta["stage"], meta["bound"]
        kinds = STAGE_DECODERS[stage]
        n_grid = len(kinds)
        grids = dict(zip(kinds, tensors[:n_grid]))
        flat_params = tensors[n_grid:]
        packed = meta["packed"]
        dev = rays_o.device
        N = rays_o.shape[0]
        gt_depth = meta["gt_depth"]
        if meta["N_importance"] == 0:
            S = meta["N_samples"] + (meta["N_surface"] if gt_depth is not None else 0)
            depth = torch.empty(N, dtype=torch.float64, device=dev)
            var = torch.empty(N, dtype=torch.float64, device=dev)
            color = torch.empty(N, 3, dtype=torch.float32, device=dev)
            ws = torch.empty(max(lib().nsb_render_workspace_bytes(N, S), 1), dtype=torch.uint8, device=dev)
            if meta.get("u") is not None:
                meta["u"] = meta["u"].to(device=dev, dtype=torch.float32).contiguous()
            # decoders whose parameters want a gradient: the forward deposits their layer inputs into the weight-gradient
            # records (NsbRenderCall.deposit_dweight_inputs), the backward then runs the tcgen05 adjoint without
            # recomputation; the scratch lives from here to the backward
            ctx.dw_scratch, ctx.dw_mask = None, 0
            if N > 0:
                call = _render_call(meta, rays_o, rays_d, grids, None, ws, depth, var, color)
                need_param = ctx.needs_input_grad[3 + n_grid:]
                if torch.is_grad_enabled() and any(need_param) and os.environ.get("NSB_DEPOSIT", "1") != "0":
                    pi, mask = 0, 0
                    for k in kinds:
                        n = lib().nsb_param_count(k)
                        if any(need_param[pi:pi + n]):
                            mask |= 1 << k
                        pi += n
                    if mask:
                        ndw = bin(mask).count("1")
                        nbytes = ndw * max(lib().nsb_dweight_scratch_bytes(k, N * S) for k in kinds if mask & (1 << k))
                        ctx.dw_scratch, ctx.dw_mask = torch.empty(nbytes, dtype=torch.uint8, device=dev), mask
                        call.scratch, call.scratch_bytes = ctx.dw_scratch.data_ptr(), nbytes
                        call.flags, call.dweight_mask = GRAD_WEIGHT, mask
                        call.deposit_dweight_inputs = 1
                check(lib().nsb_render_forward(C.byref(call), _stream()), "render_forward")
            ctx.meta, ctx.kinds, ctx.n_grid, ctx.fused_ws, ctx.S = meta, kinds, n_grid, ws, S
            ctx.param_shapes = [p.shape for p in flat_params]
            ctx.save_for_backward(rays_o, rays_d, depth, var, color, *tensors[:n_grid])
            return depth, var, color
        ctx.fused_ws = None
        z = sample_z(rays_o, rays_d, gt_depth, bound, meta["N_samples"], meta["N_surface"], meta["lindisp"],
                     meta.get("u"))
        S = z.shape[1]
        scene = make_scene(bound, grids, packed, None, meta["coarse_bound_enlarge"])
        b6 = _bound6(bound)
        depth = torch.empty(N, dtype=torch.float64, device=dev)
        var = torch.empty(N, dtype=torch.float64, device=dev)
        color = torch.empty(N, 3, dtype=torch.float32, device=dev)
        weights = torch.empty(N, S, dtype=torch.float32, device=dev) if meta["N_importance"] > 0 else None
        pts = ray_points(rays_o, rays_d, z)
        planes = decoders_forward(pts, scene, stage, N * S, dev)
        if N > 0:
            check(lib().nsb_composite_forward(C.byref(pts), b6, STAGES[stage], int(meta["occupancy"]),
                                              _ptr(planes.occ0), _ptr(planes.occ1), _ptr(planes.rgb),
                                              _ptr(depth), _ptr(var), _ptr(color), _ptr(weights), _stream()),
                  "composite_forward")
        if meta["N_importance"] > 0 and N > 0:
            Ni = meta["N_importance"]
            u = meta.get("u_importance")
            if u is None:
                u, stride = _linspace(Ni, torch.float32, dev), 0
            else:
                u = u.to(device=dev, dtype=torch.float32).contiguous()
                stride = Ni
            z2 = torch.empty(N, S + Ni, dtype=torch.float64, device=dev)
            check(lib().nsb_importance_z(_ptr(z), _ptr(weights), N, S, Ni, _ptr(u), stride, None, _ptr(z2),
                                         _stream()), "importance_z")
            z, S = z2, S + Ni
            pts = ray_points(rays_o, rays_d, z)
            planes = decoders_forward(pts, scene, stage, N * S, dev)
            check(lib().nsb_composite_forward(C.byref(pts), b6, STAGES[stage], int(meta["occupancy"]),
                                              _ptr(planes.occ0), _ptr(planes.occ1), _ptr(planes.rgb),
                                              _ptr(depth), _ptr(var), _ptr(color), None, _stream()),
                  "composite_forward")
        ctx.meta = meta
        ctx.kinds = kinds
        ctx.planes = planes
        ctx.n_grid = n_grid
        ctx.param_shapes = [p.shape for p in flat_params]
        ctx.save_for_backward(rays_o, rays_d, z, depth, *tensors[:n_grid])
        return depth, var, color

    @staticmethod
    def _backward_fused(ctx, g_depth, g_var, g_color):
        meta, kinds = ctx.meta, ctx.kinds
        rays_o, rays_d, depth, var, color = ctx.saved_tensors[:5]
        grids = dict(zip(kinds, ctx.saved_tensors[5:]))
        dev, N = rays_o.device, rays_o.shape[0]
        P = N * ctx.S
        need = ctx.needs_input_grad          # (meta, rays_o, rays_d, grids..., params...)
        out = [None] * len(need)
        if N == 0:
            return tuple(out)
        need_rays = need[1] or need[2]
        need_param = need[3 + ctx.n_grid:]
        flags, grads = 0, {}
        for i, k in enumerate(kinds):
            if need[3 + i]:
                grads[k] = torch.zeros_like(grids[k], memory_format=torch.preserve_format)
                flags |= GRAD_GRID
        call = _render_call(meta, rays_o, rays_d, grids, grads, ctx.fused_ws, depth, var, color)
        hold = []
        if g_depth is not None:
            hold.append(g_depth.double().contiguous())
            call.g_depth = hold[-1].data_ptr()
        if g_var is not None:
            hold.append(g_var.double().contiguous())
            call.g_var = hold[-1].data_ptr()
        if g_color is not None:
            hold.append(g_color.float().contiguous())
            call.g_color = hold[-1].data_ptr()
        dw, dw_mask, pi = {}, 0, 0
        for k in kinds:
            n = lib().nsb_param_count(k)
            if any(need_param[pi:pi + n]):
                dw[k] = torch.zeros(lib().nsb_flat_numel(k), dtype=torch.float32, device=dev)
                dw_mask |= 1 << k
                call.d_weight[k] = dw[k].data_ptr()
            pi += n
        if dw_mask:
            flags |= GRAD_WEIGHT
            if ctx.dw_scratch is not None and ctx.dw_mask == dw_mask:      # the forward deposited into it
                call.scratch, call.scratch_bytes = ctx.dw_scratch.data_ptr(), ctx.dw_scratch.numel()
                call.deposit_dweight_inputs = 1
            else:
                nbytes = len(dw) * max(lib().nsb_dweight_scratch_bytes(k, P) for k in dw)
                hold.append(torch.empty(nbytes, dtype=torch.uint8, device=dev))
                call.scratch, call.scratch_bytes = hold[-1].data_ptr(), nbytes
        if need_rays:
            flags |= GRAD_POINT
            g_o = torch.empty(N, 3, dtype=torch.float32, device=dev)
            g_d = torch.empty(N, 3, dtype=torch.float32, device=dev)
            call.g_rays_o, call.g_rays_d = g_o.data_ptr(), g_d.data_ptr()
            out[1] = g_o if need[1] else None
            out[2] = g_d if need[2] else None
        call.flags, call.dweight_mask = flags, dw_mask
        check(lib().nsb_render_backward(C.byref(call), _stream()), "render_backward")
        for i, k in enumerate(kinds):
            if need[3 + i]:
                out[3 + i] = grads[k]
        pi, base = 0, 3 + ctx.n_grid
        for k in kinds:
            n = lib().nsb_param_count(k)
            if k in dw:
                off = 0
                for j in range(n):
                    numel = lib().nsb_param_numel(k, j)
                    if need_param[pi + j]:
                        out[base + pi + j] = dw[k][off:off + numel].view(ctx.param_shapes[pi + j])
                    off += numel
            pi += n
        return tuple(out)

    @staticmethod
    def backward(ctx, g_depth, g_var, g_color):
        if ctx.fused_ws is not None:
            return _RenderFn._backward_fused(ctx, g_depth, g_var, g_color)
        meta, kinds, planes = ctx.meta, ctx.kinds, ctx.planes
        rays_o, rays_d, z, depth = ctx.saved_tensors[:4]
        grid_list = ctx.saved_tensors[4:]
        grids = dict(zip(kinds, grid_list))
        stage, bound = meta["stage"], meta["bound"]
        dev = rays_o.device
        N, S = z.shape
        P = N * S
        need = ctx.needs_input_grad          # (meta, rays_o, rays_d, grids..., params...)
        need_rays = need[1] or need[2]
        need_grid = {k: need[3 + i] for i, k in enumerate(kinds)}
        need_param = need[3 + ctx.n_grid:]
        out = [None, None, None] + [None] * (len(need) - 3)
        if N == 0:
            return tuple(out)
        b6 = _bound6(bound)
        pts = ray_points(rays_o, rays_d, z)
        g_depth = None if g_depth is None else g_depth.double().contiguous()
        g_var = None if g_var is None else g_var.double().contiguous()
        g_color = None if g_color is None else g_color.float().contiguous()
        d_raw = torch.empty(P, 4, dtype=torch.float32, device=dev)
        density = not meta["occupancy"]
        d_dnorm = torch.empty(N, 3, dtype=torch.float32, device=dev) if (density and need_rays) else None
        check(lib().nsb_composite_backward(C.byref(pts), b6, STAGES[stage], int(meta["occupancy"]),
                                           _ptr(planes.occ0), _ptr(planes.occ1), _ptr(planes.rgb), _ptr(depth),
                                           _ptr(g_depth), _ptr(g_var), _ptr(g_color), _ptr(d_raw), _ptr(d_dnorm),
                                           _stream()), "composite_backward")
        flags = 0
        grads = {}
        # gradient reaches: middle grid via the middle decoder, fine grid via the fine decoder,
        # colour grid via the colour decoder (the middle feature inside the fine decoder is
        # sampled under no_grad, decoder.py:199-201)
        for k in kinds:
            if need_grid[k]:
                grads[k] = torch.zeros_like(grids[k], memory_format=torch.preserve_format)
                flags |= GRAD_GRID
        d_point = None
        if need_rays:
            d_point = torch.empty(len(kinds), P, 3, dtype=torch.float32, device=dev)
            flags |= GRAD_POINT
        dw_mask = 0
        dw = {}
        pi = 0
        for k in kinds:
            n = lib().nsb_param_count(k)
            if any(need_param[pi:pi + n]):
                dw[k] = torch.zeros(lib().nsb_flat_numel(k), dtype=torch.float32, device=dev)
                dw_mask |= 1 << k
            pi += n
        scratch = None
        scratch_bytes = 0
        if dw_mask:
            flags |= GRAD_WEIGHT
            scratch_bytes = max(lib().nsb_dweight_scratch_bytes(k, P) for k in dw)
            scratch = torch.empty(max(scratch_bytes, 1), dtype=torch.uint8, device=dev)
        dw_ptrs = (C.c_void_p * 4)(*[dw[k].data_ptr() if k in dw else 0 for k in range(4)])
        scene = make_scene(bound, grids, meta["packed"], grads, meta["coarse_bound_enlarge"])
        if flags:
            check(lib().nsb_decoders_backward(C.byref(pts), C.byref(scene), STAGES[stage], _ptr(d_raw), flags,
                                              dw_mask, dw_ptrs, _ptr(d_point), _ptr(scratch), scratch_bytes,
                                              _stream()), "decoders_backward")
        if need_rays:
            g_o = torch.empty(N, 3, dtype=torch.float32, device=dev)
            g_d = torch.empty(N, 3, dtype=torch.float32, device=dev)
            check(lib().nsb_ray_grad_reduce(C.byref(pts), _ptr(d_point), len(kinds), _ptr(d_dnorm), _ptr(g_o),
                                            _ptr(g_d), _stream()), "ray_grad_reduce")
            out[1] = g_o if need[1] else None
            out[2] = g_d if need[2] else None
        for i, k in enumerate(kinds):
            if need_grid[k]:
                out[3 + i] = grads[k]
        pi = 0
        base = 3 + ctx.n_grid
        for k in kinds:
            n = lib().nsb_param_count(k)
            if k in dw:
                off = 0
                for j in range(n):
                    numel = lib().nsb_param_numel(k, j)
                    if need_param[pi + j]:
                        out[base + pi + j] = dw[k][off:off + numel].view(ctx.param_shapes[pi + j])
                    off += numel
            pi += n
        return tuple(out)


def render_batch_ray(c, decoders, rays_d, rays_o, stage, gt_depth, bound, *, N_samples, N_surface,
                     N_importance=0, lindisp=False, occupancy=True, u=None, u_importance=None,
                     coarse_bound_enlarge=2.0, keep=None):
    """Fused Renderer.render_batch_ray for the NICE decoders.  Differentiable in the grids of
    ``c``, the decoder parameters and the rays.  ``keep`` ([N] uint8, optional): fixed-shape
    alternative to compacting the batch after the box pre-filter (dropped rays are excluded
    from the batch depth max; give them zero loss weight)."""
    if stage not in STAGES:
        raise ValueError(f"unknown stage {stage!r}")
    _require_cuda(rays_o, "rays_o")
    rays_o, rays_d = _row_major_rays(rays_o), _row_major_rays(rays_d)
    if stage == "coarse":
        gt_depth = None                                      # Renderer.py:121-122
    if gt_depth is not None:
        gt_depth = gt_depth.detach().reshape(-1).float().contiguous()
    grids, packed, params = _gather_scene(stage, bound, c, decoders, coarse_bound_enlarge)
    kinds = STAGE_DECODERS[stage]
    flat_params = [p for k in kinds for p in params[k]]
    meta = dict(stage=stage, bound=torch.as_tensor(bound, dtype=torch.float64, device="cpu"),
                gt_depth=gt_depth, N_samples=int(N_samples), N_surface=int(N_surface),
                N_importance=int(N_importance), lindisp=bool(lindisp), occupancy=bool(occupancy), u=u,
                u_importance=u_importance, packed=packed, coarse_bound_enlarge=float(coarse_bound_enlarge),
                keep=keep)
    if keep is not None and N_importance > 0:
        raise NotImplementedError("keep masks are supported on the single-pass (NICE) render only")
    return _RenderFn.apply(meta, rays_o, rays_d, *[grids[k] for k in kinds], *flat_params)


class _LossFn(torch.autograd.Function):
    """The SLAM step losses (src/Mapper.py:604-612, src/Tracker.py:113-130), value and gradient
    in one kernel: nsb_mapping_loss / nsb_tracking_loss."""

    @staticmethod
    def forward(ctx, depth, var, color, gt_depth, gt_color, keep, kind, w_color, use_color, handle_dynamic):
        dev, N = depth.device, depth.shape[0]
        loss = torch.zeros(1, dtype=torch.float64, device=dev)
        g_depth = torch.zeros(N, dtype=torch.float64, device=dev)
        g_color = torch.zeros(N, 3, dtype=torch.float32, device=dev)
        if N > 0:
            depth, color = depth.detach().contiguous(), color.detach().contiguous()
            if kind == "mapping":
                check(lib().nsb_mapping_loss(_ptr(depth), _ptr(color), _ptr(gt_depth), _ptr(gt_color), _ptr(keep), N,
                                             float(w_color), int(use_color), _ptr(loss), _ptr(g_depth), _ptr(g_color),
                                             _stream()), "mapping_loss")
            else:
                var = var.detach().contiguous()
                check(lib().nsb_tracking_loss(_ptr(depth), _ptr(var), _ptr(color), _ptr(gt_depth), _ptr(gt_color),
                                              _ptr(keep), N, float(w_color), int(use_color), int(handle_dynamic),
                                              _ptr(loss), _ptr(g_depth), _ptr(g_color), _stream()), "tracking_loss")
        ctx.save_for_backward(g_depth, g_color)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        g_depth, g_color = ctx.saved_tensors
        return g_depth * g, None, g_color * g.float(), None, None, None, None, None, None, None


def _loss_args(depth, color, gt_depth, gt_color, keep):
    _require_cuda(depth, "depth")
    gt_depth = gt_depth.detach().reshape(-1).float().contiguous()
    gt_color = gt_color.detach().reshape(-1, 3).float().contiguous()
    if keep is not None:
        keep = keep.to(torch.uint8).contiguous()
    if depth.dtype != torch.float64:
        depth = depth.double()
    if color.dtype != torch.float32:
        color = color.float()
    return depth, color, gt_depth, gt_color, keep


def mapping_loss(depth, color, gt_depth, gt_color, w_color, use_color=True, keep=None):
    """sum_{gt>0} |gt - depth| + w_color sum |gt_c - color| (src/Mapper.py:604-612), fp64 scalar."""
    depth, color, gt_depth, gt_color, keep = _loss_args(depth, color, gt_depth, gt_color, keep)
    return _LossFn.apply(depth, None, color, gt_depth, gt_color, keep, "mapping", w_color, use_color, False)


def tracking_loss(depth, var, color, gt_depth, gt_color, w_color, use_color, handle_dynamic, keep=None):
    """Tracker loss with the median-based dynamic-object mask (src/Tracker.py:113-130); the
    uncertainty is detached as in the reference."""
    depth, color, gt_depth, gt_color, keep = _loss_args(depth, color, gt_depth, gt_color, keep)
    return _LossFn.apply(depth, var.double(), color, gt_depth, gt_color, keep, "tracking", w_color, use_color,
                         handle_dynamic)


STAGE_SINGLE = 16


class _EvalFn(torch.autograd.Function):
    """Decoders on explicit points: NICE.forward (decoder.py:336-370), a single MLP.forward
    (decoder.py:190-219,285-297) or Renderer.eval_points (Renderer.py:36-93, apply_oob)."""

    @staticmethod
    def forward(ctx, meta, p, *tensors):
        kinds = meta["kinds"]
        n_grid = len(kinds)
        grids = dict(zip(kinds, tensors[:n_grid]))
        dev, P = p.device, p.shape[0]
        if meta.get("aux_grid") is not None:
            grids[DEC_MIDDLE] = meta["aux_grid"]
        scene = make_scene(meta["bound"], grids, meta["packed"], None, meta["coarse_bound_enlarge"],
                           meta.get("coarse_bound"))
        pts = explicit_points(p)
        code = meta["stage_code"]
        occ0 = occ1 = rgb = None
        if any(k in (DEC_COARSE, DEC_MIDDLE) for k in kinds):
            occ0 = torch.empty(P, dtype=torch.float32, device=dev)
        if DEC_FINE in kinds:
            occ1 = torch.empty(P, dtype=torch.float32, device=dev)
        if DEC_COLOR in kinds:
            rgb = torch.empty(P, 4, dtype=torch.float32, device=dev)
        inb = None
        if P > 0:
            check(lib().nsb_decoders_forward(C.byref(pts), C.byref(scene), code, _ptr(occ0), _ptr(occ1),
                                             _ptr(rgb), _stream()), "decoders_forward")
        if meta["single"]:
            out = rgb if rgb is not None else (occ1 if occ1 is not None else occ0)
        else:
            out = torch.empty(P, 4, dtype=torch.float32, device=dev)
            if meta["apply_oob"]:
                inb = torch.empty(P, dtype=torch.uint8, device=dev)
                b6 = _bound6(meta["bound"])
            else:   # no out-of-bound rule: an all-enclosing bound
                b6 = (C.c_double * 6)(*([-float("inf"), float("inf")] * 3))
            if P > 0:
                check(lib().nsb_assemble_raw(C.byref(pts), b6, code, _ptr(occ0), _ptr(occ1), _ptr(rgb),
                                             _ptr(out), _ptr(inb), _stream()), "assemble_raw")
        ctx.meta, ctx.n_grid, ctx.inb = meta, n_grid, inb
        ctx.param_shapes = [t.shape for t in tensors[n_grid:]]
        ctx.save_for_backward(p, *tensors[:n_grid])
        return out

    @staticmethod
    def backward(ctx, g):
        meta = ctx.meta
        kinds = meta["kinds"]
        p = ctx.saved_tensors[0]
        grids = dict(zip(kinds, ctx.saved_tensors[1:]))
        dev, P = p.device, p.shape[0]
        need = ctx.needs_input_grad
        out = [None] * len(need)
        if P == 0:
            return tuple(out)
        g = g.float()
        if meta["single"]:
            if g.dim() == 1:
                d_raw = torch.zeros(P, 4, dtype=torch.float32, device=dev)
                d_raw[:, 3] = g
            else:
                d_raw = g.contiguous()
        else:
            d_raw = g.contiguous()
            if ctx.inb is not None:       # Renderer.py:87: the overwrite cuts the occupancy gradient
                d_raw = d_raw.clone()
                d_raw[:, 3] *= ctx.inb.float()
        flags, grads = 0, {}
        for i, k in enumerate(kinds):
            if need[2 + i]:
                grads[k] = torch.zeros_like(grids[k], memory_format=torch.preserve_format)
                flags |= GRAD_GRID
        d_point = None
        if need[1]:
            d_point = torch.empty(len(kinds), P, 3, dtype=torch.float32, device=dev)
            flags |= GRAD_POINT
        need_param = need[2 + ctx.n_grid:]
        dw, dw_mask, pi = {}, 0, 0
        for k in kinds:
            n = lib().nsb_param_count(k)
            if any(need_param[pi:pi + n]):
                dw[k] = torch.zeros(lib().nsb_flat_numel(k), dtype=torch.float32, device=dev)
                dw_mask |= 1 << k
            pi += n
        scratch, scratch_bytes = None, 0
        if dw_mask:
            flags |= GRAD_WEIGHT
            scratch_bytes = max(lib().nsb_dweight_scratch_bytes(k, P) for k in dw)
            scratch = torch.empty(max(scratch_bytes, 1), dtype=torch.uint8, device=dev)
        dw_ptrs = (C.c_void_p * 4)(*[dw[k].data_ptr() if k in dw else 0 for k in range(4)])
        if meta.get("aux_grid") is not None:
            grids[DEC_MIDDLE] = meta["aux_grid"]
        scene = make_scene(meta["bound"], grids, meta["packed"], grads, meta["coarse_bound_enlarge"],
                           meta.get("coarse_bound"))
        pts = explicit_points(p)
        if flags:
            check(lib().nsb_decoders_backward(C.byref(pts), C.byref(scene), meta["stage_code"], _ptr(d_raw),
                                              flags, dw_mask, dw_ptrs, _ptr(d_point), _ptr(scratch),
                                              scratch_bytes, _stream()), "decoders_backward")
        if need[1]:
            out[1] = d_point.sum(0).to(p.dtype)
        for i, k in enumerate(kinds):
            if need[2 + i]:
                out[2 + i] = grads[k]
        pi, base = 0, 2 + ctx.n_grid
        for k in kinds:
            n = lib().nsb_param_count(k)
            if k in dw:
                off = 0
                for j in range(n):
                    numel = lib().nsb_param_numel(k, j)
                    if need_param[pi + j]:
                        out[base + pi + j] = dw[k][off:off + numel].view(ctx.param_shapes[pi + j])
                    off += numel
            pi += n
        return tuple(out)


def _points_arg(p):
    p = p.reshape(-1, 3)
    _require_cuda(p, "points")
    if p.dtype not in (torch.float32, torch.float64):
        p = p.float()
    return p.contiguous()


def _eval(p, kinds, stage_code, single, apply_oob, bound, grids, packed, params, enlarge, coarse_bound=None):
    meta = dict(kinds=kinds, stage_code=stage_code, single=single, apply_oob=apply_oob,
                bound=torch.as_tensor(bound, dtype=torch.float64, device="cpu"), packed=packed,
                coarse_bound_enlarge=float(enlarge), coarse_bound=coarse_bound)
    flat_params = [q for k in kinds for q in params[k]]
    return _EvalFn.apply(meta, _points_arg(p), *[grids[k] for k in kinds], *flat_params)


def nice_forward(nice, p, c_grid, stage, coarse_bound_enlarge=2.0):
    """NICE.forward: raw [N,4] without the out-of-bound rule."""
    bound = nice.scene_bound()
    grids, packed, params = _gather_scene(stage, bound, c_grid, nice, coarse_bound_enlarge)
    cb = None
    if stage == "coarse":   # the coarse decoder's own bound is the enlarged one (NICE_SLAM.py:208)
        cb = torch.as_tensor(nice.coarse_decoder.bound, dtype=torch.float64, device="cpu")
    return _eval(p, STAGE_DECODERS[stage], STAGES[stage], False, False, bound, grids, packed, params,
                 coarse_bound_enlarge, cb)


def decoder_forward_single(dec, kind, p, c_grid):
    """One decoder on its own (MLP.forward / MLP_no_xyz.forward): [N] or [N,4]."""
    bound = torch.as_tensor(dec.bound, dtype=torch.float64, device="cpu")
    grids = {kind: grid_to_kernel_layout(c_grid[GRID_KEYS[kind]])}
    kinds = (kind,)
    if kind == DEC_FINE:
        grids[DEC_MIDDLE] = grid_to_kernel_layout(c_grid["grid_middle"]).detach()
    packed_k, params_k = packed_decoder(dec, kind)
    meta = dict(kinds=kinds, stage_code=STAGE_SINGLE + kind, single=True, apply_oob=False,
                bound=bound, packed={kind: packed_k}, coarse_bound_enlarge=1.0,
                aux_grid=grids.get(DEC_MIDDLE) if kind == DEC_FINE else None)
    return _EvalFn.apply(meta, _points_arg(p), grids[kind], *params_k)


def eval_points_autograd(p, decoders, c, stage, bound, coarse_bound_enlarge=2.0):
    """Differentiable Renderer.eval_points (raw [P,4] with the out-of-bound rule)."""
    grids, packed, params = _gather_scene(stage, bound, c, decoders, coarse_bound_enlarge)
    return _eval(p, STAGE_DECODERS[stage], STAGES[stage], False, True, bound, grids, packed, params,
                 coarse_bound_enlarge)


def eval_points(p, decoders, c, stage, bound, coarse_bound_enlarge=2.0):
    """Forward-only Renderer.eval_points (src/utils/Renderer.py:36-93): raw [P,4] fp32."""
    p = p.reshape(-1, 3)
    _require_cuda(p, "points")
    if p.dtype not in (torch.float32, torch.float64):
        p = p.float()
    p = p.detach().contiguous()
    P = p.shape[0]
    grids, packed, _ = _gather_scene(stage, bound, c, decoders, coarse_bound_enlarge)
    scene = make_scene(bound, {k: g.detach() for k, g in grids.items()}, packed, None, coarse_bound_enlarge)
    pts = explicit_points(p)
    planes = decoders_forward(pts, scene, stage, P, p.device)
    raw = torch.empty(P, 4, dtype=torch.float32, device=p.device)
    if P > 0:
        check(lib().nsb_assemble_raw(C.byref(pts), _bound6(bound), STAGES[stage], _ptr(planes.occ0),
                                     _ptr(planes.occ1), _ptr(planes.rgb), _ptr(raw), None, _stream()),
              "assemble_raw")
    return raw
