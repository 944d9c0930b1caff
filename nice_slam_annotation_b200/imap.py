"""Host side of the iMAP* path (``nice=False``: one 4 x 256 MLP, no feature grids, volume density;
configs/imap.yaml, src/conv_onet/config.py:33-37): tensors in, C-ABI calls, tensors out.

``render_batch_ray`` is the reference's two-pass renderer (src/utils/Renderer.py:95-262 with
N_importance > 0): sample depths, MLP, density compositing, importance resampling from the first
pass's weights (detached), second MLP pass over all samples, compositing.  Only the second pass
carries gradient, so the first one runs in a fixed workspace and keeps nothing.  The MLP itself is
``nsb_imap_forward`` / ``nsb_imap_backward`` (chained 3xTF32 GEMMs on tcgen05, csrc/nsb_gemm.cu).
There is no CPU path.
"""
import ctypes as C

import torch

from . import fused
from ._lib import GRAD_POINT, GRAD_WEIGHT, STAGES, NsbImapDecoder, NsbImapGrads, check, lib
from .fused import _bound6, _ptr, _require_cuda, _row_major_rays, _stream

COLOR = STAGES["color"]


def check_supported(dec):
    ok = (dec.c_dim == 0 and dec.color and dec.n_blocks == 4 and list(dec.skips) == [] and
          dec.hidden_size % 16 == 0 and 16 <= dec.hidden_size <= 256 and dec.pos_embedding_method == 'fourier')
    if not ok:
        raise NotImplementedError(
            "single-MLP decoder: the sm_100a kernels cover the iMAP* configuration of the reference factory "
            "(c_dim 0, colour head, 4 blocks, no skips, fourier embedding, hidden a multiple of 16 up to 256); got "
            f"c_dim={dec.c_dim}, color={dec.color}, n_blocks={dec.n_blocks}, skips={dec.skips}, "
            f"hidden={dec.hidden_size}, embedding={dec.pos_embedding_method}")


def canonical_params(dec):
    """pts_linears.{0..3}.weight, pts_linears.{0..3}.bias, output_linear.weight, output_linear.bias, embedder._B."""
    check_supported(dec)
    return ([dec.pts_linears[i].weight for i in range(4)] + [dec.pts_linears[i].bias for i in range(4)] +
            [dec.output_linear.weight, dec.output_linear.bias, dec.embedder._B])


def _decoder_struct(params, hidden):
    """NsbImapDecoder over the live parameter storages (+ the tensors that must stay alive)."""
    keep = []
    for p in params:
        _require_cuda(p, "decoder parameters")
        if p.dtype != torch.float32:
            raise RuntimeError(f"decoder parameters must be float32, got {p.dtype}")
        keep.append(p.detach() if p.is_contiguous() else p.detach().contiguous())
    d = NsbImapDecoder()
    for i in range(4):
        d.W[i], d.b[i] = keep[i].data_ptr(), keep[4 + i].data_ptr()
    d.Wo, d.bo, d.B = keep[8].data_ptr(), keep[9].data_ptr(), keep[10].data_ptr()
    d.hidden = hidden
    return d, keep


def mlp_forward(points, P, params, hidden, device, save):
    """raw [P,4] of an NsbPoints set; returns (raw, workspace or None)."""
    raw = torch.empty(P, 4, dtype=torch.float32, device=device)
    if P == 0:
        return raw, None
    ws = torch.empty(lib().nsb_imap_workspace_floats(P, hidden, int(save)), dtype=torch.float32, device=device)
    d, keep = _decoder_struct(params, hidden)
    check(lib().nsb_imap_forward(C.byref(points), C.byref(d), _ptr(raw), _ptr(ws), int(save), _stream()),
          "imap_forward")
    return raw, (ws if save else None)


def mlp_backward(points, P, params, hidden, ws, d_raw, need_param, need_point):
    """(list of parameter gradients in canonical order or None, d_point [P,3] or None)."""
    dev = d_raw.device
    flags = (GRAD_WEIGHT if need_param else 0) | (GRAD_POINT if need_point else 0)
    if P == 0 or not flags:
        return None, None
    d, keep = _decoder_struct(params, hidden)
    grads, g = None, NsbImapGrads()
    if need_param:
        grads = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in params]
        for i in range(4):
            g.W[i], g.b[i] = grads[i].data_ptr(), grads[4 + i].data_ptr()
        g.Wo, g.bo, g.B = grads[8].data_ptr(), grads[9].data_ptr(), grads[10].data_ptr()
    d_point = torch.empty(P, 3, dtype=torch.float32, device=dev) if need_point else None
    scratch = torch.empty(lib().nsb_imap_backward_scratch_floats(P, hidden), dtype=torch.float32, device=dev)
    check(lib().nsb_imap_backward(C.byref(points), C.byref(d), _ptr(ws), _ptr(d_raw), flags, C.byref(g),
                                  _ptr(d_point), _ptr(scratch), _stream()), "imap_backward")
    return grads, d_point


class _RenderFn(torch.autograd.Function):
    """Renderer.render_batch_ray for the single-MLP decoder as one autograd node."""

    @staticmethod
    def forward(ctx, meta, rays_o, rays_d, *params):
        dev, N = rays_o.device, rays_o.shape[0]
        bound, hidden, occ = meta["bound"], meta["hidden"], int(meta["occupancy"])
        z = fused.sample_z(rays_o, rays_d, meta["gt_depth"], bound, meta["N_samples"], meta["N_surface"],
                           meta["lindisp"], meta.get("u"))
        S, Ni = z.shape[1], meta["N_importance"]
        b6 = _bound6(bound)
        depth = torch.empty(N, dtype=torch.float64, device=dev)
        var = torch.empty(N, dtype=torch.float64, device=dev)
        color = torch.empty(N, 3, dtype=torch.float32, device=dev)
        pts = fused.ray_points(rays_o, rays_d, z)
        raw, ws = mlp_forward(pts, N * S, params, hidden, dev, save=(Ni == 0 and meta["save"]))
        weights = torch.empty(N, S, dtype=torch.float32, device=dev) if Ni > 0 else None
        if N > 0:
            check(lib().nsb_composite_forward(C.byref(pts), b6, COLOR, occ, None, None, _ptr(raw), _ptr(depth),
                                              _ptr(var), _ptr(color), _ptr(weights), _stream()), "composite_forward")
        if Ni > 0 and N > 0:
            u = meta.get("u_importance")
            if u is None:
                u, stride = fused._linspace(Ni, torch.float32, dev), 0
            else:
                u, stride = u.to(device=dev, dtype=torch.float32).contiguous(), Ni
            z2 = torch.empty(N, S + Ni, dtype=torch.float64, device=dev)
            check(lib().nsb_importance_z(_ptr(z), _ptr(weights), N, S, Ni, _ptr(u), stride, None, _ptr(z2),
                                         _stream()), "importance_z")
            z, S = z2, S + Ni
            pts = fused.ray_points(rays_o, rays_d, z)
            raw, ws = mlp_forward(pts, N * S, params, hidden, dev, save=meta["save"])
            check(lib().nsb_composite_forward(C.byref(pts), b6, COLOR, occ, None, None, _ptr(raw), _ptr(depth),
                                              _ptr(var), _ptr(color), None, _stream()), "composite_forward")
        ctx.meta, ctx.ws = meta, ws
        ctx.save_for_backward(rays_o, rays_d, z, depth, raw, *params)
        return depth, var, color

    @staticmethod
    def backward(ctx, g_depth, g_var, g_color):
        meta = ctx.meta
        rays_o, rays_d, z, depth, raw = ctx.saved_tensors[:5]
        params = ctx.saved_tensors[5:]
        dev = rays_o.device
        N, S = z.shape
        P = N * S
        need = ctx.needs_input_grad                 # (meta, rays_o, rays_d, params...)
        out = [None] * len(need)
        if N == 0:
            return tuple(out)
        need_rays, need_param = need[1] or need[2], any(need[3:])
        pts = fused.ray_points(rays_o, rays_d, z)
        g_depth = None if g_depth is None else g_depth.double().contiguous()
        g_var = None if g_var is None else g_var.double().contiguous()
        g_color = None if g_color is None else g_color.float().contiguous()
        d_raw = torch.empty(P, 4, dtype=torch.float32, device=dev)
        density = not meta["occupancy"]
        d_dnorm = torch.empty(N, 3, dtype=torch.float32, device=dev) if (density and need_rays) else None
        check(lib().nsb_composite_backward(C.byref(pts), _bound6(meta["bound"]), COLOR, int(meta["occupancy"]), None,
                                           None, _ptr(raw), _ptr(depth), _ptr(g_depth), _ptr(g_var), _ptr(g_color),
                                           _ptr(d_raw), _ptr(d_dnorm), _stream()), "composite_backward")
        grads, d_point = mlp_backward(pts, P, params, meta["hidden"], ctx.ws, d_raw, need_param, need_rays)
        if need_rays:
            g_o = torch.empty(N, 3, dtype=torch.float32, device=dev)
            g_d = torch.empty(N, 3, dtype=torch.float32, device=dev)
            check(lib().nsb_ray_grad_reduce(C.byref(pts), _ptr(d_point), 1, _ptr(d_dnorm), _ptr(g_o), _ptr(g_d),
                                            _stream()), "ray_grad_reduce")
            out[1] = g_o if need[1] else None
            out[2] = g_d if need[2] else None
        if grads is not None:
            for i, g in enumerate(grads):
                if need[3 + i]:
                    out[3 + i] = g
        return tuple(out)


def render_batch_ray(decoder, rays_d, rays_o, gt_depth, bound, *, N_samples, N_surface, N_importance, lindisp=False,
                     occupancy=False, u=None, u_importance=None):
    """Fused Renderer.render_batch_ray for the single-MLP decoder; differentiable in the decoder
    parameters and the rays."""
    _require_cuda(rays_o, "rays_o")
    rays_o, rays_d = _row_major_rays(rays_o), _row_major_rays(rays_d)
    if gt_depth is not None:
        gt_depth = gt_depth.detach().reshape(-1).float().contiguous()
    params = canonical_params(decoder)
    meta = dict(bound=torch.as_tensor(bound, dtype=torch.float64, device="cpu"), gt_depth=gt_depth,
                N_samples=int(N_samples), N_surface=int(N_surface), N_importance=int(N_importance),
                lindisp=bool(lindisp), occupancy=bool(occupancy), u=u, u_importance=u_importance,
                hidden=int(decoder.hidden_size),
                save=torch.is_grad_enabled() and (rays_o.requires_grad or rays_d.requires_grad or
                                                  any(q.requires_grad for q in params)))
    return _RenderFn.apply(meta, rays_o, rays_d, *params)


class _EvalFn(torch.autograd.Function):
    """MLP.forward on explicit points (decoder.py:190-219), optionally with the out-of-bound rule of
    Renderer.eval_points (Renderer.py:63-66,87: raw[~mask, 3] = 100)."""

    @staticmethod
    def forward(ctx, meta, p, *params):
        dev, P = p.device, p.shape[0]
        pts = fused.explicit_points(p)
        raw, ws = mlp_forward(pts, P, params, meta["hidden"], dev, save=meta["save"])
        inb = None
        if meta["bound"] is not None and P > 0:
            out = torch.empty(P, 4, dtype=torch.float32, device=dev)
            inb = torch.empty(P, dtype=torch.uint8, device=dev)
            check(lib().nsb_assemble_raw(C.byref(pts), _bound6(meta["bound"]), COLOR, None, None, _ptr(raw), _ptr(out),
                                         _ptr(inb), _stream()), "assemble_raw")
            raw = out
        ctx.meta, ctx.ws, ctx.inb = meta, ws, inb
        ctx.save_for_backward(p, *params)
        return raw

    @staticmethod
    def backward(ctx, g):
        p, params = ctx.saved_tensors[0], ctx.saved_tensors[1:]
        need = ctx.needs_input_grad
        out = [None] * len(need)
        P = p.shape[0]
        if P == 0:
            return tuple(out)
        d_raw = g.float().contiguous()
        if ctx.inb is not None:                      # the overwrite cuts the density gradient
            d_raw = d_raw.clone()
            d_raw[:, 3] *= ctx.inb.float()
        grads, d_point = mlp_backward(fused.explicit_points(p), P, params, ctx.meta["hidden"], ctx.ws, d_raw,
                                      any(need[2:]), need[1])
        if need[1]:
            out[1] = d_point.to(p.dtype)
        if grads is not None:
            for i, gr in enumerate(grads):
                if need[2 + i]:
                    out[2 + i] = gr
        return tuple(out)


def eval_points(p, decoder, bound=None):
    """raw [P,4] of explicit points; ``bound`` given: Renderer.eval_points (out-of-bound density = 100),
    ``bound`` None: the bare MLP.forward.  Differentiable when an input requires grad."""
    p = fused._points_arg(p)
    params = canonical_params(decoder)
    meta = dict(hidden=int(decoder.hidden_size),
                bound=None if bound is None else torch.as_tensor(bound, dtype=torch.float64, device="cpu"),
                save=torch.is_grad_enabled() and (p.requires_grad or any(q.requires_grad for q in params)))
    return _EvalFn.apply(meta, p, *params)
