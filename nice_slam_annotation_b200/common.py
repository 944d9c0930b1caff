"""Ray / compositing math of the hot path with the reference's function names and argument
meaning (src/common.py), evaluated by the sm_100a kernels.

``device`` arguments are kept for signature compatibility; tensors must live on a CUDA device
(there is no CPU path: the product fails loudly without the extension).  The two tiny camera
conversions (``quad2rotation`` / ``get_camera_from_tensor``, 7 numbers) stay plain torch so that
callers which differentiate through them with autograd (Mapper BA) keep working; the tracking
step uses the fused pose-gradient kernel instead (Tracker.py).
"""
import ctypes as C

import numpy as np
import torch

from . import fused
from ._lib import check, lib
from .fused import _ptr, _stream


# ---------------------------------------------------------------------------------------
# camera parameterisation (src/common.py:137-176)
# ---------------------------------------------------------------------------------------
def quad2rotation(quad):
    """Quaternion (batch, 4) -> rotation (batch, 3, 3), normalising by 2/|q|^2 (common.py:137-160)."""
    qr, qi, qj, qk = quad[:, 0], quad[:, 1], quad[:, 2], quad[:, 3]
    two_s = 2.0 / (quad * quad).sum(-1)
    rows = [
        1 - two_s * (qj ** 2 + qk ** 2), two_s * (qi * qj - qk * qr), two_s * (qi * qk + qj * qr),
        two_s * (qi * qj + qk * qr), 1 - two_s * (qi ** 2 + qk ** 2), two_s * (qj * qk - qi * qr),
        two_s * (qi * qk - qj * qr), two_s * (qj * qk + qi * qr), 1 - two_s * (qi ** 2 + qj ** 2),
    ]
    return torch.stack(rows, -1).reshape(-1, 3, 3)


def get_camera_from_tensor(inputs):
    """[quat(4), t(3)] -> [R|t] (3x4) (common.py:163-176)."""
    N = len(inputs.shape)
    if N == 1:
        inputs = inputs.unsqueeze(0)
    quad, T = inputs[:, :4], inputs[:, 4:]
    R = quad2rotation(quad)
    RT = torch.cat([R, T[:, :, None]], 2)
    if N == 1:
        RT = RT[0]
    return RT


def get_tensor_from_camera(RT, Tquad=False):
    """[R|t] -> [quat, t] (common.py:179-212).  The reference uses mathutils here; this is the
    same Shepperd conversion on the host (not on the per-iteration path)."""
    gpu_id = -1
    if isinstance(RT, torch.Tensor):
        if RT.get_device() != -1:
            gpu_id = RT.get_device()
        RT = RT.detach().cpu().numpy()
    R, T = np.asarray(RT[:3, :3], dtype=np.float64), np.asarray(RT[:3, 3], dtype=np.float64)
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    if tr > 0:
        s = np.sqrt(tr + 1.0) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = np.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = np.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = np.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    quad = np.asarray(q)
    tensor = np.concatenate([T, quad], 0) if Tquad else np.concatenate([quad, T], 0)
    tensor = torch.from_numpy(tensor).float()
    if gpu_id != -1:
        tensor = tensor.to(gpu_id)
    return tensor


def as_intrinsics_matrix(intrinsics):
    """common.py:7-16."""
    K = np.eye(3)
    K[0, 0], K[1, 1], K[0, 2], K[1, 2] = intrinsics[0], intrinsics[1], intrinsics[2], intrinsics[3]
    return K


def random_select(l, k):
    """common.py:66-71."""
    import random
    return list(np.random.permutation(np.array(range(l)))[:min(l, k)])


# ---------------------------------------------------------------------------------------
# rays
# ---------------------------------------------------------------------------------------
class _SampleRaysFn(torch.autograd.Function):
    """get_samples (common.py:92-134) as one kernel; backward = the fused pose gradient
    restricted to c2w (d(rays_o), d(rays_d) -> d(c2w[:3,:4]))."""

    @staticmethod
    def forward(ctx, c2w, indices, crop, cam, depth, color):
        H0, H1, W0, W1 = crop
        fx, fy, cx, cy = cam
        dev = depth.device
        n = indices.shape[0]
        c2w34 = c2w[:3, :4].detach().float().contiguous()
        ro = torch.empty(n, 3, dtype=torch.float32, device=dev)
        rd = torch.empty(n, 3, dtype=torch.float32, device=dev)
        gd = torch.empty(n, dtype=torch.float32, device=dev)
        gc = torch.empty(n, 3, dtype=torch.float32, device=dev)
        check(lib().nsb_sample_rays(_ptr(indices), n, H0, H1, W0, W1, depth.shape[1], fx, fy, cx, cy,
                                    _ptr(c2w34), None, _ptr(depth), _ptr(color), int(color.dtype == torch.float64),
                                    None, _ptr(ro), _ptr(rd), _ptr(gd), _ptr(gc), None, _stream()),
              "sample_rays")
        ctx.save_for_backward(indices)
        ctx.geom = (crop, cam, c2w.shape)
        ctx.mark_non_differentiable(gd, gc)
        return ro, rd, gd, gc

    @staticmethod
    def backward(ctx, g_o, g_d, _gd, _gc):
        (indices,) = ctx.saved_tensors
        (H0, H1, W0, W1), (fx, fy, cx, cy), shape = ctx.geom
        wc = W1 - W0
        i = (W0 + indices % wc).float()
        j = (H0 + indices // wc).float()
        dirs = torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)
        g = torch.zeros(shape, dtype=torch.float32, device=g_d.device)
        if g_d is not None:
            g[:3, :3] = g_d.float().t() @ dirs
        if g_o is not None:
            g[:3, 3] = g_o.float().sum(0)
        return g, None, None, None, None, None


def _as_image(t, dtype_ok):
    if t.dtype not in dtype_ok:
        t = t.float()
    return t.contiguous()


def get_samples(H0, H1, W0, W1, n, H, W, fx, fy, cx, cy, c2w, depth, color, device, indices=None):
    """n random pixels of the crop [H0:H1, W0:W1]: rays + sensor depth/colour (common.py:125-134).
    The pixel draw is torch.randint on the device, exactly as common.py:99 makes it (pass
    ``indices`` to share a draw).  Differentiable in c2w."""
    fused._require_cuda(depth, "depth image")
    if indices is None:
        indices = torch.randint((H1 - H0) * (W1 - W0), (n,), device=depth.device)
    indices = indices.to(depth.device).clamp(0, (H1 - H0) * (W1 - W0)).contiguous()
    if isinstance(c2w, np.ndarray):
        c2w = torch.from_numpy(c2w)
    c2w = c2w.to(depth.device)
    depth = _as_image(depth, (torch.float32,))
    color = _as_image(color, (torch.float32, torch.float64))
    return _SampleRaysFn.apply(c2w, indices, (int(H0), int(H1), int(W0), int(W1)),
                               (float(fx), float(fy), float(cx), float(cy)), depth, color)


def get_rays(H, W, fx, fy, cx, cy, c2w, device):
    """Rays of every pixel (common.py:259-277): rays_o, rays_d [H,W,3]."""
    if isinstance(c2w, np.ndarray):
        c2w = torch.from_numpy(c2w)
    c2w34 = c2w[:3, :4].detach().float().to(device).contiguous()
    fused._require_cuda(c2w34, "c2w (after .to(device))")
    ro = torch.empty(H, W, 3, dtype=torch.float32, device=c2w34.device)
    rd = torch.empty(H, W, 3, dtype=torch.float32, device=c2w34.device)
    check(lib().nsb_all_rays(H, W, fx, fy, cx, cy, _ptr(c2w34), _ptr(ro), _ptr(rd), _stream()), "all_rays")
    return ro, rd


def normalize_3d_coordinate(p, bound):
    """common.py:280-295 (in place on the reshaped view, like the reference)."""
    p = p.reshape(-1, 3)
    p[:, 0] = ((p[:, 0] - bound[0, 0]) / (bound[0, 1] - bound[0, 0])) * 2 - 1.0
    p[:, 1] = ((p[:, 1] - bound[1, 0]) / (bound[1, 1] - bound[1, 0])) * 2 - 1.0
    p[:, 2] = ((p[:, 2] - bound[2, 0]) / (bound[2, 1] - bound[2, 0])) * 2 - 1.0
    return p


# ---------------------------------------------------------------------------------------
# compositing / resampling
# ---------------------------------------------------------------------------------------
_INF6 = (C.c_double * 6)(*([-float("inf"), float("inf")] * 3))


class _CompositeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw, z_vals, rays_d, occupancy):
        N, S = z_vals.shape
        dev = raw.device
        raw4 = raw.reshape(N * S, 4).float().contiguous()
        z = z_vals.double().contiguous()
        rd = fused._row_major_rays(rays_d.detach())
        pts = fused.ray_points(rd, rd, z)          # origins are irrelevant without a bound
        depth = torch.empty(N, dtype=torch.float64, device=dev)
        var = torch.empty(N, dtype=torch.float64, device=dev)
        rgb = torch.empty(N, 3, dtype=torch.float32, device=dev)
        w = torch.empty(N, S, dtype=torch.float32, device=dev)
        if N > 0:
            check(lib().nsb_composite_forward(C.byref(pts), _INF6, 3, int(occupancy), None, None, _ptr(raw4),
                                              _ptr(depth), _ptr(var), _ptr(rgb), _ptr(w), _stream()),
                  "composite_forward")
        ctx.save_for_backward(raw4, z, rd, depth)
        ctx.occupancy = occupancy
        ctx.shape = raw.shape
        ctx.mark_non_differentiable(w)
        return depth, var, rgb, w

    @staticmethod
    def backward(ctx, g_depth, g_var, g_rgb, _gw):
        raw4, z, rd, depth = ctx.saved_tensors
        N, S = z.shape
        dev = raw4.device
        d_raw = torch.zeros(N * S, 4, dtype=torch.float32, device=dev)
        d_norm = torch.zeros(N, 3, dtype=torch.float32, device=dev) if not ctx.occupancy else None
        pts = fused.ray_points(rd, rd, z)
        if N > 0:
            check(lib().nsb_composite_backward(
                C.byref(pts), _INF6, 3, int(ctx.occupancy), None, None, _ptr(raw4), _ptr(depth),
                _ptr(None if g_depth is None else g_depth.double().contiguous()),
                _ptr(None if g_var is None else g_var.double().contiguous()),
                _ptr(None if g_rgb is None else g_rgb.float().contiguous()),
                _ptr(d_raw), _ptr(d_norm), _stream()), "composite_backward")
        return d_raw.view(ctx.shape), None, d_norm, None


def raw2outputs_nerf_color(raw, z_vals, rays_d, occupancy=False, device='cuda:0'):
    """Alpha compositing (common.py:215-256): depth [N] f64, depth variance [N] f64, rgb [N,3],
    weights [N,S].  Differentiable in raw (and in rays_d for the density mode)."""
    fused._require_cuda(raw, "raw")
    return _CompositeFn.apply(raw, z_vals, rays_d, bool(occupancy))


def sample_pdf(bins, weights, N_samples, det=False, device='cuda:0'):
    """Inverse-CDF resampling (common.py:19-63).  Random numbers are drawn by torch.rand on the
    CPU like the reference does (common.py:36-38)."""
    fused._require_cuda(bins, "bins")
    N, M = bins.shape
    bins = bins.double().contiguous()
    weights = weights.float().contiguous()
    if det:
        u = torch.linspace(0., 1., steps=N_samples).to(bins.device)
        stride = 0
    else:
        u = torch.rand([N, N_samples]).to(bins.device).contiguous()
        stride = N_samples
    out = torch.empty(N, N_samples, dtype=torch.float64, device=bins.device)
    check(lib().nsb_sample_pdf(_ptr(bins), _ptr(weights), N, M, N_samples, _ptr(u), stride, _ptr(out),
                               _stream()), "sample_pdf")
    return out
