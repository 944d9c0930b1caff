"""Plain torch-on-CPU restatement of the reference render hot path (see package docstring).

Every function cites the reference file:line it restates (paths relative to
/root/reference).  The arithmetic types follow the reference exactly: rays and depth in
fp32, scene bound / z values / sample points in fp64, decoder math in fp32, rendered depth
and variance in fp64.  Gradients come from torch autograd over these formulas.

Parameters of a decoder are passed as a plain ``dict`` keyed like the reference
``state_dict`` (``pts_linears.0.weight`` ... ``embedder._B``), so the oracle does not
depend on any nn.Module of the product package.
"""
import math

import torch

__all__ = [
    "quad2rotation", "get_camera_from_tensor", "rays_from_uv", "get_samples",
    "get_rays", "bbox_exit", "prefilter_mask", "sample_z", "normalize_points",
    "trilinear_border", "fourier_embed", "mlp_forward", "mlp_no_xyz_forward",
    "nice_forward", "eval_points", "composite", "sample_pdf", "render_batch_ray",
    "regulation", "tracking_loss", "mapping_loss", "init_decoder_params",
    "init_grids", "round_bound", "STAGES", "stage_of", "tracker_iteration_loss",
    "mapper_iteration_loss", "sample_grid_feature",
]

STAGES = ("coarse", "middle", "fine", "color")


# --------------------------------------------------------------------------------------
# camera
# --------------------------------------------------------------------------------------
def quad2rotation(quad):
    """Unit-normalising quaternion -> rotation, batched.  src/common.py:137-160."""
    qr, qi, qj, qk = quad[:, 0], quad[:, 1], quad[:, 2], quad[:, 3]
    two_s = 2.0 / (quad * quad).sum(-1)
    rows = [
        1 - two_s * (qj ** 2 + qk ** 2), two_s * (qi * qj - qk * qr), two_s * (qi * qk + qj * qr),
        two_s * (qi * qj + qk * qr), 1 - two_s * (qi ** 2 + qk ** 2), two_s * (qj * qk - qi * qr),
        two_s * (qi * qk - qj * qr), two_s * (qj * qk + qi * qr), 1 - two_s * (qi ** 2 + qj ** 2),
    ]
    return torch.stack(rows, -1).reshape(-1, 3, 3)


def get_camera_from_tensor(inputs):
    """[quat(4), t(3)] -> [R|t] (3x4).  src/common.py:163-176."""
    single = inputs.dim() == 1
    if single:
        inputs = inputs[None]
    R = quad2rotation(inputs[:, :4])
    RT = torch.cat([R, inputs[:, 4:, None]], 2)
    return RT[0] if single else RT


def rays_from_uv(i, j, c2w, fx, fy, cx, cy):
    """Pixel (i=column, j=row) -> un-normalised world ray.  src/common.py:74-89."""
    dirs = torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)
    rays_d = torch.sum(dirs[:, None, :] * c2w[:3, :3], -1)
    rays_o = c2w[:3, -1].expand(rays_d.shape)
    return rays_o, rays_d


def get_samples(H0, H1, W0, W1, n, fx, fy, cx, cy, c2w, depth, color, indices=None):
    """Pick n pixels of the crop [H0:H1, W0:W1]; rays + gt lookup.  src/common.py:92-134.

    ``indices`` (flat index into the crop, row-major) replaces the reference's
    ``torch.randint`` draw (src/common.py:99) so a test can share one draw between
    implementations; when None the draw is made exactly as the reference makes it.
    """
    hc, wc = H1 - H0, W1 - W0
    if indices is None:
        indices = torch.randint(hc * wc, (n,), device=depth.device)
    indices = indices.clamp(0, hc * wc)
    # the reference builds linspace(W0, W1-1, W1-W0) grids (common.py:117-118): integers in fp32
    i = (W0 + indices % wc).to(torch.float32)
    j = (H0 + indices // wc).to(torch.float32)
    d = depth[H0:H1, W0:W1].reshape(-1)[indices]
    c = color[H0:H1, W0:W1].reshape(-1, 3)[indices]
    rays_o, rays_d = rays_from_uv(i, j, c2w, fx, fy, cx, cy)
    return rays_o, rays_d, d, c


def get_rays(H, W, fx, fy, cx, cy, c2w):
    """All-pixel rays.  src/common.py:259-277."""
    jj, ii = torch.meshgrid(torch.linspace(0, H - 1, H), torch.linspace(0, W - 1, W), indexing="ij")
    dirs = torch.stack([(ii - cx) / fx, -(jj - cy) / fy, -torch.ones_like(ii)], -1)
    rays_d = torch.sum(dirs[:, :, None, :] * c2w[:3, :3], -1)
    rays_o = c2w[:3, -1].expand(rays_d.shape)
    return rays_o, rays_d


# --------------------------------------------------------------------------------------
# ray / box, depth sampling
# --------------------------------------------------------------------------------------
def bbox_exit(rays_o, rays_d, bound):
    """Slab exit parameter min_axis max(lo,hi)((bound-o)/d), fp64, no +0.01.

    src/utils/Renderer.py:139-150, src/Tracker.py:102-106, src/Mapper.py:583-590.
    """
    o = rays_o.detach()[:, :, None]
    d = rays_d.detach()[:, :, None]
    t = (bound[None].to(torch.float64) - o) / d       # fp64 - fp32 -> fp64
    return torch.min(torch.max(t, dim=2)[0], dim=1)[0]


def prefilter_mask(rays_o, rays_d, gt_depth, bound):
    """Rays whose box exit lies at or beyond the sensor depth.  src/Tracker.py:107."""
    return bbox_exit(rays_o, rays_d, bound) >= gt_depth


def sample_z(rays_o, rays_d, gt_depth, bound, N_samples, N_surface, lindisp=False, u=None):
    """Sorted fp64 sample depths [N, N_samples(+N_surface)].  src/utils/Renderer.py:116-224.

    u: optional [N, N_samples] fp32 uniforms for the stratified jitter (perturb > 0,
    Renderer.py:212-219; the reference draws them with torch.rand on the CPU).
    """
    N = rays_o.shape[0]
    far_bb = bbox_exit(rays_o, rays_d, bound)[:, None] + 0.01
    if gt_depth is None:
        N_surface = 0
        near = 0.01
        far = far_bb
    else:
        gt_depth = gt_depth.reshape(-1, 1)
        near = gt_depth.repeat(1, N_samples) * 0.01                 # fp32
        far = torch.clamp(far_bb, 0, torch.max(gt_depth * 1.2))     # fp64
    if N_surface > 0:
        ts = torch.linspace(0., 1., steps=N_surface, device=rays_o.device).double()
        nz = (gt_depth > 0).squeeze(-1)
        z_surf = torch.zeros(N, N_surface, dtype=torch.float64, device=rays_o.device)
        g = gt_depth[nz].reshape(-1, 1).repeat(1, N_surface)        # fp32
        z_surf[nz] = 0.95 * g * (1. - ts) + 1.05 * g * ts
        z_surf[~nz] = (0.001 * (1. - ts) + torch.max(gt_depth) * ts)[None]
    t = torch.linspace(0., 1., steps=N_samples, device=rays_o.device)
    if not lindisp:
        z = near * (1. - t) + far * t
    else:
        z = 1. / (1. / near * (1. - t) + 1. / far * t)
    if u is not None:
        mids = .5 * (z[..., 1:] + z[..., :-1])
        upper = torch.cat([mids, z[..., -1:]], -1)
        lower = torch.cat([z[..., :1], mids], -1)
        z = lower + (upper - lower) * u
    if N_surface > 0:
        z, _ = torch.sort(torch.cat([z, z_surf], -1), -1)
    return z.double() if z.dtype != torch.float64 else z


# --------------------------------------------------------------------------------------
# feature grids
# --------------------------------------------------------------------------------------
def normalize_points(p, bound):
    """points -> [-1,1] per axis, in the dtype of p.  src/common.py:280-295.  Each bound entry
    is used as a 0-dim tensor, as the reference does, so fp64 points stay fp64 and fp32
    points (Mesher / regulation callers) are normalised in fp32."""
    p = p.reshape(-1, 3)
    cols = [((p[:, a] - bound[a, 0]) / (bound[a, 1] - bound[a, 0])) * 2 - 1.0 for a in range(3)]
    return torch.stack(cols, -1)


class _ClipBorder(torch.autograd.Function):
    """clamp to [0, size-1] whose gradient is zero AT the limits as well as beyond them,
    which is what ATen's clip_coordinates_set_grad does (GridSampler.h, padding 'border')."""

    @staticmethod
    def forward(ctx, x, hi):
        ctx.save_for_backward((x > 0) & (x < hi))
        return torch.clamp(x, 0, hi)

    @staticmethod
    def backward(ctx, g):
        return g * ctx.saved_tensors[0].to(g.dtype), None


def trilinear_border(grid, p_nor32):
    """Explicit restatement of F.grid_sample(grid[1,C,Z,Y,X], mode='bilinear',
    padding_mode='border', align_corners=True) at P points (call sites
    src/conv_onet/models/decoder.py:186,281).  p_nor32: [P,3] fp32 in (x,y,z) order;
    x indexes the LAST grid dim.  Returns [P, C].  Differentiable in grid and coordinates.
    """
    _, C, Z, Y, X = grid.shape
    g = grid[0].permute(1, 2, 3, 0).reshape(Z * Y * X, C)      # row per voxel
    f = []
    for a, S in ((0, X), (1, Y), (2, Z)):
        fa = ((p_nor32[:, a] + 1.0) / 2) * (S - 1)
        f.append(_ClipBorder.apply(fa, float(S - 1)))
    fx, fy, fz = f
    x0, y0, z0 = fx.detach().floor(), fy.detach().floor(), fz.detach().floor()
    out = 0
    # ATen order: tnw, tne, tsw, tse, bnw, bne, bsw, bse (x fastest, then y, then z)
    for dz in (0, 1):
        for dy in (0, 1):
            for dx in (0, 1):
                wx = (fx - x0) if dx else (x0 + 1 - fx)
                wy = (fy - y0) if dy else (y0 + 1 - fy)
                wz = (fz - z0) if dz else (z0 + 1 - fz)
                xi = (x0 + dx).long().clamp(max=X - 1)           # weight is 0 when clamped
                yi = (y0 + dy).long().clamp(max=Y - 1)
                zi = (z0 + dz).long().clamp(max=Z - 1)
                out = out + g[(zi * Y + yi) * X + xi] * (wx * wy * wz)[:, None]
    return out


def sample_grid_feature(p64, grid, bound):
    """src/conv_onet/models/decoder.py:179-188 (normalise in fp64, sample in fp32)."""
    return trilinear_border(grid, normalize_points(p64.clone(), bound).float())


# --------------------------------------------------------------------------------------
# decoders
# --------------------------------------------------------------------------------------
def fourier_embed(p32, B):
    """sin(p @ B).  src/conv_onet/models/decoder.py:26-30."""
    return torch.sin(p32 @ B)


def _linear(x, prm, name):
    return torch.addmm(prm[name + ".bias"], x, prm[name + ".weight"].t())


def mlp_forward(prm, p64, c, n_blocks=5, skips=(2,)):
    """MLP.forward with fourier embedding.  src/conv_onet/models/decoder.py:190-219.

    c: [P, c_dim] sampled feature (None for iMAP, c_dim == 0)."""
    e = fourier_embed(p64.float().reshape(-1, 3), prm["embedder._B"])
    h = e
    for i in range(n_blocks):
        h = torch.relu(_linear(h, prm, f"pts_linears.{i}"))
        if c is not None:
            h = h + _linear(c, prm, f"fc_c.{i}")
        if i in skips:
            h = torch.cat([e, h], -1)
    out = _linear(h, prm, "output_linear")
    return out if out.shape[-1] == 4 else out.squeeze(-1)


def mlp_no_xyz_forward(prm, c, n_blocks=5, skips=(2,)):
    """MLP_no_xyz.forward.  src/conv_onet/models/decoder.py:285-297."""
    h = c
    for i in range(n_blocks):
        h = torch.relu(_linear(h, prm, f"pts_linears.{i}"))
        if i in skips:
            h = torch.cat([c, h], -1)
    return _linear(h, prm, "output_linear").squeeze(-1)


def _sub(prm, prefix):
    n = len(prefix)
    return {k[n:]: v for k, v in prm.items() if k.startswith(prefix)}


def nice_forward(prm, p64, grids, stage, bound, coarse_bound_enlarge=2):
    """NICE.forward stage mixing.  src/conv_onet/models/decoder.py:336-370.

    prm: flat dict with 'middle_decoder.' / 'fine_decoder.' / 'color_decoder.' /
    'coarse_decoder.' prefixes.  Returns raw [P,4]."""
    P = p64.reshape(-1, 3).shape[0]
    if stage == "coarse":
        c = sample_grid_feature(p64, grids["grid_coarse"], bound * coarse_bound_enlarge)
        occ = mlp_no_xyz_forward(_sub(prm, "coarse_decoder."), c)
        return torch.cat([torch.zeros(P, 3, device=occ.device), occ[:, None]], -1)

    def middle():
        c = sample_grid_feature(p64, grids["grid_middle"], bound)
        return mlp_forward(_sub(prm, "middle_decoder."), p64, c)

    def fine():
        c = sample_grid_feature(p64, grids["grid_fine"], bound)
        with torch.no_grad():                                   # decoder.py:199-201
            cm = sample_grid_feature(p64, grids["grid_middle"], bound)
        return mlp_forward(_sub(prm, "fine_decoder."), p64, torch.cat([c, cm], 1))

    if stage == "middle":
        return torch.cat([torch.zeros(P, 3, device=p64.device), middle()[:, None]], -1)
    if stage == "fine":
        return torch.cat([torch.zeros(P, 3, device=p64.device), (fine() + middle())[:, None]], -1)
    if stage == "color":
        f = fine()
        c = sample_grid_feature(p64, grids["grid_color"], bound)
        raw = mlp_forward(_sub(prm, "color_decoder."), p64, c)
        return torch.cat([raw[:, :3], (f + middle())[:, None]], -1)
    raise ValueError(stage)


def eval_points(p64, prm, grids, stage, bound, nice=True):
    """Renderer.eval_points.  src/utils/Renderer.py:36-93 (chunking omitted: it does not
    change values)."""
    p64 = p64.reshape(-1, 3)
    mask = torch.ones(p64.shape[0], dtype=torch.bool, device=p64.device)
    for a in range(3):          # 0-dim bound entries: the comparison runs in the dtype of p
        mask = mask & (p64[:, a] < bound[a][1]) & (p64[:, a] > bound[a][0])
    if nice:
        raw = nice_forward(prm, p64, grids, stage, bound)
    else:
        raw = mlp_forward(prm, p64, None, n_blocks=4, skips=())
    occ = torch.where(mask, raw[:, 3], torch.full_like(raw[:, 3], 100.0))   # Renderer.py:87
    return torch.cat([raw[:, :3], occ[:, None]], -1)


# --------------------------------------------------------------------------------------
# volume rendering
# --------------------------------------------------------------------------------------
def composite(raw, z_vals, rays_d, occupancy=True):
    """raw2outputs_nerf_color.  src/common.py:215-256.

    raw [N,S,4] fp32, z_vals [N,S] fp64 -> depth[N] f64, var[N] f64, rgb[N,3] f32, w[N,S] f32."""
    rgb = raw[..., :3]
    if occupancy:
        alpha = torch.sigmoid(10 * raw[..., 3])
    else:
        dists = (z_vals[..., 1:] - z_vals[..., :-1]).float()
        dists = torch.cat([dists, torch.full_like(dists[..., :1], 1e10)], -1)
        dists = dists * torch.norm(rays_d[..., None, :], dim=-1)
        alpha = 1. - torch.exp(-torch.relu(raw[..., 3]) * dists)
    ones = torch.ones((alpha.shape[0], 1), dtype=torch.float32, device=alpha.device)
    trans = torch.cumprod(torch.cat([ones, (1. - alpha + 1e-10).float()], -1), -1)[:, :-1]
    weights = alpha.float() * trans
    rgb_map = torch.sum(weights[..., None] * rgb, -2)
    depth_map = torch.sum(weights * z_vals, -1)
    tmp = z_vals - depth_map.unsqueeze(-1)
    depth_var = torch.sum(weights * tmp * tmp, dim=1)
    return depth_map, depth_var, rgb_map, weights


def sample_pdf(bins, weights, N_samples, u=None):
    """Inverse-CDF resampling.  src/common.py:19-63.  bins [N,M] f64, weights [N,M-1] f32.
    u: optional [N,N_samples] fp32 uniforms; None = deterministic linspace (det=True)."""
    weights = weights + 1e-5
    pdf = weights / torch.sum(weights, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    if u is None:
        u = torch.linspace(0., 1., steps=N_samples).to(cdf.device).expand(list(cdf.shape[:-1]) + [N_samples])
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_b, cdf_a = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    bin_b, bin_a = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - cdf_b) / denom
    return bin_b + t * (bin_a - bin_b)


def render_batch_ray(prm, grids, rays_d, rays_o, stage, gt_depth, bound, *, N_samples=32,
                     N_surface=16, N_importance=0, lindisp=False, occupancy=True, nice=True,
                     u=None, u_importance=None, return_extras=False):
    """Renderer.render_batch_ray.  src/utils/Renderer.py:95-262.

    Returns depth[N] f64, uncertainty[N] f64, color[N,3] f32 (+ dict of intermediates)."""
    if stage == "coarse":
        gt_depth = None
    N = rays_o.shape[0]
    z = sample_z(rays_o, rays_d, gt_depth, bound, N_samples, N_surface, lindisp, u)
    S = z.shape[1]
    pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]      # fp64
    raw = eval_points(pts.reshape(-1, 3), prm, grids, stage, bound, nice).reshape(N, S, 4)
    depth, var, rgb, w = composite(raw, z, rays_d, occupancy)
    extras = {"z_vals": z, "raw": raw, "weights": w}
    if N_importance > 0:
        z_mid = .5 * (z[..., 1:] + z[..., :-1])
        zs = sample_pdf(z_mid, w[..., 1:-1], N_importance, u_importance).detach()
        z, _ = torch.sort(torch.cat([z, zs], -1), -1)
        S = z.shape[1]
        pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]
        raw = eval_points(pts.reshape(-1, 3), prm, grids, stage, bound, nice).reshape(N, S, 4)
        depth, var, rgb, w = composite(raw, z, rays_d, occupancy)
        extras.update({"z_samples": zs, "z_vals2": z, "raw2": raw, "weights2": w})
    if return_extras:
        return depth, var, rgb, extras
    return depth, var, rgb


def regulation_points(rays_d, rays_o, gt_depth, u, N_samples=32):
    """Jittered depths z [N,N_samples] in [0, 0.85 d] and the points o + d z [N*N_samples,3].
    src/utils/Renderer.py:360-394.  u: [N,N_samples] uniforms (always jittered there)."""
    g = gt_depth.reshape(-1, 1).repeat(1, N_samples)
    t = torch.linspace(0., 1., steps=N_samples).to(g.device)
    z = 0.0 * (1. - t) + g * 0.85 * t
    mids = .5 * (z[..., 1:] + z[..., :-1])
    upper = torch.cat([mids, z[..., -1:]], -1)
    lower = torch.cat([z[..., :1], mids], -1)
    z = lower + (upper - lower) * u
    pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]
    return pts.reshape(-1, 3), z


def regulation(prm, grids, rays_d, rays_o, gt_depth, stage, bound, u, N_samples=32, nice=False):
    """iMAP free-space regulariser: densities at the points above.  src/utils/Renderer.py:343-409."""
    pts, _ = regulation_points(rays_d, rays_o, gt_depth, u, N_samples)
    return eval_points(pts, prm, grids, stage, bound, nice)[:, -1]


# --------------------------------------------------------------------------------------
# losses of the two SLAM steps
# --------------------------------------------------------------------------------------
def tracking_loss(depth, var, color, gt_depth, gt_color, w_color=0.5, handle_dynamic=True,
                  use_color=True):
    """src/Tracker.py:117-130."""
    var = var.detach()
    if handle_dynamic:
        tmp = torch.abs(gt_depth - depth) / torch.sqrt(var + 1e-10)
        mask = (tmp < 10 * tmp.median()) & (gt_depth > 0)
    else:
        mask = gt_depth > 0
    loss = (torch.abs(gt_depth - depth) / torch.sqrt(var + 1e-10))[mask].sum()
    if use_color:
        loss = loss + w_color * torch.abs(gt_color - color)[mask].sum()
    return loss


def mapping_loss(depth, color, gt_depth, gt_color, w_color=0.2, with_color=True):
    """src/Mapper.py:604-612."""
    m = gt_depth > 0
    loss = torch.abs(gt_depth[m] - depth[m]).sum()
    if with_color:
        loss = loss + w_color * torch.abs(gt_color - color).sum()
    return loss


# --------------------------------------------------------------------------------------
# scene construction helpers (what NICE_SLAM.__init__ does around the hot path)
# --------------------------------------------------------------------------------------
def round_bound(bound, scale=1.0, divisible=0.32):
    """src/NICE_SLAM.py:196-201: hi is pushed up to lo + k*0.32."""
    b = torch.as_tensor(bound, dtype=torch.float64) * scale
    b = b.clone()
    b[:, 1] = (((b[:, 1] - b[:, 0]) / divisible).int() + 1) * divisible + b[:, 0]
    return b


def init_grids(bound, grid_len=None, c_dim=32, coarse_bound_enlarge=2, coarse=True, gen=None):
    """src/NICE_SLAM.py:243-320: [1,C,Z,Y,X] fp32 normal-initialised grids."""
    grid_len = grid_len or {"coarse": 2, "middle": 0.32, "fine": 0.16, "color": 0.16}
    xyz = bound[:, 1] - bound[:, 0]
    out = {}
    for name, std in (("coarse", 0.01), ("middle", 0.01), ("fine", 0.0001), ("color", 0.01)):
        if name == "coarse" and not coarse:
            continue
        ext = xyz * coarse_bound_enlarge if name == "coarse" else xyz
        shp = list(map(int, (ext / grid_len[name]).tolist()))
        shp[0], shp[2] = shp[2], shp[0]
        out["grid_" + name] = torch.zeros([1, c_dim, *shp]).normal_(mean=0, std=std, generator=gen)
    return out


def _xavier(out_dim, in_dim, gain, gen):
    a = gain * math.sqrt(6.0 / (in_dim + out_dim))
    return (torch.rand(out_dim, in_dim, generator=gen) * 2 - 1) * a


def _default_linear(out_dim, in_dim, gen):
    k = 1.0 / math.sqrt(in_dim)
    w = (torch.rand(out_dim, in_dim, generator=gen) * 2 - 1) * k
    b = (torch.rand(out_dim, generator=gen) * 2 - 1) * k
    return w, b


def _init_mlp(prefix, c_dim, hidden, n_out, gen, n_blocks=5, skips=(2,), embed=93, no_xyz=False):
    prm = {}
    g = math.sqrt(2.0)
    if not no_xyz:
        if c_dim:
            for i in range(n_blocks):
                w, b = _default_linear(hidden, c_dim, gen)
                prm[f"{prefix}fc_c.{i}.weight"], prm[f"{prefix}fc_c.{i}.bias"] = w, b
        prm[f"{prefix}embedder._B"] = torch.randn(3, embed, generator=gen) * 25
        first_in = embed
        skip_extra = embed
    else:
        first_in = hidden
        skip_extra = c_dim
    dims = [first_in] + [hidden + skip_extra if (i in skips) else hidden for i in range(n_blocks - 1)]
    for i, d_in in enumerate(dims):
        prm[f"{prefix}pts_linears.{i}.weight"] = _xavier(hidden, d_in, g, gen)
        prm[f"{prefix}pts_linears.{i}.bias"] = torch.zeros(hidden)
    prm[f"{prefix}output_linear.weight"] = _xavier(n_out, hidden, 1.0, gen)
    prm[f"{prefix}output_linear.bias"] = torch.zeros(n_out)
    return prm


def init_decoder_params(nice=True, coarse=True, c_dim=32, hidden=32, seed=0):
    """Same distributions as the reference constructors (decoder.py:70-85,116-169,239-264,
    316-334; conv_onet/config.py:24-32) but its own RNG stream: tests that need the
    reference's exact weights load them from tests/golden instead."""
    gen = torch.Generator().manual_seed(seed)
    if not nice:
        return _init_mlp("", 0, 256, 4, gen, n_blocks=4, skips=())
    prm = {}
    if coarse:
        prm.update(_init_mlp("coarse_decoder.", c_dim, hidden, 1, gen, no_xyz=True))
    prm.update(_init_mlp("middle_decoder.", c_dim, hidden, 1, gen))
    prm.update(_init_mlp("fine_decoder.", 2 * c_dim, hidden, 1, gen))
    prm.update(_init_mlp("color_decoder.", c_dim, hidden, 4, gen))
    return prm


# --------------------------------------------------------------------------------------
# one iteration of the two SLAM workers (loss side; the Adam step stays in the caller)
# --------------------------------------------------------------------------------------
def stage_of(joint_iter, num_joint_iters, middle_iter_ratio=0.4, fine_iter_ratio=0.6, coarse_mapper=False):
    """Mapping stage schedule.  src/Mapper.py:502-509."""
    if coarse_mapper:
        return "coarse"
    if joint_iter <= int(num_joint_iters * middle_iter_ratio):
        return "middle"
    if joint_iter <= int(num_joint_iters * fine_iter_ratio):
        return "fine"
    return "color"


def tracker_iteration_loss(camera_tensor, prm, grids, gt_color, gt_depth, batch_size, cam, bound, *,
                           ignore_edge_H=20, ignore_edge_W=20, w_color=0.5, handle_dynamic=True,
                           use_color=True, indices=None, render_kw=None):
    """Loss of Tracker.optimize_cam_in_batch (src/Tracker.py:92-130), differentiable in
    camera_tensor.  cam = (H, W, fx, fy, cx, cy)."""
    H, W, fx, fy, cx, cy = cam
    c2w = get_camera_from_tensor(camera_tensor)
    ro, rd, gd, gc = get_samples(ignore_edge_H, H - ignore_edge_H, ignore_edge_W, W - ignore_edge_W,
                                 batch_size, fx, fy, cx, cy, c2w, gt_depth, gt_color, indices)
    with torch.no_grad():
        keep = prefilter_mask(ro, rd, gd, bound)
    ro, rd, gd, gc = ro[keep], rd[keep], gd[keep], gc[keep]
    depth, var, color = render_batch_ray(prm, grids, rd, ro, "color", gd, bound, **(render_kw or {}))
    return tracking_loss(depth, var, color, gd, gc, w_color, handle_dynamic, use_color)


def mapper_iteration_loss(prm, grids, frames, pixels_per_image, cam, bound, stage, *, w_color=0.2,
                          indices=None, nice=True, coarse_mapper=False, render_kw=None):
    """Loss of one joint iteration of Mapper.optimize_map (src/Mapper.py:543-612).
    frames: list of (c2w, gt_depth[H,W], gt_color[H,W,3]); indices: optional list of index
    tensors, one per frame (else drawn like the reference)."""
    H, W, fx, fy, cx, cy = cam
    ros, rds, gds, gcs = [], [], [], []
    for k, (c2w, gd_img, gc_img) in enumerate(frames):
        idx = None if indices is None else indices[k]
        ro, rd, gd, gc = get_samples(0, H, 0, W, pixels_per_image, fx, fy, cx, cy, c2w, gd_img, gc_img, idx)
        ros.append(ro.float()), rds.append(rd.float()), gds.append(gd.float()), gcs.append(gc.float())
    ro, rd, gd, gc = torch.cat(ros), torch.cat(rds), torch.cat(gds), torch.cat(gcs)
    if nice:
        with torch.no_grad():
            keep = prefilter_mask(ro, rd, gd, bound)
        ro, rd, gd, gc = ro[keep], rd[keep], gd[keep], gc[keep]
    depth, var, color = render_batch_ray(prm, grids, rd, ro, stage, None if coarse_mapper else gd, bound,
                                         nice=nice, **(render_kw or {}))
    return mapping_loss(depth, color, gd, gc, w_color, with_color=(not nice) or stage == "color")
