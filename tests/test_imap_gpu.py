"""GPU parity of the iMAP* path (single MLP, density compositing, importance resampling, regulariser)
against golden vectors produced by the reference itself (tests/golden/make_golden.py: ``imap`` =
hidden 64, ``imap256`` = the factory's hidden 256) and against the CPU oracle on seeded points.

Tolerances (BASELINE.json north_star): outputs 1e-4, gradients 1e-3, max|err| / max|ref|.
"""
import types

import pytest
import torch

from conftest import assert_close, sub
from oracle import nice_oracle as O

pytestmark = pytest.mark.gpu
TOL_OUT, TOL_GRAD = 1e-4, 1e-3
DEV = "cuda:0"


@pytest.fixture(scope="module")
def gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")


def _decoder(d):
    from nice_slam_annotation_b200.conv_onet.models.decoder import MLP
    hidden = d["dec.pts_linears.0.weight"].shape[0]
    dec = MLP(name="", dim=3, c_dim=0, color=True, hidden_size=hidden, skips=[], n_blocks=4,
              pos_embedding_method="fourier")
    dec.load_state_dict({k: torch.as_tensor(v) for k, v in sub(d, "dec.").items()})
    return dec.to(DEV)


def _renderer(d, dec, perturb=0.0):
    from nice_slam_annotation_b200.utils.Renderer import Renderer
    cfg = {"rendering": {"lindisp": False, "perturb": perturb, "N_samples": 32, "N_surface": 0, "N_importance": 12},
           "scale": 0.1, "occupancy": False, "model": {"coarse_bound_enlarge": 2}}
    slam = types.SimpleNamespace(nice=False, bound=torch.as_tensor(d["bound"]), H=48, W=64, fx=60.0, fy=60.0,
                                 cx=31.5, cy=23.5)
    return Renderer(cfg, None, slam)


def _rays(d):
    ro = torch.as_tensor(d["rays_o"]).to(DEV).requires_grad_(True)
    rd = torch.as_tensor(d["rays_d"]).to(DEV).requires_grad_(True)
    return ro, rd, torch.as_tensor(d["gt_depth"]).to(DEV)


@pytest.mark.parametrize("name", ["imap", "imap256"])
def test_render_forward_backward_vs_reference(gpu, golden, name):
    d = golden(name)
    dec = _decoder(d)
    r = _renderer(d, dec)
    ro, rd, gd = _rays(d)
    depth, var, color = r.render_batch_ray({}, dec, rd, ro, DEV, "color", gt_depth=gd)
    ref = sub(d, "det.")
    assert depth.dtype == torch.float64 and var.dtype == torch.float64 and color.dtype == torch.float32
    assert_close(depth.cpu(), ref["depth"], TOL_OUT, "depth")
    assert_close(var.cpu(), ref["var"], TOL_OUT, "var")
    assert_close(color.cpu(), ref["color"], TOL_OUT, "color")
    a, w = torch.as_tensor(d["coef_a"]).to(DEV), torch.as_tensor(d["coef_w"]).to(DEV)
    ((depth * a).sum() + (color * w).sum()).backward()
    assert_close(ro.grad.cpu(), ref["g_rays_o"], TOL_GRAD, "g_rays_o")
    assert_close(rd.grad.cpu(), ref["g_rays_d"], TOL_GRAD, "g_rays_d")
    n = 0
    for k, p in dec.named_parameters():
        assert p.grad is not None, k
        assert_close(p.grad.cpu(), ref["gp_" + k], TOL_GRAD, "gp_" + k)
        n += 1
    assert n == 11


@pytest.mark.parametrize("name", ["imap", "imap256"])
def test_render_perturbed_and_regulation(gpu, golden, name):
    """perturb > 0: the Renderer draws rand(N, 32) then rand(N, 12) from the CPU generator exactly
    like Renderer.py:218 / common.py:36; regulation draws rand(N, 32) (Renderer.py:385)."""
    d = golden(name)
    dec = _decoder(d)
    r = _renderer(d, dec, perturb=1.0)
    ro, rd, gd = _rays(d)
    torch.manual_seed(21)
    with torch.no_grad():
        depth, var, color = r.render_batch_ray({}, dec, rd, ro, DEV, "color", gt_depth=gd)
    ref = sub(d, "rand.")
    assert_close(depth.cpu(), ref["depth"], TOL_OUT, "depth")
    assert_close(var.cpu(), ref["var"], TOL_OUT, "var")
    assert_close(color.cpu(), ref["color"], TOL_OUT, "color")
    torch.manual_seed(31)
    with torch.no_grad():
        sig = r.regulation({}, dec, rd.detach(), ro.detach(), gd, DEV, "color")
    assert_close(sig.cpu(), d["reg.sigma"], TOL_OUT, "regulation sigma")


@pytest.mark.parametrize("n,S", [(0, 32), (1, 32), (333, 32), (5000, 32), (77, 1), (64, 48)])
def test_regulation_points_bit_exact_and_adjoint(gpu, n, S):
    """k_regulation_points: depths and points equal the eager expressions of Renderer.py:381-394 (oracle
    regulation_points) bit for bit -- fp32 throughout, including rays without sensor depth (d = 0) and a NaN
    depth; the adjoint towards the rays equals autograd through the oracle."""
    from nice_slam_annotation_b200 import fused
    g = torch.Generator().manual_seed(n * 131 + S)
    ro = (torch.randn(n, 3, generator=g) * 2).to(DEV)
    rd = torch.randn(n, 3, generator=g).to(DEV)
    gd = (torch.rand(n, generator=g) * 6).to(DEV)
    if n > 10:
        gd[3] = 0.0
        gd[5] = float("nan")
    u = torch.rand(n, S, generator=g).to(DEV)
    ro_a, rd_a = ro.clone().requires_grad_(True), rd.clone().requires_grad_(True)
    pts, z = fused.regulation_points(ro_a, rd_a, gd, S, u, return_z=True)
    ro_b, rd_b = ro.clone().requires_grad_(True), rd.clone().requires_grad_(True)
    pts0, z0 = O.regulation_points(rd_b, ro_b, gd, u, S)
    assert pts.shape == (n * S, 3) and z.shape == (n, S)
    assert torch.equal(torch.nan_to_num(z, nan=-7.0), torch.nan_to_num(z0, nan=-7.0))
    assert torch.equal(torch.nan_to_num(pts, nan=-7.0), torch.nan_to_num(pts0, nan=-7.0))
    if n == 0:
        return
    w = torch.randn(n * S, 3, generator=g).to(DEV)
    finite = torch.isfinite(pts0).all(dim=1, keepdim=True)
    (pts * w).masked_fill(~finite, 0).sum().backward()
    (pts0 * w).masked_fill(~finite, 0).sum().backward()
    ok = torch.isfinite(gd)
    assert_close(ro_a.grad[ok].cpu(), ro_b.grad[ok].cpu(), 1e-5, "d rays_o")
    assert_close(rd_a.grad[ok].cpu(), rd_b.grad[ok].cpu(), 1e-5, "d rays_d")


def test_box_prefilter_vs_reference_expression(gpu):
    """k_box_prefilter against the fp64 expression of Tracker.py:100-107 / Mapper.py:587-594, including rays
    parallel to an axis (division by zero) and depth exactly on the exit."""
    from nice_slam_annotation_b200 import fused
    g = torch.Generator().manual_seed(3)
    n = 4096
    bound = torch.tensor([[-2.9, 8.9], [-3.2, 5.5], [-3.5, 3.3]])
    ro = (torch.rand(n, 3, generator=g) * 4 - 1).to(DEV)
    rd = torch.randn(n, 3, generator=g).to(DEV)
    rd[::17, 0] = 0.0
    rd[::29, 2] = 0.0
    gd = (torch.rand(n, generator=g) * 8).to(DEV)
    b = bound.to(torch.float64).to(DEV)
    t = (b.unsqueeze(0) - ro.unsqueeze(-1)) / rd.unsqueeze(-1)
    t, _ = torch.min(torch.max(t, dim=2)[0], dim=1)
    gd[::13] = t[::13].float()                                  # ties at fp32 resolution
    want = (t >= gd).to(torch.uint8)
    got = fused.box_prefilter(ro, rd, gd, bound)
    assert torch.equal(got, want), int((got != want).sum())
    assert 0 < int(want.sum()) < n


@pytest.mark.parametrize("P", [1, 127, 5000, 300000])
def test_mlp_forward_backward_vs_oracle(gpu, golden, P):
    """MLP.forward on explicit points (ragged tile counts, a forward-only call larger than one
    workspace chunk) and its gradients against the oracle MLP on the CPU."""
    d = golden("imap256")
    dec = _decoder(d)
    prm = {k: v.clone().requires_grad_(P <= 5000) for k, v in sub(d, "dec.").items()}
    g = torch.Generator().manual_seed(P)
    p = (torch.rand(P, 3, generator=g) - 0.5) * 2.0
    if P > 5000:
        with torch.no_grad():
            out = dec(p.to(DEV))
            ref = O.mlp_forward(prm, p.double(), None, n_blocks=4, skips=())
        assert_close(out.cpu(), ref, TOL_OUT, "raw")
        return
    pg = p.to(DEV).requires_grad_(True)
    out = dec(pg)
    assert out.shape == (P, 4)
    pc = p.double().requires_grad_(True)
    ref = O.mlp_forward(prm, pc, None, n_blocks=4, skips=())
    assert_close(out.detach().cpu(), ref.detach(), TOL_OUT, "raw")
    # a ReLU unit within rounding distance of zero may take either branch, which moves the gradient of
    # that point by a whole term: such points (a handful) get zero weight in the test functional
    with torch.no_grad():
        h, margin = O.fourier_embed(p, prm["embedder._B"]), torch.full((P,), float("inf"))
        for i in range(4):
            pre = h @ prm[f"pts_linears.{i}.weight"].t() + prm[f"pts_linears.{i}.bias"]
            margin = torch.minimum(margin, pre.abs().min(1)[0])
            h = torch.relu(pre)
    safe = margin > 2e-5
    assert safe.float().mean() > 0.9
    w = torch.randn(P, 4, generator=g) * safe[:, None]
    (out * w.to(DEV)).sum().backward()
    (ref * w).sum().backward()
    assert_close(pg.grad.cpu(), pc.grad, TOL_GRAD, "d_point")
    for k, q in dec.named_parameters():
        assert_close(q.grad.cpu(), prm[k].grad, TOL_GRAD, "gp_" + k)


def test_eval_points_out_of_bound_rule(gpu, golden):
    """Renderer.eval_points: density 100 outside the bound and no gradient through it (Renderer.py:63-66,87)."""
    d = golden("imap")
    dec = _decoder(d)
    r = _renderer(d, dec)
    b = torch.as_tensor(d["bound"])
    g = torch.Generator().manual_seed(2)
    p = b[:, 0] + (b[:, 1] - b[:, 0]) * (torch.rand(400, 3, generator=g) * 1.4 - 0.2)
    inside = ((p < b[:, 1]) & (p > b[:, 0])).all(1)
    raw = r.eval_points(p.to(DEV), dec, None, "color", DEV)
    assert bool((raw[~inside.to(DEV), 3] == 100).all())
    prm = sub(d, "dec.")
    ref = O.eval_points(p.double(), prm, {}, "color", b, nice=False)
    assert_close(raw.detach().cpu(), ref, TOL_OUT, "raw")
    raw.sum().backward()
    assert dec.output_linear.bias.grad is not None
    assert abs(float(dec.output_linear.bias.grad[3]) - float(inside.sum())) < 1e-3
