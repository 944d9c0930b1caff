"""GPU parity of the CUDA render path against (a) the golden vectors produced by running the
reference itself (tests/golden/make_golden.py) and (b) the CPU oracle on seeded inputs.

Tolerances (BASELINE.json north_star): rendered depth / colour within 1e-4, gradients within
1e-3, measured as max|err| / max|ref| per tensor.
"""
import numpy as np
import pytest
import torch

from conftest import assert_close, sub
from oracle import nice_oracle as O

pytestmark = pytest.mark.gpu

TOL_OUT, TOL_GRAD = 1e-4, 1e-3


@pytest.fixture(scope="module")
def gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import gpu_util
    return gpu_util


def _render(gpu, d, stage, ro, rd, gd, grids, dec, **kw):
    from nice_slam_annotation_b200 import fused
    args = dict(N_samples=32, N_surface=16)
    args.update(kw)
    return fused.render_batch_ray(grids, dec, rd, ro, stage, gd, torch.as_tensor(d["bound"]), **args)


@pytest.mark.parametrize("stage", ["coarse", "middle", "fine", "color"])
def test_small_scene_forward_backward_vs_reference(gpu, golden, stage):
    d = golden("small_nice")
    dec = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    grids = gpu.grids_from(d, requires_grad=True)
    ro = torch.as_tensor(d["rays_o"]).to(gpu.DEV).requires_grad_(True)
    rd = torch.as_tensor(d["rays_d"]).to(gpu.DEV).requires_grad_(True)
    gd = torch.as_tensor(d["gt_depth"]).to(gpu.DEV)
    depth, var, color = _render(gpu, d, stage, ro, rd, gd, grids, dec)
    ref = sub(d, stage + ".")
    assert depth.dtype == torch.float64 and var.dtype == torch.float64 and color.dtype == torch.float32
    assert_close(depth.cpu(), ref["depth"], TOL_OUT, "depth")
    assert_close(var.cpu(), ref["var"], TOL_OUT, "var")
    if stage == "color":
        assert_close(color.cpu(), ref["color"], TOL_OUT, "color")
    else:
        assert float(color.abs().max()) == 0.0
    a = torch.as_tensor(d["coef_a"]).to(gpu.DEV)
    b = torch.as_tensor(d["coef_b"]).to(gpu.DEV)
    w = torch.as_tensor(d["coef_w"]).to(gpu.DEV)
    ((depth * a).sum() + (var * b).sum() + (color * w).sum()).backward()
    assert_close(ro.grad.cpu(), ref["g_rays_o"], TOL_GRAD, "g_rays_o")
    assert_close(rd.grad.cpu(), ref["g_rays_d"], TOL_GRAD, "g_rays_d")
    for k, g in grids.items():
        if "g_" + k in ref:
            assert g.grad is not None, k
            assert_close(g.grad.cpu(), ref["g_" + k], TOL_GRAD, "g_" + k)
        else:
            assert g.grad is None or float(g.grad.abs().max()) == 0.0, k
    _check_param_grads(dec, ref)


def _check_param_grads(dec, ref):
    """decoder parameter gradients against the reference's (gp_<state_dict name>)"""
    n = 0
    for name, p in dec.named_parameters():
        key = "gp_" + name
        if key in ref:
            assert p.grad is not None, name
            assert_close(p.grad.cpu(), ref[key], TOL_GRAD, key)
            n += 1
        else:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, name
    assert n > 0


def test_no_gt_depth(gpu, golden):
    d = golden("small_nice")
    dec = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    grids = gpu.grids_from(d)
    ro = torch.as_tensor(d["rays_o"]).to(gpu.DEV)
    rd = torch.as_tensor(d["rays_d"]).to(gpu.DEV)
    with torch.no_grad():
        depth, var, color = _render(gpu, d, "middle", ro, rd, None, grids, dec)
    ref = sub(d, "middle_nogt.")
    assert depth.shape == ref["depth"].shape
    assert_close(depth.cpu(), ref["depth"], TOL_OUT, "depth")
    assert_close(var.cpu(), ref["var"], TOL_OUT, "var")


@pytest.mark.parametrize("stage", ["coarse", "middle", "fine", "color"])
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_eval_points(gpu, golden, stage, dtype):
    from nice_slam_annotation_b200 import fused
    d = golden("small_nice")
    dec = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    grids = gpu.grids_from(d)
    p = torch.as_tensor(d["evalpts.p"])
    raw = fused.eval_points(p.to(gpu.DEV).to(dtype), dec, grids, stage, torch.as_tensor(d["bound"]))
    if dtype == torch.float64:
        ref = torch.as_tensor(d[f"evalpts.{stage}"])
    else:   # fp32 points are normalised in fp32 by the reference: oracle restates that
        prm = gpu.oracle_params(sub(d, "dec."))
        c = {k: v for k, v in sub(d, "c.").items()}
        ref = O.eval_points(p.float(), prm, c, stage, torch.as_tensor(d["bound"]))
    assert raw.shape == ref.shape
    assert_close(raw.cpu(), ref, TOL_OUT, f"raw {stage}")
    # out-of-bound points (incl. points exactly on the bound) carry occupancy 100
    assert torch.equal(raw.cpu()[:, 3] == 100.0, ref[:, 3] == 100.0)


@pytest.mark.parametrize("name,kw", [("lindisp", dict(lindisp=True)), ("perturb", dict()),
                                      ("odd", dict(N_samples=20, N_surface=7))])
def test_sampling_variants(gpu, golden, name, kw):
    d = golden("sampling_variants")
    dec = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    grids = gpu.grids_from(d)
    ro = torch.as_tensor(d[f"{name}.rays_o"]).to(gpu.DEV)
    rd = torch.as_tensor(d[f"{name}.rays_d"]).to(gpu.DEV)
    gd = torch.as_tensor(d[f"{name}.gt_depth"]).to(gpu.DEV)
    if name == "perturb":
        kw = dict(u=torch.as_tensor(d["perturb.u"]))
    with torch.no_grad():
        depth, var, color = _render(gpu, d, "color", ro, rd, gd, grids, dec, **kw)
    assert_close(depth.cpu(), d[f"{name}.depth"], TOL_OUT, "depth")
    assert_close(var.cpu(), d[f"{name}.var"], TOL_OUT, "var")
    assert_close(color.cpu(), d[f"{name}.color"], TOL_OUT, "color")


def test_sample_z_matches_oracle_exactly(gpu, golden):
    """fp64 sample depths are reproduced bit for bit (same expression order as the reference)."""
    from nice_slam_annotation_b200 import fused
    d = golden("small_nice")
    ro, rd, gd = (torch.as_tensor(d[k]) for k in ("rays_o", "rays_d", "gt_depth"))
    bound = torch.as_tensor(d["bound"])
    z_ref = O.sample_z(ro, rd, gd, bound, 32, 16)
    z = fused.sample_z(ro.to(gpu.DEV), rd.to(gpu.DEV), gd.to(gpu.DEV), bound, 32, 16)
    assert torch.equal(z.cpu(), z_ref)
    z_ref = O.sample_z(ro, rd, None, bound, 32, 16)
    z = fused.sample_z(ro.to(gpu.DEV), rd.to(gpu.DEV), None, bound, 32, 16)
    assert torch.equal(z.cpu(), z_ref)


@pytest.mark.parametrize("stage", ["middle", "fine", "color"])
def test_trained_scene_vs_reference(gpu, golden, stage):
    """F2 fixture: state produced by the reference's own mapping step (meaningful depths)."""
    d = golden("trained_nice")
    dec = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    grids = gpu.grids_from(d, requires_grad=True)
    ro = torch.as_tensor(d["r.rays_o"]).to(gpu.DEV).requires_grad_(True)
    rd = torch.as_tensor(d["r.rays_d"]).to(gpu.DEV).requires_grad_(True)
    gd = torch.as_tensor(d["r.gt_depth"]).to(gpu.DEV)
    depth, var, color = _render(gpu, d, stage, ro, rd, gd, grids, dec)
    ref = sub(d, f"r.{stage}.")
    assert_close(depth.cpu(), ref["depth"], TOL_OUT, "depth")
    assert_close(var.cpu(), ref["var"], TOL_OUT, "var")
    assert_close(color.cpu(), ref["color"], TOL_OUT, "color")
    a = torch.as_tensor(d["r.coef_a"]).to(gpu.DEV)
    w = torch.as_tensor(d["r.coef_w"]).to(gpu.DEV)
    ((depth * a).sum() + (color * w).sum()).backward()
    assert_close(ro.grad.cpu(), ref["g_rays_o"], TOL_GRAD, "g_rays_o")
    assert_close(rd.grad.cpu(), ref["g_rays_d"], TOL_GRAD, "g_rays_d")
    for k, g in grids.items():
        if "g_" + k in ref:
            assert_close(g.grad.cpu(), ref["g_" + k], TOL_GRAD, "g_" + k)
    _check_param_grads(dec, ref)
