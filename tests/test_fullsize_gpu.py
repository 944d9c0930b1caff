"""Parity and size-independent properties at BASELINE.json's FULL sizes (configs[1]: Replica room0
mapping step, 5 frames x 200 rays x 48 samples on 48.5 MB grids; configs[2]: ScanNet scene0000
tracking batch, 1000 rays on 87.9 MB grids).  The checker is the oracle restatement of the
reference's torch code run on the GPU (same fp32 ATen ops the reference runs on CUDA; ~15 ms per
step, so the full size is affordable here).  Tolerances are north_star's: outputs 1e-4, gradients
1e-3 (relative to the largest reference magnitude)."""
import pytest
import torch

from conftest import assert_close
from oracle import nice_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _scene(shape, n_frames):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from nice_slam_annotation_b200.scene import build_scene
    ns = build_scene(shape, DEV, seed=0, n_frames=n_frames)
    g = torch.Generator(device=DEV).manual_seed(5)
    for v in ns.shared_c.values():                      # moderate features: alpha not saturated everywhere
        v.add_(torch.randn(v.shape, generator=g, device=DEV) * 0.3)
    return ns


@pytest.fixture(scope="module")
def replica():
    return _scene("replica_room0", 5)


@pytest.fixture(scope="module")
def scannet():
    return _scene("scannet_scene0000", 1)


def _batch(ns, per_frame, seed, crop=0):
    from nice_slam_annotation_b200.common import get_samples
    gen = torch.Generator().manual_seed(seed)
    parts = []
    for c2w, depth, color in ns.frames:
        idx = torch.randint((ns.H - 2 * crop) * (ns.W - 2 * crop), (per_frame,), generator=gen)
        parts.append(get_samples(crop, ns.H - crop, crop, ns.W - crop, per_frame, ns.H, ns.W, ns.fx, ns.fy, ns.cx,
                                 ns.cy, c2w, depth, color, DEV, indices=idx.to(DEV)))
    ro, rd, gd, gc = (torch.cat([p[i] for p in parts]).contiguous() for i in range(4))
    keep = O.prefilter_mask(ro, rd, gd, ns.bound.to(DEV)).to(torch.uint8)
    return ro, rd, gd, gc, keep


def _oracle_state(ns, train_prefix):
    prm = {k: v.detach().clone() for k, v in ns.shared_decoders.state_dict().items()}
    for k, v in prm.items():
        v.requires_grad_(k.startswith(train_prefix))
    grids = {k: v.detach().contiguous().clone().requires_grad_(k != "grid_coarse") for k, v in ns.shared_c.items()}
    return prm, grids


def test_replica_mapping_step_vs_oracle(replica):
    """configs[1] at full size: engine step (render forward + fused mapping loss + backward into the
    gradient bucket) against the oracle: depth, colour, loss, the three grid gradients and every
    colour-decoder parameter gradient."""
    from nice_slam_annotation_b200.step import FusedStep
    ns = replica
    ro, rd, gd, gc, keep = _batch(ns, 200, seed=11)
    sel = keep.bool()
    assert int(sel.sum()) > 900
    step = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", 1000, loss="mapping", w_color=0.2,
                     use_color=True, grid_grads=("grid_middle", "grid_fine", "grid_color"),
                     decoder_grads=("color_decoder",))
    assert step.fused_loss
    step.load(ro, rd, gd, gc, keep=keep)
    loss = step.run().clone()
    step.capture()
    loss_g = step.run().clone()                                  # the replayed graph gives the same step
    torch.cuda.synchronize()
    prm, grids = _oracle_state(ns, "color_decoder.")
    d0, _, c0 = O.render_batch_ray(prm, grids, rd[sel], ro[sel], "color", gd[sel], ns.bound.to(DEV))
    l0 = O.mapping_loss(d0, c0, gd[sel], gc[sel])
    l0.backward()
    assert_close(step.depth[sel], d0.detach(), 1e-4, "depth")
    assert_close(step.color[sel], c0.detach(), 1e-4, "color")
    assert abs(float(loss) - float(l0)) <= 1e-4 * abs(float(l0))
    assert abs(float(loss_g) - float(loss)) <= 1e-9 * abs(float(loss))
    for k, key in ((1, "grid_middle"), (2, "grid_fine"), (3, "grid_color")):
        got = step.bucket.view_of(step.grids[k])
        assert float(grids[key].grad.abs().max()) > 0
        assert_close(got, grids[key].grad, 1e-3, "d " + key)
    from nice_slam_annotation_b200.fused import canonical_params
    names = [n for n, _ in ns.shared_decoders.color_decoder.named_parameters()]
    by_ptr = {p.data_ptr(): n for n, p in ns.shared_decoders.color_decoder.named_parameters()}
    for p, g in zip(canonical_params(ns.shared_decoders.color_decoder, 3), step.decoder_grads(3)):
        ref = prm["color_decoder." + by_ptr[p.data_ptr()]].grad
        assert_close(g, ref, 1e-3, "d color_decoder." + by_ptr[p.data_ptr()])
    assert len(names) == len(step.decoder_grads(3))


def test_scannet_tracking_batch_vs_oracle(scannet):
    """configs[2] at full size: render + tracking loss (median mask) + gradient to the rays."""
    from nice_slam_annotation_b200 import fused
    ns = scannet
    ro, rd, gd, gc, keep = _batch(ns, 1000, seed=12, crop=20)
    sel = keep.bool()
    ro_, rd_ = ro[sel].clone().requires_grad_(True), rd[sel].clone().requires_grad_(True)
    depth, var, color = ns.renderer.render_batch_ray(ns.shared_c, ns.shared_decoders, rd_, ro_, DEV, "color", gt_depth=gd[sel])
    loss = fused.tracking_loss(depth, var, color, gd[sel], gc[sel], 0.5, True, True)
    loss.backward()
    prm, grids = _oracle_state(ns, "none")
    ro0, rd0 = ro[sel].clone().requires_grad_(True), rd[sel].clone().requires_grad_(True)
    with torch.enable_grad():
        d0, v0, c0 = O.render_batch_ray(prm, {k: v.detach() for k, v in grids.items()}, rd0, ro0, "color", gd[sel],
                                        ns.bound.to(DEV))
        l0 = O.tracking_loss(d0, v0, c0, gd[sel], gc[sel], w_color=0.5, handle_dynamic=True)
    l0.backward()
    assert_close(depth, d0.detach(), 1e-4, "depth")
    assert_close(var, v0.detach(), 1e-4, "var")
    assert_close(color, c0.detach(), 1e-4, "color")
    assert abs(float(loss) - float(l0)) <= 1e-4 * abs(float(l0))
    assert_close(ro_.grad, ro0.grad, 1e-3, "d rays_o")
    assert_close(rd_.grad, rd0.grad, 1e-3, "d rays_d")


def test_replica_sample_depths_bit_exact_and_sorted(replica):
    from nice_slam_annotation_b200 import fused
    ns = replica
    ro, rd, gd, _, _ = _batch(ns, 200, seed=13)
    z = fused.sample_z(ro, rd, gd, ns.bound, 32, 16)
    z_ref = O.sample_z(ro.cpu(), rd.cpu(), gd.cpu(), ns.bound, 32, 16)
    assert torch.equal(z.cpu(), z_ref)
    assert bool((z[:, 1:] >= z[:, :-1]).all())
    assert int((gd == 0).sum()) > 0                              # the zero-depth branch is exercised


def test_replica_render_is_ray_permutation_equivariant(replica):
    """Each ray depends on the others only through the batch depth max: permuting the batch
    permutes the outputs bit for bit."""
    ns = replica
    ro, rd, gd, _, _ = _batch(ns, 200, seed=14)
    perm = torch.randperm(ro.shape[0], generator=torch.Generator().manual_seed(0)).to(DEV)
    with torch.no_grad():
        a = ns.renderer.render_batch_ray(ns.shared_c, ns.shared_decoders, rd, ro, DEV, "color", gt_depth=gd)
        b = ns.renderer.render_batch_ray(ns.shared_c, ns.shared_decoders, rd[perm].contiguous(), ro[perm].contiguous(),
                                         DEV, "color", gt_depth=gd[perm].contiguous())
    for x, y in zip(a, b):
        assert torch.equal(x[perm], y)


def test_replica_backward_is_linear_in_the_incoming_gradient(replica):
    """backward(a g1 + b g2) = a backward(g1) + b backward(g2) for the grid gradients (the adjoint is a
    linear map; atomics only reorder fp32 sums)."""
    ns = replica
    ro, rd, gd, _, _ = _batch(ns, 200, seed=15)
    gen = torch.Generator(device=DEV).manual_seed(3)
    w = [(torch.randn(1000, generator=gen, device=DEV, dtype=torch.float64), torch.randn(1000, 3, generator=gen, device=DEV))
         for _ in range(2)]

    def grads(cd, cc):
        c = {k: v.detach().requires_grad_(k != "grid_coarse") for k, v in ns.shared_c.items()}
        d, _, col = ns.renderer.render_batch_ray(c, ns.shared_decoders, rd, ro, DEV, "color", gt_depth=gd)
        ((d * cd).sum() + (col * cc).sum().double()).backward()
        return [c[k].grad.clone() for k in ("grid_middle", "grid_fine", "grid_color")]
    g1, g2 = grads(*w[0]), grads(*w[1])
    g12 = grads(0.7 * w[0][0] - 1.3 * w[1][0], 0.7 * w[0][1] - 1.3 * w[1][1])
    for x1, x2, x12 in zip(g1, g2, g12):
        assert_close(x12, 0.7 * x1 - 1.3 * x2, 1e-4, "linearity")


def test_sample_depths_adversarial_rays_bit_exact(replica):
    """k_sample_z ranks the S depths of a ray against each other (stable, NaNs last): rays that leave
    the box before their near plane (descending uniform list), origins outside the box, zero / huge / NaN
    sensor depths and NaN directions must still give exactly torch.sort's answer (Renderer.py:223)."""
    from nice_slam_annotation_b200 import fused
    ns = replica
    g = torch.Generator().manual_seed(21)
    n = 512
    lo, hi = ns.bound[:, 0].float(), ns.bound[:, 1].float()
    ro = lo + (hi - lo) * (torch.rand(n, 3, generator=g) * 1.6 - 0.3)          # some origins outside the box
    rd = torch.randn(n, 3, generator=g)
    rd[::7] *= 1e-3
    rd[5::11, 1] = 0.0                                                         # axis-parallel components
    gd = torch.rand(n, generator=g) * 12
    gd[::5] = 0.0
    gd[3::13] = 1e4
    gd[1::97] = float("nan")
    rd[2::101, 0] = float("nan")
    for depth in (gd, None):
        z_ref = O.sample_z(ro, rd, depth, ns.bound, 32, 16)
        z = fused.sample_z(ro.to(DEV), rd.to(DEV), None if depth is None else depth.to(DEV), ns.bound, 32, 16).cpu()
        same = (z == z_ref) | (z.isnan() & z_ref.isnan())
        assert bool(same.all()), int((~same).sum())
    # jitter + lindisp
    u = torch.rand(n, 32, generator=g)
    ok = ~(gd.isnan() | rd.isnan().any(1))
    z_ref = O.sample_z(ro[ok], rd[ok], gd[ok], ns.bound, 32, 16, lindisp=True, u=u[ok])
    z = fused.sample_z(ro[ok].to(DEV), rd[ok].to(DEV), gd[ok].to(DEV), ns.bound, 32, 16, lindisp=True, u=u[ok].to(DEV)).cpu()
    same = (z == z_ref) | (z.isnan() & z_ref.isnan())
    assert bool(same.all()), int((~same).sum())
