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


# ---- configs[3] / configs[4] at full size -------------------------------------------------------------------------
def _engine_vs_oracle(ns, per_frame, seed, tol_g=1e-3):
    """Engine step (FusedStep, fused mapping loss) against the oracle on the GPU: depth, colour, loss, the three grid
    gradients and the colour-decoder gradients."""
    from nice_slam_annotation_b200.fused import canonical_params
    from nice_slam_annotation_b200.step import FusedStep
    ro, rd, gd, gc, keep = _batch(ns, per_frame, seed=seed)
    sel = keep.bool()
    n = ro.shape[0]
    step = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", n, loss="mapping", w_color=0.2,
                     use_color=True, grid_grads=("grid_middle", "grid_fine", "grid_color"),
                     decoder_grads=("color_decoder",))
    step.load(ro, rd, gd, gc, keep=keep)
    loss = step.run().clone()
    torch.cuda.synchronize()
    prm, grids = _oracle_state(ns, "color_decoder.")
    d0, _, c0 = O.render_batch_ray(prm, grids, rd[sel], ro[sel], "color", gd[sel], ns.bound.to(DEV))
    l0 = O.mapping_loss(d0, c0, gd[sel], gc[sel])
    l0.backward()
    assert_close(step.depth[sel], d0.detach(), 1e-4, "depth")
    assert_close(step.color[sel], c0.detach(), 1e-4, "color")
    assert abs(float(loss) - float(l0)) <= 1e-4 * abs(float(l0))
    for k, key in ((1, "grid_middle"), (2, "grid_fine"), (3, "grid_color")):
        assert float(grids[key].grad.abs().max()) > 0
        assert_close(step.bucket.view_of(step.grids[k]), grids[key].grad, tol_g, "d " + key)
    by_ptr = {p.data_ptr(): nme for nme, p in ns.shared_decoders.color_decoder.named_parameters()}
    for p, g in zip(canonical_params(ns.shared_decoders.color_decoder, 3), step.decoder_grads(3)):
        assert_close(g, prm["color_decoder." + by_ptr[p.data_ptr()]].grad, tol_g, "d color_decoder." + by_ptr[p.data_ptr()])
    return int(sel.sum())


def test_apartment_5000_ray_mapping_step_vs_oracle():
    """configs[3]: Apartment bound (124.9 MB of grids), 5000 rays per iteration (configs/Apartment/apartment.yaml:21-24)."""
    ns = _scene("apartment", 5)
    assert sum(v.numel() * 4 for v in ns.shared_c.values()) > 120e6
    kept = _engine_vs_oracle(ns, 1000, seed=31)
    assert kept > 4500


def test_demo_config_step_vs_oracle():
    """configs[0]: the reference's CPU-runnable case -- Demo config (configs/Demo/demo.yaml: ScanNet frames cropped to
    460x620, bound [[0,6.5],[0,4],[0,3.5]], 5.8 MB of grids), 1000 rays, stage color, forward + backward."""
    ns = _scene("demo", 5)
    kept = _engine_vs_oracle(ns, 200, seed=30)
    assert kept > 800


def test_tum_scale_grid_mapping_step_vs_oracle():
    """TUM fr3_office grid lengths (0.16 / 0.08 / 0.08 m): 127^3-class fine / colour grids, 556 MB, far beyond L2 and
    beyond 2^31 bytes of voxel offsets x channel (the kernels index voxels with 32-bit, bytes with 64-bit arithmetic)."""
    ns = _scene("tum_fr3_office", 5)
    assert sum(v.numel() * 4 for v in ns.shared_c.values()) > 500e6
    _engine_vs_oracle(ns, 200, seed=32)


def test_replica_coarse_stage_vs_oracle(replica):
    """stage 'coarse' (MLP_no_xyz on the enlarged coarse grid, 32 samples, no sensor depth in the sampler) at the full
    Replica batch through Renderer.render_batch_ray + autograd."""
    from nice_slam_annotation_b200 import fused
    ns = replica
    ro, rd, gd, gc, keep = _batch(ns, 200, seed=33)
    sel = keep.bool()
    dec = ns.shared_decoders
    for p in dec.parameters():
        p.requires_grad_(False)
    grid = ns.shared_c["grid_coarse"].detach().clone().requires_grad_(True)
    c = dict(ns.shared_c, grid_coarse=grid)
    depth, var, color = ns.renderer.render_batch_ray(c, dec, rd[sel], ro[sel], DEV, "coarse", gt_depth=gd[sel])
    loss = fused.mapping_loss(depth, color, gd[sel], gc[sel], 0.2, use_color=False)
    loss.backward()
    prm, grids = _oracle_state(ns, "none")
    grids["grid_coarse"].requires_grad_(True)
    d0, v0, c0 = O.render_batch_ray(prm, grids, rd[sel], ro[sel], "coarse", gd[sel], ns.bound.to(DEV))
    l0 = O.mapping_loss(d0, c0, gd[sel], gc[sel], with_color=False)
    l0.backward()
    assert_close(depth, d0.detach(), 1e-4, "depth")
    assert_close(var, v0.detach(), 1e-4, "var")
    assert abs(float(loss) - float(l0)) <= 1e-4 * abs(float(l0))
    assert float(grids["grid_coarse"].grad.abs().max()) > 0
    assert_close(grid.grad, grids["grid_coarse"].grad, 1e-3, "d grid_coarse")


def test_imap_5000_ray_mapping_step_vs_oracle():
    """configs[4], the iMAP* stress case at its configured size (configs/imap.yaml: 5000 pixels, 32 + 12 samples, density
    compositing, regulariser): render forward + mapping loss + 5e-4 x regulation + backward to all 11 parameter tensors
    of the 4 x 256 MLP against the oracle on the GPU (fp32 torch ops)."""
    import bench
    from nice_slam_annotation_b200 import fused
    from nice_slam_annotation_b200.common import get_samples
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    sc = bench.imap_scene(DEV)
    n = 5000
    idx = torch.randint(sc.H * sc.W, (n,), generator=torch.Generator().manual_seed(3))
    ro, rd, gd, gc = get_samples(0, sc.H, 0, sc.W, n, sc.H, sc.W, sc.fx, sc.fy, sc.cx, sc.cy, sc.c2w, sc.depth, sc.color,
                                 DEV, indices=idx)
    ro, rd = ro.contiguous(), rd.contiguous()
    bound = sc.r.bound
    torch.manual_seed(7)
    u = torch.rand(n, 32)                                     # the draw Renderer.regulation makes (CPU generator)
    torch.manual_seed(7)
    depth, _, color = sc.r.render_batch_ray({}, sc.dec, rd, ro, DEV, "color", gt_depth=gd)
    sigma = sc.r.regulation({}, sc.dec, rd, ro, gd, DEV, "color")
    loss = fused.mapping_loss(depth, color, gd, gc, 0.05, use_color=True) + 0.0005 * sigma.abs().sum()
    loss.backward()
    prm = {k: v.detach().clone().requires_grad_(True) for k, v in sc.dec.state_dict().items()}
    d0, _, c0 = O.render_batch_ray(prm, {}, rd, ro, "color", gd, bound.to(DEV), N_samples=32, N_surface=0, N_importance=12,
                                   occupancy=False, nice=False)
    s0 = O.regulation(prm, {}, rd, ro, gd, "color", bound.to(DEV), u.to(DEV), N_samples=32, nice=False)
    l0 = O.mapping_loss(d0, c0, gd, gc, w_color=0.05) + 0.0005 * s0.abs().sum()
    l0.backward()
    assert_close(depth, d0.detach(), 1e-4, "depth")
    assert_close(color, c0.detach(), 1e-4, "color")
    assert_close(sigma, s0.detach(), 1e-4, "regulation sigma")
    assert abs(float(loss) - float(l0)) <= 1e-4 * abs(float(l0))
    for k, p in sc.dec.named_parameters():
        assert_close(p.grad, prm[k].grad, 1e-3, "d " + k)
