"""The autograd-free engine step (step.FusedStep: render forward + loss + render backward as
three library calls, optionally one CUDA graph) against the autograd path through
Renderer.render_batch_ray, and the fused losses against the oracle's restatement of
src/Mapper.py:604-612 / src/Tracker.py:113-130."""
import pytest
import torch

from conftest import assert_close
from oracle import nice_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def scene():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from nice_slam_annotation_b200 import synth
    from nice_slam_annotation_b200.scene import build_scene
    synth.SHAPES["step_test"] = (48, 64, 60.0, 60.0, 31.5, 23.5, [[-0.6, 0.9], [-0.5, 0.7], [-0.4, 0.5]])
    ns = build_scene("step_test", DEV, seed=0, n_frames=2)
    g = torch.Generator().manual_seed(5)
    for v in ns.shared_c.values():
        v.add_((torch.randn(v.shape, generator=g) * 0.3).to(v.device))
    return ns


def _batch(ns, n, seed):
    from nice_slam_annotation_b200.common import get_samples
    gen = torch.Generator().manual_seed(seed)
    parts = []
    for c2w, depth, color in ns.frames:
        idx = torch.randint(ns.H * ns.W, (n // len(ns.frames),), generator=gen)
        parts.append(get_samples(0, ns.H, 0, ns.W, idx.numel(), ns.H, ns.W, ns.fx, ns.fy, ns.cx, ns.cy, c2w, depth,
                                 color, DEV, indices=idx))
    ro, rd, gd, gc = (torch.cat([p[i] for p in parts]) for i in range(4))
    keep = O.prefilter_mask(ro.cpu(), rd.cpu(), gd.cpu(), ns.bound).to(DEV).to(torch.uint8)
    return ro, rd, gd, gc, keep


def _autograd_reference(ns, batch, kind, trainable):
    from nice_slam_annotation_b200 import fused
    ro, rd, gd, gc, keep = batch
    dec = ns.shared_decoders
    for name, p in dec.named_parameters():
        p.requires_grad_(name.startswith(trainable))
        p.grad = None
    grids = {k: v.detach().clone(memory_format=torch.preserve_format).requires_grad_(k != "grid_coarse")
             for k, v in ns.shared_c.items()}
    ro_, rd_ = ro.clone().requires_grad_(kind == "tracking"), rd.clone().requires_grad_(kind == "tracking")
    depth, var, color = ns.renderer.render_batch_ray(grids, dec, rd_, ro_, DEV, "color", gt_depth=gd, keep=keep)
    if kind == "mapping":
        loss = fused.mapping_loss(depth, color, gd, gc, 0.2, True, keep=keep)
    else:
        loss = fused.tracking_loss(depth, var, color, gd, gc, 0.5, True, True, keep=keep)
    loss.backward()
    return loss.detach(), grids, (ro_.grad, rd_.grad), depth.detach(), var.detach(), color.detach()


@pytest.mark.parametrize("graph", [False, True])
def test_mapping_step_matches_autograd_path(scene, graph):
    from nice_slam_annotation_b200.step import FusedStep
    ns = scene
    batch = _batch(ns, 128, seed=3)
    loss0, grids0, _, depth0, _, color0 = _autograd_reference(ns, batch, "mapping", "color_decoder")
    pgrads0 = {n: p.grad.clone() for n, p in ns.shared_decoders.named_parameters() if p.grad is not None}
    step = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", 160, loss="mapping", w_color=0.2,
                     grid_grads=("grid_middle", "grid_fine", "grid_color"), decoder_grads=("color_decoder",))
    step.load(*batch[:4], keep=batch[4])       # 128 rays in a 160-ray step: the tail is masked off
    step.run()
    if graph:
        step.capture()
        step.load(*batch[:4], keep=batch[4])
        step.run()
    torch.cuda.synchronize()
    assert_close(step.loss.cpu(), loss0.cpu().reshape(1), 1e-6, "loss")
    assert_close(step.depth[:128].cpu(), depth0.cpu(), 1e-6, "depth")
    assert_close(step.color[:128].cpu(), color0.cpu(), 1e-6, "color")
    for k in ("grid_middle", "grid_fine", "grid_color"):
        g = step.bucket.view_of(step.grids[{"grid_middle": 1, "grid_fine": 2, "grid_color": 3}[k]])
        assert g.shape == grids0[k].grad.shape
        assert_close(g.cpu(), grids0[k].grad.cpu(), 1e-5, "g_" + k)
    step.assign_grads(ns.shared_c)
    n = 0
    for name, p in ns.shared_decoders.named_parameters():
        if name in pgrads0:
            assert_close(p.grad.cpu(), pgrads0[name].cpu(), 1e-5, "gp_" + name)
            n += 1
    assert n == 23


def test_tracking_step_matches_autograd_path(scene):
    from nice_slam_annotation_b200.step import FusedStep
    ns = scene
    batch = _batch(ns, 96, seed=11)
    loss0, _, (g_o0, g_d0), depth0, var0, _ = _autograd_reference(ns, batch, "tracking", "none")
    step = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", 96, loss="tracking", w_color=0.5,
                     use_color=True, handle_dynamic=True, ray_grads=True)
    step.load(*batch[:4], keep=batch[4])
    step.run()
    torch.cuda.synchronize()
    assert_close(step.loss.cpu(), loss0.cpu().reshape(1), 1e-6, "loss")
    assert_close(step.var.cpu(), var0.cpu(), 1e-6, "var")
    assert_close(step.g_rays_o.cpu(), g_o0.cpu(), 1e-5, "g_rays_o")
    assert_close(step.g_rays_d.cpu(), g_d0.cpu(), 1e-5, "g_rays_d")


@pytest.mark.parametrize("handle_dynamic", [False, True])
def test_fused_losses_vs_oracle(scene, handle_dynamic):
    """value and gradient of the two losses against the oracle's torch restatement"""
    from nice_slam_annotation_b200 import fused
    g = torch.Generator().manual_seed(2)
    n = 301
    depth = (torch.rand(n, generator=g, dtype=torch.float64) * 3 + 0.2)
    var = torch.rand(n, generator=g, dtype=torch.float64) * 0.1
    color = torch.rand(n, 3, generator=g)
    gd = (torch.rand(n, generator=g) * 3 + 0.2)
    gd[::7] = 0.0
    gc = torch.rand(n, 3, generator=g)
    d0, c0 = depth.clone().requires_grad_(True), color.clone().requires_grad_(True)
    l0 = O.tracking_loss(d0, var, c0, gd, gc, w_color=0.5, use_color=True, handle_dynamic=handle_dynamic)
    l0.backward()
    d1, c1 = depth.to(DEV).requires_grad_(True), color.to(DEV).requires_grad_(True)
    l1 = fused.tracking_loss(d1, var.to(DEV), c1, gd.to(DEV), gc.to(DEV), 0.5, True, handle_dynamic)
    l1.backward()
    assert_close(l1.detach().cpu().reshape(1), l0.detach().reshape(1), 1e-7, "tracking loss")
    assert_close(d1.grad.cpu(), d0.grad, 1e-7, "d loss/d depth")
    assert_close(c1.grad.cpu(), c0.grad, 1e-6, "d loss/d color")
    d0, c0 = depth.clone().requires_grad_(True), color.clone().requires_grad_(True)
    l0 = O.mapping_loss(d0, c0, gd, gc, w_color=0.2)
    l0.backward()
    d1, c1 = depth.to(DEV).requires_grad_(True), color.to(DEV).requires_grad_(True)
    l1 = fused.mapping_loss(d1, c1, gd.to(DEV), gc.to(DEV), 0.2, True)
    l1.backward()
    assert_close(l1.detach().cpu().reshape(1), l0.detach().reshape(1), 1e-6, "mapping loss")
    assert_close(d1.grad.cpu(), d0.grad, 1e-7, "d loss/d depth")
    assert_close(c1.grad.cpu(), c0.grad, 1e-6, "d loss/d color")


@pytest.mark.parametrize("n", [1, 2, 33, 1000, 4096, 5000])
def test_tracking_loss_median_edge_cases(scene, n):
    """The dynamic-object threshold is 10 x torch.median of the live residuals (Tracker.py:118-121): ties, a
    NaN residual (the median is NaN then), masked rays and both median paths (rank counting <= 4096 rays,
    sort above)."""
    from nice_slam_annotation_b200 import fused
    g = torch.Generator().manual_seed(n)
    depth = torch.rand(n, generator=g, dtype=torch.float64) * 3 + 0.2
    var = torch.rand(n, generator=g, dtype=torch.float64) * 0.1
    gd = torch.rand(n, generator=g) * 3 + 0.2
    if n > 8:
        depth[5:n // 2] = depth[4]; var[5:n // 2] = var[4]; gd[5:n // 2] = gd[4]     # ties around the median
        gd[-3:] += 40.0                                                                # above the threshold
    color = torch.rand(n, 3, generator=g)
    gc = torch.rand(n, 3, generator=g)
    keep = torch.rand(n, generator=g) > 0.2 if n > 2 else torch.ones(n, dtype=torch.bool)
    for case, kp in enumerate((None, keep, None)):
        sel = slice(None) if kp is None else kp
        if case == 2:
            depth[n // 2] = float("nan")        # torch.median is NaN then: nothing passes the threshold
        d0 = depth.clone().requires_grad_(True)
        l0 = O.tracking_loss(d0[sel], var[sel], color[sel], gd[sel], gc[sel], w_color=0.5, use_color=True,
                             handle_dynamic=True)
        l0.backward()
        d1 = depth.to(DEV).requires_grad_(True)
        l1 = fused.tracking_loss(d1, var.to(DEV), color.to(DEV), gd.to(DEV), gc.to(DEV), 0.5, True, True,
                                 keep=None if kp is None else kp.to(DEV))
        l1.backward()
        assert_close(l1.detach().cpu().reshape(1), l0.detach().reshape(1), 1e-7, "tracking loss")
        g1, g0 = torch.nan_to_num(d1.grad.cpu()), torch.nan_to_num(d0.grad)
        assert torch.equal(g1 != 0, g0 != 0), "threshold mask"
        assert_close(g1, g0, 1e-12, "d loss/d depth")


def test_render_img_vs_oracle(scene):
    """f3: Renderer.render_img (Renderer.py:264-340) = get_rays + render_batch_ray over the whole
    frame (one 100k-ray batch here), against the oracle on every pixel."""
    ns = scene
    c2w, depth, color = ns.frames[0]
    H, W = ns.renderer.H, ns.renderer.W
    d, v, col = ns.renderer.render_img(ns.shared_c, ns.shared_decoders, c2w, "cuda:0", "color", gt_depth=depth)
    assert d.shape == (H, W) and d.dtype == torch.float64 and v.dtype == torch.float64 and col.shape == (H, W, 3)
    prm = {k: t.detach().cpu() for k, t in ns.shared_decoders.state_dict().items()}
    cg = {k: t.detach().cpu().contiguous() for k, t in ns.shared_c.items()}
    ro, rd = O.get_rays(H, W, ns.fx, ns.fy, ns.cx, ns.cy, c2w.cpu())
    with torch.no_grad():
        d0, v0, c0 = O.render_batch_ray(prm, cg, rd.reshape(-1, 3), ro.reshape(-1, 3), "color", depth.cpu().reshape(-1),
                                        ns.bound)
    assert_close(d.cpu().reshape(-1), d0, 1e-4, "depth")
    assert_close(v.cpu().reshape(-1), v0, 1e-4, "var")
    assert_close(col.cpu().reshape(-1, 3), c0, 1e-4, "color")


def test_step_from_host_matches_load_and_run(scene):
    """capture_host / step_from_host: the whole host-to-host step as one graph gives the loss and the gradients of
    load() + run(), also after a batch with fewer rays (the tail is masked off by the staged keep bytes)."""
    from nice_slam_annotation_b200.step import FusedStep
    ns = scene
    kw = dict(loss="mapping", w_color=0.2, grid_grads=("grid_middle", "grid_fine", "grid_color"),
              decoder_grads=("color_decoder",))
    a = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", 160, **kw)
    b = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", 160, **kw)
    big, small = _batch(ns, 160, seed=5), _batch(ns, 96, seed=6)
    b.load(*big[:4]); b.run(); b.capture_host()
    for batch in (big, small, big):
        a.load(*batch[:4])
        l0 = float(a.run().item())
        g0 = a.bucket.flat.clone()
        l1 = b.step_from_host(FusedStep.pack_batch(*batch[:4]))
        assert abs(l1 - l0) <= 1e-9 * abs(l0)
        assert_close(b.bucket.flat.cpu(), g0.cpu(), 1e-5, "bucket")


@pytest.mark.parametrize("chunk", [96, 250])
def test_chunked_step_equals_whole_batch(scene, chunk, monkeypatch):
    """A large mapping batch runs as ray chunks (forward + loss + backward per chunk, step.FusedStep.chunk); with the
    batch depth max computed once over the whole batch every value must equal the single-call step: outputs bit for
    bit, loss to rounding, gradients to atomic-order noise.  Ragged last chunk, dropped rays, graph replay."""
    from nice_slam_annotation_b200.step import FusedStep
    ns = scene
    n = 600
    ro, rd, gd, gc, keep = _batch(ns, n, seed=13)
    kw = dict(loss="mapping", w_color=0.2, use_color=True, grid_grads=("grid_middle", "grid_fine", "grid_color"),
              decoder_grads=("color_decoder",), ray_grads=True)
    whole = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", n, **kw)
    assert whole.nchunk == 1
    monkeypatch.setenv("NSB_STEP_CHUNK_RAYS", str(chunk))
    parts = FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, "color", n, **kw)
    assert parts.nchunk == (n + chunk - 1) // chunk and parts.chunk == chunk
    for st in (whole, parts):
        st.load(ro, rd, gd, gc, keep=keep)
        st.run()
    parts.capture()
    parts.run()
    torch.cuda.synchronize()
    for name in ("depth", "var", "color", "g_rays_o", "g_rays_d"):
        assert torch.equal(getattr(parts, name), getattr(whole, name)), name
    assert abs(float(parts.loss) - float(whole.loss)) <= 1e-12 * abs(float(whole.loss))
    assert float(whole.bucket.flat.abs().max()) > 0
    assert_close(parts.bucket.flat, whole.bucket.flat, 1e-5, "gradient bucket")
