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


# ---- the two SLAM steps through the reference-facing classes -------------------------------------
def _slam_ns(gpu, d, cfg_over=None):
    """cfg + the SimpleNamespace NICE_SLAM hands to Renderer / Tracker / Mapper."""
    import types
    from nice_slam_annotation_b200.utils.Renderer import Renderer
    H, W, fx, fy, cx, cy = [float(v) for v in d["cam"]]
    cfg = {
        "scale": 1, "coarse": True, "occupancy": True, "sync_method": "strict",
        "rendering": {"N_samples": 32, "N_surface": 16, "N_importance": 0, "lindisp": False, "perturb": 0.0},
        "model": {"coarse_bound_enlarge": 2},
        "tracking": {"lr": 1e-3, "device": gpu.DEV, "iters": 10, "gt_camera": False, "pixels": 150,
                     "seperate_LR": False, "w_color_loss": 0.5, "ignore_edge_W": 4, "ignore_edge_H": 4,
                     "handle_dynamic": True, "use_color_in_tracking": True, "const_speed_assumption": True},
        "mapping": {"device": gpu.DEV, "every_frame": 5, "no_vis_on_first_frame": True, "fix_fine": True,
                    "BA_cam_lr": 0.001, "fix_color": False, "pixels": 400, "iters": 60, "color_refine": True,
                    "w_color_loss": 0.2, "keyframe_every": 50, "fine_iter_ratio": 0.6, "middle_iter_ratio": 0.4,
                    "mapping_window_size": 5, "frustum_feature_selection": False,
                    "keyframe_selection_method": "overlap", "save_selected_keyframes_info": False,
                    "stage": {"coarse": {"decoders_lr": 0.0, "coarse_lr": 0.001, "middle_lr": 0.0, "fine_lr": 0.0, "color_lr": 0.0},
                              "middle": {"decoders_lr": 0.0, "coarse_lr": 0.0, "middle_lr": 0.1, "fine_lr": 0.0, "color_lr": 0.0},
                              "fine": {"decoders_lr": 0.0, "coarse_lr": 0.0, "middle_lr": 0.005, "fine_lr": 0.005, "color_lr": 0.0},
                              "color": {"decoders_lr": 0.005, "coarse_lr": 0.0, "middle_lr": 0.005, "fine_lr": 0.005, "color_lr": 0.005}}},
    }
    for k, v in (cfg_over or {}).items():
        cfg[k].update(v)
    ns = types.SimpleNamespace(nice=True, bound=torch.as_tensor(d["bound"]), H=int(H), W=int(W), fx=fx, fy=fy,
                               cx=cx, cy=cy)
    ns.renderer = Renderer(cfg, None, ns)
    return cfg, ns


def test_tracker_iteration_vs_reference(gpu, golden):
    """Tracker.optimize_cam_in_batch: loss, camera gradient and Adam step against the
    reference's own step on the same pixel draw (golden 't.*')."""
    from nice_slam_annotation_b200.Tracker import Tracker
    d = golden("trained_nice")
    cfg, ns = _slam_ns(gpu, d)
    ns.shared_decoders = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    ns.shared_c = gpu.grids_from(d)
    tr = Tracker(cfg, None, ns)
    tr.update_para_from_mapping()
    cam = torch.as_tensor(d["t.cam0"]).to(gpu.DEV).requires_grad_(True)
    opt = torch.optim.Adam([cam], lr=1e-3)
    grads = {}
    step0 = opt.step

    def spy(*a, **k):
        grads["g"] = cam.grad.detach().clone()
        return step0(*a, **k)
    opt.step = spy
    depth = torch.as_tensor(d["depth_img"]).to(gpu.DEV)
    color = torch.as_tensor(d["color_img"]).to(gpu.DEV)
    loss = tr.optimize_cam_in_batch(cam, color, depth, 150, opt, indices=torch.as_tensor(d["t.indices"]))
    assert abs(loss - float(d["t.loss"])) <= 1e-4 * abs(float(d["t.loss"])), (loss, float(d["t.loss"]))
    assert_close(grads["g"].cpu(), d["t.grad"], TOL_GRAD, "camera gradient")
    assert_close((cam.detach().cpu() - torch.as_tensor(d["t.cam0"])), d["t.cam1"] - d["t.cam0"], 1e-2, "Adam step")


def test_tracker_api_follows_frame_changes(gpu, golden):
    """The engine behind optimize_cam_in_batch caches the loaded frame: a new frame tensor, and an in-place edit of
    the same tensor, must both be picked up (checked against a fresh Tracker on each frame), and the random pixel
    draw inside the graph must follow torch's CUDA generator like common.py:99."""
    from nice_slam_annotation_b200.Tracker import Tracker
    d = golden("trained_nice")
    cfg, ns = _slam_ns(gpu, d)
    ns.shared_decoders = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    ns.shared_c = gpu.grids_from(d)
    depth = torch.as_tensor(d["depth_img"]).to(gpu.DEV)
    color = torch.as_tensor(d["color_img"]).to(gpu.DEV)
    idx = torch.as_tensor(d["t.indices"])

    def one(tr, col, dep, indices=idx):
        cam = torch.as_tensor(d["t.cam0"]).to(gpu.DEV).requires_grad_(True)
        return tr.optimize_cam_in_batch(cam, col, dep, 150, torch.optim.SGD([cam], lr=0.0), indices=indices)

    def fresh():
        tr = Tracker(cfg, None, ns)
        tr.update_para_from_mapping()
        return tr
    tr = fresh()
    l0 = one(tr, color, depth)
    depth2, color2 = (depth * 1.07).contiguous(), (color * 0.9).contiguous()
    l1 = one(tr, color2, depth2)                         # another tensor
    depth2.mul_(0.95)
    l2 = one(tr, color2, depth2)                         # the same tensor, edited in place
    assert l1 != l0 and l2 != l1
    assert l1 == one(fresh(), color2, depth2 / 0.95) or abs(l1 - one(fresh(), color2, depth2 / 0.95)) <= 1e-5 * abs(l1)
    ref2 = one(fresh(), color2, depth2)
    assert abs(l2 - ref2) <= 1e-6 * abs(ref2), (l2, ref2)
    assert abs(one(tr, color, depth) - l0) <= 1e-6 * abs(l0)
    # random draw: same generator state -> same loss; advancing state -> another draw
    torch.manual_seed(11)
    a = one(tr, color, depth, indices=None)
    b = one(tr, color, depth, indices=None)
    torch.manual_seed(11)
    a2 = one(tr, color, depth, indices=None)
    assert a == a2 and a != b, (a, a2, b)


def test_fused_tracking_loop_vs_api_loop(gpu, golden):
    """Tracker.optimize_cam_fused (pixel -> rays -> render -> loss -> pose gradient -> Adam ->
    candidate rule as one replayed CUDA graph) follows the same trajectory as the reference-shaped
    loop of optimize_cam_in_batch + torch.optim.Adam + loss.item() bookkeeping (Tracker.py:244-275)
    on the same pixel draw; the first iteration is the one the reference golden pins."""
    from nice_slam_annotation_b200.Tracker import Tracker
    from nice_slam_annotation_b200.step import TrackingStep
    d = golden("trained_nice")
    for separate in (False, True):
        cfg, ns = _slam_ns(gpu, d, {"tracking": {"seperate_LR": separate}})
        ns.shared_decoders = gpu.nice_from_state(sub(d, "dec."), d["bound"])
        ns.shared_c = gpu.grids_from(d)
        tr = Tracker(cfg, None, ns)
        tr.update_para_from_mapping()
        depth = torch.as_tensor(d["depth_img"]).to(gpu.DEV)
        color = torch.as_tensor(d["color_img"]).to(gpu.DEV)
        idx = torch.as_tensor(d["t.indices"]).to(gpu.DEV)
        cam0 = torch.as_tensor(d["t.cam0"]).to(gpu.DEV)
        iters = 6
        # reference-shaped loop (Tracker.py:226-275)
        if separate:
            T = cam0[-3:].clone().requires_grad_(True)
            quad = cam0[:4].clone().requires_grad_(True)
            opt = torch.optim.Adam([{"params": [T], "lr": 1e-3}, {"params": [quad], "lr": 1e-3 * 0.2}])
            cam_of = lambda: torch.cat([quad, T], 0)
        else:
            cam = cam0.clone().requires_grad_(True)
            opt = torch.optim.Adam([cam], lr=1e-3)
            cam_of = lambda: cam
        best, cand, losses = 1e10, None, []
        for _ in range(iters):
            loss = tr.optimize_cam_in_batch(cam_of(), color, depth, 150, opt, indices=idx)
            losses.append(loss)
            if loss < best:
                best, cand = loss, cam_of().detach().clone()
        # fused loop on the same pixels
        ts = TrackingStep(tr, 150)
        ts.draw = False
        ts.indices.copy_(idx)
        ts.load_frame(color, depth)
        for graph in (False, True):
            ts.reset(cam0)
            if graph:
                ts.capture()
            ts.iterate(iters)
            bc, bl = ts.result()
            assert abs(bl - best) <= 1e-4 * abs(best), (separate, graph, bl, best)
            assert_close((ts.cam - cam0).cpu(), (cam_of().detach() - cam0).cpu(), 1e-2, "camera after the loop")
            assert_close((bc - cam0).cpu(), (cand - cam0).cpu(), 1e-2, "candidate camera")
        if not separate:
            assert abs(losses[0] - float(d["t.loss"])) <= 1e-4 * abs(float(d["t.loss"]))
    # the public entry point: random pixel draw inside the graph, loss must not increase much
    cam_c, loss_c = tr.optimize_cam_fused(cam0, color, depth, 150, num_iters=10)
    assert cam_c.shape == (7,) and loss_c > 0 and loss_c < 2 * best


@pytest.mark.parametrize("fused_loop", [True, False])
def test_mapper_iterations_vs_reference(gpu, golden, fused_loop):
    """Three Mapper.optimize_map iterations (middle, middle, color schedule of a 3-iteration run)
    from the trained state reproduce the reference's updated grids and decoders (golden 'm.*'),
    through the device-resident loop (map_loop.MappingLoop) and through the autograd path."""
    from nice_slam_annotation_b200.Mapper import Mapper
    d = golden("trained_nice")
    cfg, ns = _slam_ns(gpu, d)
    ns.shared_decoders = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    ns.shared_c = gpu.grids_from(d)
    mp = Mapper(cfg, None, ns)
    mp.fused_loop = fused_loop
    mp.index_source = iter([torch.as_tensor(r) for r in d["m.indices"]])
    depth = torch.as_tensor(d["depth_img"]).to(gpu.DEV)
    color = torch.as_tensor(d["color_img"]).to(gpu.DEV)
    c2w = torch.as_tensor(d["c2w"]).to(gpu.DEV)
    before = {k: v.detach().clone() for k, v in ns.shared_c.items()}
    mp.optimize_map(3, 1.0, 5, color, depth, c2w, [], [], c2w)
    def mostly_close(actual, ref, what, tol=2e-3, frac=0.999):
        # Adam's first steps are lr * sign(g): an element whose gradient is rounding noise may step
        # the other way, so require all but a sliver of the elements to agree
        ref = torch.as_tensor(ref).double()
        err = (actual.double() - ref).abs()
        ok = (err <= tol * ref.abs().max()).double().mean().item()
        assert ok >= frac, f"{what}: only {ok:.5f} of the elements within {tol:g}"

    for k, ref in sub(d, "m.c_after.").items():
        delta_ref = ref - torch.as_tensor(d["c." + k])
        delta = (mp.c[k].detach() - before[k]).cpu()
        if float(delta_ref.abs().max()) == 0.0:
            assert float(delta.abs().max()) == 0.0, k
        else:
            assert torch.equal(delta_ref != 0, delta != 0) or \
                ((delta_ref != 0) != (delta != 0)).double().mean().item() < 1e-3, "touched voxels differ: " + k
            mostly_close(mp.c[k].detach().cpu(), ref, "grid after 3 iterations: " + k)
    for k, ref in sub(d, "m.dec_after.").items():
        mostly_close(mp.decoders.state_dict()[k].cpu(), ref, "decoder after 3 iterations: " + k, frac=0.99)


@pytest.mark.parametrize("frustum,ba", [(True, False), (False, True), (True, True)])
def test_fused_mapping_loop_vs_autograd_loop(gpu, golden, frustum, ba):
    """Mapper.optimize_map with frustum feature selection (masked in-place Adam instead of the
    compact val[mask] copies, Mapper.py:393-404,633-642) and local BA (pose gradient + camera
    Adam, Mapper.py:430-458,644-667): the device-resident loop follows the autograd path (which
    the reference golden pins) on the same keyframes and pixel draws."""
    import numpy as np
    from nice_slam_annotation_b200.Mapper import Mapper
    d = golden("trained_nice")
    depth = torch.as_tensor(d["depth_img"]).to(gpu.DEV)
    color = torch.as_tensor(d["color_img"]).to(gpu.DEV)
    c2w = torch.as_tensor(d["c2w"]).to(gpu.DEV).float()
    if c2w.shape[0] == 3:
        c2w = torch.cat([c2w, torch.tensor([[0, 0, 0, 1.]], device=gpu.DEV)], 0)
    iters, pix = 5, 400
    g = torch.Generator().manual_seed(5)
    H, W = depth.shape
    draws = [torch.randint(H * W, (pix // 3,), generator=g) for _ in range(3 * iters)]

    def run(fused_loop):
        cfg, ns = _slam_ns(gpu, d, {"mapping": {"frustum_feature_selection": frustum, "keyframe_selection_method": "global"}})
        ns.shared_decoders = gpu.nice_from_state(sub(d, "dec."), d["bound"])
        ns.shared_c = gpu.grids_from(d)
        mp = Mapper(cfg, None, ns)
        mp.fused_loop, mp.BA = fused_loop, ba
        kf = []
        for i in range(2):
            pose = c2w.clone()
            pose[:3, 3] += 0.02 * (i + 1)
            kf.append({"gt_c2w": pose.clone(), "idx": i, "color": color, "depth": depth, "est_c2w": pose.clone()})
        mp.keyframe_dict = kf
        mp.index_source = iter([x.to(gpu.DEV) for x in draws])
        np.random.seed(0)
        out = mp.optimize_map(iters, 1.0, 7, color, depth, c2w, kf, [0, 1], c2w.clone())
        return mp, kf, out

    ref_mp, ref_kf, ref_out = run(False)
    mp, kf, out = run(True)

    def mostly_close(a, b, what, tol=2e-3, frac=0.999):
        err = (a.double() - b.double()).abs()
        ok = (err <= tol * b.abs().max().clamp_min(1e-30)).double().mean().item()
        assert ok >= frac, f"{what}: only {ok:.5f} of the elements within {tol:g}"

    for k in ("grid_middle", "grid_fine", "grid_color"):
        moved_ref = (ref_mp.c[k].detach() != torch.as_tensor(d["c." + k]).to(gpu.DEV))
        moved = (mp.c[k].detach() != torch.as_tensor(d["c." + k]).to(gpu.DEV))
        assert moved_ref.any() and (moved_ref != moved).double().mean().item() < 1e-3, "touched voxels differ: " + k
        mostly_close(mp.c[k].detach(), ref_mp.c[k].detach(), "grid " + k)
    sd, sd_ref = mp.decoders.state_dict(), ref_mp.decoders.state_dict()
    for k in sd:
        mostly_close(sd[k], sd_ref[k], "decoder " + k, frac=0.99)
    if ba:
        assert_close(out.cpu() - c2w.cpu(), ref_out.cpu() - c2w.cpu(), 5e-2, "BA: current pose update")
        assert float((out - c2w).abs().max()) > 0
        for a, b in zip(kf, ref_kf):
            assert torch.allclose(a["est_c2w"], b["est_c2w"], atol=1e-4), "BA: keyframe pose"
    else:
        assert out is None and ref_out is None


@pytest.mark.parametrize("fused_loop", [True, False])
@pytest.mark.parametrize("name,frustum,ba", [("frustum", True, False), ("ba", False, True), ("frustum_ba", True, True)])
def test_mapping_loop_vs_reference_frustum_and_ba(gpu, golden, name, frustum, ba, fused_loop):
    """f1 pinned to the reference itself: Mapper.optimize_map with frustum feature selection (the reference's compact
    val[mask] Adam + write-back, Mapper.py:393-404,491-500,633-642) and with local BA (Mapper.py:430-458,644-667),
    10 iterations from the trained state on the reference's recorded pixel draws, against the grids / colour decoder
    / poses the reference's own Mapper produced (tests/golden/mapper_variants.npz)."""
    import numpy as np
    from nice_slam_annotation_b200.Mapper import Mapper
    d, v = golden("trained_nice"), golden("mapper_variants")
    iters, pix = int(v["iters"]), int(v["pixels"])
    depth = torch.as_tensor(d["depth_img"]).to(gpu.DEV)
    color = torch.as_tensor(d["color_img"]).to(gpu.DEV)
    c2w = torch.as_tensor(d["c2w"]).to(gpu.DEV).float()
    if c2w.shape[0] == 3:
        c2w = torch.cat([c2w, torch.tensor([[0, 0, 0, 1.]], device=gpu.DEV)], 0)
    cfg, ns = _slam_ns(gpu, d, {"mapping": {"frustum_feature_selection": frustum, "keyframe_selection_method": "global",
                                            "pixels": pix}})
    ns.shared_decoders = gpu.nice_from_state(sub(d, "dec."), d["bound"])
    ns.shared_c = gpu.grids_from(d)
    mp = Mapper(cfg, None, ns)
    mp.fused_loop, mp.BA = fused_loop, ba
    kf = []
    for i in range(2):
        pose = c2w.clone()
        pose[:3, 3] += 0.02 * (i + 1)
        kf.append({"gt_c2w": pose.clone(), "idx": i, "color": color, "depth": depth, "est_c2w": pose.clone()})
    mp.keyframe_dict = kf
    mp.index_source = iter([torch.as_tensor(x).to(gpu.DEV) for x in v[name + ".indices"]])
    np.random.seed(0)
    out = mp.optimize_map(iters, 1.0, 7, color, depth, c2w, kf, [0, 1], c2w.clone())

    def mostly_close(a, b, what, tol=2e-3, frac=0.999):
        b = torch.as_tensor(b).double()
        err = (a.double().cpu() - b).abs()
        ok = (err <= tol * b.abs().max().clamp_min(1e-30)).double().mean().item()
        assert ok >= frac, f"{what}: only {ok:.5f} of the elements within {tol:g}"
    for k in ("grid_middle", "grid_fine", "grid_color"):
        ref = torch.as_tensor(v[f"{name}.c_after.{k}"])
        start = torch.as_tensor(d["c." + k])
        moved_ref, moved = ref != start, mp.c[k].detach().cpu() != start
        assert moved_ref.any()
        assert (moved_ref != moved).double().mean().item() < 1e-3, "optimised voxel sets differ: " + k
        mostly_close(mp.c[k].detach(), ref, "grid " + k)
    sd = mp.decoders.state_dict()
    for k, ref in sub(v, name + ".dec_after.").items():
        mostly_close(sd[k], ref, "decoder " + k, frac=0.99)
    kf_ref, kf0 = torch.as_tensor(v[name + ".kf_c2w"]), torch.as_tensor(v[name + ".kf_c2w0"])
    if ba:
        ref_cur = torch.as_tensor(v[name + ".cur_c2w"])
        assert float((ref_cur - c2w.cpu()).abs().max()) > 0
        assert_close(out.cpu() - c2w.cpu(), ref_cur - c2w.cpu(), 5e-2, "BA: current pose update")
        for i, k_ in enumerate(kf):
            assert torch.allclose(k_["est_c2w"].cpu(), kf_ref[i], atol=2e-4), "BA: keyframe pose"
        assert float((kf_ref[1] - kf0[1]).abs().max()) > 0      # the reference did move the newer keyframe
    else:
        assert out is None
        for i, k_ in enumerate(kf):
            assert torch.equal(k_["est_c2w"].cpu(), kf0[i])
