"""The drop-in seam of INTEGRATION.md section 1, run for real: the reference's OWN source of
``Tracker.optimize_cam_in_batch`` (src/Tracker.py:78-135) and ``Mapper.optimize_map`` (src/Mapper.py:295-669),
unmodified from ``baseline/_ref`` (baseline/make_ref.py; the copy travels to the GPU box), executing on top of THIS
package's ``Renderer`` and ``NICE`` decoders -- and the results compared with the golden vectors the all-reference
run produced (tests/golden/trained_nice.npz ``t.*`` / ``m.*``).  The grids are handed over the way the reference
allocates them (plain contiguous [1,32,Z,Y,X], ``NICE_SLAM.grid_init``), i.e. through the layout conversion of
``fused.grid_to_kernel_layout``.

The only thing patched is the random pixel draw (``torch.randint`` inside ``src.common.get_samples``): the golden
run drew on the CPU generator, this run is on the GPU, so the recorded indices are replayed.
"""
import contextlib

import pytest
import torch

from conftest import assert_close, sub

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ref():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from baseline import ref_harness as RH
    if not RH.available():
        pytest.skip("baseline/_ref is missing (python baseline/make_ref.py in the build container)")
    RH.load()
    return RH


@contextlib.contextmanager
def replay_randint(draws):
    """torch.randint(high, (n,), device=...) -> the next recorded draw."""
    it = iter(draws)
    real = torch.randint

    def fake(*a, **k):
        idx = torch.as_tensor(next(it))
        size = a[1] if len(a) > 1 else k.get("size")
        assert tuple(size) == tuple(idx.shape), (size, idx.shape)
        return idx.to(k.get("device", "cpu"))
    torch.randint = fake
    try:
        yield
    finally:
        torch.randint = real


def _seam_scene(ref, golden):
    """Product Renderer + product NICE decoders + reference-layout grids inside the namespace the reference's
    Tracker / Mapper read."""
    import test_gpu_parity as T
    import gpu_util
    d = golden("trained_nice")
    cfg, ns = T._slam_ns(gpu_util, d)
    cfg["data"] = {"dim": 3}
    ns.cfg = cfg
    ns.shared_decoders = gpu_util.nice_from_state(sub(d, "dec."), d["bound"])
    ns.shared_c = {k: v.to(DEV).contiguous() for k, v in sub(d, "c.").items()}      # channels-FIRST, as grid_init makes them
    ns.bound = torch.as_tensor(d["bound"])
    return d, cfg, ns


def test_reference_tracker_source_over_product_renderer(ref, golden):
    d, cfg, ns = _seam_scene(ref, golden)
    slam = ref.reference_scene(ns, DEV, renderer=ns.renderer, decoders=ns.shared_decoders)
    assert type(slam.renderer).__module__.startswith("nice_slam_annotation_b200")
    tr = ref.make_tracker(slam, cfg, DEV)
    assert type(tr).__module__ == "src.Tracker"
    cam = torch.as_tensor(d["t.cam0"]).to(DEV).requires_grad_(True)
    opt = torch.optim.Adam([cam], lr=1e-3)
    grads, step0 = {}, opt.step

    def spy(*a, **k):
        grads["g"] = cam.grad.detach().clone()
        return step0(*a, **k)
    opt.step = spy
    depth = torch.as_tensor(d["depth_img"]).to(DEV)
    color = torch.as_tensor(d["color_img"]).to(DEV)
    with replay_randint([d["t.indices"]]):
        loss = tr.optimize_cam_in_batch(cam, color, depth, 150, opt)
    assert abs(loss - float(d["t.loss"])) <= 1e-4 * abs(float(d["t.loss"])), (loss, float(d["t.loss"]))
    assert_close(grads["g"].cpu(), d["t.grad"], 1e-3, "camera gradient")
    assert_close((cam.detach().cpu() - torch.as_tensor(d["t.cam0"])), d["t.cam1"] - d["t.cam0"], 1e-2, "Adam step")
    # the tracker's grids do not change between mapper updates: the layout conversion is cached per grid
    from nice_slam_annotation_b200 import fused
    before = dict(fused._layout_cache)
    with replay_randint([d["t.indices"]]):
        tr.optimize_cam_in_batch(cam, color, depth, 150, opt)
    assert dict(fused._layout_cache).keys() == before.keys()


def test_reference_mapper_source_over_product_renderer(ref, golden):
    d, cfg, ns = _seam_scene(ref, golden)
    slam = ref.reference_scene(ns, DEV, renderer=ns.renderer, decoders=ns.shared_decoders)
    mp = ref.make_mapper(slam, cfg, DEV, frustum=False)
    assert type(mp).__module__ == "src.Mapper"
    depth = torch.as_tensor(d["depth_img"]).to(DEV)
    color = torch.as_tensor(d["color_img"]).to(DEV)
    c2w = torch.as_tensor(d["c2w"]).to(DEV)
    before = {k: v.detach().clone() for k, v in slam.shared_c.items()}
    with replay_randint(list(d["m.indices"])):
        mp.optimize_map(3, 1.0, 5, color, depth, c2w, [], [], c2w)

    def mostly_close(actual, ref_, what, tol=2e-3, frac=0.999):
        ref_ = torch.as_tensor(ref_).double()
        err = (actual.double() - ref_).abs()
        ok = (err <= tol * ref_.abs().max()).double().mean().item()
        assert ok >= frac, f"{what}: only {ok:.5f} of the elements within {tol:g}"
    for k, r in sub(d, "m.c_after.").items():
        delta_ref = r - torch.as_tensor(d["c." + k])
        delta = (mp.c[k].detach() - before[k]).cpu()
        if float(delta_ref.abs().max()) == 0.0:
            assert float(delta.abs().max()) == 0.0, k
        else:
            assert ((delta_ref != 0) != (delta != 0)).double().mean().item() < 1e-3, "touched voxels differ: " + k
            mostly_close(mp.c[k].detach().cpu(), r, "grid after 3 iterations: " + k)
    for k, r in sub(d, "m.dec_after.").items():
        mostly_close(mp.decoders.state_dict()[k].cpu(), r, "decoder after 3 iterations: " + k, frac=0.99)
