"""Own memory-safety and race checks for the step kernels.  compute-sanitizer (memcheck / racecheck) is closed on the
GPU pool this repo is measured on (profiles/r02p_sanitizer_closed.txt), so the two properties it would have checked
are tested with the shipping library instead:

* out-of-bounds writes: every device buffer the engine hands to the C ABI (workspace, weight-gradient scratch,
  outputs, ray gradients, packed decoders, the gradient bucket and its line marks) is allocated between two 4 KiB
  guard bands filled with a pattern; after forward + backward steps at ragged sizes the bands must be untouched;
* shared-memory / TMEM / named-barrier races: the step is replayed 30 times on identical inputs; everything that is
  not an atomic sum (depth, variance, colour, the loss, the ray gradients) must be bit-identical run to run, and
  the atomically accumulated gradients must agree to 5e-5 (reordering noise is ~4e-7) -- a race on an operand image or
  an accumulator shows up as run-to-run differences far above that.
"""
import contextlib

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GUARD = 4096
PATTERN = 0xA5


class _Guards:
    def __init__(self):
        self.bufs = []

    def wrap(self, fn, zero):
        def alloc(*a, **k):
            dev = k.get("device")
            if dev is None or torch.device(dev).type != "cuda":
                return fn(*a, **k)
            proto = fn(*a, **{**k, "device": "meta"})
            nbytes = proto.numel() * proto.element_size()
            pad = (-nbytes) % 256
            raw = torch.full((GUARD + nbytes + pad + GUARD,), PATTERN, dtype=torch.uint8, device=dev)
            mid = raw[GUARD:GUARD + nbytes]
            if zero:
                mid.zero_()
            self.bufs.append((raw, nbytes))
            return mid.view(proto.dtype).reshape(proto.shape)
        return alloc

    def check(self):
        bad = []
        for i, (raw, nbytes) in enumerate(self.bufs):
            lo, hi = raw[:GUARD], raw[GUARD + nbytes:]
            if not (bool((lo == PATTERN).all()) and bool((hi == PATTERN).all())):
                bad.append((i, nbytes, int((lo != PATTERN).sum()), int((hi != PATTERN).sum())))
        return bad


@contextlib.contextmanager
def guarded_allocations():
    g = _Guards()
    empty, zeros = torch.empty, torch.zeros
    torch.empty, torch.zeros = g.wrap(empty, False), g.wrap(zeros, True)
    try:
        yield g
    finally:
        torch.empty, torch.zeros = empty, zeros


@pytest.fixture(scope="module")
def scene():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from nice_slam_annotation_b200.scene import build_scene
    ns = build_scene("replica_room0", DEV, seed=0, n_frames=2)
    g = torch.Generator().manual_seed(5)
    with torch.no_grad():
        for v in ns.shared_c.values():
            v.add_((torch.randn(v.shape, generator=g) * 0.3).to(DEV))
    return ns


def _batch(ns, n, seed):
    from nice_slam_annotation_b200.common import get_samples
    from oracle import nice_oracle as O
    gen = torch.Generator().manual_seed(seed)
    c2w, depth, color = ns.frames[seed % len(ns.frames)]
    idx = torch.randint(ns.H * ns.W, (n,), generator=gen)
    ro, rd, gd, gc = get_samples(0, ns.H, 0, ns.W, n, ns.H, ns.W, ns.fx, ns.fy, ns.cx, ns.cy, c2w, depth, color, DEV,
                                 indices=idx.to(DEV))
    keep = O.prefilter_mask(ro, rd, gd, ns.bound.to(DEV)).to(torch.uint8)
    return ro.contiguous(), rd.contiguous(), gd.contiguous(), gc.contiguous(), keep


def _make_step(ns, stage, n, ray_grads, dec_grads):
    from nice_slam_annotation_b200.step import FusedStep
    grids = {"middle": ("grid_middle",), "fine": ("grid_middle", "grid_fine"),
             "color": ("grid_middle", "grid_fine", "grid_color")}[stage]
    return FusedStep(ns.renderer, ns.shared_decoders, ns.shared_c, stage, n, loss="mapping", w_color=0.2,
                     use_color=stage == "color", grid_grads=grids, decoder_grads=dec_grads, ray_grads=ray_grads)


@pytest.mark.parametrize("stage,n,ray_grads,dec_grads", [
    ("color", 1000, False, ("color_decoder",)),          # the bench step (tcgen05 forward + backward, k_dweight)
    ("color", 777, True, ("color_decoder",)),            # ragged last tile, ray gradients
    ("color", 1, True, ()),                              # one ray: one partly filled tile
    ("color", 129, True, ("middle_decoder", "fine_decoder", "color_decoder")),
    ("fine", 333, False, ()),
    ("middle", 61, True, ()),
])
def test_guard_bands_stay_intact(scene, stage, n, ray_grads, dec_grads):
    ns = scene
    with guarded_allocations() as g:
        step = _make_step(ns, stage, n, ray_grads, dec_grads)
    assert len(g.bufs) >= 12
    for seed in (1, 2):
        step.load(*_batch(ns, n, seed)[:4], keep=_batch(ns, n, seed)[4])
        step.run()
    step.capture()
    step.run()
    torch.cuda.synchronize()
    assert torch.isfinite(step.loss).all()
    assert g.check() == []


def test_step_is_deterministic_run_to_run(scene):
    ns = scene
    n = 1000
    step = _make_step(ns, "color", n, True, ("color_decoder",))
    ro, rd, gd, gc, keep = _batch(ns, n, 7)
    step.load(ro, rd, gd, gc, keep=keep)
    step.run()
    step.capture()
    ref = None
    worst = 0.0
    for _ in range(30):
        step.run()
        torch.cuda.synchronize()
        cur = [t.clone() for t in (step.depth, step.var, step.color, step.loss, step.g_rays_o, step.g_rays_d)]
        flat = step.bucket.flat.clone()
        if ref is None:
            ref, flat0 = cur, flat
            assert float(flat0.abs().max()) > 0
            continue
        for name, a, b in zip(("depth", "var", "color", "loss", "g_rays_o", "g_rays_d"), cur, ref):
            if name == "loss":                                    # a sum over rays by atomics on a double
                assert abs(float(a) - float(b)) <= 1e-9 * abs(float(b)), name
            else:
                assert torch.equal(a, b), name
        worst = max(worst, float((flat - flat0).abs().max() / flat0.abs().max()))
    # fp32 atomics reorder the sums: 1e-7 .. 4e-7 in 5500 replays (scripts/det_probe*.py); one replay in ~4000 showed
    # 7.5e-6 that did not reproduce.  A race on an operand image or an accumulator is orders above either
    # (a lost 16-point record alone is 5e-4), so the bound sits between: 5e-5.
    assert worst <= 5e-5, worst


def test_guard_mechanism_catches_an_overrun():
    """Negative control: one byte written past a guarded buffer is reported."""
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    with guarded_allocations() as g:
        t = torch.empty(100, dtype=torch.float32, device=DEV)
        z = torch.zeros(7, 3, device=DEV)
    assert float(z.abs().sum()) == 0 and g.check() == []
    raw, nbytes = g.bufs[0]
    assert t.data_ptr() == raw.data_ptr() + GUARD and nbytes == 400
    raw[GUARD + nbytes + 256 + 3] = 0                        # past the 256-byte rounding, inside the upper band
    assert [b[0] for b in g.check()] == [0]
