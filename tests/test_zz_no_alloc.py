"""nsb_composite_mapping_static: the allocation-free variant of nsb_composite_mapping (compositing + mapping loss +
adjoint, src/common.py:215-256 and src/Mapper.py:604-612 in one kernel) whose scratch is a file-scope static device
array of the library.  The entry point has no counterpart in the reference; it was added
for callers that cannot call cudaMalloc.  Checked here through the C ABI against the existing entry point
(caller-provided scratch) on the same inputs, on two calls in a row of different size (the shared array keeps the
first call's contents), plus the size limit and the no-allocation promise as far as torch can see it.

This file is named so that pytest collects it LAST: a failure here must not
hide any other test.  It passes alone in a fresh process and after the other tests (it shares no state with them)."""
import ctypes as C

import pytest
import torch

from conftest import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL_OUT, TOL_GRAD = 1e-4, 1e-3          # the tolerances of the parity tests (north_star): outputs / gradients
STAGE_COLOR, S = 3, 48


def _inputs(n, seed):
    g = torch.Generator().manual_seed(seed)
    ro = (torch.rand(n, 3, generator=g) - 0.5).to(DEV)
    rd = torch.randn(n, 3, generator=g).to(DEV)
    z = torch.sort(torch.rand(n, S, generator=g, dtype=torch.float64) * 4.0 + 0.05, -1)[0].to(DEV)
    P = n * S
    occ0 = (torch.randn(P, generator=g) * 0.3).to(DEV)
    occ1 = (torch.randn(P, generator=g) * 0.3).to(DEV)
    rgb = torch.rand(P, 4, generator=g).to(DEV)
    gd = (torch.rand(n, generator=g) * 4.0).to(DEV)
    gd[::7] = 0.0                                          # zero sensor depth: no depth term for these rays
    gc = torch.rand(n, 3, generator=g).to(DEV)
    keep = (torch.rand(n, generator=g) > 0.1).to(torch.uint8).to(DEV)
    return ro, rd, z, occ0, occ1, rgb, gd, gc, keep


def _outputs(n):
    return (torch.full((n,), float("nan"), dtype=torch.float64, device=DEV),
            torch.full((n,), float("nan"), dtype=torch.float64, device=DEV),
            torch.full((n, 3), float("nan"), device=DEV), torch.full((n * S, 4), float("nan"), device=DEV),
            torch.full((1,), float("nan"), dtype=torch.float64, device=DEV))


def _run(static, inp, out, bound6):
    from nice_slam_annotation_b200 import _lib, fused
    L = _lib.lib()
    ro, rd, z, occ0, occ1, rgb, gd, gc, keep = inp
    depth, var, col, d_raw, loss = out
    pts = fused.ray_points(ro, rd, z)
    p, st = fused._ptr, fused._stream()
    head = (C.byref(pts), bound6, STAGE_COLOR, 1, p(occ0), p(occ1), p(rgb), p(gd), p(gc), p(keep), 0.2, 1,
            p(depth), p(var), p(col), p(d_raw))
    if static:
        rc = L.nsb_composite_mapping_static(*head, p(loss), st)
    else:
        n = ro.shape[0]
        partials = torch.empty(int(L.nsb_composite_mapping_partials(n)), dtype=torch.float64, device=DEV)
        ticket = torch.zeros(1, dtype=torch.int32, device=DEV)
        rc = L.nsb_composite_mapping(*head, p(partials), p(ticket), p(loss), st)
    _lib.check(rc, "composite_mapping" + ("_static" if static else ""))
    torch.cuda.synchronize()


def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from nice_slam_annotation_b200 import fused
    bound6 = fused._bound6([[-3.0, 3.0], [-3.0, 3.0], [-3.0, 3.0]])
    # second call smaller than the first: the array still holds the first call's partials beyond its own blocks
    for n, seed in ((1000, 1), (301, 2)):
        inp = _inputs(n, seed)
        ref, got = _outputs(n), _outputs(n)
        _run(False, inp, ref, bound6)
        before = torch.cuda.memory_allocated()
        _run(True, inp, got, bound6)
        assert torch.cuda.memory_allocated() == before      # nothing allocated through torch by the call
        for a, b, tol, what in zip(got, ref, (TOL_OUT, TOL_OUT, TOL_OUT, TOL_GRAD, TOL_OUT),
                                   ("depth", "var", "color", "d_raw", "loss")):
            assert torch.isfinite(b).all(), what
            assert float(b.abs().max()) > 0, what
            assert_close(a, b, tol, f"{what} (n = {n})")


def test_static_scratch_entry_point_rejects_batches_beyond_its_limit():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from nice_slam_annotation_b200 import _lib, fused
    L = _lib.lib()
    inp = _inputs(8, 3)
    ro, rd, z, occ0, occ1, rgb, gd, gc, keep = inp
    depth, var, col, d_raw, loss = _outputs(8)
    pts = fused.ray_points(ro, rd, z)
    pts.N = 65536 + 1                      # NSB_STATIC_MAX_RAYS + 1: rejected before anything is launched or read
    p = fused._ptr
    rc = L.nsb_composite_mapping_static(C.byref(pts), fused._bound6([[-3.0, 3.0]] * 3), STAGE_COLOR, 1, p(occ0), p(occ1),
                                        p(rgb), p(gd), p(gc), p(keep), 0.2, 1, p(depth), p(var), p(col), p(d_raw),
                                        p(loss), fused._stream())
    assert rc == _lib.NSB_EINVAL
    msg = L.nsb_last_error().decode()
    assert "65536" in msg and "limit" in msg, msg
    torch.cuda.synchronize()
    assert torch.isnan(loss).all()          # nothing ran
