"""f2: frustum feature selection (Mapper.get_mask_from_c2w, src/Mapper.py:100-187) and
the keyframe overlap scores (Mapper.keyframe_selection_overlap, :189-293).

CPU: the oracle restatement (oracle/frustum_oracle.py, cv2-free) against vectors produced by the
reference itself (tests/golden/frustum.npz, cv2.remap inside).  GPU: the kernels (nsb_frustum_mask,
nsb_keyframe_overlap) through the C ABI against the golden vectors and, at Replica scale, the oracle.
Masks and counts are integer work: the bar is bit-exact."""
import os

import numpy as np
import pytest
import torch

from oracle import frustum_oracle as FO

GOLD = os.path.join(os.path.dirname(__file__), "golden", "frustum.npz")
KEYS = ("grid_coarse", "grid_middle", "grid_fine")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _unpack(g, i, key):
    shape = tuple(int(v) for v in g[f"f{i}.{key}.shape"])
    n = shape[0] * shape[1] * shape[2]
    return np.unpackbits(g[f"f{i}.{key}.mask"])[:n].astype(bool).reshape(shape[::-1]), shape


def test_remap_emulation_matches_cv2_when_available():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    img = rng.random((37, 53), dtype=np.float32) * 5
    img[rng.random((37, 53)) < 0.1] = 0
    mx = (rng.random(20000, dtype=np.float32) * 70 - 8).astype(np.float32)
    my = (rng.random(20000, dtype=np.float32) * 50 - 6).astype(np.float32)
    mx[:5] = [np.nan, np.inf, -np.inf, 3e9, -3e9]
    ref = cv2.remap(img, mx, my, interpolation=cv2.INTER_LINEAR)[:, 0]
    got = FO.remap_bilinear(img, mx, my)
    assert np.array_equal(ref, got)


@pytest.mark.parametrize("i", [0, 2, 5])
def test_oracle_frustum_mask_vs_reference(gold, i):
    H, W, fx, fy, cx, cy = gold["cam"]
    for key in KEYS:
        ref, shape = _unpack(gold, i, key)
        got = FO.frustum_mask(gold[f"f{i}.c2w"], key, shape, gold[f"f{i}.depth"], gold["bound"], int(H), int(W),
                              fx, fy, cx, cy)
        assert got.shape == ref.shape
        assert np.array_equal(got, ref), (key, int((got != ref).sum()))


def test_oracle_keyframe_scores_vs_reference(gold):
    from oracle import nice_oracle as O
    H, W, fx, fy, cx, cy = gold["cam"]
    H, W = int(H), int(W)
    ro, rd, gd, _ = O.get_samples(0, H, 0, W, 100, fx, fy, cx, cy, torch.from_numpy(gold["k.c2w"]),
                                  torch.from_numpy(gold["k.depth"]), torch.from_numpy(gold["k.color"]),
                                  indices=torch.from_numpy(gold["k.indices"]))
    pts = FO.keyframe_overlap_points(ro.numpy(), rd.numpy(), gd.numpy(), 16)
    got = FO.keyframe_overlap(pts, gold["k.kf_c2w"], H, W, fx, fy, cx, cy)
    assert np.array_equal(got, gold["k.scores"]), (got, gold["k.scores"])


# ---- GPU: the kernels through the C ABI -----------------------------------------------------------
def _mapper(gold_or_none, dev, H, W, fx, fy, cx, cy, bound):
    from nice_slam_annotation_b200.Mapper import Mapper
    m = object.__new__(Mapper)
    m.device, m.H, m.W, m.fx, m.fy, m.cx, m.cy = dev, int(H), int(W), float(fx), float(fy), float(cx), float(cy)
    m.bound = torch.as_tensor(bound)
    return m


@pytest.mark.gpu
@pytest.mark.parametrize("i", [0, 2, 5])
def test_kernel_frustum_mask_vs_reference(gold, i):
    m = _mapper(gold, "cuda:0", *gold["cam"], gold["bound"])
    for key in KEYS:
        ref, shape = _unpack(gold, i, key)
        got = m.get_mask_from_c2w(torch.from_numpy(gold[f"f{i}.c2w"]), key, shape, gold[f"f{i}.depth"])
        assert got.dtype == np.bool_ and got.shape == ref.shape
        assert np.array_equal(got, ref), (key, int((got != ref).sum()))
        zyx = m._frustum_mask(torch.from_numpy(gold[f"f{i}.c2w"]), key, shape, gold[f"f{i}.depth"]).cpu().numpy()
        assert np.array_equal(zyx, ref.transpose(2, 1, 0))


@pytest.mark.gpu
def test_kernel_keyframe_scores_vs_reference(gold):
    from nice_slam_annotation_b200.common import get_samples
    H, W, fx, fy, cx, cy = gold["cam"]
    H, W = int(H), int(W)
    dev = "cuda:0"
    m = _mapper(gold, dev, *gold["cam"], gold["bound"])
    ro, rd, gd, _ = get_samples(0, H, 0, W, 100, H, W, fx, fy, cx, cy, torch.from_numpy(gold["k.c2w"]).to(dev),
                                torch.from_numpy(gold["k.depth"]).to(dev), torch.from_numpy(gold["k.color"]).to(dev), dev,
                                indices=torch.from_numpy(gold["k.indices"]).to(dev))
    got = m.keyframe_overlap_scores(ro, rd, gd, [torch.from_numpy(c) for c in gold["k.kf_c2w"]])
    assert np.array_equal(got, gold["k.scores"]), (got, gold["k.scores"])


@pytest.mark.gpu
def test_kernel_frustum_and_overlap_replica_scale_vs_oracle():
    """Replica room0 shape (680x1200, fine grid 43x56x74 = 178k voxels; middle 21x28x37): bit-exact with
    the oracle, many keyframes, depth holes."""
    from nice_slam_annotation_b200 import synth
    H, W, fx, fy, cx, cy, bound = synth.SHAPES["replica_room0"]
    bound_t = torch.tensor(bound, dtype=torch.float64)
    room = torch.stack([bound_t[:, 0] + 0.5, bound_t[:, 1] - 0.5], 1)
    poses = synth.orbit_poses(room, 12, seed=1)
    depth, color = synth.make_frame(H, W, fx, fy, cx, cy, room, poses[3], zero_frac=0.05, seed=5)
    dev = "cuda:0"
    m = _mapper(None, dev, H, W, fx, fy, cx, cy, bound)
    for shape in ((21, 28, 37), (43, 56, 74)):
        ref = FO.frustum_mask(poses[3].float().numpy(), "grid_fine", shape, depth.numpy(), bound, H, W, fx, fy, cx, cy)
        got = m.get_mask_from_c2w(poses[3].float(), "grid_fine", shape, depth)
        assert 0.02 < ref.mean() < 0.98
        assert np.array_equal(got, ref), int((got != ref).sum())
    g = torch.Generator().manual_seed(0)
    idx = torch.randint(H * W, (100,), generator=g)
    from nice_slam_annotation_b200.common import get_samples
    ro, rd, gd, _ = get_samples(0, H, 0, W, 100, H, W, fx, fy, cx, cy, poses[3].float().to(dev), depth.to(dev),
                                color.to(dev), dev, indices=idx.to(dev))
    pts = FO.keyframe_overlap_points(ro.cpu().numpy(), rd.cpu().numpy(), gd.cpu().numpy(), 16)
    ref = FO.keyframe_overlap(pts, [p.float().numpy() for p in poses], H, W, fx, fy, cx, cy)
    got = m.keyframe_overlap_scores(ro, rd, gd, [p.float() for p in poses])
    assert (ref > 0).sum() >= 2
    assert np.array_equal(got, ref), (got, ref)
