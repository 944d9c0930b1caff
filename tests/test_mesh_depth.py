"""mesh depth images for the 2-D reconstruction metric (src/tools/eval_recon.py:131-210).
The reference's GL rasteriser is unavailable on this box; the oracle (oracle/mesh_oracle.py) states the image
as a ray cast, the kernel must equal it bit for bit, and both must equal the analytic depth of a box room."""
import numpy as np
import pytest
import torch

from oracle import mesh_oracle as MO

H, W, FOCAL = 48, 64, 40.0
CX, CY = W / 2.0 - 0.5, H / 2.0 - 0.5


def quad(z, half=1.0, flip=False):
    v = np.array([[-half, -half, z], [half, -half, z], [half, half, z], [-half, half, z]], dtype=np.float32)
    t = np.array([[0, 1, 2], [0, 2, 3]], dtype=np.int32)
    return v, (t[:, ::-1].copy() if flip else t)


def box_room(lo, hi, n):
    """the six faces of an axis-aligned box, n x n quads each"""
    lo, hi = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)
    verts, tris = [], []
    s = np.linspace(0.0, 1.0, n + 1)
    for axis in range(3):
        a, b = [k for k in range(3) if k != axis]
        for side in (lo[axis], hi[axis]):
            base = sum(len(v) for v in verts)
            ga, gb = np.meshgrid(lo[a] + s * (hi[a] - lo[a]), lo[b] + s * (hi[b] - lo[b]), indexing="ij")
            p = np.zeros((n + 1, n + 1, 3))
            p[..., axis], p[..., a], p[..., b] = side, ga, gb
            verts.append(p.reshape(-1, 3))
            i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
            q = (base + i * (n + 1) + j).reshape(-1)
            tris.append(np.stack([q, q + 1, q + n + 2], 1))
            tris.append(np.stack([q, q + n + 2, q + n + 1], 1))
    return np.concatenate(verts).astype(np.float32), np.concatenate(tris).astype(np.int32)


def random_pose(rng, lo, hi, margin=0.4):
    pos = rng.uniform(np.asarray(lo) + margin, np.asarray(hi) - margin)
    target = rng.uniform(-1, 1, 3)
    c2w = np.eye(4)
    c2w[:3] = MO.viewmatrix(target, np.array([0.0, 0.0, -1.0]), pos)
    return c2w


def box_depth(c2w, lo, hi, H, W, fx, fy, cx, cy):
    """analytic z-depth of the inside of a box: the ray parameter of dir = (.., .., 1) is the z-depth"""
    u, v = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    d = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], -1) @ c2w[:3, :3].T
    o = c2w[:3, 3]
    with np.errstate(divide="ignore"):
        t = np.maximum((np.asarray(lo) - o) / d, (np.asarray(hi) - o) / d)
    return t.min(-1)


def test_oracle_known_answers():
    eye = np.eye(4)
    v, t = quad(2.0)
    d = MO.mesh_depth(v, t, eye, H, W, FOCAL, FOCAL, CX, CY)
    # the quad covers |x| <= 1 at z = 2: pixels with |u - cx| <= 20 and |v - cy| <= 20
    uu, vv = np.meshgrid(np.arange(W), np.arange(H))
    inside = (np.abs(uu - CX) < 19.9) & (np.abs(vv - CY) < 19.9)
    outside = (np.abs(uu - CX) > 20.1) | (np.abs(vv - CY) > 20.1)
    assert np.all(d[inside] == np.float32(2.0)) and np.all(d[outside] == 0)
    # back faces are shown (mesh_show_back_face = True, eval_recon.py:157)
    vb, tb = quad(2.0, flip=True)
    assert np.array_equal(MO.mesh_depth(vb, tb, eye, H, W, FOCAL, FOCAL, CX, CY), d)
    # the nearer surface wins; beyond z far / behind the camera nothing is seen
    v2, t2 = quad(1.5, half=0.3)
    both = MO.mesh_depth(np.concatenate([v, v2]), np.concatenate([t, t2 + 4]), eye, H, W, FOCAL, FOCAL, CX, CY)
    assert both[H // 2, W // 2] == np.float32(1.5) and both[H // 2, 15] == np.float32(2.0)
    assert not MO.mesh_depth(*quad(25.0, half=30.0), eye, H, W, FOCAL, FOCAL, CX, CY).any()
    assert not MO.mesh_depth(*quad(-2.0), eye, H, W, FOCAL, FOCAL, CX, CY).any()
    assert MO.depth_l1(both, d) == pytest.approx(np.abs(both - d).mean())


def test_oracle_box_room_and_pixel_box():
    rng = np.random.default_rng(0)
    lo, hi = [-2.0, -1.5, -2.5], [2.0, 1.5, 2.5]
    v, t = box_room(lo, hi, 3)
    for _ in range(3):
        c2w = random_pose(rng, lo, hi)
        w2c = np.linalg.inv(c2w)
        d = MO.mesh_depth(v, t, w2c, H, W, FOCAL, FOCAL, CX, CY)
        assert np.array_equal(d, MO.mesh_depth(v, t, w2c, H, W, FOCAL, FOCAL, CX, CY, box=False))
        ref = box_depth(c2w, lo, hi, H, W, FOCAL, FOCAL, CX, CY)
        assert (d > 0).all() and np.abs(d - ref).max() < 2e-5 * ref.max()


def test_check_proj_and_viewmatrix():
    c2w = np.eye(4)
    c2w[:3] = MO.viewmatrix([0.0, 1.0, 0.0], np.array([0.0, 0.0, -1.0]), [0.5, 0.0, 0.2])
    assert np.allclose(c2w[:3, :3].T @ c2w[:3, :3], np.eye(3), atol=1e-12) and np.allclose(c2w[:3, 2], [0, 1, 0])
    ahead = c2w[:3, 3] + 3.0 * c2w[:3, 2]
    behind = c2w[:3, 3] - 3.0 * c2w[:3, 2]
    assert MO.check_proj(ahead[None], 500, 500, 300.0, 300.0, 249.5, 249.5, c2w)
    assert not MO.check_proj(behind[None], 500, 500, 300.0, 300.0, 249.5, 249.5, c2w)
    far_off = c2w[:3, 3] + 1.0 * c2w[:3, 2] + 5.0 * c2w[:3, 0]
    assert not MO.check_proj(far_off[None], 500, 500, 300.0, 300.0, 249.5, 249.5, c2w)


@pytest.mark.gpu
def test_kernel_matches_oracle_bit_exact():
    from nice_slam_annotation_b200.tools.eval_recon import render_mesh_depth
    rng = np.random.default_rng(1)
    lo, hi = [-2.0, -1.5, -2.5], [2.0, 1.5, 2.5]
    v, t = box_room(lo, hi, 4)
    # clutter: small random triangles, slivers, a degenerate one, one with a NaN vertex, one far outside
    n0 = len(v)
    extra = rng.uniform(-1.2, 1.2, (300, 3)).astype(np.float32)
    centres = extra[:, None, :] + rng.normal(0, 0.08, (300, 3, 3)).astype(np.float32)
    cv = centres.reshape(-1, 3)
    ct = (n0 + np.arange(900).reshape(300, 3)).astype(np.int32)
    cv[3:6] = cv[3]                                  # degenerate
    cv[7, 1] = np.nan
    cv[9:12] += 100.0
    v = np.concatenate([v, cv])
    t = np.concatenate([t, ct])
    c2ws = np.stack([random_pose(rng, lo, hi, margin=0.3) for _ in range(5)])
    d = render_mesh_depth(v, t, c2ws, H, W, FOCAL, FOCAL, CX, CY).cpu().numpy()
    assert d.shape == (5, H, W)
    for k in range(5):
        ref = MO.mesh_depth(v, t, np.linalg.inv(c2ws[k]).astype(np.float32), H, W, FOCAL, FOCAL, CX, CY)
        assert np.array_equal(d[k], ref), f"view {k}: {np.abs(d[k] - ref).max()} on {(d[k] != ref).sum()} pixels"
    one = render_mesh_depth(torch.from_numpy(v), torch.from_numpy(t), c2ws[0], H, W, FOCAL, FOCAL, CX, CY)
    assert one.shape == (H, W) and np.array_equal(one.cpu().numpy(), d[0])
    # z near / far window
    ref = MO.mesh_depth(v, t, np.linalg.inv(c2ws[1]).astype(np.float32), H, W, FOCAL, FOCAL, CX, CY, 0.8, 2.0)
    got = render_mesh_depth(v, t, c2ws[1], H, W, FOCAL, FOCAL, CX, CY, z_near=0.8, z_far=2.0).cpu().numpy()
    assert np.array_equal(got, ref)


@pytest.mark.gpu
def test_metric_size_box_room_analytic():
    """eval_recon's own size (500 x 500, focal 300, eval_recon.py:137-144) on a 235 k-triangle room: every pixel
    equals the analytic depth; the metric of a mesh against itself is 0 and against a room whose walls moved
    by 1 cm it is ~1 cm x the mean of z-depth / wall distance."""
    from nice_slam_annotation_b200.tools.eval_recon import render_mesh_depth, depth_l1
    rng = np.random.default_rng(2)
    lo, hi = np.array([-2.0, -1.5, -2.5]), np.array([2.0, 1.5, 2.5])
    v, t = box_room(lo, hi, 140)
    assert len(t) == 12 * 140 * 140
    c2ws = np.stack([random_pose(rng, lo, hi) for _ in range(8)])
    Hm = Wm = 500
    d = render_mesh_depth(v, t, c2ws, Hm, Wm, 300.0, 300.0, 249.5, 249.5).cpu().numpy()
    for k in range(8):
        ref = box_depth(c2ws[k], lo, hi, Hm, Wm, 300.0, 300.0, 249.5, 249.5)
        assert (d[k] > 0).all()
        assert np.abs(d[k] - ref).max() < 3e-5 * ref.max()
    m0, _ = depth_l1((v, t), (v, t), c2ws)
    assert m0 == 0.0
    v2, t2 = box_room(lo - 0.01, hi + 0.01, 97)
    m1, per_view = depth_l1((v, t), (v2, t2), c2ws)
    ref1 = np.mean([np.abs(box_depth(c, lo - 0.01, hi + 0.01, Hm, Wm, 300.0, 300.0, 249.5, 249.5)
                           - box_depth(c, lo, hi, Hm, Wm, 300.0, 300.0, 249.5, 249.5)).mean() for c in c2ws])
    assert per_view.shape == (8,) and abs(m1 - ref1) < 1e-5


@pytest.mark.gpu
def test_mesh_depth_edge_cases():
    from nice_slam_annotation_b200.tools.eval_recon import render_mesh_depth
    eye = np.eye(4)
    empty = render_mesh_depth(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32), eye, H, W, FOCAL, FOCAL, CX, CY)
    assert empty.shape == (H, W) and not empty.any()
    # one huge triangle through the camera plane (vertices behind the camera): the whole-image path
    v = np.array([[-50, -40, -1.0], [50, -40, -1.0], [0, 60, 9.0]], dtype=np.float32)
    t = np.array([[0, 1, 2]], dtype=np.int32)
    got = render_mesh_depth(v, t, eye, H, W, FOCAL, FOCAL, CX, CY).cpu().numpy()
    assert np.array_equal(got, MO.mesh_depth(v, t, eye, H, W, FOCAL, FOCAL, CX, CY)) and got.any()
    with pytest.raises(ValueError):
        render_mesh_depth(v, np.array([[0, 1, 3]]), eye, H, W, FOCAL, FOCAL, CX, CY)
    with pytest.raises(RuntimeError):
        render_mesh_depth(v, t, eye, H, W, FOCAL, FOCAL, CX, CY, z_near=0.0)


def test_product_path_needs_cuda():
    """no CPU fallback: without a CUDA device the product call raises instead of routing through the oracle"""
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from nice_slam_annotation_b200.tools.eval_recon import render_mesh_depth
    v, t = quad(2.0)
    with pytest.raises(RuntimeError):
        render_mesh_depth(v, t, np.eye(4), H, W, FOCAL, FOCAL, CX, CY)
