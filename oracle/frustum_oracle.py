"""CPU restatement (numpy) of the two per-mapped-frame callers of the hot path:
``Mapper.get_mask_from_c2w`` (src/Mapper.py:100-187) and the scoring part of
``Mapper.keyframe_selection_overlap`` (src/Mapper.py:189-293).  Test infrastructure only.

The reference looks the depth image up with ``cv2.remap(..., INTER_LINEAR)``: OpenCV (4.x,
modules/imgproc/src/imgwarp.cpp, remapBilinear / initInterTab2D) rounds the float map to 1/32 pixel
(INTER_BITS = 5, ``cvRound(x * 32)``), takes the four float weights ``(1-fy)(1-fx), (1-fy)fx, fy(1-fx),
fy fx`` from a table of float products and sums the four taps left to right in float, with
BORDER_CONSTANT 0 outside the image.  ``remap_bilinear`` restates exactly that, so the oracle needs no
OpenCV on the GPU box; it is pinned against the reference (which does call cv2) by
tests/golden/frustum.npz.
"""
import numpy as np

__all__ = ["remap_bilinear", "project", "frustum_mask", "keyframe_overlap_points", "keyframe_overlap"]

INTER_BITS = 5
INTER_TAB = 1 << INTER_BITS


def remap_bilinear(img, mx, my):
    """cv2.remap(img, mx, my, INTER_LINEAR) for a float32 single-channel image and float32 maps
    (borderMode BORDER_CONSTANT, borderValue 0)."""
    img = np.ascontiguousarray(img, dtype=np.float32)
    H, W = img.shape
    mx = np.asarray(mx, dtype=np.float32)
    my = np.asarray(my, dtype=np.float32)

    def fix(m):
        v = (m * np.float32(INTER_TAB)).astype(np.float32)
        bad = ~(np.abs(v) < 2.0 ** 31)                       # NaN / overflow: cvtss2si gives INT_MIN
        r = np.rint(np.where(bad, 0, v)).astype(np.int64)    # round half to even, like cvRound
        return np.where(bad, -(1 << 31), r)
    sx, sy = fix(mx), fix(my)
    ix = np.clip(sx >> INTER_BITS, -32768, 32767)            # the integer part is stored as short
    iy = np.clip(sy >> INTER_BITS, -32768, 32767)
    fx = ((sx & (INTER_TAB - 1)).astype(np.float32) * np.float32(1.0 / INTER_TAB)).astype(np.float32)
    fy = ((sy & (INTER_TAB - 1)).astype(np.float32) * np.float32(1.0 / INTER_TAB)).astype(np.float32)
    one = np.float32(1.0)
    wx0, wx1, wy0, wy1 = one - fx, fx, one - fy, fy
    w = [wy0 * wx0, wy0 * wx1, wy1 * wx0, wy1 * wx1]        # float32 products (initInterTab2D)

    def tap(x, y):
        ok = (x >= 0) & (x < W) & (y >= 0) & (y < H)
        return np.where(ok, img[np.clip(y, 0, H - 1), np.clip(x, 0, W - 1)], np.float32(0))
    acc = (tap(ix, iy) * w[0]).astype(np.float32)
    acc = (acc + tap(ix + 1, iy) * w[1]).astype(np.float32)
    acc = (acc + tap(ix, iy + 1) * w[2]).astype(np.float32)
    acc = (acc + tap(ix + 1, iy + 1) * w[3]).astype(np.float32)
    return acc


def project(points, w2c, fx, fy, cx, cy):
    """world points [P,3] f32, w2c [4,4] f32 -> uv [P,2] f32, z [P] f64 (camera z + 1e-5).
    Mapper.py:129-146: float32 rigid transform, x negated, float64 intrinsics product."""
    p = np.asarray(points, dtype=np.float32)
    w = np.asarray(w2c, dtype=np.float32)
    cam = [((p[:, 0] * w[k, 0] + p[:, 1] * w[k, 1]).astype(np.float32) + p[:, 2] * w[k, 2]).astype(np.float32)
           + w[k, 3] for k in range(3)]
    x, y, z = (-cam[0]).astype(np.float64), cam[1].astype(np.float64), cam[2].astype(np.float64)
    u_h = fx * x + 0.0 * y + cx * z
    v_h = 0.0 * x + fy * y + cy * z
    zz = z + 1e-5
    uv = np.stack([u_h / zz, v_h / zz], -1).astype(np.float32)
    return uv, zz


def grid_points(bound, val_shape):
    """Voxel centres in the reference's order (X outermost, Z innermost), float32 torch.linspace values."""
    import torch
    b = [[float(bound[a][0]), float(bound[a][1])] for a in range(3)]
    xs = torch.linspace(b[0][0], b[0][1], int(val_shape[2])).numpy()
    ys = torch.linspace(b[1][0], b[1][1], int(val_shape[1])).numpy()
    zs = torch.linspace(b[2][0], b[2][1], int(val_shape[0])).numpy()
    X, Y, Z = np.meshgrid(xs, ys, zs, indexing="ij")
    return np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)


def frustum_mask(c2w, key, val_shape, depth, bound, H, W, fx, fy, cx, cy):
    """get_mask_from_c2w (Mapper.py:100-187): bool [X,Y,Z].  val_shape = grid.shape[2:] = (Z,Y,X)."""
    if key == "grid_coarse":
        return np.ones(tuple(val_shape)[::-1], dtype=bool)
    c2w = np.asarray(c2w, dtype=np.float32)
    if c2w.shape[0] == 3:
        c2w = np.concatenate([c2w, np.array([[0, 0, 0, 1]], dtype=np.float32)], 0)
    w2c = np.linalg.inv(c2w)
    pts = grid_points(bound, val_shape)
    uv, z = project(pts, w2c, fx, fy, cx, cy)
    depths = remap_bilinear(depth, uv[:, 0], uv[:, 1])
    mask = (uv[:, 0] < W) & (uv[:, 0] > 0) & (uv[:, 1] < H) & (uv[:, 1] > 0)
    depths = np.where(depths == 0, depths.max(), depths)
    mask = mask & (0 <= -z) & (-z <= (depths + np.float32(0.5)))
    d = pts - c2w[:3, 3][None].astype(np.float32)
    d2 = ((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32) + d[:, 2] * d[:, 2]).astype(np.float32)
    mask = mask | (d2 < np.float32(0.25))
    return mask.reshape(int(val_shape[2]), int(val_shape[1]), int(val_shape[0]))


def keyframe_overlap_points(rays_o, rays_d, gt_depth, N_samples=16):
    """Sample points of keyframe_selection_overlap (Mapper.py:213-222), float32 [N*N_samples, 3]."""
    import torch
    o, d = np.asarray(rays_o, np.float32), np.asarray(rays_d, np.float32)
    g = np.asarray(gt_depth, np.float32).reshape(-1, 1)
    t = torch.linspace(0., 1., steps=N_samples).numpy()
    near, far = (g * np.float32(0.8)).astype(np.float32), (g + np.float32(0.5)).astype(np.float32)
    z = (near * (np.float32(1.) - t) + far * t).astype(np.float32)
    return (o[:, None, :] + d[:, None, :] * z[:, :, None]).astype(np.float32).reshape(-1, 3)


def keyframe_overlap(points, keyframe_c2ws, H, W, fx, fy, cx, cy, edge=20):
    """percent_inside of every keyframe (Mapper.py:224-252)."""
    out = []
    for c2w in keyframe_c2ws:
        c2w = np.asarray(c2w, dtype=np.float32)
        if c2w.shape[0] == 3:
            c2w = np.concatenate([c2w, np.array([[0, 0, 0, 1]], dtype=np.float32)], 0)
        uv, z = project(points, np.linalg.inv(c2w), fx, fy, cx, cy)
        m = (uv[:, 0] < W - edge) & (uv[:, 0] > edge) & (uv[:, 1] < H - edge) & (uv[:, 1] > edge) & (z < 0)
        out.append(m.sum() / uv.shape[0])
    return np.array(out)
