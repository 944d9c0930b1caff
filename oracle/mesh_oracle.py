"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the mesh depth images of the 2-D reconstruction metric
(reference: src/tools/eval_recon.py:131-210, calc_2d_metric).

The reference renders the ground-truth and the reconstructed mesh through an Open3D GL window
(vis.add_geometry / convert_from_pinhole_camera_parameters / capture_depth_float_buffer(True), back faces
shown, z far 20) and averages |gt_depth - rec_depth| over the image (eval_recon.py:205).  The GL path is
unavailable on this box, so parity with the reference's rasteriser is UNPINNED; this file states the image as
a ray cast (pixel (u, v) sees the nearest triangle along ((u - cx) / fx, (v - cy) / fy, 1) in the camera frame
of the extrinsic inv(c2w), eval_recon.py:182-187) in numpy float32, operation by operation as
csrc/nsb_mesh.cu evaluates it, and the kernel is checked against it bit for bit.  Only tests/ and
__graft_entry__.smoke() may import this module."""
import numpy as np

F = np.float32
BARY_EPS = F(1e-4)


def to_cam(w2c, p):
    """rows 0..2 of the extrinsic applied to points [n,3], float32, ((m0 x + m1 y) + m2 z) + m3"""
    M = np.asarray(w2c, dtype=F).reshape(4, 4)
    p = np.asarray(p, dtype=F)
    return np.stack([((M[k, 0] * p[:, 0] + M[k, 1] * p[:, 1]) + M[k, 2] * p[:, 2]) + M[k, 3] for k in range(3)], 1)


def ray_triangle(a, e1, e2, dx, dy, z_near, z_far):
    """Moeller-Trumbore from the camera centre; a, e1, e2: [3] float32; dx, dy: arrays.  Returns depth, < 0 = miss."""
    with np.errstate(all="ignore"):
        px = dy * e2[2] - e2[1]
        py = e2[0] - dx * e2[2]
        pz = dx * e2[1] - dy * e2[0]
        det = (e1[0] * px + e1[1] * py) + e1[2] * pz
        ok = np.abs(det) > F(1e-20)
        inv = F(1.0) / det
        tx, ty, tz = -a[0], -a[1], -a[2]
        bu = ((tx * px + ty * py) + tz * pz) * inv
        ok &= (bu >= -BARY_EPS) & (bu <= F(1.0) + BARY_EPS)
        qx = ty * e1[2] - tz * e1[1]
        qy = tz * e1[0] - tx * e1[2]
        qz = tx * e1[1] - ty * e1[0]
        bv = ((dx * qx + dy * qy) + qz) * inv
        ok &= (bv >= -BARY_EPS) & ((bu + bv) <= F(1.0) + BARY_EPS)
        t = ((e2[0] * qx + e2[1] * qy) + e2[2] * qz) * inv
        ok &= (t >= F(z_near)) & (t <= F(z_far))
    return np.where(ok, t, F(-1.0)).astype(F)


def mesh_depth(vertices, triangles, w2c, H, W, fx, fy, cx, cy, z_near=0.01, z_far=20.0, box=True):
    """depth [H,W] float32 of one view; 0 where no triangle is hit within [z_near, z_far].
    box = True tests only the pixels around a triangle's projection (float64, a few pixels of slack) -- the
    result is the same as testing every pixel, it is only faster."""
    cam = to_cam(w2c, vertices)
    tri = np.asarray(triangles, dtype=np.int64)
    fx, fy, cx, cy = F(fx), F(fy), F(cx), F(cy)
    us, vs = np.arange(W, dtype=F), np.arange(H, dtype=F)
    dxs, dys = (us - cx) / fx, (vs - cy) / fy
    out = np.full((H, W), np.inf, dtype=F)
    for i0, i1, i2 in tri:
        c = cam[[i0, i1, i2]]
        zmax, zmin = np.nanmax(c[:, 2]), np.nanmin(c[:, 2])
        if not (zmax >= F(z_near)) or not (zmin <= F(z_far)):
            continue
        u0, u1, v0, v1 = 0, W - 1, 0, H - 1
        if box and zmin > 1e-6 and np.isfinite(c).all():
            u = np.float64(fx) * c[:, 0].astype(np.float64) / c[:, 2] + np.float64(cx)
            v = np.float64(fy) * c[:, 1].astype(np.float64) / c[:, 2] + np.float64(cy)
            e = np.abs(c[1] - c[0]).sum() + np.abs(c[2] - c[0]).sum()
            pad = 2 + float(BARY_EPS) * float(e) * float(max(fx, fy)) / max(float(zmin), float(z_near))
            if not np.isfinite(pad) or pad > 4 * max(H, W):
                pad = 4 * max(H, W)
            u0, u1 = int(max(np.floor(u.min() - pad), 0)), int(min(np.ceil(u.max() + pad), W - 1))
            v0, v1 = int(max(np.floor(v.min() - pad), 0)), int(min(np.ceil(v.max() + pad), H - 1))
            if u0 > u1 or v0 > v1:
                continue
        a, e1, e2 = c[0], c[1] - c[0], c[2] - c[0]
        dx, dy = np.meshgrid(dxs[u0:u1 + 1], dys[v0:v1 + 1])
        t = ray_triangle(a, e1, e2, dx, dy, z_near, z_far)
        sub = out[v0:v1 + 1, u0:u1 + 1]
        np.minimum(sub, np.where(t >= 0, t, np.inf).astype(F), out=sub)
    out[np.isinf(out)] = 0
    return out


def depth_l1(gt_depth, rec_depth):
    """eval_recon.py:205: np.abs(gt_depth - ours_depth).mean() per image"""
    return np.abs(np.asarray(gt_depth, dtype=F) - np.asarray(rec_depth, dtype=F)).mean(axis=(-2, -1))


def viewmatrix(z, up, pos):
    """eval_recon.py:15-21"""
    def normalize(x):
        return x / np.linalg.norm(x)
    vec2 = normalize(np.asarray(z, dtype=np.float64))
    vec0 = normalize(np.cross(up, vec2))
    vec1 = normalize(np.cross(vec2, vec0))
    return np.stack([vec0, vec1, vec2, np.asarray(pos, dtype=np.float64)], 1)


def check_proj(points, W, H, fx, fy, cx, cy, c2w):
    """eval_recon.py:60-86: is any of the (unseen) points inside the view?  The y and z columns of c2w are
    negated first, x of the camera coordinates after, and a point counts when 0 <= -z and its pixel is
    strictly inside the image (float32 like the reference's .float() tensors)."""
    c2w = np.array(c2w, dtype=np.float64)
    c2w[:3, 1] *= -1.0
    c2w[:3, 2] *= -1.0
    w2c = np.linalg.inv(c2w).astype(F)
    pts = np.asarray(points).astype(F)
    hom = np.concatenate([pts, np.ones_like(pts[:, :1])], 1)
    cam = (hom @ w2c.T)[:, :3]
    cam[:, 0] *= -1
    K = np.array([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]], dtype=F)
    uv = cam @ K.T
    z = uv[:, 2] + F(1e-5)
    u, v = uv[:, 0] / z, uv[:, 1] / z
    m = (0 <= -z) & (u < W) & (u > 0) & (v < H) & (v > 0)
    return bool(m.sum() > 0)
