"""Depth images of triangle meshes and the 2-D reconstruction metric (reference: src/tools/eval_recon.py).

The reference's ``calc_2d_metric`` (eval_recon.py:131-210) opens an Open3D GL window, renders the ground-truth
and the reconstructed mesh from 1000 random views inside the room and averages ``|gt_depth - rec_depth|``.
GL is unavailable on a headless CUDA box, so the depth buffer is ray-cast by ``nsb_mesh_depth``
(csrc/nsb_mesh.cu): same pinhole convention (extrinsic = inv(c2w), camera looks down +z, y down), back faces
shown, z far 20, 0 where nothing is seen.  Mesh loading, ICP alignment and the free-space sampling of
``get_cam_position`` use open3d / trimesh in the reference and stay with the caller: pass arrays."""
import ctypes as C

import numpy as np
import torch

from .. import _lib


def _f32(x, device):
    return torch.as_tensor(np.asarray(x, dtype=np.float32) if not torch.is_tensor(x) else x,
                           dtype=torch.float32, device=device).contiguous()


def render_mesh_depth(vertices, triangles, c2w, H, W, fx, fy, cx, cy, z_near=0.01, z_far=20.0, device="cuda:0"):
    """Depth images [K,H,W] fp32 (device) of a mesh from K pinhole views.

    vertices [V,3], triangles [T,3] (numpy or torch); c2w [4,4] or [K,4,4] camera-to-world as eval_recon.py:167-171
    builds them (the extrinsic handed to Open3D is inv(c2w), :183).  One view: returns [H,W]."""
    if not torch.cuda.is_available():
        raise RuntimeError("render_mesh_depth needs a CUDA device (there is no CPU path)")
    L = _lib.lib()
    dev = torch.device(device)
    v = _f32(vertices, dev).reshape(-1, 3)
    t = torch.as_tensor(np.asarray(triangles) if not torch.is_tensor(triangles) else triangles).to(
        device=dev, dtype=torch.int32).reshape(-1, 3).contiguous()
    if t.numel() and (int(t.min()) < 0 or int(t.max()) >= v.shape[0]):
        raise ValueError("triangle index out of range")
    c2w = np.asarray(c2w.cpu() if torch.is_tensor(c2w) else c2w, dtype=np.float64)
    single = c2w.ndim == 2
    c2w = c2w.reshape(-1, 4, 4)
    w2c = _f32(np.linalg.inv(c2w), dev).reshape(-1, 16)
    K = w2c.shape[0]
    depth = torch.empty(K, H, W, dtype=torch.float32, device=dev)
    ws_bytes = int(L.nsb_mesh_depth_workspace_bytes(t.shape[0], K))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.nsb_mesh_depth(v.data_ptr(), v.shape[0], t.data_ptr(), t.shape[0], w2c.data_ptr(), K, H, W,
                                    C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                                    C.c_float(z_near), C.c_float(z_far), ws.data_ptr(), ws_bytes,
                                    depth.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return depth[0] if single else depth


def depth_l1(gt_mesh, rec_mesh, c2ws, H=500, W=500, focal=300.0, z_near=0.01, z_far=20.0, device="cuda:0",
             views_per_call=64):
    """The 2-D metric of calc_2d_metric for given views: mean over views of mean |gt_depth - rec_depth| (metres;
    the reference prints it x 100, eval_recon.py:208-210).  gt_mesh / rec_mesh: (vertices, triangles);
    intrinsics as eval_recon.py:137-144 (cx = H/2 - 0.5, cy = W/2 - 0.5)."""
    cx, cy = H / 2.0 - 0.5, W / 2.0 - 0.5
    c2ws = np.asarray(c2ws, dtype=np.float64).reshape(-1, 4, 4)
    errs = []
    for i in range(0, len(c2ws), views_per_call):
        kw = dict(z_near=z_near, z_far=z_far, device=device)
        dg = render_mesh_depth(*gt_mesh, c2ws[i:i + views_per_call], H, W, focal, focal, cx, cy, **kw)
        dr = render_mesh_depth(*rec_mesh, c2ws[i:i + views_per_call], H, W, focal, focal, cx, cy, **kw)
        errs.append((dg - dr).abs().mean(dim=(1, 2)))
    errs = torch.cat(errs)
    return float(errs.mean()), errs


def viewmatrix(z, up, pos):
    """eval_recon.py:15-21: camera-to-world [3,4] looking along z."""
    z = np.asarray(z, dtype=np.float64)
    vec2 = z / np.linalg.norm(z)
    vec0 = np.cross(up, vec2)
    vec0 = vec0 / np.linalg.norm(vec0)
    vec1 = np.cross(vec2, vec0)
    vec1 = vec1 / np.linalg.norm(vec1)
    return np.stack([vec0, vec1, vec2, np.asarray(pos, dtype=np.float64)], 1)
