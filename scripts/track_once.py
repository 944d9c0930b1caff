"""A few fused tracking iterations, un-graphed (for an ncu launch list of one tracking iteration)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nice_slam_annotation_b200.Tracker import Tracker
from nice_slam_annotation_b200.common import get_tensor_from_camera
from nice_slam_annotation_b200.scene import build_scene
from nice_slam_annotation_b200.step import TrackingStep
dev = "cuda:0"
ns = build_scene("scannet_scene0000", dev, seed=0, n_frames=2)
ns.cfg["tracking"].update({"pixels": 1000, "iters": 50, "lr": 5e-4, "seperate_LR": True, "ignore_edge_W": 20, "ignore_edge_H": 20})
tr = Tracker(ns.cfg, None, ns)
tr.update_para_from_mapping()
c2w, depth, color = ns.frames[1]
cam0 = get_tensor_from_camera(ns.frames[0][0].cpu()).to(dev).float()
ts = TrackingStep(tr, 1000)
ts.load_frame(color, depth)
ts.reset(cam0)
ts.iterate(6)
torch.cuda.synchronize()
print("done", ts.result()[1])
