"""Engine tracking iterations / s (the graph-replayed TrackingStep of bench.tracking_bench), for tuning runs:
    NSB_BWD_SHARES=a,b,c NSB_FWD_SHARES=a,b,c python scripts/track_time.py [iters]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nice_slam_annotation_b200.Tracker import Tracker
from nice_slam_annotation_b200.common import get_tensor_from_camera
from nice_slam_annotation_b200.scene import build_scene
from nice_slam_annotation_b200.step import TrackingStep
dev = "cuda:0"
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
ns = build_scene("scannet_scene0000", dev, seed=0, n_frames=2)
ns.cfg["tracking"].update({"pixels": 1000, "iters": 50, "lr": 5e-4, "seperate_LR": True, "ignore_edge_W": 20, "ignore_edge_H": 20})
tr = Tracker(ns.cfg, None, ns)
tr.update_para_from_mapping()
c2w, depth, color = ns.frames[1]
cam0 = get_tensor_from_camera(ns.frames[0][0].cpu()).to(dev).float()
ts = TrackingStep(tr, 1000)
ts.load_frame(color, depth)
ts.reset(cam0); ts.iterate(5); ts.capture(); ts.reset(cam0); ts.iterate(5)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ts.iterate(iters); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(os.environ.get("NSB_BWD_SHARES", "-"), os.environ.get("NSB_FWD_SHARES", "-"), "ms/iter %.4f  iters/s %.0f" % (ms, 1000 / ms))
