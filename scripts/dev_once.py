"""Developer probe: a few fwd+bwd iterations of one stage (for ncu launch lists)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nice_slam_annotation_b200 import fused, synth
from nice_slam_annotation_b200.conv_onet.models.decoder import NICE
from oracle import nice_oracle as O

dev = "cuda:0"
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
stage = sys.argv[2] if len(sys.argv) > 2 else "color"
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
rays_grad = len(sys.argv) > 4 and sys.argv[4] == "rays"
H, W, fx, fy, cx, cy, bnd = synth.SHAPES["replica_room0"]
bound = O.round_bound(bnd)
torch.manual_seed(0)
dec = NICE(coarse=True)
for d in (dec.middle_decoder, dec.fine_decoder, dec.color_decoder):
    d.bound = bound
dec.coarse_decoder.bound = bound * 2
dec = dec.to(dev)
for p in dec.parameters():
    p.requires_grad_(False)
grids = {k: v.to(dev).contiguous(memory_format=torch.channels_last_3d).requires_grad_(not rays_grad)
         for k, v in O.init_grids(bound).items()}
room = synth.room_of(bound)
c2w = synth.orbit_poses(room, 4)[0]
depth, color = synth.make_frame(H, W, fx, fy, cx, cy, room, c2w)
idx = torch.randint(H * W, (N,))
ro, rd, gd, gc = O.get_samples(0, H, 0, W, N, fx, fy, cx, cy, c2w, depth, color, idx)
ro, rd, gd, gc = ro.contiguous().to(dev), rd.to(dev), gd.to(dev), gc.to(dev)
for _ in range(iters):
    r_o = ro.clone().requires_grad_(rays_grad)
    d, v, c = fused.render_batch_ray(grids, dec, rd, r_o, stage, gd, bound, N_samples=32, N_surface=16)
    loss = (d - gd).abs().sum() + 0.2 * (c - gc).abs().sum()
    loss.backward()
torch.cuda.synchronize()
print("done")
