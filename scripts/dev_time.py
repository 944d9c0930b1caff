"""Developer timing probe (not the benchmark): fwd / fwd+bwd of the fused render at Replica shape."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nice_slam_annotation_b200 import fused, synth
from nice_slam_annotation_b200.conv_onet.models.decoder import NICE
from oracle import nice_oracle as O

dev = "cuda:0"
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
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
grids = {k: v.to(dev).contiguous(memory_format=torch.channels_last_3d).requires_grad_(True)
         for k, v in O.init_grids(bound).items()}
print({k: tuple(v.shape) for k, v in grids.items()})
room = synth.room_of(bound)
c2w = synth.orbit_poses(room, 4)[0]
depth, color = synth.make_frame(H, W, fx, fy, cx, cy, room, c2w)
idx = torch.randint(H * W, (N,))
ro, rd, gd, gc = O.get_samples(0, H, 0, W, N, fx, fy, cx, cy, c2w, depth, color, idx)
ro, rd, gd, gc = ro.contiguous().to(dev), rd.to(dev), gd.to(dev), gc.to(dev)

def step(stage, bwd, rays_grad=False):
    r_o = ro.clone().requires_grad_(rays_grad)
    d, v, c = fused.render_batch_ray(grids, dec, rd, r_o, stage, gd, bound, N_samples=32, N_surface=16)
    if bwd:
        loss = (d - gd).abs().sum() + 0.2 * (c - gc).abs().sum()
        loss.backward()
        for g in grids.values():
            g.grad = None

for stage in ("middle", "fine", "color"):
    for bwd in (False, True):
        for _ in range(3):
            step(stage, bwd)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        K = 20
        for _ in range(K):
            step(stage, bwd)
        e1.record()
        torch.cuda.synchronize()
        print(f"N={N} stage={stage} bwd={bwd}: gpu {e0.elapsed_time(e1)/K*1e3:.1f} us/iter  wall {(time.perf_counter()-t0)/K*1e6:.1f} us/iter")
