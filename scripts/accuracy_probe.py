"""Developer probe: outputs of Renderer.render_batch_ray at the bench workload against the oracle on the same GPU."""
import sys, torch
sys.path.insert(0, ".")
import bench
from nice_slam_annotation_b200.scene import build_scene
from oracle import nice_oracle as O
ns = bench.desaturate(build_scene("replica_room0", "cuda:0", seed=0, n_frames=5))
ro, rd, gd, gc = bench.make_batches(ns, 1000, 1, seed=100)[0]
with torch.no_grad():
    d, v, c = ns.renderer.render_batch_ray(ns.shared_c, ns.shared_decoders, rd, ro, "cuda:0", "color", gt_depth=gd)
    prm = {k: x.detach() for k, x in ns.shared_decoders.state_dict().items()}
    grids = {k: x.detach().contiguous() for k, x in ns.shared_c.items()}
    d0, v0, c0 = O.render_batch_ray(prm, grids, rd, ro, "color", gd, ns.bound.to("cuda:0"))
rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max())
print("ours vs fp32 oracle : depth %.2e var %.2e color %.2e" % (rel(d, d0), rel(v, v0), rel(c, c0)))
