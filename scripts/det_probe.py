import sys, torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import test_guard_bands_gpu as T
from nice_slam_annotation_b200.scene import build_scene
ns = build_scene("replica_room0", "cuda:0", seed=0, n_frames=2)
g = torch.Generator().manual_seed(5)
with torch.no_grad():
    for v in ns.shared_c.values():
        v.add_((torch.randn(v.shape, generator=g) * 0.3).to("cuda:0"))
step = T._make_step(ns, "color", 1000, True, ("color_decoder",))
ro, rd, gd, gc, keep = T._batch(ns, 1000, 7)
step.load(ro, rd, gd, gc, keep=keep)
step.run(); step.capture()
ng = sum(step.grids[k].numel() for k in step.grid_grad_kinds)
ref = None
wg = wd = 0.0
import os
spikes = []
for it in range(int(os.environ.get('DET_ITERS', '30'))):
    step.run(); torch.cuda.synchronize()
    f = step.bucket.flat.clone()
    if ref is None:
        ref = f; continue
    eg = float((f[:ng] - ref[:ng]).abs().max() / ref[:ng].abs().max())
    ed = float((f[ng:] - ref[ng:]).abs().max() / ref[ng:].abs().max())
    if eg > 2e-6 or ed > 2e-6:
        spikes.append((it, eg, ed, int((f[:ng] - ref[:ng]).abs().argmax()), int((f[ng:] - ref[ng:]).abs().argmax())))
    wg, wd = max(wg, eg), max(wd, ed)
print("spikes", spikes[:10], len(spikes))
print("grids", wg, "decoder", wd, "max grid", float(ref[:ng].abs().max()), "max dec", float(ref[ng:].abs().max()))
