"""Developer probe: the guard-band test cases followed by a long run-to-run comparison of the gradient bucket, several
times in one process; reports every deviation above 2e-6 with the segment and index it occurred at."""
import os, sys, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import test_guard_bands_gpu as T
from nice_slam_annotation_b200.scene import build_scene
ns = build_scene("replica_room0", "cuda:0", seed=0, n_frames=2)
g = torch.Generator().manual_seed(5)
with torch.no_grad():
    for v in ns.shared_c.values():
        v.add_((torch.randn(v.shape, generator=g) * 0.3).to("cuda:0"))
cases = [("color", 1000, False, ("color_decoder",)), ("color", 777, True, ("color_decoder",)), ("color", 1, True, ()),
         ("color", 129, True, ("middle_decoder", "fine_decoder", "color_decoder")), ("fine", 333, False, ()), ("middle", 61, True, ())]
for rep in range(int(os.environ.get("DET_REPS", "4"))):
    for c in cases:
        T.test_guard_bands_stay_intact(ns, *c)
    step = T._make_step(ns, "color", 1000, True, ("color_decoder",))
    ro, rd, gd, gc, keep = T._batch(ns, 1000, 7)
    step.load(ro, rd, gd, gc, keep=keep)
    step.run(); step.capture()
    ng = sum(step.grids[k].numel() for k in step.grid_grad_kinds)
    ref, spikes, wg, wd = None, [], 0.0, 0.0
    for it in range(int(os.environ.get("DET_ITERS", "300"))):
        step.run(); torch.cuda.synchronize()
        f = step.bucket.flat.clone()
        if ref is None:
            ref = f; continue
        dg, dd = (f[:ng] - ref[:ng]).abs(), (f[ng:] - ref[ng:]).abs()
        eg, ed = float(dg.max() / ref.abs().max()), float(dd.max() / ref.abs().max())
        if eg > 2e-6 or ed > 2e-6:
            spikes.append((it, eg, ed, int(dg.argmax()), int(dd.argmax()), int((dg > 1e-4).sum()), int((dd > 1e-4).sum())))
        wg, wd = max(wg, eg), max(wd, ed)
    print("rep", rep, "worst grids", wg, "decoder", wd, "spikes", len(spikes), spikes[:6], flush=True)
