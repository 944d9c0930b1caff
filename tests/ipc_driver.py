"""Driver of tests/test_shared_memory_gpu.py, run as its own interpreter (the 'NICE_SLAM main process'): it creates
the shared objects, spawns the mapper process on them, then plays the tracker.  A separate interpreter keeps every
CUDA-IPC handle out of the pytest process (torch's IPC teardown at interpreter exit is fragile: the in-process
version of this test passed and then crashed pytest at exit), and this one leaves through os._exit once the result
is printed.

    python tests/ipc_driver.py prepared|unprepared      -> prints IPC_OK or raises
"""
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
DEV = "cuda:0"


def mapper_process(ns, iters, done, release):
    """Child: what the reference's mapping process does with the shared objects (one optimize_map call)."""
    from nice_slam_annotation_b200.Mapper import Mapper
    mp = Mapper(ns.cfg, None, ns)
    kf = [{"gt_c2w": c2w, "idx": i, "color": color, "depth": depth, "est_c2w": c2w}
          for i, (c2w, depth, color) in enumerate(ns.frames[:-1])]
    mp.keyframe_dict = kf
    c2w, depth, color = ns.frames[-1]
    mp.optimize_map(iters, 1.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)
    torch.cuda.synchronize()
    done.put("ok")
    release.wait(timeout=300)       # keep the process (and its IPC mappings) alive until the parent has looked
    os._exit(0)


def main(prepared):
    import torch.multiprocessing as mp
    from nice_slam_annotation_b200 import synth
    from nice_slam_annotation_b200.Tracker import Tracker
    from nice_slam_annotation_b200.scene import build_scene
    from nice_slam_annotation_b200.shared import prepare_shared
    synth.SHAPES["ipc"] = (48, 64, 60.0, 60.0, 31.5, 23.5, [[-0.6, 0.9], [-0.5, 0.7], [-0.4, 0.5]])
    ns = build_scene("ipc", DEV, seed=0, n_frames=3)
    ns.cfg["grid_len"]["coarse"] = 0.8
    ns.cfg["mapping"].update({"pixels": 300, "frustum_feature_selection": False, "keyframe_selection_method": "global"})
    with torch.no_grad():
        for v in ns.shared_c.values():
            v.add_(torch.randn_like(v) * 0.3)
    if prepared:
        prepare_shared(ns)
    else:
        # what the reference's own grid_init allocates: plain contiguous channels-first grids
        ns.shared_c = {k: v.contiguous() for k, v in ns.shared_c.items()}
        assert not ns.shared_c["grid_fine"].is_contiguous(memory_format=torch.channels_last_3d)
    ns.shared_decoders.share_memory()                      # NICE_SLAM.py:134 (no-op for CUDA tensors, kept for the contract)
    ns.mapping_idx = None
    grids0 = {k: v.detach().clone() for k, v in ns.shared_c.items()}
    dec0 = {k: v.detach().clone() for k, v in ns.shared_decoders.state_dict().items()}
    ptrs0 = {k: v.data_ptr() for k, v in ns.shared_c.items()}
    ctx = mp.get_context("spawn")
    done, release = ctx.Queue(), ctx.Event()
    child = ctx.Process(target=mapper_process, args=(ns, 12, done, release))
    child.start()
    try:
        assert done.get(timeout=300) == "ok"
        torch.cuda.synchronize()
        # the parent's handles still point at the storages it created ...
        assert {k: v.data_ptr() for k, v in ns.shared_c.items()} == ptrs0
        # ... and they hold what the child trained: this is what Tracker.update_para_from_mapping copies
        tr = Tracker(ns.cfg, None, ns)
        tr.update_para_from_mapping()
        for k in ("grid_middle", "grid_fine", "grid_color"):
            assert not torch.equal(tr.c[k], grids0[k]), f"{k}: the tracker still sees the untrained grid"
        assert torch.equal(tr.c["grid_coarse"], grids0["grid_coarse"])
        sd = tr.decoders.state_dict()
        changed = [k for k in sd if k.startswith("color_decoder.") and not torch.equal(sd[k], dec0[k])]
        assert len(changed) >= 20, changed
        assert all(torch.equal(sd[k], dec0[k]) for k in sd if k.startswith("middle_decoder."))
    finally:
        release.set()
        child.join(timeout=60)
        if child.is_alive():
            child.terminate()
    assert child.exitcode == 0, child.exitcode
    print("IPC_OK", flush=True)
    os._exit(0)


if __name__ == "__main__":
    main(sys.argv[1] == "prepared")
