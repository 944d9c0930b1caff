#!/usr/bin/env python
"""Ray-count / scene-size sweep of the engine step (BASELINE configs[3], configs[4] on one GPU).

    python scripts/sweep.py [--shapes replica_room0,apartment] [--rays 1000,5000,20000,100000] > profiles/rNN_sweep.json

Same step as bench.py (stage color, fwd + mapping loss + bwd to 3 grids and the colour decoder, CUDA-graph
replay, 256 MB L2 flush between steps); prints one JSON object with a row per (shape, rays)."""
import argparse
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402  (make_batches, constants)
from nice_slam_annotation_b200 import _lib  # noqa: E402
from nice_slam_annotation_b200.scene import build_scene  # noqa: E402
from nice_slam_annotation_b200.step import FusedStep  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="replica_room0,apartment")
    ap.add_argument("--rays", default="1000,5000,20000,100000")
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    dev = "cuda:0"
    L = _lib.lib()
    flush = torch.empty(bench.L2_FLUSH_BYTES // 4, device=dev)
    rows = []
    for shape in a.shapes.split(","):
        ns = build_scene(shape, dev, seed=0, n_frames=bench.FRAMES)
        dec, c = ns.shared_decoders, ns.shared_c
        grids_mb = sum(v.numel() * 4 for k, v in c.items() if k != "grid_coarse") / 1e6
        for n in [int(x) for x in a.rays.split(",")]:
            batches = bench.make_batches(ns, n, 4, seed=7)
            step = FusedStep(ns.renderer, dec, c, "color", n, loss="mapping", w_color=0.2, use_color=True,
                             grid_grads=("grid_middle", "grid_fine", "grid_color"), decoder_grads=("color_decoder",))
            kept = sum(b[0].shape[0] for b in batches) / len(batches)

            def run(i):
                step.load(*batches[i % len(batches)])
                step.run()
            for i in range(3):
                run(i)
            torch.cuda.synchronize()
            L.nsb_timing(1)
            for i in range(4):
                flush.fill_(float(i))
                run(i)
            torch.cuda.synchronize()
            km = []
            for which in range(3):
                ms, cnt = C.c_double(), C.c_int()
                L.nsb_timing_read_kernel(which, C.byref(ms), C.byref(cnt))
                km.append(ms.value / max(cnt.value, 1))
            L.nsb_timing(0)
            step.capture()
            evs = []
            for i in range(a.steps):
                flush.fill_(float(i))
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                run(i)
                e1.record()
                evs.append((e0, e1))
            torch.cuda.synchronize()
            ms = sum(x.elapsed_time(y) for x, y in evs) / a.steps
            pts = n * 48
            rows.append({"shape": shape, "grids_mb": round(grids_mb, 1), "rays": n, "kept": kept, "ms_per_step": ms,
                         "rays_per_s": kept / (ms * 1e-3), "fwd_ms": km[0], "bwd_ms": km[1], "dweight_ms": km[2],
                         "bwd_request_gbs": bench.BYTES_PER_POINT_BWD * pts / (km[1] * 1e-3) / 1e9,
                         "fwd_request_gbs": bench.BYTES_PER_POINT_FWD * pts / (km[0] * 1e-3) / 1e9})
            print(json.dumps(rows[-1]), file=sys.stderr, flush=True)
            del step
            torch.cuda.empty_cache()
    print(json.dumps({"what": "engine step sweep, 1 B200, stage color, fwd+loss+bwd, graph replay, L2 flushed",
                      "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
