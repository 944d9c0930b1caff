#!/usr/bin/env python
"""Developer probe: per-kernel times of the bench step with / without the colour decoder's weight gradient
(what the record writes + embedding gradient cost the backward), on the de-saturated headline fixture."""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nice_slam_annotation_b200 import _lib  # noqa: E402
from nice_slam_annotation_b200.scene import build_scene  # noqa: E402
from nice_slam_annotation_b200.step import FusedStep  # noqa: E402

dev = "cuda:0"
L = _lib.lib()
ns = bench.desaturate(build_scene("replica_room0", dev, seed=0, n_frames=bench.FRAMES))
bench.train_only_color_decoder(ns.shared_decoders)
batches = bench.make_batches(ns, 1000, 4, seed=100)
flush = torch.empty(bench.L2_FLUSH_BYTES // 4, device=dev)
out = {}
for name, kw in (("with_dw", bench.STEP_KW), ("no_dw", dict(bench.STEP_KW, decoder_grads=()))):
    step = FusedStep(ns.renderer, ns.shared_decoders, bench.grids_of(ns), "color", 1000, **kw)
    for i in range(3):
        step.load(*batches[i % 4]); step.run()
    torch.cuda.synchronize()
    L.nsb_timing(1)
    for i in range(8):
        flush.fill_(float(i)); step.load(*batches[i % 4]); step.run()
    torch.cuda.synchronize()
    km = []
    for w in range(3):
        ms, n = C.c_double(), C.c_int()
        L.nsb_timing_read_kernel(w, C.byref(ms), C.byref(n))
        km.append(ms.value / max(n.value, 1))
    L.nsb_timing(0)
    out[name] = {"fwd_ms": km[0], "bwd_ms": km[1], "dweight_ms": km[2]}
print(json.dumps(out))
