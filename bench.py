#!/usr/bin/env python
"""Benchmark of the render hot path: mapping rays/s, forward + backward, Replica room0 shape.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A *step* is one mapping iteration body (src/Mapper.py:598-624 without the optimiser): render a
batch of rays at stage ``color`` (32 uniform + 16 surface samples, middle + fine + colour
decoders), the mapping loss, and the backward pass to the three feature grids and the colour
decoder.  ``value`` times it with inputs resident in HBM; ``e2e`` times the same step with the ray batch in
pinned HOST memory (H2D copies and the D2H read of the loss inside the timed region) through the
engine call ``Mapper.optimize_map`` makes (``FusedStep``), ``e2e_autograd`` through
``Renderer.render_batch_ray`` + autograd.  ``--impl reference`` times the reference's CPU
implementation of the same step (the oracle restatement: /root/reference is not on the GPU box)
on the host cores.  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

SHAPE = "replica_room0"
STAGE = "color"
N_SAMPLES, N_SURFACE = 32, 16
FRAMES = 5
# algorithmic bytes (request level: 8 corners x 128 B per level per point)
BYTES_PER_POINT_BWD = 3 * 1024 + 3 * 2048          # recompute gather of 3 levels + scatter RMW of 3 levels
BYTES_PER_POINT_FWD = 3 * 1024
L2_FLUSH_BYTES = 256 << 20


def measured_peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm = sorted(float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def make_batches(ns, n_rays, n_batches, seed):
    """Ray batches as Mapper.optimize_map builds them (Mapper.py:543-596): n_rays // FRAMES pixels
    from each of FRAMES frames, concatenated, pre-filtered against the scene box."""
    from nice_slam_annotation_b200.common import get_samples
    g = torch.Generator(device="cpu").manual_seed(seed)
    out = []
    b = ns.bound.to(ns.frames[0][1].device)
    for _ in range(n_batches):
        parts = []
        for c2w, depth, color in ns.frames:
            idx = torch.randint(ns.H * ns.W, (n_rays // len(ns.frames),), generator=g)
            parts.append(get_samples(0, ns.H, 0, ns.W, idx.numel(), ns.H, ns.W, ns.fx, ns.fy, ns.cx, ns.cy, c2w,
                                     depth, color, depth.device, indices=idx))
        ro, rd, gd, gc = (torch.cat([p[i] for p in parts]) for i in range(4))
        t = (b.unsqueeze(0) - ro.unsqueeze(-1)) / rd.unsqueeze(-1)
        keep = torch.min(torch.max(t, dim=2)[0], dim=1)[0] >= gd
        out.append(tuple(x[keep].contiguous() for x in (ro, rd, gd, gc)))
    return out


def make_batches_cpu(ns, n_rays, n_batches, seed):
    """make_batches on the host through the oracle (reference arm: nothing of the product runs)."""
    from oracle import nice_oracle as O
    g = torch.Generator(device="cpu").manual_seed(seed)
    out = []
    for _ in range(n_batches):
        parts = []
        for c2w, depth, color in ns.frames:
            idx = torch.randint(ns.H * ns.W, (n_rays // len(ns.frames),), generator=g)
            parts.append(O.get_samples(0, ns.H, 0, ns.W, idx.numel(), ns.fx, ns.fy, ns.cx, ns.cy, c2w, depth, color, idx))
        ro, rd, gd, gc = (torch.cat([p[i] for p in parts]) for i in range(4))
        keep = O.prefilter_mask(ro, rd, gd, ns.bound)
        out.append(tuple(x[keep].contiguous() for x in (ro, rd, gd, gc)))
    return out


W_COLOR = 0.2            # mapping w_color_loss (configs/nice_slam.yaml)


# ---------------------------------------------------------------------------------------
def run_ours(args):
    import ctypes as C
    import torch.distributed as dist
    from nice_slam_annotation_b200 import _lib, fused
    from nice_slam_annotation_b200.scene import build_scene
    from nice_slam_annotation_b200.step import FusedStep

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    L = _lib.lib()

    ns = build_scene(SHAPE, dev, seed=0, n_frames=FRAMES)
    dec, c = ns.shared_decoders, ns.shared_c
    # what the mapper optimises at stage color: the three grids and the colour decoder
    for p in dec.parameters():
        p.requires_grad_(False)
    for p in dec.color_decoder.parameters():
        p.requires_grad_(True)
    grids = {k: (v.requires_grad_(True) if k != "grid_coarse" else v) for k, v in c.items()}
    leaves = [grids["grid_middle"], grids["grid_fine"], grids["grid_color"]] + list(dec.color_decoder.parameters())
    batches = make_batches(ns, args.rays, 8, seed=100 + rank)       # weak scaling: every rank its own rays
    n_rays = sum(b[0].shape[0] for b in batches) / len(batches)
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.float32, device=dev)

    # ---- engine path (value): forward + loss + backward as 3 library calls, CUDA-graph replay,
    #      gradients in one flat bucket, ONE NCCL all-reduce per step when sharded
    #      gradient exchange when sharded: the bucket lives in symmetric memory and ONE kernel
    #      (nsb_bucket_allreduce: multimem.ld_reduce / multimem.st through the NVSwitch, or peer ld/st) runs
    #      inside the same CUDA graph; --nccl-exchange (or no symmetric memory) falls back to a NCCL call
    peers, exchange = None, "none (1 GPU)"
    if world > 1:
        exchange = "NCCL all-reduce of the flat gradient bucket, one call per step"
        if not args.nccl_exchange:
            try:
                from nice_slam_annotation_b200.peers import PeerGroup
                peers = PeerGroup(device=torch.device(dev))
                exchange = ("nsb_bucket_allreduce: one kernel over symmetric memory ("
                            + ("NVLS multimem.ld_reduce / multimem.st" if (peers.multicast and world > 2)
                               else "peer ld / st over NVLink")
                            + "), inside the step's CUDA graph")
            except Exception as e:                                   # noqa: BLE001 - report and use NCCL
                exchange += f" (symmetric memory unavailable: {type(e).__name__}: {e})"
                peers = None
        flag = torch.tensor([1 if peers is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if peers is not None and int(flag.item()) == 0:
            peers = None
    step_kw = dict(loss="mapping", w_color=W_COLOR, use_color=True,
                   grid_grads=("grid_middle", "grid_fine", "grid_color"), decoder_grads=("color_decoder",))
    step = FusedStep(ns.renderer, dec, grids, STAGE, args.rays, peers=peers, **step_kw)
    if peers is not None and step.bucket.sparse:
        exchange = ("nsb_bucket_allreduce_sparse: one kernel over symmetric memory (touched-line words pushed to the "
                    "slice owner, only voxel lines some rank touched are reduced and stored back, "
                    + ("NVLS multimem.ld_reduce / multimem.st" if peers.multicast else "peer ld / st over NVLink")
                    + "), inside the step's CUDA graph; depth max over ranks: nsb_depth_max_peers in the same graph")

    def step_device(b):
        step.load(*b)
        return step.run()

    # ---- API path (e2e): Renderer.render_batch_ray + loss + backward through autograd, host batches
    def step_e2e(hb):
        ro, rd, gd, gc = (x.to(dev, non_blocking=True) for x in hb)
        depth, _, color = ns.renderer.render_batch_ray(grids, dec, rd, ro, dev, STAGE, gt_depth=gd)
        loss = fused.mapping_loss(depth, color, gd, gc, W_COLOR, use_color=True)
        loss.backward()
        if world > 1:
            flat = torch.cat([p.grad.reshape(-1) if p.grad.dim() != 5 else p.grad.permute(0, 2, 3, 4, 1).reshape(-1)
                              for p in leaves])
            dist.all_reduce(flat)
        for p in leaves:
            p.grad = None
        return loss.item()                                          # device -> host read of the result

    def step_engine_host(hb):
        step.load(*hb)                                              # pinned host -> static device buffers
        return step.run().item()

    def step_engine_packed(pb):
        if step.graph_host is not None:                             # H2D + batch load + iteration + D2H as ONE graph
            return step.step_from_host(pb)
        step.load_packed(pb)                                        # ONE pinned [n,10] array -> one H2D copy + one kernel
        return step.run().item()

    def timed(fn, items, steps, warmup, after_warmup=None):
        for i in range(warmup):
            fn(items[i % len(items)])
        torch.cuda.synchronize()
        if after_warmup:
            after_warmup()
        if world > 1:
            dist.barrier()
        evs = []
        for i in range(steps):
            flush.fill_(float(i))                                   # evict the grids from L2, untimed
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(items[(warmup + i) % len(items)])
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item() / steps

    def kernel_ms(which):
        ms, n = C.c_double(), C.c_int()
        _lib.check(L.nsb_timing_read_kernel(which, C.byref(ms), C.byref(n)))
        return ms.value / max(n.value, 1), n.value

    with ClockSampler(local) as clk:
        # (1) per-kernel durations: the same step, un-graphed, CUDA events around the decoder kernels
        def start_timing():
            L.nsb_timing(1)
        timed(step_device, batches, max(args.steps // 2, 3), args.warmup, start_timing)
        (fwd_ms, _), (bwd_ms, _), (dw_ms, _) = kernel_ms(0), kernel_ms(1), kernel_ms(2)
        x_kernel_ms = dmax_kernel_ms = None
        if world > 1:       # includes the wait for the slowest rank at the kernels' opening barriers
            x_kernel_ms, dmax_kernel_ms = kernel_ms(3)[0] or None, kernel_ms(4)[0] or None
        L.nsb_timing(0)
        # (2) value: graph replay
        L.nsb_launch_count(1)
        if not args.no_graph:
            step.capture()
        launches_per_step = L.nsb_launch_count(0) if not args.no_graph else None
        L.nsb_launch_count(1)
        ms_dev = timed(step_device, batches, args.steps, args.warmup)
        if launches_per_step is not None:
            launches = launches_per_step * args.steps               # kernels inside the replayed graph
        else:
            launches = int(L.nsb_launch_count(0) * args.steps / (args.steps + args.warmup))
        host_batches = [tuple(x.cpu().pin_memory() for x in b) for b in batches]
        ms_nccl = x_err = None
        x_err = None
        if peers is not None:                            # the same step with the NCCL all-reduce, for comparison
            step_peer, step = step, FusedStep(ns.renderer, dec, grids, STAGE, args.rays, **step_kw)
            if not args.no_graph:
                step.load(*batches[0]); step.run(); step.capture()
            ms_nccl = timed(step_device, batches, args.steps, args.warmup)
            # the two exchanges must agree (atomic scatter order differs between runs: not bit-exact)
            step.load(*batches[0]); step.run()
            step_peer.load(*batches[0]); step_peer.run()
            torch.cuda.synchronize()
            x_err = float((step.bucket.flat - step_peer.bucket.flat).abs().max() / step.bucket.flat.abs().max())
            peers.check_error()
            step = step_peer
        ms_e2e = timed(step_e2e, host_batches, args.steps, args.warmup)
        ms_eng4 = timed(step_engine_host, host_batches, args.steps, args.warmup)
        packed_ok = world == 1 or step._peer_max
        if packed_ok:
            packed_batches = [FusedStep.pack_batch(*b) for b in batches]
            if not args.no_graph:
                step.capture_host()
            ms_eng = timed(step_engine_packed, packed_batches, args.steps, args.warmup)
        else:
            ms_eng = ms_eng4

    peaks, peak_src = measured_peaks()
    traffic = args.traffic
    if traffic is None:                                  # dram bytes of the last committed ncu --set full capture
        try:
            t = json.load(open(os.path.join(REPO, "profiles", "ncu_traffic.json")))["k_decoders_backward"]
            traffic = t["dram_bytes_read"] + t["dram_bytes_write"]
        except (OSError, KeyError, ValueError):
            traffic = None
    S = N_SAMPLES + N_SURFACE
    pts = args.rays * S
    achieved = BYTES_PER_POINT_BWD * pts / (bwd_ms * 1e-3) / 1e9
    h2d = sum(x.numel() * x.element_size() for x in host_batches[0])
    out = {
        "metric": "mapping rays/s fwd+bwd at Replica shape", "value": world * n_rays / (ms_dev * 1e-3),
        "unit": "rays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"Replica room0 shape (680x1200) mapping step: {args.rays} rays from {FRAMES} frames "
                               f"(~{n_rays:.0f} kept by the box pre-filter) x ({N_SAMPLES} uniform + {N_SURFACE} surface) "
                               f"samples, stage color (middle+fine+color decoders), fwd + loss + bwd to 3 grids and "
                               f"the colour decoder; optimiser excluded",
                   "shape": SHAPE, "rays_per_gpu": args.rays, "samples_per_ray": S, "stage": STAGE,
                   "grids_mb": sum(v.numel() * 4 for v in c.values()) / 1e6,
                   "l2": "256 MB buffer written between timed steps (outside the per-step CUDA events)",
                   "engine": "FusedStep: render forward + loss + render backward as 3 library calls"
                             + ("" if args.no_graph else ", replayed as one CUDA graph"),
                   "parallelism": f"rays sharded over {world} GPU(s), grids+decoders replicated"
                                  + ("; gradient exchange: " + exchange if world > 1 else ""),
                   "exchange": exchange, "exchange_kernel_ms": x_kernel_ms, "depth_max_peers_kernel_ms": dmax_kernel_ms,
                   "ms_per_step_with_nccl_allreduce": ms_nccl,
                   "exchange_max_rel_diff_vs_nccl": x_err},
        "e2e": {"value": world * n_rays / (ms_eng * 1e-3), "unit": "rays/s", "ms_per_step": ms_eng,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8,
                "api": (("FusedStep.step_from_host(ray batch as one pinned [n,10] host array): memcpy into the pinned staging "
                         "buffer, then H2D copy + nsb_load_batch_packed + the iteration + D2H copy of the loss replayed as ONE "
                         "CUDA graph, stream synchronise, loss returned as a float"
                         if step.graph_host is not None else
                         "FusedStep.load_packed(ray batch as one pinned [n,10] host array: 1 H2D copy + nsb_load_batch_packed) "
                         "+ run() + loss.item()")
                        if packed_ok else "FusedStep.load(4 pinned host tensors) + run() + loss.item()")
                       + ": the engine call Mapper.optimize_map makes per iteration (2 C-ABI calls: "
                         "nsb_render_forward, nsb_render_backward with the fused mapping loss), loss read back every step",
                "ms_per_step_from_4_host_tensors": ms_eng4},
        "e2e_autograd": {"value": world * n_rays / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_step": ms_e2e,
                         "api": "Renderer.render_batch_ray + mapping loss + loss.backward() (torch autograd around the "
                                "same kernels), ray batch in pinned host memory"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_decoders_backward", "achieved": achieved,
                     "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                     "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else "fallback",
                     "traffic": traffic, "kernel_ms": bwd_ms, "algorithmic_bytes_per_launch": BYTES_PER_POINT_BWD * pts,
                     "forward_kernel_ms": fwd_ms,
                     "forward_achieved_gbs": BYTES_PER_POINT_FWD * pts / (fwd_ms * 1e-3) / 1e9,
                     "dweight_kernel_ms": dw_ms,
                     "how": "CUDA events around each launch on the launching stream, un-graphed pass of the same step"},
        "clocks": clk.summary(),
    }
    if rank == 0 and world == 1 and not args.no_tracking:
        out["tracking"] = tracking_bench(dev, iters=max(args.steps, 20) * 5, warmup=args.warmup)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["torch_cuda_baseline"] = torch_cuda_port(ns, batches, dev)
        out["cpu_baseline"] = cpu_baseline(ns, batches, budget_s=12.0)
    if rank == 0 and world == 1 and not args.no_tracking:
        out["bulk_forward"] = bulk_forward_bench(ns, dev)
        out["mesh_depth"] = mesh_depth_bench(dev)
        out["optimize_map"] = optimize_map_bench(ns, dev)              # last: it trains the scene
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------
def oracle_step_fn(ns, batches, device="cpu"):
    """The same step through the oracle restatement of the reference's torch code, on the CPU (or,
    for the torch-CUDA comparison line, the same torch ops on the GPU)."""
    from oracle import nice_oracle as O
    dec = ns.shared_decoders
    prm = {k: v.detach().to(device).clone() for k, v in dec.state_dict().items()}
    for k, v in prm.items():
        v.requires_grad_(k.startswith("color_decoder."))
    grids = {k: v.detach().to(device).contiguous().clone().requires_grad_(k != "grid_coarse")
             for k, v in ns.shared_c.items()}
    bound = ns.bound.to(device)
    cpu_batches = [tuple(x.to(device) for x in b) for b in batches]

    def step(i):
        ro, rd, gd, gc = cpu_batches[i % len(cpu_batches)]
        depth, _, color = O.render_batch_ray(prm, grids, rd, ro, STAGE, gd, bound, N_samples=N_SAMPLES,
                                             N_surface=N_SURFACE)
        loss = O.mapping_loss(depth, color, gd, gc)
        loss.backward()
        for t in list(prm.values()) + list(grids.values()):
            t.grad = None
        return float(loss), ro.shape[0]
    return step


def cpu_baseline(ns, batches, budget_s):
    step = oracle_step_fn(ns, batches)
    step(0)                                   # warm-up
    t0, n, rays = time.perf_counter(), 0, 0
    while True:
        _, r = step(n)
        n += 1
        rays += r
        if time.perf_counter() - t0 > budget_s or n >= 200:
            break
    dt = time.perf_counter() - t0
    return {"value": rays / dt, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{n} steps of the same workload (oracle/nice_oracle.py, torch CPU fp32, {dt:.1f} s)"}


def torch_cuda_port(ns, batches, dev, steps=20, warmup=3):
    """north_star's 50x bar is quoted against the reference's torch-CUDA render_batch_ray fwd+bwd.
    /root/reference is not on this box, so the line times the oracle restatement (the same ATen
    op sequence) on the GPU: a reported comparison, not a product path."""
    try:
        step = oracle_step_fn(ns, batches, device=dev)
        for i in range(warmup):
            step(i)
        torch.cuda.synchronize()
        ms, rays = [], 0
        for i in range(steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rays += step(warmup + i)[1]
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort()
        med = ms[len(ms) // 2]
        return {"value": rays / steps / (med * 1e-3), "unit": "rays/s", "ms_per_step": med, "kind": "port",
                "sample": f"median of {steps} steps, oracle/nice_oracle.py (torch {torch.__version__}, fp32, "
                          f"allow_tf32 off) on the same GPU, loss read back each step like the reference"}
    except Exception as e:                                   # a comparison line must not sink the bench
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


def tracking_bench(dev, iters, warmup):
    """BASELINE metric 2, configs[2]: tracking iterations/s at ScanNet shape, 1000 pixels, stage
    color, pose-gradient backward through the fused render, Adam on the 7 camera parameters.
    engine = Tracker.optimize_cam_fused (whole iteration replayed as one CUDA graph, no host read
    per iteration); api = Tracker.optimize_cam_in_batch + torch.optim.Adam + loss.item() per
    iteration, the reference's own loop shape (Tracker.py:244-275)."""
    from nice_slam_annotation_b200.Tracker import Tracker
    from nice_slam_annotation_b200.common import get_tensor_from_camera
    from nice_slam_annotation_b200.scene import build_scene
    from nice_slam_annotation_b200.step import TrackingStep
    ns = build_scene("scannet_scene0000", dev, seed=0, n_frames=2)
    cfg = ns.cfg
    cfg["tracking"].update({"pixels": 1000, "iters": 50, "lr": 5e-4, "seperate_LR": True,
                            "ignore_edge_W": 20, "ignore_edge_H": 20})
    tr = Tracker(cfg, None, ns)
    tr.update_para_from_mapping()
    c2w, depth, color = ns.frames[1]
    cam0 = get_tensor_from_camera(ns.frames[0][0].cpu()).to(dev).float()       # previous frame's pose as the start
    ts = TrackingStep(tr, 1000)
    ts.load_frame(color, depth)
    ts.reset(cam0)
    ts.iterate(warmup)
    ts.capture()
    ts.reset(cam0)
    ts.iterate(warmup)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ts.iterate(iters)
    e1.record()
    torch.cuda.synchronize()
    ms_engine = e0.elapsed_time(e1) / iters
    _, best = ts.result()
    # API loop
    T = cam0[-3:].clone().requires_grad_(True)
    quad = cam0[:4].clone().requires_grad_(True)
    opt = torch.optim.Adam([{"params": [T], "lr": 5e-4}, {"params": [quad], "lr": 1e-4}])
    for _ in range(warmup):
        tr.optimize_cam_in_batch(torch.cat([quad, T], 0), color, depth, 1000, opt)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        tr.optimize_cam_in_batch(torch.cat([quad, T], 0), color, depth, 1000, opt)
    e1.record()
    torch.cuda.synchronize()
    ms_api = e0.elapsed_time(e1) / iters
    return {"metric": "tracking iters/s", "value": 1e3 / ms_engine, "unit": "iters/s", "ms_per_iter": ms_engine,
            "api": {"value": 1e3 / ms_api, "unit": "iters/s", "ms_per_iter": ms_api,
                    "path": "Tracker.optimize_cam_in_batch + torch.optim.Adam + loss.item() per iteration"},
            "config": {"workload": "ScanNet scene0000 shape (460x620, 87.9 MB grids) tracking iteration: 1000 pixels x 48 "
                                   "samples, stage color, tracking loss with the dynamic-object median mask, backward "
                                   "to the rays, pose gradient, Adam(7 dof)", "iters_timed": iters,
                       "engine": "Tracker.optimize_cam_fused / step.TrackingStep: one CUDA graph per iteration, "
                                 "pixel draw + candidate rule on the device"},
            "final_loss": best}


def bulk_forward_bench(ns, dev):
    """SURVEY.md 8 f3, the bulk forward consumers: Renderer.render_img (Renderer.py:264-340, the whole
    680x1200 frame in 100k-ray batches, what the visualiser calls) and Renderer.eval_points on a
    256^3 query lattice in 500k-point batches (what Mesher.get_mesh issues, Mesher.py:563-575)."""
    c2w, depth, color = ns.frames[-1]
    r, dec, c = ns.renderer, ns.shared_decoders, ns.shared_c
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    ms_img = timed(lambda: r.render_img(c, dec, c2w, dev, "color", gt_depth=depth), 3)
    n = 256
    b = torch.as_tensor(ns.bound, dtype=torch.float32)
    ax = [torch.linspace(float(b[a][0]) + 0.01, float(b[a][1]) - 0.01, n, device=dev) for a in range(3)]
    pts = torch.stack(torch.meshgrid(*ax, indexing="ij"), -1).reshape(-1, 3)
    with torch.no_grad():
        ms_pts = timed(lambda: r.eval_points(pts, dec, c, "color", dev), 2)
    H, W = r.H, r.W
    return {"render_img": {"rays": H * W, "ms": ms_img, "rays_per_s": H * W / (ms_img * 1e-3),
                           "what": "Renderer.render_img, stage color, gt depth given, 100k-ray batches, forward only"},
            "eval_points": {"points": n ** 3, "ms": ms_pts, "points_per_s": n ** 3 / (ms_pts * 1e-3),
                            "what": "Renderer.eval_points, 256^3 lattice inside the bound, 500k-point batches, stage color"}}


def mesh_depth_bench(dev, n=200, views=64):
    """f4: depth images of a triangle mesh at the size of eval_recon.calc_2d_metric (500 x 500, focal 300,
    eval_recon.py:137-144): a box room of 12 n^2 triangles from ``views`` random poses inside it."""
    import numpy as np
    from nice_slam_annotation_b200.tools.eval_recon import render_mesh_depth, viewmatrix
    lo, hi = np.array([-2.0, -1.5, -2.5]), np.array([2.0, 1.5, 2.5])
    verts, tris, s = [], [], np.linspace(0.0, 1.0, n + 1)
    for axis in range(3):
        a, b = [k for k in range(3) if k != axis]
        for side in (lo[axis], hi[axis]):
            base = sum(len(v) for v in verts)
            ga, gb = np.meshgrid(lo[a] + s * (hi[a] - lo[a]), lo[b] + s * (hi[b] - lo[b]), indexing="ij")
            p = np.zeros((n + 1, n + 1, 3))
            p[..., axis], p[..., a], p[..., b] = side, ga, gb
            verts.append(p.reshape(-1, 3))
            i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
            q = (base + i * (n + 1) + j).reshape(-1)
            tris += [np.stack([q, q + 1, q + n + 2], 1), np.stack([q, q + n + 2, q + n + 1], 1)]
    v = torch.as_tensor(np.concatenate(verts), dtype=torch.float32, device=dev)
    t = torch.as_tensor(np.concatenate(tris), dtype=torch.int32, device=dev)
    rng = np.random.default_rng(0)
    c2ws = np.tile(np.eye(4), (views, 1, 1))
    for k in range(views):
        c2ws[k, :3] = viewmatrix(rng.uniform(-1, 1, 3), np.array([0.0, 0.0, -1.0]), rng.uniform(lo + 0.4, hi - 0.4))
    fn = lambda: render_mesh_depth(v, t, c2ws, 500, 500, 300.0, 300.0, 249.5, 249.5, device=dev)
    d = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        d = fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    return {"triangles": int(t.shape[0]), "views": views, "image": "500x500", "ms": ms,
            "images_per_s": views / (ms * 1e-3), "covered": float((d > 0).float().mean()),
            "what": "tools.eval_recon.render_mesh_depth (nsb_mesh_depth): headless ray-cast of the depth buffer "
                    "calc_2d_metric reads from an Open3D GL window (unavailable on this box), incl. the host-side "
                    "pose inversion and upload"}


def optimize_map_bench(ns, dev, iters=60):
    """Whole Mapper.optimize_map calls at the bench shape (Mapper.py:295-669: keyframe selection,
    frustum masks, then ``iters`` iterations of sampling 5 x 200 pixels, render fwd + bwd, Adam on the
    masked grids and the colour decoder).  loop = the device-resident MappingLoop (default),
    autograd = the same call with ``fused_loop = False`` (Renderer + autograd + torch.optim.Adam)."""
    from nice_slam_annotation_b200.Mapper import Mapper
    out = {}
    state = {k: v.detach().clone() for k, v in ns.shared_c.items()}
    dec_state = {k: v.detach().clone() for k, v in ns.shared_decoders.state_dict().items()}
    for name, fused_loop in (("loop", True), ("autograd", False)):
        for k, v in state.items():
            ns.shared_c[k] = v.clone().contiguous(memory_format=torch.channels_last_3d)
        ns.shared_decoders.load_state_dict(dec_state)
        mp = Mapper(ns.cfg, None, ns)
        mp.fused_loop = fused_loop
        kf = [{"gt_c2w": c2w, "idx": i, "color": color, "depth": depth, "est_c2w": c2w}
              for i, (c2w, depth, color) in enumerate(ns.frames[:-1])]
        mp.keyframe_dict = kf
        c2w, depth, color = ns.frames[-1]
        n_it = iters if fused_loop else max(iters // 3, 6)
        mp.optimize_map(n_it, 1.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)     # warm-up call
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mp.optimize_map(n_it, 1.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out[name] = {"ms_per_call": dt * 1e3, "iters": n_it, "ms_per_iter": dt * 1e3 / n_it,
                     "rays_per_s": ns.cfg["mapping"]["pixels"] * n_it / dt}
    out["config"] = ("Mapper.optimize_map, Replica room0 shape, 4 keyframes + current frame, 1000 pixels/iteration, "
                     "frustum feature selection on, stage schedule middle/fine/color, optimiser included; wall clock "
                     "around the call incl. keyframe selection and frustum masks")
    return out


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port) on the host cores, rank 0 only."""
    if int(os.environ.get("RANK", 0)) != 0:
        return
    from nice_slam_annotation_b200.scene import build_scene
    if "OMP_NUM_THREADS" in os.environ and int(os.environ.get("WORLD_SIZE", 1)) > 1:
        # torchrun pins OMP_NUM_THREADS=1 for its workers; this arm runs on rank 0 alone and may use the host:
        # the CPUs it is allowed on, which is what torch picks by itself at N=1
        try:
            torch.set_num_threads(max(1, min(len(os.sched_getaffinity(0)), 64)))
        except (AttributeError, OSError):
            pass
    ns = build_scene(SHAPE, "cpu", seed=0, n_frames=FRAMES)        # host only: no kernel of ours on this arm
    batches = make_batches_cpu(ns, args.rays, 4, seed=100)
    step = oracle_step_fn(ns, batches)
    for i in range(args.warmup):
        step(i)
    t0, rays = time.perf_counter(), 0
    for i in range(args.steps):
        rays += step(args.warmup + i)[1]
    dt = time.perf_counter() - t0
    v = rays / dt
    S = N_SAMPLES + N_SURFACE
    out = {"impl": "reference", "metric": "mapping rays/s fwd+bwd at Replica shape", "value": v, "unit": "rays/s",
           "n_gpus": int(os.environ.get("WORLD_SIZE", 1)), "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic",
           "config": {"workload": f"Replica room0 shape (680x1200) mapping step: {args.rays} rays from {FRAMES} frames x "
                                  f"({N_SAMPLES} uniform + {N_SURFACE} surface) samples, stage color, fwd + loss + bwd "
                                  f"to 3 grids and the colour decoder; optimiser excluded",
                      "shape": SHAPE, "rays_per_gpu": args.rays, "samples_per_ray": S, "stage": STAGE},
           "cpu_baseline": {"value": v, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "port",
                            "sample": f"{args.steps} full steps ({dt:.1f} s), oracle/nice_oracle.py on torch CPU"},
           "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rays", type=int, default=1000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tracking", action="store_true", help="skip the tracking iters/s section")
    ap.add_argument("--no-graph", action="store_true", help="launch the engine step kernel by kernel")
    ap.add_argument("--nccl-exchange", action="store_true",
                    help="N > 1: all-reduce the gradient bucket with NCCL instead of the peer-memory kernel")
    ap.add_argument("--traffic", type=float, default=None,
                    help="ncu dram bytes per launch of the dominant kernel (profiles/), reported as roofline.traffic")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
