#!/usr/bin/env python
"""Benchmark of the render hot path: mapping rays/s, forward + backward, Replica room0 shape.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A *step* is one mapping iteration body (src/Mapper.py:598-624 without the optimiser): render a
batch of rays at stage ``color`` (32 uniform + 16 surface samples, middle + fine + colour
decoders), the mapping loss, and the backward pass to the three feature grids and the colour
decoder.  ``value`` times it with inputs resident in HBM; ``e2e`` times the same step with the ray batch in
pinned HOST memory (H2D copies and the D2H read of the loss inside the timed region) through the
engine call ``Mapper.optimize_map`` makes (``FusedStep``), ``e2e_autograd`` through
``Renderer.render_batch_ray`` + autograd.

Headline fixture: reference-init decoders and grids, every grid de-saturated with seeded N(0, 0.3^2)
noise (the fixture of the full-size parity tests): every sample carries gradient, so the backward gets
no zero-tile discount.  ``fixtures`` reports the same step on the saturated reference-init state (F1)
and on a trained-like state (F2: F1 + mapping iterations on the synthetic room) with the share of live
tiles of each.  Further sections (one JSON line in total, rank 0): ``tracking`` (configs[2]),
``apartment`` (configs[3]: 5000 rays on 125 MB grids, strong-scaled over the ranks), ``imap`` and
``sweep`` (configs[4]: iMAP* tensor-core case, 1K-1M rays), ``roofline``, ``cpu_baseline``,
``torch_cuda_baseline`` (the reference's own files from baseline/_ref on the same GPU).

``--impl reference`` times the reference's own CPU implementation of the same step (baseline/_ref, the
unmodified reference files; the oracle restatement only if that copy is missing) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import types

import torch

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

SHAPE = "replica_room0"
STAGE = "color"
N_SAMPLES, N_SURFACE = 32, 16
S_TOTAL = N_SAMPLES + N_SURFACE
FRAMES = 5
W_COLOR = 0.2            # mapping w_color_loss (configs/nice_slam.yaml)
# algorithmic work per point (request level: 8 corners x 128 B per level per point, no cache credit)
BYTES_PER_POINT_FWD = 3 * 1024                 # trilinear gather of 3 levels
BYTES_PER_POINT_SCATTER = 3 * 2048             # scatter read-modify-write of 3 levels (mapping backward)
BYTES_PER_POINT_REGATHER = 3 * 1024            # coordinate-gradient re-gather (tracking / BA backward only)
BYTES_PER_POINT_BWD = BYTES_PER_POINT_SCATTER  # the mapping backward does no gather (ReLU bits + deposited inputs)
FLOP_PER_POINT_FWD = 103306                    # 2 x MAC, stage color
FLOP_PER_POINT_FWD_BWD = 309918                # mapping as the reference does it (dX + dW for all three decoders)
FLOP_PER_POINT_TRACK = 206612                  # dX only
IMAP_FLOP_PER_POINT_FWD = 2 * (3 * 93 + 93 * 256 + 3 * 256 * 256 + 256 * 4)     # 443 438
IMAP_N_SAMPLES, IMAP_N_IMPORTANCE = 32, 12
# reference-equivalent evaluations per ray: 32 forward-only (first pass) + 44 x 3 (second pass,
# fwd + dX + dW) + 32 x 3 (regulariser)
IMAP_FLOP_PER_RAY = IMAP_FLOP_PER_POINT_FWD * (IMAP_N_SAMPLES + 3 * (IMAP_N_SAMPLES + IMAP_N_IMPORTANCE) + 3 * IMAP_N_SAMPLES)
L2_FLUSH_BYTES = 256 << 20
SWEEP_RAYS = (1000, 4000, 16000, 64000, 256000, 1000000)
APARTMENT_RAYS = 5000
FIXTURE = ("reference-init decoders (config.get_model, seed 0) and grids (grid_init), every grid + N(0, 0.3^2) noise from "
           "torch.Generator().manual_seed(5) on the CPU: de-saturated, every sample carries gradient")


def measured_peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1590.0}, "fallback"


def workload_config(rays, grids_mb=None):
    """The ``config`` object of the JSON line: the workload only, identical for both arms (--impl ours /
    --impl reference), so that the driver's same_config check compares like with like."""
    return {"workload": f"Replica room0 shape (680x1200) mapping step (configs[1]): {rays} rays from {FRAMES} frames x "
                        f"({N_SAMPLES} uniform + {N_SURFACE} surface) samples, stage color (middle+fine+color decoders), "
                        f"fwd + mapping loss + bwd to 3 grids and the colour decoder; optimiser excluded",
            "shape": SHAPE, "rays_per_gpu": rays, "frames": FRAMES, "samples_per_ray": S_TOTAL, "stage": STAGE,
            "fixture": FIXTURE,
            "l2": "GPU arm: 256 MB buffer written between timed steps (outside the per-step CUDA events); CPU arm: n/a"}


def desaturate(ns, seed=5, std=0.3):
    """The headline fixture: add seeded N(0, std^2) noise to every grid (CPU generator, so that the GPU arm and the
    CPU reference arm hold bit-identical states)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k in ("grid_coarse", "grid_middle", "grid_fine", "grid_color"):
            v = ns.shared_c.get(k)
            if v is None:
                continue
            v.add_((torch.randn(v.shape, generator=g) * std).to(v.device))
    return ns


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


def make_batches(ns, n_rays, n_batches, seed, frames=None):
    """Ray batches as Mapper.optimize_map builds them (Mapper.py:543-596): n_rays // FRAMES pixels
    from each of FRAMES frames, concatenated, pre-filtered against the scene box."""
    from nice_slam_annotation_b200.common import get_samples
    g = torch.Generator(device="cpu").manual_seed(seed)
    out = []
    frames = ns.frames if frames is None else frames
    b = ns.bound.to(frames[0][1].device)
    for _ in range(n_batches):
        parts = []
        for c2w, depth, color in frames:
            idx = torch.randint(ns.H * ns.W, (max(n_rays // len(frames), 1),), generator=g)
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


STEP_KW = dict(loss="mapping", w_color=W_COLOR, use_color=True,
               grid_grads=("grid_middle", "grid_fine", "grid_color"), decoder_grads=("color_decoder",))


def grids_of(ns):
    return {k: (v.requires_grad_(True) if k != "grid_coarse" else v) for k, v in ns.shared_c.items()}


def train_only_color_decoder(dec):
    for p in dec.parameters():
        p.requires_grad_(False)
    for p in dec.color_decoder.parameters():
        p.requires_grad_(True)


class Ctx:
    """Process-wide state of one bench run."""

    def __init__(self, args):
        import torch.distributed as dist
        from nice_slam_annotation_b200 import _lib
        self.args = args
        self.rank = int(os.environ.get("RANK", 0))
        self.world = int(os.environ.get("WORLD_SIZE", 1))
        self.local = int(os.environ.get("LOCAL_RANK", 0))
        torch.cuda.set_device(self.local)
        self.dev = f"cuda:{self.local}"
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device(self.dev))
        self.dist = dist
        self.L = _lib.lib()
        self._lib = _lib
        self.flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.float32, device=self.dev)
        self.peers, self.exchange = None, "none (1 GPU)"
        self.alive = []      # sharded steps: their symmetric buffers are released at exit, never between graph captures
        if self.world > 1:
            self.exchange = "NCCL all-reduce of the flat gradient bucket, one call per step"
            if not args.nccl_exchange:
                try:
                    from nice_slam_annotation_b200.peers import PeerGroup
                    self.peers = PeerGroup(device=torch.device(self.dev))
                except Exception as e:                                   # noqa: BLE001 - report and use NCCL
                    self.exchange += f" (symmetric memory unavailable: {type(e).__name__}: {e})"
                    self.peers = None
            flag = torch.tensor([1 if self.peers is not None else 0], device=self.dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if self.peers is not None and int(flag.item()) == 0:
                self.peers = None

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, x):
        t = torch.tensor([x], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.item()

    def sum_over_ranks(self, x):
        t = torch.tensor([x], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.item()

    def timed(self, fn, items, steps, warmup, after_warmup=None, flush=True):
        """ms per step: W warm-up calls, then K calls each bracketed by CUDA events on the launching stream, an L2
        flush (untimed) before each; barrier + synchronize on both sides; MAX over ranks."""
        for i in range(warmup):
            fn(items[i % len(items)])
        torch.cuda.synchronize()
        if after_warmup:
            after_warmup()
        self.barrier()
        evs = []
        for i in range(steps):
            if flush:
                self.flush.fill_(float(i))                          # evict the grids from L2, untimed
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(items[(warmup + i) % len(items)])
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        self.barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        return self.max_over_ranks(ms) / steps

    def kernel_ms(self, which):
        import ctypes as C
        ms, n = C.c_double(), C.c_int()
        self._lib.check(self.L.nsb_timing_read_kernel(which, C.byref(ms), C.byref(n)))
        return ms.value / max(n.value, 1), n.value


def tf32_matmul_peak(dev, n=8192, reps=10):
    """Dense TF32 tensor-core throughput of this GPU, measured like MEASURED_PEAKS.json's bf16 number (torch.matmul,
    n^3, best of ``reps``, CUDA events) with allow_tf32 on: the denominator of the tensor-bound roofline fractions."""
    old = torch.backends.cuda.matmul.allow_tf32
    try:
        torch.backends.cuda.matmul.allow_tf32 = True
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        torch.matmul(a, b)
        torch.cuda.synchronize()
        best = float("inf")
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    except Exception:                                                    # noqa: BLE001
        return None
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def ncu_traffic():
    """dram bytes / tensor-pipe share per launch from the last committed ``ncu --set full`` capture (profiles/)."""
    try:
        return json.load(open(os.path.join(REPO, "profiles", "ncu_traffic.json")))
    except (OSError, ValueError):
        return {}


# ---------------------------------------------------------------------------------------
def headline(ctx, ns):
    from nice_slam_annotation_b200 import fused
    from nice_slam_annotation_b200.step import FusedStep
    args, dev, world, rank, dist = ctx.args, ctx.dev, ctx.world, ctx.rank, ctx.dist
    L = ctx.L
    dec, c = ns.shared_decoders, ns.shared_c
    train_only_color_decoder(dec)                                       # what the mapper optimises at stage color
    grids = grids_of(ns)
    leaves = [grids["grid_middle"], grids["grid_fine"], grids["grid_color"]] + list(dec.color_decoder.parameters())
    batches = make_batches(ns, args.rays, 8, seed=100 + rank)           # weak scaling: every rank its own rays
    n_rays = sum(b[0].shape[0] for b in batches) / len(batches)

    # ---- engine path (value): forward + loss + backward as 2 library calls, CUDA-graph replay, gradients in one
    #      flat bucket; sharded: the bucket lives in symmetric memory and ONE kernel (nsb_bucket_allreduce*) runs
    #      inside the same graph; --nccl-exchange (or no symmetric memory) falls back to a NCCL call
    peers, exchange = ctx.peers, ctx.exchange
    step = FusedStep(ns.renderer, dec, grids, STAGE, args.rays, peers=peers, **STEP_KW)
    if peers is not None:
        mc = "NVLS multimem.ld_reduce / multimem.st" if peers.multicast else "peer ld / st over NVLink"
        if step.bucket.sparse:
            exchange = ("nsb_bucket_allreduce_sparse: one kernel over symmetric memory (touched-line words pushed to the "
                        "slice owner, only voxel lines some rank touched are reduced and stored back, " + mc +
                        "), inside the step's CUDA graph; depth max over ranks: nsb_depth_max_peers in the same graph")
        else:
            exchange = "nsb_bucket_allreduce: one kernel over symmetric memory (" + mc + "), inside the step's CUDA graph"

    def step_device(b):
        step.load(*b)
        return step.run()

    # ---- API path (e2e_autograd): Renderer.render_batch_ray + loss + backward through autograd, host batches
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

    with ClockSampler(ctx.local) as clk:
        # (1) per-kernel durations: the same step, un-graphed, CUDA events around the decoder kernels
        ctx.timed(step_device, batches, max(args.steps // 2, 3), args.warmup, lambda: L.nsb_timing(1))
        (fwd_ms, _), (bwd_ms, _), (dw_ms, _) = ctx.kernel_ms(0), ctx.kernel_ms(1), ctx.kernel_ms(2)
        x_kernel_ms = dmax_kernel_ms = None
        if world > 1:       # includes the wait for the slowest rank at the kernels' opening barriers
            x_kernel_ms, dmax_kernel_ms = ctx.kernel_ms(3)[0] or None, ctx.kernel_ms(4)[0] or None
        L.nsb_timing(0)
        live16, live128 = step.live_tiles(16), step.live_tiles(128)
        # (2) value: graph replay
        L.nsb_launch_count(1)
        if not args.no_graph:
            step.capture()
        launches_per_step = L.nsb_launch_count(0) if not args.no_graph else None
        L.nsb_launch_count(1)
        ms_dev = ctx.timed(step_device, batches, args.steps, args.warmup)
        if launches_per_step is not None:
            launches = launches_per_step * args.steps               # kernels inside the replayed graph
        else:
            launches = int(L.nsb_launch_count(0) * args.steps / (args.steps + args.warmup))
        host_batches = [tuple(x.cpu().pin_memory() for x in b) for b in batches]
        ms_nccl = x_err = None
        if peers is not None:                            # the same step with the NCCL all-reduce, for comparison
            step_peer, step = step, FusedStep(ns.renderer, dec, grids, STAGE, args.rays, **STEP_KW)
            if not args.no_graph:
                step.load(*batches[0]); step.run(); step.capture()
            ms_nccl = ctx.timed(step_device, batches, args.steps, args.warmup)
            # the two exchanges must agree (atomic scatter order differs between runs: not bit-exact)
            step.load(*batches[0]); step.run()
            step_peer.load(*batches[0]); step_peer.run()
            torch.cuda.synchronize()
            x_err = float((step.bucket.flat - step_peer.bucket.flat).abs().max() / step.bucket.flat.abs().max())
            peers.check_error()
            step = step_peer
        ms_e2e = ctx.timed(step_e2e, host_batches, args.steps, args.warmup)
        ms_eng4 = ctx.timed(step_engine_host, host_batches, args.steps, args.warmup)
        packed_ok = world == 1 or step._peer_max
        if packed_ok:
            packed_batches = [FusedStep.pack_batch(*b) for b in batches]
            if not args.no_graph:
                step.capture_host()
            ms_eng = ctx.timed(step_engine_packed, packed_batches, args.steps, args.warmup)
        else:
            ms_eng = ms_eng4

    total_rays = ctx.sum_over_ranks(n_rays)
    grids_mb = sum(v.numel() * 4 for v in c.values()) / 1e6
    h2d = sum(x.numel() * x.element_size() for x in host_batches[0])
    api = (("FusedStep.step_from_host(ray batch as one pinned [n,10] host array): memcpy into the pinned staging buffer, then "
            "H2D copy + nsb_load_batch_packed + the iteration + D2H copy of the loss replayed as ONE CUDA graph, stream "
            "synchronise, loss returned as a float" if step.graph_host is not None else
            "FusedStep.load_packed(ray batch as one pinned [n,10] host array: 1 H2D copy + nsb_load_batch_packed) + run() + "
            "loss.item()") if packed_ok else "FusedStep.load(4 pinned host tensors) + run() + loss.item()")
    out = {
        "metric": "mapping rays/s fwd+bwd at Replica shape", "value": total_rays / (ms_dev * 1e-3),
        "unit": "rays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.rays),
        "detail": {"rays_kept_by_prefilter_per_gpu": n_rays, "grids_mb": grids_mb,
                   "engine": "FusedStep: render forward + fused mapping loss + render backward as 2 library calls"
                             + ("" if args.no_graph else ", replayed as one CUDA graph"),
                   "arithmetic": "3xTF32 tensor-core products with fp32 accumulation (fp32-class accuracy), fp32 Fourier "
                                 "embedding, fp64 sample depths / depth / variance like the reference",
                   "parallelism": f"rays sharded over {world} GPU(s), grids+decoders replicated"
                                  + ("; gradient exchange: " + exchange if world > 1 else ""),
                   "exchange": exchange, "exchange_kernel_ms": x_kernel_ms, "depth_max_peers_kernel_ms": dmax_kernel_ms,
                   "ms_per_step_with_nccl_allreduce": ms_nccl, "exchange_max_rel_diff_vs_nccl": x_err,
                   "live_tiles": {"16_points": live16, "128_points": live128,
                                  "what": "share of tiles whose incoming gradient is not exactly zero (the decoder backward "
                                          "skips the others) on this fixture"}},
        "e2e": {"value": total_rays / (ms_eng * 1e-3), "unit": "rays/s", "ms_per_step": ms_eng,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8,
                "api": api + ": the engine call Mapper.optimize_map makes per iteration (2 C-ABI calls: nsb_render_forward, "
                             "nsb_render_backward with the fused mapping loss), loss read back every step",
                "ms_per_step_from_4_host_tensors": ms_eng4},
        "e2e_autograd": {"value": total_rays / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_step": ms_e2e,
                         "api": "Renderer.render_batch_ray + mapping loss + loss.backward() (torch autograd around the "
                                "same kernels), ray batch in pinned host memory"},
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
    }
    out["roofline"] = roofline_block(ctx, args.rays * S_TOTAL, fwd_ms, bwd_ms, dw_ms, ms_dev)
    return out, step, batches, x_err


def roofline_block(ctx, pts, fwd_ms, bwd_ms, dw_ms, step_ms):
    """The dominant kernel is k_decoders_forward_tc.  SURVEY.md 8d classifies the path as HBM-bound at request level;
    measured (ncu, profiles/), the 48.5 MB of grids live in L2 and DRAM traffic is a few percent of peak, so the
    binding resource is tensor-core issue / latency: ``bound`` is 'tensor', achieved = algorithmic FLOP/s of the
    forward kernel, peak = dense TF32 matmul throughput measured on this GPU.  The HBM views (request level as 8d
    defines it, and ncu DRAM bytes) are given beside it per kernel and for the whole step."""
    peaks, peak_src = measured_peaks()
    tf32 = tf32_matmul_peak(ctx.dev)
    hbm = peaks["hbm_gbs"]
    tr = ncu_traffic()

    def dram(kernel):
        t = tr.get(kernel)
        return None if not t else t.get("dram_bytes_read", 0) + t.get("dram_bytes_write", 0)

    def hbm_view(bytes_alg, ms, kernel):
        d = dram(kernel) if kernel else None
        return {"algorithmic_bytes_per_launch": bytes_alg, "kernel_ms": ms,
                "request_level_gbs": bytes_alg / (ms * 1e-3) / 1e9, "request_level_frac_of_hbm": bytes_alg / (ms * 1e-3) / 1e9 / hbm,
                "dram_bytes_per_launch_ncu": d,
                "dram_gbs": None if d is None else d / (ms * 1e-3) / 1e9,
                "dram_frac_of_hbm": None if d is None else d / (ms * 1e-3) / 1e9 / hbm}
    peak_tc = tf32 if tf32 else peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"]) / 2
    fwd_tf = FLOP_PER_POINT_FWD * pts / (fwd_ms * 1e-3) / 1e12
    step_tf = FLOP_PER_POINT_FWD_BWD * pts / (step_ms * 1e-3) / 1e12
    bf16 = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
    return {
        "bound": "tensor", "kernel": "k_decoders_forward_tc", "achieved": fwd_tf, "peak": peak_tc, "unit": "TFLOP/s",
        "frac": fwd_tf / peak_tc,
        "peak_source": ("dense TF32 torch.matmul 8192^3 measured in this run (best of 10, CUDA events)" if tf32 else
                        "fallback: half of MEASURED_PEAKS.json bf16_tflops_sustained") +
                       f"; MEASURED_PEAKS.json ({peak_src}) bf16 sustained {bf16} TF, hbm {hbm} GB/s",
        "traffic": dram("k_decoders_forward_tc"),
        "tensor_pipe_active_pct_ncu": (tr.get("k_decoders_forward_tc") or {}).get("sm__pipe_tensor_cycles_active_pct"),
        "algorithmic_flop_per_launch": FLOP_PER_POINT_FWD * pts, "kernel_ms": fwd_ms,
        "frac_of_bf16_sustained": fwd_tf / bf16,
        "note": "algorithmic fp32 FLOPs (2 x MAC, SURVEY.md 8d); the kernels issue 3x as many tensor-core FLOPs (3xTF32)",
        "step": {"algorithmic_flop": FLOP_PER_POINT_FWD_BWD * pts, "ms": step_ms, "tflops": step_tf,
                 "frac_of_tf32_peak": step_tf / peak_tc, "frac_of_bf16_sustained": step_tf / bf16},
        "hbm": {"peak_gbs": hbm,
                "k_decoders_forward_tc": hbm_view(BYTES_PER_POINT_FWD * pts, fwd_ms, "k_decoders_forward_tc"),
                "k_decoders_backward_tc": hbm_view(BYTES_PER_POINT_BWD * pts, bwd_ms, "k_decoders_backward_tc"),
                "k_dweight_tc": {"kernel_ms": dw_ms, "algorithmic_bytes_per_launch": 0,
                                 "dram_bytes_per_launch_ncu": dram("k_dweight_tc") or dram("k_dweight"),
                                 "note": "SURVEY.md 8d assigns the weight gradient no bytes; what it streams is the "
                                         "16-point records (47.6 KB each, live ones only)"},
                "step": hbm_view((BYTES_PER_POINT_FWD + BYTES_PER_POINT_BWD) * pts, step_ms, None),
                "note": "request level = 8 corners x 128 B per level per point, no cache credit (SURVEY.md 8d): forward "
                        "3072 B/pt (gather), mapping backward 6144 B/pt (scatter RMW; it does not re-gather); the grids "
                        "(48.5 MB) are L2-resident, so DRAM traffic (ncu) is far below the request-level figure"},
        "how": "CUDA events around each launch on the launching stream, un-graphed pass of the same step",
    }


# ---------------------------------------------------------------------------------------
def fixture_figures(ctx, ns_head, head_value):
    """The same engine step on the other two fixtures of SURVEY.md 8c: F1 (reference init: saturated, most tiles
    behind the first opaque sample carry exactly zero gradient, which the backward skips) and F2 (trained-like: F1
    followed by mapping iterations of Mapper.optimize_map on the synthetic room with the first-frame schedule)."""
    from nice_slam_annotation_b200.scene import build_scene
    from nice_slam_annotation_b200.step import FusedStep
    args, dev = ctx.args, ctx.dev
    out = {}

    def figure(ns):
        dec = ns.shared_decoders
        train_only_color_decoder(dec)
        grids = grids_of(ns)
        step = FusedStep(ns.renderer, dec, grids, STAGE, args.rays, **STEP_KW)
        batches = make_batches(ns, args.rays, 8, seed=100)
        n = sum(b[0].shape[0] for b in batches) / len(batches)

        def fn(b):
            step.load(*b)
            return step.run()
        fn(batches[0])
        ctx.L.nsb_timing(1)
        for i in range(4):
            fn(batches[i])
        torch.cuda.synchronize()
        km = [ctx.kernel_ms(w)[0] for w in range(3)]
        ctx.L.nsb_timing(0)
        live = step.live_tiles(16)
        sel = (step.gt_depth > 0) & step.keep.bool()
        l1 = float((step.depth[sel] - step.gt_depth[sel].double()).abs().mean())
        step.capture()
        ms = ctx.timed(fn, batches, args.steps, args.warmup)
        return {"rays_per_s": n / (ms * 1e-3), "ms_per_step": ms, "fwd_ms": km[0], "bwd_ms": km[1], "dweight_ms": km[2],
                "live_tiles_16": live, "rendered_depth_l1_m": l1}

    out["F1_reference_init"] = figure(build_scene(SHAPE, dev, seed=0, n_frames=FRAMES))
    ns2 = build_scene(SHAPE, dev, seed=0, n_frames=FRAMES)
    try:
        out["F2_trained_like"] = dict(train_fixture(ns2, dev), **figure(ns2))
    except Exception as e:                                               # noqa: BLE001 - a side figure must not sink the bench
        out["F2_trained_like"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    out["headline_fixture_value"] = head_value
    out["what"] = ("engine step (graph replay, L2 flushed) at the headline workload on the other fixtures; live_tiles_16 = "
                   "share of 16-point tiles with a non-zero incoming gradient")
    return out


def train_fixture(ns, dev, iters=160, pixels=1000):
    """F2: ``iters`` iterations of Mapper.optimize_map (middle -> fine -> color schedule, lr x 5 as on the first
    frame, Mapper.py:740-760 / configs/nice_slam.yaml:77) over the 5 synthetic frames, all voxels trainable."""
    import random

    import numpy as np
    from nice_slam_annotation_b200.Mapper import Mapper
    ns.cfg["mapping"].update({"pixels": pixels, "frustum_feature_selection": False, "keyframe_selection_method": "global",
                              "mapping_window_size": FRAMES})
    mp = Mapper(ns.cfg, None, ns)
    kf = [{"gt_c2w": c2w, "idx": i, "color": color, "depth": depth, "est_c2w": c2w}
          for i, (c2w, depth, color) in enumerate(ns.frames[:-1])]
    mp.keyframe_dict = kf
    c2w, depth, color = ns.frames[-1]
    random.seed(0), np.random.seed(0), torch.manual_seed(0), torch.cuda.manual_seed(0)
    t0 = time.perf_counter()
    mp.optimize_map(iters, 5.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)
    torch.cuda.synchronize()
    for k in list(ns.shared_c):
        ns.shared_c[k] = mp.c[k].detach()
    return {"recipe": f"F1 + {iters} iterations of Mapper.optimize_map ({pixels} pixels from {FRAMES} frames, lr factor 5, "
                      f"no frustum mask)", "train_s": time.perf_counter() - t0}


# ---------------------------------------------------------------------------------------
def oracle_step_fn(ns, batches, device="cpu"):
    """The same step through the oracle restatement of the reference's torch code (fallback when baseline/_ref is
    missing), on the CPU or, for the torch-CUDA comparison line, the same torch ops on the GPU."""
    from oracle import nice_oracle as O
    dec = ns.shared_decoders
    prm = {k: v.detach().to(device).clone() for k, v in dec.state_dict().items()}
    for k, v in prm.items():
        v.requires_grad_(k.startswith("color_decoder."))
    grids = {k: v.detach().to(device).contiguous().clone().requires_grad_(k != "grid_coarse")
             for k, v in ns.shared_c.items()}
    bound = ns.bound.to(device)
    dev_batches = [tuple(x.to(device) for x in b) for b in batches]

    def step(i):
        ro, rd, gd, gc = dev_batches[i % len(dev_batches)]
        depth, _, color = O.render_batch_ray(prm, grids, rd, ro, STAGE, gd, bound, N_samples=N_SAMPLES,
                                             N_surface=N_SURFACE)
        loss = O.mapping_loss(depth, color, gd, gc)
        loss.backward()
        for t in list(prm.values()) + list(grids.values()):
            t.grad = None
        return float(loss), ro.shape[0]
    return step, "port", "oracle/nice_oracle.py (restatement of the reference's torch code; baseline/_ref is missing)"


def reference_step_fn(ns, batches, device="cpu"):
    """The same step through the UNMODIFIED reference files (baseline/_ref: src.utils.Renderer.render_batch_ray on
    src.conv_onet decoders, the loss lines of Mapper.py:604-612, loss.backward()); falls back to the oracle port."""
    try:
        from baseline import ref_harness as RH
        if not RH.available():
            raise FileNotFoundError("baseline/_ref is missing")
        slam = RH.reference_scene(ns, device)
        fn = RH.mapping_step(slam, device, STAGE, W_COLOR)
        dev_batches = [tuple(x.to(device) for x in b) for b in batches]

        def step(i):
            ro, rd, gd, gc = dev_batches[i % len(dev_batches)]
            return float(fn(ro, rd, gd, gc).detach()), ro.shape[0]           # the reference reads loss.item() too
        return step, "reference", "baseline/_ref (unmodified src/utils/Renderer.py, src/conv_onet, src/common.py)"
    except Exception as e:                                                   # noqa: BLE001
        step, kind, what = oracle_step_fn(ns, batches, device)
        return step, kind, what + f" [{type(e).__name__}: {e}]"[:160]


def cpu_baseline(ns, batches, budget_s):
    step, kind, what = reference_step_fn(ns, batches, "cpu")
    step(0)                                   # warm-up
    t0, n, rays = time.perf_counter(), 0, 0
    while True:
        _, r = step(n)
        n += 1
        rays += r
        if time.perf_counter() - t0 > budget_s or n >= 200:
            break
    dt = time.perf_counter() - t0
    return {"value": rays / dt, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": kind,
            "sample": f"{n} steps of the same workload ({what}, torch CPU fp32, {dt:.1f} s)"}


def torch_cuda_baseline(ns, batches, dev, steps=20, warmup=3):
    """north_star's 50x bar is quoted against the reference's torch-CUDA render_batch_ray fwd+bwd: the reference's
    own files on the same GPU (a reported comparison, not a product path)."""
    try:
        step, kind, what = reference_step_fn(ns, batches, dev)
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
        return {"value": rays / steps / (med * 1e-3), "unit": "rays/s", "ms_per_step": med, "kind": kind,
                "sample": f"median of {steps} steps, {what} (torch {torch.__version__}, fp32, allow_tf32 off) on the same "
                          f"GPU, loss read back each step like the reference"}
    except Exception as e:                                   # a comparison line must not sink the bench
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


def tracking_bench(dev, iters, warmup):
    """BASELINE metric 2, configs[2]: tracking iterations/s at ScanNet shape, 1000 pixels, stage
    color, pose-gradient backward through the fused render, Adam on the 7 camera parameters.
    engine = Tracker.optimize_cam_fused (whole iteration replayed as one CUDA graph, no host read
    per iteration); api = Tracker.optimize_cam_in_batch + torch.optim.Adam + loss.item() per
    iteration, the reference's own loop shape (Tracker.py:244-275); torch_cuda_baseline = the reference's own
    Tracker.optimize_cam_in_batch (baseline/_ref) on the same GPU and state."""
    from nice_slam_annotation_b200.Tracker import Tracker
    from nice_slam_annotation_b200.common import get_tensor_from_camera
    from nice_slam_annotation_b200.scene import build_scene
    from nice_slam_annotation_b200.step import TrackingStep
    ns = build_scene("scannet_scene0000", dev, seed=0, n_frames=2)
    desaturate(ns)
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

    def api_loop(tracker, n):
        T = cam0[-3:].clone().requires_grad_(True)
        quad = cam0[:4].clone().requires_grad_(True)
        opt = torch.optim.Adam([{"params": [T], "lr": 5e-4}, {"params": [quad], "lr": 1e-4}])
        for _ in range(warmup):
            tracker.optimize_cam_in_batch(torch.cat([quad, T], 0), color, depth, 1000, opt)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            tracker.optimize_cam_in_batch(torch.cat([quad, T], 0), color, depth, 1000, opt)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    ms_api = api_loop(tr, iters)
    out = {"metric": "tracking iters/s", "value": 1e3 / ms_engine, "unit": "iters/s", "ms_per_iter": ms_engine,
           "api": {"value": 1e3 / ms_api, "unit": "iters/s", "ms_per_iter": ms_api,
                   "path": "Tracker.optimize_cam_in_batch + torch.optim.Adam + loss.item() per iteration"},
           "config": {"workload": "ScanNet scene0000 shape (460x620, 87.9 MB grids) tracking iteration (configs[2]): 1000 "
                                  "pixels x 48 samples, stage color, tracking loss with the dynamic-object median mask, "
                                  "backward to the rays, pose gradient, Adam(7 dof)", "iters_timed": iters,
                      "fixture": FIXTURE,
                      "engine": "Tracker.optimize_cam_fused / step.TrackingStep: one CUDA graph per iteration, "
                                "pixel draw + candidate rule on the device"},
           "roofline": {"algorithmic_bytes_per_iter": (BYTES_PER_POINT_FWD + BYTES_PER_POINT_REGATHER) * 1000 * S_TOTAL,
                        "algorithmic_flop_per_iter": FLOP_PER_POINT_TRACK * 1000 * S_TOTAL,
                        "tflops": FLOP_PER_POINT_TRACK * 1000 * S_TOTAL / (ms_engine * 1e-3) / 1e12},
           "final_loss": best}
    try:
        from baseline import ref_harness as RH
        if not RH.available():
            raise FileNotFoundError("baseline/_ref is missing")
        slam = RH.reference_scene(ns, dev)
        rt = RH.make_tracker(slam, cfg, dev)
        ms_ref = api_loop(rt, 20)
        out["torch_cuda_baseline"] = {"value": 1e3 / ms_ref, "unit": "iters/s", "ms_per_iter": ms_ref, "kind": "reference",
                                      "sample": "20 iterations of the reference's own Tracker.optimize_cam_in_batch "
                                                "(baseline/_ref, unmodified) on the same GPU, frame and start pose"}
    except Exception as e:                                               # noqa: BLE001
        out["torch_cuda_baseline"] = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    return out


# ---------------------------------------------------------------------------------------
def imap_scene(dev):
    """iMAP* (configs/imap.yaml): one 4 x 256 MLP, no grids, scale 0.1, density compositing, 32 + 12 samples."""
    from nice_slam_annotation_b200 import synth
    from nice_slam_annotation_b200.conv_onet.models.decoder import MLP
    from nice_slam_annotation_b200.scene import load_bound
    from nice_slam_annotation_b200.utils.Renderer import Renderer
    torch.manual_seed(0)
    H, W, fx, fy, cx, cy, bnd = synth.SHAPES[SHAPE]
    scale = 0.1
    bound = load_bound(bnd, scale)
    dec = MLP(name="", dim=3, c_dim=0, color=True, hidden_size=256, skips=[], n_blocks=4,
              pos_embedding_method="fourier").to(dev)
    with torch.no_grad():
        dec.output_linear.bias.add_(torch.tensor([0.3, 0.5, 0.4, 2.0], device=dev))
    dec.bound = bound
    cfg = {"rendering": {"lindisp": False, "perturb": 0.0, "N_samples": IMAP_N_SAMPLES, "N_surface": 0,
                         "N_importance": IMAP_N_IMPORTANCE},
           "scale": scale, "occupancy": False, "model": {"coarse_bound_enlarge": 2}}
    slam = types.SimpleNamespace(nice=False, bound=bound, H=H, W=W, fx=fx, fy=fy, cx=cx, cy=cy)
    r = Renderer(cfg, None, slam)
    room = synth.room_of(bound, margin=0.05)
    c2w = synth.orbit_poses(room, 1, seed=0)[0].to(dev)
    depth, color = synth.make_frame(H, W, fx, fy, cx, cy, room, c2w, zero_frac=0.05, seed=0, device=dev)
    return types.SimpleNamespace(dec=dec, r=r, H=H, W=W, fx=fx, fy=fy, cx=cx, cy=cy, c2w=c2w, depth=depth, color=color)


def imap_step_fn(ctx, sc, n_rays, chunk=65536):
    """One iMAP* mapping iteration body (Mapper.py:598-624 with nice = False): Renderer.render_batch_ray (two passes,
    importance resampling), the mapping loss with w_color 0.05, Renderer.regulation * 5e-4, backward to the MLP.
    Batches above ``chunk`` rays run as micro-batches with gradient accumulation (the saved activations of the
    256-wide MLP are 4.5 KB per point).  Sharded: every rank renders n_rays / world rays, one NCCL all-reduce of the
    222 747 parameter gradients."""
    from nice_slam_annotation_b200 import fused
    from nice_slam_annotation_b200.common import get_samples
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    n_local = max(n_rays // world, 1)
    idx = torch.randint(sc.H * sc.W, (n_local,), generator=torch.Generator().manual_seed(n_rays + rank))
    ro, rd, gd, gc = get_samples(0, sc.H, 0, sc.W, n_local, sc.H, sc.W, sc.fx, sc.fy, sc.cx, sc.cy, sc.c2w, sc.depth,
                                 sc.color, dev, indices=idx)
    ro, rd = ro.contiguous(), rd.contiguous()
    params = list(sc.dec.parameters())

    def step(_):
        for a in range(0, n_local, chunk):
            o, d, g, c = ro[a:a + chunk], rd[a:a + chunk], gd[a:a + chunk], gc[a:a + chunk]
            depth, _, color = sc.r.render_batch_ray({}, sc.dec, d, o, dev, "color", gt_depth=g)
            loss = fused.mapping_loss(depth, color, g, c, 0.05, use_color=True)
            sigma = sc.r.regulation({}, sc.dec, d, o, g, dev, "color")
            (loss + 0.0005 * sigma.abs().sum()).backward()
        if world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            ctx.dist.all_reduce(flat)
        for p in params:
            p.grad = None
    return step, n_local


def imap_bench(ctx, rays=5000):
    sc = imap_scene(ctx.dev)
    step, n_local = imap_step_fn(ctx, sc, rays)
    ms = ctx.timed(step, [None], max(ctx.args.steps // 2, 5), 3, flush=False)
    n_total = n_local * ctx.world
    tf32 = tf32_matmul_peak(ctx.dev)
    peaks, _ = measured_peaks()
    tfl = n_total * IMAP_FLOP_PER_RAY / (ms * 1e-3) / 1e12
    # the MLP alone: forward + backward over the same number of points, no rendering around it
    P = n_local * (IMAP_N_SAMPLES + IMAP_N_IMPORTANCE)
    pts = (torch.rand(P, 3, device=ctx.dev) - 0.5) * 0.5

    def mlp(_):
        sc.dec(pts).sum().backward()
        for q in sc.dec.parameters():
            q.grad = None
    ms_mlp = ctx.timed(mlp, [None], 5, 3, flush=False)
    mlp_tf = 3 * IMAP_FLOP_PER_POINT_FWD * P * ctx.world / (ms_mlp * 1e-3) / 1e12
    peak = tf32 or peaks.get("bf16_tflops_sustained", 1400.0) / 2
    return {"metric": "iMAP* mapping rays/s fwd+bwd", "value": n_total / (ms * 1e-3), "unit": "rays/s", "ms_per_step": ms,
            "rays": n_total, "rays_per_gpu": n_local, "scaling": "strong",
            "config": {"workload": f"iMAP* mapping iteration (configs[4], configs/imap.yaml): {n_total} rays x (32 + 12 "
                                   f"importance) samples through Renderer.render_batch_ray (two passes), density "
                                   f"compositing, mapping loss (w_color 0.05) + 5e-4 x Renderer.regulation (32 jittered "
                                   f"samples per ray), backward to the 4x256 MLP (222 747 parameters); random-init MLP, "
                                   f"Replica room0 camera at scale 0.1",
                       "inputs_larger_than_l2": "activations of one step (4.5 KB per point) exceed L2; no flush"},
            "roofline": {"bound": "tensor", "achieved": tfl, "peak": peak, "unit": "TFLOP/s", "frac": tfl / peak,
                         "peak_source": "dense TF32 torch.matmul 8192^3 measured in this run" if tf32 else "fallback",
                         "algorithmic_flop_per_ray": IMAP_FLOP_PER_RAY,
                         "mlp_fwd_bwd_only": {"points": P * ctx.world, "ms": ms_mlp, "tflops": mlp_tf, "frac": mlp_tf / peak},
                         "frac_of_bf16_sustained": tfl / peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"]),
                         "note": "algorithmic fp32 FLOPs; k_gemm3x issues 3x as many tensor-core FLOPs (3xTF32)"}}


# ---------------------------------------------------------------------------------------
def engine_figure(ctx, ns, rays_total, steps, warmup, seed, kernel_times=True):
    """Engine step on ``rays_total`` rays split over the ranks (strong scaling): ms per step (max over ranks),
    rays/s over all ranks, per-kernel durations of rank 0."""
    from nice_slam_annotation_b200.step import FusedStep
    per = max(rays_total // ctx.world, FRAMES)
    dec = ns.shared_decoders
    train_only_color_decoder(dec)
    grids = grids_of(ns)
    step = FusedStep(ns.renderer, dec, grids, STAGE, per, peers=ctx.peers, **STEP_KW)
    nb = 4 if per <= 100000 else 2
    batches = make_batches(ns, per, nb, seed=seed + ctx.rank)
    kept = sum(b[0].shape[0] for b in batches) / len(batches)

    def fn(b):
        step.load(*b)
        return step.run()
    km = [None, None, None]
    for i in range(2):
        fn(batches[i % nb])
    torch.cuda.synchronize()
    if kernel_times:
        ctx.barrier()
        ctx.L.nsb_timing(1)
        for i in range(2):
            fn(batches[i % nb])
        torch.cuda.synchronize()
        km = [ctx.kernel_ms(w) for w in range(3)]
        km = [m * max(n, 1) / 2 for m, n in km]                     # ms per STEP (a step may run the kernel in chunks)
        ctx.L.nsb_timing(0)
    step.capture()
    ms = ctx.timed(fn, batches, steps, warmup)
    total = ctx.sum_over_ranks(kept)
    pts = per * S_TOTAL
    row = {"rays": rays_total, "rays_per_gpu": per, "kept_all_gpus": total, "ms_per_step": ms,
           "rays_per_s": total / (ms * 1e-3), "fwd_ms": km[0], "bwd_ms": km[1], "dweight_ms": km[2]}
    if kernel_times:
        row["fwd_request_gbs"] = BYTES_PER_POINT_FWD * pts / (km[0] * 1e-3) / 1e9
        row["bwd_request_gbs"] = BYTES_PER_POINT_BWD * pts / (km[1] * 1e-3) / 1e9
        row["step_tflops"] = FLOP_PER_POINT_FWD_BWD * total * S_TOTAL / (ms * 1e-3) / 1e12
    if ctx.peers is not None:
        ctx.peers.check_error()
        ctx.alive.append(step)
    del step
    return row


def apartment_bench(ctx):
    """configs[3]: Apartment-scale bound (124.9 MB of grids, beyond L2 together with the gradient bucket), 5000 rays
    per iteration (configs/Apartment/apartment.yaml:21-24), the rays sharded over the ranks: STRONG scaling."""
    from nice_slam_annotation_b200.scene import build_scene
    ns = desaturate(build_scene("apartment", ctx.dev, seed=0, n_frames=FRAMES))
    row = engine_figure(ctx, ns, APARTMENT_RAYS, ctx.args.steps, ctx.args.warmup, seed=300)
    peaks, _ = measured_peaks()
    row.update({"metric": "mapping rays/s fwd+bwd at Apartment shape", "value": row["rays_per_s"], "unit": "rays/s",
                "scaling": "strong", "n_gpus": ctx.world,
                "grids_mb": sum(v.numel() * 4 for v in ns.shared_c.values()) / 1e6,
                "config": {"workload": f"Apartment shape (720x1280, bound [[-5.8,11.3],[-4,4.5],[-7.9,4.9]]) mapping step "
                                       f"(configs[3]): {APARTMENT_RAYS} rays from {FRAMES} frames x 48 samples, stage color, "
                                       f"fwd + loss + bwd to 3 grids and the colour decoder, rays split over {ctx.world} GPU(s)",
                           "fixture": FIXTURE, "l2": "256 MB buffer written between timed steps"}})
    if row.get("fwd_request_gbs"):
        row["fwd_request_frac_of_hbm"] = row["fwd_request_gbs"] / peaks["hbm_gbs"]
        row["bwd_request_frac_of_hbm"] = row["bwd_request_gbs"] / peaks["hbm_gbs"]
    del ns
    torch.cuda.empty_cache()
    return row


def sweep_bench(ctx, ns):
    """configs[4]: ray-count sweep 1K-1M on the headline scene (NICE, stage color) and at the iMAP* configuration;
    the rays of every point are split over the ranks (strong scaling)."""
    rows, irows = [], []
    for n in SWEEP_RAYS:
        steps = max(3, min(ctx.args.steps, int(2e6 // n)))
        try:
            rows.append(engine_figure(ctx, ns, n, steps, 3, seed=400))
        except Exception as e:                                           # noqa: BLE001
            rows.append({"rays": n, "unavailable": f"{type(e).__name__}: {e}"[:200]})
        torch.cuda.empty_cache()
    sc = imap_scene(ctx.dev)
    for n in SWEEP_RAYS:
        try:
            step, n_local = imap_step_fn(ctx, sc, n)
            ms = ctx.timed(step, [None], max(2, min(10, int(3e5 // n))), 2, flush=False)
            tot = n_local * ctx.world
            irows.append({"rays": tot, "rays_per_gpu": n_local, "ms_per_step": ms, "rays_per_s": tot / (ms * 1e-3),
                          "tflops": tot * IMAP_FLOP_PER_RAY / (ms * 1e-3) / 1e12})
        except Exception as e:                                           # noqa: BLE001
            irows.append({"rays": n, "unavailable": f"{type(e).__name__}: {e}"[:200]})
        torch.cuda.empty_cache()
    return {"what": f"engine step vs ray count on {ctx.world} GPU(s), rays split over the ranks (strong scaling), graph "
                    f"replay, L2 flushed between steps; kernel times are per step on rank 0; above 4 194 304 points per call "
                    f"the backward runs in chunks (weight-gradient scratch)", "nice_stage_color": rows, "imap": irows}


def sharded_parity(ctx, ns):
    """Driver-side evidence for the ray-sharded path: ONE global batch, every rank renders its slice
    through the sharded engine step (peer exchange or NCCL) -> summed gradients; the same batch through an unsharded
    step on every rank.  max|a - b| / max|b| over the whole gradient bucket, max over ranks."""
    from nice_slam_annotation_b200.step import FusedStep, shard_slice
    dec = ns.shared_decoders
    grids = grids_of(ns)
    n = ctx.args.rays
    batch = make_batches(ns, n, 1, seed=999)[0]                        # same seed on every rank: the same global batch
    k = batch[0].shape[0]
    sl = shard_slice(k, ctx.world, ctx.rank)
    sharded = FusedStep(ns.renderer, dec, grids, STAGE, n, peers=ctx.peers, **STEP_KW)
    single = FusedStep(ns.renderer, dec, grids, STAGE, n, sharded=False, **STEP_KW)
    sharded.load(*(x[sl].contiguous() for x in batch))
    ls = sharded.run().clone()
    single.load(*batch)
    l1 = single.run().clone()
    torch.cuda.synchronize()
    if ctx.world > 1:
        ctx.dist.all_reduce(ls)
    ref = single.bucket.flat
    err = float((sharded.bucket.flat - ref).abs().max() / ref.abs().max())
    err = ctx.max_over_ranks(err)
    if ctx.peers is not None:
        ctx.peers.check_error()
        ctx.alive.append(sharded)
    return {"sharded_vs_single_max_rel": err, "loss_rel_diff": abs(float(ls) - float(l1)) / abs(float(l1)),
            "rays": k, "what": "gradient bucket (3 grids + colour decoder) of one global batch: rays sharded over the ranks + "
                               "exchange vs the whole batch on one GPU; fp32 atomics reorder the sums"}


# ---------------------------------------------------------------------------------------
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
    fn = lambda: render_mesh_depth(v, t, c2ws, 500, 500, 300.0, 300.0, 249.5, 249.5, device=dev)    # noqa: E731
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
    autograd = the same call with ``fused_loop = False`` (Renderer + autograd + torch.optim.Adam),
    torch_cuda_baseline = the reference's own Mapper.optimize_map (baseline/_ref) on the same GPU and state."""
    from nice_slam_annotation_b200.Mapper import Mapper
    out = {}
    state = {k: v.detach().clone() for k, v in ns.shared_c.items()}
    dec_state = {k: v.detach().clone() for k, v in ns.shared_decoders.state_dict().items()}

    def restore():
        for k, v in state.items():
            ns.shared_c[k] = v.clone().contiguous(memory_format=torch.channels_last_3d)
        ns.shared_decoders.load_state_dict(dec_state)

    def keyframes():
        return [{"gt_c2w": c2w, "idx": i, "color": color, "depth": depth, "est_c2w": c2w}
                for i, (c2w, depth, color) in enumerate(ns.frames[:-1])]
    c2w, depth, color = ns.frames[-1]
    for name, fused_loop in (("loop", True), ("autograd", False)):
        restore()
        mp = Mapper(ns.cfg, None, ns)
        mp.fused_loop = fused_loop
        kf = keyframes()
        mp.keyframe_dict = kf
        n_it = iters if fused_loop else max(iters // 3, 6)
        mp.optimize_map(n_it, 1.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)     # warm-up call
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mp.optimize_map(n_it, 1.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out[name] = {"ms_per_call": dt * 1e3, "iters": n_it, "ms_per_iter": dt * 1e3 / n_it,
                     "rays_per_s": ns.cfg["mapping"]["pixels"] * n_it / dt}
    try:
        from baseline import ref_harness as RH
        if not RH.available():
            raise FileNotFoundError("baseline/_ref is missing")
        restore()
        slam = RH.reference_scene(ns, dev)
        rm = RH.make_mapper(slam, ns.cfg, dev, frustum=True)
        kf = keyframes()
        rm.keyframe_dict = kf
        n_it = 6
        rm.optimize_map(n_it, 1.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rm.optimize_map(n_it, 1.0, len(kf), color, depth, c2w, kf, list(range(len(kf))), c2w)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out["torch_cuda_baseline"] = {"ms_per_call": dt * 1e3, "iters": n_it, "ms_per_iter": dt * 1e3 / n_it,
                                      "rays_per_s": ns.cfg["mapping"]["pixels"] * n_it / dt, "kind": "reference",
                                      "sample": "the reference's own Mapper.optimize_map (baseline/_ref, unmodified), same "
                                                "frames, frustum feature selection on"}
    except Exception as e:                                               # noqa: BLE001
        out["torch_cuda_baseline"] = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    restore()
    out["config"] = ("Mapper.optimize_map, Replica room0 shape, 4 keyframes + current frame, 1000 pixels/iteration, "
                     "frustum feature selection on, stage schedule middle/fine/color, optimiser included; wall clock "
                     "around the call incl. keyframe selection and frustum masks")
    return out


# ---------------------------------------------------------------------------------------
def run_ours(args):
    from nice_slam_annotation_b200.scene import build_scene
    ctx = Ctx(args)
    rank, world, dev = ctx.rank, ctx.world, ctx.dev
    ns = desaturate(build_scene(SHAPE, dev, seed=0, n_frames=FRAMES))
    out, step, batches, x_err = headline(ctx, ns)
    failed = []
    if x_err is not None and not x_err <= 1e-5:
        failed.append(f"peer exchange differs from the NCCL all-reduce: max rel {x_err:.3e} > 1e-5")

    def section(name, fn, *a, only_rank0=False, **k):
        """Sections run on every rank (they contain collectives); a failure is recorded, not fatal."""
        if args.only and name not in args.only.split(","):
            return
        try:
            out[name] = fn(*a, **k)
        except Exception as e:                                           # noqa: BLE001
            import traceback
            traceback.print_exc(file=sys.stderr)
            out[name] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.empty_cache()

    if world > 1:
        section("sharded_parity", sharded_parity, ctx, ns)
        sp = out.get("sharded_parity", {}).get("sharded_vs_single_max_rel")
        if sp is not None and not sp <= 1e-4:
            failed.append(f"sharded step differs from the single-GPU step: max rel {sp:.3e} > 1e-4")
    if not args.quick:
        section("apartment", apartment_bench, ctx)
        section("imap", imap_bench, ctx)
        section("sweep", sweep_bench, ctx, ns)
    if world == 1 and not args.quick:
        section("fixtures", fixture_figures, ctx, ns, out["value"])
        if not args.no_tracking:
            section("tracking", tracking_bench, dev, iters=max(args.steps, 20) * 5, warmup=args.warmup)
        if not args.no_cpu_baseline:
            section("torch_cuda_baseline", torch_cuda_baseline, ns, batches, dev)
            section("cpu_baseline", cpu_baseline, ns, batches, budget_s=12.0)
        if not args.no_tracking:
            section("bulk_forward", bulk_forward_bench, ns, dev)
            section("mesh_depth", mesh_depth_bench, dev)
            section("optimize_map", optimize_map_bench, ns, dev)              # last: it trains the scene
    elif world == 1 and not args.no_cpu_baseline:
        section("cpu_baseline", cpu_baseline, ns, batches, budget_s=12.0)
    if failed:
        out["parity_failed"] = failed
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        ctx.dist.destroy_process_group()
    if failed:
        sys.exit(3)


def run_reference(args):
    """--impl reference: the reference's own CPU implementation (baseline/_ref) on the host cores, rank 0 only."""
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
    ns = desaturate(build_scene(SHAPE, "cpu", seed=0, n_frames=FRAMES))    # host only: no kernel of ours on this arm
    batches = make_batches_cpu(ns, args.rays, 4, seed=100)
    step, kind, what = reference_step_fn(ns, batches, "cpu")
    for i in range(args.warmup):
        step(i)
    t0, rays = time.perf_counter(), 0
    for i in range(args.steps):
        rays += step(args.warmup + i)[1]
    dt = time.perf_counter() - t0
    v = rays / dt
    out = {"impl": "reference", "metric": "mapping rays/s fwd+bwd at Replica shape", "value": v, "unit": "rays/s",
           "n_gpus": int(os.environ.get("WORLD_SIZE", 1)), "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": workload_config(args.rays),
           "cpu_baseline": {"value": v, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": kind,
                            "sample": f"{args.steps} full steps ({dt:.1f} s), {what} on torch CPU"},
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
    ap.add_argument("--no-tracking", action="store_true", help="skip the tracking / bulk / optimize_map sections")
    ap.add_argument("--quick", action="store_true", help="headline only (no apartment / imap / sweep / fixtures sections)")
    ap.add_argument("--only", default="", help="comma-separated section names to run besides the headline")
    ap.add_argument("--no-graph", action="store_true", help="launch the engine step kernel by kernel")
    ap.add_argument("--nccl-exchange", action="store_true",
                    help="N > 1: all-reduce the gradient bucket with NCCL instead of the peer-memory kernel")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
