"""Peer-memory gradient exchange on N GPUs of one box: correctness against NCCL + timings.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 scripts/exchange_check.py [--json out.json]

Checks (every rank; rank 0 prints): nsb_bucket_allreduce (multicast and peer load/store paths) ==
dist.all_reduce on ragged sizes; nsb_bucket_reduce_adam == nsb_adam on the all-reduced gradient with
identical parameters on every rank; replay inside a CUDA graph.  Then times the Replica-size bucket
(48.5 MB) through NCCL and through the kernels.
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nice_slam_annotation_b200._lib import check, lib          # noqa: E402
from nice_slam_annotation_b200.peers import PeerGroup           # noqa: E402


def ev_time(fn, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=None)
    ap.add_argument("--mb", type=float, default=48.480768)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    dist.init_process_group("nccl", device_id=dev)
    L = lib()
    s = lambda: torch.cuda.current_stream().cuda_stream
    pg = PeerGroup(device=dev)
    out = {"world": world, "multicast": pg.multicast, "checks": {}}
    paths = [("multicast", True)] if pg.multicast else []
    paths.append(("peer_ld_st", False))

    # ---- all-reduce vs NCCL on ragged sizes --------------------------------------------------
    sizes = [64, 4 * 1001, 65536 + 4 * 3, 1 << 20]
    n_max = max(sizes)
    buf = pg.alloc(n_max)
    for name, mc in paths:
        px = pg.struct(buf, use_multicast=mc)
        worst = 0.0
        for n in sizes:
            g = torch.Generator(device=dev).manual_seed(1000 * rank + n)
            x = torch.randn(n, device=dev, generator=g)
            ref = x.clone()
            dist.all_reduce(ref)
            buf[:n].copy_(x)
            buf[n:].fill_(7.0)                          # outside the range: must stay untouched
            torch.cuda.synchronize()
            dist.barrier()
            check(L.nsb_bucket_allreduce(C.byref(px), 0, n, s()), "allreduce")
            torch.cuda.synchronize()
            dist.barrier()
            err = (buf[:n] - ref).abs().max().item() / ref.abs().max().item()
            worst = max(worst, err)
            assert err < 1e-6, (name, n, err)
            assert bool((buf[n:] == 7.0).all()), (name, n, "wrote outside the range")
        # sub-range with an offset
        buf.fill_(float(rank + 1))
        torch.cuda.synchronize(); dist.barrier()
        check(L.nsb_bucket_allreduce(C.byref(px), 128, 4 * 999, s()), "allreduce")
        torch.cuda.synchronize(); dist.barrier()
        tot = float(sum(range(1, world + 1)))
        assert bool((buf[128:128 + 4 * 999] == tot).all()) and bool((buf[:128] == rank + 1).all()) \
            and bool((buf[128 + 4 * 999:] == rank + 1).all()), name
        out["checks"]["allreduce_" + name] = {"max_rel_err": worst}
    pg.check_error()

    # ---- sparse all-reduce (touched 128-byte lines only) vs NCCL --------------------------------
    n_sp = 32 * 40013 + 4 * 5                            # ragged: partial last line, partial last unit
    sp = pg.alloc(n_sp)
    tw = pg.alloc_touched(n_sp)
    for name, mc in paths:
        px = pg.struct(sp, use_multicast=mc, touched=tw)
        for frac in (0.0, 0.05, 0.4, 1.0):
            g = torch.Generator(device=dev).manual_seed(31 * rank + int(frac * 100))
            lines = (torch.rand((n_sp + 31) // 32, device=dev, generator=g) < frac)
            x = torch.randn(n_sp, device=dev, generator=g) * lines.repeat_interleave(32)[:n_sp]
            ref = x.clone()
            dist.all_reduce(ref)
            for rep in range(2):                          # twice: the scratch words are reused
                sp.copy_(x)
                torch.cuda.synchronize(); dist.barrier()
                check(L.nsb_bucket_allreduce_sparse(C.byref(px), 0, n_sp, s()), "allreduce_sparse")
                torch.cuda.synchronize(); dist.barrier()
                err = (sp - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30)
                assert err < 1e-6, (name, frac, rep, err)
        # a sub-range at a line-aligned offset leaves the rest alone
        sp.fill_(float(rank + 1))
        torch.cuda.synchronize(); dist.barrier()
        check(L.nsb_bucket_allreduce_sparse(C.byref(px), 64, 4 * 2999, s()), "allreduce_sparse")
        torch.cuda.synchronize(); dist.barrier()
        tot = float(sum(range(1, world + 1)))
        assert bool((sp[64:64 + 4 * 2999] == tot).all()) and bool((sp[:64] == rank + 1).all()) \
            and bool((sp[64 + 4 * 2999:] == rank + 1).all()), name
        out["checks"]["allreduce_sparse_" + name] = {"max_rel_err": err}
    pg.check_error()

    # ---- fused reduce + Adam vs all-reduce + nsb_adam -----------------------------------------
    n = 32 * 4099                                        # 4099 voxel lines: ragged over the ranks
    grad, param = pg.alloc(n), pg.alloc(n)
    gen = torch.Generator(device=dev).manual_seed(5)
    p0 = torch.randn(n, device=dev, generator=gen)
    mask = (torch.rand(n // 32, device=dev, generator=gen) < 0.7).to(torch.uint8)
    lo = C.c_int64()
    n_own = L.nsb_exchange_slice(rank, world, n, C.byref(lo))
    for name, mc in paths:
        px = pg.struct(grad, param, use_multicast=mc)
        for use_mask in (False, True):
            mk = mask if use_mask else None
            param.copy_(p0)
            p_ref = p0.clone()
            m_ref, v_ref = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
            st_ref = torch.zeros(2, dtype=torch.int32, device=dev)
            m_own, v_own = torch.zeros(max(n_own, 4), device=dev), torch.zeros(max(n_own, 4), device=dev)
            st = torch.zeros(2, dtype=torch.int32, device=dev)
            lr = torch.full((1,), 1e-2, device=dev)
            for it in range(3):
                g = torch.Generator(device=dev).manual_seed(77 * rank + it)
                x = torch.randn(n, device=dev, generator=g)
                ref = x.clone()
                dist.all_reduce(ref)
                grad.copy_(x)
                torch.cuda.synchronize(); dist.barrier()
                check(L.nsb_bucket_reduce_adam(C.byref(px), 0, n, 0, m_own.data_ptr(), v_own.data_ptr(),
                                               mk.data_ptr() if mk is not None else None, st.data_ptr(), lr.data_ptr(),
                                               0.9, 0.999, 1e-8, s()), "reduce_adam")
                check(L.nsb_adam(p_ref.data_ptr(), ref.data_ptr(), m_ref.data_ptr(), v_ref.data_ptr(), n,
                                 mk.data_ptr() if mk is not None else None, st_ref.data_ptr(), lr.data_ptr(),
                                 0.9, 0.999, 1e-8, s()), "adam")
                torch.cuda.synchronize(); dist.barrier()
            err = (param - p_ref).abs().max().item()
            assert err < 1e-5, (name, use_mask, err)
            assert int(st[0]) == 3 and int(st[1]) == 0
            same = param.clone()
            dist.broadcast(same, 0)
            assert bool((same == param).all()), "parameters differ between ranks"
            out["checks"][f"reduce_adam_{name}_{'masked' if use_mask else 'dense'}"] = {"max_abs_err": err}
    pg.check_error()

    # ---- graph replay -------------------------------------------------------------------------
    px = pg.struct(buf)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        buf.fill_(1.0)
        check(L.nsb_bucket_allreduce(C.byref(px), 0, n_max, s()), "allreduce")          # warm-up
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            buf.fill_(1.0)
            check(L.nsb_bucket_allreduce(C.byref(px), 0, n_max, s()), "allreduce")
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
    assert bool((buf == float(world)).all())
    out["checks"]["graph_replay"] = True
    pg.check_error()

    # ---- timings at the Replica bucket size ----------------------------------------------------
    nb = int(a.mb * 1e6 / 4) // 64 * 64
    big = pg.alloc(nb)
    pbig = pg.alloc(nb)
    plain = torch.zeros(nb, device=dev)
    t = {"bucket_mb": nb * 4 / 1e6, "nccl_allreduce_ms": ev_time(lambda: dist.all_reduce(plain))}
    for name, mc in paths:
        px = pg.struct(big, pbig, use_multicast=mc)
        t[f"exchange_{name}_ms"] = ev_time(lambda: check(L.nsb_bucket_allreduce(C.byref(px), 0, nb, s())))
        n_own = L.nsb_exchange_slice(rank, world, nb, None)
        m_own, v_own = torch.zeros(n_own, device=dev), torch.zeros(n_own, device=dev)
        st = torch.zeros(2, dtype=torch.int32, device=dev)
        lr = torch.full((1,), 1e-3, device=dev)
        t[f"reduce_adam_{name}_ms"] = ev_time(lambda: check(L.nsb_bucket_reduce_adam(
            C.byref(px), 0, nb, 0, m_own.data_ptr(), v_own.data_ptr(), None, st.data_ptr(), lr.data_ptr(),
            0.9, 0.999, 1e-8, s())))
    # sparse exchange at the touched fractions of a mapping step (scripts/bucket_sparsity.py: 16 % of the
    # lines per 1000 rays; the union over ranks grows towards 39 % at 8000 rays)
    twb = pg.alloc_touched(nb)
    for frac in (0.16, 0.39, 1.0):
        g = torch.Generator(device=dev).manual_seed(7 + rank)
        keep_l = (torch.rand(nb // 32, device=dev, generator=g) < (frac / world if frac < 1 else 1.0) * 1.0)
        src = torch.randn(nb, device=dev, generator=g) * keep_l.repeat_interleave(32)
        for name, mc in paths:
            px = pg.struct(big, use_multicast=mc, touched=twb)

            def run():
                big.copy_(src)
                check(L.nsb_bucket_allreduce_sparse(C.byref(px), 0, nb, s()))
            t_copy = ev_time(lambda: big.copy_(src))
            t[f"sparse_{name}_union{int(frac * 100)}pct_ms"] = ev_time(run) - t_copy
    # what the fused kernel replaces: all-reduce, then Adam over the whole bucket on every rank
    m_f, v_f, p_f = torch.zeros(nb, device=dev), torch.zeros(nb, device=dev), torch.zeros(nb, device=dev)
    st = torch.zeros(2, dtype=torch.int32, device=dev)
    lr = torch.full((1,), 1e-3, device=dev)
    t["adam_full_ms"] = ev_time(lambda: check(L.nsb_adam(p_f.data_ptr(), plain.data_ptr(), m_f.data_ptr(), v_f.data_ptr(), nb,
                                                         None, st.data_ptr(), lr.data_ptr(), 0.9, 0.999, 1e-8, s())))
    small = 16000 // 64 * 64
    t["small_floats"] = small
    t["nccl_allreduce_small_ms"] = ev_time(lambda: dist.all_reduce(plain[:small]))
    px = pg.struct(big)
    t["exchange_small_ms"] = ev_time(lambda: check(L.nsb_bucket_allreduce(C.byref(px), 0, small, s())))
    pg.check_error()
    out["timing"] = t
    if rank == 0:
        print(json.dumps(out))
        if a.json:
            with open(a.json, "w") as fh:
                json.dump(out, fh, indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
