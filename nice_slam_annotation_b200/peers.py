"""Symmetric (peer-mapped) device memory for the sharded mapping step.

torch.distributed._symmetric_memory is used for what it is: plumbing.  It allocates a buffer of the
same size on every rank of one box, exchanges the handles, and maps every rank's copy -- and, on an
NVSwitch fabric with NVLS, one multicast alias of all copies -- into this process.  The raw
addresses go into an ``NsbPeers`` struct; the reduction itself is csrc/nsb_exchange.cu
(``nsb_bucket_allreduce``, ``nsb_bucket_reduce_adam``).
"""
import ctypes as C

import torch

from ._lib import NSB_MAX_PEERS, NsbPeers, check, lib


def _group_name(group):
    import torch.distributed as dist
    g = group if group is not None else dist.group.WORLD
    return g.group_name


class PeerGroup:
    """The ranks of one box that exchange gradients through peer memory.

    ``alloc(n)`` returns a flat fp32 tensor of n elements in symmetric memory (collective: every
    rank must call it with the same n, in the same order) and remembers its peer mappings.
    """

    def __init__(self, group=None, device=None):
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("PeerGroup needs an initialised torch.distributed process group")
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > NSB_MAX_PEERS:
            raise RuntimeError(f"peer exchange supports up to {NSB_MAX_PEERS} ranks of one box, got {self.world}")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self._handles = {}
        words = int(lib().nsb_exchange_signal_words())
        self.signal = self._symmetric(words, torch.int32)
        self.scalar = self._symmetric(64, torch.float32)
        self.error = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.multicast = bool(self._handles[self.signal.data_ptr()].multicast_ptr)

    def _symmetric(self, n, dtype):
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(n, dtype=dtype, device=self.device)
        h = symm.rendezvous(t, _group_name(self.group))
        t.zero_()
        torch.cuda.synchronize(self.device)
        h.barrier()                      # every copy is zero before any rank uses it
        self._handles[t.data_ptr()] = h
        return t

    def alloc(self, n):
        return self._symmetric(int(n), torch.float32)

    def _ptrs(self, t):
        """(peer pointers of the allocation ``t`` is a view of, multicast pointer or 0), shifted to t."""
        for base, h in self._handles.items():
            off = t.data_ptr() - base
            if 0 <= off < h.buffer_size:
                return [p + off for p in h.buffer_ptrs], (h.multicast_ptr + off if h.multicast_ptr else 0)
        raise RuntimeError("tensor does not live in this group's symmetric memory")

    def alloc_touched(self, n):
        """Symmetric scratch of the sparse exchange for a bucket of n floats (one bit per 128-byte line)."""
        return self._symmetric(int(lib().nsb_exchange_touched_words(int(n))), torch.int32)

    def struct(self, bucket, param=None, use_multicast=None, touched=None):
        """NsbPeers for a gradient bucket (and optionally the parameter buffer the fused optimiser
        writes), both flat tensors from ``alloc``.  use_multicast=None picks the faster path: with two
        ranks the in-switch reduction moves MORE bytes per link than plain peer loads / stores (both
        copies of a slice travel to the switch) and measures slower on B200 (48.5 MB: 0.155 ms vs
        0.102 ms, NCCL 0.114 ms); from four ranks on it moves fewer."""
        if use_multicast is None:
            # the sparse exchange is request-bound, not byte-bound: one multimem.ld_reduce + one multimem.st per
            # touched line beats world loads + world stores at every rank count (2 ranks, 16 % / 39 % of the
            # lines touched: 0.055 / 0.082 ms vs 0.095 / 0.100 ms)
            use_multicast = self.world > 2 or (touched is not None and self.multicast)
        s = NsbPeers()
        s.rank, s.world = self.rank, self.world
        bp, bmc = self._ptrs(bucket)
        sp, _ = self._ptrs(self.signal)
        for r in range(self.world):
            s.bucket[r], s.signal[r] = bp[r], sp[r]
        s.bucket_multicast = bmc if use_multicast else 0
        if param is not None:
            pp, pmc = self._ptrs(param)
            for r in range(self.world):
                s.param[r] = pp[r]
            s.param_multicast = pmc if use_multicast else 0
        cp, _ = self._ptrs(self.scalar)
        for r in range(self.world):
            s.scalar[r] = cp[r]
        if touched is not None:
            tp, _ = self._ptrs(touched)
            for r in range(self.world):
                s.touched[r] = tp[r]
        s.error = self.error.data_ptr()
        return s

    def check_error(self):
        if int(self.error.item()):
            raise RuntimeError("peer exchange: a rank did not reach the barrier (timeout)")


def allreduce(peers_struct, offset, n, stream):
    check(lib().nsb_bucket_allreduce(C.byref(peers_struct), int(offset), int(n), stream), "bucket_allreduce")
