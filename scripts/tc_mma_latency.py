"""Prints the tcgen05 MMA issue / completion cycles measured by nsb_tc_mma_latency (B200 only)."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nice_slam_annotation_b200._lib import check, lib

out = torch.zeros(2, dtype=torch.int64, device="cuda")
print("N count nacc issue_cycles total_cycles cycles_per_mma")
for N in (16, 32, 64, 160, 256):
    for count in (12, 48, 96):
        for nacc in (1, 2, 3, 6):
            if 16 + nacc * N > 512:
                continue
            for _ in range(2):
                check(lib().nsb_tc_mma_latency(C.c_void_p(out.data_ptr()), N, count, nacc, None), "lat")
                torch.cuda.synchronize()
            a, b = out.tolist()
            print(N, count, nacc, a, b, round(b / count, 1))
