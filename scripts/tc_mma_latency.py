"""Prints the tcgen05 MMA issue / completion cycles measured by nsb_tc_mma_latency (B200 only).
mode 0: A in tensor memory, B K-major; 1: A and B K-major in shared memory; 2: both MN-major; 3: A K-major, B MN-major;
4: A MN-major, B K-major."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nice_slam_annotation_b200._lib import check, lib

out = torch.zeros(2, dtype=torch.int64, device="cuda")
print("mode N count nacc issue_cycles total_cycles cycles_per_mma")
for mode in (0, 1, 2, 3, 4):
    for N in (16, 32, 64, 128, 160, 256):
        if mode >= 2 and N % 32:
            continue
        for count in (12, 96):
            for nacc in (1, 2):
                if 16 + nacc * N > 512:
                    continue
                for _ in range(2):
                    check(lib().nsb_tc_mma_latency(C.c_void_p(out.data_ptr()), N, count, nacc, mode, None), "lat")
                    torch.cuda.synchronize()
                a, b = out.tolist()
                print(mode, N, count, nacc, a, b, round(b / count, 1))
