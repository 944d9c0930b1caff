#!/usr/bin/env python
"""Counts of the SASS mnemonics that show which hardware paths the built library uses, per kernel
(/opt/skills/guides/B200_PROFILING.md "What proves a Blackwell-native kernel").  Runs without a GPU.

    python scripts/sass_summary.py [nice_slam_annotation_b200/csrc/libnsb200.so] > profiles/rNN_sass_summary.txt

UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit, UTMALDG / UTMASTG / UBLKCP = TMA / bulk
copies, HMMA = legacy mma.sync, LDGSTS = cp.async, REDG / ATOMG = global reductions / atomics, LDGMC / STGMC =
multimem.ld_reduce / multimem.st (NVLS).  The library embeds SASS only (code=sm_100a), so there is no PTX to count."""
import collections
import os
import re
import subprocess
import sys

LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "nice_slam_annotation_b200", "csrc", "libnsb200.so")
PATTERNS = [("UTC*MMA (tcgen05.mma)", r"\bUTC[A-Z]*MMA"), ("LDTM (tcgen05.ld)", r"\bLDTM"), ("STTM (tcgen05.st)", r"\bSTTM"),
            ("UTCBAR (tcgen05.commit)", r"\bUTCBAR"), ("UTMALDG/UTMASTG/UBLKCP (TMA, bulk copy)", r"\b(UTMALDG|UTMASTG|UBLKCP)"),
            ("SYNCS (mbarrier)", r"\bSYNCS"), ("HMMA (mma.sync, legacy)", r"\bHMMA"), ("LDGSTS (cp.async)", r"\bLDGSTS"),
            ("REDG (red.global, incl. .v4)", r"\bREDG"), ("ATOMG (global atomic)", r"\bATOMG"), ("ATOMS (shared atomic)", r"\bATOMS"),
            ("LDGMC / STGMC / REDGMC (multimem.*, NVLS)", r"\b(LDGMC|STGMC|REDGMC)"),
            ("STG.E.ENL2.256 / 256-bit stores", r"\bSTG\.[A-Z0-9.]*256")]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    per, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name.replace("(anonymous namespace)::", "")                 # k_gemm3x<mode> lives in one
            cur = per.setdefault(re.sub(r"\(.*", "", name), collections.Counter())
            continue
        if cur is None:
            continue
        for label, pat in PATTERNS:
            if re.search(pat, line):
                cur[label] += 1
    total = collections.Counter()
    for c in per.values():
        total.update(c)
    print(f"# {os.path.relpath(LIB)}: SASS mnemonic counts (cuobjdump -sass), sm_100a")
    print("\n## whole library")
    for label, _ in PATTERNS:
        print(f"  {label:45s} {total[label]:6d}")
    print("\n## per kernel (non-zero counts only)")
    for name, c in per.items():
        if sum(c.values()):
            print(f"  {name}: " + ", ".join(f"{k.split(' ')[0]} {v}" for k, v in c.items() if v))


if __name__ == "__main__":
    main()
