#!/usr/bin/env python
"""Hot source lines of one kernel of an ncu report (needs -lineinfo + --import-source on).

    python scripts/ncu_hot_lines.py gpurun_out/prof.ncu-rep k_decoders_backward [top] [launch index among the matches]
"""
import csv
import io
import subprocess
import sys


def main(path, kernel, top=40, launch=None):
    sel = [] if launch is None else ["--launch-skip", str(launch), "--launch-count", "1"]
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", kernel, *sel], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname, hdr, data, seen_fn = "?", None, [], 0
    for r in rows:
        if not r:
            continue
        if r[0] in ("File Path", "File Name"):
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or not r[0].isdigit():
            continue
        d = dict(zip(hdr[4:], r[4:]))
        try:
            data.append((int(d["# Samples"]), fname, int(r[0]), r[1].strip()[:100], int(d["Instructions Executed"]),
                         {k[6:]: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v) > 0}))
        except (KeyError, ValueError):
            pass
    tot = sum(x[0] for x in data) or 1
    print(f"# {kernel}: {tot} stall samples; top {top} source lines")
    for s, f, ln, src, ie, st in sorted(data, key=lambda x: x[:3], reverse=True)[:top]:
        top3 = ",".join(f"{k}:{v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{100 * s / tot:5.1f}% {f}:{ln:<5d} inst={ie:<9d} [{top3}]  {src}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40, int(sys.argv[4]) if len(sys.argv) > 4 else None)
