#!/usr/bin/env python
"""Summarise an ncu report (``ncu --set full``) into the text committed under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_<what>.txt

Reads the report with ``ncu -i ... --page raw --csv`` (works without a GPU) and prints, per
profiled launch, the metrics the roofline discussion in DESIGN.md uses."""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum",
    "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
    "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "smsp__inst_executed.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {path}: {len(rows) - 2} profiled launch(es); ncu --set full --clock-control none")
    for r in rows[2:]:
        print(f"\n== {r[hdr.index('Kernel Name')]}  (launch id {r[0]})")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:90s} {r[i]:>16s} {units[i]}")


def traffic_json(path, source):
    """profiles/ncu_traffic.json (read by bench.py's roofline block): per kernel, DRAM bytes and pipe utilisation of
    its last profiled launch."""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]

    def val(r, k, scale_bytes=False):
        i = hdr.index(k)
        v = float(r[i].replace(",", ""))
        if scale_bytes:
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
        return v
    res = {"source": source}
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0].split("<")[0]
        t = val(r, "gpu__time_duration.sum")
        t *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(units[hdr.index("gpu__time_duration.sum")], 1.0)
        res[name] = {"dram_bytes_read": int(val(r, "dram__bytes_read.sum", True)),
                     "dram_bytes_write": int(val(r, "dram__bytes_write.sum", True)),
                     "sm__pipe_tensor_cycles_active_pct": val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                     "issue_active_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                     "lts_hit_rate_pct": val(r, "lts__t_sector_hit_rate.pct"),
                     "gpu_time_us": t}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "--traffic-json":
        traffic_json(sys.argv[1], sys.argv[3] if len(sys.argv) > 3 else sys.argv[1])
    else:
        main(sys.argv[1])
