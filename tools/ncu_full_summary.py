#!/usr/bin/env python
"""Summarise an `ncu --set full` report (.ncu-rep) per captured launch: duration, grid, registers, DRAM bytes,
throughput percentages, tensor-pipe activity, issue rate and the top warp-stall reasons.
Usage: python tools/ncu_full_summary.py report.ncu-rep [> profiles/rN_kernel.txt]   (needs `ncu` on PATH, no GPU)"""
import csv
import io
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem/block"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % active"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots % busy"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__inst_executed.sum", "warp instructions"),
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# {path}: {len(rows) - 2} launch(es) captured with --set full --clock-control none")
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        print(f"\n## {name[:120]}")
        for key, label in WANT:
            if key in col and r[col[key]] != "":
                print(f"  {label:24s} {r[col[key]]:>16s} {units[col[key]]}")
        if "dram__bytes_read.sum" in col:
            def mb(k):
                v = float(r[col[k]].replace(",", "")); u = units[col[k]].lower()
                return v * {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1.0)
            tot = mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum")
            t = float(r[col["gpu__time_duration.sum"]].replace(",", "")); tu = units[col["gpu__time_duration.sum"]]
            t_us = t * {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(tu, 1.0)
            print(f"  {chr(100)}ram traffic             {tot:16.3f} MB  -> {tot / t_us:8.2f} TB/s under ncu (cold, serialised)")
        stalls = []
        for h, i in col.items():
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and r[i] != "":
                stalls.append((float(r[i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
        stalls.sort(reverse=True)
        print("  top stalls (warps per issue): " + ", ".join(f"{n} {v:.2f}" for v, n in stalls[:6]))


if __name__ == "__main__":
    for p in sys.argv[1:]:
        main(p)
