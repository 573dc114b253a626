#!/usr/bin/env python
"""profiles/traffic.json from an ncu launch list of the headline decode step (the file bench.py's roofline.traffic comes from), stamped
with the commit and the kernel it was taken on, so a stale capture is recognisable (bench.py drops it when the kernel name differs):
    python bench.py --steps 8 --warmup 3 --pp 0 --no-cpu-baseline --configs none > plain.log &&
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:gemv_chain -c 16 --csv \\
        --log-file gpurun_out/r2_chain_launches.csv python bench.py --steps 8 --warmup 3 --pp 0 --no-cpu-baseline --configs none
    python tools/make_traffic.py gpurun_out/r2_chain_launches.csv <commit> > profiles/traffic.json"""
import collections
import csv
import json
import sys


def main(path, commit):
    lines = [l for l in open(path) if not l.startswith("==")]
    per = collections.defaultdict(dict)
    names = {}
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"].lower()
        m = row["Metric Name"]
        if m.startswith("dram__bytes"):
            v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        else:
            v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(u, 1)
        per[row["ID"]][m] = v
        names[row["ID"]] = row["Kernel Name"]
    ids = sorted(per, key=int)
    kern = collections.Counter(names[i].split("(")[0] for i in ids).most_common(1)[0][0]
    sel = [i for i in ids if names[i].split("(")[0] == kern]
    n = len(sel)
    out = {"round": 2, "commit": commit, "workload": "llama-3-8b-q4_k_m tg N=1", "launches_per_token": 1, "launches_captured": n,
           "kernel": kern.replace("void ", "").replace("<unnamed>::", "").replace("(anonymous namespace)::", ""),
           "us_per_launch_under_ncu": sum(per[i]["gpu__time_duration.sum"] for i in sel) / n,
           "dram_bytes_read_per_token": sum(per[i]["dram__bytes_read.sum"] for i in sel) / n,
           "dram_bytes_write_per_token": sum(per[i]["dram__bytes_write.sum"] for i in sel) / n,
           "algorithmic_bytes_per_token": 4616331264,
           "source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none over the chain launches of bench.py (tools/make_traffic.py)"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "unknown")
