#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel family count, total, share."""
import collections, csv, re, sys

def main(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0.0
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", "")); u = row["Metric Unit"]
        v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
        name = row["Kernel Name"]
        mt = re.search(r"(\w+)<\(b200_type\)(\d+)", name) or re.search(r"(\w+)<(\d+)", name)
        key = f"{mt.group(1)}<type={mt.group(2)}> grid={row.get('Grid Size','?')}" if mt else re.sub(r"\(.*", "", name)[:50]
        agg[key][0] += 1; agg[key][1] += v; tot += v
    print(f"# {path}: {sum(n for n, _ in agg.values())} launches, {tot:.1f} us serialised (cold-cache: compare SHARES)")
    print(f"{'kernel':58s} {'n':>5s} {'total us':>10s} {'avg us':>8s} {'share':>6s}")
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:58s} {n:5d} {t:10.1f} {t / n:8.2f} {t / tot:6.3f}")

if __name__ == "__main__":
    for p in sys.argv[1:]:
        main(p)
