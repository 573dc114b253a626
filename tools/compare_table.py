"""Parse the perf logs tools/compare_ggml_cuda.sh wrote (reference test-backend-ops perf lines) into one table: us per run through the
reference's CUDA backend and through this plugin, and the ratio (> 1: this backend is faster)."""
import re
import sys
from pathlib import Path

LINE = re.compile(r"^\s*(MUL_MAT(?:_ID)?)\((.*?)\):\s+(\d+) runs -\s+([\d.]+) us/run")


def parse(path):
    out = {}
    if not path.exists():
        return out
    for ln in path.read_text(errors="replace").splitlines():
        ln = re.sub(r"\x1b\[[0-9;]*m", "", ln)
        m = LINE.match(ln)
        if m:
            out[(m.group(1), m.group(2))] = float(m.group(4))
    return out


def short(op, v):
    kv = dict(p.split("=", 1) for p in re.findall(r"(\w+=\[[^\]]*\]|\w+=[^,]*)", v))
    if op == "MUL_MAT":
        return f"{kv.get('type_a')} m={kv.get('m')} n={kv.get('n')} k={kv.get('k')}"
    return f"{kv.get('type_a')} n_mats={kv.get('n_mats')} n_used={kv.get('n_used')} m={kv.get('m')} n={kv.get('n')} k={kv.get('k')}"


def main():
    d = Path(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out")
    rows = []
    for op, f in (("MUL_MAT", "mul_mat"), ("MUL_MAT_ID", "mul_mat_id")):
        cu, us = parse(d / f"cuda_{f}.txt"), parse(d / f"b200_{f}.txt")
        for key in cu.keys() | us.keys():
            if key[0] == op and "FUSION" not in key[1]:
                rows.append((op, short(op, key[1]), cu.get(key), us.get(key)))
    rows.sort(key=lambda r: (r[0], r[1].split()[0], int(re.search(r" n=(\d+)", r[1]).group(1)), r[1]))
    print(f"| op | case | ggml-cuda us/run | this backend us/run | ggml-cuda / this |")
    print("|---|---|---|---|---|")
    for op, case, a, b in rows:
        ratio = f"{a / b:.2f}" if a and b else "-"
        print(f"| {op} | {case} | {a if a is not None else 'n/a'} | {b if b is not None else 'n/a'} | {ratio} |")
    both = [(a, b) for _, _, a, b in rows if a and b]
    if both:
        wins = sum(1 for a, b in both if b <= a)
        print(f"\n{wins} of {len(both)} cases: this backend at least as fast as ggml-cuda")


if __name__ == "__main__":
    main()
