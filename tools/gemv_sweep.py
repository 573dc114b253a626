"""GPU micro-benchmark: decode GEMV bandwidth per weight format / shape / (rows-per-warp, unroll).
Weights rotate over >= 512 MB of distinct buffers so nothing is L2-resident (B200 L2 ~126 MB).
Usage: python tools/gemv_sweep.py [--quick]    -> table on stdout (+ gpurun_out/gemv_sweep.json)"""
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_pkg  # noqa: E402

b = load_pkg()
T = b.Types
PEAK = 6557.4
try:
    PEAK = json.load(open(ROOT / "MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    pass


def rand_weights(t, m, k, copies):
    n = m * b.row_size(t, k)
    w = torch.randint(0, 256, (copies, n), dtype=torch.uint8, device="cuda")
    return w


def bench(t, m, k, tune, flags, iters=24):
    nbytes = m * b.row_size(t, k)
    copies = max(2, int(np.ceil(512e6 / nbytes)))
    w = rand_weights(t, m, k, copies)
    x = torch.randn((1, k), device="cuda")
    y = torch.empty((1, m), device="cuda")
    ws = torch.empty(b.mul_mat_workspace(t, 1, k) + 256, dtype=torch.uint8, device="cuda")
    layout = b.LAYOUT_PLANES
    b.mul_mat(t, w[0], x, m, layout=layout, y=y, workspace=ws, tune=tune)       # quantises X once
    fl = flags | b.MM_X_PREQUANTISED
    for i in range(4):
        b.mul_mat(t, w[i % copies], x, m, layout=layout, y=y, workspace=ws, tune=tune, flags=fl)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(iters):
        b.mul_mat(t, w[i % copies], x, m, layout=layout, y=y, workspace=ws, tune=tune, flags=fl)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    return us, nbytes / us * 1e-3


def main():
    quick = "--quick" in sys.argv
    shapes = [(T.Q4_K, 4096, 4096), (T.Q4_K, 1024, 4096), (T.Q4_K, 14336, 4096), (T.Q4_K, 4096, 14336),
              (T.Q6_K, 4096, 14336), (T.Q6_K, 1024, 4096), (T.Q6_K, 128256, 4096),
              (T.Q8_0, 14336, 4096), (T.Q4_0, 8192, 2048), (T.IQ4_XS, 14336, 4096)]
    tunes = [0] + [r | (u << 8) for r in (1, 2, 4) for u in (1, 2, 4) if not (r == 4 and u == 4)]
    if quick:
        shapes, tunes = shapes[:4], [0, 0x0202]
    rows = []
    print(f"{'type':7s} {'m':>6s} {'k':>6s} {'MB':>7s} {'R':>2s} {'U':>2s} {'pdl':>3s} {'us':>8s} {'GB/s':>8s} {'frac':>5s}")
    for (t, m, k) in shapes:
        for tune in tunes:
            for pdl in ((0, b.MM_PDL) if tune in (0, 0x0202) else (0,)):
                try:
                    us, gbs = bench(t, m, k, tune, pdl)
                except Exception as e:  # noqa: BLE001
                    print(f"{T.NAMES[t]} {m} {k} tune={tune:#x}: {e}")
                    continue
                r = dict(type=T.NAMES[t], m=m, k=k, mb=m * b.row_size(t, k) / 1e6, R=tune & 0xff, U=tune >> 8, pdl=bool(pdl), us=us, gbs=gbs, frac=gbs / PEAK)
                rows.append(r)
                print(f"{r['type']:7s} {m:6d} {k:6d} {r['mb']:7.1f} {r['R']:2d} {r['U']:2d} {int(r['pdl']):3d} {us:8.2f} {gbs:8.1f} {r['frac']:5.2f}", flush=True)
    out = ROOT / "gpurun_out"
    out.mkdir(exist_ok=True)
    json.dump(rows, open(out / "gemv_sweep.json", "w"), indent=1)


if __name__ == "__main__":
    main()
