"""GPU probe: per-launch time of the prefill GEMM (tcgen05) per weight type and shape, inside a CUDA graph,
with and without the fp16 activation prologue (X_PREQUANTISED re-uses the workspace of the previous call).
Usage: python tools/gemm_probe.py [--n 512] [--quick]   (env B200_GEMM_KSPLIT forces the k split)"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_pkg  # noqa: E402

b = load_pkg()
from zyk_llama_cpp_b200.workloads import random_weight  # noqa: E402

T = b.Types


def probe(t, m, n, k, prequant, reps=8):
    gen = torch.Generator(device="cuda"); gen.manual_seed(1)
    w = b.repack(t, random_weight(t, m, k, "cuda", gen), m, k)
    x = torch.randn((n, k), device="cuda")
    y = torch.empty((n, m), device="cuda")
    wsp = torch.empty(b.mul_mat_workspace(t, n, k) + 256, dtype=torch.uint8, device="cuda")
    s = b.Stream(0)
    torch.cuda.synchronize()
    b.mul_mat(t, w, x, m, layout=b.LAYOUT_PLANES, y=y, workspace=wsp, algo=2, stream=s)   # fills the fp16 X in the workspace
    s.synchronize()
    flags = b.MM_X_PREQUANTISED if prequant else 0
    with b.Graph(s) as g:
        for _ in range(reps):
            b.mul_mat(t, w, x, m, layout=b.LAYOUT_PLANES, y=y, workspace=wsp, algo=2, flags=flags, stream=s)
    for _ in range(3):
        g.launch()
    s.synchronize()
    e0, e1 = b.Event(0), b.Event(0)
    e0.record(s)
    for _ in range(5):
        g.launch()
    e1.record(s)
    s.synchronize()
    us = e0.elapsed_ms(e1) * 1e3 / (5 * reps)
    return us, 2.0 * m * n * k / us * 1e-6


def main():
    n = int(sys.argv[sys.argv.index("--n") + 1]) if "--n" in sys.argv else 512
    shapes = [(T.Q4_K, 4096, 4096), (T.Q4_K, 6144, 4096), (T.Q4_K, 28672, 4096), (T.Q4_K, 4096, 14336), (T.Q6_K, 4096, 14336),
              (T.Q6_K, 1024, 4096), (T.Q6_K, 128256, 4096), (T.Q8_0, 14336, 4096), (T.Q4_0, 4096, 14336), (T.IQ4_XS, 4096, 14336)]
    if "--quick" in sys.argv:
        shapes = shapes[:5]
    if "--shape" in sys.argv:
        i = int(sys.argv[sys.argv.index("--shape") + 1])
        shapes = shapes[i:i + 1]
    names = {v: k for k, v in vars(T).items() if isinstance(v, int)}
    print(f"n={n}")
    print(f"{'type':7s} {'m':>7s} {'k':>6s} | {'with prologue':>22s} | {'GEMM only':>22s}")
    for t, m, k in shapes:
        cells = []
        for pre in (False, True):
            us, tf = probe(t, m, n, k, pre)
            cells.append(f"{us:9.1f} us {tf:7.1f} TF")
        print(f"{names[t]:7s} {m:7d} {k:6d} | " + " | ".join(cells), flush=True)


if __name__ == "__main__":
    main()
