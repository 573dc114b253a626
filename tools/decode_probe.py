"""GPU probe: per-launch time of the decode kernels inside a CUDA graph (as bench.py runs them), per shape,
kernel (1 = register GEMV + quantiser kernel, 3 = streaming GEMV) and PDL on/off. Weights rotate over > 400 MB.
Usage: python tools/decode_probe.py [--tune "KB,per_sm"]"""
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_pkg  # noqa: E402

b = load_pkg()
from zyk_llama_cpp_b200.workloads import random_weight, row_bytes  # noqa: E402

T = b.Types
PEAK = 6553.0


def probe(t, m, k, algo, pdl, reps=48):
    nbytes = m * row_bytes(t, k)
    copies = max(2, int(420e6 // nbytes) + 1)
    gen = torch.Generator(device="cuda"); gen.manual_seed(1)
    ws = [b.repack(t, random_weight(t, m, k, "cuda", gen), m, k) for _ in range(copies)]
    x = torch.randn((1, k), device="cuda")
    y = torch.empty((1, m), device="cuda")
    wsp = torch.empty(b.mul_mat_workspace(t, 1, k) + 256, dtype=torch.uint8, device="cuda")
    s = b.Stream(0)
    torch.cuda.synchronize()
    flags = b.MM_PDL if pdl else 0
    with b.Graph(s) as g:
        for i in range(reps):
            b.mul_mat(t, ws[i % copies], x, m, layout=b.LAYOUT_PLANES, y=y, workspace=wsp, algo=algo, flags=flags, stream=s)
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
    return us, nbytes / us * 1e-3


def main():
    if "--debug" in sys.argv:
        os.environ["B200_STREAM_DEBUG"] = sys.argv[sys.argv.index("--debug") + 1]
    if "--tune" in sys.argv:
        os.environ["B200_STREAM_TUNE"] = sys.argv[sys.argv.index("--tune") + 1]
    shapes = [(T.Q4_K, 32768, 2048), (T.Q6_K, 32768, 2048), (T.Q8_0, 32768, 2048), (T.Q4_K, 4096, 4096), (T.Q4_K, 6144, 4096), (T.Q4_K, 28672, 4096), (T.Q4_K, 4096, 14336), (T.Q6_K, 4096, 14336),
              (T.Q6_K, 1024, 4096), (T.Q6_K, 128256, 4096), (T.Q8_0, 14336, 4096), (T.Q4_0, 8192, 2048), (T.IQ4_XS, 14336, 4096)]
    if "--quick" in sys.argv:
        shapes = shapes[:3] + [(T.Q6_K, 128256, 4096)]
    print(f"tune={os.environ.get('B200_STREAM_TUNE', 'default')} debug={os.environ.get('B200_STREAM_DEBUG', '0')}")
    print(f"{'type':7s} {'m':>7s} {'k':>6s} {'MB':>7s} | " + " | ".join(f"{n:>16s}" for n in ("reg", "reg+pdl", "stream", "stream+pdl")))
    for t, m, k in shapes:
        cells = []
        for algo, pdl in ((1, 0), (1, 1), (3, 0), (3, 1)):
            us, gbs = probe(t, m, k, algo, pdl)
            cells.append(f"{us:6.1f}us {gbs / PEAK:5.2f}")
        print(f"{T.NAMES[t]:7s} {m:7d} {k:6d} {m * row_bytes(t, k) / 1e6:7.1f} | " + " | ".join(f"{c:>16s}" for c in cells), flush=True)


if __name__ == "__main__":
    main()
