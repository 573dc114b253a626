"""Measured dense tensor-core peaks of this B200 next to the prefill GEMM's operand choice ("to be measured on box"):
cuBLASLt int8 x int8 -> int32 (torch._int_mm), fp16 and bf16 (torch.matmul) at 8192^3, best of 10 and sustained over ~2 s, CUDA events.
Library GEMMs are only the yardstick here (the product's GEMM is gemm.cu).   python tools/tensor_peaks.py > profiles/r2_tensor_peaks.json"""
import json
import time

import torch


def bench(fn, flops, burst=10, sustain_s=2.0):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 0.0
    for _ in range(burst):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = max(best, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time(); e0.record()
    while time.time() - t0 < sustain_s:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    return best, flops * n / (e0.elapsed_time(e1) * 1e-3) / 1e12


def main():
    N = 8192
    out = {"gpu": torch.cuda.get_device_name(0), "shape": f"{N}^3", "unit": "T(FL)OP/s = 2 N^3 / time"}
    a8 = torch.randint(-127, 127, (N, N), dtype=torch.int8, device="cuda")
    b8 = torch.randint(-127, 127, (N, N), dtype=torch.int8, device="cuda")
    try:
        burst, sus = bench(lambda: torch._int_mm(a8, b8), 2.0 * N ** 3)
        out["int8_int32_cublaslt"] = {"burst": burst, "sustained": sus}
    except Exception as e:  # noqa: BLE001
        out["int8_int32_cublaslt"] = {"error": str(e)[:200]}
    for name, dt in (("fp16", torch.float16), ("bf16", torch.bfloat16)):
        a = torch.randn((N, N), dtype=dt, device="cuda"); b = torch.randn((N, N), dtype=dt, device="cuda")
        burst, sus = bench(lambda: torch.matmul(a, b), 2.0 * N ** 3)
        out[name + "_cublas"] = {"burst": burst, "sustained": sus}
    try:
        a = torch.randn((N, N), device="cuda").to(torch.float8_e4m3fn); b = torch.randn((N, N), device="cuda").to(torch.float8_e4m3fn).t()
        one = torch.ones((), device="cuda")
        burst, sus = bench(lambda: torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16), 2.0 * N ** 3)
        out["fp8_e4m3_cublaslt"] = {"burst": burst, "sustained": sus}
    except Exception as e:  # noqa: BLE001
        out["fp8_e4m3_cublaslt"] = {"error": str(e)[:200]}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
