#!/usr/bin/env python
"""bench.py -- the headline benchmark of the quantised MUL_MAT hot path (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--model llama-3-8b-q4_k_m] [--mode tg|pp]
    python bench.py --impl reference ...      # the reference's own CPU path on the host cores

A "step" is one pass of the hot path over one batch: every quantised MUL_MAT of one forward pass of
the named model (225 nodes for Llama-3-8B Q4_K_M) at N = 1 (tg: one generated token) -- the part of a
llama-bench tg128 token this repository replaces. `value` = tokens/s with everything resident in
HBM, timed with CUDA events on the launching stream; `e2e` = the same through the C ABI with HOST
buffers (pinned x in, logits out, copies inside the timed region, host sync per token as a decode
loop needs). `pp512` (N = 512) rides along in the same JSON line. Weights are random-init blocks of
the model's shapes/types (4.6 GB >> the 126 MB L2, so nothing is cache resident between steps).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


def _peaks():
    try:
        return json.load(open(ROOT / "MEASURED_PEAKS.json")), "measured"
    except Exception:  # noqa: BLE001
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 <= ts <= t1 + 0.2] or [r for _, r in self.rows]
        sm, mx, reasons = [], None, set()
        for r in rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# the reference's CPU path (oracle/_ref): cpu_baseline leg and --impl reference
# ------------------------------------------------------------------------------------------------
class CpuChain:
    """The reference CPU backend's MUL_MAT (N = 1) on one matrix of every distinct (type, m, k) of the model,
    extrapolated to a full token: t_token = sum_nodes t(type, m, k). Weights are generated and the per-shape
    iteration counts calibrated once; every step() re-times all shapes within `budget_s` of compute."""

    def __init__(self, model: str, threads: int):
        import numpy as np
        from oracle.oracle import Ref, Oracle, random_blocks, ref_variant   # checker / baseline only
        self.w = _load_pkg_module("workloads")
        self.ops = self.w.model_ops(model)
        self.threads = threads
        shapes: dict[tuple, int] = {}
        for o in self.ops:
            shapes[(o.type, o.m, o.k)] = shapes.get((o.type, o.m, o.k), 0) + (o.n_used if o.n_expert else 1)
        rng = np.random.default_rng(1)
        self.kind = "reference" if ref_variant() else "port"
        self.ref = Ref() if self.kind == "reference" else None
        self.orc = Oracle()
        self.items = []
        self.paths = set()
        for (t, m, k), cnt in sorted(shapes.items(), key=lambda kv: -kv[0][1] * kv[0][2]):
            mm = min(m, 16384)                       # bounded sample: at most 16384 rows of the matrix, scaled
            w = random_blocks(t, mm, k, rng)
            x = rng.standard_normal((1, k)).astype(np.float32)
            path, s1 = "oracle-port", None
            if self.ref is not None:
                path = "repack"
                try:
                    _, s1 = self.ref.mul_mat(t, w, x, n_threads=threads, repack=True, iters=5)
                except RuntimeError:
                    path = "generic"
                    _, s1 = self.ref.mul_mat(t, w, x, n_threads=threads, repack=False, iters=5)
            self.paths.add(path)
            self.items.append({"t": t, "m": m, "mm": mm, "k": k, "cnt": cnt, "w": w, "x": x, "path": path, "s1": s1})

    def step(self, budget_s: float) -> float:
        per_shape = budget_s / len(self.items)
        total = 0.0
        for it in self.items:
            if self.ref is not None:
                iters = max(3, min(200, int(per_shape / max(it["s1"], 1e-6))))
                _, secs = self.ref.mul_mat(it["t"], it["w"], it["x"], n_threads=self.threads, repack=(it["path"] == "repack"), iters=iters)
                it["s1"] = secs
            else:
                t0 = time.perf_counter(); self.orc.mul_mat(it["t"], it["w"], it["x"]); secs = time.perf_counter() - t0
            total += secs * it["m"] / it["mm"] * it["cnt"]
        return 1.0 / total

    def sample(self) -> str:
        return (f"one N=1 MUL_MAT per distinct (type,m,k) [{len(self.items)} shapes, <=16384 rows each, scaled], summed over the "
                f"{len(self.ops)} nodes of a token; paths={sorted(self.paths)}; {self.threads} threads")


def cpu_tokens_per_s(model: str, budget_s: float, threads: int):
    c = CpuChain(model, threads)
    return c.step(budget_s), c.kind, c.sample()


def _load_pkg_module(name):
    import importlib
    from __graft_entry__ import load_pkg, PKG_NAME
    try:
        load_pkg()
        return importlib.import_module(f"{PKG_NAME}.{name}")
    except ImportError:
        # the reference arm must run even where the CUDA library is absent: workloads.py is pure host logic
        import importlib.util
        spec = importlib.util.spec_from_file_location(f"_b200_{name}", ROOT / "zyk-llama.cpp_b200" / f"{name}.py")
        mod = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = mod
        spec.loader.exec_module(mod)
        return mod


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    chain = CpuChain(args.model, threads)
    # a step = one bounded sample of the token's MUL_MATs; the whole run is sized to about two minutes of CPU work
    budget = min(8.0, max(0.25, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        chain.step(budget)
    vals = []
    t0 = time.time()
    for _ in range(args.steps):
        vals.append(chain.step(budget))
    dt = time.time() - t0
    v = len(vals) / sum(1.0 / x for x in vals)       # steps / total extrapolated seconds
    kind, sample = chain.kind, chain.sample() + f"; {budget:.2f} s of compute per step"
    line = {"impl": "reference", "metric": "tg tokens/s (quantised MUL_MAT chain, N=1)", "value": v, "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int8 x int8 -> int32, f32 scales", "data": "synthetic",
            "config": {"workload": f"{args.model} tg (N=1), every quantised MUL_MAT of one token", "host": "reference CPU backend (oracle/_ref)"},
            "cpu_baseline": {"value": v, "unit": "tokens/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": dt}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama-3-8b-q4_k_m")
    ap.add_argument("--pp", type=int, default=512, help="prefill ubatch (0 = skip the pp leg)")
    ap.add_argument("--pp-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="row split (N > 1): Y slices stored into every rank's arena by the kernels, or one NCCL all-gather per node")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--per-node", action="store_true",
                    help="tg on one GPU: one launch per MUL_MAT group (CUDA graph) instead of the persistent chain kernel (b200_chain_*)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_pkg
    b = load_pkg()                       # raises if libb200mm.so is missing: no fallback
    from zyk_llama_cpp_b200 import workloads as W
    from zyk_llama_cpp_b200.chain import MatMulChain

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (the product has no CPU path)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    peaks, peak_src = _peaks()
    ops = W.model_ops(args.model)
    warmup = max(3, args.warmup)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- row split exchange: one all-gather of the Y slices per node (rowsplit.py) --------
    rowsplit = None
    if world > 1:
        from zyk_llama_cpp_b200.rowsplit import RowSplit
        rowsplit = RowSplit(world, rank, b.row_split)

    def time_chain(n, steps, wu):
        chain = MatMulChain(ops, n, device=local, rowsplit=rowsplit, fused_exchange=args.exchange == "fused", persistent=False if args.per_node else "auto")
        if chain.capturable:
            chain.capture()
        e0, e1 = b.Event(local), b.Event(local)
        for _ in range(wu):
            chain.step()
        chain.stream.synchronize()
        barrier()
        sampler = ClockSampler(local) if rank == 0 else None
        t0 = time.time()
        e0.record(chain.stream)
        for _ in range(steps):
            chain.step()
        e1.record(chain.stream)
        chain.stream.synchronize()
        barrier()
        t1 = time.time()
        ms = torch.tensor([e0.elapsed_ms(e1)], device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        clocks = sampler.stop(t0, t1) if sampler else None
        return chain, float(ms.item()), clocks

    # ---- tg: device-resident ---------------------------------------------------------------------------------
    chain, ms_total, clocks = time_chain(1, args.steps, warmup)
    graph_used = chain.graph is not None
    persistent = chain.decode_chain is not None
    chain_info = chain.decode_chain.info() if persistent else None
    ms_step = ms_total / args.steps
    tg = 1e3 / ms_step
    launches = (chain.kernels_per_step or len(chain.calls) * 2) * args.steps
    bytes_tok = W.tg_bytes(ops)
    achieved = bytes_tok / world / (ms_step * 1e-3) / 1e9       # per-GPU GB/s of algorithmic weight bytes

    # ---- e2e: pinned host x -> device, chain, logits -> host, host sync per token ------------------------
    x_dev, logits_dev = chain.bufs["x"], chain.bufs["logits"]
    x_host = torch.randn((1, x_dev.shape[1])).pin_memory()
    logits_host = torch.empty((1, logits_dev.shape[1])).pin_memory()
    h2d, d2h = x_host.numel() * 4, logits_host.numel() * 4

    def e2e_step():
        b.lib.b200_memcpy_h2d(x_dev.data_ptr(), x_host.data_ptr(), h2d, chain.stream.handle)
        chain.step()
        b.lib.b200_memcpy_d2h(logits_host.data_ptr(), logits_dev.data_ptr(), d2h, chain.stream.handle)
        chain.stream.synchronize()
    for _ in range(warmup):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_tg = args.steps / float(e2e_s.item())
    chain.close()
    del chain
    torch.cuda.empty_cache()

    # ---- pp: N = 512 ubatch through the same nodes -----------------------------------------------------------
    pp = None
    if args.pp > 0:
        try:
            chain_pp, ms_pp, _ = time_chain(args.pp, args.pp_steps, 3)
            ms_pp /= args.pp_steps
            pp_ops = W.pp_ops(ops) * args.pp
            pp = {"value": args.pp / (ms_pp * 1e-3), "unit": "tokens/s", "n": args.pp, "ms_per_ubatch": ms_pp,
                  "int8_tops": pp_ops / (ms_pp * 1e-3) / 1e12 / world, "steps": args.pp_steps,
                  # fp16 tcgen05 path: the same op count against the measured cuBLAS bf16 throughput (sustained: this is a long step)
                  "tensor_roofline": {"bound": "tensor", "achieved": pp_ops / (ms_pp * 1e-3) / 1e12 / world, "peak": peaks.get("bf16_tflops_sustained"),
                                      "unit": "TFLOP/s", "frac": pp_ops / (ms_pp * 1e-3) / 1e12 / world / peaks.get("bf16_tflops_sustained", 1400.0)}}
            chain_pp.close()
            del chain_pp
        except Exception as e:  # noqa: BLE001
            pp = {"error": str(e)[:200]}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    cpu = None
    if not args.no_cpu_baseline:
        try:
            v, kind, sample = cpu_tokens_per_s(args.model, args.cpu_seconds, os.cpu_count() or 1)
            cpu = {"value": v, "unit": "tokens/s", "cores": os.cpu_count() or 1, "kind": kind, "sample": sample}
        except Exception as e:  # noqa: BLE001
            cpu = {"value": None, "unit": "tokens/s", "cores": os.cpu_count() or 1, "kind": "port", "sample": f"failed: {e}"[:200]}

    traffic = None
    tf = ROOT / "profiles" / "traffic.json"
    if tf.exists():
        try:
            tj = json.load(open(tf))     # ncu dram__bytes_read.sum + dram__bytes_write.sum over the launches of one token
            traffic = (tj["dram_bytes_read_per_token"] + tj["dram_bytes_write_per_token"]) / max(1, world)
        except Exception:  # noqa: BLE001
            pass
    line = {
        "metric": "tg tokens/s (quantised MUL_MAT chain, N=1)", "value": tg, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int8 x int8 -> int32, f32 scales", "data": "synthetic",
        "config": {"workload": f"{args.model} tg (N=1): every quantised MUL_MAT of one token ({len(ops)} nodes)",
                   "weights": "random-init blocks, plane layout (load-time repack)", "weight_gb_per_token": bytes_tok / 1e9,
                   "l2": "inputs larger than L2 (weights streamed once per step)", "parallelism": f"row-split x{world}" if world > 1 else "single GPU",
                   "cuda_graph": graph_used, "exchange": (args.exchange if world > 1 else None),
                   "decode_kernel": ("persistent chain: one launch per token, %d dependent MUL_MAT groups" % chain_info["n_groups"]) if persistent
                                    else "one launch per MUL_MAT group"},
        "e2e": {"value": e2e_tg, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                     "traffic": traffic, "peak_source": peak_src, "frac_of_8tbs": achieved / 8000.0,
                     "algorithmic_bytes": bytes_tok / world,
                     "kernel": ("gemv_chain_kernel<1>: ONE launch = one decode step (every MUL_MAT group of the token), bytes and time per launch; "
                                if persistent else
                                "gemv_fused_kernel<1> (+ act_quant_kernel / gemv_kernel for the k = 14336 nodes): all launches of one decode step, "
                                "bytes and time per step (= per token); ") + "traffic from profiles/traffic.json"},
        "pp512": pp, "cpu_baseline": cpu, "clocks": clocks,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
