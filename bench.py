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


def _load_workloads():
    """workloads.py is pure host logic (shape tables): loaded BY PATH so that the reference arm never imports the product package,
    hence never maps zyk-llama.cpp_b200/lib/libb200mm.so"""
    import importlib.util
    name = "_b200_workloads_by_path"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, ROOT / "zyk-llama.cpp_b200" / "workloads.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def _load_pkg_module(name):
    assert name == "workloads"
    return _load_workloads()


def plugin_e2e(steps: int, warmup: int, pp: int, pp_steps: int):
    """tg (and pp) of the headline model through libggml-b200.so driven by the reference's own ggml runtime (GGML_BACKEND_PATH), as a
    subprocess: oracle/_ref/<variant>/plugin_bench. None where the reference libraries cannot run on this host."""
    try:
        from oracle.oracle import ref_dir
        d = ref_dir()
        exe = d / "plugin_bench" if d else None
        plugin = ROOT / "zyk-llama.cpp_b200" / "lib" / "libggml-b200.so"
        if exe is None or not exe.exists() or not plugin.exists():
            return {"unavailable": "oracle/_ref/*/plugin_bench or libggml-b200.so not built for this host"}
        env = dict(os.environ, GGML_BACKEND_PATH=str(plugin))
        out = {}
        for key, n, k, w in (("tg", 1, steps, warmup), ("pp%d" % pp, pp, pp_steps, 2)):
            if n <= 0:
                continue
            r = subprocess.run([str(exe), "--n", str(n), "--steps", str(k), "--warmup", str(w)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
            line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
            if r.returncode != 0 or not line:
                out[key] = {"error": (r.stderr or r.stdout)[-300:]}
                continue
            out[key] = json.loads(line[-1])
        return out
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[:300]}


def llama_bench_leg(model: str, reps: int = 5):
    """The reference's own llama-bench binary (compiled from /root/reference into oracle/_ref, unmodified) on a random-init GGUF of the headline
    model (tools/make_gguf.py), with this repository's plugin as its GPU device: the WHOLE llama graph (matmuls, norms, rope, KV cache, attention)
    on the B200, pp512 and tg128 as llama-bench defines them. The file lives in a temporary directory for the duration of the call."""
    import tempfile
    try:
        from oracle.oracle import ref_dir
        d = ref_dir()
        exe = d / "llama-bench" if d else None
        plugin = ROOT / "zyk-llama.cpp_b200" / "lib" / "libggml-b200.so"
        if exe is None or not exe.exists() or not plugin.exists():
            return {"unavailable": "oracle/_ref/*/llama-bench or libggml-b200.so not built for this host"}
        with tempfile.TemporaryDirectory() as tmp:
            gguf = Path(tmp) / "m.gguf"
            r = subprocess.run([sys.executable, str(ROOT / "tools" / "make_gguf.py"), "--model", model, "--out", str(gguf)],
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
            if r.returncode != 0:
                return {"error": "make_gguf: " + r.stdout[-300:]}
            env = dict(os.environ, GGML_BACKEND_PATH=str(plugin))
            r = subprocess.run([str(exe), "-m", str(gguf), "-p", "512", "-n", "128", "-r", str(reps), "-ngl", "99", "-o", "json"],
                               env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
        if r.returncode != 0:
            return {"error": (r.stderr or r.stdout)[-300:]}
        start = r.stdout.find("[")
        rows = json.loads(r.stdout[start:])
        out = {"through": "reference llama-bench (oracle/_ref, unmodified) -> reference loader, scheduler and llama_decode -> this plugin (GGML_BACKEND_PATH), -ngl 99",
               "model": model, "data": "random-init GGUF (tools/make_gguf.py)", "repetitions": reps}
        for row in rows:
            key = "pp%d" % row["n_prompt"] if row.get("n_prompt") else "tg%d" % row["n_gen"]
            out[key] = {"value": row["avg_ts"], "stddev": row["stddev_ts"], "unit": "tokens/s"}
            out.setdefault("backend", row.get("backends"))
        return out
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[:300]}


def workload_string(model: str, n_nodes: int) -> str:
    """the same string in both arms' config.workload"""
    return f"{model} tg (N=1): every quantised MUL_MAT of one token ({n_nodes} nodes)"


METRIC = "tg tokens/s (quantised MUL_MAT chain, N=1)"


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
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / v, "ms_per_step_is": "extrapolated token time (bounded sample scaled to the 225 nodes), not a measured step",
            "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int8 x int8 -> int32, f32 scales", "data": "synthetic",
            "config": {"workload": workload_string(args.model, len(chain.ops)), "host": "reference CPU backend (oracle/_ref)"},
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
    ap.add_argument("--pp-steps", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-llama-bench", action="store_true", help="skip the llama_bench key (the reference's llama-bench binary on the plugin, N = 1 only)")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="row split (N > 1): Y slices stored into every rank's arena by the kernels, or one NCCL all-gather per node")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--per-node", action="store_true",
                    help="tg on one GPU: one launch per MUL_MAT group (CUDA graph) instead of the persistent chain kernel (b200_chain_*)")
    ap.add_argument("--configs", default="auto",
                    help="extra BASELINE.json configs reported as extra keys: 'auto' (all at N = 1; 70B and Mixtral at N > 1), 'none', or a comma list of model names")
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
    warmup = max(3, args.warmup)
    dev = f"cuda:{local}"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rowsplit = None
    if world > 1:
        from zyk_llama_cpp_b200.rowsplit import RowSplit
        rowsplit = RowSplit(world, rank, b.row_split)

    def make_chain(ops, n, persistent="auto"):
        return MatMulChain(ops, n, device=local, rowsplit=rowsplit, fused_exchange=args.exchange == "fused",
                           persistent=False if args.per_node else persistent)

    def time_chain(chain, steps, wu, sample_clocks=False):
        if chain.capturable and chain.graph is None:
            chain.capture()
        e0, e1 = b.Event(local), b.Event(local)
        for _ in range(wu):
            chain.step()
        chain.stream.synchronize()
        barrier()
        sampler = ClockSampler(local) if (rank == 0 and sample_clocks) else None
        t0 = time.time()
        e0.record(chain.stream)
        for _ in range(steps):
            chain.step()
        e1.record(chain.stream)
        chain.stream.synchronize()
        barrier()
        t1 = time.time()
        ms = torch.tensor([e0.elapsed_ms(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        clocks = sampler.stop(t0, t1) if sampler else None
        return float(ms.item()) / steps, clocks

    # ---- parity of the N-rank row split against the single-GPU per-node chain (same seeded weights and inputs) ------------------
    def split_parity(ops, n, chain):
        """every activation buffer after one pass, NMSE against a single-GPU chain of per-node launches built on THIS rank from the
        same seed; max over buffers and ranks"""
        ref = MatMulChain(ops, n, device=local, persistent=False)
        x0 = ref.bufs["x"].clone()
        ref.step(); ref.stream.synchronize()
        worst, names = 0.0, []
        chain.bufs["x"].copy_(x0)
        barrier()
        chain.issue(); chain.stream.synchronize()
        barrier()
        for name, want in ref.bufs.items():
            got = chain.bufs[name].double(); w = want.double()
            den = float((w * w).sum())
            if den == 0.0 or not bool(torch.isfinite(w).all()):
                continue                      # a buffer nobody writes (static input): nothing to compare
            worst = max(worst, float(((got - w) ** 2).sum()) / den)
            names.append(name)
        ref.close()
        del ref
        torch.cuda.empty_cache()
        t = torch.tensor([worst], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), names

    parity = None
    PARITY_TOL = 1e-9        # decode: integer arithmetic and fixed summation orders, the row split changes nothing but which GPU adds
    PARITY_TOL_PP = 1e-6     # prefill: fp32 sums in a different tile / k-split order, then every node rounds its input to fp16 (11 bits): differences
                             # of one ulp of fp16 in a few activations travel down 225 dependent nodes (measured 6e-9 at N = 2)

    def bench_model(model, steps, wu, pp_steps, headline=False):
        """tg (N = 1) and pp (N = args.pp) of one model; returns the keys of its block"""
        nonlocal parity
        ops = W.model_ops(model)
        out = {"workload": workload_string(model, len(ops)), "weight_gb_per_token": W.tg_bytes(ops) / 1e9}
        chain = make_chain(ops, 1)
        if world > 1:
            nm, names = split_parity(ops, 1, chain)
            pr = {"nmse_max": nm, "checked_buffers": names, "against": "single-GPU per-node chain, same seeded weights and x", "tolerance": PARITY_TOL, "n": 1}
            out["parity"] = pr
            if headline:
                parity = {"tg": pr}
            if nm > PARITY_TOL:
                raise RuntimeError(f"{model}: row-split parity FAILED at world {world}: nmse {nm:.3e} > {PARITY_TOL}")
        ms_step, clocks = time_chain(chain, steps, wu, sample_clocks=headline)
        persistent = chain.decode_chain is not None
        info = chain.decode_chain.info() if persistent else None
        bytes_tok = W.tg_bytes(ops)
        achieved = bytes_tok / world / (ms_step * 1e-3) / 1e9
        out.update({"tg": {"value": 1e3 / ms_step, "unit": "tokens/s", "ms_per_step": ms_step, "steps": steps,
                           "decode_kernel": ("persistent chain: one launch per token, %d dependent MUL_MAT groups" % info["n_groups"]) if persistent
                                            else "gemv_chain_kernel, one launch per MUL_MAT group (chains of a single group; CUDA graph replay)" if getattr(chain, "group_chains", None) is not None
                                            else "one launch per MUL_MAT group (CUDA graph replay)" if chain.graph is not None else "one launch per MUL_MAT group (eager)"},
                    "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                                 "frac_of_8tbs": achieved / 8000.0, "algorithmic_bytes": bytes_tok / world, "peak_source": peak_src}})
        extra = {"chain": chain, "ops": ops, "ms_step": ms_step, "clocks": clocks, "persistent": persistent, "info": info,
                 "launches": (chain.kernels_per_step or len(chain.calls) * 2) * steps, "graph_used": chain.graph is not None}
        if not headline:
            chain.close()
            del chain
            extra["chain"] = None
            torch.cuda.empty_cache()
            if args.pp > 0 and pp_steps > 0:
                out["pp%d" % args.pp] = bench_pp(ops, model, pp_steps)
        return out, extra

    def bench_pp(ops, model, pp_steps):
        try:
            chain_pp = make_chain(ops, args.pp)
            res = {}
            if world > 1:
                nm, names = split_parity(ops, args.pp, chain_pp)
                res["parity"] = {"nmse_max": nm, "checked_buffers": names, "tolerance": PARITY_TOL_PP, "n": args.pp}
                if nm > PARITY_TOL_PP:
                    raise RuntimeError(f"{model}: row-split pp parity FAILED at world {world}: nmse {nm:.3e}")
            ms_pp, _ = time_chain(chain_pp, pp_steps, 3)
            pp_ops = W.pp_ops(ops) * args.pp
            tf = pp_ops / (ms_pp * 1e-3) / 1e12 / world
            res.update({"value": args.pp / (ms_pp * 1e-3), "unit": "tokens/s", "n": args.pp, "ms_per_ubatch": ms_pp, "steps": pp_steps,
                        "dtype": "fp16 x fp16 -> f32 (tcgen05.mma.kind::f16; weights dequantised in-kernel, activations scaled per row by a power of two)",
                        "fp16_tflops": tf,
                        # the same op count against the measured cuBLAS bf16 throughput (sustained: this is a long step) and against the nominal peaks
                        "tensor_roofline": {"bound": "tensor", "achieved": tf, "peak": peaks.get("bf16_tflops_sustained"), "unit": "TFLOP/s",
                                            "frac": tf / peaks.get("bf16_tflops_sustained", 1400.0), "peak_source": peak_src + " cuBLAS bf16, sustained",
                                            "frac_of_nominal_fp16_peak": tf / 2250.0, "frac_of_nominal_int8_peak": tf / 4500.0}})
            chain_pp.close()
            del chain_pp
            torch.cuda.empty_cache()
            if world > 1 and W.tg_bytes(ops) < 60e9:
                # prefill of INDEPENDENT ubatches shards by tokens, not by rows: every rank holds the whole model (it fits) and runs its own
                # ubatch with no exchange at all; the row-split figure above is what ONE ubatch gets from N GPUs
                try:
                    own = MatMulChain(ops, args.pp, device=local, persistent=False)
                    ms_own, _ = time_chain(own, pp_steps, 3)
                    res["token_parallel"] = {"value": world * args.pp / (ms_own * 1e-3), "unit": "tokens/s", "ms_per_ubatch": ms_own,
                                             "ubatches_in_flight": world, "sharding": "one ubatch per rank, weights replicated, no collective",
                                             "scaling": "weak"}
                    own.close()
                    del own
                    torch.cuda.empty_cache()
                except Exception as e:  # noqa: BLE001
                    res["token_parallel"] = {"error": str(e)[:200]}
            return res
        except RuntimeError:
            raise
        except Exception as e:  # noqa: BLE001
            return {"error": str(e)[:200]}

    # ---- headline: Llama-3-8B Q4_K_M tg, device resident -------------------------------------------------------------------------
    head, hx = bench_model(args.model, args.steps, warmup, args.pp_steps, headline=True)
    chain, ops, ms_step, clocks = hx["chain"], hx["ops"], hx["ms_step"], hx["clocks"]
    tg = 1e3 / ms_step
    bytes_tok = W.tg_bytes(ops)

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
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_tg = args.steps / float(e2e_s.item())
    chain.close()
    del chain
    hx["chain"] = None
    torch.cuda.empty_cache()

    # ---- the path a per-node caller reaches today (the ggml plugin on a real llama.cpp graph: MUL_MAT groups are never adjacent there) --
    tg_per_group = None
    if world == 1 and hx["persistent"]:
        ch2 = MatMulChain(ops, 1, device=local, persistent=False)
        ms2, _ = time_chain(ch2, args.steps, warmup)
        ach2 = bytes_tok / (ms2 * 1e-3) / 1e9
        tg_per_group = {"value": 1e3 / ms2, "unit": "tokens/s", "ms_per_step": ms2, "decode_kernel": "one launch per MUL_MAT group (gemv_fused_kernel / act_quant + gemv_kernel), CUDA graph replay",
                        "roofline_frac": ach2 / peaks["hbm_gbs"], "note": "the per-node GEMV kernels, one launch per group of nodes that share X"}
        ch2.close()
        del ch2
        torch.cuda.empty_cache()
        # what the plugin does on a real llama.cpp graph since round 2: every group is its own ONE-group chain-kernel launch
        try:
            ch3 = MatMulChain(ops, 1, device=local, persistent="groups")
            ms3, _ = time_chain(ch3, args.steps, warmup)
            tg_per_group["group_chains"] = {"value": 1e3 / ms3, "unit": "tokens/s", "ms_per_step": ms3, "roofline_frac": bytes_tok / (ms3 * 1e-3) / 1e9 / peaks["hbm_gbs"],
                                            "decode_kernel": "gemv_chain_kernel, one launch per MUL_MAT group (a chain of a single group), CUDA graph replay",
                                            "note": "what graph_compute reaches on a real llama.cpp graph (groups separated by non-MUL_MAT ops)"}
            ch3.close()
            del ch3
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001
            tg_per_group["group_chains"] = {"error": str(e)[:200]}

    # ---- pp: N = 512 ubatch through the same nodes -----------------------------------------------------------
    pp = bench_pp(ops, args.model, args.pp_steps) if args.pp > 0 else None
    if world > 1 and pp and "parity" in pp:
        parity["pp"] = pp["parity"]

    # ---- the other BASELINE.json configs, each with its own roofline ---------------------------------------------------------------
    extras = {}
    if args.configs != "none":
        if args.configs == "auto":
            names = (["llama-3-70b-q4_k_m", "mixtral-8x7b-q4_k_m", "llama-3-8b-q8_0", "llama-3-8b-iq4_xs", "llama-3.2-1b-q4_0"] if world == 1
                     else ["llama-3-70b-q4_k_m"] + (["mixtral-8x7b-q4_k_m"] if world == 2 else []))
        else:
            names = [n for n in args.configs.split(",") if n]
        for name in names:
            if name == args.model:
                continue
            try:
                big = "70b" in name or "mixtral" in name
                blk, _ = bench_model(name, 32 if big else 64, 3, 3 if big else 5)
                extras[name] = blk
            except RuntimeError:
                raise
            except Exception as e:  # noqa: BLE001
                extras[name] = {"error": str(e)[:300]}
            torch.cuda.empty_cache()

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
        # configs[0]: the fork's ZYK Q4_0 CPU kernels (CPU_REPACK buffer) on the 1B model, next to the GPU number of the same model
        if "llama-3.2-1b-q4_0" in extras and "error" not in extras["llama-3.2-1b-q4_0"]:
            try:
                v, kind, sample = cpu_tokens_per_s("llama-3.2-1b-q4_0", 5.0, os.cpu_count() or 1)
                extras["llama-3.2-1b-q4_0"]["cpu_baseline"] = {"value": v, "unit": "tokens/s", "cores": os.cpu_count() or 1, "kind": kind, "sample": sample}
            except Exception as e:  # noqa: BLE001
                extras["llama-3.2-1b-q4_0"]["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"[:200]}

    # ---- e2e through the ggml plugin: the same 225-node graph built with the reference's API (oracle/plugin_bench.cpp, linked against the
    #      reference's libggml from oracle/_ref; the plugin is loaded through GGML_BACKEND_PATH), host buffers in and out ----------------------
    e2e_plugin = None
    if world == 1 and args.model == "llama-3-8b-q4_k_m":
        e2e_plugin = plugin_e2e(args.steps, warmup, args.pp, max(3, args.pp_steps // 2))

    # ---- the whole llama graph: the reference's llama-bench binary with the plugin as its GPU device (N = 1; the real-model counterpart of `value`) ----
    llama_bench = None
    if world == 1 and not args.no_llama_bench and args.model in ("llama-3-8b-q4_k_m",):
        llama_bench = llama_bench_leg(args.model)

    # DRAM traffic of the dominant kernel from the ncu capture committed under profiles/ -- only when it is a capture of the kernel that ran
    traffic, traffic_src = None, None
    tf = ROOT / "profiles" / "traffic.json"
    kernel_name = "gemv_chain_kernel<1>" if hx["persistent"] else "gemv_fused_kernel<1>"
    if tf.exists():
        try:
            tj = json.load(open(tf))     # ncu dram__bytes_read.sum + dram__bytes_write.sum over the launches of one token
            if tj.get("kernel", "").startswith(kernel_name.split("<")[0]) and world == 1:
                traffic = tj["dram_bytes_read_per_token"] + tj["dram_bytes_write_per_token"]
                traffic_src = f"profiles/traffic.json (round {tj.get('round')}, commit {tj.get('commit', 'n/a')})"
        except Exception:  # noqa: BLE001
            pass
    persistent, chain_info = hx["persistent"], hx["info"]
    roof = head["roofline"]
    roof.update({"traffic": traffic, "traffic_source": traffic_src,
                 "kernel": ("gemv_chain_kernel<1>: ONE launch = one decode step (every MUL_MAT group of the token), bytes and time per launch"
                            if persistent else
                            "gemv_fused_kernel<1> (+ act_quant_kernel / gemv_kernel for the k = 14336 nodes): all launches of one decode step, "
                            "bytes and time per step (= per token)")})
    line = {
        "metric": METRIC, "value": tg, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int8 x int8 -> int32, f32 scales", "data": "synthetic",
        "config": {"workload": workload_string(args.model, len(ops)),
                   "weights": "random-init blocks, plane layout (load-time repack)", "weight_gb_per_token": bytes_tok / 1e9,
                   "l2": "inputs larger than L2 (weights streamed once per step)", "parallelism": f"row-split x{world}" if world > 1 else "single GPU",
                   "cuda_graph": hx["graph_used"], "exchange": (args.exchange if world > 1 else None),
                   "decode_kernel": head["tg"]["decode_kernel"],
                   "value_is": ("the persistent decode chain: reachable when the caller hands over consecutive MUL_MAT groups (C ABI b200_chain_*, plugin on "
                                "MUL_MAT-only graphs); tg_per_group is the per-node path") if persistent else "one launch per MUL_MAT group"},
        "e2e": {"value": e2e_tg, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "through": "C ABI (ctypes -> libb200mm.so): pinned host x in, logits out, host sync per token"},
        "e2e_plugin": e2e_plugin,
        "llama_bench": llama_bench,
        "gpu_launches": hx["launches"],
        "roofline": roof,
        "tg_per_group": tg_per_group,
        "pp512" if args.pp == 512 else "pp%d" % args.pp: pp,
        "parity": parity,
        "configs": extras,
        "cpu_baseline": cpu, "clocks": clocks,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
