"""Multi-GPU parity check (run under torchrun, one rank per GPU): the row-split chain -- exchange fused into the kernels
(b200_mul_mat_gather) or one NCCL all-gather per node -- must leave in every activation buffer of every rank what the
single-GPU chain computes from the same seeded weights. Decode (n = 1, CUDA graph replay included) and prefill (n = 64).
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/rowsplit_check.py"""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_pkg  # noqa: E402

b = load_pkg()
from zyk_llama_cpp_b200 import workloads as W  # noqa: E402
from zyk_llama_cpp_b200.chain import MatMulChain  # noqa: E402
from zyk_llama_cpp_b200.rowsplit import RowSplit  # noqa: E402


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    rs = RowSplit(world, rank, b.row_split)
    failures = 0
    for model in ("tiny-q4_k_m", "tiny-moe-q4_k_m"):
        ops = W.model_ops(model)
        for n in (1, 64):
            ref = MatMulChain(ops, n, device=local, persistent=False)      # the reference is the per-node single-GPU path
            x0 = ref.bufs["x"].clone()
            ref.step(); ref.stream.synchronize()
            want = {k: v.clone() for k, v in ref.bufs.items()}
            for name, w in want.items():
                if not bool(torch.isfinite(w).all()) or float(w.abs().max()) == 0.0:
                    failures += 1
                    print(f"[rank {rank}] reference buffer {name} of {model} n={n} is not finite / all zero: the check would be vacuous", flush=True)
            for fused, persistent in ((True, True), (True, False), (False, False)):
                if persistent and (n > 4 or "moe" in model):
                    continue
                ch = MatMulChain(ops, n, device=local, rowsplit=rs, fused_exchange=fused, persistent=persistent)
                if persistent and ch.decode_chain is None:
                    failures += 1
                    print(f"[rank {rank}] {model} n={n}: the persistent row-split chain was not created: {b.lib.b200_last_error().decode()}", flush=True)
                modes = ["eager", "eager"] + (["graph", "graph", "graph"] if ch.capturable else [])
                for i, mode in enumerate(modes):
                    if mode == "graph" and ch.graph is None:
                        ch.capture()
                    ch.bufs["x"].copy_(x0)                       # the chain overwrites x: every pass starts from the same input
                    torch.cuda.synchronize(); dist.barrier()
                    ch.step(); ch.stream.synchronize(); torch.cuda.synchronize(); dist.barrier()
                    for name, w in want.items():
                        got = ch.bufs[name]
                        ok = torch.allclose(got, w, rtol=2e-4, atol=1e-4 * float(w.abs().max()), equal_nan=True)
                        if not ok:
                            failures += 1
                            err = float((got - w).abs().max())
                            print(f"[rank {rank}] MISMATCH {model} n={n} exchange={'fused' if fused else 'nccl'} pass {i} ({mode}) buffer {name}: max|d|={err:.3e}", flush=True)
                if rank == 0:
                    print(f"{model} n={n} exchange={'chain' if ch.decode_chain is not None else 'fused' if fused else 'nccl'} world={world}: {len(modes)} passes "
                          f"({'graph replay included' if ch.capturable else 'eager only'}) checked", flush=True)
                ch.close()
                del ch
            del ref
    t = torch.tensor([failures], device=f"cuda:{local}")
    dist.all_reduce(t)
    if rank == 0:
        print("ROWSPLIT_CHECK", "FAILED" if int(t.item()) else "OK", flush=True)
    dist.destroy_process_group()
    return 1 if int(t.item()) else 0


if __name__ == "__main__":
    sys.exit(main())
