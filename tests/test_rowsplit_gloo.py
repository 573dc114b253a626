"""CPU, world_size 2 and 3 over gloo: the host side of the row split.

Each rank owns rows [lo, hi) of W (b200_row_split through the C ABI), computes its slice of Y = W . X with the
oracle (the checker stands in for the CUDA kernel here -- no GPU in this test), and RowSplit.gather must leave the
reference's full MUL_MAT result on every rank. Mirrors how `ggml_cuda_op_mul_mat` splits rows and copies the slices
back (ggml-cuda/ggml-cuda.cu:790-803, 1725-1744).
"""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

Q4_K, Q6_K, Q8_0 = 12, 14, 8


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, errors):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from __graft_entry__ import load_pkg
        b = load_pkg()
        from zyk_llama_cpp_b200.rowsplit import RowSplit
        from oracle.oracle import Oracle, random_blocks
        orc = Oracle()
        rs = RowSplit(world, rank, b.row_split)

        # (1) plain tensors: every (n, m) class the chain produces -- even split, uneven tail, prefill (n > 1), m < tile
        for n, m in [(1, 4096), (1, 128256), (1, 1000), (5, 14336), (3, 384), (2, 100)]:
            g = torch.Generator().manual_seed(1000 * n + m)
            full = torch.randn((n, m + 7), generator=g)                     # 7 columns of padding must stay untouched
            y = torch.full_like(full, -1.0)
            lo, hi = rs.mine(m)
            y[:, lo:hi] = full[:, lo:hi]
            rs.gather(y, m)
            assert torch.equal(y[:, :m], full[:, :m]), (n, m, rank)
            assert torch.all(y[:, m:] == -1.0), (n, m, rank)

        # (2) a quantised MUL_MAT split by rows: slices computed independently, gathered == the full oracle result
        rng = np.random.default_rng(7)                                       # same seed on every rank: same W and X
        for t, m, k, n in [(Q4_K, 512, 512, 1), (Q6_K, 300, 256, 2), (Q8_0, 640, 128, 3)]:
            w = random_blocks(t, m, k, rng).reshape(m, -1)
            x = rng.standard_normal((n, k)).astype(np.float32)
            want = orc.mul_mat(t, w, x)
            lo, hi = rs.mine(m)
            y = torch.zeros((n, m))
            if hi > lo:
                y[:, lo:hi] = torch.from_numpy(orc.mul_mat(t, np.ascontiguousarray(w[lo:hi]), x))
            rs.gather(y, m)
            assert np.array_equal(y.numpy(), want), (t, m, k, n, rank)      # same arithmetic per row: bit-exact
        dist.barrier()
        dist.destroy_process_group()
    except BaseException as e:  # noqa: BLE001
        import traceback
        errors.put(f"rank {rank}: {e!r}\n{traceback.format_exc()}")
        raise


@pytest.mark.parametrize("world", [2, 3])
def test_row_split_gather_gloo(world):
    ctx = mp.get_context("spawn")
    errors = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, errors)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    alive = [p for p in procs if p.is_alive()]
    for p in alive:
        p.kill()
    msgs = []
    while not errors.empty():
        msgs.append(errors.get())
    assert not alive and not msgs and all(p.exitcode == 0 for p in procs), "\n".join(msgs) or f"exit codes {[p.exitcode for p in procs]}"


def test_row_split_bounds_are_tile_aligned():
    from __graft_entry__ import load_pkg
    b = load_pkg()
    from zyk_llama_cpp_b200.rowsplit import RowSplit
    for world in (1, 2, 4, 8):
        for m in (128, 1000, 1024, 4096, 14336, 28672, 128256):
            rs = RowSplit(world, 0, b.row_split)
            bounds = rs.bounds(m)
            assert bounds[0][0] == 0 and bounds[-1][1] == m
            for r, (lo, hi) in enumerate(bounds):
                assert lo <= hi and lo % 128 == 0, (world, m, bounds)
                if r + 1 < world:
                    assert hi == bounds[r + 1][0]


def test_fused_exchange_wait_plan():
    """host logic of the fused exchange: calls that are not gather calls get one wait per run, for every gather call before them"""
    from __graft_entry__ import load_pkg
    load_pkg()
    from zyk_llama_cpp_b200.rowsplit import plan_mid_waits
    G, O = True, False
    assert plan_mid_waits([G, G, G, G]) == {}                                  # dense model: gather calls wait by themselves
    assert plan_mid_waits([O, O, G]) == {}                                     # nothing gathered yet: nothing to wait for
    # MoE layer: qkv, wo | gate, up, down (MUL_MAT_ID) | next layer's qkv, wo | gate ...
    assert plan_mid_waits([G, G, O, O, O, G, G, O, O, O]) == {2: 2, 7: 4}
    assert plan_mid_waits([G, O, G, O]) == {1: 1, 3: 2}
