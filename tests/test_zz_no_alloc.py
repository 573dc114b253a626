"""GPU: b200_mul_mat_noalloc (include/b200_mm.h) -- the decode MUL_MAT whose scratch is a fixed file-scope device array
(no cudaMalloc anywhere in the call) -- gives what b200_mul_mat gives with a caller workspace. No counterpart in the
reference, so the oracle here is the existing entry point. Collected last
on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(b, fn, t, layout, w, x, m, y):
    n, k = x.shape
    a = b.MulMatArgs(t, layout, m, n, k, w.data_ptr(), 0, x.data_ptr(), x.stride(0), y.data_ptr(), y.stride(0),
                     0, 0, b.Q8_X86, 1, 0, 0)
    return a, fn(C.byref(a), None)


@pytest.mark.parametrize("tname", ["Q4_K", "Q5_K", "Q6_K", "Q8_0", "Q4_0", "IQ4_XS"])
def test_noalloc_matches_workspace_entry_point_twice_in_a_row(b200, orc, tname):
    import torch
    from oracle.oracle import random_blocks
    import oracle.oracle as O
    b = b200
    t = getattr(O, tname)
    rng = np.random.default_rng(11)
    dev = torch.device("cuda:0")
    m, k = 384, 2048
    w = torch.from_numpy(random_blocks(t, m, k, rng)).to(dev).reshape(-1)
    wp = b.repack(t, w, m, k)
    # two calls in a row on the same stream share the array: different n and X, the second must not see the first's leftovers
    for n, layout, wt in ((3, b.LAYOUT_PLANES, wp), (1, b.LAYOUT_PLANES, wp), (5, b.LAYOUT_CANONICAL, w)):
        x = torch.from_numpy(rng.standard_normal((n, k)).astype(np.float32)).to(dev)
        want = b.mul_mat(t, wt, x, m, layout=layout, algo=1)
        got = torch.full((n, m), float("nan"), device=dev)
        _, rc = _call(b, b.lib.b200_mul_mat_noalloc, t, layout, wt, x, m, got)
        assert rc == 0, b.lib.b200_last_error().decode()
        torch.cuda.synchronize()
        # same kernels, same arithmetic, only the scratch differs: bit-identical (tolerance of the decode tests: NMSE < 1e-10)
        assert torch.equal(got, want), f"{tname} n={n}: max |diff| {(got - want).abs().max().item():.3e}"
        ref = orc.mul_mat(t, w.cpu().numpy(), x.cpu().numpy())
        err = float(np.sum((got.cpu().numpy() - ref) ** 2) / np.sum(ref ** 2))
        assert err < 1e-10, f"{tname} n={n}: nmse {err:.3e} vs oracle"


def test_noalloc_rejects_shapes_beyond_the_fixed_scratch(b200):
    import torch
    import oracle.oracle as O
    b = b200
    dev = torch.device("cuda:0")
    cap = int(b.lib.b200_noalloc_scratch_max())
    assert cap == 1 << 20
    k, n = 8192, 256                      # 256 x 8192 int8 alone is 2 MiB
    assert int(b.lib.b200_mul_mat_noalloc_scratch(O.Q4_K, n, k)) > cap
    w = torch.zeros(16 * (k // 256) * 144, dtype=torch.uint8, device=dev)
    x = torch.zeros((n, k), device=dev)
    y = torch.zeros((n, 16), device=dev)
    _, rc = _call(b, b.lib.b200_mul_mat_noalloc, O.Q4_K, b.LAYOUT_CANONICAL, w, x, 16, y)
    assert rc != 0
    msg = b.lib.b200_last_error().decode()
    assert "scratch" in msg and str(cap) in msg
