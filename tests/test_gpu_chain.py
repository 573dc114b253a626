"""GPU (-m gpu): the persistent decode chain (b200_chain_*, gemv_chain.cu) through the C ABI.

A chain is a sequence of DEPENDENT MUL_MAT groups in one launch: group g + 1 reads what group g wrote. Every node's output is
checked against the CPU oracle applied to the input that node actually saw on the GPU (so rounding differences do not compound),
at the tolerance of the decode path (the integer arithmetic is the CPU's, only the fp32 summation order differs), and against
the per-node path (b200_mul_mat_group) on the same buffers."""
import numpy as np
import pytest
import torch

from conftest import nmse
from oracle.oracle import Q4_0, Q8_0, Q4_K, Q5_K, Q6_K, IQ4_XS, TYPE_NAMES, random_blocks

pytestmark = pytest.mark.gpu
DECODE_NMSE = 1e-10


def _mk(b200, t, m, k, rng):
    w = random_blocks(t, m, k, rng)
    wd = torch.from_numpy(w).cuda().reshape(-1)
    return w, b200.repack(t, wd, m, k)


def _args(b200, t, wp, x, y, m, col0=0):
    """MulMatArgs writing y[:, col0:col0+m]"""
    n, k = x.shape
    return b200.MulMatArgs(t, b200.LAYOUT_PLANES, m, n, k, wp.data_ptr(), 0, x.data_ptr(), x.stride(0),
                           y.data_ptr() + 4 * col0, y.stride(0), 0, 0, b200.Q8_X86, 0, 0, 0)


def _run_chain(b200, groups, stream=None):
    ch = b200.DecodeChain(0, groups)
    ch.launch(stream)
    torch.cuda.synchronize()
    return ch


# a small "model": x0 -> {A, B, C} -> h1 (concatenated) -> D -> h2 -> {E, F} -> h3 -> G -> out ; every group reads the previous output
CASES = [
    # (name, k0, [(types and row counts of group 1)], m2, [(group 3)], m4)
    ("kquants", 2048, [(Q4_K, 1536), (Q4_K, 256), (Q6_K, 256)], 4096, [(Q4_K, 3584), (Q4_K, 3584)], 2048, Q6_K),
    ("q6k_k14336", 4096, [(Q6_K, 14336)], 4096, [(Q6_K, 1024), (Q4_K, 1024)], 2048, Q4_K),
    ("q6k_70b_shapes", 8192, [(Q6_K, 1024), (Q4_K, 27648)], 2048, [(Q6_K, 512)], 256, Q4_K),      # 32 and 112 units per row: one / two rows per tensor copy
    ("q5k_70b_attn", 8192, [(Q4_K, 2048), (Q4_K, 256), (Q5_K, 256)], 2048, [(Q5_K, 1536), (Q4_K, 512)], 4096, Q5_K),   # 70B recipe: attn_v in Q5_K next to Q4_K
    ("iq4_xs", 1024, [(IQ4_XS, 2048), (IQ4_XS, 512), (Q6_K, 512)], 2048, [(IQ4_XS, 4096)], 1024, IQ4_XS),
    ("q4_0", 2048, [(Q4_0, 2048), (Q4_0, 512), (Q4_0, 512)], 2048, [(Q4_0, 8192), (Q4_0, 8192)], 2048, Q8_0),
    ("q8_0", 1024, [(Q8_0, 1024), (Q8_0, 256)], 512, [(Q8_0, 3072)], 1024, Q4_0),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("n", [1, 2, 4])
def test_chain_matches_oracle_node_by_node(b200, orc, case, n):
    name, k0, g1, m2, g3, m4, t_last = case
    rng = np.random.default_rng(abs(hash(name)) % 2**32 + n)
    t_mid = g1[0][0]
    h1w = sum(m for _, m in g1)
    h3w = sum(m for _, m in g3)
    x0 = torch.from_numpy((rng.standard_normal((n, k0)) * 3).astype(np.float32)).cuda()
    h1 = torch.zeros((n, h1w), device="cuda"); h2 = torch.zeros((n, m2), device="cuda")
    h3 = torch.zeros((n, h3w), device="cuda"); out = torch.zeros((n, m4), device="cuda")
    nodes = []     # (type, canonical w, x tensor, y tensor, col0, m)
    groups = []
    keep = []
    col = 0; grp = []
    for t, m in g1:
        w, wp = _mk(b200, t, m, k0, rng); keep.append(wp)
        grp.append(_args(b200, t, wp, x0, h1, m, col)); nodes.append((t, w, x0, h1, col, m)); col += m
    groups.append(grp)
    w, wp = _mk(b200, t_mid, m2, h1w, rng); keep.append(wp)
    groups.append([_args(b200, t_mid, wp, h1, h2, m2)]); nodes.append((t_mid, w, h1, h2, 0, m2))
    col = 0; grp = []
    for t, m in g3:
        w, wp = _mk(b200, t, m, m2, rng); keep.append(wp)
        grp.append(_args(b200, t, wp, h2, h3, m, col)); nodes.append((t, w, h2, h3, col, m)); col += m
    groups.append(grp)
    w, wp = _mk(b200, t_last, m4, h3w, rng); keep.append(wp)
    groups.append([_args(b200, t_last, wp, h3, out, m4)]); nodes.append((t_last, w, h3, out, 0, m4))

    if name == "q6k_70b_shapes" and n == 4:
        # 4 x 28672 quantised activations leave no ring stage that holds one 23 KB row: refused up front, create agrees
        assert not b200.DecodeChain.supported(groups)
        with pytest.raises(b200.B200Error):
            b200.DecodeChain(0, groups)
        return
    assert b200.DecodeChain.supported(groups)
    ch = _run_chain(b200, groups)
    info = ch.info()
    assert info["n_groups"] == 4 and info["ctas"] >= 1
    for t, w, x, y, c0, m in nodes:
        want = orc.mul_mat(t, w, x.cpu().numpy())
        got = y[:, c0:c0 + m].cpu().numpy()
        assert np.isfinite(got).all()
        e = nmse(got, want)
        assert e < DECODE_NMSE, f"{name} n={n} node {TYPE_NAMES[t]} m={m} k={x.shape[1]}: nmse {e:.3e}"

    # replay (same chain object, counter reset by the launch) is deterministic, bit for bit
    first = [t.clone() for t in (h1, h2, h3, out)]
    for t in (h1, h2, h3, out):
        t.zero_()
    ch.launch(); torch.cuda.synchronize()
    for a, c in zip(first, (h1, h2, h3, out)):
        assert torch.equal(a, c)
    ch.close()


def test_chain_in_cuda_graph_and_vs_per_node_path(b200):
    """8B-shaped layer slice: chain replayed from a CUDA graph == the per-node launches (b200_mul_mat_group) to fp32 rounding"""
    from zyk_llama_cpp_b200 import workloads as W
    from zyk_llama_cpp_b200.chain import MatMulChain
    ops = W.llama_dense(n_layer=2, n_embd=4096, n_ff=14336, n_head=32, n_kv=8, n_vocab=32000, ftype="q4_k_m")
    a = MatMulChain(ops, 1, persistent=True)
    assert a.decode_chain is not None
    b = MatMulChain(ops, 1, persistent=False)
    assert b.decode_chain is None
    a.capture(); assert a.kernels_per_step == 1
    for _ in range(3):
        a.step(); b.step()
    a.stream.synchronize(); b.stream.synchronize()
    for name in ("q", "k", "v", "x2", "gate", "up", "x", "logits"):
        ya, yb = a.bufs[name].cpu().numpy(), b.bufs[name].cpu().numpy()
        assert np.isfinite(ya).all()
        assert nmse(ya, yb) < 1e-9, name


def test_full_size_8b_token_chain_vs_per_node_and_replay(b200):
    """BASELINE.json configs[1] at full size (225 nodes, 4.6 GB of weights): every activation buffer the persistent chain leaves
    equals what the per-node launches leave (fp32 summation order only), and a replay is bit-identical"""
    from zyk_llama_cpp_b200 import workloads as W
    from zyk_llama_cpp_b200.chain import MatMulChain
    ops = W.model_ops("llama-3-8b-q4_k_m")
    a = MatMulChain(ops, 1, persistent=True)
    assert a.decode_chain is not None and a.decode_chain.info()["n_groups"] == 129
    a.capture()
    a.step(); a.stream.synchronize()
    first = {k: v.clone() for k, v in a.bufs.items()}
    a.step(); a.stream.synchronize()
    # x is rewritten by the last ffn_down before the output node reads it, so a second pass sees the same inputs everywhere
    # except layer 0's attn_q/k/v, whose results are overwritten by the later layers: every buffer must repeat exactly
    for k, v in a.bufs.items():
        assert torch.equal(v, first[k]), k
    a.close(); del a
    torch.cuda.empty_cache()
    b = MatMulChain(ops, 1, persistent=False)
    b.step(); b.stream.synchronize()
    for name in ("q", "k", "v", "x2", "gate", "up", "x", "logits"):
        ya, yb = first[name].cpu().numpy(), b.bufs[name].cpu().numpy()
        assert np.isfinite(ya).all() and float(np.abs(ya).max()) > 0
        assert nmse(ya, yb) < 1e-9, name


def test_chain_rejects_what_it_cannot_run(b200):
    x = torch.zeros((1, 300), device="cuda"); y = torch.zeros((1, 64), device="cuda")
    w = torch.zeros(64 * 300, dtype=torch.uint8, device="cuda")
    bad_k = [[b200.MulMatArgs(Q8_0, b200.LAYOUT_PLANES, 64, 1, 288, w.data_ptr(), 0, x.data_ptr(), 300, y.data_ptr(), 64, 0, 0, 1, 0, 0, 0)]]
    assert not b200.DecodeChain.supported(bad_k)
    with pytest.raises(Exception):
        b200.DecodeChain(0, bad_k)
    x8 = torch.zeros((8, 256), device="cuda"); y8 = torch.zeros((8, 64), device="cuda")
    big_n = [[b200.MulMatArgs(Q4_K, b200.LAYOUT_PLANES, 64, 8, 256, w.data_ptr(), 0, x8.data_ptr(), 256, y8.data_ptr(), 64, 0, 0, 1, 0, 0, 0)]]
    assert not b200.DecodeChain.supported(big_n)
