"""GPU (-m gpu): the CUDA path through the C ABI against the CPU oracle and the committed golden
vectors of the reference. Bit-exact for activation quantisation and repack; MUL_MAT within the
reference's own tolerance (NMSE <= 5e-4, tests/test-backend-ops.cpp:3716-3718, 3852-3854) -- and
in fact ~1e-12 for the decode path, because the integer arithmetic is the CPU's."""
from pathlib import Path

import numpy as np
import pytest
import torch

from conftest import nmse
from oracle.oracle import (WEIGHT_TYPES, TYPE_NAMES, Q8_0, Q8_1, Q8_K, Q4_K, Q5_K, Q6_K, Q8_REF, Q8_X86, random_blocks)

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
MAX_NMSE = 5e-4          # the reference's tolerance for MUL_MAT / MUL_MAT_ID
DECODE_NMSE = 1e-10      # what the dp4a decode path actually achieves against the CPU arithmetic


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _acts(rng, n, k):
    x = (rng.standard_normal((n, k)) * rng.uniform(0.01, 30, (n, 1))).astype(np.float32)
    if n > 2 and k >= 256:
        x[0, :256] = 0
        x[1] = np.round(x[1] * 4) / 4
        x[2, :32] = np.arange(32, dtype=np.float32) + 0.5
        x[2, 0] = 127.0
        x[2, 40] = -x[2, 33]
    return x


# ---- (3) activation quantiser: bit-exact -----------------------------------------------------
@pytest.mark.parametrize("q8", [Q8_0, Q8_1, Q8_K])
def test_quantise_golden_bit_exact(b200, golden, q8):
    name = TYPE_NAMES[q8]
    x = dev(golden["x"])
    assert np.array_equal(b200.quantize_rows(q8, x, b200.Q8_X86).cpu().numpy(), golden[f"xq_cpu_{name}"])
    if f"xq_ref_{name}" in golden:
        assert np.array_equal(b200.quantize_rows(q8, x, b200.Q8_REF).cpu().numpy(), golden[f"xq_ref_{name}"])


@pytest.mark.parametrize("q8", [Q8_0, Q8_1, Q8_K])
@pytest.mark.parametrize("shape", [(1, 256), (7, 4096), (33, 14336), (512, 4096)])
def test_quantise_vs_oracle_bit_exact(b200, orc, q8, shape):
    rng = np.random.default_rng(hash((q8,) + shape) % 2**32)
    x = _acts(rng, *shape)
    for fl in (Q8_X86, Q8_REF):
        got = b200.quantize_rows(q8, dev(x), fl).cpu().numpy()
        assert np.array_equal(got, orc.quantize(q8, x, fl)), f"{TYPE_NAMES[q8]} flavour {fl} shape {shape}"


@pytest.mark.parametrize("q8", [Q8_0, Q8_1])
def test_quantise_ragged_k_and_strided_rows(b200, orc, q8):
    rng = np.random.default_rng(5)
    big = _acts(rng, 9, 1000)                       # rows of 1000 floats, only the first 992/32 used
    for k in (32, 96, 992):
        xs = dev(big)[:, :k]                        # row stride 1000: unaligned, non-contiguous
        got = b200.quantize_rows(q8, xs, Q8_X86).cpu().numpy()
        assert np.array_equal(got, orc.quantize(q8, big[:, :k].copy(), Q8_X86))


def test_quantise_empty(b200):
    x = torch.empty((0, 256), dtype=torch.float32, device="cuda")
    assert b200.quantize_rows(Q8_K, x).shape == (0, 292)


# ---- (5) repack: bit-exact round trip, same bytes as the oracle's layout ------------------------
@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("shape", [(1, 256), (16, 512), (96, 4096), (1024, 14336)])
def test_repack_bit_exact(b200, orc, t, shape):
    m, k = shape
    rng = np.random.default_rng(t * 1000 + m)
    w = random_blocks(t, m, k, rng)
    wd = dev(w).reshape(-1)
    p = b200.repack(t, wd, m, k)
    assert np.array_equal(p.cpu().numpy(), orc.b200_repack(t, w.reshape(-1), m, k))
    assert torch.equal(b200.unrepack(t, p, m, k), wd)


def test_repack_refuses_unaligned_k(b200):
    assert not b200.repack_supported(Q8_0, 16, 96)
    w = torch.zeros(16 * 3 * 34, dtype=torch.uint8, device="cuda")
    with pytest.raises(b200.B200Error):
        b200.repack(Q8_0, w, 16, 96)


# ---- (1) decode MUL_MAT ----------------------------------------------------------------------------
@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("tag", ["quant", "rand"])
def test_mul_mat_golden(b200, golden, t, tag):
    name = TYPE_NAMES[t]
    w, x, want = golden[f"w_{name}_{tag}"], golden["x"], golden[f"y_{name}_{tag}"]
    m, k = w.shape[0], x.shape[1]
    wd = dev(w).reshape(-1)
    got_c = b200.mul_mat(t, wd, dev(x), m).cpu().numpy()
    got_p = b200.mul_mat(t, b200.repack(t, wd, m, k), dev(x), m, layout=b200.LAYOUT_PLANES).cpu().numpy()
    assert nmse(got_c, want) < DECODE_NMSE and nmse(got_p, want) < DECODE_NMSE


@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 6, 7, 8, 9, 17])
def test_mul_mat_decode_all_batches(b200, orc, t, n):
    rng = np.random.default_rng(t * 100 + n)
    m, k = 131, 1024                                   # m not a multiple of the rows-per-block
    w = random_blocks(t, m, k, rng)
    x = _acts(rng, n, k)
    want = orc.mul_mat(t, w, x)
    wd = dev(w).reshape(-1)
    for layout, wl in ((b200.LAYOUT_CANONICAL, wd), (b200.LAYOUT_PLANES, b200.repack(t, wd, m, k))):
        got = b200.mul_mat(t, wl, dev(x), m, layout=layout, algo=1).cpu().numpy()
        assert nmse(got, want) < DECODE_NMSE, f"{TYPE_NAMES[t]} n={n} layout={layout}"


@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("tune", [0x0101, 0x0201, 0x0401, 0x0102, 0x0202, 0x0402, 0x0104, 0x0204])
def test_mul_mat_every_rows_unroll_variant(b200, orc, t, tune):
    rng = np.random.default_rng(tune)
    m, k = 77, 2304                                    # 9 super-blocks: ragged against U and the warp step
    w = random_blocks(t, m, k, rng)
    x = _acts(rng, 1, k)
    want = orc.mul_mat(t, w, x)
    wp = b200.repack(t, dev(w).reshape(-1), m, k)
    got = b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES, tune=tune).cpu().numpy()
    assert nmse(got, want) < DECODE_NMSE


@pytest.mark.parametrize("t,k", [(2, 32), (2, 96), (8, 32), (8, 160), (12, 256), (13, 256), (14, 256), (23, 256), (2, 8192), (8, 576)])
def test_mul_mat_small_and_ragged_k(b200, orc, t, k):
    rng = np.random.default_rng(k + t)
    m = 16
    w = random_blocks(t, m, k, rng)
    x = rng.standard_normal((2, k)).astype(np.float32)
    got = b200.mul_mat(t, dev(w).reshape(-1), dev(x), m).cpu().numpy()
    assert nmse(got, orc.mul_mat(t, w, x)) < DECODE_NMSE


@pytest.mark.parametrize("t", WEIGHT_TYPES)
def test_mul_mat_strided_operands(b200, orc, t):
    """W as a row-strided view (test-backend-ops k_v cases), X and Y with padded rows."""
    rng = np.random.default_rng(t + 9)
    m, k, kv, n = 24, 512, 768, 3
    wbig = random_blocks(t, m, kv, rng)
    rs_k, rs_kv = b200.row_size(t, k), b200.row_size(t, kv)
    w = wbig[:, :rs_k].copy()
    x = rng.standard_normal((n, k + 64)).astype(np.float32)
    y = torch.full((n, m + 5), 7.0, device="cuda")
    b200.mul_mat(t, dev(wbig).reshape(-1), dev(x)[:, :k], m, w_row_stride=rs_kv, y=y[:, :m])
    assert nmse(y[:, :m].cpu().numpy(), orc.mul_mat(t, w, x[:, :k].copy())) < DECODE_NMSE
    assert torch.all(y[:, m:] == 7.0)                   # nothing written past a row (sentinel check)


@pytest.mark.parametrize("t", WEIGHT_TYPES)
def test_mul_mat_within_reference_tolerance_of_f64_truth(b200, orc, t):
    rng = np.random.default_rng(t)
    m, k, n = 64, 4096, 4
    w = random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)   # test-backend-ops' input distribution
    got = b200.mul_mat(t, b200.repack(t, dev(w).reshape(-1), m, k), dev(x), m, layout=b200.LAYOUT_PLANES).cpu().numpy()
    assert nmse(got, orc.mul_mat_f64(t, w, x)) < MAX_NMSE


# ---- (4) MUL_MAT_ID ---------------------------------------------------------------------------------
@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("nb1", [1, 2])
def test_mul_mat_id_golden(b200, golden, t, nb1):
    name = TYPE_NAMES[t]
    as_, ids, b = golden[f"id_as_{name}"], golden[f"id_ids_{name}"], golden[f"id_b{nb1}_{name}"]
    m = as_.shape[1]
    got = b200.mul_mat_id(t, dev(as_).reshape(as_.shape[0], -1), dev(b), dev(ids), m).cpu().numpy()
    assert nmse(got, golden[f"id_y{nb1}_{name}"]) < DECODE_NMSE


@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("cfg", [(4, 1, 1), (8, 2, 5), (8, 4, 17), (8, 2, 129)])
def test_mul_mat_id_vs_oracle(b200, orc, t, cfg):
    n_as, n_used, n_tok = cfg
    rng = np.random.default_rng(sum(cfg) + t)
    m, k = 40, 512
    as_ = np.stack([random_blocks(t, m, k, rng) for _ in range(n_as)])
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    for nb1 in (1, n_used):
        b = rng.standard_normal((n_tok, nb1, k)).astype(np.float32)
        want = orc.mul_mat_id(t, as_, b, ids)
        flat = dev(as_).reshape(n_as, -1)
        got = b200.mul_mat_id(t, flat, dev(b), dev(ids), m, algo=1).cpu().numpy()
        assert nmse(got, want) < DECODE_NMSE
        planes = torch.stack([b200.repack(t, flat[e].contiguous(), m, k) for e in range(n_as)])
        got = b200.mul_mat_id(t, planes, dev(b), dev(ids), m, layout=b200.LAYOUT_PLANES, algo=1).cpu().numpy()
        assert nmse(got, want) < DECODE_NMSE
        # auto: the per-pair GEMV (CPU integer arithmetic) or, from eight pairs per expert on, the expert-gathered tensor-core GEMM
        got = b200.mul_mat_id(t, planes, dev(b), dev(ids), m, layout=b200.LAYOUT_PLANES).cpu().numpy()
        gemm = n_tok * n_used > 8 and n_tok * n_used >= 8 * min(n_as, n_tok * n_used)
        assert nmse(got, want) < (MAX_NMSE if gemm else DECODE_NMSE)


def _id_truth(orc, t, as_, b, ids, k):
    """exact arithmetic on the dequantised expert weights (float64)"""
    n_tok, n_used = ids.shape
    n_as, m = as_.shape[0], as_.shape[1]
    wf = [orc.dequantize(t, as_[e], k).reshape(m, k).astype(np.float64) for e in range(n_as)]
    y = np.zeros((n_tok, n_used, m))
    for tk in range(n_tok):
        for sl in range(n_used):
            e = int(ids[tk, sl])
            if 0 <= e < n_as:
                y[tk, sl] = wf[e] @ b[tk, sl % b.shape[1]].astype(np.float64)
    return y


# the reference's MUL_MAT_ID matrix (tests/test-backend-ops.cpp:7781-7811: n_used 1 / 2 / 4, b with one row per token or one per slot,
# n in 1 .. 129) plus the prefill ubatch, forced onto the expert-gathered tensor-core GEMM (algo = 2)
@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("n_used", [1, 2, 4])
@pytest.mark.parametrize("n_tok", [1, 4, 5, 17, 32, 129, 512])
def test_mul_mat_id_gemm_vs_oracle(b200, orc, t, n_used, n_tok):
    n_as = 8
    rng = np.random.default_rng(1000 * t + 10 * n_tok + n_used)
    m, k = (300, 512) if n_tok < 512 else (136, 256)          # 300 rows: CTA pairs with a ragged last tile; 136: keeps the 2048-pair oracle run short
    as_ = np.stack([random_blocks(t, m, k, rng) for _ in range(n_as)])
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    planes = torch.stack([b200.repack(t, dev(as_[e]).reshape(-1), m, k) for e in range(n_as)])
    for nb1 in sorted({1, n_used}):
        b = _acts(rng, n_tok * nb1, k).reshape(n_tok, nb1, k)
        got = b200.mul_mat_id(t, planes, dev(b), dev(ids), m, layout=b200.LAYOUT_PLANES, algo=2).cpu().numpy()
        assert np.isfinite(got).all()
        assert nmse(got, orc.mul_mat_id(t, as_, b, ids)) < MAX_NMSE
        assert nmse(got, _id_truth(orc, t, as_, b, ids, k)) < GEMM_VS_TRUTH


def test_mul_mat_id_gemm_single_cta_tiles_skewed_routing_and_invalid_ids(b200, orc):
    """m <= 128 (no CTA pair), every token routed to the same two experts (one expert needs several n-tiles, six have no rows), ids out
    of range (their Y rows are zero, as on the GEMV path), Y / b / ids with padded strides"""
    t, n_as, n_used, n_tok, m, k = Q4_K, 8, 2, 300, 96, 512
    rng = np.random.default_rng(5)
    as_ = np.stack([random_blocks(t, m, k, rng) for _ in range(n_as)])
    ids = np.tile(np.array([[5, 2]], np.int32), (n_tok, 1))
    ids[7, 0] = -1; ids[11, 1] = n_as + 3; ids[299, 0] = 0
    b = _acts(rng, n_tok * n_used, k).reshape(n_tok, n_used, k)
    planes = torch.stack([b200.repack(t, dev(as_[e]).reshape(-1), m, k) for e in range(n_as)])
    got = b200.mul_mat_id(t, planes, dev(b), dev(ids), m, layout=b200.LAYOUT_PLANES, algo=2).cpu().numpy()
    want = orc.mul_mat_id(t, as_, b, ids)
    assert nmse(got, want) < MAX_NMSE
    assert not got[7, 0].any() and not got[11, 1].any()
    assert nmse(got, _id_truth(orc, t, as_, b, ids, k)) < GEMM_VS_TRUTH
    # the same values whichever path computes them (up to the GEMM's rounding): GEMV vs GEMM
    ref = b200.mul_mat_id(t, planes, dev(b), dev(ids), m, layout=b200.LAYOUT_PLANES, algo=1).cpu().numpy()
    assert nmse(got, ref) < MAX_NMSE


def test_mul_mat_id_gemm_full_size_mixtral_sampled_rows(b200, orc):
    """BASELINE configs[3]: Mixtral expert matrices (8 x 14336 x 4096 Q4_K), pp512 ubatch with 2 experts per token, sampled rows vs the oracle"""
    t, n_as, n_used, n_tok, m, k = Q4_K, 8, 2, 512, 14336, 4096
    from zyk_llama_cpp_b200 import workloads as W
    gen = torch.Generator(device="cuda"); gen.manual_seed(3)
    planes = torch.stack([W.random_weight(t, m, k, torch.device("cuda"), gen) for _ in range(n_as)])     # canonical Q4_K is its own plane layout
    rng = np.random.default_rng(9)
    ids = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    b = rng.standard_normal((n_tok, 1, k)).astype(np.float32)
    got = b200.mul_mat_id(t, planes, dev(b), dev(ids), m, layout=b200.LAYOUT_PLANES).cpu().numpy()          # auto: 1024 pairs over 8 experts -> GEMM
    rows = rng.choice(m, 24, replace=False)
    rb = k // 256 * 144
    for e in range(n_as):
        sub = planes[e].view(m, rb)[torch.from_numpy(rows).cuda()].cpu().numpy().reshape(-1)
        toks, slots = np.nonzero(ids == e)
        if len(toks) == 0:
            continue
        pick = slice(0, min(len(toks), 40))
        want = orc.mul_mat_f64(t, sub, np.ascontiguousarray(b[toks[pick], 0]))
        assert nmse(got[toks[pick], slots[pick]][:, rows], want) < GEMM_VS_TRUTH


# ---- full-size properties (BASELINE.json shapes; the oracle would take minutes here) -----------------
@pytest.mark.parametrize("t,m,k", [(Q4_K, 14336, 4096), (Q4_K, 4096, 14336), (Q6_K, 4096, 14336), (Q6_K, 8192, 4096)])
def test_full_size_linearity_and_sampled_rows(b200, orc, t, m, k):
    rng = np.random.default_rng(m + k)
    w = random_blocks(t, m, k, rng)
    wp = b200.repack(t, dev(w).reshape(-1), m, k)
    x = rng.standard_normal((2, k)).astype(np.float32)
    y = b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES).cpu().numpy()
    rows = rng.choice(m, 48, replace=False)
    assert nmse(y[:, rows], orc.mul_mat(t, w[rows], x)) < DECODE_NMSE
    # positive homogeneity by a power of two is exact through the quantiser: Y(4x) == 4 Y(x) bit for bit
    y4 = b200.mul_mat(t, wp, dev(x * 4), m, layout=b200.LAYOUT_PLANES).cpu().numpy()
    assert np.array_equal(y4, y * 4)
    # canonical and plane layouts run different kernels over the same integers
    yc = b200.mul_mat(t, dev(w).reshape(-1), dev(x), m).cpu().numpy()
    assert nmse(yc, y) < DECODE_NMSE


def test_cuda_graph_replay_and_pdl(b200, orc):
    rng = np.random.default_rng(11)
    m, k = 512, 4096
    w = random_blocks(Q4_K, m, k, rng)
    x = rng.standard_normal((1, k)).astype(np.float32)
    s = b200.Stream(0)
    wd, xd = dev(w).reshape(-1), dev(x)
    y = torch.zeros((1, m), device="cuda")
    ws = torch.empty(b200.mul_mat_workspace(Q4_K, 1, k) + 256, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    with b200.Graph(s) as g:
        for _ in range(3):
            b200.mul_mat(Q4_K, wd, xd, m, y=y, workspace=ws, flags=b200.MM_PDL, stream=s)
    for _ in range(2):
        g.launch()
    s.synchronize()
    assert nmse(y.cpu().numpy(), orc.mul_mat(Q4_K, w, x)) < DECODE_NMSE


# ---- (2) prefill GEMM (tcgen05): within the reference's tolerance of the CPU path, and much closer to the f64 truth ------
GEMM_VS_TRUTH = 2e-6     # fp16 operands, fp32 accumulation in TMEM: what the kernel actually achieves vs exact arithmetic


@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("n", [9, 16, 33, 128, 256, 300, 512])
def test_gemm_vs_oracle(b200, orc, t, n):
    rng = np.random.default_rng(t * 1000 + n)
    m, k = 256, 1024
    w = random_blocks(t, m, k, rng)
    x = _acts(rng, n, k)
    wp = b200.repack(t, dev(w).reshape(-1), m, k)
    got = b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES, algo=2).cpu().numpy()
    assert np.isfinite(got).all()
    assert nmse(got, orc.mul_mat(t, w, x)) < MAX_NMSE          # vs the CPU's quantised-activation arithmetic
    assert nmse(got, orc.mul_mat_f64(t, w, x)) < GEMM_VS_TRUTH  # vs exact arithmetic on the dequantised weights


@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("shape", [(131, 256, 40), (128, 2304, 17), (1000, 512, 77), (384, 4096, 512)])
def test_gemm_ragged_shapes_and_strides(b200, orc, t, shape):
    m, k, n = shape
    rng = np.random.default_rng(m + k + n + t)
    w = random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k + 32)).astype(np.float32)     # padded activation rows (row stride != k)
    y = torch.full((n, m + 3), 7.0, device="cuda")
    wp = b200.repack(t, dev(w).reshape(-1), m, k)
    b200.mul_mat(t, wp, dev(x)[:, :k], m, layout=b200.LAYOUT_PLANES, algo=2, y=y[:, :m])
    assert nmse(y[:, :m].cpu().numpy(), orc.mul_mat_f64(t, w, x[:, :k].copy())) < GEMM_VS_TRUTH
    assert torch.all(y[:, m:] == 7.0)                           # sentinel columns untouched


def test_gemm_is_the_default_for_prefill_batches(b200, orc):
    rng = np.random.default_rng(3)
    m, k, n = 512, 2048, 64
    w = random_blocks(Q4_K, m, k, rng)
    x = rng.standard_normal((n, k)).astype(np.float32)
    wd = dev(w).reshape(-1)
    auto = b200.mul_mat(Q4_K, wd, dev(x), m).cpu().numpy()                 # canonical Q4_K is its own plane layout
    gemv = b200.mul_mat(Q4_K, wd, dev(x), m, algo=1).cpu().numpy()
    assert not np.array_equal(auto, gemv)                                   # different kernels ...
    assert nmse(auto, gemv) < MAX_NMSE                                      # ... same answer within tolerance
    assert nmse(auto, orc.mul_mat_f64(Q4_K, w, x)) < GEMM_VS_TRUTH


def test_gemm_huge_activation_magnitudes_do_not_overflow_fp16(b200, orc):
    rng = np.random.default_rng(4)
    m, k, n = 128, 512, 32
    w = random_blocks(Q6_K, m, k, rng)
    x = (rng.standard_normal((n, k)) * np.logspace(-20, 20, n)[:, None]).astype(np.float32)
    got = b200.mul_mat(Q6_K, b200.repack(Q6_K, dev(w).reshape(-1), m, k), dev(x), m, layout=b200.LAYOUT_PLANES, algo=2).cpu().numpy()
    want = orc.mul_mat_f64(Q6_K, w, x)
    assert np.isfinite(got).all()
    for r in range(n):
        assert nmse(got[r], want[r]) < GEMM_VS_TRUTH


# ---- (1) streaming decode GEMV (persistent CTAs, bulk-copy ring, fused activation quantiser) ------------------------
@pytest.mark.parametrize("t", WEIGHT_TYPES)
@pytest.mark.parametrize("n", [1, 2, 3, 4])
@pytest.mark.parametrize("mk", [(77, 2048), (1000, 4096), (2500, 6144)])
def test_stream_gemv_vs_oracle(b200, orc, t, n, mk):
    m, k = mk
    rng = np.random.default_rng(t * 31 + n * 7 + m)
    w = random_blocks(t, m, k, rng)
    x = _acts(rng, n, k)
    want = orc.mul_mat(t, w, x)
    wp = b200.repack(t, dev(w).reshape(-1), m, k)
    if t == Q5_K:       # not in the fused kernel (register GEMV only): a forced request is refused, never served by something else
        with pytest.raises(b200.B200Error):
            b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES, algo=3)
        assert nmse(b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES, algo=1).cpu().numpy(), want) < DECODE_NMSE
        return
    got = b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES, algo=3).cpu().numpy()
    assert nmse(got, want) < DECODE_NMSE
    # the register GEMV runs the same integer arithmetic: the two kernels agree to summation order
    reg = b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES, algo=1).cpu().numpy()
    assert nmse(got, reg) < DECODE_NMSE


def test_stream_gemv_group_mixed_types(b200, orc):
    """wq (Q4_K) + wk (Q4_K) + wv (Q6_K) against one X in one launch, as in a Q4_K_M attention block."""
    rng = np.random.default_rng(21)
    k = 4096
    ts, ms = [Q4_K, Q4_K, Q6_K], [4096, 1024, 1016]
    ws = [random_blocks(t, m, k, rng) for t, m in zip(ts, ms)]
    for n in (1, 3):
        x = _acts(rng, n, k) if n > 2 else rng.standard_normal((n, k)).astype(np.float32)
        wps = [b200.repack(t, dev(w).reshape(-1), m, k) for t, w, m in zip(ts, ws, ms)]
        before = b200.launch_count()
        ys = b200.mul_mat_group(ts, wps, dev(x), ms)
        # n * k <= 8192: ONE kernel for the three nodes, quantiser included; beyond that one quantiser launch shared by three GEMVs
        assert b200.launch_count() - before == (1 if n * k <= 8192 else 4)
        rows = rng.choice(1016, 64, replace=False)
        for t, w, y in zip(ts, ws, ys):
            assert nmse(y.cpu().numpy()[:, rows], orc.mul_mat(t, w[rows], x)) < DECODE_NMSE


@pytest.mark.parametrize("n", [33, 512])
def test_gemm_group_one_launch_per_weight_type(b200, orc, n):
    """prefill: wq + wk (Q4_K) share ONE tensor-core GEMM launch (one tile list over both matrices), wv (Q6_K) takes its own; ragged
    row counts (1000, 72) and a k-split tail included. Compared with the CPU arithmetic at the reference's tolerance and with
    the one-node-per-call results."""
    rng = np.random.default_rng(23 + n)
    k = 2048
    ts, ms = [Q4_K, Q4_K, Q6_K, Q4_K], [1000, 4096, 520, 72]
    ws = [random_blocks(t, m, k, rng) for t, m in zip(ts, ms)]
    x = _acts(rng, n, k)
    wps = [b200.repack(t, dev(w).reshape(-1), m, k) for t, w, m in zip(ts, ws, ms)]
    before = b200.launch_count()
    ys = b200.mul_mat_group(ts, wps, dev(x), ms)
    launches = b200.launch_count() - before
    assert 3 <= launches <= 3 + len(ts)               # fp16 prologue + one GEMM per weight type (+ a zero-fill launch per matrix with k-split tiles)
    for t, w, wp, m, y in zip(ts, ws, wps, ms, ys):
        got = y.cpu().numpy()
        assert np.isfinite(got).all()
        rows = rng.choice(m, min(m, 48), replace=False)
        assert nmse(got[:, rows], orc.mul_mat(t, w[rows], x)) < MAX_NMSE
        single = b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES).cpu().numpy()
        assert nmse(got, single) < 1e-10          # same tiles, same arithmetic; only the k-split partition may differ


def test_stream_gemv_leaves_quantised_x_for_prequantised_calls(b200, orc):
    rng = np.random.default_rng(22)
    m, k = 512, 4096
    w = random_blocks(Q4_K, m, k, rng)
    x = rng.standard_normal((1, k)).astype(np.float32)
    wd, xd = dev(w).reshape(-1), dev(x)
    ws = torch.zeros(b200.mul_mat_workspace(Q4_K, 1, k) + 256, dtype=torch.uint8, device="cuda")
    y1 = b200.mul_mat(Q4_K, wd, xd, m, workspace=ws, algo=3).cpu().numpy()
    y2 = b200.mul_mat(Q4_K, wd, xd, m, workspace=ws, algo=1, flags=b200.MM_X_PREQUANTISED).cpu().numpy()
    want = orc.mul_mat(Q4_K, w, x)
    assert nmse(y1, want) < DECODE_NMSE and nmse(y2, want) < DECODE_NMSE


def test_stream_gemv_full_size_baseline_shapes(b200, orc):
    rng = np.random.default_rng(23)
    for t, m, k in [(Q4_K, 14336, 4096), (Q6_K, 4096, 14336), (Q6_K, 128256, 4096)]:
        w = random_blocks(t, m, k, rng)
        x = rng.standard_normal((1, k)).astype(np.float32)
        y = b200.mul_mat(t, b200.repack(t, dev(w).reshape(-1), m, k), dev(x), m, layout=b200.LAYOUT_PLANES, algo=3).cpu().numpy()
        rows = np.concatenate([rng.choice(m, 40, replace=False), [0, m - 1]])
        assert nmse(y[:, rows], orc.mul_mat(t, w[rows], x)) < DECODE_NMSE


def test_gemm_full_size_baseline_shapes_sampled_rows(b200, orc):
    """The launches that carry the pp512 number (BASELINE configs[1]), at full size, n = 512: ffn_gate + ffn_up as ONE grouped launch on
    CTA pairs (2 x 14336 rows, k-split tail tiles), ffn_down (k = 14336, Q6_K), and the output matrix (128256 rows). Sampled rows and
    activation rows against exact arithmetic on the dequantised weights and against the CPU's quantised-activation arithmetic."""
    from zyk_llama_cpp_b200 import workloads as W
    gen = torch.Generator(device="cuda"); gen.manual_seed(11)
    rng = np.random.default_rng(31)
    n = 512
    cols = np.concatenate([rng.choice(n, 6, replace=False), [0, n - 1]])

    def check(t, wdev, m, k, x, y):
        rows = np.concatenate([rng.choice(m, 30, replace=False), [0, m - 1, (m // 256) * 256 - 1, min(m - 1, (m // 256) * 256)]])
        rb = W.row_bytes(t, k)
        sub = wdev.view(m, rb)[torch.from_numpy(rows).cuda()].cpu().numpy().reshape(-1)
        got = y[torch.from_numpy(cols).cuda()][:, torch.from_numpy(rows).cuda()].cpu().numpy()
        xs = np.ascontiguousarray(x[cols])
        assert np.isfinite(got).all()
        assert nmse(got, orc.mul_mat_f64(t, sub, xs)) < GEMM_VS_TRUTH, (t, m, k)
        assert nmse(got, orc.mul_mat(t, sub, xs)) < MAX_NMSE, (t, m, k)

    # gate + up grouped (the same X, one tile list over both matrices)
    k = 4096
    x = rng.standard_normal((n, k)).astype(np.float32)
    wg, wu = (W.random_weight(Q4_K, 14336, k, torch.device("cuda"), gen) for _ in range(2))       # canonical Q4_K is its own plane layout
    yg, yu = b200.mul_mat_group([Q4_K, Q4_K], [wg, wu], dev(x), [14336, 14336])
    check(Q4_K, wg, 14336, k, x, yg)
    check(Q4_K, wu, 14336, k, x, yu)
    del wg, wu, yg, yu
    # ffn_down of a use_more_bits layer, and the output matrix
    for t, m, k in ((Q6_K, 4096, 14336), (Q6_K, 128256, 4096), (Q4_K, 4096, 14336)):
        x = rng.standard_normal((n, k)).astype(np.float32)
        w = W.random_weight(t, m, k, torch.device("cuda"), gen)
        wp = b200.repack(t, w, m, k)
        y = b200.mul_mat(t, wp, dev(x), m, layout=b200.LAYOUT_PLANES)
        check(t, w, m, k, x, y)
        del w, wp, y


@pytest.mark.parametrize("ks", [2, 3, 8])
def test_gemm_split_k(b200, orc, monkeypatch, ks):
    """k split over gridDim.z, partial tiles reduced into a pre-zeroed Y (what small-m prefill shapes use to fill 148 SMs)."""
    monkeypatch.setenv("B200_GEMM_KSPLIT", str(ks))
    rng = np.random.default_rng(ks)
    m, k, n = 200, 2048, 70
    for t in (Q4_K, Q6_K):
        w = random_blocks(t, m, k, rng)
        x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
        y = torch.full((n, m + 2), 5.0, device="cuda")
        b200.mul_mat(t, b200.repack(t, dev(w).reshape(-1), m, k), dev(x), m, layout=b200.LAYOUT_PLANES, algo=2, y=y[:, :m])
        assert nmse(y[:, :m].cpu().numpy(), orc.mul_mat_f64(t, w, x)) < GEMM_VS_TRUTH
        assert torch.all(y[:, m:] == 5.0)


# ------------------------------------------------------------------------------------------------
# row split over two GPUs: only where the box has them
# ------------------------------------------------------------------------------------------------
def test_row_split_two_gpus_matches_single_gpu_chain(b200):
    """tools/rowsplit_check.py under torchrun: fused exchange and NCCL all-gather, eager and graph replay, decode and prefill,
    dense and MoE, every activation buffer against the single-GPU chain built from the same seeded weights."""
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", str(ROOT / "tools" / "rowsplit_check.py")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "ROWSPLIT_CHECK OK" in r.stdout, r.stdout[-3000:]


def test_gemm_group_two_weight_types_captured_in_a_graph(b200, orc):
    """prefill group with two weight types: the second bucket runs on the library's side stream (fork after the fp16 prologue,
    join before the call returns). The call must be capturable and the replay must give what the eager call gave."""
    rng = np.random.default_rng(77)
    n, k = 96, 2048
    ts, ms = [Q4_K, Q6_K, Q4_K], [640, 384, 200]
    ws = [random_blocks(t, m, k, rng) for t, m in zip(ts, ms)]
    x = dev(_acts(rng, n, k))
    wps = [b200.repack(t, dev(w).reshape(-1), m, k) for t, w, m in zip(ts, ws, ms)]
    ys = [torch.zeros((n, m), dtype=torch.float32, device="cuda") for m in ms]
    need = max(b200.mul_mat_workspace(t, n, k) for t in ts) + 256
    workspace = torch.empty(need, dtype=torch.uint8, device="cuda")
    wsp = (workspace.data_ptr() + 255) & ~255
    arr = (b200.MulMatArgs * len(ts))()
    for i, (t, wp, m, y) in enumerate(zip(ts, wps, ms, ys)):
        arr[i] = b200.MulMatArgs(t, b200.LAYOUT_PLANES, m, n, k, wp.data_ptr(), 0, x.data_ptr(), x.stride(0), y.data_ptr(), y.stride(0),
                                 wsp, need - (wsp - workspace.data_ptr()), b200.Q8_X86, 0, b200.MM_PDL, 0)
    s = b200.Stream(0)
    torch.cuda.synchronize()
    assert b200.lib.b200_mul_mat_group(arr, len(ts), s.handle) == 0
    s.synchronize()
    eager = [y.clone() for y in ys]
    for t, w, y in zip(ts, ws, eager):
        assert nmse(y.cpu().numpy(), orc.mul_mat(t, w, x.cpu().numpy())) < MAX_NMSE
    with b200.Graph(s) as g:
        assert b200.lib.b200_mul_mat_group(arr, len(ts), s.handle) == 0
    for _ in range(2):
        for y in ys:
            y.zero_()
        torch.cuda.synchronize()
        g.launch()
        s.synchronize()
        for y, e in zip(ys, eager):
            assert nmse(y.cpu().numpy(), e.cpu().numpy()) < 1e-10
