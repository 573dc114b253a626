"""CPU: pin oracle/ggml_oracle.c against the UNMODIFIED reference compiled into oracle/_ref
(live, on seeded inputs larger than the golden fixtures). Skipped where oracle/_ref is absent."""
import numpy as np
import pytest

from conftest import nmse
from oracle.oracle import WEIGHT_TYPES, TYPE_NAMES, Q8_0, Q8_1, Q8_K, Q4_0, Q4_K, Q8_REF, Q8_X86, random_blocks


def _acts(rng, n, k):
    x = (rng.standard_normal((n, k)) * rng.uniform(0.01, 30, (n, 1))).astype(np.float32)
    x[0, :256] = 0
    x[1] = np.round(x[1] * 4) / 4
    x[2, :32] = np.arange(32, dtype=np.float32) + 0.5   # exact .5 ties at id == 1
    x[2, 0] = 127.0
    return x


def test_fp16_conversions_exhaustive(orc):
    h = np.arange(65536, dtype=np.uint16)
    f = h.view(np.float16).astype(np.float32)
    mine = np.array([orc.lib.orc_fp16_to_fp32(int(v)) for v in h], dtype=np.float32)
    ok = ~np.isnan(f)
    assert np.array_equal(mine.view(np.uint32)[ok], f.view(np.uint32)[ok])
    back = np.array([orc.lib.orc_fp32_to_fp16(float(v)) for v in f[ok]], dtype=np.uint16)
    assert np.array_equal(back, h[ok])
    ties = np.array([65519.99, 65520.0, 2.9802322e-8, 2.98e-8, 5.96e-8, 6.1e-5, 1.0009766, 1.00048828125], np.float32)
    with np.errstate(over="ignore"):
        want = ties.astype(np.float16).view(np.uint16)
    got = np.array([orc.lib.orc_fp32_to_fp16(float(v)) for v in ties], dtype=np.uint16)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("q8", [Q8_0, Q8_1, Q8_K])
def test_activation_quantisers_bit_exact(orc, ref, q8):
    x = _acts(np.random.default_rng(1), 96, 2048)
    assert np.array_equal(orc.quantize(q8, x, Q8_X86), ref.quantize(q8, x, cpu_variant=True))
    r = ref.quantize(q8, x, cpu_variant=False)
    if r is not None:
        assert np.array_equal(orc.quantize(q8, x, Q8_REF), r)


def test_q8_0_cpu_and_ref_definitions_differ(ref):
    # why enum b200_q8_flavour exists: the reference holds two non-identical Q8_0 quantisers
    x = _acts(np.random.default_rng(2), 4, 256)
    x[2, :32] = np.arange(32, dtype=np.float32) + 0.5   # amax 127 => id == 1: exact .5 ties, RNE vs roundf
    x[2, 0] = 127.0
    assert not np.array_equal(ref.quantize(Q8_0, x, True), ref.quantize(Q8_0, x, False))


@pytest.mark.parametrize("t", WEIGHT_TYPES)
def test_dequant_vecdot_mulmat(orc, ref, t):
    rng = np.random.default_rng(100 + t)
    k, m, n = 1024, 40, 5
    x = _acts(rng, n, k)
    wq = ref.quantize_weights(t, (rng.standard_normal((m, k)) * 0.02).astype(np.float32))
    for w in (wq, random_blocks(t, m, k, rng)):
        a, b = orc.dequantize(t, w, k), ref.dequantize(t, w, k)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        xq = ref.quantize(orc.vec_dot_type(t), x, True)
        for i in range(4):
            got, want = orc.vec_dot(t, k, w[i], xq[i]), ref.vec_dot(t, k, w[i], xq[i])
            assert abs(got - want) <= 2e-5 * max(1e-3, abs(want))
        assert nmse(orc.mul_mat(t, w, x), ref.mul_mat(t, w, x, n_threads=4)) < 1e-12


@pytest.mark.parametrize("t", WEIGHT_TYPES)
def test_mul_mat_id(orc, ref, t):
    rng = np.random.default_rng(200 + t)
    k, m, n_as, n_used, n_tok = 512, 24, 8, 2, 7
    as_ = np.stack([random_blocks(t, m, k, rng) for _ in range(n_as)])
    ids = rng.integers(0, n_as, (n_tok, n_used)).astype(np.int32)
    for nb1 in (1, n_used):
        b = rng.standard_normal((n_tok, nb1, k)).astype(np.float32)
        assert nmse(orc.mul_mat_id(t, as_, b, ids), ref.mul_mat_id(t, as_, b, ids, n_threads=2)) < 1e-12


def test_cpu_repack_layouts_bit_exact(orc, ref):
    rng = np.random.default_rng(3)
    w = random_blocks(Q4_K, 64, 1024, rng)
    d = ref.repack_dump(Q4_K, w, 64, 1024)
    assert d is not None and np.array_equal(orc.cpu_repack_q4_K_8x8(w, 64, 1024), d)
    w = random_blocks(Q4_0, 128, 512, rng)
    d = ref.repack_dump(Q4_0, w, 128, 512)
    assert d is not None and np.array_equal(orc.cpu_repack_q4_0_zyk(w, 128, 512), d)


@pytest.mark.parametrize("t", [Q4_0, Q4_K])
def test_fork_repack_kernels_agree_with_generic_path(ref, t):
    """The fork's CPU_REPACK kernels (ZYK Q4_0 with bf16-scaled q8_2 activations, 8x8 Q4_K) are not
    covered by any reference test; cross-check them against the plain CPU path here."""
    rng = np.random.default_rng(4)
    k, m, n = 1024, 256, 6
    w = ref.quantize_weights(t, (rng.standard_normal((m, k)) * 0.02).astype(np.float32))
    x = rng.standard_normal((n, k)).astype(np.float32)
    plain = ref.mul_mat(t, w, x, n_threads=4)
    repacked = ref.mul_mat(t, w, x, n_threads=4, repack=True)
    assert nmse(repacked, plain) < 5e-4   # test-backend-ops' MUL_MAT tolerance (tests/test-backend-ops.cpp:3716-3718)


# ---- the fork's activation formats of the repack path -----------------------------------------------
def test_q8_Kx4_interleave_bit_exact(orc, ref):
    """block_q8_Kx4 (four rows quantised at once, 8-byte runs interleaved; repack.h:55-61): the restatement is bit for bit what the CPU
    backend runs on x86, ggml_quantize_mat_q8_K_4x8 (arch/x86/repack.cpp:774-1000, exported by the reference's libggml-cpu). The scalar
    ggml_quantize_mat_q8_K_4x8_generic (repack.cpp:180-229) computes nearest_int(x * iscale) in one expression, which this build of the
    reference contracts into an fma: it differs from the SIMD variant (and from us) by one unit on a handful of exact ties."""
    for seed, k in ((12, 256), (13, 2048), (14, 4096)):
        x = _acts(np.random.default_rng(seed), 4, k)
        mine = orc.quantize_mat_q8_K_4x8(x)
        assert np.array_equal(mine, ref.quantize_mat_q8_K_4x8(x, generic=False))
        gen = ref.quantize_mat_q8_K_4x8(x, generic=True).reshape(-1, 1168)
        mine = mine.reshape(-1, 1168)
        assert np.array_equal(gen[:, :16], mine[:, :16])                       # d
        dq = gen[:, 16:16 + 1024].view(np.int8).astype(np.int32) - mine[:, 16:16 + 1024].view(np.int8).astype(np.int32)
        assert np.abs(dq).max() <= 1 and np.count_nonzero(dq) <= 1e-3 * dq.size
        # layout: row r, element 8 g + e of a block sits at qs[32 g + 8 r + e]
        q = mine[0, 16:16 + 1024].view(np.int8).reshape(32, 4, 8)
        rows = orc.quantize(Q8_K, x, Q8_X86).reshape(4, k // 256, 292)
        for r in range(4):
            qs_row = rows[r, 0, 4:260].view(np.int8)
            if np.abs(qs_row.astype(np.int32)).max() < 127:                    # (the row quantiser clamps to 127, the 4x8 one does not)
                assert np.array_equal(q[:, r, :].reshape(-1), qs_row)


def test_zyk_q4_0_pipeline_bit_exact_and_q8_2_pinned(orc, ref):
    """The fork's ZYK Q4_0 CPU path (q8_2_x4 activations with a bf16 scale, unsigned nibbles with the -8 * sum bias, one fma per block):
    quantize_row_q8_2_x4 is a static function of the reference, so its restatement is pinned through the matmul that consumes it --
    bit for bit against the reference's CPU_REPACK buffer path, which differs measurably from the generic Q8_0 path."""
    rng = np.random.default_rng(77)
    m, k = 256, 1024
    w = random_blocks(Q4_0, m, k, rng)
    for n in (1, 3):
        x = _acts(rng, 4, k)[:n] if n <= 4 else None
        got_ref = ref.mul_mat(Q4_0, w, x, n_threads=1, repack=True)
        mine = orc.mul_mat_q4_0_zyk(w, x)
        assert np.array_equal(mine.view(np.uint32), got_ref.view(np.uint32)), float(np.abs(mine - got_ref).max())
        generic = ref.mul_mat(Q4_0, w, x, n_threads=1, repack=False)
        assert not np.array_equal(generic, got_ref)          # bf16 activation scales: a different function from the Q8_0 path
        assert nmse(generic, got_ref) < 1e-3
