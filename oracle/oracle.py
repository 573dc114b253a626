"""ctypes bindings for the CPU oracle -- TEST INFRASTRUCTURE ONLY.

Two libraries live behind this module:

* ``libggml_oracle.so``  -- our plain-C restatement (oracle/ggml_oracle.c), class :class:`Oracle`.
* ``_ref/<variant>/libref_harness.so`` -- the UNMODIFIED reference compiled by oracle/Makefile,
  class :class:`Ref`. It only exists where oracle/_ref was built (this container, and the GPU box
  because the built files travel with the snapshot).

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline / ``--impl reference`` leg may
import this module. The product (zyk-llama.cpp_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent

F32, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, Q8_K, IQ4_XS = 0, 2, 8, 9, 12, 13, 14, 15, 23
WEIGHT_TYPES = (Q4_0, Q8_0, Q4_K, Q5_K, Q6_K, IQ4_XS)
TYPE_NAMES = {F32: "f32", Q4_0: "q4_0", Q8_0: "q8_0", Q8_1: "q8_1", Q4_K: "q4_K", Q5_K: "q5_K", Q6_K: "q6_K",
              Q8_K: "q8_K", IQ4_XS: "iq4_xs"}
Q8_REF, Q8_X86 = 0, 1


def _ptr(a: np.ndarray):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def build_oracle() -> Path:
    so = HERE / "libggml_oracle.so"
    src = HERE / "ggml_oracle.c"
    if not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE), "oracle"], check=True, capture_output=True)
    return so


class Oracle:
    """Our C restatement (oracle/ggml_oracle.c)."""

    def __init__(self):
        self.lib = C.CDLL(str(build_oracle()))
        L = self.lib
        L.orc_fp16_to_fp32.restype = C.c_float
        L.orc_fp16_to_fp32.argtypes = [C.c_uint16]
        L.orc_fp32_to_fp16.restype = C.c_uint16
        L.orc_fp32_to_fp16.argtypes = [C.c_float]
        L.orc_block_elems.argtypes = [C.c_int]
        L.orc_block_bytes.restype = C.c_size_t
        L.orc_block_bytes.argtypes = [C.c_int]
        L.orc_row_size.restype = C.c_size_t
        L.orc_row_size.argtypes = [C.c_int, C.c_int64]
        L.orc_vec_dot_type.argtypes = [C.c_int]
        L.orc_dequantize_row.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_quantize_row.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        L.orc_quantize_row_q8_2_x4.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.orc_vec_dot.restype = C.c_float
        L.orc_vec_dot.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_mul_mat.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_mul_mat_f64.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_mul_mat_id.argtypes = [C.c_int] + [C.c_int64] * 6 + [C.c_void_p] * 4 + [C.c_int]
        L.orc_b200_repack_supported.argtypes = [C.c_int, C.c_int64, C.c_int64]
        L.orc_b200_repack.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_b200_unrepack.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_cpu_repack_q4_K_8x8.argtypes = [C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_cpu_repack_q4_0_zyk.argtypes = [C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]

    def row_size(self, t, k):
        return int(self.lib.orc_row_size(t, k))

    def block_elems(self, t):
        return int(self.lib.orc_block_elems(t))

    def block_bytes(self, t):
        return int(self.lib.orc_block_bytes(t))

    def vec_dot_type(self, t):
        return int(self.lib.orc_vec_dot_type(t))

    def dequantize(self, t, w: np.ndarray, k: int) -> np.ndarray:
        """w: uint8 [rows, row_size] -> f32 [rows, k]."""
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, self.row_size(t, k))
        out = np.empty((w.shape[0], k), np.float32)
        self.lib.orc_dequantize_row(t, _ptr(w), _ptr(out), w.shape[0] * k)
        return out

    def quantize(self, q8t, x: np.ndarray, flavour=Q8_X86) -> np.ndarray:
        """x: f32 [rows, k] -> uint8 [rows, row_size(q8t, k)] canonical ggml blocks."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.shape
        out = np.zeros((rows, self.row_size(q8t, k)), np.uint8)
        for r in range(rows):
            self.lib.orc_quantize_row(q8t, _ptr(x[r]), _ptr(out[r]), k, flavour)
        return out

    def quantize_q8_2_x4(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.shape
        out = np.zeros((rows, k // 32 * 36), np.uint8)
        for r in range(rows):
            self.lib.orc_quantize_row_q8_2_x4(_ptr(x[r]), _ptr(out[r]), k)
        return out

    def quantize_mat_q8_K_4x8(self, x: np.ndarray) -> np.ndarray:
        """x: f32 [4, k] -> block_q8_Kx4 bytes uint8 [k / 256 * 1168]"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        assert x.shape[0] == 4 and x.shape[1] % 256 == 0
        out = np.zeros(x.shape[1] // 256 * 1168, np.uint8)
        self.lib.orc_quantize_mat_q8_K_4x8(_ptr(x), _ptr(out), C.c_int64(x.shape[1]))
        return out

    def mul_mat_q4_0_zyk(self, w: np.ndarray, x: np.ndarray) -> np.ndarray:
        """the fork's ZYK Q4_0 path: w canonical q4_0 bytes [m, k / 32 * 18], x f32 [n, k] -> f32 [n, m]"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        n, k = x.shape
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, k // 32 * 18)
        y = np.empty((n, w.shape[0]), np.float32)
        self.lib.orc_mul_mat_q4_0_zyk(C.c_int64(w.shape[0]), C.c_int64(n), C.c_int64(k), _ptr(w), _ptr(x), _ptr(y))
        return y

    def vec_dot(self, t, k, w_row: np.ndarray, xq_row: np.ndarray) -> float:
        return float(self.lib.orc_vec_dot(t, k, _ptr(w_row), _ptr(xq_row)))

    def mul_mat(self, t, w: np.ndarray, x: np.ndarray, flavour=Q8_X86) -> np.ndarray:
        """w: uint8 [m, row_size], x: f32 [n, k] -> f32 [n, m]."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        n, k = x.shape
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, self.row_size(t, k))
        m = w.shape[0]
        y = np.empty((n, m), np.float32)
        self.lib.orc_mul_mat(t, m, n, k, _ptr(w), _ptr(x), _ptr(y), flavour)
        return y

    def mul_mat_f64(self, t, w, x) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        n, k = x.shape
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, self.row_size(t, k))
        m = w.shape[0]
        y = np.empty((n, m), np.float32)
        self.lib.orc_mul_mat_f64(t, m, n, k, _ptr(w), _ptr(x), _ptr(y))
        return y

    def mul_mat_id(self, t, as_: np.ndarray, b: np.ndarray, ids: np.ndarray, flavour=Q8_X86) -> np.ndarray:
        """as_: uint8 [n_as, m, row_size]; b: f32 [n_tokens, nb1, k]; ids: i32 [n_tokens, n_used]
        -> f32 [n_tokens, n_used, m]."""
        b = np.ascontiguousarray(b, dtype=np.float32)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n_tokens, nb1, k = b.shape
        n_used = ids.shape[1]
        as_ = np.ascontiguousarray(as_, dtype=np.uint8)
        n_as, m = as_.shape[0], as_.shape[1]
        y = np.empty((n_tokens, n_used, m), np.float32)
        self.lib.orc_mul_mat_id(t, m, k, n_as, n_used, nb1, n_tokens, _ptr(as_), _ptr(b), _ptr(ids), _ptr(y), flavour)
        return y

    def b200_repack_supported(self, t, m, k) -> bool:
        return bool(self.lib.orc_b200_repack_supported(t, m, k))

    def b200_repack(self, t, w: np.ndarray, m, k) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8)
        out = np.empty_like(w)
        self.lib.orc_b200_repack(t, m, k, _ptr(w), _ptr(out))
        return out

    def b200_unrepack(self, t, p: np.ndarray, m, k) -> np.ndarray:
        p = np.ascontiguousarray(p, dtype=np.uint8)
        out = np.empty_like(p)
        self.lib.orc_b200_unrepack(t, m, k, _ptr(p), _ptr(out))
        return out

    def cpu_repack_q4_K_8x8(self, w: np.ndarray, m, k) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8)
        out = np.empty_like(w)
        assert self.lib.orc_cpu_repack_q4_K_8x8(m, k, _ptr(w), _ptr(out)) == 0
        return out

    def cpu_repack_q4_0_zyk(self, w: np.ndarray, m, k) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8)
        out = np.empty_like(w)
        assert self.lib.orc_cpu_repack_q4_0_zyk(m, k, _ptr(w), _ptr(out)) == 0
        return out


# ------------------------------------------------------------------------------------------------
# the real reference (oracle/_ref), when it has been built
# ------------------------------------------------------------------------------------------------
def _cpu_flags() -> set:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def ref_variant(prefer_amx: bool = False) -> str | None:
    """Pick the oracle/_ref CPU variant this host can execute (oracle/Makefile header)."""
    flags = _cpu_flags()
    need512 = {"avx2", "fma", "f16c", "bmi2", "avx512f", "avx512cd", "avx512bw", "avx512dq", "avx512vl", "avx512_vnni"}
    if not need512 <= flags:
        return None
    if prefer_amx and {"amx_tile", "amx_int8"} <= flags and (HERE / "_ref" / "amx" / "libref_harness.so").exists():
        return "amx"
    if (HERE / "_ref" / "avx512vnni" / "libref_harness.so").exists():
        return "avx512vnni"
    return None


def ref_dir(variant: str | None = None) -> Path | None:
    variant = variant or ref_variant()
    return None if variant is None else HERE / "_ref" / variant


class Ref:
    """The unmodified reference through oracle/ref_harness.cpp."""

    def __init__(self, variant: str | None = None):
        d = ref_dir(variant)
        if d is None:
            raise RuntimeError("oracle/_ref is not built for a CPU variant this host can run")
        self.variant = d.name
        self.lib = C.CDLL(str(d / "libref_harness.so"))
        L = self.lib
        L.ref_row_size.restype = C.c_size_t
        L.ref_row_size.argtypes = [C.c_int, C.c_int64]
        L.ref_vec_dot_type.argtypes = [C.c_int]
        L.ref_quantize_weights.restype = C.c_size_t
        L.ref_quantize_weights.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.ref_dequantize.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_from_float_cpu.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_from_float_ref.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.ref_vec_dot.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_mul_mat.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.ref_mul_mat_id.argtypes = [C.c_int] + [C.c_int64] * 6 + [C.c_void_p] * 4 + [C.c_int]
        L.ref_repack_dump.restype = C.c_int64
        L.ref_repack_dump.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]

    def row_size(self, t, k):
        return int(self.lib.ref_row_size(t, k))

    def quantize_mat_q8_K_4x8(self, x: np.ndarray, generic: bool = False) -> np.ndarray:
        """the reference's own ggml_quantize_mat_q8_K_4x8 (SIMD, arch/x86/repack.cpp:774-1000) or _generic (repack.cpp:180-229),
        called directly in its libggml-cpu.so"""
        cpu = C.CDLL(str(ref_dir(self.variant) / "libggml-cpu.so"))
        fn = cpu.ggml_quantize_mat_q8_K_4x8_generic if generic else cpu.ggml_quantize_mat_q8_K_4x8
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.zeros(x.shape[1] // 256 * 1168, np.uint8)
        fn(_ptr(x), _ptr(out), x.shape[1])
        return out

    def quantize_weights(self, t, w: np.ndarray) -> np.ndarray:
        """f32 [m, k] -> canonical quantised bytes uint8 [m, row_size] (ggml_quantize_chunk)."""
        w = np.ascontiguousarray(w, dtype=np.float32)
        m, k = w.shape
        out = np.zeros((m, self.row_size(t, k)), np.uint8)
        n = self.lib.ref_quantize_weights(t, _ptr(w), _ptr(out), m, k)
        assert n == out.size
        return out

    def dequantize(self, t, w: np.ndarray, k) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, self.row_size(t, k))
        out = np.empty((w.shape[0], k), np.float32)
        self.lib.ref_dequantize(t, _ptr(w), _ptr(out), w.shape[0] * k)
        return out

    def quantize(self, q8t, x: np.ndarray, cpu_variant=True) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        rows, k = x.shape
        out = np.zeros((rows, self.row_size(q8t, k)), np.uint8)
        for r in range(rows):
            if cpu_variant:
                self.lib.ref_from_float_cpu(q8t, _ptr(x[r]), _ptr(out[r]), k)
            elif not self.lib.ref_from_float_ref(q8t, _ptr(x[r]), _ptr(out[r]), k):
                return None  # the base traits have no *_ref quantiser for this type (Q8_K)
        return out

    def vec_dot(self, t, k, w_row, xq_row) -> float:
        s = C.c_float(0)
        self.lib.ref_vec_dot(t, k, C.byref(s), _ptr(w_row), _ptr(xq_row))
        return float(s.value)

    def mul_mat(self, t, w, x, n_threads=1, repack=False, iters=0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        n, k = x.shape
        w = np.ascontiguousarray(w, dtype=np.uint8).reshape(-1, self.row_size(t, k))
        m = w.shape[0]
        y = np.empty((n, m), np.float32)
        secs = C.c_double(0)
        rc = self.lib.ref_mul_mat(t, m, n, k, _ptr(w), _ptr(x), _ptr(y), n_threads, int(repack), iters, C.byref(secs))
        if rc != 0:
            raise RuntimeError(f"ref_mul_mat rc={rc}")
        return (y, secs.value) if iters else y

    def mul_mat_id(self, t, as_, b, ids, n_threads=1):
        b = np.ascontiguousarray(b, dtype=np.float32)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n_tokens, nb1, k = b.shape
        n_used = ids.shape[1]
        as_ = np.ascontiguousarray(as_, dtype=np.uint8)
        n_as, m = as_.shape[0], as_.shape[1]
        y = np.empty((n_tokens, n_used, m), np.float32)
        rc = self.lib.ref_mul_mat_id(t, m, k, n_as, n_used, nb1, n_tokens, _ptr(as_), _ptr(b), _ptr(ids), _ptr(y), n_threads)
        if rc != 0:
            raise RuntimeError(f"ref_mul_mat_id rc={rc}")
        return y

    def repack_dump(self, t, w, m, k):
        w = np.ascontiguousarray(w, dtype=np.uint8)
        out = np.empty_like(w)
        rc = self.lib.ref_repack_dump(t, m, k, _ptr(w), _ptr(out))
        return None if rc < 0 else out


# ------------------------------------------------------------------------------------------------
# synthetic weights: random but VALID blocks (finite fp16 scales of sane magnitude, every bit
# pattern of the packed fields exercised). All byte patterns except NaN/Inf scales are legal input
# for MUL_MAT, so this covers more of the format than quantising gaussian weights does.
# ------------------------------------------------------------------------------------------------
def random_blocks(t: int, m: int, k: int, rng: np.random.Generator, scale: float = 0.02) -> np.ndarray:
    elems = {Q4_0: 32, Q8_0: 32, Q4_K: 256, Q5_K: 256, Q6_K: 256, IQ4_XS: 256}[t]
    bsz = {Q4_0: 18, Q8_0: 34, Q4_K: 144, Q5_K: 176, Q6_K: 210, IQ4_XS: 136}[t]
    nb = m * (k // elems)
    blk = rng.integers(0, 256, size=(nb, bsz), dtype=np.uint8)

    def put_half(col, vals):
        blk[:, col:col + 2] = np.asarray(vals, dtype=np.float16).view(np.uint8).reshape(nb, 2)

    mag = lambda div: (scale / div) * (0.25 + rng.random(nb)) * rng.choice([-1.0, 1.0], nb)
    if t == Q4_0:
        put_half(0, mag(8))
    elif t == Q8_0:
        put_half(0, mag(127))
    elif t == Q4_K:
        put_half(0, np.abs(mag(15 * 40)))
        put_half(2, np.abs(mag(40)))
    elif t == Q5_K:
        put_half(0, np.abs(mag(31 * 40)))
        put_half(2, np.abs(mag(40)))
    elif t == Q6_K:
        put_half(208, mag(32 * 64))
    elif t == IQ4_XS:
        put_half(0, mag(127 * 20))
    return blk.reshape(m, (k // elems) * bsz)
