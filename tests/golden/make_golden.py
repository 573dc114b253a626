"""Generate tests/golden/ref_vectors.npz from the UNMODIFIED reference (oracle/_ref).

Run in the build container (needs oracle/_ref, i.e. /root/reference):  python tests/golden/make_golden.py
Every array below was produced by reference code through oracle/ref_harness.cpp; the inputs are
seeded so the file is reproducible. The committed .npz travels to the GPU box, the reference does not.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle.oracle import Ref, WEIGHT_TYPES, TYPE_NAMES, Q8_0, Q8_1, Q8_K, Q4_K, Q5_K, Q4_0, random_blocks  # noqa: E402


def main():
    ref = Ref("avx512vnni")
    rng = np.random.default_rng(20261104)
    out = {}
    k, m, n = 512, 16, 4
    x = (rng.standard_normal((n, k)) * rng.uniform(0.05, 8.0, (n, 1))).astype(np.float32)
    x[1, :256] = 0.0                      # an all-zero Q8_K block
    x[2, 5] = 37.5; x[2, 9] = -37.5       # a sign tie at the block maximum (Q8_K "first max" rule)
    x[3] = np.round(x[3] * 2) / 2         # exact .5 products exercise round-half rules
    out["x"] = x
    for q8 in (Q8_0, Q8_1, Q8_K):
        out[f"xq_cpu_{TYPE_NAMES[q8]}"] = ref.quantize(q8, x, cpu_variant=True)
        r = ref.quantize(q8, x, cpu_variant=False)
        if r is not None:
            out[f"xq_ref_{TYPE_NAMES[q8]}"] = r
    def per_type(t):
        name = TYPE_NAMES[t]
        wf = (rng.standard_normal((m, k)) * 0.02).astype(np.float32)
        wq = ref.quantize_weights(t, wf)
        wr = random_blocks(t, m, k, rng)
        for tag, w in (("quant", wq), ("rand", wr)):
            out[f"w_{name}_{tag}"] = w
            out[f"deq_{name}_{tag}"] = ref.dequantize(t, w, k)
            out[f"y_{name}_{tag}"] = ref.mul_mat(t, w, x, n_threads=1)
        # MUL_MAT_ID: 4 experts, 2 used, 3 tokens, both b shapes (nb1 = n_used and nb1 = 1)
        n_as, n_used, n_tok = 4, 2, 3
        as_ = np.stack([random_blocks(t, m, k, rng) for _ in range(n_as)])
        ids = np.array([[0, 3], [2, 2], [1, 0]], dtype=np.int32)
        b2 = rng.standard_normal((n_tok, n_used, k)).astype(np.float32)
        b1 = rng.standard_normal((n_tok, 1, k)).astype(np.float32)
        out[f"id_as_{name}"] = as_
        out[f"id_ids_{name}"] = ids
        out[f"id_b2_{name}"] = b2
        out[f"id_b1_{name}"] = b1
        out[f"id_y2_{name}"] = ref.mul_mat_id(t, as_, b2, ids)
        out[f"id_y1_{name}"] = ref.mul_mat_id(t, as_, b1, ids)

    first = [t for t in WEIGHT_TYPES if t != Q5_K]      # types added later draw from the generator last: earlier vectors keep their bytes
    for t in first:
        per_type(t)
    # the reference CPU's own load-time repacks (CPU_REPACK buffer bytes)
    w4k = random_blocks(Q4_K, 16, 512, rng)
    d = ref.repack_dump(Q4_K, w4k, 16, 512)
    if d is not None:
        out["cpu_repack_q4_K_in"] = w4k
        out["cpu_repack_q4_K_out"] = d
    w40 = random_blocks(Q4_0, 64, 256, rng)
    d = ref.repack_dump(Q4_0, w40, 64, 256)
    if d is not None:
        out["cpu_repack_q4_0_in"] = w40
        out["cpu_repack_q4_0_out"] = d
    per_type(Q5_K)
    path = Path(__file__).with_name("ref_vectors.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path} ({path.stat().st_size} bytes, {len(out)} arrays; reference variant {ref.variant})")


if __name__ == "__main__":
    main()
