"""MUL_MAT workloads of the BASELINE.json configs: which quantised matmuls one llama-bench token
(tg, N = 1) or one ubatch (pp, N = 512) issues, with the per-tensor types of the reference's own
llama-quantize recipe (/root/reference/src/llama-quant.cpp:185-187 use_more_bits, :207-226 output
tensor -> Q6_K, :302-303 attn_v, :358-364 ffn_down), and random-but-valid quantised weights for
them (there is no network for checkpoints: "random-init GGUF weights of the named shapes").

Pure host logic + torch for device memory; no arithmetic of the hot path lives here.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch

F32, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, Q8_K, IQ4_XS = 0, 2, 8, 9, 12, 13, 14, 15, 23
BLOCK = {Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210), IQ4_XS: (256, 136)}


def row_bytes(t: int, k: int) -> int:
    e, b = BLOCK[t]
    return k // e * b


@dataclass(frozen=True)
class MatMul:
    """One GGML_OP_MUL_MAT node: W[m, k] (type) applied to activation buffer `src` -> buffer `dst`."""
    name: str
    type: int
    m: int
    k: int
    src: str          # activation buffer read (matmuls sharing `src` share one quantisation of X)
    dst: str
    n_expert: int = 0  # > 0: MUL_MAT_ID over this many expert matrices
    n_used: int = 0

    @property
    def weight_bytes(self) -> int:
        return self.m * row_bytes(self.type, self.k) * max(1, self.n_expert)

    @property
    def streamed_bytes(self) -> int:
        """algorithmic bytes one decode token streams from HBM for this node"""
        return self.m * row_bytes(self.type, self.k) * (self.n_used if self.n_expert else 1)


def use_more_bits(i: int, n: int) -> bool:
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def llama_dense(n_layer: int, n_embd: int, n_ff: int, n_head: int, n_kv: int, n_vocab: int, ftype: str,
                n_expert: int = 0, n_used: int = 0, recipe_layers: int = 0, recipe_offset: int = 0, exact: bool = False) -> list[MatMul]:
    """The quantised MUL_MAT / MUL_MAT_ID nodes of one llama-arch forward pass, in graph order
    (src/models/llama.cpp: wq, wk, wv, wo, ffn_gate, ffn_up, ffn_down; output once).

    exact=True gives every tensor the type llama-quantize (no imatrix) really writes, pinned against the reference's quantiser in
    tests/test_quant_recipe.py. The default differs from it in three places, all Q5_K tensors the round-1 scope (five formats) left out
    and the measurements of round 2 were taken without: IQ4_XS files keep attn_v (n_gqa >= 4, llama-quant.cpp:299-301) and the ffn_down of
    the first n_layer / 8 layers (:366-368) in IQ4_XS instead of Q5_K, and 8-expert files keep attn_output in Q4_K instead of Q5_K (:383-390)."""
    exact = exact or os.environ.get("B200_EXACT_RECIPE", "0") == "1"      # flips every caller (bench.py configs, make_gguf.py) at once for an A/B run
    hd = n_embd // n_head
    kvd = hd * n_kv
    rl = recipe_layers or n_layer          # a file cut to a few layers (tests) keeps the recipe of the full model it was cut from
    base = {"q4_k_m": Q4_K, "q8_0": Q8_0, "iq4_xs": IQ4_XS, "q4_0": Q4_0}[ftype]
    out_t = Q8_0 if ftype == "q8_0" else Q6_K
    ops: list[MatMul] = []
    for i in range(n_layer):
        more = ftype == "q4_k_m" and use_more_bits(i + recipe_offset, rl)      # (recipe_offset: layer i of a cut file is layer i + offset of the full model)
        tv = Q6_K if more else base
        td = Q6_K if more else base
        tk = base
        if rl == 80 and tv == Q4_K:            # llama-quant.cpp:305-310: the 70B model (8 heads share attn_v) takes Q5_K where Q4_K was due
            tv = Q5_K
        to = base
        if exact and ftype == "iq4_xs":
            if n_head // n_kv >= 4:
                tv = Q5_K
            if i + recipe_offset < rl // 8:
                td = Q5_K
        if n_expert == 8 and ftype == "q4_k_m":     # llama-quant.cpp:311-322: 8-expert models keep attn_k/attn_v in Q8_0
            tv = tk = Q8_0
            if exact:
                to = Q5_K
        p = f"blk.{i}."
        ops += [MatMul(p + "attn_q", base, n_embd, n_embd, "x", "q"),
                MatMul(p + "attn_k", tk, kvd, n_embd, "x", "k"),
                MatMul(p + "attn_v", tv, kvd, n_embd, "x", "v"),
                MatMul(p + "attn_output", to, n_embd, n_embd, "attn", "x2")]
        if n_expert:
            ops += [MatMul(p + "ffn_gate_exps", base, n_ff, n_embd, "x2", "gate", n_expert, n_used),
                    MatMul(p + "ffn_up_exps", base, n_ff, n_embd, "x2", "up", n_expert, n_used),
                    MatMul(p + "ffn_down_exps", td, n_embd, n_ff, "act", "x", n_expert, n_used)]
        else:
            ops += [MatMul(p + "ffn_gate", base, n_ff, n_embd, "x2", "gate"),
                    MatMul(p + "ffn_up", base, n_ff, n_embd, "x2", "up"),
                    MatMul(p + "ffn_down", td, n_embd, n_ff, "act", "x")]
    ops.append(MatMul("output", out_t, n_vocab, n_embd, "x", "logits"))
    return ops


MODELS = {
    # BASELINE.json configs[0..4]
    "llama-3.2-1b-q4_0": dict(n_layer=16, n_embd=2048, n_ff=8192, n_head=32, n_kv=8, n_vocab=128256, ftype="q4_0"),
    "llama-3-8b-q4_k_m": dict(n_layer=32, n_embd=4096, n_ff=14336, n_head=32, n_kv=8, n_vocab=128256, ftype="q4_k_m"),
    "llama-3-8b-q8_0": dict(n_layer=32, n_embd=4096, n_ff=14336, n_head=32, n_kv=8, n_vocab=128256, ftype="q8_0"),
    "llama-3-8b-iq4_xs": dict(n_layer=32, n_embd=4096, n_ff=14336, n_head=32, n_kv=8, n_vocab=128256, ftype="iq4_xs"),
    "mixtral-8x7b-q4_k_m": dict(n_layer=32, n_embd=4096, n_ff=14336, n_head=32, n_kv=8, n_vocab=32000, ftype="q4_k_m",
                                n_expert=8, n_used=2),
    # 70B: attn_v would be Q5_K in llama-quantize's recipe (llama-quant.cpp:305-310); the synthetic file
    # pins it to Q6_K/Q4_K via --tensor-type so that only the five named formats occur
    # small shapes for the multi-GPU parity check (tools/rowsplit_check.py): uneven vocab split, Q6_K rows, k = 3584
    "tiny-q4_k_m": dict(n_layer=4, n_embd=1024, n_ff=3584, n_head=8, n_kv=2, n_vocab=5000, ftype="q4_k_m"),
    "tiny-moe-q4_k_m": dict(n_layer=2, n_embd=1024, n_ff=1792, n_head=8, n_kv=2, n_vocab=3000, ftype="q4_k_m", n_expert=8, n_used=2),
    "llama-3-70b-q4_k_m": dict(n_layer=80, n_embd=8192, n_ff=28672, n_head=64, n_kv=8, n_vocab=128256, ftype="q4_k_m"),
}


def model_ops(name: str) -> list[MatMul]:
    return llama_dense(**MODELS[name])


def buffer_widths(ops: list[MatMul]) -> dict[str, int]:
    """element count of every activation buffer the op list names"""
    w: dict[str, int] = {}
    for o in ops:
        w[o.src] = max(w.get(o.src, 0), o.k)
        w[o.dst] = max(w.get(o.dst, 0), o.m)
    return w


def tg_bytes(ops: list[MatMul]) -> int:
    return sum(o.streamed_bytes for o in ops)


def pp_ops(ops: list[MatMul]) -> int:
    """integer multiply-accumulate ops (x2) per prompt token"""
    return sum(2 * o.m * o.k * (o.n_used if o.n_expert else 1) for o in ops)


# ------------------------------------------------------------------------------------------------
# random but VALID blocks: every packed bit pattern is legal MUL_MAT input except NaN/Inf in the
# fp16 scale fields, so the payload is uniform random bytes and the scale fields are patched with
# finite values of the magnitude a N(0, 0.02^2) weight matrix would produce.
# ------------------------------------------------------------------------------------------------
def random_weight(t: int, m: int, k: int, device, gen: torch.Generator, scale: float = 0.02, chunk_rows: int = 8192) -> torch.Tensor:
    """uint8 [m * row_bytes] canonical ggml blocks on `device`."""
    e, bsz = BLOCK[t]
    nbr = k // e
    out = torch.empty((m, nbr, bsz), dtype=torch.uint8, device=device)

    def half_field(nb, div, positive=False):
        mag = (0.25 + torch.rand((nb,), device=device, generator=gen)) * (scale / div)
        if not positive:
            mag = mag * (torch.randint(0, 2, (nb,), device=device, generator=gen).float() * 2 - 1)
        return mag.to(torch.float16).view(torch.uint8).reshape(nb, 2)

    for r0 in range(0, m, chunk_rows):
        r1 = min(m, r0 + chunk_rows)
        nb = (r1 - r0) * nbr
        blk = torch.randint(0, 256, (nb, bsz), dtype=torch.uint8, device=device, generator=gen)
        if t == Q4_0:
            blk[:, 0:2] = half_field(nb, 8)
        elif t == Q8_0:
            blk[:, 0:2] = half_field(nb, 127)
        elif t == Q4_K:
            blk[:, 0:2] = half_field(nb, 15 * 40, True)
            blk[:, 2:4] = half_field(nb, 40, True)
        elif t == Q5_K:
            blk[:, 0:2] = half_field(nb, 31 * 40, True)
            blk[:, 2:4] = half_field(nb, 40, True)
        elif t == Q6_K:
            blk[:, 208:210] = half_field(nb, 32 * 64)
        elif t == IQ4_XS:
            blk[:, 0:2] = half_field(nb, 127 * 20)
        out[r0:r1] = blk.reshape(r1 - r0, nbr, bsz)
    return out.reshape(-1)
