#!/usr/bin/env python
"""Random-init GGUF writer for the BASELINE.json configs (there is no network for checkpoints): a llama-arch model file with the real
hyper-parameters of the named model, per-tensor quantisation types of the reference's own llama-quantize recipe (workloads.py) and
random-but-valid quantised blocks, which the reference's loader (src/llama-model-loader.cpp, gguf.cpp) reads like any other model.
Self-contained GGUF v3 writer (format: ggml/src/gguf.cpp / gguf-py): no dependency on the reference tree at run time, so it also runs
on the GPU box. Vocabulary: tokenizer.ggml.model = "no_vocab" with llama.vocab_size set (src/llama-vocab.cpp:1706-1720) -- llama-bench
only needs the vocabulary size.

    python tools/make_gguf.py --model llama-3-8b-q4_k_m --out /tmp/llama-3-8b-q4_k_m.gguf [--layers N]
Prints the per-tensor type histogram and the byte model of the file ("computed from the actual file")."""
from __future__ import annotations

import argparse
import importlib.util
import struct
import sys
from collections import Counter
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
GGML_TYPE_NAMES = {0: "f32", 2: "q4_0", 8: "q8_0", 12: "q4_K", 13: "q5_K", 14: "q6_K", 23: "iq4_xs"}
F32, Q4_0, Q8_0, Q4_K, Q5_K, Q6_K, IQ4_XS = 0, 2, 8, 12, 13, 14, 23
BLOCK = {Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210), IQ4_XS: (256, 136)}
ALIGN = 32
# gguf value types
T_U32, T_F32, T_STR = 4, 6, 8


def _workloads():
    spec = importlib.util.spec_from_file_location("_b200_workloads_gguf", ROOT / "zyk-llama.cpp_b200" / "workloads.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod


def gstr(s: str) -> bytes:
    b = s.encode()
    return struct.pack("<Q", len(b)) + b


def kv(key: str, vtype: int, value) -> bytes:
    out = gstr(key) + struct.pack("<I", vtype)
    if vtype == T_U32:
        return out + struct.pack("<I", value)
    if vtype == T_F32:
        return out + struct.pack("<f", value)
    if vtype == T_STR:
        return out + gstr(value)
    raise ValueError(vtype)


def random_blocks(t: int, rows: int, k: int, rng: np.random.Generator, scale: float = 0.02) -> np.ndarray:
    """uint8 [rows * row_bytes]: uniform random payload, finite fp16 scale fields of the magnitude of a N(0, scale^2) matrix"""
    e, bsz = BLOCK[t]
    nb = rows * (k // e)
    blk = rng.integers(0, 256, (nb, bsz), dtype=np.uint8)

    def half(div, positive=False):
        mag = (0.25 + rng.random(nb, dtype=np.float32)) * np.float32(scale / div)
        if not positive:
            mag *= rng.integers(0, 2, nb).astype(np.float32) * 2 - 1
        return mag.astype(np.float16).view(np.uint8).reshape(nb, 2)

    if t == Q4_0:
        blk[:, 0:2] = half(8)
    elif t == Q8_0:
        blk[:, 0:2] = half(127)
    elif t == Q4_K:
        blk[:, 0:2] = half(15 * 40, True); blk[:, 2:4] = half(40, True)
    elif t == Q5_K:
        blk[:, 0:2] = half(31 * 40, True); blk[:, 2:4] = half(40, True)
    elif t == Q6_K:
        blk[:, 208:210] = half(32 * 64)
    elif t == IQ4_XS:
        blk[:, 0:2] = half(127 * 20)
    return blk.reshape(-1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3-8b-q4_k_m")
    ap.add_argument("--out", required=True)
    ap.add_argument("--layers", type=int, default=0, help="truncate to this many layers (0 = the model's own count)")
    ap.add_argument("--layer-offset", type=int, default=0, help="with --layers: the file's layer i takes the per-tensor types of layer i + offset of the full model")
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--exact-recipe", action="store_true", help="per-tensor types exactly as llama-quantize writes them (workloads.llama_dense(exact=True): Q5_K in the IQ4_XS and Mixtral files)")
    ap.add_argument("--tiny", action="store_true", help="same layer count, head ratio and expert count, but n_embd = n_ff = n_vocab = 256 (input of tests/test_quant_recipe.py)")
    ap.add_argument("--f32", action="store_true", help="every tensor in f32, general.file_type 0: an unquantised source for the reference's llama-quantize")
    a = ap.parse_args()
    W = _workloads()
    cfg = dict(W.MODELS[a.model])
    if a.tiny:
        cfg.update(n_embd=256, n_ff=256, n_vocab=256, n_head=cfg["n_head"] // cfg["n_kv"], n_kv=1)
    if a.exact_recipe:
        cfg["exact"] = True
    if a.layers:
        cfg["recipe_layers"] = cfg["n_layer"]          # the per-tensor types of the full model (70B: Q5_K attn_v), on fewer layers
        cfg["n_layer"] = a.layers
        cfg["recipe_offset"] = a.layer_offset
    ops = W.llama_dense(**cfg)
    n_embd, n_ff, n_head, n_kv, n_vocab, n_layer = cfg["n_embd"], cfg["n_ff"], cfg["n_head"], cfg["n_kv"], cfg["n_vocab"], cfg["n_layer"]
    n_expert, n_used = cfg.get("n_expert", 0), cfg.get("n_used", 0)
    base = {"q4_k_m": Q4_K, "q8_0": Q8_0, "iq4_xs": IQ4_XS, "q4_0": Q4_0}[cfg["ftype"]]
    ftype_id = {"q4_k_m": 15, "q8_0": 7, "iq4_xs": 30, "q4_0": 2}[cfg["ftype"]]

    # tensor list: (name, ggml dims [ne0, ne1, ...], type)
    tensors = [("token_embd.weight", [n_embd, n_vocab], base)]
    for o in ops:
        if o.name == "output":
            continue
        name = o.name + ".weight"
        tensors.append((name, [o.k, o.m] + ([o.n_expert] if o.n_expert else []), o.type))
    for i in range(n_layer):
        tensors.append((f"blk.{i}.attn_norm.weight", [n_embd], F32))
        tensors.append((f"blk.{i}.ffn_norm.weight", [n_embd], F32))
        if n_expert:
            tensors.append((f"blk.{i}.ffn_gate_inp.weight", [n_embd, n_expert], F32))
    tensors.append(("output_norm.weight", [n_embd], F32))
    out_op = [o for o in ops if o.name == "output"][0]
    tensors.append(("output.weight", [out_op.k, out_op.m], out_op.type))
    if a.f32:
        tensors, ftype_id = [(name, dims, F32) for name, dims, _ in tensors], 0

    meta = [kv("general.architecture", T_STR, "llama"), kv("general.name", T_STR, f"random-init {a.model}"),
            kv("general.file_type", T_U32, ftype_id), kv("general.quantization_version", T_U32, 2), kv("general.alignment", T_U32, ALIGN),
            kv("llama.block_count", T_U32, n_layer), kv("llama.context_length", T_U32, 8192), kv("llama.embedding_length", T_U32, n_embd),
            kv("llama.feed_forward_length", T_U32, n_ff), kv("llama.attention.head_count", T_U32, n_head),
            kv("llama.attention.head_count_kv", T_U32, n_kv), kv("llama.rope.dimension_count", T_U32, n_embd // n_head),
            kv("llama.rope.freq_base", T_F32, 500000.0), kv("llama.attention.layer_norm_rms_epsilon", T_F32, 1e-5),
            kv("llama.vocab_size", T_U32, n_vocab), kv("tokenizer.ggml.model", T_STR, "no_vocab")]
    if n_expert:
        meta += [kv("llama.expert_count", T_U32, n_expert), kv("llama.expert_used_count", T_U32, n_used)]

    def nbytes(dims, t):
        n = int(np.prod(dims))
        if t == F32:
            return 4 * n
        e, b = BLOCK[t]
        return n // e * b

    infos, off = b"", 0
    offsets = []
    for name, dims, t in tensors:
        infos += gstr(name) + struct.pack("<I", len(dims)) + b"".join(struct.pack("<Q", d) for d in dims) + struct.pack("<IQ", t, off)
        offsets.append(off)
        off += (nbytes(dims, t) + ALIGN - 1) // ALIGN * ALIGN
    header = struct.pack("<IIQQ", 0x46554747, 3, len(tensors), len(meta)) + b"".join(meta) + infos
    pad = (-len(header)) % ALIGN
    rng = np.random.default_rng(a.seed)
    hist, by_type = Counter(), Counter()
    with open(a.out, "wb") as f:
        f.write(header + b"\0" * pad)
        for (name, dims, t), o in zip(tensors, offsets):
            nb = nbytes(dims, t)
            if t == F32:
                data = (np.ones(int(np.prod(dims)), np.float32) if "norm" in name else (rng.standard_normal(int(np.prod(dims))) * 0.02).astype(np.float32)).view(np.uint8)
                f.write(data.tobytes())
            else:
                rows = int(np.prod(dims[1:]))
                for r0 in range(0, rows, 16384):                     # bounded host memory
                    f.write(random_blocks(t, min(16384, rows - r0), dims[0], rng).tobytes())
            f.write(b"\0" * ((-nb) % ALIGN))
            hist[GGML_TYPE_NAMES[t]] += 1
            by_type[GGML_TYPE_NAMES[t]] += nb
    streamed = sum(o.streamed_bytes for o in ops)
    print(f"{a.out}: {len(tensors)} tensors, {off / 1e9:.3f} GB of tensor data; types {dict(hist)}; bytes by type {{{', '.join(f'{k}: {v / 1e9:.3f} GB' for k, v in by_type.items())}}}")
    print(f"byte model: quantised MUL_MAT weights streamed per decode token = {streamed / 1e9:.3f} GB ({len(ops)} MUL_MAT nodes)")


if __name__ == "__main__":
    main()
