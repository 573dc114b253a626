"""The per-tensor quantisation recipes of zyk-llama.cpp_b200/workloads.py (what bench.py streams and what tools/make_gguf.py writes) against the
rules of the reference's llama-quantize (src/llama-quant.cpp: use_more_bits :185-187, attn_v :299-322, ffn_down :358-364, output :207-...) and
the byte model of SURVEY.md 8d. CPU only."""
import importlib.util
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def W():
    spec = importlib.util.spec_from_file_location("b200_workloads", ROOT / "zyk-llama.cpp_b200" / "workloads.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["b200_workloads"] = mod
    spec.loader.exec_module(mod)
    return mod


def _by_name(ops):
    return {o.name: o for o in ops}


def test_use_more_bits_is_the_reference_rule(W):
    # llama-quant.cpp:185-187: i < n/8 || i >= 7n/8 || (i - n/8) % 3 == 2
    n = 32
    want = [i for i in range(n) if i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2]
    assert [i for i in range(n) if W.use_more_bits(i, n)] == want
    assert len(want) == 16            # "16 of 32" layers of the 8B Q4_K_M file carry Q6_K attn_v / ffn_down


def test_8b_q4_k_m_recipe_and_bytes(W):
    ops = W.llama_dense(**W.MODELS["llama-3-8b-q4_k_m"])
    assert len(ops) == 32 * 7 + 1
    d = _by_name(ops)
    for i in range(32):
        more = W.use_more_bits(i, 32)
        assert d[f"blk.{i}.attn_v"].type == (W.Q6_K if more else W.Q4_K)
        assert d[f"blk.{i}.ffn_down"].type == (W.Q6_K if more else W.Q4_K)
        for t in ("attn_q", "attn_k", "attn_output", "ffn_gate", "ffn_up"):
            assert d[f"blk.{i}.{t}"].type == W.Q4_K
    assert d["output"].type == W.Q6_K and d["output"].m == 128256
    assert abs(W.tg_bytes(ops) / 1e9 - 4.616) < 0.001          # 4.616 GB per token


def test_70b_keeps_attn_v_in_q5_k(W):
    # llama-quant.cpp:305-310: in the 70B model attn_v that would be Q4_K becomes Q5_K; the use_more_bits layers stay Q6_K
    ops = W.llama_dense(**W.MODELS["llama-3-70b-q4_k_m"])
    d = _by_name(ops)
    n5 = n6 = 0
    for i in range(80):
        t = d[f"blk.{i}.attn_v"].type
        assert t == (W.Q6_K if W.use_more_bits(i, 80) else W.Q5_K)
        n5 += t == W.Q5_K
        n6 += t == W.Q6_K
        assert d[f"blk.{i}.attn_k"].type == W.Q4_K
    assert n5 + n6 == 80 and n5 > 0 and n6 > 0
    assert W.BLOCK[W.Q5_K] == (256, 176)
    assert abs(W.tg_bytes(ops) / 1e9 - 41.9) < 0.1


def test_mixtral_recipe(W):
    # llama-quant.cpp:311-322: 8-expert models keep attn_k / attn_v in Q8_0; experts: 2 of 8 streamed per token
    cfg = W.MODELS["mixtral-8x7b-q4_k_m"]
    ops = W.llama_dense(**cfg)
    d = _by_name(ops)
    for i in range(cfg["n_layer"]):
        assert d[f"blk.{i}.attn_k"].type == W.Q8_0 and d[f"blk.{i}.attn_v"].type == W.Q8_0
        g = d[f"blk.{i}.ffn_gate_exps"]
        assert g.n_expert == 8 and g.n_used == 2 and g.streamed_bytes == 2 * g.m * W.row_bytes(g.type, g.k)
        assert d[f"blk.{i}.ffn_down_exps"].type == (W.Q6_K if W.use_more_bits(i, cfg["n_layer"]) else W.Q4_K)


@pytest.mark.parametrize("model,gb", [("llama-3.2-1b-q4_0", 0.763), ("llama-3-8b-q8_0", 7.974), ("llama-3-8b-iq4_xs", 4.139)])
def test_byte_model_of_the_other_configs(W, model, gb):
    assert abs(W.tg_bytes(W.llama_dense(**W.MODELS[model])) / 1e9 - gb) < 0.002          # SURVEY.md 8d figures
