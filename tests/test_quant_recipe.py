"""The per-tensor quantisation recipe of zyk-llama.cpp_b200/workloads.py (which bench.py and tools/make_gguf.py build their random-init
models from) pinned against the reference's own quantiser: an f32 GGUF with the layer count, head ratio and expert count of each
BASELINE.json model (tiny widths) goes through llama_model_quantize of the unmodified reference (oracle/_ref, built by
oracle/Makefile.llama from /root/reference/src/llama-quant.cpp), and the type it gave every tensor is compared with
llama_dense(exact=True). CPU only; skipped where the reference could not be compiled."""
import importlib.util
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _tool():
    """the reference quantiser driver of the CPU variant this host can run (oracle.ref_dir), or None"""
    sys.path.insert(0, str(ROOT))
    from oracle.oracle import ref_dir
    d = ref_dir()
    exe = d / "llama_quantize_types" if d else None
    return exe if exe is not None and exe.exists() else None


TOOL = _tool()
FTYPE_ID = {"q4_0": 2, "q8_0": 7, "q4_k_m": 15, "iq4_xs": 30}          # enum llama_ftype (include/llama.h)
TYPE_NAME = {0: "f32", 2: "q4_0", 8: "q8_0", 12: "q4_K", 13: "q5_K", 14: "q6_K", 23: "iq4_xs"}
BASELINE_MODELS = ["llama-3.2-1b-q4_0", "llama-3-8b-q4_k_m", "llama-3-8b-q8_0", "llama-3-8b-iq4_xs", "llama-3-70b-q4_k_m", "mixtral-8x7b-q4_k_m"]


def _workloads():
    spec = importlib.util.spec_from_file_location("_b200_workloads_recipe", ROOT / "zyk-llama.cpp_b200" / "workloads.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod


def _tiny_cfg(W, model):
    cfg = dict(W.MODELS[model])
    cfg.update(n_embd=256, n_ff=256, n_vocab=256, n_head=cfg["n_head"] // cfg["n_kv"], n_kv=1)     # what make_gguf.py --tiny does
    return cfg


def _reference_types(tmp_path, model, ftype):
    src, dst = tmp_path / "f32.gguf", tmp_path / "quantised.gguf"
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "make_gguf.py"), "--model", model, "--tiny", "--f32", "--out", str(src)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout
    env = dict(os.environ)
    env.pop("GGML_BACKEND_PATH", None)
    r = subprocess.run([str(TOOL), str(src), str(dst), str(FTYPE_ID[ftype]), "8"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    return dict(line.split() for line in r.stdout.splitlines() if line.strip())


@pytest.mark.skipif(TOOL is None, reason="reference quantiser not built (oracle/Makefile.llama)")
@pytest.mark.parametrize("model", BASELINE_MODELS)
def test_exact_recipe_is_what_the_reference_quantiser_writes(tmp_path, model):
    W = _workloads()
    cfg = _tiny_cfg(W, model)
    ref = _reference_types(tmp_path, model, cfg["ftype"])
    ops = W.llama_dense(**cfg, exact=True)
    assert len(ops) == 7 * cfg["n_layer"] + 1
    wrong = {o.name: (TYPE_NAME[o.type], ref.get(o.name + ".weight")) for o in ops if ref.get(o.name + ".weight") != TYPE_NAME[o.type]}
    assert not wrong, f"{model}: (recipe, reference) per tensor: {wrong}"
    # what is not a MUL_MAT weight: norms and the router stay f32, the embedding table takes the base type of the file
    base = {"q4_0": "q4_0", "q8_0": "q8_0", "q4_k_m": "q4_K", "iq4_xs": "iq4_xs"}[cfg["ftype"]]
    assert ref["token_embd.weight"] == base
    assert all(t == "f32" for n, t in ref.items() if "norm" in n or "ffn_gate_inp" in n)
    assert len(ref) == len(ops) + 1 + 2 * cfg["n_layer"] + 1 + (cfg["n_layer"] if cfg.get("n_expert") else 0)


@pytest.mark.parametrize("model", BASELINE_MODELS)
def test_default_recipe_differs_from_the_exact_one_only_where_documented(model, monkeypatch):
    """the recipe the round's measurements were taken on: identical to llama-quantize's for the Q4_0, Q8_0 and dense Q4_K_M files (8B, 70B);
    IQ4_XS without its Q5_K attn_v / first-eighth ffn_down, Mixtral without its Q5_K attn_output (workloads.llama_dense docstring)"""
    monkeypatch.delenv("B200_EXACT_RECIPE", raising=False)
    W = _workloads()
    cfg = dict(W.MODELS[model])
    default, exact = W.llama_dense(**cfg), W.llama_dense(**cfg, exact=True)
    diff = {(d.name.split(".")[-1], TYPE_NAME[d.type], TYPE_NAME[e.type]) for d, e in zip(default, exact) if d.type != e.type}
    expect = {"llama-3-8b-iq4_xs": {("attn_v", "iq4_xs", "q5_K"), ("ffn_down", "iq4_xs", "q5_K")},
              "mixtral-8x7b-q4_k_m": {("attn_output", "q4_K", "q5_K")}}.get(model, set())
    assert diff == expect
    # and the difference is small in streamed bytes: below 1.5 % of a decode token
    assert abs(W.tg_bytes(exact) - W.tg_bytes(default)) <= 0.015 * W.tg_bytes(default)
