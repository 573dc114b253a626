"""GPU (-m gpu): the ggml backend plugin (libggml-b200.so) driven through the reference's own public API.

* the reference's unmodified conformance harness (tests/test-backend-ops.cpp, built by oracle/Makefile into oracle/_ref) on a
  slice of its MUL_MAT cases, loaded the way the reference loads any out-of-tree backend (GGML_BACKEND_PATH);
* a multi-node decode graph ({wq,wk,wv} -> wo -> {gate,up} -> down) on the reference CPU backend and on the plugin
  (oracle/plugin_graph_check.cpp): graph_compute's grouping (b200_mul_mat_group) and decode-chain (b200_chain_*) paths.
Skipped where oracle/_ref was not built for the host CPU (the reference binaries need avx512_vnni)."""
import os
import subprocess
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
VAR = ROOT / "oracle" / "_ref" / "avx512vnni"


def _need(binary):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    if not (VAR / binary).exists() or "avx512_vnni" not in open("/proc/cpuinfo").read():
        pytest.skip("oracle/_ref not built for this host CPU")
    plugin = ROOT / "zyk-llama.cpp_b200" / "lib" / "libggml-b200.so"
    assert plugin.exists(), "the plugin is not built (no fallback)"
    return dict(os.environ, GGML_BACKEND_PATH=str(plugin))


CASES = [({}, 1), ({}, 3), ({}, 64), ({"GGML_B200_CHAIN": "0"}, 1), ({"GGML_B200_GROUP": "0"}, 1)]


@pytest.mark.parametrize("mode,n", CASES, ids=["chain-n1", "chain-n3", "gemm-n64", "groups-n1", "per-node-n1"])
def test_multi_node_graph_matches_reference_cpu_backend(mode, n):
    env = _need("plugin_graph_check")
    env.update(mode)
    r = subprocess.run([str(VAR / "plugin_graph_check"), str(n)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "PLUGIN_GRAPH_CHECK OK" in r.stdout, r.stdout[-3000:]


@pytest.mark.parametrize("op", ["MUL_MAT", "MUL_MAT_ID"])
def test_reference_conformance_harness_all_types(op):
    """The reference's unmodified test-backend-ops, every case of the op (all weight types; tests/test-backend-ops.cpp:7616-7729,
    7768-7820): the cases supports_op declines are reported "not supported" by the harness, every case it accepts must pass."""
    env = _need("test-backend-ops")
    r = subprocess.run([str(VAR / "test-backend-ops"), "test", "-o", op], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    tail = r.stdout[-2000:]
    assert r.returncode == 0 and "tests passed" in tail and "FAIL" not in r.stdout, tail
    import re
    m = re.search(r"(\d+)/(\d+) tests passed", tail)
    assert m and m.group(1) == m.group(2) and int(m.group(2)) >= 100, tail
    # every one of the six block formats of the path was actually exercised (not merely declined)
    for t in ("q4_0", "q8_0", "q4_K", "q5_K", "q6_K", "iq4_xs"):
        ok = [ln for ln in r.stdout.splitlines() if f"type_a={t}," in ln and "OK" in ln]
        assert ok, f"no passing {op} case with type_a={t}"


# ---- a whole llama graph through the reference's loader and llama_decode: plugin vs the reference CPU backend ------------------------------
LLAMA_MODES = [({}, "default"), ({"GGML_B200_FUSE": "0", "GGML_B200_HOIST": "0"}, "no-fusion"), ({"GGML_B200_GRAPHS": "0"}, "no-cuda-graphs")]


@pytest.fixture(scope="module")
def tiny_gguf(tmp_path_factory):
    """4 layers of Llama-3.2-1B shapes (Q4_0 + Q6_K output), random-init, written by tools/make_gguf.py"""
    import sys
    path = tmp_path_factory.mktemp("gguf") / "tiny-1b-q4_0.gguf"
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "make_gguf.py"), "--model", "llama-3.2-1b-q4_0", "--layers", "4", "--out", str(path)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    return path


@pytest.mark.parametrize("mode,name", LLAMA_MODES, ids=[m[1] for m in LLAMA_MODES])
def test_whole_llama_graph_logits_match_reference_cpu_backend(tiny_gguf, mode, name):
    """oracle/llama_logits_check.cpp: prefill of 40 tokens + 6 decoded tokens, logits of every step, NMSE <= 2e-3 (the reference's own
    whole-model tolerance). Exercises the loader into B200 buffers (repack), every glue op, the KV cache on the device, the fused norm /
    attention kernels, the hoisted q / k / v group and the whole-graph CUDA graph replay."""
    env = _need("llama_logits_check")
    env.update(mode)
    r = subprocess.run([str(VAR / "llama_logits_check"), str(tiny_gguf), "40", "6"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0 and "LLAMA_LOGITS_CHECK OK" in r.stdout, r.stdout[-3000:]


def _gpu_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("graphs", ["1", "0"], ids=["cuda-graphs", "eager"])
def test_row_split_buffer_type_logits_match_reference_cpu_backend(tiny_gguf, graphs):
    """`-sm row` through the reference's loader: llama-model.cpp:388-410 asks the registry for "ggml_backend_split_buffer_type" and puts the
    MUL_MAT weights there. One GPU: every row stays on the main device, the split buffer's init_tensor / set_tensor / compute path is
    still the one that runs. Two or more GPUs (gpurun --gpus N): rows are spread, Y arrives by peer stores. Same logits bar as above."""
    env = _need("llama_logits_check")
    env.update({"LLAMA_LOGITS_SPLIT": "row", "GGML_B200_GRAPHS": graphs})
    r = subprocess.run([str(VAR / "llama_logits_check"), str(tiny_gguf), "40", "6"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0 and "LLAMA_LOGITS_CHECK OK" in r.stdout, r.stdout[-3000:]
    assert "_Split" in r.stdout, "the loader did not place any tensor in the split buffer type:\n" + r.stdout[-3000:]


@pytest.fixture(scope="module")
def tiny_kquant_gguf(tmp_path_factory):
    """layers 3..5 of Llama-3-8B in the Q4_K_M recipe (one use_more_bits layer with Q6_K attn_v / ffn_down, two without; Q6_K output), random-init"""
    import sys
    path = tmp_path_factory.mktemp("gguf") / "tiny-8b-q4_k_m.gguf"
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "make_gguf.py"), "--model", "llama-3-8b-q4_k_m", "--layers", "3", "--layer-offset", "3", "--out", str(path)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:]
    return path


KQ_MODES = [({}, "default"), ({"GGML_B200_CHAIN_SINGLE": "0"}, "per-node-gemv"), ({"GGML_B200_FUSE": "0"}, "no-fusion"), ({"GGML_B200_PDL": "0", "B200_GLUE_PDL": "0"}, "no-pdl")]


@pytest.mark.parametrize("mode,name", KQ_MODES, ids=[m[1] for m in KQ_MODES])
def test_whole_kquant_graph_logits_match_reference_cpu_backend(tiny_kquant_gguf, mode, name):
    """The K-quant decode path of the plugin in a real graph: every MUL_MAT group is one chain-kernel launch (a chain of a single group), GLU goes
    into the quantiser of ffn_down, rope + KV-cache writes are one launch, prefill attention is the fused tensor-core kernel -- logits of
    llama_decode (40-token prompt + 6 decoded tokens) against the reference CPU backend, and the same with each of those switched off."""
    env = _need("llama_logits_check")
    env.update(mode)
    r = subprocess.run([str(VAR / "llama_logits_check"), str(tiny_kquant_gguf), "40", "6"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert r.returncode == 0 and "LLAMA_LOGITS_CHECK OK" in r.stdout, r.stdout[-3000:]


@pytest.fixture(scope="module")
def tiny_moe_gguf(tmp_path_factory):
    """2 layers of Mixtral-8x7B shapes (8 experts, 2 used; Q4_K_M recipe: Q8_0 attn_k / attn_v, Q6_K on the use_more_bits layers), random-init"""
    import sys
    path = tmp_path_factory.mktemp("gguf") / "tiny-mixtral.gguf"
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "make_gguf.py"), "--model", "mixtral-8x7b-q4_k_m", "--layers", "2", "--layer-offset", "3", "--out", str(path)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:]
    return path


def test_whole_mixtral_graph_logits_match_reference_cpu_backend(tiny_moe_gguf):
    """build_moe_ffn on the device: router matmul (f32), SOFT_MAX, ARGSORT + view (top-k), GET_ROWS of the probabilities, SUM_ROWS / CLAMP / DIV,
    MUL_MAT_ID x 3, SWIGLU, MUL by the weights, the ADDs over the used experts -- logits of llama_decode against the reference CPU backend"""
    env = _need("llama_logits_check")
    r = subprocess.run([str(VAR / "llama_logits_check"), str(tiny_moe_gguf), "40", "6"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert r.returncode == 0 and "LLAMA_LOGITS_CHECK OK" in r.stdout, r.stdout[-3000:]


GLUE_OPS = ["GET_ROWS", "RMS_NORM", "RMS_NORM_MUL_ADD", "ADD", "MUL", "DIV", "SCALE", "CLAMP", "SUM_ROWS", "ARGSORT", "ROPE", "SET_ROWS", "CPY", "CONT", "DUP", "SOFT_MAX",
            "SWIGLU", "GEGLU", "REGLU"]


@pytest.mark.parametrize("op", GLUE_OPS)
def test_reference_conformance_harness_glue_ops(op):
    """The ops of a llama-arch graph between the quantised matmuls (csrc/glue_ops.cu): the reference's unmodified test-backend-ops on every
    variant supports_op claims (the others are reported "not supported" and stay with the CPU backend)."""
    env = _need("test-backend-ops")
    r = subprocess.run([str(VAR / "test-backend-ops"), "test", "-o", op, "-b", "B200_0"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    import re
    m = re.search(r"(\d+)/(\d+) tests passed", r.stdout)
    assert r.returncode == 0 and m and m.group(1) == m.group(2) and int(m.group(2)) >= 1 and "FAIL" not in r.stdout, r.stdout[-2500:]
