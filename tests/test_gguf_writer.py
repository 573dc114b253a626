"""tools/make_gguf.py against the reference's own loader: the random-init GGUFs that llama-bench, the plugin tests and bench.py use must load and run in
the unmodified reference (llama-bench from oracle/_ref on its CPU backend, no plugin, no GPU). Covers the 70B recipe with its Q5_K attn_v."""
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _bench_bin():
    from oracle.oracle import ref_dir
    d = ref_dir()
    exe = d / "llama-bench" if d else None
    if exe is None or not exe.exists():
        pytest.skip("oracle/_ref/*/llama-bench not built for this host CPU")
    return exe


@pytest.mark.parametrize("model,layers,offset,expect,extra", [("llama-3.2-1b-q4_0", 2, 0, "q4_0", []), ("llama-3-70b-q4_k_m", 1, 10, "q5_K", []),
                                                             ("mixtral-8x7b-q4_k_m", 1, 0, "q8_0", []), ("llama-3-8b-iq4_xs", 1, 0, "q5_K", ["--exact-recipe"])])
def test_reference_loader_runs_the_random_init_gguf(tmp_path, ref, model, layers, offset, expect, extra):
    exe = _bench_bin()
    gguf = tmp_path / "m.gguf"
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "make_gguf.py"), "--model", model, "--layers", str(layers), "--layer-offset", str(offset), "--out", str(gguf)] + extra,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:]
    assert f"'{expect}'" in r.stdout, r.stdout[-1500:]          # the type histogram the script prints
    env = {k: v for k, v in __import__("os").environ.items() if k != "GGML_BACKEND_PATH"}
    r = subprocess.run([str(exe), "-m", str(gguf), "-p", "8", "-n", "2", "-r", "1", "-ngl", "0", "-t", "8"], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    gguf.unlink()
    assert r.returncode == 0 and "pp8" in r.stdout and "tg2" in r.stdout, r.stdout[-2500:]
