"""What is compiled into the product library, read from its SASS (no GPU needed): sm_100a cubins only, and the Blackwell paths DESIGN.md
names are the ones in the code -- tcgen05 MMA with TMEM loads and TMA tensor loads in the prefill GEMM, TMA / bulk copies and integer dot
products in the decode chain, warp-level tensor-core MMA in the attention kernels. The mnemonics are those of
/opt/skills/guides/B200_PROFILING.md; tools/sass_histogram.sh writes the full per-kernel table (profiles/r2_sass_histogram.txt)."""
import re
import shutil
import subprocess
from collections import Counter
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "zyk-llama.cpp_b200" / "lib" / "libb200mm.so"
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
needs_tools = pytest.mark.skipif(not LIB.exists() or not Path(CUOBJDUMP).exists(), reason="built library or cuobjdump not present")


@needs_tools
def test_only_sm_100a_code_is_in_the_library():
    out = subprocess.run([CUOBJDUMP, "--list-elf", str(LIB)], stdout=subprocess.PIPE, text=True, timeout=120).stdout
    elfs = re.findall(r"ELF file\s+\d+:\s+(\S+)", out)
    assert len(elfs) >= 11, out
    assert all(e.endswith(".sm_100a.cubin") for e in elfs), elfs
    ptx = subprocess.run([CUOBJDUMP, "--list-ptx", str(LIB)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120).stdout
    assert all("sm_100a" in line for line in ptx.splitlines() if line.startswith("PTX file")), ptx          # (none embedded: code=sm_100a only)


def _opcodes(tmp_path, unit):
    """opcode histogram of one translation unit's cubin (a full dump of the library takes most of a minute)"""
    name = f"{unit}.sm_100a.cubin"
    r = subprocess.run([CUOBJDUMP, "-xelf", name, str(LIB)], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert (tmp_path / name).exists(), r.stdout
    sass = subprocess.run([CUOBJDUMP, "-sass", name], cwd=tmp_path, stdout=subprocess.PIPE, text=True, timeout=600).stdout
    ops = Counter(re.findall(r"\b(UTCHMMA|UTCIMMA|UTCBAR|UTMALDG|UBLKCP|LDTM|HMMA\.16816|IDP\.4A|ACQBULK|SYNCS)\b", sass))
    funcs = re.findall(r"Function : (\S+)", sass)
    return ops, funcs, sass


@needs_tools
def test_prefill_gemm_is_tcgen05_with_tma_and_tmem(tmp_path):
    ops, funcs, sass = _opcodes(tmp_path, "gemm")
    assert ops["UTCHMMA"] > 0 and ops["UTMALDG"] > 0 and ops["LDTM"] > 0 and ops["UTCBAR"] > 0, ops
    assert ops["HMMA.16816"] == 0, "the GEMM must not fall back to warp-level mma.sync"
    assert "UTCHMMA.2CTA" in sass, "CTA-pair (cta_group::2) MMA expected in gemm_pair_kernel"
    assert any("gemm_pair_kernel" in f for f in funcs) and any("gemm_moe" in f for f in funcs), funcs[:8]


@needs_tools
def test_decode_chain_streams_with_bulk_copies_and_integer_dot_products(tmp_path):
    ops, funcs, _ = _opcodes(tmp_path, "gemv_chain")
    assert ops["UBLKCP"] + ops["UTMALDG"] > 0 and ops["IDP.4A"] > 0 and ops["SYNCS"] > 0, ops
    for ks in range(4):                                   # the four item-kind sets of the persistent kernel
        assert any(f"gemv_chain_kernelILi1ELi{ks}E" in f for f in funcs), (ks, [f for f in funcs if "gemv_chain_kernel" in f][:8])


@needs_tools
def test_attention_kernels_use_tensor_core_mma_and_glue_kernels_programmatic_launch(tmp_path):
    ops, funcs, _ = _opcodes(tmp_path, "glue_ops")
    assert ops["HMMA.16816"] > 0 and ops["ACQBULK"] > 0, ops
    assert any("attn_prefill_kernel" in f for f in funcs) and any("mm_mma_kernel" in f for f in funcs), funcs[:8]
