#!/bin/bash
# The reference's own llama-bench (tools/llama-bench/llama-bench.cpp, unmodified, built into oracle/_ref by oracle/Makefile.llama) on
# random-init GGUFs of the BASELINE.json shapes (tools/make_gguf.py):
#   1. Llama-3.2-1B Q4_0 on the reference CPU backend (configs[0]: the fork's ZYK / repack kernels, all host threads)
#   2. Llama-3-8B Q4_K_M with this repository's plugin as the GPU device (GGML_BACKEND_PATH): every quantised MUL_MAT on the B200, the other
#      ops on the reference CPU backend (-nkvo 1: the KV cache stays on the host, this backend has no SET_ROWS / CPY)
#   3. the same file through the reference's CUDA backend (oracle/_ref/cuda), if it was built
#   tools/run_llama_bench.sh [outdir] [-r reps]
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="${1:-$ROOT/gpurun_out/llama_bench}"
mkdir -p "$OUT"
VAR=avx512vnni
grep -q avx512_vnni /proc/cpuinfo || { echo "host CPU lacks avx512_vnni: the prebuilt reference cannot run here"; exit 3; }
BIN="$ROOT/oracle/_ref/$VAR/llama-bench"
[ -x "$BIN" ] || { echo "$BIN missing (make -C oracle -f Makefile.llama)"; exit 4; }
TMP="${TMPDIR:-/tmp}"
T=$(nproc)
python "$ROOT/tools/make_gguf.py" --model llama-3.2-1b-q4_0 --out "$TMP/llama-3.2-1b-q4_0.gguf" > "$OUT/gguf_1b.txt" 2>&1 || { cat "$OUT/gguf_1b.txt"; exit 5; }
python "$ROOT/tools/make_gguf.py" --model llama-3-8b-q4_k_m --out "$TMP/llama-3-8b-q4_k_m.gguf" > "$OUT/gguf_8b.txt" 2>&1 || { cat "$OUT/gguf_8b.txt"; exit 5; }
cat "$OUT/gguf_1b.txt" "$OUT/gguf_8b.txt"
# ONLY_PLUGIN=1: legs 2 and 2b only (the plugin), for iterating on the kernels
if [ -z "${ONLY_PLUGIN:-}" ]; then
echo "== 1. reference CPU backend, Llama-3.2-1B Q4_0, $T threads"
env -u GGML_BACKEND_PATH "$BIN" -m "$TMP/llama-3.2-1b-q4_0.gguf" -p 512 -n 128 -r 2 -t "$T" -ngl 0 2>&1 | tee "$OUT/cpu_1b_q4_0.txt" | tail -6
fi
echo "== 2. this plugin as the GPU device, Llama-3-8B Q4_K_M (MUL_MAT on the B200, the rest on the CPU backend, KV cache on the host)"
GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so" timeout 900 "$BIN" -m "$TMP/llama-3-8b-q4_k_m.gguf" -p 512 -n 128 -r 5 -t "$T" -ngl 99 2>&1 | tee "$OUT/b200_8b_q4_k_m.txt" | tail -6
echo "== 2b. this plugin, Llama-3.2-1B Q4_0"
GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so" timeout 900 "$BIN" -m "$TMP/llama-3.2-1b-q4_0.gguf" -p 512 -n 128 -r 5 -ngl 99 2>&1 | tee "$OUT/b200_1b_q4_0.txt" | tail -4
if [ -z "${ONLY_PLUGIN:-}" ] && [ -f "$ROOT/oracle/_ref/cuda/libggml-cuda.so" ]; then
    echo "== 3. the reference's CUDA backend, Llama-3-8B Q4_K_M"
    GGML_BACKEND_PATH="$ROOT/oracle/_ref/cuda/libggml-cuda.so" timeout 900 "$BIN" -m "$TMP/llama-3-8b-q4_k_m.gguf" -p 512 -n 128 -r 3 -ngl 99 2>&1 | tee "$OUT/cuda_8b_q4_k_m.txt" | tail -6
    GGML_BACKEND_PATH="$ROOT/oracle/_ref/cuda/libggml-cuda.so" timeout 900 "$BIN" -m "$TMP/llama-3.2-1b-q4_0.gguf" -p 512 -n 128 -r 3 -ngl 99 2>&1 | tee "$OUT/cuda_1b_q4_0.txt" | tail -4
fi
rm -f "$TMP/llama-3.2-1b-q4_0.gguf" "$TMP/llama-3-8b-q4_k_m.gguf"
