#!/bin/bash
# The reference's llama-bench on a random-init GGUF of ANY model of zyk-llama.cpp_b200/workloads.py (tools/make_gguf.py), through this repository's
# plugin and, where it was built, through the reference's CUDA backend:
#   tools/run_llama_bench_model.sh <model> [outdir] [layers]        e.g. mixtral-8x7b-q4_k_m, llama-3-70b-q4_k_m, llama-3-8b-iq4_xs
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
MODEL="${1:?model name}"
OUT="${2:-$ROOT/gpurun_out/llama_bench_$MODEL}"
LAYERS="${3:-0}"
mkdir -p "$OUT"
BIN="$ROOT/oracle/_ref/avx512vnni/llama-bench"
[ -x "$BIN" ] || { echo "$BIN missing (make -C oracle -f Makefile.llama)"; exit 4; }
TMP="${TMPDIR:-/tmp}"
df -h "$TMP" | tail -1
python "$ROOT/tools/make_gguf.py" --model "$MODEL" --layers "$LAYERS" --out "$TMP/m.gguf" > "$OUT/gguf.txt" 2>&1 || { tail -5 "$OUT/gguf.txt"; exit 5; }
tail -3 "$OUT/gguf.txt"
echo "== plugin, $MODEL"
GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so" timeout 1200 "$BIN" -m "$TMP/m.gguf" -p 512 -n 128 -r 2 -ngl 99 2>&1 | tee "$OUT/b200.txt" | grep -E "^\| |error|failed" | tail -4
if [ -z "${ONLY_PLUGIN:-}" ] && [ -f "$ROOT/oracle/_ref/cuda/libggml-cuda.so" ]; then
    echo "== reference CUDA backend, $MODEL"
    GGML_BACKEND_PATH="$ROOT/oracle/_ref/cuda/libggml-cuda.so" timeout 1200 "$BIN" -m "$TMP/m.gguf" -p 512 -n 128 -r 2 -ngl 99 2>&1 | tee "$OUT/cuda.txt" | grep -E "^\| |error|failed" | tail -4
fi
rm -f "$TMP/m.gguf"
