#!/bin/bash
# llama-bench on an N-GPU box (gpurun --gpus N): Llama-3-8B Q4_K_M with -sm none / layer / row through this repository's plugin and through the
# reference's CUDA backend (oracle/_ref/cuda), plus the whole-graph logits check under -sm row (oracle/llama_logits_check.cpp).
#   tools/run_llama_bench_multi.sh [outdir] [model]
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="${1:-$ROOT/gpurun_out/llama_bench_multi}"
MODEL="${2:-llama-3-8b-q4_k_m}"
mkdir -p "$OUT"
VAR=avx512vnni
BIN="$ROOT/oracle/_ref/$VAR/llama-bench"
CHK="$ROOT/oracle/_ref/$VAR/llama_logits_check"
[ -x "$BIN" ] || { echo "$BIN missing (make -C oracle -f Makefile.llama)"; exit 4; }
TMP="${TMPDIR:-/tmp}"
nvidia-smi -L | tee "$OUT/gpus.txt"
python "$ROOT/tools/make_gguf.py" --model llama-3.2-1b-q4_0 --layers 4 --out "$TMP/tiny.gguf" > /dev/null 2>&1 || exit 5
python "$ROOT/tools/make_gguf.py" --model "$MODEL" --out "$TMP/m.gguf" > "$OUT/gguf.txt" 2>&1 || { cat "$OUT/gguf.txt"; exit 5; }
PLUG="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so"
echo "== logits check, -sm row over all GPUs (plugin vs reference CPU backend)"
for g in 1 0; do
    GGML_BACKEND_PATH="$PLUG" LLAMA_LOGITS_SPLIT=row GGML_B200_GRAPHS=$g timeout 600 "$CHK" "$TMP/tiny.gguf" 40 6 2>&1 | grep -E "Split|LLAMA_LOGITS_CHECK|nmse|failed|error" | tail -8 | tee -a "$OUT/logits_row.txt"
done
for sm in none layer row; do
    echo "== plugin, $MODEL, -sm $sm"
    GGML_BACKEND_PATH="$PLUG" timeout 900 "$BIN" -m "$TMP/m.gguf" -p 512 -n 128 -r 2 -ngl 99 -sm $sm 2>&1 | tee "$OUT/b200_$sm.txt" | tail -4
done
if [ -f "$ROOT/oracle/_ref/cuda/libggml-cuda.so" ]; then
    for sm in layer row; do
        echo "== reference CUDA backend, $MODEL, -sm $sm"
        GGML_BACKEND_PATH="$ROOT/oracle/_ref/cuda/libggml-cuda.so" timeout 900 "$BIN" -m "$TMP/m.gguf" -p 512 -n 128 -r 2 -ngl 99 -sm $sm 2>&1 | tee "$OUT/cuda_$sm.txt" | tail -4
    done
fi
rm -f "$TMP/m.gguf" "$TMP/tiny.gguf"
