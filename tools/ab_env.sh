#!/bin/bash
# A/B of one environment switch on llama-bench tg128 / pp512 (8B Q4_K_M, plugin): tools/ab_env.sh VAR=VALUE [pp]
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
BIN="$ROOT/oracle/_ref/avx512vnni/llama-bench"
python "$ROOT/tools/make_gguf.py" --model "${MODEL:-llama-3-8b-q4_k_m}" --out /tmp/m.gguf > /dev/null 2>&1
P=0; N=128; [ "${2:-}" = "pp" ] && { P=512; N=0; }
for i in 1 2; do
  echo "base";  GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so" "$BIN" -m /tmp/m.gguf -p $P -n $N -r 3 -ngl 99 2>&1 | grep -E "tg128|pp512"
  echo "$1";    env "$1" GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so" "$BIN" -m /tmp/m.gguf -p $P -n $N -r 3 -ngl 99 2>&1 | grep -E "tg128|pp512"
done
rm -f /tmp/m.gguf
