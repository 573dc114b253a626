#!/bin/bash
# The reference's own perf harness (tests/test-backend-ops.cpp perf mode, unmodified, built into oracle/_ref) over the MUL_MAT and
# MUL_MAT_ID perf cases of the five block formats of this path -- once through the reference's CUDA backend (oracle/_ref/cuda/
# libggml-cuda.so, built natively for sm_100 by oracle/Makefile.cuda: SURVEY.md §8 a18, "the number to beat on the same box") and
# once through this repository's plugin. Both are loaded the way the reference loads any out-of-tree backend (GGML_BACKEND_PATH).
#   tools/compare_ggml_cuda.sh [outdir]      -> outdir/{cuda,b200}_{mul_mat,mul_mat_id}.txt + table on stdout (tools/compare_table.py)
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="${1:-$ROOT/gpurun_out}"
mkdir -p "$OUT"
BIN="$ROOT/oracle/_ref/avx512vnni/test-backend-ops"
grep -q avx512_vnni /proc/cpuinfo || { echo "host CPU lacks avx512_vnni: the prebuilt reference cannot run here"; exit 3; }
TYPES='type_a=(q4_0|q8_0|q4_K|q6_K|iq4_xs),'
# ONLY=b200: re-run this plugin only and reuse cuda_*.txt already in the output directory (the reference does not change)
for be in ${ONLY:-cuda b200}; do
    if [ $be = cuda ]; then lib="$ROOT/oracle/_ref/cuda/libggml-cuda.so"; else lib="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so"; fi
    [ -f "$lib" ] || { echo "$lib missing"; exit 4; }
    GGML_BACKEND_PATH="$lib" "$BIN" perf -o MUL_MAT -p "$TYPES" > "$OUT/${be}_mul_mat.txt" 2>&1
    GGML_BACKEND_PATH="$lib" "$BIN" perf -o MUL_MAT_ID -p "$TYPES" > "$OUT/${be}_mul_mat_id.txt" 2>&1
done
python "$ROOT/tools/compare_table.py" "$OUT"
