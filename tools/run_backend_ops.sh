#!/bin/bash
# Drive the UNMODIFIED reference conformance harness (tests/test-backend-ops.cpp, built from
# /root/reference by oracle/Makefile into oracle/_ref) against the plugin, loaded the way the
# reference loads any out-of-tree backend: GGML_BACKEND_PATH (ggml-backend-reg.cpp:627-631).
#   tools/run_backend_ops.sh test|perf|support [extra test-backend-ops args...]
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
VAR=avx512vnni
grep -q avx512_vnni /proc/cpuinfo || { echo "host CPU lacks avx512_vnni: the prebuilt reference cannot run here"; exit 3; }
export GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so"
MODE="${1:-test}"; shift || true
exec "$ROOT/oracle/_ref/$VAR/test-backend-ops" "$MODE" "$@"
