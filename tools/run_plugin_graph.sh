#!/bin/bash
# Multi-node decode graph ({wq,wk,wv} -> wo -> {gate,up} -> down) through the reference's public API on the reference CPU backend and on
# the plugin (oracle/plugin_graph_check.cpp, built by oracle/Makefile into oracle/_ref): exercises graph_compute's grouping and
# decode-chain paths. Runs the three plugin modes: chain, groups only, one call per node.
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
VAR=avx512vnni
grep -q avx512_vnni /proc/cpuinfo || { echo "host CPU lacks avx512_vnni: the prebuilt reference cannot run here"; exit 3; }
export GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so"
rc=0
for n in 1 3 64; do
  echo "== n=$n chain";  "$ROOT/oracle/_ref/$VAR/plugin_graph_check" $n || rc=1
done
echo "== n=1 groups only"; GGML_B200_CHAIN=0 "$ROOT/oracle/_ref/$VAR/plugin_graph_check" 1 || rc=1
echo "== n=1 per node";    GGML_B200_GROUP=0 "$ROOT/oracle/_ref/$VAR/plugin_graph_check" 1 || rc=1
exit $rc
