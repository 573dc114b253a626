#!/bin/bash
# ncu launch lists (gpu__time_duration, no clock control) of one llama-bench pass through the plugin: a 4-layer cut of Llama-3-8B Q4_K_M, prefill
# (pp512) and decode (tg16) separately. Output: <outdir>/llama_pp_launches.csv, llama_tg_launches.csv and a per-kernel summary of each.
set -u
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="${1:-$ROOT/gpurun_out/llama_launches}"
mkdir -p "$OUT"
BIN="$ROOT/oracle/_ref/avx512vnni/llama-bench"
TMP="${TMPDIR:-/tmp}"
python "$ROOT/tools/make_gguf.py" --model "${MODEL:-llama-3-8b-q4_k_m}" --layers "${LAYERS:-4}" --out "$TMP/cut.gguf" > /dev/null 2>&1 || exit 5
export GGML_BACKEND_PATH="$ROOT/zyk-llama.cpp_b200/lib/libggml-b200.so"
# plain runs first (the profiled command must have exited 0 without ncu)
"$BIN" -m "$TMP/cut.gguf" -p 512 -n 0 -r 1 -ngl 99 > "$OUT/plain_pp.txt" 2>&1 || { tail -5 "$OUT/plain_pp.txt"; exit 6; }
"$BIN" -m "$TMP/cut.gguf" -p 0 -n 16 -r 1 -ngl 99 > "$OUT/plain_tg.txt" 2>&1 || { tail -5 "$OUT/plain_tg.txt"; exit 6; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file "$OUT/llama_pp_launches.csv" "$BIN" -m "$TMP/cut.gguf" -p 512 -n 0 -r 1 -ngl 99 > "$OUT/ncu_pp.txt" 2>&1
GGML_B200_GRAPHS=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file "$OUT/llama_tg_launches.csv" "$BIN" -m "$TMP/cut.gguf" -p 0 -n 16 -r 1 -ngl 99 > "$OUT/ncu_tg.txt" 2>&1
for f in pp tg; do
python - "$OUT/llama_${f}_launches.csv" > "$OUT/llama_${f}_summary.txt" <<'PY'
import csv, sys, collections, re
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
tot = collections.Counter(); cnt = collections.Counter()
for r in rows[1:]:
    name = re.sub(r"<.*", "", r[ki]); v = float(r[vi].replace(",", "")); v = v / 1e3 if r[ui] == "ns" else v
    tot[name] += v; cnt[name] += 1
s = sum(tot.values())
print(f"total {s:.1f} us over {sum(cnt.values())} launches")
for k, v in tot.most_common(25): print(f"{v:10.1f} us {100 * v / s:5.1f} %  n={cnt[k]:5d}  {k}")
PY
cat "$OUT/llama_${f}_summary.txt"
done
rm -f "$TMP/cut.gguf"
