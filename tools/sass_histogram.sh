#!/bin/bash
# SASS opcode histogram per kernel of the built library (evidence that the Blackwell paths are what runs): tcgen05 MMA (UTCHMMA / UTCIMMA /
# UTCQMMA ...), TMA (UTMALDG, UBLKCP), TMEM loads (LDTM), commit barriers (UTCBAR), integer dot products (IDP), warp-level MMA of the attention kernels (HMMA), programmatic dependent launch (ACQBULK), byte permutes of the table-lookup formats (PRMT).
#   tools/sass_histogram.sh > profiles/r2_sass_histogram.txt
LIB="$(cd "$(dirname "$0")/.." && pwd)/zyk-llama.cpp_b200/lib/libb200mm.so"
echo "# cuobjdump -sass $LIB | per-function counts of the opcodes of interest ($(date -u +%Y-%m-%d), commit $(git -C "$(dirname "$LIB")/../.." rev-parse --short HEAD))"
/usr/local/cuda/bin/cuobjdump -sass "$LIB" | awk '
/Function :/ { fn=$3; next }
{ if (match($0, /(UTC[A-Z]+(\.[A-Z0-9_]+)*|LDTM(\.[A-Za-z0-9_]+)*|UTMALDG(\.[A-Z0-9_]+)*|UBLKCP(\.[A-Z0-9_]+)*|IDP(\.[A-Z0-9_]+)*|SYNCS(\.[A-Z0-9_]+)*|UTMAPF[A-Z0-9_.]*|REDUX[A-Z0-9_.]*|HMMA(\.[A-Z0-9_]+)*|ACQBULK|PRMT)/)) { op=substr($0, RSTART, RLENGTH); c[fn" "op]++ } }
END { for (k in c) print c[k], k }' | sort -k2,2 -k1,1nr | c++filt | awk '{ n=$1; $1=""; printf "%6d %s\n", n, $0 }' | sed -E 's/\(anonymous namespace\):://g' | cut -c1-200
