#!/bin/bash
# chain kernel with one 256-element unit per lane for IQ4_XS / Q4_0 against the eight-lanes-per-block fragments and the per-node launches
set -u
cd "$(dirname "$0")/.."
python -m pytest tests/test_gpu_chain.py -m gpu -x -q -k "iq4_xs or q4_0 or q5k" 2>&1 | tail -4
for m in llama-3-8b-iq4_xs llama-3.2-1b-q4_0 llama-3-8b-q8_0; do
    echo "== $m"
    python tools/chain_probe.py --model $m --steps 48 2>&1 | tail -12
    echo "-- B200_CHAIN_LANES8=1"
    B200_CHAIN_LANES8=1 python tools/chain_probe.py --model $m --steps 48 --only chain 2>&1 | tail -2
done
