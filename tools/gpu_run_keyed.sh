#!/bin/bash
# keyed chain and window median: their tests, then the two rows of the per-config table
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -m gpu -q -x --timeout 300 -k "median or keyed or consec or rolling or a7 or fullsize or heavy or known" 2>&1 | tail -3
for row in "C2 3-consecutive, 100 K keys, dense" "N3 10 s"; do MEASURE_ONLY="$row" timeout -s KILL 300 python tools/measure_configs.py 2>&1 | grep "^| [CN]" | sort -u | cut -c1-120; done
