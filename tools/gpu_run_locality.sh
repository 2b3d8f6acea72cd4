#!/bin/bash
# hash-range split in front of the ingest (tables much larger than the L2): full-size parity, then C4 / C5 with and without it
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -x --timeout 600 > gpurun_out/loc_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/loc_pytest.log
for mb in 96 0; do
  for row in "C4 60 s" "C5 Zipf"; do
    echo "FMQ_LOCALITY_MIN_MB=$mb  $row"
    FMQ_LOCALITY_MIN_MB=$mb MEASURE_ONLY="$row" timeout -s KILL 300 python tools/measure_configs.py 2>&1 | grep "^| C" | sort -u | cut -c1-200
  done
done 2>&1 | tee gpurun_out/loc_measure.log
