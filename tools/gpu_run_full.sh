#!/bin/bash
# one GPU: smoke + whole GPU test suite + bench (gpu_run_b.sh), then the per-config table (measure_configs.py)
cd "$(dirname "$0")/.."
bash tools/gpu_run_b.sh || exit 1
timeout -s KILL 600 python tools/measure_configs.py > gpurun_out/configs.log 2>&1; echo "measure rc=$?"
sed -n '/^| config/,$p' gpurun_out/configs.log | cut -c1-200
