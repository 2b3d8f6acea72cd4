#!/bin/bash
# ncu --set full of ONE launch of a kernel inside a row of tools/measure_configs.py
# ROW="substring of the row name"  KERNEL="regex"  SKIP=launches of that kernel to skip  OUT=name
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
MEASURE_ONLY="$ROW" timeout -s KILL 300 python tools/measure_configs.py > gpurun_out/row_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/row_plain.log; exit 1; }
MEASURE_ONLY="$ROW" timeout -s KILL 900 ncu --set full --import-source on --clock-control none -k regex:"$KERNEL" --launch-skip ${SKIP:-4} --launch-count 1 -f -o gpurun_out/${OUT:-row_kernel} \
   python tools/measure_configs.py > gpurun_out/row_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/row_ncu_full.log; ls -la gpurun_out/${OUT:-row_kernel}.ncu-rep
