#!/bin/bash
# launch list (this library's kernels only) and one full capture of the ingest and fire kernels of the default bench command (1 GPU)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
S="--steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-variants --no-clocks"
timeout -s KILL 300 python bench.py $S > gpurun_out/n_plain1.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"ingest|fire_kernel|plan_kernel|commit_kernel|xs_|scan_|keyed|ps_|resolve|retire|cmb_|widen|median|txt_|dict_" -c 60 --csv --log-file gpurun_out/n_launches.csv python bench.py $S > gpurun_out/n_ncu_list.log 2>&1
if [ "$FULL" = "1" ]; then
timeout -s KILL 300 python bench.py $S > gpurun_out/n_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"ingest_ring|fire_kernel" -s 6 -c 2 -o gpurun_out/n_prof python bench.py $S > gpurun_out/n_ncu_full.log 2>&1
fi
ls -la gpurun_out | grep " n_"
