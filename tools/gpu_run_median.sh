cd /root/repo; mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -m gpu -q -x --timeout 300 -k "median" > gpurun_out/med_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/med_pytest.log
MEASURE_ONLY=median timeout -s KILL 300 python tools/measure_configs.py > gpurun_out/med_measure.log 2>&1; tail -4 gpurun_out/med_measure.log
