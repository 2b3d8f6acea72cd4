#!/bin/bash
# single-GPU run: smoke, GPU test suite, bench of the default build and of every library under build/variants
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; rm -f gpurun_out/b_*
timeout -s KILL 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/b_smoke.log 2>&1 || { echo "SMOKE FAILED"; tail -30 gpurun_out/b_smoke.log; exit 1; }
echo "smoke ok"
if [ "$SKIP_TESTS" != "1" ]; then
timeout -s KILL 1800 python -m pytest tests -m gpu -q --timeout 900 ${PYTEST_ARGS:--x} > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?"; tail -${PYTEST_TAIL:-15} gpurun_out/b_pytest.log
fi
B="--steps 20 --warmup 3"
Q="$B --no-e2e --no-cpu-baseline --no-variants"
if [ "$FULL_BENCH" = "1" ]; then timeout -s KILL 600 python bench.py $B > gpurun_out/b_bench_n1.json 2> gpurun_out/b_bench_n1.err; echo "bench rc=$?";
else timeout -s KILL 300 python bench.py $Q > gpurun_out/b_bench_default.json 2>&1; fi
shopt -s nullglob
for f in build/variants/libfmq_*.so; do
  v=$(basename $f .so); v=${v#libfmq_}
  FMQ_LIB_PATH=$PWD/$f timeout -s KILL 300 python bench.py $Q $VARIANT_ARGS > gpurun_out/b_bench_$v.json 2>&1
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/b_bench_*.json")):
    try:
        l=json.loads(open(f).read().strip().splitlines()[-1]); r=l["roofline"]
        print(f, "value %.2f G ev/s  step %.4f ms  ingest %.4f ms  frac %.3f  fire %.4f" % (l["value"]/1e9, l["ms_per_step"], r["avg_launch_ms"], r["frac"], r["fire_chain_ms"]))
        if l.get("variants"):
            for k,v in l["variants"].items(): print("   ", k[:60], {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ("value","ms_per_step","ingest_kernel_ms","roofline_frac")})
        if l.get("e2e"): print("    e2e %.3f G ev/s, packed %.3f" % (l["e2e"]["value"]/1e9, (l.get("e2e_packed") or {"value":0})["value"]/1e9))
    except Exception as e: print(f, "unreadable", e)
PY
