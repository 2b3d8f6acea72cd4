#!/bin/bash
# Round-2 single-GPU run A: smoke, full GPU test suite, bench (new ingest kernel, legacy A/B, tile variants), launch list.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/a_smoke.log 2>&1 || { echo "SMOKE FAILED"; tail -30 gpurun_out/a_smoke.log; exit 1; }
echo "smoke ok"
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/a_pytest.log
B="--steps 20 --warmup 3"
timeout -s KILL 600 python bench.py $B > gpurun_out/a_bench_n1.json 2> gpurun_out/a_bench_n1.err; echo "bench rc=$?"
Q="$B --no-e2e --no-cpu-baseline --no-variants"
FMQ_INGEST_LEGACY=1 timeout -s KILL 300 python bench.py $Q > gpurun_out/a_bench_legacy.json 2>&1
for v in w4s2b5 w8s2b3 w4s2b6; do
  FMQ_LIB_PATH=$PWD/build/variants/libfmq_$v.so FMQ_SKIP_BUILD=1 timeout -s KILL 300 python bench.py $Q > gpurun_out/a_bench_$v.json 2>&1
done
timeout -s KILL 300 python bench.py $Q --key-mode hashed > gpurun_out/a_bench_hashed.json 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/a_bench_*.json")):
    try:
        l=json.loads(open(f).read().strip().splitlines()[-1]); r=l["roofline"]
        print(f, "value %.2f G ev/s  step %.4f ms  ingest %.4f ms  frac %.3f  fire %.4f" % (l["value"]/1e9, l["ms_per_step"], r["avg_launch_ms"], r["frac"], r["fire_chain_ms"]))
        if l.get("variants"):
            for k,v in l["variants"].items(): print("   ", k[:60], {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ("value","ms_per_step","ingest_kernel_ms","roofline_frac")})
        if l.get("e2e"): print("    e2e %.3f G ev/s, packed %.3f" % (l["e2e"]["value"]/1e9, (l.get("e2e_packed") or {"value":0})["value"]/1e9))
    except Exception as e: print(f, "unreadable", e)
PY
# launch list of the same short command (after it has exited 0 plain)
S="--steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-variants --no-clocks"
timeout -s KILL 300 python bench.py $S > gpurun_out/a_plain_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/a_launches.csv python bench.py $S > gpurun_out/a_ncu_list.log 2>&1
timeout -s KILL 300 python bench.py $S > gpurun_out/a_plain_short2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:ingest_ring -s 4 -c 1 -o gpurun_out/a_prof_ring python bench.py $S > gpurun_out/a_ncu_full.log 2>&1
ls -la gpurun_out | tail -20
