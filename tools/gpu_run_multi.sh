#!/bin/bash
# multi-GPU run: N = number of GPUs on the box.  Multi-process parity tests (log kept), then bench.py at N.
cd "$(dirname "$0")/.."
N=${N:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout -s KILL 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/m${N}_smoke.log 2>&1 || { echo "SMOKE FAILED"; tail -30 gpurun_out/m${N}_smoke.log; exit 1; }
if [ "$SKIP_TESTS" != "1" ]; then
if [ -n "$PYTEST_IDS_FILE" ]; then mapfile -t IDS < "$PYTEST_IDS_FILE"; else IDS=(tests); fi
timeout -s KILL 1500 python -m pytest "${IDS[@]}" -m gpu -v --timeout 600 -rs ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/m${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/m${N}_pytest.log
fi
run_bench () {   # $1 = tag, rest = bench args
  tag=$1; shift
  timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N "$@" > gpurun_out/m${N}_bench_$tag.json 2> gpurun_out/m${N}_bench_$tag.err
  echo "bench $tag rc=$?"; tail -3 gpurun_out/m${N}_bench_$tag.err
}
run_bench main --steps 20 --warmup 3 $BENCH_ARGS
if [ "$EXTRA_ROUTES" = "1" ]; then
  run_bench exchange --steps 20 --warmup 3 --multi exchange --no-variants --no-e2e
fi
python - <<'PY'
import json,glob,os
N=os.environ.get("N","2")
for f in sorted(glob.glob(f"gpurun_out/m{N}_bench_*.json")):
    try:
        l=json.loads(open(f).read().strip().splitlines()[-1]); r=l["roofline"]
        print(f, "value %.2f G ev/s  step %.4f ms  ingest %.4f ms  fire %.4f  issue %.3f" % (l["value"]/1e9, l["ms_per_step"], r["avg_launch_ms"], r["fire_chain_ms"], r["host_issue_ms_per_step"]))
        for k,v in (l.get("variants") or {}).items(): print("   ", k[:70], {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ("value","ms_per_step","nvlink_GBps_per_gpu_each_direction")})
        if l.get("e2e"): print("    e2e %.3f G ev/s" % (l["e2e"]["value"]/1e9))
        print("    verify:", l.get("verify_vs_oracle"))
    except Exception as e: print(f, "unreadable", e)
PY
