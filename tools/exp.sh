timeout 200 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for f in "" "--key-mode dense" "--key-mode dense --no-efilter"; do
echo "== $f"; timeout 120 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-e2e $f 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(round(d['value']/1e9,2),'Gev/s ingest_ms',round(r['avg_launch_ms'],3),'fire_ms',round(r['fire_chain_ms'],3))"
done
