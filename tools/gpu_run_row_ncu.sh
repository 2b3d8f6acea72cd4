#!/bin/bash
# launch list (ncu, gpu__time_duration) of the last push of one row of tools/measure_configs.py: ROW="substring of the row name"
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
MEASURE_ONLY="$ROW" timeout -s KILL 300 python tools/measure_configs.py > gpurun_out/row_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/row_plain.log; exit 1; }
grep "^| " gpurun_out/row_plain.log | sort -u | cut -c1-160
MEASURE_ONLY="$ROW" timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fmq|median|ps_|scan|ingest|fire_kernel|plan_kernel|commit|keyed|resolve|xs_|dir_|set_mlog" -c 600 --csv \
   --log-file gpurun_out/row_launches.csv python tools/measure_configs.py > gpurun_out/row_ncu.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/row_launches.csv", errors="replace")) if len(r) > 5]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r); H = rows[hdr]; kn, mv = H.index("Kernel Name"), H.index("Metric Value")
seq = [(r[kn].split("(")[0][:70], float(r[mv].replace(",", ""))) for r in rows[hdr + 1:] if r[mv].replace(",", "").replace(".", "").isdigit()]
# the last push = the launches after the last plan kernel's predecessor chain: cut at the last "plan_kernel", walk back to the previous one
plans = [i for i, (k, _) in enumerate(seq) if k.startswith("plan_kernel")]
lo = plans[-2] + 1 if len(plans) > 1 else 0
while lo < len(seq) and (seq[lo][0].startswith("void fire") or seq[lo][0].startswith("commit") or "median" in seq[lo][0] or seq[lo][0].startswith("set_mlog") or seq[lo][0].startswith("fire")): lo += 1
agg = collections.OrderedDict()
for k, v in seq[lo:]:
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
for k, (c, v) in agg.items(): print(f"{k:72s} x{c:3d} {v / 1e3:10.1f} us")
print("total of the last push: %.1f us" % (sum(v for _, v in seq[lo:]) / 1e3))
PY
