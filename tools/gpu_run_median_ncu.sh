#!/bin/bash
# launch list (ncu, gpu__time_duration) of the window-median step; run only after tools/gpu_run_median.sh exited 0
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
MEASURE_ONLY=median timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"median|ps_|scan|ingest|fire_kernel|plan_kernel|commit|set_mlog" -c 400 --csv \
   --log-file gpurun_out/med_launches.csv python tools/measure_configs.py > gpurun_out/med_ncu.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/med_launches.csv", errors="replace")) if len(r) > 5]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r); H = rows[hdr]; kn, mv = H.index("Kernel Name"), H.index("Metric Value")
seq = [(r[kn].split("(")[0][:60], float(r[mv].replace(",", ""))) for r in rows[hdr + 1:] if r[mv].replace(",", "").replace(".", "").isdigit()]
# the last push = the launches after the last ingest kernel
last = max(i for i, (k, _) in enumerate(seq) if "ingest" in k)
agg = collections.OrderedDict()
for k, v in seq[last:]:
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
for k, (c, v) in agg.items(): print(f"{k:62s} x{c:3d} {v / 1e3:10.1f} us")
print("total of the last push: %.1f us" % (sum(v for _, v in seq[last:]) / 1e3))
PY
